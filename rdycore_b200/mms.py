"""Method-of-manufactured-solutions harness for the SWE path (config B of BASELINE.json).

Host-side INPUT GENERATION only, as in the reference's MMS driver (src/rdymms.c): the analytic forms of
driver/tests/swe_roe/mms_conv_study.yaml:12-46 evaluated with numpy (muParser there), the source terms of
RDyMMSComputeSourceTerms (src/rdymms.c:560-582), the Dirichlet values of RDyMMSEnforceBoundaryConditions (:708-745),
the error norms of RDyMMSComputeErrorNorms (:850-903) and the rate fit of RDyMMSEstimateConvergenceRates (:973-995).
The right-hand side itself runs on the GPU (or, in tests, in the CPU oracle) — nothing here is on the hot path.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

GRAVITY = 9.806  # src/swe/swe_types_petsc.h:7


@dataclass
class Forms:
    """mms.constants of mms_conv_study.yaml"""
    H: float = 0.005
    T: float = 20.0
    U: float = 0.025
    V: float = 0.025
    N: float = 0.01
    Z: float = 0.0025
    K: float = 0.6283185307179586

    # mms.swe: functions of x, y, t
    def fields(self, x, y, t):
        H, T, U, V, K = self.H, self.T, self.U, self.V, self.K
        sx, cx, sy, cy, et = np.sin(K * x), np.cos(K * x), np.sin(K * y), np.cos(K * y), np.exp(t / T)
        return dict(
            h=H * (1 + sx * sy) * et, dhdx=H * K * sy * cx * et, dhdy=H * K * sx * cy * et, dhdt=H / T * (1 + sx * sy) * et,
            u=U * cx * sy * et, dudx=-U * K * sx * sy * et, dudy=U * K * cx * cy * et, dudt=U / T * cx * sy * et,
            v=V * sx * cy * et, dvdx=K * V * cx * cy * et, dvdy=-K * V * sx * sy * et, dvdt=V / T * sx * cy * et,
        )

    def z(self, x, y):
        return self.Z * np.sin(self.K * x) * np.sin(self.K * y)

    def dzdx(self, x, y):
        return self.Z * self.K * np.cos(self.K * x) * np.sin(self.K * y)

    def dzdy(self, x, y):
        return self.Z * self.K * np.sin(self.K * x) * np.cos(self.K * y)

    def mannings(self, x, y):
        return self.N * (1 + np.sin(self.K * x) * np.sin(self.K * y))

    def solution(self, x, y, t) -> np.ndarray:
        """[n, 3] (h, hu, hv): RDyMMSComputeSolution, src/rdymms.c:352-487"""
        f = self.fields(x, y, t)
        return np.stack([f["h"], f["h"] * f["u"], f["h"] * f["v"]], axis=1)

    def sources(self, x, y, t) -> np.ndarray:
        """[n, 3] source terms at the cell centroids, src/rdymms.c:560-582"""
        f = self.fields(x, y, t)
        h, u, v = f["h"], f["u"], f["v"]
        n = self.mannings(x, y)
        g = GRAVITY
        Cd = g * n * n * np.power(h, -1.0 / 3.0)
        speed = np.sqrt(u * u + v * v)
        sh = f["dhdt"] + u * f["dhdx"] + h * f["dudx"] + v * f["dhdy"] + h * f["dvdy"]
        shu = u * f["dhdt"] + h * f["dudt"]
        shu = shu + 2.0 * u * h * f["dudx"] + u * u * f["dhdx"] + g * h * f["dhdx"]
        shu = shu + u * h * f["dvdy"] + v * h * f["dudy"] + u * v * f["dhdy"]
        shu = shu + self.dzdx(x, y) * g * h
        shu = shu + Cd * u * speed
        shv = v * f["dhdt"] + h * f["dvdt"]
        shv = shv + u * h * f["dvdx"] + v * h * f["dudx"] + u * v * f["dhdx"]
        shv = shv + v * v * f["dhdy"] + 2.0 * v * h * f["dvdy"] + g * h * f["dhdy"]
        shv = shv + self.dzdy(x, y) * g * h
        shv = shv + Cd * v * speed
        return np.stack([sh, shu, shv], axis=1)


def error_norms(u: np.ndarray, exact: np.ndarray, areas: np.ndarray):
    """(L1[3], L2[3], Linf[3]) over owned cells, src/rdymms.c:866-892"""
    e = np.asarray(u).reshape(-1, 3) - exact
    a = areas[:, None]
    return (np.abs(e) * a).sum(axis=0), np.sqrt((e * e * a).sum(axis=0)), np.abs(e).max(axis=0)


def convergence_rates(num_cells, norms) -> np.ndarray:
    """norms: [levels, 3]; rate = -dim * slope of log10(err) against log10(N), dim = 2 (src/rdymms.c:973-995)"""
    x = np.log10(np.asarray(num_cells, dtype=float))
    y = np.log10(np.asarray(norms, dtype=float))
    return np.array([-2.0 * np.polyfit(x, y[:, c], 1)[0] for c in range(y.shape[1])])


# mms.convergence.expected_rates of mms_conv_study.yaml:49-64 (a rate must EXCEED its threshold, src/rdymms.c:1005-1008)
EXPECTED_RATES = {"h": (0.94, 0.95, 0.94), "hu": (0.91, 0.93, 0.77), "hv": (0.91, 0.93, 0.77)}
# driver/tests/swe_roe/mms_conv_study_second_order.yaml:57-69
EXPECTED_RATES_SECOND_ORDER = {"h": (1.40, 1.20, 0.80), "hu": (1.30, 1.30, 0.85), "hv": (1.30, 1.30, 0.85)}
