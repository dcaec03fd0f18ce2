/* device_layout.c — builds the device mesh layout (see device_layout.h) on the host, in C.
 *
 * Reference quantities consumed: RDyCells / RDyEdges members listed in RdybMeshView
 * (include/private/rdymeshimpl.h:26-140), RDyBoundary.edge_ids (include/private/rdyboundaryimpl.h:9-16).
 * Accumulation order per cell follows the reference operator: interior edges in internal_edge_ids order
 * (src/swe/swe_petsc.c:244-310), then boundary 0..B-1 in edge_ids order (src/operator_fluxes_petsc.c:28-51).
 */
#include "device_layout.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "rdyb_error.h"

#define LALLOC(ptr, n)                                                                   \
  do {                                                                                   \
    (ptr) = calloc((size_t)((n) > 0 ? (n) : 1), sizeof(*(ptr)));                         \
    if (!(ptr)) { ierr = rdyb_fail(RDYB_ERR_MEM, "device layout: out of memory (%s)", #ptr); goto fail; } \
  } while (0)

/* Hilbert index of (x,y) on a 2^order x 2^order grid */
static uint32_t hilbert_xy2d(uint32_t order, uint32_t x, uint32_t y) {
  const uint32_t n = 1u << order;
  uint32_t d = 0;
  for (uint32_t s = n >> 1; s > 0; s >>= 1) {
    uint32_t rx = (x & s) ? 1u : 0u, ry = (y & s) ? 1u : 0u;
    d += s * s * ((3u * rx) ^ ry);
    if (!ry) { /* rotate the quadrant; bits above s are already consumed, so flipping them is harmless */
      if (rx) { x = n - 1 - x; y = n - 1 - y; }
      uint32_t t = x; x = y; y = t;
    }
  }
  return d;
}

/* stable LSD radix sort of (key,idx) pairs packed as key<<32|idx */
static int radix_sort_u64(uint64_t *a, int64_t n) {
  uint64_t *b = malloc(sizeof(uint64_t) * (size_t)(n > 0 ? n : 1));
  if (!b) return rdyb_fail(RDYB_ERR_MEM, "device layout: out of memory (sort)");
  for (int pass = 0; pass < 4; ++pass) {
    int shift = 32 + 8 * pass;
    int64_t count[257] = {0};
    for (int64_t i = 0; i < n; ++i) ++count[((a[i] >> shift) & 0xff) + 1];
    for (int k = 0; k < 256; ++k) count[k + 1] += count[k];
    for (int64_t i = 0; i < n; ++i) b[count[(a[i] >> shift) & 0xff]++] = a[i];
    uint64_t *t = a; a = b; b = t;
  }
  free(b); /* 4 passes: the data ended up back in the caller's buffer */
  return RDYB_SUCCESS;
}

void rdyb_layout_free(RdybLayout *L) {
  if (!L) return;
  free(L->dev_to_local); free(L->local_to_dev); free(L->owned_to_dev); free(L->dev_to_owned); free(L->inv_area); free(L->area);
  free(L->dzdx); free(L->dzdy); free(L->zc); free(L->ell_nbr); free(L->ell_sn); free(L->ell_cn); free(L->ell_slen); free(L->ell_edge);
  free(L->b_offsets); free(L->b_type); free(L->b_cell); free(L->peer_ranks); free(L->send_offsets); free(L->send_cells);
  free(L->recv_offsets);
  free(L->ell_gcx); free(L->ell_gcy); free(L->ell_dxs); free(L->ell_dys); free(L->ell_dxn); free(L->ell_dyn);
  memset(L, 0, sizeof(*L));
}

int rdyb_layout_build(const RdybMeshView *mv, int32_t nb, const RdybBoundaryView *bv, const RdybHaloView *halo, int reorder,
                      RdybLayout *L) {
  int ierr = RDYB_SUCCESS;
  uint64_t *keys = NULL;
  int8_t   *frontier = NULL, *placed = NULL;
  int32_t  *fill = NULL;
  RDYB_CHECK(mv && L, RDYB_ERR_ARG, "device layout: NULL argument");
  memset(L, 0, sizeof(*L));
  const int32_t nc = mv->num_cells, N = mv->num_owned_cells;
  RDYB_CHECK(nc > 0 && N > 0 && N <= nc, RDYB_ERR_USER, "device layout: mesh has %d cells, %d owned", nc, N);
  RDYB_CHECK(mv->cell_is_owned && mv->cell_local_to_owned && mv->cell_areas && mv->edge_cell_ids && mv->edge_lengths && mv->edge_cn &&
                 mv->edge_sn && (mv->internal_edge_ids || mv->num_internal_edges == 0),
             RDYB_ERR_ARG, "device layout: mesh view is missing a required array");
  RDYB_CHECK(nb == 0 || bv, RDYB_ERR_ARG, "device layout: %d boundaries but no boundary array", nb);

  L->n_owned = N; L->n_local = nc; L->num_view_cells = nc;
  LALLOC(L->dev_to_local, nc); LALLOC(L->local_to_dev, nc); LALLOC(L->owned_to_dev, N); LALLOC(L->dev_to_owned, N);
  LALLOC(frontier, nc); LALLOC(placed, nc);
  {
    int32_t count = 0;
    for (int32_t c = 0; c < nc; ++c) {
      L->local_to_dev[c] = -1;
      if (mv->cell_is_owned[c]) {
        int32_t o = mv->cell_local_to_owned[c];
        if (o < 0 || o >= N) { ierr = rdyb_fail(RDYB_ERR_USER, "device layout: local_to_owned[%d] = %d out of range", c, o); goto fail; }
        ++count;
      }
    }
    if (count != N) { ierr = rdyb_fail(RDYB_ERR_USER, "device layout: is_owned marks %d cells, num_owned_cells = %d", count, N); goto fail; }
  }
  /* frontier = owned cells sharing an edge with a non-owned cell */
  for (int32_t i = 0; i < mv->num_internal_edges; ++i) {
    int32_t e = mv->internal_edge_ids[i], l = mv->edge_cell_ids[2 * e], r = mv->edge_cell_ids[2 * e + 1];
    if (r < 0) continue;
    if (mv->cell_is_owned[l] && !mv->cell_is_owned[r]) frontier[l] = 1;
    if (mv->cell_is_owned[r] && !mv->cell_is_owned[l]) frontier[r] = 1;
  }
  /* order owned cells: (frontier flag, Hilbert key, local id) */
  LALLOC(keys, N);
  {
    double xmin = 1e300, xmax = -1e300, ymin = 1e300, ymax = -1e300;
    int use_curve = reorder && mv->cell_xc && mv->cell_yc;
    if (use_curve) {
      for (int32_t c = 0; c < nc; ++c) {
        if (!mv->cell_is_owned[c]) continue;
        if (mv->cell_xc[c] < xmin) xmin = mv->cell_xc[c];
        if (mv->cell_xc[c] > xmax) xmax = mv->cell_xc[c];
        if (mv->cell_yc[c] < ymin) ymin = mv->cell_yc[c];
        if (mv->cell_yc[c] > ymax) ymax = mv->cell_yc[c];
      }
    }
    double span = fmax(xmax - xmin, ymax - ymin);
    double scale = (use_curve && span > 0) ? 32767.0 / span : 0.0; /* 15-bit grid -> 30-bit curve index, bit 31 = frontier */
    int32_t k = 0;
    for (int32_t c = 0; c < nc; ++c) {
      if (!mv->cell_is_owned[c]) continue;
      uint32_t key = 0;
      if (use_curve) key = hilbert_xy2d(15, (uint32_t)((mv->cell_xc[c] - xmin) * scale), (uint32_t)((mv->cell_yc[c] - ymin) * scale));
      if (frontier[c]) key |= 0x80000000u;
      keys[k++] = ((uint64_t)key << 32) | (uint32_t)c;
    }
    if ((ierr = radix_sort_u64(keys, N))) goto fail;
    L->n_interior = 0;
    for (int32_t d = 0; d < N; ++d) {
      int32_t c = (int32_t)(keys[d] & 0xffffffffu);
      L->dev_to_local[d] = c; L->local_to_dev[c] = d;
      L->owned_to_dev[mv->cell_local_to_owned[c]] = d; L->dev_to_owned[d] = mv->cell_local_to_owned[c];
      if (!frontier[c]) ++L->n_interior;
    }
    /* the last (< 8) interior cells are treated as frontier so that both ranges start on a multiple of 8 cells
     * (16-byte alignment of every per-cell slice a tile fetches with a bulk async copy) */
    if (L->n_interior < N) L->n_interior &= ~15; /* 16-byte aligned slices of the per-cell uint8 / uint16 / double arrays for bulk copies */
  }
  /* ghosts: per peer in recv order, then the remaining non-owned cells in ascending local id */
  {
    int32_t d = N;
    if (halo && halo->num_peers > 0) {
      int32_t np = halo->num_peers;
      L->num_peers = np;
      LALLOC(L->peer_ranks, np); LALLOC(L->send_offsets, np + 1); LALLOC(L->recv_offsets, np + 1);
      LALLOC(L->send_cells, halo->send_offsets[np]);
      for (int32_t p = 0; p < np; ++p) {
        L->peer_ranks[p] = halo->peer_ranks[p];
        L->recv_offsets[p] = d - N;
        for (int32_t k = halo->recv_offsets[p]; k < halo->recv_offsets[p + 1]; ++k) {
          int32_t c = halo->recv_cells[k];
          if (c < 0 || c >= nc || mv->cell_is_owned[c] || placed[c]) {
            ierr = rdyb_fail(RDYB_ERR_USER, "device layout: halo recv cell %d (peer %d) is not a distinct ghost cell", c, halo->peer_ranks[p]);
            goto fail;
          }
          placed[c] = 1; L->dev_to_local[d] = c; L->local_to_dev[c] = d; ++d;
        }
      }
      L->recv_offsets[np] = d - N;
      for (int32_t p = 0; p <= np; ++p) L->send_offsets[p] = halo->send_offsets[p];
      for (int32_t k = 0; k < halo->send_offsets[np]; ++k) {
        int32_t c = halo->send_cells[k];
        if (c < 0 || c >= nc || !mv->cell_is_owned[c]) { ierr = rdyb_fail(RDYB_ERR_USER, "device layout: halo send cell %d is not owned", c); goto fail; }
        L->send_cells[k] = L->local_to_dev[c];
      }
    } else {
      LALLOC(L->peer_ranks, 1); LALLOC(L->send_offsets, 1); LALLOC(L->recv_offsets, 1); LALLOC(L->send_cells, 1);
    }
    for (int32_t c = 0; c < nc; ++c)
      if (!mv->cell_is_owned[c] && !placed[c]) { L->dev_to_local[d] = c; L->local_to_dev[c] = d; ++d; }
  }
  L->stride = (nc + 63) & ~63;

  /* per-cell geometry */
  LALLOC(L->inv_area, L->stride); LALLOC(L->area, L->stride); LALLOC(L->dzdx, L->stride); LALLOC(L->dzdy, L->stride);
  if (mv->cell_zc) LALLOC(L->zc, L->stride);
  for (int32_t d = 0; d < nc; ++d) {
    int32_t c = L->dev_to_local[d];
    L->area[d] = mv->cell_areas[c];
    L->inv_area[d] = 1.0 / mv->cell_areas[c];
    if (d < N) { L->dzdx[d] = mv->cell_dz_dx ? mv->cell_dz_dx[c] : 0.0; L->dzdy[d] = mv->cell_dz_dy ? mv->cell_dz_dy[c] : 0.0; }
    if (mv->cell_zc) L->zc[d] = mv->cell_zc[c];
  }

  /* boundary-edge table */
  L->num_boundaries = nb;
  LALLOC(L->b_offsets, nb + 1);
  for (int32_t b = 0; b < nb; ++b) {
    if (bv[b].num_edges < 0 || (bv[b].num_edges > 0 && !bv[b].edge_ids)) { ierr = rdyb_fail(RDYB_ERR_ARG, "device layout: boundary %d has no edge list", b); goto fail; }
    L->b_offsets[b + 1] = L->b_offsets[b] + bv[b].num_edges;
  }
  L->num_boundary_edges = L->b_offsets[nb];
  LALLOC(L->b_type, L->num_boundary_edges); LALLOC(L->b_cell, L->num_boundary_edges);

  /* ELL slots */
  {
    size_t tot = (size_t)RDYB_ELL_WIDTH * (size_t)L->stride;
    LALLOC(L->ell_nbr, tot); LALLOC(L->ell_sn, tot); LALLOC(L->ell_cn, tot); LALLOC(L->ell_slen, tot); LALLOC(L->ell_edge, tot);
    for (size_t q = 0; q < tot; ++q) { L->ell_nbr[q] = RDYB_NO_EDGE; L->ell_edge[q] = -1; }
    LALLOC(fill, N);
    int32_t maxfill = 0;
#define PUSH(dcell, nbrval, e, signedlen)                                                                        \
  do {                                                                                                           \
    int32_t k_ = fill[dcell]++;                                                                                  \
    if (k_ >= RDYB_ELL_WIDTH) { ierr = rdyb_fail(RDYB_ERR_USER, "device layout: cell %d has more than %d edges", L->dev_to_local[dcell], RDYB_ELL_WIDTH); goto fail; } \
    size_t q_ = (size_t)k_ * (size_t)L->stride + (size_t)(dcell);                                                \
    L->ell_nbr[q_] = (nbrval); L->ell_sn[q_] = mv->edge_sn[e]; L->ell_cn[q_] = mv->edge_cn[e];                   \
    L->ell_slen[q_] = (signedlen); L->ell_edge[q_] = (e);                                                        \
    if (k_ + 1 > maxfill) maxfill = k_ + 1;                                                                      \
  } while (0)
    for (int32_t i = 0; i < mv->num_internal_edges; ++i) {
      int32_t e = mv->internal_edge_ids[i];
      if (e < 0 || e >= mv->num_edges) { ierr = rdyb_fail(RDYB_ERR_USER, "device layout: internal edge id %d out of range", e); goto fail; }
      int32_t l = mv->edge_cell_ids[2 * e], r = mv->edge_cell_ids[2 * e + 1];
      if (r < 0) continue; /* src/swe/swe_petsc.c:251 */
      if (l < 0 || l >= nc || r >= nc) { ierr = rdyb_fail(RDYB_ERR_USER, "device layout: edge %d references a cell out of range", e); goto fail; }
      if (mv->cell_is_owned[l]) PUSH(L->local_to_dev[l], L->local_to_dev[r], e, -mv->edge_lengths[e]);
      if (mv->cell_is_owned[r]) PUSH(L->local_to_dev[r], L->local_to_dev[l], e, mv->edge_lengths[e]);
    }
    for (int32_t b = 0; b < nb; ++b) {
      for (int32_t k = 0; k < bv[b].num_edges; ++k) {
        int32_t e = bv[b].edge_ids[k], j = L->b_offsets[b] + k;
        if (e < 0 || e >= mv->num_edges) { ierr = rdyb_fail(RDYB_ERR_USER, "device layout: boundary %d edge id %d out of range", b, e); goto fail; }
        int32_t l = mv->edge_cell_ids[2 * e];
        L->b_type[j] = bv[b].bc_type;
        L->b_cell[j] = -1;
        if (l < 0 || l >= nc) { ierr = rdyb_fail(RDYB_ERR_USER, "device layout: boundary edge %d has no left cell", e); goto fail; }
        if (!mv->cell_is_owned[l]) continue; /* src/swe/swe_petsc.c:593: only owned left cells receive the flux */
        L->b_cell[j] = L->local_to_dev[l];
        PUSH(L->local_to_dev[l], -(1 + j), e, -mv->edge_lengths[e]);
      }
    }
#undef PUSH
    L->width = maxfill <= 3 ? 3 : 4;
  }
  free(keys); free(frontier); free(placed); free(fill);
  return RDYB_SUCCESS;
fail:
  free(keys); free(frontier); free(placed); free(fill);
  rdyb_layout_free(L);
  return ierr;
}

/* ---------------- second-order (MUSCL) slot geometry ---------------- */
/* Weighted least-squares gradient weights and face offsets per ELL slot.  Follows PrecomputeLSGradCoeffs
 * (src/operator_fluxes_ceed.c:884-975): weight 1/d per neighbour, per-cell normal matrix summed over the local internal edges in
 * internal_edge_ids order, analytic 2x2 inverse, |det| < 1e-15 => zero gradient; the left cell's weights are M_l^-1 (w dx, w dy),
 * the right cell's M_r^-1 (w dx, w dy) with (dx, dy) = x_right - x_left for both. */
int rdyb_layout_build_muscl(const RdybMeshView *mv, RdybLayout *L) {
  RDYB_CHECK(mv->cell_xc && mv->cell_yc && mv->edge_xm && mv->edge_ym, RDYB_ERR_USER,
             "second-order reconstruction needs cell centroids and edge midpoints in the mesh view");
  const int32_t nc = mv->num_cells;
  const size_t tot = (size_t)RDYB_ELL_WIDTH * (size_t)L->stride;
  double *M = calloc(3 * (size_t)(nc > 0 ? nc : 1), sizeof(double)), *inv = calloc(4 * (size_t)(nc > 0 ? nc : 1), sizeof(double));
  L->ell_gcx = calloc(tot, sizeof(double)); L->ell_gcy = calloc(tot, sizeof(double));
  L->ell_dxs = calloc(tot, sizeof(double)); L->ell_dys = calloc(tot, sizeof(double));
  L->ell_dxn = calloc(tot, sizeof(double)); L->ell_dyn = calloc(tot, sizeof(double));
  if (!M || !inv || !L->ell_gcx || !L->ell_gcy || !L->ell_dxs || !L->ell_dys || !L->ell_dxn || !L->ell_dyn) {
    free(M); free(inv);
    return rdyb_fail(RDYB_ERR_MEM, "device layout: out of memory (MUSCL slot geometry)");
  }
  for (int32_t i = 0; i < mv->num_internal_edges; ++i) {
    const int32_t e = mv->internal_edge_ids[i], l = mv->edge_cell_ids[2 * e], r = mv->edge_cell_ids[2 * e + 1];
    if (r < 0) continue;
    const double dx = mv->cell_xc[r] - mv->cell_xc[l], dy = mv->cell_yc[r] - mv->cell_yc[l];
    const double d = sqrt(dx * dx + dy * dy), w = d > 0.0 ? 1.0 / d : 0.0;
    M[3 * l] += w * dx * dx; M[3 * l + 1] += w * dx * dy; M[3 * l + 2] += w * dy * dy;
    M[3 * r] += w * dx * dx; M[3 * r + 1] += w * dx * dy; M[3 * r + 2] += w * dy * dy;
  }
  for (int32_t c = 0; c < nc; ++c) {
    const double m00 = M[3 * c], m01 = M[3 * c + 1], m11 = M[3 * c + 2], det = m00 * m11 - m01 * m01;
    if (fabs(det) < 1e-15) continue; /* degenerate: zero gradient */
    const double id = 1.0 / det;
    inv[4 * c] = m11 * id; inv[4 * c + 1] = -m01 * id; inv[4 * c + 2] = -m01 * id; inv[4 * c + 3] = m00 * id;
  }
  for (int32_t dcell = 0; dcell < L->n_owned; ++dcell) {
    const int32_t self = L->dev_to_local[dcell];
    for (int k = 0; k < RDYB_ELL_WIDTH; ++k) {
      const size_t q = (size_t)k * (size_t)L->stride + (size_t)dcell;
      if (L->ell_nbr[q] < 0) continue; /* empty or boundary slot: first-order boundary flux, no gradient contribution */
      const int32_t e = L->ell_edge[q], l = mv->edge_cell_ids[2 * e], r = mv->edge_cell_ids[2 * e + 1];
      const int32_t other = self == l ? r : l;
      const double dx = mv->cell_xc[r] - mv->cell_xc[l], dy = mv->cell_yc[r] - mv->cell_yc[l];
      const double d = sqrt(dx * dx + dy * dy), w = d > 0.0 ? 1.0 / d : 0.0, wdx = w * dx, wdy = w * dy;
      /* the side is folded into the stored weight (an exact sign flip): the left cell sees q_right - q_left = q_nbr - q_self, the right
       * cell -(q_nbr - q_self), so that the kernels need neither the slot's signed length nor a select */
      const double sgn = self == l ? 1.0 : -1.0;
      L->ell_gcx[q] = sgn * (inv[4 * self] * wdx + inv[4 * self + 1] * wdy);
      L->ell_gcy[q] = sgn * (inv[4 * self + 2] * wdx + inv[4 * self + 3] * wdy);
      L->ell_dxs[q] = mv->edge_xm[e] - mv->cell_xc[self];  L->ell_dys[q] = mv->edge_ym[e] - mv->cell_yc[self];
      L->ell_dxn[q] = mv->edge_xm[e] - mv->cell_xc[other]; L->ell_dyn[q] = mv->edge_ym[e] - mv->cell_yc[other];
    }
  }
  free(M); free(inv);
  return RDYB_SUCCESS;
}

/* ---------------- tiled layout ---------------- */
void rdyb_tiled_layout_free(RdybTiledLayout *T) {
  if (!T) return;
  free(T->tile_cell_begin); free(T->halo_offsets); free(T->halo_cells); free(T->edge_offsets); free(T->edge_lr);
  free(T->edge_bidx); free(T->edge_sn); free(T->edge_cn); free(T->edge_len); free(T->cell_edges); free(T->edge_id); free(T->edge_cell); free(T->tile_desc);
  free(T->edge_off); free(T->halo2_cells); free(T->halo_nbr);
  memset(T, 0, sizeof(*T));
}

static int cmp_i32(const void *a, const void *b) {
  int32_t x = *(const int32_t *)a, y = *(const int32_t *)b;
  return x < y ? -1 : x > y;
}
static int32_t lower_bound_i32(const int32_t *a, int32_t n, int32_t key) {
  int32_t lo = 0, hi = n;
  while (lo < hi) { int32_t mid = (lo + hi) / 2; if (a[mid] < key) lo = mid + 1; else hi = mid; }
  return lo;
}

#define GROW(ptr, cap, need)                                                                      \
  do {                                                                                            \
    if ((need) > (cap)) {                                                                         \
      int64_t nc_ = (cap) + (cap) / 2 + 1024;                                                     \
      if (nc_ < (need)) nc_ = (need);                                                             \
      void *np_ = realloc((ptr), (size_t)nc_ * sizeof(*(ptr)));                                   \
      if (!np_) { ierr = rdyb_fail(RDYB_ERR_MEM, "tiled layout: out of memory (%s)", #ptr); goto fail; } \
      (ptr) = np_;                                                                                \
      (cap) = nc_;                                                                                \
    }                                                                                             \
  } while (0)

static int edges_reserve(RdybTiledLayout *T, int64_t *cap, int64_t need) {
  if (need <= *cap) return RDYB_SUCCESS;
  int64_t ncap = *cap + *cap / 2 + 4096;
  if (ncap < need) ncap = need;
  void *p;
#define RES(field, type)                                                                      \
  p = realloc(T->field, (size_t)ncap * sizeof(type));                                         \
  if (!p) return rdyb_fail(RDYB_ERR_MEM, "tiled layout: out of memory (edges)");              \
  T->field = p;
  RES(edge_lr, uint32_t) RES(edge_bidx, int32_t) RES(edge_sn, double) RES(edge_cn, double) RES(edge_len, double) RES(edge_id, int32_t)
  RES(edge_cell, int32_t)
#undef RES
  *cap = ncap;
  return RDYB_SUCCESS;
}

/* ---- bank-aware order of a tile's edges ----
 * K4 reads shared memory through two kinds of gathers: phase 2 loads the staged state of an edge's left and right cell
 * (index = tile-local cell), phase 3 loads the flux of each of a cell's edges (index = edge slot).  A 64-bit shared load is
 * served per half-warp, one wavefront per distinct 8-byte word and bank pair, so 16 consecutive threads whose indices
 * collide modulo 16 are replayed: with the edges in cell-major order a gather costs 3.8 (right cells) to 4.3 (fluxes)
 * wavefronts per warp instead of 2 (scratch/bank/bankmodel.c reproduces the ncu counters: 3.9 and 4 - 4.7).  The order of a
 * tile's edges is free (each cell keeps ITS edges in the reference's accumulation order, so results are bit-identical), so
 * it is chosen here: slot = 16 * group + lane with
 *   - the LANE (slot mod 16) of the edges gathered together by 16 consecutive cells in phase 3 pairwise distinct (a greedy
 *     edge colouring: an edge belongs to at most two such sets, its left and its right cell's), and
 *   - within a GROUP (the 16 edges of one half-warp in phase 2) left cells, and right cells, in distinct bank pairs
 *     (assignment lane by lane, cheapest free group first, then pairwise swaps).
 * Modelled wavefronts per warp instruction on the floodplain mesh: left cells 2.69 -> 3.12, right cells 3.81 -> 3.16, fluxes
 * 4.30 -> 2.60, i.e. 18 % fewer gather wavefronts per tile.  MEASURED on the B200 (A/B, same box, 300 steps, twice each):
 * 0.3364 / 0.3363 ms with, 0.3400 / 0.3364 ms without -- no difference, so bank conflicts are not what limits K4 and the pass
 * (0.4 ms of host time per tile) is OFF unless RDYB_EDGE_SCHEDULE is set. */
#define ES_MAX_EDGES 2048
#define ES_MAX_GROUPS (ES_MAX_EDGES / 16)
/* perm[old slot] = new slot; returns 0 when the tile is left as it is */
static int tile_edge_schedule(int32_t ne, const uint32_t *lr, int32_t ncell, int32_t W, const uint16_t *cell_edges, size_t S, int32_t cb, int32_t *perm) {
  if (ne < 32 || ne > ES_MAX_EDGES) return 0;
  const int32_t G = (ne + 15) / 16, rem = ne - 16 * (G - 1), CG = (ncell + 15) / 16, nsets = W * CG;
  int32_t (*eset)[2] = malloc(sizeof(int32_t[2]) * (size_t)ne);
  uint8_t *nset = calloc((size_t)ne, 1);
  uint8_t (*setcnt)[16] = calloc((size_t)nsets, 16);
  if (!eset || !nset || !setcnt) { free(eset); free(nset); free(setcnt); return 0; }
  uint8_t lane[ES_MAX_EDGES];
  int16_t order[ES_MAX_EDGES], grp[ES_MAX_EDGES];
  int ok = 1;
  for (int32_t k = 0; k < W; ++k)
    for (int32_t t = 0; t < ncell; ++t) {
      const uint16_t ce = cell_edges[(size_t)k * S + cb + t];
      if (ce == 0xffffu) continue;
      const int32_t e = ce & 0x7fff;
      if (e < ne && nset[e] < 2 && !(nset[e] == 1 && eset[e][0] == k * CG + t / 16)) eset[e][nset[e]++] = k * CG + t / 16; /* both cells in one set: a broadcast */
    }
#define LCOST(e, c) ((nset[e] > 0 ? setcnt[eset[e][0]][c] : 0) + (nset[e] > 1 ? setcnt[eset[e][1]][c] : 0))
#define LADD(e, c, d) do { if (nset[e] > 0) setcnt[eset[e][0]][c] += (d); if (nset[e] > 1) setcnt[eset[e][1]][c] += (d); } while (0)
  /* ---- lanes: most constrained edges first, the free-est lane on ties ---- */
  int32_t no = 0, cap[16];
  for (int pass = 2; pass >= 0; --pass)
    for (int32_t e = 0; e < ne; ++e) if (nset[e] == pass) order[no++] = (int16_t)e;
  for (int c = 0; c < 16; ++c) cap[c] = c < rem ? G : G - 1;
  for (int32_t i = 0; i < ne; ++i) {
    const int32_t e = order[i];
    int best = -1, bcost = 0, bcap = 0;
    for (int c = 0; c < 16; ++c) {
      if (!cap[c]) continue;
      const int cost = LCOST(e, c);
      if (best < 0 || cost < bcost || (cost == bcost && cap[c] > bcap)) { best = c; bcost = cost; bcap = cap[c]; }
    }
    lane[e] = (uint8_t)best; cap[best]--;
    LADD(e, best, 1);
  }
  /* repair: an edge whose lane collides trades lanes with any edge when the two together collide less afterwards */
  for (int sweep = 0; sweep < 3; ++sweep) {
    int changed = 0;
    for (int32_t e = 0; e < ne; ++e) {
      int ce_ = LCOST(e, lane[e]) - nset[e];
      if (ce_ <= 0) continue;
      for (int32_t f = 0; f < ne && ce_ > 0; ++f) {
        const int a = lane[e], b = lane[f];
        if (b == a || LCOST(e, b) >= ce_) continue; /* e would not be better off in f's lane */
        const int before = ce_ + LCOST(f, b) - nset[f];
        LADD(e, a, -1); LADD(f, b, -1);
        const int after = LCOST(e, b) + LCOST(f, a);
        if (after < before) { lane[e] = (uint8_t)b; lane[f] = (uint8_t)a; LADD(e, b, 1); LADD(f, a, 1); changed = 1; ce_ = LCOST(e, b) - nset[e]; }
        else { LADD(e, a, 1); LADD(f, b, 1); }
      }
    }
    if (!changed) break;
  }
#undef LCOST
#undef LADD
  /* ---- groups, lane by lane: each lane has exactly one edge in each of its groups ---- */
  uint8_t cl[ES_MAX_GROUPS][16], cr[ES_MAX_GROUPS][16];
  uint16_t fl[ES_MAX_GROUPS][16], fr[ES_MAX_GROUPS][16]; /* the first cell in each bank pair: the same cell again is a broadcast, not a conflict */
  memset(cl, 0, sizeof(cl[0]) * (size_t)G); memset(cr, 0, sizeof(cr[0]) * (size_t)G);
#define EL(e) (lr[e] & 0xffffu)
#define ER(e) (lr[e] >> 16)
#define SIDE(cnt, first, g, x) ((cnt)[g][(x) & 15u] == 0 || ((cnt)[g][(x) & 15u] == 1 && (first)[g][(x) & 15u] == (x)) ? 0 : (cnt)[g][(x) & 15u])
#define GCOST(e, g) (SIDE(cl, fl, g, EL(e)) + (ER(e) == 0xffffu ? 0 : SIDE(cr, fr, g, ER(e))))
#define GADD(cnt, first, g, x) do { uint8_t *c_ = &(cnt)[g][(x) & 15u]; if (*c_ == 0) { (first)[g][(x) & 15u] = (uint16_t)(x); *c_ = 1; } else if (!(*c_ == 1 && (first)[g][(x) & 15u] == (x))) ++*c_; } while (0)
  for (int c = 0; c < 16 && ok; ++c) {
    const int32_t ng = c < rem ? G : G - 1;
    int16_t members[ES_MAX_GROUPS];
    uint8_t gfree[ES_MAX_GROUPS];
    int32_t nm = 0;
    for (int32_t e = 0; e < ne; ++e) if (lane[e] == c) { if (nm < ES_MAX_GROUPS) members[nm] = (int16_t)e; nm++; }
    if (nm != ng) { ok = 0; break; }
    memset(gfree, 1, (size_t)ng);
    for (int32_t i = 0; i < nm; ++i) {
      const int32_t e = members[i];
      int32_t bg = -1, bc = 0;
      for (int32_t g = 0; g < ng; ++g) {
        if (!gfree[g]) continue;
        const int cst = GCOST(e, g);
        if (bg < 0 || cst < bc) { bg = g; bc = cst; if (!cst) break; }
      }
      grp[e] = (int16_t)bg; gfree[bg] = 0;
    }
    for (int sweep = 0; sweep < 2; ++sweep)
      for (int32_t i = 0; i < nm; ++i) {
        const int32_t e = members[i];
        if (!GCOST(e, grp[e])) continue;
        for (int32_t j = 0; j < nm; ++j) {
          const int32_t f = members[j];
          if (GCOST(e, grp[f]) + GCOST(f, grp[e]) < GCOST(e, grp[e]) + GCOST(f, grp[f])) { const int16_t x = grp[e]; grp[e] = grp[f]; grp[f] = x; if (!GCOST(e, grp[e])) break; }
        }
      }
    for (int32_t i = 0; i < nm; ++i) {
      const int32_t e = members[i];
      GADD(cl, fl, grp[e], EL(e));
      if (ER(e) != 0xffffu) GADD(cr, fr, grp[e], ER(e));
      perm[e] = 16 * grp[e] + c;
    }
  }
#undef GCOST
#undef GADD
#undef SIDE
#undef EL
#undef ER
  free(eset); free(nset); free(setcnt);
  return ok;
}

int rdyb_tiled_layout_build(const RdybLayout *L, int32_t tile_cells, RdybTiledLayout *T) {
  int ierr = RDYB_SUCCESS;
  int32_t *tmp_halo = NULL, *slot_edge = NULL;
  int32_t eperm[ES_MAX_EDGES];
  double escratch[ES_MAX_EDGES];
  const int schedule = getenv("RDYB_EDGE_SCHEDULE") != NULL; /* opt-in: measured, no gain */
  RDYB_CHECK(L && T, RDYB_ERR_ARG, "tiled layout: NULL argument");
  RDYB_CHECK(tile_cells >= 128 && tile_cells <= 8192 && tile_cells % 8 == 0, RDYB_ERR_ARG, "tiled layout: tile_cells (%d) must be a multiple of 8 in [128, 8192]", tile_cells);
  memset(T, 0, sizeof(*T));
  const int32_t N = L->n_owned, NI = L->n_interior, W = L->width, TC = tile_cells;
  const size_t S = (size_t)L->stride;
  /* frontier cells (neighbours of ghosts) are cut into SHORT tiles: a strip along a partition border has up to `width` outside
   * neighbours per cell, so a full-size tile of it would need several hundred halo cells and could only run on the general
   * kernel in a launch of its own.  With at most 80 / width cells (a multiple of 8, for the 16-byte slices) a frontier tile has
   * <= 80 halo cells and <= 80 edges and fits every compiled tile shape: interior and frontier run in ONE launch (K4 waits for the
   * neighbours' border cells inside the kernel, swe_tiled_fixed.cuh).  There are few of them (~0.05 % of the cells at 10 M). */
  int32_t TCF = (80 / (W > 0 ? W : 1)) & ~7;
  if (TCF < 8) TCF = 8;
  if (TCF > TC) TCF = TC;
  const int32_t nt_int = (NI + TC - 1) / TC, nt_fr = (N - NI + TCF - 1) / TCF, nt = nt_int + nt_fr;
  T->tile_cells = TC; T->num_tiles = nt; T->num_interior_tiles = nt_int;
  LALLOC(T->tile_cell_begin, nt + 1); LALLOC(T->halo_offsets, nt + 1); LALLOC(T->edge_offsets, nt + 1);
  LALLOC(T->tile_desc, 8 * (size_t)nt);
  LALLOC(T->cell_edges, (size_t)RDYB_ELL_WIDTH * S);
  memset(T->cell_edges, 0xff, sizeof(uint16_t) * (size_t)RDYB_ELL_WIDTH * S);
  LALLOC(tmp_halo, (size_t)TC * RDYB_ELL_WIDTH); LALLOC(slot_edge, (size_t)TC * RDYB_ELL_WIDTH);
  int64_t hcap = 0, ecap = 0, nh = 0, ne = 0;
  for (int32_t t = 0; t < nt; ++t) {
    const int32_t cb = t < nt_int ? t * TC : NI + (t - nt_int) * TCF;
    const int32_t lim = t < nt_int ? NI : N, span = t < nt_int ? TC : TCF;
    const int32_t ce = cb + span < lim ? cb + span : lim, ncell = ce - cb;
    /* tile starts are padded to 4 entries (16 bytes for the int32/uint32 arrays, 32 for the doubles) so that a tile's slice
     * of every per-tile array can be fetched with one bulk async copy */
    while (nh & 3) { GROW(T->halo_cells, hcap, nh + 1); T->halo_cells[nh++] = 0; }
    if (ne & 3) {
      int64_t pad = 4 - (ne & 3);
      if (ne + pad > ecap) { ierr = edges_reserve(T, &ecap, ne + pad); if (ierr) goto fail; }
      for (int64_t g = ne; g < ne + pad; ++g) {
        T->edge_lr[g] = 0; T->edge_bidx[g] = -1; T->edge_sn[g] = 0.0; T->edge_cn[g] = 1.0; T->edge_len[g] = 0.0; T->edge_id[g] = -1; T->edge_cell[g] = -1;
      }
      ne += pad;
    }
    T->tile_cell_begin[t] = cb; T->halo_offsets[t] = (int32_t)nh; T->edge_offsets[t] = (int32_t)ne;
    /* halo = neighbours outside [cb,ce), sorted unique */
    int32_t nth = 0;
    for (int32_t i = cb; i < ce; ++i)
      for (int32_t k = 0; k < W; ++k) {
        int32_t nbr = L->ell_nbr[(size_t)k * S + i];
        if (nbr >= 0 && (nbr < cb || nbr >= ce)) tmp_halo[nth++] = nbr;
      }
    qsort(tmp_halo, (size_t)nth, sizeof(int32_t), cmp_i32);
    int32_t nuniq = 0;
    for (int32_t j = 0; j < nth; ++j) if (j == 0 || tmp_halo[j] != tmp_halo[j - 1]) tmp_halo[nuniq++] = tmp_halo[j];
    nth = nuniq;
    if (TC + nth > 65534) { ierr = rdyb_fail(RDYB_ERR_SUP, "tiled layout: tile %d needs %d staged cells (> 65534)", t, ncell + nth); goto fail; }
    GROW(T->halo_cells, hcap, nh + nth);
    memcpy(T->halo_cells + nh, tmp_halo, sizeof(int32_t) * (size_t)nth);
    /* pass 1: edges owned by the tile = slots where the cell is the left cell, or the neighbour is outside, or a boundary */
    int32_t nte = 0;
    for (int32_t i = cb; i < ce; ++i)
      for (int32_t k = 0; k < W; ++k) {
        const size_t q = (size_t)k * S + i;
        const int32_t nbr = L->ell_nbr[q];
        slot_edge[(i - cb) * RDYB_ELL_WIDTH + k] = -1;
        if (nbr == RDYB_NO_EDGE) continue;
        const int in_tile = nbr >= cb && nbr < ce;
        const int me_left = L->ell_slen[q] < 0.0;
        if (nbr >= 0 && in_tile && !me_left) continue; /* created by the left cell */
        const int64_t g = ne + nte;
        if (g + 1 > ecap) { ierr = edges_reserve(T, &ecap, g + 1); if (ierr) goto fail; }
        uint32_t me = (uint32_t)(i - cb), other = 0xFFFFu;
        int32_t report_cell = i;
        if (nbr >= 0) {
          other = in_tile ? (uint32_t)(nbr - cb) : (uint32_t)(TC + lower_bound_i32(tmp_halo, nth, nbr)); /* halo cells sit at TC + t */
          const int32_t lc = me_left ? i : nbr, rc = me_left ? nbr : i;
          report_cell = L->area[lc] < L->area[rc] ? lc : rc; /* src/swe/swe_petsc.c:289-296 */
        }
        const uint32_t left = me_left ? me : other, right = me_left ? other : me;
        T->edge_lr[g] = left | (right << 16);
        T->edge_bidx[g] = nbr < 0 ? -(nbr + 1) : -1;
        T->edge_sn[g] = L->ell_sn[q]; T->edge_cn[g] = L->ell_cn[q]; T->edge_len[g] = fabs(L->ell_slen[q]);
        T->edge_id[g] = L->ell_edge[q]; T->edge_cell[g] = report_cell;
        slot_edge[(i - cb) * RDYB_ELL_WIDTH + k] = nte++;
      }
    if (nte > 32767) { ierr = rdyb_fail(RDYB_ERR_SUP, "tiled layout: tile %d has %d edges (> 32767)", t, nte); goto fail; }
    /* pass 2: per-cell edge slots, in the ELL (reference accumulation) order */
    for (int32_t i = cb; i < ce; ++i)
      for (int32_t k = 0; k < W; ++k) {
        const size_t q = (size_t)k * S + i;
        const int32_t nbr = L->ell_nbr[q];
        if (nbr == RDYB_NO_EDGE) continue;
        int32_t e = slot_edge[(i - cb) * RDYB_ELL_WIDTH + k];
        const int me_left = L->ell_slen[q] < 0.0;
        if (e < 0) { /* in-tile edge created by the left cell nbr: find its slot with the same edge id */
          for (int32_t k2 = 0; k2 < W && e < 0; ++k2)
            if (L->ell_nbr[(size_t)k2 * S + nbr] == i && L->ell_edge[(size_t)k2 * S + nbr] == L->ell_edge[q]) e = slot_edge[(nbr - cb) * RDYB_ELL_WIDTH + k2];
          if (e < 0) { ierr = rdyb_fail(RDYB_ERR_USER, "tiled layout: inconsistent edge lists at cell %d", L->dev_to_local[i]); goto fail; }
        }
        T->cell_edges[q] = (uint16_t)((uint32_t)e | (me_left ? 0u : 0x8000u));
      }
    /* pass 3: bank-aware edge order (see tile_edge_schedule) */
    if (schedule && nte <= ES_MAX_EDGES && tile_edge_schedule(nte, T->edge_lr + ne, ncell, W, T->cell_edges, S, cb, eperm)) {
#define PERMUTE(field, type)                                                              \
      do {                                                                                \
        type *a_ = T->field + ne, *b_ = (type *)escratch;                                  \
        for (int32_t e = 0; e < nte; ++e) b_[eperm[e]] = a_[e];                           \
        memcpy(a_, b_, sizeof(type) * (size_t)nte);                                       \
      } while (0)
      PERMUTE(edge_lr, uint32_t); PERMUTE(edge_bidx, int32_t); PERMUTE(edge_id, int32_t); PERMUTE(edge_cell, int32_t);
      PERMUTE(edge_sn, double); PERMUTE(edge_cn, double); PERMUTE(edge_len, double);
#undef PERMUTE
      for (int32_t k = 0; k < W; ++k)
        for (int32_t i = cb; i < ce; ++i) {
          const uint16_t v = T->cell_edges[(size_t)k * S + i];
          if (v != 0xffffu) T->cell_edges[(size_t)k * S + i] = (uint16_t)((uint32_t)eperm[v & 0x7fffu] | (v & 0x8000u));
        }
    }
    { int32_t *d = &T->tile_desc[8 * (size_t)t]; d[0] = cb; d[1] = ncell; d[2] = (int32_t)ne; d[3] = nte; d[4] = (int32_t)nh; d[5] = nth; d[6] = 0; d[7] = 0; }
    nh += nth; ne += nte;
    if (nth > T->max_halo) T->max_halo = nth;
    if (nte > T->max_edges) T->max_edges = nte;
    if (t < nt_int && nth > T->max_halo_interior) T->max_halo_interior = nth;
    if (t < nt_int && nte > T->max_edges_interior) T->max_edges_interior = nte;
    if (ne > INT32_MAX - 65536 || nh > INT32_MAX - 65536) { ierr = rdyb_fail(RDYB_ERR_SUP, "tiled layout: too many tile edges for int32 offsets"); goto fail; }
  }
  T->tile_cell_begin[nt] = N; T->halo_offsets[nt] = (int32_t)nh; T->edge_offsets[nt] = (int32_t)ne;
  /* slack so that bulk copies rounded up to 16 bytes stay inside the arrays */
  GROW(T->halo_cells, hcap, nh + 16);
  for (int j = 0; j < 16; ++j) T->halo_cells[nh + j] = 0;
  if ((ierr = edges_reserve(T, &ecap, ne + 16))) goto fail;
  for (int64_t g = ne; g < ne + 16; ++g) { T->edge_lr[g] = 0; T->edge_bidx[g] = -1; T->edge_sn[g] = 0.0; T->edge_cn[g] = 1.0; T->edge_len[g] = 0.0; T->edge_id[g] = -1; T->edge_cell[g] = -1; }
  T->total_edges = ne + 16; T->total_halo = nh + 16;
  free(tmp_halo); free(slot_edge);
  return RDYB_SUCCESS;
fail:
  free(tmp_halo); free(slot_edge);
  rdyb_tiled_layout_free(T);
  return ierr;
}

/* Second-order face offsets per tile edge, derived from the per-slot MUSCL geometry (rdyb_layout_build_muscl): every tile edge is
 * listed in cell_edges by at least one tile cell, whose slot holds midpoint - own centroid (dxs, dys) and midpoint - neighbour
 * centroid (dxn, dyn); bit 15 of the cell_edges entry says whether that cell is the edge's right cell.  Both cells of an in-tile
 * edge write the same doubles (the left cell's dxn IS the right cell's dxs), so the result does not depend on the visiting order
 * nor on the optional edge schedule.  Layout: four planes of edge_off_plane doubles {dxl, dyl, dxr, dyr}; boundary and padding
 * edges keep 0 (boundary fluxes stay first order, ReconstructFaceValues only visits internal edges). */
int rdyb_tiled_layout_build_muscl(const RdybLayout *L, RdybTiledLayout *T) {
  RDYB_CHECK(L && T && T->tile_desc, RDYB_ERR_ARG, "tiled MUSCL layout: NULL argument");
  RDYB_CHECK(L->ell_dxs && L->ell_dys && L->ell_dxn && L->ell_dyn, RDYB_ERR_USER, "tiled MUSCL layout: build the MUSCL slot geometry first");
  free(T->edge_off);
  const size_t plane = ((size_t)T->total_edges + 3) & ~(size_t)3, S = (size_t)L->stride;
  T->edge_off_plane = (int64_t)plane;
  T->edge_off = calloc(4 * (plane > 0 ? plane : 1), sizeof(double));
  if (!T->edge_off) return rdyb_fail(RDYB_ERR_MEM, "tiled MUSCL layout: out of memory (face offsets)");
  for (int32_t t = 0; t < T->num_tiles; ++t) {
    const int32_t *d = &T->tile_desc[8 * (size_t)t];
    const int32_t cb = d[0], ncell = d[1], eo = d[2], nedge = d[3];
    for (int32_t i = cb; i < cb + ncell; ++i)
      for (int32_t k = 0; k < L->width; ++k) {
        const size_t q = (size_t)k * S + (size_t)i;
        const uint16_t ce = T->cell_edges[q];
        if (ce == 0xffffu || L->ell_nbr[q] < 0) continue;
        const int32_t e = ce & 0x7fff;
        if (e >= nedge) return rdyb_fail(RDYB_ERR_USER, "tiled MUSCL layout: inconsistent edge slot at cell %d", L->dev_to_local[i]);
        const size_t g = (size_t)eo + (size_t)e;
        const int right = (ce & 0x8000u) != 0;
        T->edge_off[(right ? 2 : 0) * plane + g] = L->ell_dxs[q]; T->edge_off[(right ? 3 : 1) * plane + g] = L->ell_dys[q];
        T->edge_off[(right ? 0 : 2) * plane + g] = L->ell_dxn[q]; T->edge_off[(right ? 1 : 3) * plane + g] = L->ell_dyn[q];
      }
  }
  return rdyb_tiled_layout_build_two_ring(L, T);
}

/* The second ring of every tile and the staged position of each halo cell's neighbours (device_layout.h): what K4M needs to compute the
 * least-squares gradient of a HALO cell inside the tile kernel with the arithmetic (slot order, weights) of the cell's own tile, so
 * that both tiles of a shared edge reconstruct bit-identical face values and the scheme stays conservative without a gradient array. */
int rdyb_tiled_layout_build_two_ring(const RdybLayout *L, RdybTiledLayout *T) {
  int ierr = RDYB_SUCCESS;
  int32_t *cand = NULL;
  int64_t cap2 = 0, n2 = 0;
  RDYB_CHECK(L && T && T->tile_desc, RDYB_ERR_ARG, "two-ring layout: NULL argument");
  free(T->halo2_cells); free(T->halo_nbr);
  T->halo2_cells = NULL; T->halo_nbr = NULL; T->max_halo2 = 0; T->total_halo2 = 0; T->two_ring = 0;
  const int32_t W = L->width, TC = T->tile_cells, N = L->n_owned;
  const size_t S = (size_t)L->stride;
  for (int64_t j = 0; j < T->total_halo; ++j)
    if (T->halo_cells[j] >= N) return RDYB_SUCCESS; /* a ghost in the first ring: no two-ring staging on this rank */
  T->halo_nbr = malloc(sizeof(uint16_t) * 4 * (size_t)(T->total_halo > 0 ? T->total_halo : 1));
  cand = malloc(sizeof(int32_t) * (size_t)RDYB_ELL_WIDTH * (size_t)(T->max_halo > 0 ? T->max_halo : 1));
  if (!T->halo_nbr || !cand) { ierr = rdyb_fail(RDYB_ERR_MEM, "two-ring layout: out of memory"); goto fail; }
  memset(T->halo_nbr, 0xff, sizeof(uint16_t) * 4 * (size_t)(T->total_halo > 0 ? T->total_halo : 1));
  for (int32_t t = 0; t < T->num_tiles; ++t) {
    int32_t *d = &T->tile_desc[8 * (size_t)t];
    const int32_t cb = d[0], ce = d[0] + d[1], ho = d[4], nh = d[5];
    const int32_t *H = T->halo_cells + ho;
    while (n2 & 3) { GROW(T->halo2_cells, cap2, n2 + 1); T->halo2_cells[n2++] = 0; }
    /* candidates: neighbours of the halo cells outside tile and halo */
    int32_t nc = 0;
    for (int32_t j = 0; j < nh; ++j)
      for (int32_t k = 0; k < W; ++k) {
        const int32_t nbr = L->ell_nbr[(size_t)k * S + (size_t)H[j]];
        if (nbr < 0 || (nbr >= cb && nbr < ce)) continue;
        const int32_t pos = lower_bound_i32(H, nh, nbr);
        if (pos < nh && H[pos] == nbr) continue;
        if (nbr >= N) goto fail; /* a ghost in the second ring: ierr is still RDYB_SUCCESS, the layout simply has no two-ring staging */
        cand[nc++] = nbr;
      }
    qsort(cand, (size_t)nc, sizeof(int32_t), cmp_i32);
    int32_t nu = 0;
    for (int32_t j = 0; j < nc; ++j) if (j == 0 || cand[j] != cand[j - 1]) cand[nu++] = cand[j];
    if (nu > 32766 || TC + nh > 32766) { ierr = rdyb_fail(RDYB_ERR_SUP, "two-ring layout: tile %d stages too many cells for 15-bit positions", t); goto fail; }
    GROW(T->halo2_cells, cap2, n2 + nu);
    memcpy(T->halo2_cells + n2, cand, sizeof(int32_t) * (size_t)nu);
    for (int32_t j = 0; j < nh; ++j)
      for (int32_t k = 0; k < W; ++k) {
        const int32_t nbr = L->ell_nbr[(size_t)k * S + (size_t)H[j]];
        uint16_t v = 0xffffu;
        if (nbr >= 0) {
          if (nbr >= cb && nbr < ce) v = (uint16_t)(nbr - cb);
          else {
            const int32_t pos = lower_bound_i32(H, nh, nbr);
            if (pos < nh && H[pos] == nbr) v = (uint16_t)(TC + pos);
            else v = (uint16_t)(0x8000u | (uint32_t)lower_bound_i32(cand, nu, nbr));
          }
        }
        T->halo_nbr[4 * (size_t)(ho + j) + (size_t)k] = v;
      }
    d[6] = (int32_t)n2; d[7] = nu;
    n2 += nu;
    if (nu > T->max_halo2) T->max_halo2 = nu;
    if (n2 > INT32_MAX - 65536) { ierr = rdyb_fail(RDYB_ERR_SUP, "two-ring layout: too many second-ring cells for int32 offsets"); goto fail; }
  }
  GROW(T->halo2_cells, cap2, n2 + 16);
  for (int j = 0; j < 16; ++j) T->halo2_cells[n2 + j] = 0;
  T->total_halo2 = n2 + 16;
  T->two_ring = 1;
  free(cand);
  return RDYB_SUCCESS;
fail:
  free(cand);
  free(T->halo2_cells); free(T->halo_nbr);
  T->halo2_cells = NULL; T->halo_nbr = NULL; T->max_halo2 = 0; T->two_ring = 0;
  return ierr;
}
