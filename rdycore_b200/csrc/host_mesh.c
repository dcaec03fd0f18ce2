/* host_mesh.c — host-side FV mesh builder for the B200 SWE path (see rdyb200_mesh.h).
 *
 * What RDycore derives from DMPlex in src/rdymesh.c is derived here from plain
 * vertex / cell arrays.  Formulas follow the reference so that the oracle and the
 * CUDA path see the same geometry:
 *   sn, cn and edge orientation     src/rdymesh.c:628-688
 *   bed slopes                      src/rdymesh.c:747-784 (triangle), 822-858 (quad)
 *   2-D projection (HR)             src/rdymesh.c:1478-1509
 * DMPlexComputeCellGeometryFVM itself is PETSc code that is not under the
 * reference tree ("parity unpinned"): areas/centroids use the polygon formulas
 * below (3-D area of the planar polygon; triangle-fan centroid), edge lengths
 * are 3-D vertex distances, edge centroids are midpoints.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "rdyb200_mesh.h"
#include "rdyb_error.h"

struct RdybHostMesh_ {
  int32_t nv, nc, nowned, ne, nie, nbe;
  double  *xyz;                                /* [3*nv] */
  int32_t *cell_offsets, *cell_verts;          /* CSR */
  int32_t *cell_is_owned, *local_to_owned, *owned_to_local;
  int64_t *cell_gid;
  double  *area, *dzdx, *dzdy, *zc, *xc, *yc, *zcen;
  int32_t *edge_cells;                         /* [2*ne] */
  int32_t *edge_verts;                         /* [2*ne] */
  double  *len, *cn, *sn, *exc, *eyc, *ezc;
  int32_t *internal_edge_ids, *boundary_edge_ids;
  int64_t *edge_gid;
  /* halo plan (local meshes only) */
  int32_t  npeers;
  int32_t *peer_ranks, *send_off, *send_cells, *recv_off, *recv_cells;
  int      projected2d;
};

#define ALLOC(ptr, n)                                                                        \
  do {                                                                                       \
    (ptr) = calloc((size_t)((n) > 0 ? (n) : 1), sizeof(*(ptr)));                             \
    if (!(ptr)) return rdyb_fail(RDYB_ERR_MEM, "host mesh: out of memory (%s)", #ptr);       \
  } while (0)

static double sq(double x) { return x * x; }

int rdyb_hostmesh_destroy(RdybHostMesh *pm) {
  if (!pm || !*pm) return RDYB_SUCCESS;
  RdybHostMesh m = *pm;
  free(m->xyz); free(m->cell_offsets); free(m->cell_verts); free(m->cell_is_owned); free(m->local_to_owned);
  free(m->owned_to_local); free(m->cell_gid); free(m->area); free(m->dzdx); free(m->dzdy); free(m->zc);
  free(m->xc); free(m->yc); free(m->zcen); free(m->edge_cells); free(m->edge_verts); free(m->len);
  free(m->cn); free(m->sn); free(m->exc); free(m->eyc); free(m->ezc); free(m->internal_edge_ids);
  free(m->boundary_edge_ids); free(m->edge_gid); free(m->peer_ranks); free(m->send_off); free(m->send_cells);
  free(m->recv_off); free(m->recv_cells);
  free(m);
  *pm = NULL;
  return RDYB_SUCCESS;
}

/* ---------------- edge construction ---------------- */
/* open-addressing hash: key (vmin,vmax) -> edge id, in first-appearance order over cells */
typedef struct {
  uint64_t *keys;
  int32_t  *vals;
  uint64_t  mask;
} EdgeHash;

static uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}

static int build_edges(RdybHostMesh m) {
  int64_t nhalf = m->cell_offsets[m->nc];
  uint64_t cap = 16;
  while (cap < (uint64_t)nhalf * 2) cap <<= 1;
  EdgeHash h;
  h.mask = cap - 1;
  h.keys = malloc(cap * sizeof(uint64_t));
  h.vals = malloc(cap * sizeof(int32_t));
  if (!h.keys || !h.vals) { free(h.keys); free(h.vals); return rdyb_fail(RDYB_ERR_MEM, "host mesh: edge hash allocation failed"); }
  memset(h.keys, 0xff, cap * sizeof(uint64_t));
  int32_t *ecells = malloc((size_t)nhalf * 2 * sizeof(int32_t));
  int32_t *everts = malloc((size_t)nhalf * 2 * sizeof(int32_t));
  if (!ecells || !everts) { free(h.keys); free(h.vals); free(ecells); free(everts); return rdyb_fail(RDYB_ERR_MEM, "host mesh: edge allocation failed"); }
  int32_t ne = 0;
  for (int32_t c = 0; c < m->nc; ++c) {
    int32_t o = m->cell_offsets[c], n = m->cell_offsets[c + 1] - o;
    for (int32_t k = 0; k < n; ++k) {
      int32_t a = m->cell_verts[o + k], b = m->cell_verts[o + (k + 1) % n];
      uint64_t key = a < b ? ((uint64_t)(uint32_t)a << 32) | (uint32_t)b : ((uint64_t)(uint32_t)b << 32) | (uint32_t)a;
      uint64_t slot = mix64(key) & h.mask;
      while (h.keys[slot] != key && h.keys[slot] != ~0ULL) slot = (slot + 1) & h.mask;
      if (h.keys[slot] == ~0ULL) {
        h.keys[slot] = key;
        h.vals[slot] = ne;
        ecells[2 * ne] = c;
        ecells[2 * ne + 1] = -1;
        everts[2 * ne] = a;
        everts[2 * ne + 1] = b;
        ++ne;
      } else {
        int32_t e = h.vals[slot];
        if (ecells[2 * e + 1] != -1) {
          free(h.keys); free(h.vals); free(ecells); free(everts);
          return rdyb_fail(RDYB_ERR_USER, "host mesh: edge (%d,%d) is shared by more than two cells", a, b);
        }
        ecells[2 * e + 1] = c;
      }
    }
  }
  free(h.keys); free(h.vals);
  m->ne = ne;
  m->edge_cells = realloc(ecells, (size_t)(ne > 0 ? ne : 1) * 2 * sizeof(int32_t));
  m->edge_verts = realloc(everts, (size_t)(ne > 0 ? ne : 1) * 2 * sizeof(int32_t));
  return RDYB_SUCCESS;
}

/* ---------------- geometry ---------------- */
/* src/rdymesh.c:720-739 */
static int ccw(const double *p0, const double *p1, const double *p2) {
  return (p1[1] - p0[1]) * (p2[0] - p1[0]) - (p2[1] - p1[1]) * (p1[0] - p0[0]) < 0;
}
/* src/rdymesh.c:747-784 */
static void tri_slopes(const double *p0, const double *pa, const double *pb, double *dzdx, double *dzdy) {
  const double *p1 = pa, *p2 = pb;
  if (!ccw(p0, pa, pb)) { p1 = pb; p2 = pa; }
  double x0 = p0[0], y0 = p0[1], z0 = p0[2], x1 = p1[0], y1 = p1[1], z1 = p1[2], x2 = p2[0], y2 = p2[1], z2 = p2[2];
  double num = (y2 - y0) * (z1 - z0) - (y1 - y0) * (z2 - z0);
  double den = (y2 - y0) * (x1 - x0) - (y1 - y0) * (x2 - x0);
  *dzdx = num / den;
  num = (x2 - x0) * (z1 - z0) - (x1 - x0) * (z2 - z0);
  den = (x2 - x0) * (y1 - y0) - (x1 - x0) * (y2 - y0);
  *dzdy = num / den;
}
/* src/rdymesh.c:786-796 */
static double tri_area2d(const double *a, const double *b, const double *c) {
  double cz = (b[0] - a[0]) * (c[1] - a[1]) - (b[1] - a[1]) * (c[0] - a[0]);
  return 0.5 * fabs(cz);
}

static void cell_geometry(RdybHostMesh m, int32_t c) {
  int32_t o = m->cell_offsets[c], n = m->cell_offsets[c + 1] - o;
  const double *p0 = &m->xyz[3 * m->cell_verts[o]];
  /* triangle fan from vertex 0: vector area and area-weighted centroid */
  double ax = 0, ay = 0, az = 0, cx = 0, cy = 0, cz = 0, asum = 0;
  for (int32_t k = 1; k + 1 < n; ++k) {
    const double *p1 = &m->xyz[3 * m->cell_verts[o + k]], *p2 = &m->xyz[3 * m->cell_verts[o + k + 1]];
    double e1[3] = {p1[0] - p0[0], p1[1] - p0[1], p1[2] - p0[2]}, e2[3] = {p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2]};
    double nx = e1[1] * e2[2] - e1[2] * e2[1], ny = e1[2] * e2[0] - e1[0] * e2[2], nz = e1[0] * e2[1] - e1[1] * e2[0];
    double a = 0.5 * sqrt(nx * nx + ny * ny + nz * nz);
    ax += 0.5 * nx; ay += 0.5 * ny; az += 0.5 * nz;
    cx += a * (p0[0] + p1[0] + p2[0]) / 3.0;
    cy += a * (p0[1] + p1[1] + p2[1]) / 3.0;
    cz += a * (p0[2] + p1[2] + p2[2]) / 3.0;
    asum += a;
  }
  m->area[c] = sqrt(ax * ax + ay * ay + az * az);
  m->xc[c] = cx / asum; m->yc[c] = cy / asum; m->zcen[c] = cz / asum;
  if (m->projected2d) { /* src/rdymesh.c:1496-1506 */
    double twice = 0.0;
    for (int32_t k = 0; k < n; ++k) {
      const double *pc = &m->xyz[3 * m->cell_verts[o + k]], *pn = &m->xyz[3 * m->cell_verts[o + (k + 1) % n]];
      twice += pc[0] * pn[1] - pn[0] * pc[1];
    }
    m->area[c] = fabs(twice) / 2.0;
  }
  /* slopes, src/rdymesh.c:808-859 */
  if (n == 3) {
    tri_slopes(p0, &m->xyz[3 * m->cell_verts[o + 1]], &m->xyz[3 * m->cell_verts[o + 2]], &m->dzdx[c], &m->dzdy[c]);
  } else {
    double cen[3] = {m->xc[c], m->yc[c], m->zcen[c]}, sx = 0.0, sy = 0.0, total = 0.0;
    for (int32_t k = 0; k < 4; ++k) {
      const double *a = &m->xyz[3 * m->cell_verts[o + k]], *b = &m->xyz[3 * m->cell_verts[o + (k + 1) % 4]];
      double ak = tri_area2d(a, b, cen), dx, dy;
      tri_slopes(a, b, cen, &dx, &dy);
      sx += ak * dx; sy += ak * dy; total += ak;
    }
    m->dzdx[c] = sx / total;
    m->dzdy[c] = sy / total;
  }
}

/* src/rdymesh.c:590-689 */
static void edge_geometry(RdybHostMesh m, int32_t e) {
  int32_t l = m->edge_cells[2 * e], r = m->edge_cells[2 * e + 1];
  int32_t v1 = m->edge_verts[2 * e], v2 = m->edge_verts[2 * e + 1];
  const double *p1 = &m->xyz[3 * v1], *p2 = &m->xyz[3 * v2];
  double par[2] = {p2[0] - p1[0], p2[1] - p1[1]}, l2r[2];
  if (r >= 0) { l2r[0] = m->xc[r] - m->xc[l]; l2r[1] = m->yc[r] - m->yc[l]; }
  else { l2r[0] = (p2[0] + p1[0]) / 2.0 - m->xc[l]; l2r[1] = (p2[1] + p1[1]) / 2.0 - m->yc[l]; }
  double perp[2] = {par[1], -par[0]};
  if (l2r[0] * perp[0] + l2r[1] * perp[1] < 0.0) {
    m->edge_verts[2 * e] = v2; m->edge_verts[2 * e + 1] = v1;
    p1 = &m->xyz[3 * v2]; p2 = &m->xyz[3 * v1];
  }
  double dx = p2[0] - p1[0], dy = p2[1] - p1[1], dz = p2[2] - p1[2];
  double ds = sqrt(sq(dx) + sq(dy));
  m->sn[e] = -dx / ds;
  m->cn[e] = dy / ds;
  m->len[e] = m->projected2d ? ds : sqrt(sq(dx) + sq(dy) + sq(dz));
  m->exc[e] = 0.5 * (p1[0] + p2[0]); m->eyc[e] = 0.5 * (p1[1] + p2[1]); m->ezc[e] = 0.5 * (p1[2] + p2[2]);
}

static int compute_geometry(RdybHostMesh m) {
  for (int32_t c = 0; c < m->nc; ++c) cell_geometry(m, c);
  for (int32_t e = 0; e < m->ne; ++e) edge_geometry(m, e);
  return RDYB_SUCCESS;
}

static int alloc_geometry(RdybHostMesh m) {
  ALLOC(m->area, m->nc); ALLOC(m->dzdx, m->nc); ALLOC(m->dzdy, m->nc); ALLOC(m->xc, m->nc); ALLOC(m->yc, m->nc);
  ALLOC(m->zcen, m->nc); ALLOC(m->len, m->ne); ALLOC(m->cn, m->ne); ALLOC(m->sn, m->ne); ALLOC(m->exc, m->ne);
  ALLOC(m->eyc, m->ne); ALLOC(m->ezc, m->ne);
  return RDYB_SUCCESS;
}

/* internal / boundary edge id lists, src/rdymesh.c:692-710 */
static int classify_edges(RdybHostMesh m) {
  m->nie = m->nbe = 0;
  for (int32_t e = 0; e < m->ne; ++e) (m->edge_cells[2 * e + 1] >= 0) ? ++m->nie : ++m->nbe;
  ALLOC(m->internal_edge_ids, m->nie); ALLOC(m->boundary_edge_ids, m->nbe);
  int32_t i = 0, b = 0;
  for (int32_t e = 0; e < m->ne; ++e) {
    if (m->edge_cells[2 * e + 1] >= 0) m->internal_edge_ids[i++] = e;
    else m->boundary_edge_ids[b++] = e;
  }
  return RDYB_SUCCESS;
}

/* ---------------- construction ---------------- */
static int finish_serial_ownership(RdybHostMesh m) {
  ALLOC(m->cell_is_owned, m->nc); ALLOC(m->local_to_owned, m->nc); ALLOC(m->owned_to_local, m->nc);
  ALLOC(m->cell_gid, m->nc);
  for (int32_t c = 0; c < m->nc; ++c) {
    m->cell_is_owned[c] = 1; m->local_to_owned[c] = c; m->owned_to_local[c] = c; m->cell_gid[c] = c;
  }
  m->nowned = m->nc;
  return RDYB_SUCCESS;
}

static int create_from_arrays(int32_t nv, const double *xyz, int32_t nc, const int32_t *offs, const int32_t *verts, RdybHostMesh *out) {
  RDYB_CHECK(out && xyz && offs && verts, RDYB_ERR_ARG, "host mesh: NULL argument");
  RDYB_CHECK(nv > 0 && nc > 0, RDYB_ERR_USER, "host mesh: need at least one vertex and one cell");
  RdybHostMesh m = calloc(1, sizeof(*m));
  if (!m) return rdyb_fail(RDYB_ERR_MEM, "host mesh: out of memory");
  *out = m;
  m->nv = nv; m->nc = nc;
  ALLOC(m->xyz, 3 * (size_t)nv);
  memcpy(m->xyz, xyz, sizeof(double) * 3 * (size_t)nv);
  ALLOC(m->cell_offsets, nc + 1);
  memcpy(m->cell_offsets, offs, sizeof(int32_t) * ((size_t)nc + 1));
  ALLOC(m->cell_verts, offs[nc]);
  memcpy(m->cell_verts, verts, sizeof(int32_t) * (size_t)offs[nc]);
  for (int32_t c = 0; c < nc; ++c) {
    int32_t n = offs[c + 1] - offs[c];
    /* src/rdymesh.c:811 */
    RDYB_CHECK(n == 3 || n == 4, RDYB_ERR_USER, "Cell has %d vertices (must be 3 or 4)", n);
    for (int32_t k = offs[c]; k < offs[c + 1]; ++k)
      RDYB_CHECK(verts[k] >= 0 && verts[k] < nv, RDYB_ERR_USER, "host mesh: cell %d references vertex %d out of range", c, verts[k]);
  }
  return RDYB_SUCCESS;
}

int rdyb_hostmesh_create(int32_t nv, const double *xyz, int32_t nc, const int32_t *offs, const int32_t *verts, RdybHostMesh *out) {
  int ierr = create_from_arrays(nv, xyz, nc, offs, verts, out);
  if (!ierr) ierr = build_edges(*out);
  if (!ierr) ierr = alloc_geometry(*out);
  if (!ierr) ierr = compute_geometry(*out);
  if (!ierr) ierr = classify_edges(*out);
  if (!ierr) ierr = finish_serial_ownership(*out);
  if (!ierr) {
    RdybHostMesh m = *out;
    m->edge_gid = calloc((size_t)(m->ne > 0 ? m->ne : 1), sizeof(int64_t));
    if (!m->edge_gid) ierr = rdyb_fail(RDYB_ERR_MEM, "host mesh: out of memory");
    else for (int32_t e = 0; e < m->ne; ++e) m->edge_gid[e] = e; /* src/rdymesh.c:519-524 */
  }
  if (ierr && out) rdyb_hostmesh_destroy(out);
  return ierr;
}

int rdyb_hostmesh_override_2d_projection(RdybHostMesh m) {
  RDYB_CHECK(m, RDYB_ERR_ARG, "host mesh: NULL mesh");
  m->projected2d = 1;
  return compute_geometry(m);
}

int rdyb_hostmesh_set_zc(RdybHostMesh m, const double *zc) {
  RDYB_CHECK(m, RDYB_ERR_ARG, "host mesh: NULL mesh");
  if (!m->zc) ALLOC(m->zc, m->nc);
  for (int32_t c = 0; c < m->nc; ++c) {
    if (zc) { m->zc[c] = zc[c]; continue; }
    /* src/swe/swe_petsc.c:1217-1222 */
    double zsum = 0.0;
    int32_t n = m->cell_offsets[c + 1] - m->cell_offsets[c];
    for (int32_t k = m->cell_offsets[c]; k < m->cell_offsets[c + 1]; ++k) zsum += m->xyz[3 * m->cell_verts[k] + 2];
    m->zc[c] = zsum / (double)n;
  }
  return RDYB_SUCCESS;
}

int rdyb_hostmesh_set_vertex_z(RdybHostMesh m, const double *z) {
  RDYB_CHECK(m && z, RDYB_ERR_ARG, "host mesh: NULL argument");
  for (int32_t v = 0; v < m->nv; ++v) m->xyz[3 * v + 2] = z[v];
  /* orientation was fixed with x,y only, so recomputing geometry keeps edge_verts as they are */
  RDYB_CALL(compute_geometry(m));
  if (m->zc) RDYB_CALL(rdyb_hostmesh_set_zc(m, NULL));
  return RDYB_SUCCESS;
}

int rdyb_hostmesh_vertices(RdybHostMesh m, const double **xyz, int32_t *nv) {
  RDYB_CHECK(m, RDYB_ERR_ARG, "host mesh: NULL mesh");
  if (xyz) *xyz = m->xyz;
  if (nv) *nv = m->nv;
  return RDYB_SUCCESS;
}

int rdyb_hostmesh_view(RdybHostMesh m, RdybMeshView *v) {
  RDYB_CHECK(m && v, RDYB_ERR_ARG, "host mesh: NULL argument");
  memset(v, 0, sizeof(*v));
  v->num_cells = m->nc; v->num_owned_cells = m->nowned; v->num_edges = m->ne; v->num_internal_edges = m->nie;
  v->cell_is_owned = m->cell_is_owned; v->cell_local_to_owned = m->local_to_owned; v->cell_global_ids = m->cell_gid;
  v->cell_areas = m->area; v->cell_dz_dx = m->dzdx; v->cell_dz_dy = m->dzdy; v->cell_zc = m->zc;
  v->cell_xc = m->xc; v->cell_yc = m->yc;
  v->edge_cell_ids = m->edge_cells; v->edge_lengths = m->len; v->edge_cn = m->cn; v->edge_sn = m->sn;
  v->internal_edge_ids = m->internal_edge_ids; v->edge_global_ids = m->edge_gid;
  v->edge_xm = m->exc; v->edge_ym = m->eyc; /* midpoints 0.5*(v0+v1) */
  return RDYB_SUCCESS;
}

int rdyb_hostmesh_halo(RdybHostMesh m, RdybHaloView *h) {
  RDYB_CHECK(m && h, RDYB_ERR_ARG, "host mesh: NULL argument");
  h->num_peers = m->npeers; h->peer_ranks = m->peer_ranks; h->send_offsets = m->send_off; h->send_cells = m->send_cells;
  h->recv_offsets = m->recv_off; h->recv_cells = m->recv_cells;
  return RDYB_SUCCESS;
}

int rdyb_hostmesh_sizes(RdybHostMesh m, int32_t out[6]) {
  RDYB_CHECK(m && out, RDYB_ERR_ARG, "host mesh: NULL argument");
  out[0] = m->nv; out[1] = m->nc; out[2] = m->nowned; out[3] = m->ne; out[4] = m->nie; out[5] = m->nbe;
  return RDYB_SUCCESS;
}

int rdyb_hostmesh_array(RdybHostMesh m, const char *name, const void **ptr, int64_t *length) {
  RDYB_CHECK(m && name && ptr && length, RDYB_ERR_ARG, "host mesh: NULL argument");
#define ARR(nm, p, n) if (!strcmp(name, nm)) { *ptr = (p); *length = (n); return RDYB_SUCCESS; }
  ARR("boundary_edge_ids", m->boundary_edge_ids, m->nbe)
  ARR("edge_xc", m->exc, m->ne) ARR("edge_yc", m->eyc, m->ne) ARR("edge_zc", m->ezc, m->ne)
  ARR("cell_xc", m->xc, m->nc) ARR("cell_yc", m->yc, m->nc) ARR("cell_zcen", m->zcen, m->nc)
  ARR("edge_vertex_ids", m->edge_verts, 2 * (int64_t)m->ne)
  ARR("owned_to_local", m->owned_to_local, m->nowned)
  ARR("cell_global_ids", m->cell_gid, m->nc)
  ARR("cell_offsets", m->cell_offsets, (int64_t)m->nc + 1)
  ARR("cell_vertex_ids", m->cell_verts, m->cell_offsets[m->nc])
#undef ARR
  return rdyb_fail(RDYB_ERR_ARG, "host mesh: unknown array '%s'", name);
}

/* ---------------- partitioning ---------------- */
typedef struct { double key; int32_t idx; } KeyIdx;
static int keyidx_less(const KeyIdx *a, const KeyIdx *b) { return a->key < b->key || (a->key == b->key && a->idx < b->idx); }

/* quickselect: after the call items [0,k) are <= items [k,n) in (key, idx) order */
static void select_k(KeyIdx *a, int64_t n, int64_t k) {
  int64_t lo = 0, hi = n - 1;
  while (lo < hi) {
    KeyIdx p = a[lo + (hi - lo) / 2];
    int64_t i = lo, j = hi;
    while (i <= j) {
      while (keyidx_less(&a[i], &p)) ++i;
      while (keyidx_less(&p, &a[j])) --j;
      if (i <= j) { KeyIdx t = a[i]; a[i] = a[j]; a[j] = t; ++i; --j; }
    }
    if (k <= j) hi = j; else if (k >= i) lo = i; else return;
  }
}

static void rcb(RdybHostMesh m, KeyIdx *items, int64_t n, int32_t p0, int32_t np, int32_t *part) {
  if (np == 1) { for (int64_t i = 0; i < n; ++i) part[items[i].idx] = p0; return; }
  double xmin = 1e300, xmax = -1e300, ymin = 1e300, ymax = -1e300;
  for (int64_t i = 0; i < n; ++i) {
    int32_t c = items[i].idx;
    if (m->xc[c] < xmin) xmin = m->xc[c];
    if (m->xc[c] > xmax) xmax = m->xc[c];
    if (m->yc[c] < ymin) ymin = m->yc[c];
    if (m->yc[c] > ymax) ymax = m->yc[c];
  }
  const double *coord = (xmax - xmin >= ymax - ymin) ? m->xc : m->yc;
  for (int64_t i = 0; i < n; ++i) items[i].key = coord[items[i].idx];
  int32_t npl = np / 2;
  int64_t k = (int64_t)((double)n * (double)npl / (double)np + 0.5);
  select_k(items, n, k);
  rcb(m, items, k, p0, npl, part);
  rcb(m, items + k, n - k, p0 + npl, np - npl, part);
}

int rdyb_hostmesh_partition_rcb(RdybHostMesh m, int32_t nparts, int32_t *part) {
  RDYB_CHECK(m && part, RDYB_ERR_ARG, "partition: NULL argument");
  RDYB_CHECK(nparts >= 1 && nparts <= m->nc, RDYB_ERR_USER, "partition: nparts (%d) must be in [1, num_cells]", nparts);
  KeyIdx *items = malloc(sizeof(KeyIdx) * (size_t)m->nc);
  if (!items) return rdyb_fail(RDYB_ERR_MEM, "partition: out of memory");
  for (int32_t c = 0; c < m->nc; ++c) { items[c].idx = c; items[c].key = 0.0; }
  rcb(m, items, m->nc, 0, nparts, part);
  free(items);
  return RDYB_SUCCESS;
}

typedef struct { int64_t gid; int32_t owner, cell; } GhostRec;
static int ghost_cmp(const void *pa, const void *pb) {
  const GhostRec *a = pa, *b = pb;
  if (a->owner != b->owner) return a->owner < b->owner ? -1 : 1;
  if (a->gid != b->gid) return a->gid < b->gid ? -1 : 1;
  return 0;
}

int rdyb_hostmesh_extract_local(RdybHostMesh G, const int32_t *part, int32_t nparts, int32_t rank, RdybHostMesh *out) {
  RDYB_CHECK(G && part && out, RDYB_ERR_ARG, "extract_local: NULL argument");
  RDYB_CHECK(rank >= 0 && rank < nparts, RDYB_ERR_USER, "extract_local: rank %d out of range", rank);
  int ierr = RDYB_SUCCESS;
  int32_t nown = 0;
  for (int32_t c = 0; c < G->nc; ++c) nown += (part[c] == rank);
  RDYB_CHECK(nown > 0, RDYB_ERR_USER, "extract_local: rank %d owns no cells", rank);
  GhostRec *own = malloc(sizeof(GhostRec) * (size_t)nown), *gh = NULL, *snd = NULL;
  int32_t *g2l = malloc(sizeof(int32_t) * (size_t)G->nc), *vmap = NULL, *offs = NULL, *verts = NULL;
  double  *xyz = NULL;
  RdybHostMesh L = NULL;
  if (!own || !g2l) { ierr = rdyb_fail(RDYB_ERR_MEM, "extract_local: out of memory"); goto done; }
  for (int32_t c = 0, k = 0; c < G->nc; ++c) {
    g2l[c] = -1;
    if (part[c] == rank) { own[k].gid = G->cell_gid[c]; own[k].owner = rank; own[k].cell = c; ++k; }
  }
  qsort(own, (size_t)nown, sizeof(GhostRec), ghost_cmp);
  for (int32_t k = 0; k < nown; ++k) g2l[own[k].cell] = k;
  /* ghosts = face neighbours owned elsewhere; sends = (peer, my cell) pairs across partition edges */
  int32_t ncut = 0;
  for (int32_t e = 0; e < G->ne; ++e) {
    int32_t l = G->edge_cells[2 * e], r = G->edge_cells[2 * e + 1];
    if (r >= 0 && ((part[l] == rank) != (part[r] == rank))) ++ncut;
  }
  gh  = malloc(sizeof(GhostRec) * (size_t)(ncut > 0 ? ncut : 1));
  snd = malloc(sizeof(GhostRec) * (size_t)(ncut > 0 ? ncut : 1));
  if (!gh || !snd) { ierr = rdyb_fail(RDYB_ERR_MEM, "extract_local: out of memory"); goto done; }
  int32_t ng = 0, ns = 0;
  for (int32_t e = 0; e < G->ne; ++e) {
    int32_t l = G->edge_cells[2 * e], r = G->edge_cells[2 * e + 1];
    if (r < 0 || ((part[l] == rank) == (part[r] == rank))) continue;
    int32_t mine = part[l] == rank ? l : r, other = part[l] == rank ? r : l;
    gh[ng].gid = G->cell_gid[other]; gh[ng].owner = part[other]; gh[ng].cell = other; ++ng;
    snd[ns].gid = G->cell_gid[mine]; snd[ns].owner = part[other]; snd[ns].cell = mine; ++ns;
  }
  qsort(gh, (size_t)ng, sizeof(GhostRec), ghost_cmp);
  qsort(snd, (size_t)ns, sizeof(GhostRec), ghost_cmp);
  int32_t nguniq = 0, nsuniq = 0;
  for (int32_t k = 0; k < ng; ++k) if (k == 0 || ghost_cmp(&gh[k], &gh[k - 1]) != 0) gh[nguniq++] = gh[k];
  for (int32_t k = 0; k < ns; ++k) if (k == 0 || ghost_cmp(&snd[k], &snd[k - 1]) != 0) snd[nsuniq++] = snd[k];
  ng = nguniq; ns = nsuniq;
  for (int32_t k = 0; k < ng; ++k) g2l[gh[k].cell] = nown + k;
  int32_t nloc = nown + ng;

  /* local vertices in first-use order, local cells */
  vmap = malloc(sizeof(int32_t) * (size_t)G->nv);
  offs = malloc(sizeof(int32_t) * ((size_t)nloc + 1));
  if (!vmap || !offs) { ierr = rdyb_fail(RDYB_ERR_MEM, "extract_local: out of memory"); goto done; }
  for (int32_t v = 0; v < G->nv; ++v) vmap[v] = -1;
  int32_t nvl = 0, nconn = 0;
  offs[0] = 0;
  for (int32_t k = 0; k < nloc; ++k) {
    int32_t c = k < nown ? own[k].cell : gh[k - nown].cell;
    nconn += G->cell_offsets[c + 1] - G->cell_offsets[c];
    offs[k + 1] = nconn;
  }
  verts = malloc(sizeof(int32_t) * (size_t)nconn);
  xyz   = malloc(sizeof(double) * 3 * (size_t)nconn);
  if (!verts || !xyz) { ierr = rdyb_fail(RDYB_ERR_MEM, "extract_local: out of memory"); goto done; }
  for (int32_t k = 0; k < nloc; ++k) {
    int32_t c = k < nown ? own[k].cell : gh[k - nown].cell;
    for (int32_t j = G->cell_offsets[c], q = offs[k]; j < G->cell_offsets[c + 1]; ++j, ++q) {
      int32_t v = G->cell_verts[j];
      if (vmap[v] < 0) { memcpy(&xyz[3 * nvl], &G->xyz[3 * v], 3 * sizeof(double)); vmap[v] = nvl++; }
      verts[q] = vmap[v];
    }
  }
  ierr = create_from_arrays(nvl, xyz, nloc, offs, verts, &L);
  if (ierr) goto done;
  L->projected2d = G->projected2d;
  ierr = build_edges(L);
  if (ierr) goto done;
  /* keep only edges touching an owned cell; orient left = smaller global id so that both
   * ranks sharing an edge evaluate the identical flux (swe_petsc.c:272-274 relies on that) */
  {
    int32_t keep = 0;
    for (int32_t e = 0; e < L->ne; ++e) {
      int32_t l = L->edge_cells[2 * e], r = L->edge_cells[2 * e + 1];
      if (!(l < nown || (r >= 0 && r < nown))) continue;
      if (r >= 0) {
        int64_t gl = l < nown ? own[l].gid : gh[l - nown].gid, gr = r < nown ? own[r].gid : gh[r - nown].gid;
        if (gl > gr) { int32_t t = l; l = r; r = t; }
      }
      L->edge_cells[2 * keep] = l; L->edge_cells[2 * keep + 1] = r;
      L->edge_verts[2 * keep] = L->edge_verts[2 * e]; L->edge_verts[2 * keep + 1] = L->edge_verts[2 * e + 1];
      ++keep;
    }
    L->ne = keep;
  }
  if ((ierr = alloc_geometry(L)) || (ierr = compute_geometry(L)) || (ierr = classify_edges(L))) goto done;
  L->nowned = nown;
  L->cell_is_owned = calloc((size_t)nloc, sizeof(int32_t)); L->local_to_owned = calloc((size_t)nloc, sizeof(int32_t));
  L->owned_to_local = calloc((size_t)nown, sizeof(int32_t)); L->cell_gid = calloc((size_t)nloc, sizeof(int64_t));
  L->edge_gid = calloc((size_t)(L->ne > 0 ? L->ne : 1), sizeof(int64_t));
  L->peer_ranks = calloc((size_t)nparts, sizeof(int32_t)); L->send_off = calloc((size_t)nparts + 1, sizeof(int32_t));
  L->recv_off = calloc((size_t)nparts + 1, sizeof(int32_t));
  L->send_cells = calloc((size_t)(ns > 0 ? ns : 1), sizeof(int32_t)); L->recv_cells = calloc((size_t)(ng > 0 ? ng : 1), sizeof(int32_t));
  if (!L->cell_is_owned || !L->local_to_owned || !L->owned_to_local || !L->cell_gid || !L->edge_gid || !L->peer_ranks || !L->send_off ||
      !L->recv_off || !L->send_cells || !L->recv_cells) { ierr = rdyb_fail(RDYB_ERR_MEM, "extract_local: out of memory"); goto done; }
  for (int32_t k = 0; k < nloc; ++k) {
    L->cell_is_owned[k] = k < nown; L->local_to_owned[k] = k; /* owned first, ghosts after: src/rdymesh.c:159-177 */
    L->cell_gid[k] = k < nown ? own[k].gid : gh[k - nown].gid;
    if (k < nown) L->owned_to_local[k] = k;
  }
  for (int32_t e = 0; e < L->ne; ++e) L->edge_gid[e] = e;
  /* peers = union of ghost owners (== send targets, adjacency is symmetric) */
  {
    int32_t np = 0, ig = 0, is = 0;
    while (ig < ng || is < ns) {
      int32_t p = (ig < ng && (is >= ns || gh[ig].owner <= snd[is].owner)) ? gh[ig].owner : snd[is].owner;
      L->peer_ranks[np] = p; L->recv_off[np] = ig; L->send_off[np] = is;
      while (ig < ng && gh[ig].owner == p) { L->recv_cells[ig] = nown + ig; ++ig; }
      while (is < ns && snd[is].owner == p) { L->send_cells[is] = g2l[snd[is].cell]; ++is; }
      ++np;
    }
    L->recv_off[np] = ng; L->send_off[np] = ns; L->npeers = np;
  }
  if (G->zc) {
    L->zc = calloc((size_t)nloc, sizeof(double));
    if (!L->zc) { ierr = rdyb_fail(RDYB_ERR_MEM, "extract_local: out of memory"); goto done; }
    for (int32_t k = 0; k < nloc; ++k) L->zc[k] = G->zc[k < nown ? own[k].cell : gh[k - nown].cell];
  }
  *out = L;
  L = NULL;
done:
  free(own); free(gh); free(snd); free(g2l); free(vmap); free(offs); free(verts); free(xyz);
  if (L) rdyb_hostmesh_destroy(&L);
  return ierr;
}

/* ---------------- generators ---------------- */
int rdyb_hostmesh_generate_quads(int32_t nx, int32_t ny, double lx, double ly, RdybHostMesh *out) {
  RDYB_CHECK(nx > 0 && ny > 0 && lx > 0 && ly > 0, RDYB_ERR_USER, "generate_quads: sizes must be positive");
  int64_t nv = ((int64_t)nx + 1) * (ny + 1), nc = (int64_t)nx * ny;
  RDYB_CHECK(4 * nc < INT32_MAX, RDYB_ERR_USER, "generate_quads: mesh too large for int32 local ids");
  double  *xyz = malloc(sizeof(double) * 3 * (size_t)nv);
  int32_t *offs = malloc(sizeof(int32_t) * ((size_t)nc + 1)), *verts = malloc(sizeof(int32_t) * 4 * (size_t)nc);
  if (!xyz || !offs || !verts) { free(xyz); free(offs); free(verts); return rdyb_fail(RDYB_ERR_MEM, "generate_quads: out of memory"); }
  for (int32_t j = 0; j <= ny; ++j)
    for (int32_t i = 0; i <= nx; ++i) {
      int64_t v = (int64_t)j * (nx + 1) + i;
      xyz[3 * v] = lx * (double)i / (double)nx; xyz[3 * v + 1] = ly * (double)j / (double)ny; xyz[3 * v + 2] = 0.0;
    }
  for (int32_t j = 0; j < ny; ++j)
    for (int32_t i = 0; i < nx; ++i) {
      int64_t c = (int64_t)j * nx + i;
      int32_t v0 = j * (nx + 1) + i;
      offs[c] = (int32_t)(4 * c);
      verts[4 * c] = v0; verts[4 * c + 1] = v0 + 1; verts[4 * c + 2] = v0 + nx + 2; verts[4 * c + 3] = v0 + nx + 1;
    }
  offs[nc] = (int32_t)(4 * nc);
  int ierr = rdyb_hostmesh_create((int32_t)nv, xyz, (int32_t)nc, offs, verts, out);
  free(xyz); free(offs); free(verts);
  return ierr;
}

/* counter-based uniform(0,1): xorshift64* scrambling of (seed, counter) */
static double u01(uint64_t seed, uint64_t ctr) {
  uint64_t x = mix64(seed ^ (ctr * 0x9e3779b97f4a7c15ULL + 0x632be59bd9b4e019ULL));
  if (!x) x = 0x2545F4914F6CDD1DULL;
  x ^= x >> 12; x ^= x << 25; x ^= x >> 27;
  x *= 0x2545F4914F6CDD1DULL;
  return (double)(x >> 11) * (1.0 / 9007199254740992.0);
}

static void block_range(int32_t n, int32_t p, int32_t b, int32_t *lo, int32_t *hi) {
  *lo = (int32_t)((int64_t)n * b / p);
  *hi = (int32_t)((int64_t)n * (b + 1) / p);
}
static int32_t block_of(int32_t n, int32_t p, int32_t i) {
  int32_t b = (int32_t)(((int64_t)i * p) / n);
  int32_t lo, hi;
  for (;;) {
    block_range(n, p, b, &lo, &hi);
    if (i < lo) --b; else if (i >= hi) ++b; else return b;
  }
}

int rdyb_hostmesh_generate_floodplain(int32_t nx, int32_t ny, double lx, double ly, double jitter, double zamp, uint64_t seed, int32_t px,
                                      int32_t py, int32_t rank, RdybHostMesh *out) {
  RDYB_CHECK(nx > 0 && ny > 0 && px > 0 && py > 0 && px <= nx && py <= ny, RDYB_ERR_USER, "generate_floodplain: bad sizes");
  RDYB_CHECK(rank >= 0 && rank < px * py, RDYB_ERR_USER, "generate_floodplain: rank %d out of range for %dx%d blocks", rank, px, py);
  int32_t bi = rank % px, bj = rank / px, i0, i1, j0, j1;
  block_range(nx, px, bi, &i0, &i1);
  block_range(ny, py, bj, &j0, &j1);
  /* extended block: one ring of quads (clipped), enough to contain every face-neighbour triangle */
  int32_t ei0 = i0 > 0 ? i0 - 1 : 0, ei1 = i1 < nx ? i1 + 1 : nx, ej0 = j0 > 0 ? j0 - 1 : 0, ej1 = j1 < ny ? j1 + 1 : ny;
  int32_t enx = ei1 - ei0, eny = ej1 - ej0;
  int64_t nv = ((int64_t)enx + 1) * (eny + 1), nc = 2 * (int64_t)enx * eny;
  RDYB_CHECK(3 * nc < INT32_MAX, RDYB_ERR_USER, "generate_floodplain: block too large for int32 local ids");
  const double dx = lx / nx, dy = ly / ny, twopi = 6.283185307179586;
  double  *xyz = malloc(sizeof(double) * 3 * (size_t)nv);
  int32_t *offs = malloc(sizeof(int32_t) * ((size_t)nc + 1)), *verts = malloc(sizeof(int32_t) * 3 * (size_t)nc);
  int32_t *part = malloc(sizeof(int32_t) * (size_t)nc);
  int64_t *gids = malloc(sizeof(int64_t) * (size_t)nc);
  RdybHostMesh E = NULL;
  int ierr = RDYB_SUCCESS;
  if (!xyz || !offs || !verts || !part || !gids) { ierr = rdyb_fail(RDYB_ERR_MEM, "generate_floodplain: out of memory"); goto done; }
  for (int32_t j = 0; j <= eny; ++j)
    for (int32_t i = 0; i <= enx; ++i) {
      int32_t gi = ei0 + i, gj = ej0 + j;
      uint64_t gv = (uint64_t)gj * ((uint64_t)nx + 1) + (uint64_t)gi;
      double x = dx * gi, y = dy * gj;
      if (gi > 0 && gi < nx && gj > 0 && gj < ny) {
        x += (2.0 * u01(seed, 2 * gv) - 1.0) * jitter * dx;
        y += (2.0 * u01(seed, 2 * gv + 1) - 1.0) * jitter * dy;
      }
      int64_t v = (int64_t)j * (enx + 1) + i;
      xyz[3 * v] = x; xyz[3 * v + 1] = y;
      xyz[3 * v + 2] = 20.0 * (x / lx) + zamp * 5.0 * sin(twopi * x / 3000.0) * cos(twopi * y / 2000.0) + 2.0 * u01(seed + 11111, gv);
    }
  for (int32_t j = 0; j < eny; ++j)
    for (int32_t i = 0; i < enx; ++i) {
      int32_t gi = ei0 + i, gj = ej0 + j;
      int64_t q = (int64_t)j * enx + i;
      int32_t v00 = j * (enx + 1) + i, v10 = v00 + 1, v01 = v00 + enx + 1, v11 = v01 + 1;
      int32_t *t0 = &verts[6 * q], *t1 = t0 + 3;
      if (((gi + gj) & 1) == 0) { t0[0] = v00; t0[1] = v10; t0[2] = v11; t1[0] = v00; t1[1] = v11; t1[2] = v01; }
      else { t0[0] = v00; t0[1] = v10; t0[2] = v01; t1[0] = v10; t1[1] = v11; t1[2] = v01; }
      offs[2 * q] = (int32_t)(6 * q); offs[2 * q + 1] = (int32_t)(6 * q + 3);
      int32_t owner = block_of(ny, py, gj) * px + block_of(nx, px, gi);
      part[2 * q] = part[2 * q + 1] = owner;
      gids[2 * q] = 2 * ((int64_t)gj * nx + gi); gids[2 * q + 1] = gids[2 * q] + 1;
    }
  offs[nc] = (int32_t)(3 * nc);
  ierr = rdyb_hostmesh_create((int32_t)nv, xyz, (int32_t)nc, offs, verts, &E);
  if (ierr) goto done;
  memcpy(E->cell_gid, gids, sizeof(int64_t) * (size_t)nc);
  if (px * py == 1) { *out = E; E = NULL; goto done; }
  ierr = rdyb_hostmesh_extract_local(E, part, px * py, rank, out);
done:
  free(xyz); free(offs); free(verts); free(part); free(gids);
  if (E) rdyb_hostmesh_destroy(&E);
  return ierr;
}

/* regular 1->4 refinement of a triangle mesh */
int rdyb_hostmesh_refine_triangles(RdybHostMesh C, RdybHostMesh *out) {
  RDYB_CHECK(C && out, RDYB_ERR_ARG, "refine: NULL argument");
  RDYB_CHECK(C->cell_offsets[C->nc] == 3 * C->nc, RDYB_ERR_USER, "refine: only all-triangle meshes are supported");
  uint64_t cap = 16;
  while (cap < (uint64_t)C->ne * 2) cap <<= 1;
  uint64_t *keys = malloc(cap * sizeof(uint64_t));
  int32_t  *vals = malloc(cap * sizeof(int32_t));
  int64_t   nv = (int64_t)C->nv + C->ne, nc = 4 * (int64_t)C->nc;
  double   *xyz = malloc(sizeof(double) * 3 * (size_t)nv);
  int32_t  *offs = malloc(sizeof(int32_t) * ((size_t)nc + 1)), *verts = malloc(sizeof(int32_t) * 3 * (size_t)nc);
  if (!keys || !vals || !xyz || !offs || !verts) {
    free(keys); free(vals); free(xyz); free(offs); free(verts);
    return rdyb_fail(RDYB_ERR_MEM, "refine: out of memory");
  }
  memset(keys, 0xff, cap * sizeof(uint64_t));
  memcpy(xyz, C->xyz, sizeof(double) * 3 * (size_t)C->nv);
  for (int32_t e = 0; e < C->ne; ++e) {
    int32_t a = C->edge_verts[2 * e], b = C->edge_verts[2 * e + 1];
    uint64_t key = a < b ? ((uint64_t)(uint32_t)a << 32) | (uint32_t)b : ((uint64_t)(uint32_t)b << 32) | (uint32_t)a;
    uint64_t slot = mix64(key) & (cap - 1);
    while (keys[slot] != ~0ULL) slot = (slot + 1) & (cap - 1);
    keys[slot] = key; vals[slot] = C->nv + e;
    for (int d = 0; d < 3; ++d) xyz[3 * ((int64_t)C->nv + e) + d] = 0.5 * (C->xyz[3 * a + d] + C->xyz[3 * b + d]);
  }
  for (int32_t c = 0; c < C->nc; ++c) {
    int32_t v[3], mid[3];
    for (int k = 0; k < 3; ++k) v[k] = C->cell_verts[3 * c + k];
    for (int k = 0; k < 3; ++k) {
      int32_t a = v[k], b = v[(k + 1) % 3];
      uint64_t key = a < b ? ((uint64_t)(uint32_t)a << 32) | (uint32_t)b : ((uint64_t)(uint32_t)b << 32) | (uint32_t)a;
      uint64_t slot = mix64(key) & (cap - 1);
      while (keys[slot] != key) slot = (slot + 1) & (cap - 1);
      mid[k] = vals[slot];
    }
    int32_t *t = &verts[12 * (int64_t)c];
    t[0] = v[0]; t[1] = mid[0]; t[2] = mid[2];
    t[3] = mid[0]; t[4] = v[1]; t[5] = mid[1];
    t[6] = mid[2]; t[7] = mid[1]; t[8] = v[2];
    t[9] = mid[0]; t[10] = mid[1]; t[11] = mid[2];
    for (int k = 0; k < 4; ++k) offs[4 * (int64_t)c + k] = (int32_t)(12 * (int64_t)c + 3 * k);
  }
  offs[nc] = (int32_t)(3 * nc);
  int ierr = rdyb_hostmesh_create((int32_t)nv, xyz, (int32_t)nc, offs, verts, out);
  free(keys); free(vals); free(xyz); free(offs); free(verts);
  return ierr;
}
