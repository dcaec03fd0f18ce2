/* device_layout.c — builds the device mesh layout (see device_layout.h) on the host, in C.
 *
 * Reference quantities consumed: RDyCells / RDyEdges members listed in RdybMeshView
 * (include/private/rdymeshimpl.h:26-140), RDyBoundary.edge_ids (include/private/rdyboundaryimpl.h:9-16).
 * Accumulation order per cell follows the reference operator: interior edges in internal_edge_ids order
 * (src/swe/swe_petsc.c:244-310), then boundary 0..B-1 in edge_ids order (src/operator_fluxes_petsc.c:28-51).
 */
#include "device_layout.h"

#include <math.h>
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

void rdyb_tiled_layout_free(RdybTiledLayout *T) {
  if (!T) return;
  free(T->tile_cell_begin); free(T->halo_offsets); free(T->halo_cells); free(T->edge_offsets); free(T->edge_left); free(T->edge_right);
  free(T->edge_bidx); free(T->edge_sn); free(T->edge_cn); free(T->edge_len); free(T->cell_edges);
  memset(T, 0, sizeof(*T));
}

int rdyb_tiled_layout_build(const RdybLayout *L, int32_t tile_cells, RdybTiledLayout *T) {
  (void)L; (void)tile_cells; (void)T;
  return rdyb_fail(RDYB_ERR_SUP, "tiled layout not built yet");
}
