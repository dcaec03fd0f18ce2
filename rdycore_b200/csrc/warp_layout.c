/* warp_layout.c — host-side construction of the warp-tile layout (kernel K5) on top of an RdybLayout.
 *
 * The reference walks mesh->edges.internal_edge_ids and scatters into cells (src/swe/swe_petsc.c:244-310); here the
 * same edges are regrouped per 32-cell mini-tile so that one warp owns everything its cells need.  Accumulation order
 * per cell stays the reference's (ascending edge number = ELL slot order).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "device_layout.h"
#include "rdyb_error.h"

void rdyb_warp_layout_free(RdybWarpLayout *W) {
  if (!W) return;
  free(W->tile_desc); free(W->halo_cells); free(W->wdesc); free(W->whalo); free(W->edge_lr); free(W->edge_bidx); free(W->edge_id);
  free(W->edge_cell); free(W->edge_sn); free(W->edge_cn); free(W->edge_len); free(W->cell_edges);
  memset(W, 0, sizeof(*W));
}

static int cmp_i32w(const void *a, const void *b) {
  int32_t x = *(const int32_t *)a, y = *(const int32_t *)b;
  return x < y ? -1 : x > y;
}
static int32_t lower_bound_w(const int32_t *a, int32_t n, int32_t key) {
  int32_t lo = 0, hi = n;
  while (lo < hi) { int32_t mid = (lo + hi) / 2; if (a[mid] < key) lo = mid + 1; else hi = mid; }
  return lo;
}
static int32_t sort_unique(int32_t *a, int32_t n) {
  qsort(a, (size_t)n, sizeof(int32_t), cmp_i32w);
  int32_t m = 0;
  for (int32_t j = 0; j < n; ++j) if (j == 0 || a[j] != a[j - 1]) a[m++] = a[j];
  return m;
}

#define WGROW(ptr, cap, need)                                                                             \
  do {                                                                                                    \
    if ((int64_t)(need) > (cap)) {                                                                        \
      int64_t nc_ = (cap) + (cap) / 2 + 4096;                                                             \
      if (nc_ < (int64_t)(need)) nc_ = (need);                                                            \
      void *np_ = realloc((ptr), (size_t)nc_ * sizeof(*(ptr)));                                           \
      if (!np_) { ierr = rdyb_fail(RDYB_ERR_MEM, "warp layout: out of memory (%s)", #ptr); goto fail; }   \
      (ptr) = np_;                                                                                        \
      (cap) = nc_;                                                                                        \
    }                                                                                                     \
  } while (0)

/* grow all per-edge arrays together */
static int wedges_reserve(RdybWarpLayout *W, int64_t *cap, int64_t need) {
  if (need <= *cap) return RDYB_SUCCESS;
  int64_t ncap = *cap + *cap / 2 + 4096;
  if (ncap < need) ncap = need;
  void *p;
#define RES(field, type)                                                                 \
  p = realloc(W->field, (size_t)ncap * sizeof(type));                                    \
  if (!p) return rdyb_fail(RDYB_ERR_MEM, "warp layout: out of memory (edges)");          \
  W->field = p;
  RES(edge_lr, uint16_t) RES(edge_bidx, int32_t) RES(edge_sn, double) RES(edge_cn, double) RES(edge_len, double) RES(edge_id, int32_t)
  RES(edge_cell, int32_t)
#undef RES
  *cap = ncap;
  return RDYB_SUCCESS;
}

static void wedge_pad(RdybWarpLayout *W, int64_t g) {
  W->edge_lr[g] = 0; W->edge_bidx[g] = -1; W->edge_sn[g] = 0.0; W->edge_cn[g] = 1.0; W->edge_len[g] = 0.0; W->edge_id[g] = -1; W->edge_cell[g] = -1;
}

int rdyb_warp_layout_build(const RdybLayout *L, int32_t tile_cells, RdybWarpLayout *W) {
  int ierr = RDYB_SUCCESS;
  int32_t *tile_halo = NULL, *warp_halo = NULL;
  RDYB_CHECK(L && W, RDYB_ERR_ARG, "warp layout: NULL argument");
  RDYB_CHECK(tile_cells >= 64 && tile_cells <= 1024 && tile_cells % 64 == 0, RDYB_ERR_ARG, "warp layout: tile_cells (%d) must be a multiple of 64 in [64, 1024]", tile_cells);
  memset(W, 0, sizeof(*W));
  const int32_t N = L->n_owned, NI = L->n_interior, WD = L->width, TC = tile_cells, WC = RDYB_WARP_CELLS, WPT = TC / WC;
  const size_t S = (size_t)L->stride;
  const int32_t nt_int = (NI + TC - 1) / TC, nt_fr = (N - NI + TC - 1) / TC, nt = nt_int + nt_fr;
  W->tile_cells = TC; W->num_tiles = nt; W->num_interior_tiles = nt_int;
#define WALLOC(ptr, n)                                                                                      \
  do {                                                                                                      \
    (ptr) = calloc((size_t)((n) > 0 ? (n) : 1), sizeof(*(ptr)));                                            \
    if (!(ptr)) { ierr = rdyb_fail(RDYB_ERR_MEM, "warp layout: out of memory (%s)", #ptr); goto fail; }     \
  } while (0)
  WALLOC(W->tile_desc, 8 * (size_t)nt);
  WALLOC(W->wdesc, 4 * (size_t)WPT * (size_t)nt + 64);
  WALLOC(W->cell_edges, (size_t)RDYB_ELL_WIDTH * S);
  memset(W->cell_edges, 0xff, (size_t)RDYB_ELL_WIDTH * S);
  WALLOC(tile_halo, (size_t)TC * RDYB_ELL_WIDTH);
  WALLOC(warp_halo, (size_t)WC * RDYB_ELL_WIDTH);
  int64_t hcap = 0, ecap = 0, wcap = 0, nh = 0, ne = 0, nw = 0;
  int32_t slot_edge[RDYB_WARP_CELLS * RDYB_ELL_WIDTH];
  for (int32_t t = 0; t < nt; ++t) {
    const int32_t cb = t < nt_int ? t * TC : NI + (t - nt_int) * TC;
    const int32_t lim = t < nt_int ? NI : N;
    const int32_t ce = cb + TC < lim ? cb + TC : lim, ncell = ce - cb;
    /* tile blocks start on multiples of 8 entries: 16-byte aligned slices of the uint16 arrays (64 bytes of the doubles) */
    while (nh & 7) { WGROW(W->halo_cells, hcap, nh + 1); W->halo_cells[nh++] = 0; }
    while (nw & 7) { WGROW(W->whalo, wcap, nw + 1); W->whalo[nw++] = 0; }
    if (ne & 7) {
      const int64_t pad = 8 - (ne & 7);
      if ((ierr = wedges_reserve(W, &ecap, ne + pad))) goto fail;
      for (int64_t g = ne; g < ne + pad; ++g) wedge_pad(W, g);
      ne += pad;
    }
    const int64_t tile_e0 = ne, tile_w0 = nw, tile_h0 = nh;
    /* tile halo = neighbours outside [cb,ce), sorted unique; staged at index TC + position */
    int32_t nth = 0;
    for (int32_t i = cb; i < ce; ++i)
      for (int32_t k = 0; k < WD; ++k) {
        const int32_t nbr = L->ell_nbr[(size_t)k * S + i];
        if (nbr >= 0 && (nbr < cb || nbr >= ce)) tile_halo[nth++] = nbr;
      }
    nth = sort_unique(tile_halo, nth);
    if (TC + nth > 65534) { ierr = rdyb_fail(RDYB_ERR_SUP, "warp layout: tile %d stages %d cells (> 65534)", t, TC + nth); goto fail; }
    WGROW(W->halo_cells, hcap, nh + nth);
    memcpy(W->halo_cells + nh, tile_halo, sizeof(int32_t) * (size_t)nth);
    nh += nth;
    for (int32_t w = 0; w < WPT; ++w) {
      uint16_t *wd = &W->wdesc[4 * ((size_t)t * WPT + w)];
      const int32_t wb = cb + w * WC, we = wb + WC < ce ? wb + WC : ce;
      if (wb >= ce) { wd[0] = (uint16_t)(ne - tile_e0); wd[1] = 0; wd[2] = (uint16_t)(nw - tile_w0); wd[3] = 0; continue; }
      /* mini-tile halo */
      int32_t nwh = 0;
      for (int32_t i = wb; i < we; ++i)
        for (int32_t k = 0; k < WD; ++k) {
          const int32_t nbr = L->ell_nbr[(size_t)k * S + i];
          if (nbr >= 0 && (nbr < wb || nbr >= we)) warp_halo[nwh++] = nbr;
        }
      nwh = sort_unique(warp_halo, nwh);
      if (nwh > 222) { ierr = rdyb_fail(RDYB_ERR_SUP, "warp layout: a mini-tile reads %d outside cells (> 222)", nwh); goto fail; }
      WGROW(W->whalo, wcap, nw + nwh);
      for (int32_t j = 0; j < nwh; ++j) {
        const int32_t c = warp_halo[j];
        W->whalo[nw + j] = (uint16_t)((c >= cb && c < ce) ? c - cb : TC + lower_bound_w(tile_halo, nth, c));
      }
      /* edges: interior first (pass 0), boundary last (pass 1) */
      const int64_t e0 = ne;
      int32_t nint = 0, nbnd = 0;
      for (int32_t q = 0; q < WC * RDYB_ELL_WIDTH; ++q) slot_edge[q] = -1;
      for (int pass = 0; pass < 2; ++pass)
        for (int32_t i = wb; i < we; ++i)
          for (int32_t k = 0; k < WD; ++k) {
            const size_t q = (size_t)k * S + i;
            const int32_t nbr = L->ell_nbr[q];
            if (nbr == RDYB_NO_EDGE || (pass == 0) != (nbr >= 0)) continue;
            const int in_warp = nbr >= wb && nbr < we;
            const int me_left = L->ell_slen[q] < 0.0;
            if (nbr >= 0 && in_warp && !me_left) continue; /* listed by the left cell */
            if ((ierr = wedges_reserve(W, &ecap, ne + 1))) goto fail;
            const uint32_t me = (uint32_t)(i - wb);
            uint32_t other = 0xFFu;
            int32_t report_cell = i;
            if (nbr >= 0) {
              other = in_warp ? (uint32_t)(nbr - wb) : (uint32_t)(WC + lower_bound_w(warp_halo, nwh, nbr));
              const int32_t lc = me_left ? i : nbr, rc = me_left ? nbr : i;
              report_cell = L->area[lc] < L->area[rc] ? lc : rc; /* src/swe/swe_petsc.c:289-296 */
            }
            const uint32_t left = me_left ? me : other, right = me_left ? other : me;
            W->edge_lr[ne] = (uint16_t)(left | (right << 8));
            W->edge_bidx[ne] = nbr < 0 ? -(nbr + 1) : -1;
            W->edge_sn[ne] = L->ell_sn[q]; W->edge_cn[ne] = L->ell_cn[q]; W->edge_len[ne] = fabs(L->ell_slen[q]);
            W->edge_id[ne] = L->ell_edge[q]; W->edge_cell[ne] = report_cell;
            slot_edge[(i - wb) * RDYB_ELL_WIDTH + k] = (int32_t)(ne - e0);
            ++ne;
            if (nbr >= 0) ++nint; else ++nbnd;
          }
      if (nint + nbnd > 126 || nint > 255 || nbnd > 255) { ierr = rdyb_fail(RDYB_ERR_SUP, "warp layout: a mini-tile has %d edges (> 126)", nint + nbnd); goto fail; }
      /* per-cell edge slots, in the ELL (reference accumulation) order */
      for (int32_t i = wb; i < we; ++i)
        for (int32_t k = 0; k < WD; ++k) {
          const size_t q = (size_t)k * S + i;
          const int32_t nbr = L->ell_nbr[q];
          if (nbr == RDYB_NO_EDGE) continue;
          int32_t e = slot_edge[(i - wb) * RDYB_ELL_WIDTH + k];
          const int me_left = L->ell_slen[q] < 0.0;
          if (e < 0) { /* in-warp edge listed by the left cell nbr: its slot with the same edge id */
            for (int32_t k2 = 0; k2 < WD && e < 0; ++k2)
              if (L->ell_nbr[(size_t)k2 * S + nbr] == i && L->ell_edge[(size_t)k2 * S + nbr] == L->ell_edge[q]) e = slot_edge[(nbr - wb) * RDYB_ELL_WIDTH + k2];
            if (e < 0) { ierr = rdyb_fail(RDYB_ERR_USER, "warp layout: inconsistent edge lists at cell %d", L->dev_to_local[i]); goto fail; }
          }
          W->cell_edges[q] = (uint8_t)((uint32_t)e | (me_left ? 0u : 0x80u));
        }
      if (e0 - tile_e0 > 65535 || nw - tile_w0 > 65535) { ierr = rdyb_fail(RDYB_ERR_SUP, "warp layout: tile %d is too large for 16-bit offsets", t); goto fail; }
      wd[0] = (uint16_t)(e0 - tile_e0); wd[1] = (uint16_t)(nint | (nbnd << 8)); wd[2] = (uint16_t)(nw - tile_w0); wd[3] = (uint16_t)nwh;
      nw += nwh;
      if (nint + nbnd > W->max_warp_edges) W->max_warp_edges = nint + nbnd;
      if (nwh > W->max_warp_halo) W->max_warp_halo = nwh;
    }
    int32_t *d = &W->tile_desc[8 * (size_t)t];
    d[0] = cb; d[1] = ncell; d[2] = (int32_t)tile_e0; d[3] = (int32_t)(ne - tile_e0); d[4] = (int32_t)tile_h0; d[5] = nth; d[6] = (int32_t)tile_w0; d[7] = (int32_t)(nw - tile_w0);
    if (nth > W->max_halo) W->max_halo = nth;
    if (d[3] > W->max_tile_edges) W->max_tile_edges = d[3];
    if (d[7] > W->max_tile_whalo) W->max_tile_whalo = d[7];
    if (ne > INT32_MAX - 65536 || nh > INT32_MAX - 65536 || nw > INT32_MAX - 65536) { ierr = rdyb_fail(RDYB_ERR_SUP, "warp layout: too many edges for int32 offsets"); goto fail; }
  }
  /* slack so that bulk copies rounded up to 16 bytes stay inside the arrays */
  if ((ierr = wedges_reserve(W, &ecap, ne + 16))) goto fail;
  for (int64_t g = ne; g < ne + 16; ++g) wedge_pad(W, g);
  WGROW(W->halo_cells, hcap, nh + 16);
  WGROW(W->whalo, wcap, nw + 16);
  for (int j = 0; j < 16; ++j) { W->halo_cells[nh + j] = 0; W->whalo[nw + j] = 0; }
  W->total_edges = ne + 16; W->total_halo = nh + 16; W->total_whalo = nw + 16;
  free(tile_halo); free(warp_halo);
  return RDYB_SUCCESS;
fail:
  free(tile_halo); free(warp_halo);
  rdyb_warp_layout_free(W);
  return ierr;
}
