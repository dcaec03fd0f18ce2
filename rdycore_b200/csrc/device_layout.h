/* device_layout.h — host-side (C) construction of the device mesh layout of the B200 SWE path.
 *
 * The reference keeps RDyMesh as SoA arrays in DMPlex local numbering (include/private/rdymeshimpl.h:26-140)
 * and loops over edges, scattering into cells (src/swe/swe_petsc.c:244-310).  The device layout turns that
 * inside out: owned cells are renumbered along a Hilbert curve (interior cells first, frontier cells that touch
 * a ghost last, ghosts after owned grouped by source rank exactly like local_to_owned in src/rdymesh.c:159-177),
 * and every owned cell carries a fixed-width (ELL) list of its edges with the neighbour's device index, the
 * edge normal and a signed length, so a thread gathers its own flux sum without atomics.
 *
 * Two layouts are derived from the same renumbering:
 *   (1) the cell-centric ELL layout (kernel "gather"): every interior edge is evaluated from both sides;
 *   (2) the tiled layout (kernel "tiled"): contiguous runs of cells form tiles; each tile lists the halo cells
 *       it reads and its edges once, so a block stages states in shared memory and evaluates every tile edge once.
 */
#ifndef RDYB_DEVICE_LAYOUT_H
#define RDYB_DEVICE_LAYOUT_H
#include "rdyb200.h"

#ifdef __cplusplus
extern "C" {
#endif

#define RDYB_NO_EDGE INT32_MIN /* empty ELL slot */
#define RDYB_ELL_WIDTH 4       /* cells have 3 or 4 edges, src/rdymesh.c:811 */

typedef struct {
  int32_t n_owned, n_local, n_interior; /* device order: [0,n_interior) interior, [n_interior,n_owned) frontier, then ghosts */
  int32_t stride;                       /* padded leading dimension of every per-cell array (multiple of 32) */
  int32_t width;                        /* ELL width actually used (3 for all-triangle meshes, else 4) */
  int32_t num_view_cells;
  int32_t *dev_to_local;   /* [n_local]  device index -> mesh-view local cell id */
  int32_t *local_to_dev;   /* [num_view_cells] (-1 for view cells that are neither owned nor face-adjacent ghosts) */
  int32_t *owned_to_dev;   /* [n_owned]  owned index (cells.local_to_owned) -> device index */
  int32_t *dev_to_owned;   /* [n_owned]  device index -> owned index */
  /* per device cell */
  double *inv_area;        /* [n_local] */
  double *area;            /* [n_local] */
  double *dzdx, *dzdy;     /* [n_owned] */
  double *zc;              /* [n_local] or NULL */
  /* ELL, slot-major: x[k*stride + i].  nbr >= 0: device index of the neighbour; nbr = -(1+j): boundary edge j of the
   * flattened boundary-edge table; RDYB_NO_EDGE: empty.  slen < 0: this cell is the edge's left cell (normal points out of
   * it, flux scale -L/A); slen > 0: it is the right cell (flux scale +L/A). */
  int32_t *ell_nbr;
  double  *ell_sn, *ell_cn, *ell_slen;
  int32_t *ell_edge;       /* local edge id of the slot (diagnostics) */
  /* second-order (MUSCL) geometry per ELL slot, built by rdyb_layout_build_muscl (NULL otherwise):
   *   gcx, gcy : this cell's least-squares gradient weights for the slot's edge with the side folded in, so that
   *              grad q += (gcx, gcy) * (q_neighbour - q_self)   [= w * (q_right - q_left) with w the reference's weight]
   *              (PrecomputeLSGradCoeffs / ComputeLeastSquaresGradients, src/operator_fluxes_ceed.c:884-1038)
   *   dxs, dys : edge midpoint - this cell's centroid;  dxn, dyn : edge midpoint - the neighbour's centroid
   *              (ReconstructFaceValues, src/operator_fluxes_ceed.c:1162-1172) */
  double *ell_gcx, *ell_gcy, *ell_dxs, *ell_dys, *ell_dxn, *ell_dyn;
  /* flattened boundary-edge table: boundary b owns entries [b_offsets[b], b_offsets[b+1]) in RDyBoundary.edge_ids order */
  int32_t  num_boundaries, num_boundary_edges;
  int32_t *b_offsets;      /* [num_boundaries+1] */
  int32_t *b_type;         /* [num_boundary_edges] RDYB_CONDITION_* */
  int32_t *b_cell;         /* [num_boundary_edges] device index of the adjacent owned cell, -1 if its left cell is not owned */
  /* halo plan in device indices */
  int32_t  num_peers;
  int32_t *peer_ranks;     /* [num_peers] */
  int32_t *send_offsets;   /* [num_peers+1] */
  int32_t *send_cells;     /* device indices of owned cells to pack, per peer */
  int32_t *recv_offsets;   /* [num_peers+1]; peer p's ghosts are device cells [n_owned+recv_offsets[p], n_owned+recv_offsets[p+1]) */
} RdybLayout;

/* tiled layout (built on top of an RdybLayout) */
typedef struct {
  int32_t tile_cells;        /* cells per tile (last tile of a range may be short) */
  int32_t num_tiles;
  int32_t num_interior_tiles; /* tiles [0,num_interior_tiles) cover interior cells only */
  int32_t max_halo, max_edges; /* per-tile maxima (shared-memory sizing) */
  int32_t max_halo_interior, max_edges_interior; /* the same over the interior tiles only (frontier tiles are thin strips with many more outside neighbours) */
  int32_t *tile_cell_begin;  /* [num_tiles+1] device cell ranges */
  int32_t *tile_desc;        /* [8*num_tiles] per tile {cell_begin, ncell, edge_begin, nedge, halo_begin, nhalo, 0, 0}; begins are padded to 4 */
  int32_t *halo_offsets;     /* [num_tiles+1] into halo_cells */
  int32_t *halo_cells;       /* device indices of non-tile cells a tile reads (ghosts included) */
  int32_t *edge_offsets;     /* [num_tiles+1] into the per-tile-edge arrays */
  /* per tile edge: endpoints as tile-local indices (0..ncell-1 tile cells, ncell.. halo cells); right = 0xFFFF-coded boundary
   * uses edge_bidx */
  uint32_t *edge_lr;         /* left | right << 16; right == 0xFFFF => boundary edge, see edge_bidx */
  int32_t  *edge_bidx;       /* flattened boundary-edge index or -1 */
  int32_t  *edge_id, *edge_cell; /* host only: local edge id and the device cell the Courant diagnostic reports (al < ar ? l : r) */
  double   *edge_sn, *edge_cn, *edge_len;
  /* per cell: up to `width` tile-edge slots in the reference's accumulation order; value = tile-local edge index,
   * bit 15 set => the cell is the edge's right cell; 0xFFFF => empty */
  uint16_t *cell_edges;      /* [width*stride], slot-major like the ELL arrays */
  int64_t   total_edges, total_halo;
  /* second order only (rdyb_tiled_layout_build_muscl, NULL otherwise): [4][edge_off_plane] face offsets per tile edge,
   * edge midpoint - centroid of the left cell (dxl, dyl) and of the right cell (dxr, dyr); edge_off_plane = total_edges rounded up
   * to a multiple of 4, so that a tile's slice of every plane starts 16-byte aligned (bulk copies) */
  double   *edge_off;
  int64_t   edge_off_plane;
  /* second order, two-ring staging (K4M: the gradients of a tile's cells AND of its halo cells are computed inside the tile kernel):
   * a tile's SECOND ring = the interior neighbours of its halo cells that are neither tile nor halo cells (sorted, begins padded to 4;
   * tile_desc[6], [7] = begin, count), and per halo entry the staged position of each of its `width` neighbours in ELL slot order:
   *   0 .. tile_cells-1 tile cell, tile_cells + k the k-th halo cell, 0x8000 | k the k-th second-ring cell, 0xFFFF no interior neighbour.
   * two_ring = 0 when a halo cell is a ghost of another rank (its own neighbours are not all here): such layouts exchange gradients. */
  int32_t  *halo2_cells;
  uint16_t *halo_nbr;        /* [total_halo][4] */
  int32_t   max_halo2, two_ring;
  int64_t   total_halo2;
} RdybTiledLayout;

int  rdyb_layout_build(const RdybMeshView *mesh, int32_t num_boundaries, const RdybBoundaryView *boundaries, const RdybHaloView *halo,
                       int reorder, RdybLayout *out);
int  rdyb_layout_build_muscl(const RdybMeshView *mesh, RdybLayout *layout);
void rdyb_layout_free(RdybLayout *layout);
int  rdyb_tiled_layout_build(const RdybLayout *layout, int32_t tile_cells, RdybTiledLayout *out);
void rdyb_tiled_layout_free(RdybTiledLayout *tiled);
int  rdyb_tiled_layout_build_muscl(const RdybLayout *layout, RdybTiledLayout *tiled);
int  rdyb_tiled_layout_build_two_ring(const RdybLayout *layout, RdybTiledLayout *tiled); /* called by rdyb_tiled_layout_build_muscl */

#ifdef __cplusplus
}
#endif
#endif
