/* rdyb200_mesh.h — host-side finite-volume mesh for the B200 SWE path.
 *
 * Produces exactly the RDyMesh quantities the kernels read, from vertices and
 * cell->vertex lists, with the reference's formulas:
 *   edge orientation, sn/cn           src/rdymesh.c:628-688
 *   internal / boundary edge lists    src/rdymesh.c:692-710
 *   triangle / quad bed slopes        src/rdymesh.c:747-784, 799-862
 *   2-D projection for HR             src/rdymesh.c:1478-1509
 *   owned-first local_to_owned map    src/rdymesh.c:159-177
 *   HR cell bed elevation zc          src/swe/swe_petsc.c:1208-1224
 * In RDycore these come out of PETSc DMPlex (src/rdydm.c:82-188); here DMPlex is
 * replaced by a plain C builder, a coordinate-bisection partitioner with one
 * layer of face-neighbour ghosts (DMPlexDistributeOverlap(dm,1), src/rdydm.c:145-157)
 * and generators for the synthetic benchmark meshes.
 */
#ifndef RDYB200_MESH_H
#define RDYB200_MESH_H

#include "rdyb200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct RdybHostMesh_ *RdybHostMesh;

/* Build from vertices xyz[3*nv] and cells given as CSR (cell_offsets[nc+1], cell_verts);
 * cells have 3 or 4 vertices (src/rdymesh.c:811).  All cells are owned (serial mesh). */
int rdyb_hostmesh_create(int32_t num_vertices, const double *xyz, int32_t num_cells, const int32_t *cell_offsets, const int32_t *cell_verts,
                         RdybHostMesh *mesh);
int rdyb_hostmesh_destroy(RdybHostMesh *mesh);

/* RDyMeshOverride2DProjection (HR): 2-D edge lengths and shoelace areas */
int rdyb_hostmesh_override_2d_projection(RdybHostMesh mesh);
/* HR bed elevation: mean of vertex z, or the given per-cell values (cell_elevation file case) */
int rdyb_hostmesh_set_zc(RdybHostMesh mesh, const double *zc_or_null);

/* borrowed view for rdyb_operator_create(); valid until the mesh is destroyed */
int rdyb_hostmesh_view(RdybHostMesh mesh, RdybMeshView *view);
/* halo plan of a local (partitioned) mesh; num_peers = 0 for a serial mesh */
int rdyb_hostmesh_halo(RdybHostMesh mesh, RdybHaloView *halo);

/* sizes: out[0]=num_vertices, [1]=num_cells, [2]=num_owned_cells, [3]=num_edges, [4]=num_internal_edges, [5]=num_boundary_edges */
int rdyb_hostmesh_sizes(RdybHostMesh mesh, int32_t out[6]);
/* borrowed array pointers; names: "boundary_edge_ids" (int32), "edge_xc","edge_yc","edge_zc","cell_xc","cell_yc","cell_zcen" (double),
 * "edge_vertex_ids" (int32 [2e]), "owned_to_local" (int32), "cell_global_ids" (int64) */
int rdyb_hostmesh_array(RdybHostMesh mesh, const char *name, const void **ptr, int64_t *length);

/* ---- partitioning (replaces DMPlex distribute + overlap 1) ---- */
/* recursive coordinate bisection on cell centroids into nparts balanced parts; part[num_cells] out */
int rdyb_hostmesh_partition_rcb(RdybHostMesh mesh, int32_t nparts, int32_t *part);
/* local mesh of `rank`: owned cells (ascending global id) then face-neighbour ghosts grouped by owner rank
 * (ascending global id inside each group), all edges that touch an owned cell, and the halo plan */
int rdyb_hostmesh_extract_local(RdybHostMesh global_mesh, const int32_t *part, int32_t nparts, int32_t rank, RdybHostMesh *local_mesh);

/* ---- generators ---- */
/* nx x ny quads on [0,lx]x[0,ly], z = 0 (DMPlexCreateBoxMesh-like); cell (i,j) -> j*nx+i */
int rdyb_hostmesh_generate_quads(int32_t nx, int32_t ny, double lx, double ly, RdybHostMesh *mesh);
/* Synthetic floodplain of SURVEY §8(d) config C/D/E: nx x ny quads on an lx x ly rectangle, each split into 2 triangles
 * along alternating diagonals, interior vertices jittered by U(-jitter,jitter)*dx (xorshift64*, seed), bathymetry
 * z = 20*(x/lx) + 5*sin(2*pi*x/3000)*cos(2*pi*y/2000)*zamp + 2*U(0,1).
 * Only the block of quads [i0,i1)x[j0,j1) is owned; with overlap != 0 the face-neighbour triangles across the block
 * border are added as ghosts and the halo plan is filled for a px x py block decomposition (rank = bj*px+bi). */
int rdyb_hostmesh_generate_floodplain(int32_t nx, int32_t ny, double lx, double ly, double jitter, double zamp, uint64_t seed, int32_t px,
                                      int32_t py, int32_t rank, RdybHostMesh *mesh);
/* uniform 1->4 refinement of a triangle mesh (DMPlex -dm_refine regular refinement) */
int rdyb_hostmesh_refine_triangles(RdybHostMesh coarse, RdybHostMesh *fine);
/* overwrite vertex z (MMS SnapVerticesToBathymetry, src/rdymms.c:108-134) and recompute geometry */
int rdyb_hostmesh_set_vertex_z(RdybHostMesh mesh, const double *z);
int rdyb_hostmesh_vertices(RdybHostMesh mesh, const double **xyz, int32_t *num_vertices);

#ifdef __cplusplus
}
#endif
#endif
