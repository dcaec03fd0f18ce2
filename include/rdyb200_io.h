/* rdyb200_io.h — readers for the on-disk formats either side of the SWE path.
 *
 * Host-side C, no PETSc / ExodusII / NetCDF library: the reference reads these files through DMPlexCreateFromFile (Exodus II,
 * src/rdydm.c:96-121) and VecLoad of PETSc binary viewers (initial conditions src/rdysetup.c:804-836, Manning n and other
 * material properties :881-907, cell elevations for hydrostatic reconstruction src/swe/swe_petsc.c:1208-1224).  Here the two
 * formats are parsed directly:
 *   - Exodus II stored as NetCDF-3 classic (CDF-1) or 64-bit offset (CDF-2): nodes, element blocks (QUAD / SHELL4 / TRI / TRI3),
 *     side sets;
 *   - PETSc binary Vec, 32- or 64-bit index builds (the *.int32.bin / *.int64.bin fixtures of share/).
 * Every function returns RDYB_SUCCESS or an error code with the message in rdyb_last_error().
 */
#ifndef RDYB200_IO_H
#define RDYB200_IO_H
#include <stdint.h>

#include "rdyb200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* An Exodus II surface mesh.  Elements are numbered across blocks in file order, as DMPlex numbers the cells it creates from the
 * file (and as the reference's natural-order IC / material files are indexed in a serial run). */
typedef struct {
  int32_t num_dim;          /* coordinates stored in the file (2 or 3) */
  int32_t num_nodes, num_elems, num_blocks, num_side_sets;
  double *x, *y, *z;        /* [num_nodes]; z is all zero for a 2-D file */
  int32_t *elem_offsets;    /* [num_elems + 1] into elem_nodes */
  int32_t *elem_nodes;      /* 0-based node ids, counter-clockwise as stored */
  int32_t *elem_block;      /* [num_elems] id of the element's block (eb_prop1: the deck's grid_region_id) */
  int32_t *block_ids;       /* [num_blocks] */
  int32_t *side_set_ids;    /* [num_side_sets] (ss_prop1: the deck's grid_boundary_id) */
  int32_t *side_set_offsets; /* [num_side_sets + 1] into the two arrays below */
  int32_t *side_elems;      /* 0-based element of each side */
  int32_t *side_edges;      /* local edge e of that element: between its nodes e and (e + 1) % n.  Exodus numbers the sides of
                               2-D QUAD / TRI elements 1..n and those of shell elements (SHELL4, TRI3 in a 3-D file) 3..n+2,
                               sides 1 and 2 being the two faces; a face side is reported as -1 */
} RdybExodusMesh;

/* replaces DMPlexCreateFromFile for *.exo (src/rdydm.c:96-121) */
int rdyb_exodus_read(const char *path, RdybExodusMesh *mesh);
void rdyb_exodus_free(RdybExodusMesh *mesh);

/* replaces PetscViewerBinaryOpen + VecLoad (src/rdysetup.c:804-836, 881-907): *values is malloc'ed, the caller frees it with
 * rdyb_free().  index_bytes (may be NULL) receives 4 or 8, the PetscInt size of the build that wrote the file. */
int rdyb_petsc_vec_read(const char *path, int64_t *n, double **values, int32_t *index_bytes);
/* VecView to a binary viewer (checkpoint-style output of a natural-order Vec); index_bytes = 4 or 8 */
int rdyb_petsc_vec_write(const char *path, int64_t n, const double *values, int32_t index_bytes);
void rdyb_free(void *p);

/* ---- checkpoints (binary format): WriteCheckpointFile / ReadCheckpointFile, src/checkpoint.c:153-212, 258-320 ----
 * A checkpoint is a PETSc binary stream of two records: the PetscBag "metadata" {nproc, t, dt, step} (CreateMetadata, :57-76) and
 * the solution Vec in NATURAL cell order, block size = number of components (DMPlexGlobalToNatural + VecView, :190-205), next to a
 * "<file>.info" holding "-vecload_block_size <ndof>" as the binary viewer leaves it.  No PETSc here: the bag record follows
 * PetscBagView's binary layout (PETSc 3.24 src/sys/classes/bag/bag.c: classid 1211219, a deprecated size word, the item count,
 * name[64], help[128], then per item {offset, PetscDataType, name[64], help[128], msize, value}).  The reference tree holds no
 * checkpoint file to pin the bag bytes against ("parity unpinned" for the bag; the Vec record is the format of the share/
 * fixtures, which is pinned) and the item offsets are build-specific in PETSc (PetscBagLoad does not use them). */
typedef struct {
  int32_t nproc; /* number of MPI tasks that wrote the file */
  double t, dt;  /* simulation time and time step, in seconds */
  int32_t step;
} RdybCheckpointMetadata;

/* "<directory>/<prefix>.rdycore.r.<index, zero-padded to the digits of max_index>" (GenerateE3SMCheckpointFilenameBase, :141-152);
 * the binary format appends ".bin" */
int rdyb_checkpoint_filename_base(const char *directory, const char *prefix, int64_t index, int64_t max_index, char *out, int32_t capacity);
/* Each caller writes the rows natural_ids[i] (NULL: rows 0..n_local-1) of the natural-order solution: values[ndof * i + c].  One
 * process per GPU calls it with its owned cells and their global ids: the rows land at their natural offsets with pwrite, so no
 * gather to rank 0 is needed; the call with write_header != 0 (rank 0) also writes the two record headers and the .info file.
 * index_bytes = 4 or 8 (the PetscInt of the build that is to read the file). */
int rdyb_checkpoint_write(const char *path, const RdybCheckpointMetadata *metadata, int64_t num_global_cells, int32_t ndof, int32_t index_bytes,
                          int64_t n_local, const int64_t *natural_ids, const double *values, int32_t write_header);
/* *values (malloc'ed, rdyb_free) receives the natural-order solution, *n_values its length (cells * ndof) */
int rdyb_checkpoint_read(const char *path, RdybCheckpointMetadata *metadata, int64_t *n_values, double **values, int32_t *index_bytes);

#ifdef __cplusplus
}
#endif
#endif
