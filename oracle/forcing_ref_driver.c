/* forcing_ref_driver.c — drives the reference's OWN forcing-data code ("reference" build of forcing_oracle.h).
 * TEST INFRASTRUCTURE ONLY.  Linked with /root/reference/src/forcing/rdyforcing_dataset.c and rdyforcing_map.c, compiled where
 * they lie against oracle/ref_shim/.  This file only fills the reference's dataset structs from plain arrays, calls the
 * reference function and copies the result out; the PETSc / RDy entry points those two files merely reference (file viewers,
 * centroid getters) are stubbed in forcing_ref_stubs.c — none of them is on the paths used here. */
#include <private/rdyforcingimpl.h>

#include "forcing_oracle.h"

const char *forcing_oracle_kind(void) { return "reference"; }

int forcing_oracle_current_data(const double *pairs, int32_t ndata, double t, int32_t interpolate, int32_t *cur_idx, double *cur) {
  PetscInt idx = *cur_idx;
  PetscErrorCode e = RDyForcingGetCurrentData((PetscScalar *)pairs, ndata, t, interpolate ? PETSC_TRUE : PETSC_FALSE, &idx, cur);
  *cur_idx = idx;
  return e;
}

int forcing_oracle_map_unstructured(int32_t n, const double *mx, const double *my, int32_t nd, const double *dx, const double *dy, int32_t *out) {
  RDyUnstructuredDataset d;
  memset(&d, 0, sizeof(d));
  d.mesh_nelements = n; d.ndata = nd;
  d.mesh_xc = (PetscReal *)mx; d.mesh_yc = (PetscReal *)my; d.data_xc = (PetscReal *)dx; d.data_yc = (PetscReal *)dy;
  PetscErrorCode e = RDyForcingCreateUnstructuredDatasetMap(&d);
  if (e) return e;
  for (int32_t i = 0; i < n; ++i) out[i] = d.data2mesh_idx[i];
  free(d.data2mesh_idx);
  return 0;
}

int forcing_oracle_map_raster(int32_t n, const double *mx, const double *my, int32_t ncols, int32_t nrows, double xlc, double ylc, double cs, int32_t *out) {
  RDyRasterDataset d;
  memset(&d, 0, sizeof(d));
  d.ncols = ncols; d.nrows = nrows; d.xlc = xlc; d.ylc = ylc; d.cellsize = cs;
  d.mesh_ncells_local = n; d.mesh_xc = (PetscReal *)mx; d.mesh_yc = (PetscReal *)my;
  /* the pixel centres as DoPostprocessForSourceRasterDataset lays them out (static in src/forcing/rdyforcing.c:253-260, which needs
   * the whole RDy object and is not part of this build): */
  d.data_xc = calloc((size_t)ncols * nrows, sizeof(PetscReal)); d.data_yc = calloc((size_t)ncols * nrows, sizeof(PetscReal));
  PetscInt k = 0;
  for (PetscInt r = 0; r < nrows; ++r)
    for (PetscInt c = 0; c < ncols; ++c, ++k) {
      d.data_xc[k] = xlc + c * cs + cs / 2.0;
      d.data_yc[k] = ylc + (nrows - 1 - r) * cs + cs / 2.0;
    }
  d.data2mesh_idx = calloc((size_t)(n > 0 ? n : 1), sizeof(PetscInt)); /* PetscCalloc1 in the same function, :262 */
  PetscErrorCode e = RDyForcingCreateRasterDatasetMapping(NULL, &d);
  for (int32_t i = 0; i < n; ++i) out[i] = d.data2mesh_idx[i];
  free(d.data_xc); free(d.data_yc); free(d.data2mesh_idx);
  return e;
}

int forcing_oracle_set_raster(const double *vec, int32_t nvec, int32_t off, int32_t n, const int32_t *idx, double *rain) {
  struct _shim_Vec v = {(PetscScalar *)vec, nvec, 1, NULL};
  RDyRasterDataset d;
  memset(&d, 0, sizeof(d));
  d.data_vec = &v; d.header_offset = off; d.data2mesh_idx = (PetscInt *)idx;
  d.ndata_file = 1 << 20; d.dtime_in_hour = 1.0; /* the hour is never over: no next file is opened */
  return RDyForcingSetRasterData(&d, 0.0, n, rain);
}

int forcing_oracle_set_unstructured(const double *vec, int32_t nvec, int32_t stride, int32_t n, const int32_t *idx, double *values) {
  struct _shim_Vec v = {(PetscScalar *)vec, nvec, 1, NULL};
  RDyUnstructuredDataset d;
  memset(&d, 0, sizeof(d));
  d.data_vec = &v; d.stride = stride; d.mesh_nelements = n; d.data2mesh_idx = (PetscInt *)idx;
  d.ndata_file = 1 << 20; d.dtime_in_hour = 1.0;
  return RDyForcingSetUnstructuredData(&d, 0.0, values);
}

int forcing_oracle_homogeneous(const double *pairs, int32_t ndata, double t, int32_t interpolate, int32_t boundary, int32_t n, double *values, int32_t *cur_idx,
                               int32_t *prev_idx) {
  struct _shim_Vec v = {(PetscScalar *)pairs, 2 * ndata, 1, NULL};
  RDyHomogeneousDataset d;
  memset(&d, 0, sizeof(d));
  d.data_vec = &v; d.ndata = ndata; d.temporally_interpolate = interpolate ? PETSC_TRUE : PETSC_FALSE; d.cur_idx = *cur_idx; d.prev_idx = *prev_idx;
  PetscErrorCode e = boundary ? RDyForcingSetHomogeneousBoundary(&d, t, n, values) : RDyForcingSetHomogeneousData(&d, t, n, values);
  *cur_idx = d.cur_idx; *prev_idx = d.prev_idx;
  return e;
}
