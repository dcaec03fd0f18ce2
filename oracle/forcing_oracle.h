/* forcing_oracle.h — CPU oracle of the forcing-data step either side of the SWE path.  TEST INFRASTRUCTURE ONLY:
 * may be used by tests/, never by the product package.
 *
 * Two builds export this same API:
 *   oracle/_ref/libforcing_oracle_ref.so  ("reference"): the reference's OWN src/forcing/rdyforcing_dataset.c and
 *       src/forcing/rdyforcing_map.c compiled where they lie against oracle/ref_shim/, driven by forcing_ref_driver.c;
 *   oracle/libforcing_oracle_port.so      ("port"): forcing_oracle.c, a plain-C restatement (travels to the GPU box).
 * Pin: the two are compared bit for bit on the reference's own forcing fixtures (tests/test_forcing.py). */
#ifndef FORCING_ORACLE_H
#define FORCING_ORACLE_H
#include <stdint.h>

const char *forcing_oracle_kind(void);
/* RDyForcingGetCurrentData, src/forcing/rdyforcing_dataset.c:32-66: pairs = {t0, v0, t1, v1, ...} */
int forcing_oracle_current_data(const double *pairs, int32_t ndata, double cur_time, int32_t interpolate, int32_t *cur_idx, double *cur_value);
/* RDyForcingCreateUnstructuredDatasetMap, src/forcing/rdyforcing_map.c:78-106 */
int forcing_oracle_map_unstructured(int32_t n, const double *mesh_xc, const double *mesh_yc, int32_t ndata, const double *data_xc, const double *data_yc,
                                    int32_t *data2mesh_idx);
/* raster pixel centres of DoPostprocessForSourceRasterDataset (src/forcing/rdyforcing.c:253-260) + RDyForcingCreateRasterDatasetMapping
 * (src/forcing/rdyforcing_map.c:111-138) */
int forcing_oracle_map_raster(int32_t n, const double *mesh_xc, const double *mesh_yc, int32_t ncols, int32_t nrows, double xlc, double ylc, double cellsize,
                              int32_t *data2mesh_idx);
/* RDyForcingSetRasterData, src/forcing/rdyforcing_dataset.c:295-315 (mm/h -> m/s); vec = the whole data Vec, header included */
int forcing_oracle_set_raster(const double *vec, int32_t nvec, int32_t header_offset, int32_t n, const int32_t *data2mesh_idx, double *rain);
/* RDyForcingSetUnstructuredData, src/forcing/rdyforcing_dataset.c:350-374; vec = {ndata, stride, values...} */
int forcing_oracle_set_unstructured(const double *vec, int32_t nvec, int32_t stride, int32_t n, const int32_t *data2mesh_idx, double *values);
/* RDyForcingSetHomogeneousData :321-344 (boundary = 0: values[n]) and RDyForcingSetHomogeneousBoundary :380-406 (boundary = 1:
 * values[3n] = {v, 0, 0}); cur_idx / prev_idx carry the dataset's state between calls */
int forcing_oracle_homogeneous(const double *pairs, int32_t ndata, double cur_time, int32_t interpolate, int32_t boundary, int32_t n, double *values,
                               int32_t *cur_idx, int32_t *prev_idx);
#endif
