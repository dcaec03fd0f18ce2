/* rdyb200_forcing.h — forcing and boundary DATASETS applied on the device.
 *
 * In the reference every RDyAdvance starts with RDyApplyForcing (src/forcing/rdyforcing.c:688-770): a host loop maps the current
 * rainfall / boundary dataset onto the owned cells (or the edges of a Dirichlet boundary) through a nearest-neighbour index map and
 * hands a full-size host array to RDySetRegionalWaterSource / RDySetFlowDirichletBoundaryValues — on a GPU run that is one host
 * array of num_owned doubles over PCIe per coupling interval.  Here the index map lives on the device: only the (small) dataset
 * itself is uploaded and one gather kernel writes the operator's source / boundary-value arrays in place.
 *
 *   reference                                                          here
 *   RDyForcingGetCurrentData        rdyforcing_dataset.c:32-66         rdyb_forcing_current_data (host, scalar)
 *   RDyForcingCreateUnstructuredDatasetMap   rdyforcing_map.c:78-106   rdyb_dataset_map_create_unstructured (CUDA, brute force)
 *   RDyForcingCreateRasterDatasetMapping     rdyforcing_map.c:111-138  rdyb_dataset_map_create_raster (CUDA, 5x5 window scan)
 *   RDyForcingReadSpatialMap / WriteMap      rdyforcing_map.c:143-195  rdyb_dataset_map_create_from_indices / rdyb_dataset_map_get
 *   RDyForcingSetRasterData + RDySetRegionalWaterSource   :295-315     rdyb_operator_set_source_from_dataset(offset 5, stride 1, mm/h)
 *   RDyForcingSetUnstructuredData (rain)                  :350-374     rdyb_operator_set_source_from_dataset(offset 2, stride 1, 1.0)
 *   RDyForcingSetUnstructuredData (boundary, stride 3)
 *       + RDySetFlowDirichletBoundaryValues                            rdyb_operator_set_boundary_values_from_dataset(offset 2)
 *   RDyForcingSetHomogeneousData / SetConstantRainfall    :282-344     rdyb_operator_fill_source
 *   RDyForcingSetHomogeneousBoundary                      :380-406     rdyb_operator_fill_boundary_values
 * Index maps are bit-exact with the reference's (same scan order, strict '<', unfused arithmetic); values are verbatim doubles.
 */
#ifndef RDYB200_FORCING_H
#define RDYB200_FORCING_H
#include <stdint.h>

#include "rdyb200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct RdybDatasetMap_ *RdybDatasetMap;

/* time_value_pairs = {t0, v0, t1, v1, ...} (ndata pairs, seconds): the bracket [t_i, t_i+1) holding cur_time, its left value or the
 * linear interpolant; past the last time: the last value and cur_idx = ndata - 1 */
int rdyb_forcing_current_data(const double *time_value_pairs, int32_t ndata, double cur_time, int32_t temporally_interpolate, int32_t *cur_idx,
                              double *cur_value);

/* mesh_xc, mesh_yc [n]: centroids of the owned cells (natural owned order) or of the edges of one boundary (that boundary's edge
 * order).  For every element the index of the nearest data point; the first one wins a tie. */
int rdyb_dataset_map_create_unstructured(RdybOperator op, int32_t n, const double *mesh_xc, const double *mesh_yc, int32_t ndata, const double *data_xc,
                                         const double *data_yc, RdybDatasetMap *map);
/* raster of ncols x nrows square pixels, (xlc, ylc) = lower-left corner, row 0 = the northernmost row; index = row * ncols + col.
 * An element farther than (max(ncols, nrows) + 1) * cellsize from every pixel centre maps to pixel 0, as in the reference. */
int rdyb_dataset_map_create_raster(RdybOperator op, int32_t n, const double *mesh_xc, const double *mesh_yc, int32_t ncols, int32_t nrows, double xlc,
                                   double ylc, double cellsize, RdybDatasetMap *map);
/* a map computed elsewhere (the reference's -*_read_map_file options) */
int rdyb_dataset_map_create_from_indices(RdybOperator op, int32_t n, const int32_t *data2mesh_idx, int32_t ndata, RdybDatasetMap *map);
int rdyb_dataset_map_get(RdybDatasetMap map, int32_t n, int32_t *data2mesh_idx);
int rdyb_dataset_map_destroy(RdybDatasetMap *map);

/* external source `component` (0 water, 1 / 2 momenta) of owned cell i <- data[offset + map[i] * stride] * scale.
 * data: host array of ndata_values doubles (the dataset Vec as loaded, header included); the map must cover the owned cells. */
int rdyb_operator_set_source_from_dataset(RdybOperator op, int32_t component, RdybDatasetMap map, const double *data, int32_t ndata_values, int32_t offset,
                                          int32_t stride, double scale);
/* Dirichlet values (h, hu, hv) of edge e of boundary b <- data[offset + 3 * map[e] + {0, 1, 2}]; the map must cover that boundary */
int rdyb_operator_set_boundary_values_from_dataset(RdybOperator op, int32_t b, RdybDatasetMap map, const double *data, int32_t ndata_values, int32_t offset);
/* spatially homogeneous datasets */
int rdyb_operator_fill_source(RdybOperator op, int32_t component, double value);
int rdyb_operator_fill_boundary_values(RdybOperator op, int32_t b, double h, double hu, double hv);
/* Dirichlet values of boundary b as the operator holds them: values[3 * e + c] */
int rdyb_operator_get_boundary_values(RdybOperator op, int32_t b, int32_t num_edges, double *values);


/* ---- time series (src/time_series.c) ----
 * (h, hu, hv) at n observation sites (owned cell ids), values[3 i + c]: WriteObservations' VecScatter, :400-445, without moving the
 * whole solution to the host */
int rdyb_operator_get_state_at(RdybOperator op, int32_t n, const int32_t *owned_cell_ids, double *values);
/* zero the accumulated boundary fluxes once they have been written: ResetAccumulatedBoundaryFluxes, :508-527 */
int rdyb_operator_reset_boundary_fluxes(RdybOperator op);

#ifdef __cplusplus
}
#endif
#endif
