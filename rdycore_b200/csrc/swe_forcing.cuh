/* swe_forcing.cuh — forcing / boundary datasets applied on the device (include/rdyb200_forcing.h).
 * Included at the end of swe_device.cu (needs the operator's internals: source and boundary-value arrays, cell permutation,
 * stream).  All kernels are gathers or short scans: HBM / latency bound, a few hundred microseconds at 10 M cells. */
#ifndef RDYB_SWE_FORCING_CUH
#define RDYB_SWE_FORCING_CUH
#include "rdyb200_forcing.h"

struct RdybDatasetMap_ {
  int device = 0;
  int32_t n = 0, ndata = 0;
  int32_t *idx = nullptr; /* device, natural element order */
};

/* RDyForcingCreateUnstructuredDatasetMap, src/forcing/rdyforcing_map.c:78-106: thread per mesh element, data points staged through
 * shared memory in the reference's scan order; dist = sqrt(dx*dx + dy*dy) with separately rounded operations (no FMA), strict '<' */
#define RDYB_MAP_TILE 512
__global__ void dataset_map_unstructured_kernel(int32_t n, const double *__restrict__ mx, const double *__restrict__ my, int32_t ndata,
                                                const double *__restrict__ dxc, const double *__restrict__ dyc, int32_t *__restrict__ out) {
  __shared__ double sx[RDYB_MAP_TILE], sy[RDYB_MAP_TILE];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const double x = i < n ? mx[i] : 0.0, y = i < n ? my[i] : 0.0;
  double best = 0.0;
  int32_t arg = 0;
  for (int32_t k0 = 0; k0 < ndata; k0 += RDYB_MAP_TILE) {
    const int32_t m = min(RDYB_MAP_TILE, ndata - k0);
    __syncthreads();
    for (int t = threadIdx.x; t < m; t += blockDim.x) { sx[t] = dxc[k0 + t]; sy[t] = dyc[k0 + t]; }
    __syncthreads();
    for (int32_t t = 0; t < m; ++t) {
      const double dx = __dsub_rn(x, sx[t]), dy = __dsub_rn(y, sy[t]);
      const double dist = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
      if (k0 + t == 0 || dist < best) { best = dist; arg = k0 + t; }
    }
  }
  if (i < n) out[i] = arg;
}

/* RDyForcingCreateRasterDatasetMapping, src/forcing/rdyforcing_map.c:111-138, with the pixel centres of
 * DoPostprocessForSourceRasterDataset (src/forcing/rdyforcing.c:253-260).  The reference scans all pixels; the nearest centre (and
 * anything that could tie with it after rounding) lies within two pixels of the pixel containing the clamped point, so the same
 * comparison sequence over that 5x5 window — in the same row-major order — selects the same index. */
__global__ void dataset_map_raster_kernel(int32_t n, const double *__restrict__ mx, const double *__restrict__ my, int32_t ncols, int32_t nrows, double xlc,
                                          double ylc, double cs, int32_t *__restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = mx[i], y = my[i];
  double best = __dmul_rn((double)(max(ncols, nrows) + 1), cs);
  int32_t arg = 0;
  const double half = cs / 2.0;
  int cj = (int)floor((x - xlc) / cs), rj = (int)floor((y - ylc) / cs); /* column, and row counted from the south */
  cj = min(max(cj, 0), ncols - 1);
  rj = min(max(rj, 0), nrows - 1);
  const int r_lo = max(nrows - 1 - rj - 2, 0), r_hi = min(nrows - 1 - rj + 2, nrows - 1);
  const int c_lo = max(cj - 2, 0), c_hi = min(cj + 2, ncols - 1);
  for (int r = r_lo; r <= r_hi; ++r) {
    const double py = __dadd_rn(__dadd_rn(ylc, __dmul_rn((double)(nrows - 1 - r), cs)), half);
    const double dy = __dsub_rn(y, py);
    for (int c = c_lo; c <= c_hi; ++c) {
      const double px = __dadd_rn(__dadd_rn(xlc, __dmul_rn((double)c, cs)), half);
      const double dx = __dsub_rn(x, px);
      const double dist = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
      if (dist < best) { best = dist; arg = r * ncols + c; }
    }
  }
  out[i] = arg;
}

/* src[d] = data[offset + map[dev_to_owned[d]] * stride] * scale over the owned device cells */
__global__ void dataset_gather_source_kernel(int32_t n, const int32_t *__restrict__ dev_to_owned, const int32_t *__restrict__ map,
                                             const double *__restrict__ data, int32_t offset, int32_t stride, double scale, double *__restrict__ dst) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d < n) dst[d] = __dmul_rn(data[offset + (int64_t)map[dev_to_owned[d]] * stride], scale);
}
/* bval[c][e] = data[offset + 3 map[e] + c] */
__global__ void dataset_gather_boundary_kernel(int32_t ne, const int32_t *__restrict__ map, const double *__restrict__ data, int32_t offset,
                                               double *__restrict__ bval, int64_t comp_stride) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= ne) return;
  const int64_t k = offset + 3 * (int64_t)map[e];
  bval[e] = data[k]; bval[comp_stride + e] = data[k + 1]; bval[2 * comp_stride + e] = data[k + 2];
}
/* out[3 i + c] = state[c][sites[i]] : observation sites */
__global__ void sample_state_kernel(int32_t n, const int32_t *__restrict__ sites, const double *__restrict__ state, int64_t stride, double *__restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t d = sites[i];
  out[3 * (int64_t)i] = state[d]; out[3 * (int64_t)i + 1] = state[stride + d]; out[3 * (int64_t)i + 2] = state[2 * stride + d];
}
__global__ void fill_kernel(int64_t n, double v, double *__restrict__ dst) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = v;
}

/* ---------------- C-ABI ---------------- */
extern "C" int rdyb_forcing_current_data(const double *p, int32_t ndata, double t, int32_t interpolate, int32_t *cur_idx, double *cur) {
  RDYB_CHECK(p && cur_idx && cur && ndata > 0, RDYB_ERR_ARG, "bad argument");
  /* src/forcing/rdyforcing_dataset.c:32-66 */
  for (int32_t i = 0; i < ndata - 1; ++i) {
    const double t_dn = p[2 * i], d_dn = p[2 * i + 1], t_up = p[2 * i + 2], d_up = p[2 * i + 3];
    if (t >= t_dn && t < t_up) {
      *cur_idx = i;
      *cur = interpolate ? (t - t_dn) / (t_up - t_dn) * (d_up - d_dn) + d_dn : d_dn;
      return RDYB_SUCCESS;
    }
  }
  *cur_idx = ndata - 1;
  *cur = p[2 * ndata - 1];
  return RDYB_SUCCESS;
}

static int dataset_map_new(RdybOperator op, int32_t n, int32_t ndata, RdybDatasetMap *map) {
  RDYB_CHECK(op && map && n >= 0, RDYB_ERR_ARG, "bad argument");
  CU(cudaSetDevice(op->device));
  RdybDatasetMap m = new (std::nothrow) RdybDatasetMap_();
  RDYB_CHECK(m, RDYB_ERR_MEM, "out of memory");
  m->device = op->device; m->n = n; m->ndata = ndata;
  const int rc = dalloc(&m->idx, (size_t)(n > 0 ? n : 1));
  if (rc) { delete m; return rc; }
  *map = m;
  return RDYB_SUCCESS;
}

/* two host coordinate arrays -> device scratch */
static int upload_xy(const double *x, const double *y, int32_t n, double **dx, double **dy) {
  *dx = *dy = nullptr;
  RDYB_CALL(upload(dx, x, (size_t)n));
  const int rc = upload(dy, y, (size_t)n);
  if (rc) { cudaFree(*dx); *dx = nullptr; }
  return rc;
}

extern "C" int rdyb_dataset_map_create_unstructured(RdybOperator op, int32_t n, const double *mesh_xc, const double *mesh_yc, int32_t ndata,
                                                    const double *data_xc, const double *data_yc, RdybDatasetMap *map) {
  RDYB_CHECK(op && map && (n == 0 || (mesh_xc && mesh_yc)) && data_xc && data_yc, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(ndata > 0, RDYB_ERR_USER, "The unstructured dataset has no data points");
  RDYB_CALL(dataset_map_new(op, n, ndata, map));
  if (n == 0) return RDYB_SUCCESS;
  double *mx, *my, *px, *py;
  RDYB_CALL(upload_xy(mesh_xc, mesh_yc, n, &mx, &my));
  int rc = upload_xy(data_xc, data_yc, ndata, &px, &py);
  if (!rc) {
    dataset_map_unstructured_kernel<<<grid_for(n, 256), 256, 0, op->stream>>>(n, mx, my, ndata, px, py, (*map)->idx);
    LAUNCHED();
    cudaError_t e = cudaStreamSynchronize(op->stream);
    if (e != cudaSuccess) rc = rdyb_fail(RDYB_ERR_CUDA, "CUDA error %s in the dataset map kernel", cudaGetErrorName(e));
    cudaFree(px); cudaFree(py);
  }
  cudaFree(mx); cudaFree(my);
  if (rc) rdyb_dataset_map_destroy(map);
  return rc;
}

extern "C" int rdyb_dataset_map_create_raster(RdybOperator op, int32_t n, const double *mesh_xc, const double *mesh_yc, int32_t ncols, int32_t nrows,
                                              double xlc, double ylc, double cellsize, RdybDatasetMap *map) {
  RDYB_CHECK(op && map && (n == 0 || (mesh_xc && mesh_yc)), RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(ncols > 0 && nrows > 0 && cellsize > 0.0, RDYB_ERR_USER, "Invalid raster: %d columns x %d rows, cell size %g", ncols, nrows, cellsize);
  RDYB_CALL(dataset_map_new(op, n, ncols * nrows, map));
  if (n == 0) return RDYB_SUCCESS;
  double *mx, *my;
  int rc = upload_xy(mesh_xc, mesh_yc, n, &mx, &my);
  if (!rc) {
    dataset_map_raster_kernel<<<grid_for(n, 256), 256, 0, op->stream>>>(n, mx, my, ncols, nrows, xlc, ylc, cellsize, (*map)->idx);
    LAUNCHED();
    cudaError_t e = cudaStreamSynchronize(op->stream);
    if (e != cudaSuccess) rc = rdyb_fail(RDYB_ERR_CUDA, "CUDA error %s in the raster map kernel", cudaGetErrorName(e));
    cudaFree(mx); cudaFree(my);
  }
  if (rc) rdyb_dataset_map_destroy(map);
  return rc;
}

extern "C" int rdyb_dataset_map_create_from_indices(RdybOperator op, int32_t n, const int32_t *idx, int32_t ndata, RdybDatasetMap *map) {
  RDYB_CHECK(op && map && (n == 0 || idx), RDYB_ERR_ARG, "NULL argument");
  for (int32_t i = 0; i < n; ++i) RDYB_CHECK(idx[i] >= 0 && idx[i] < ndata, RDYB_ERR_USER, "Map entry %d = %d is outside the dataset (%d points)", i, idx[i], ndata);
  RDYB_CALL(dataset_map_new(op, n, ndata, map));
  if (n) CU(cudaMemcpy((*map)->idx, idx, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice));
  return RDYB_SUCCESS;
}

extern "C" int rdyb_dataset_map_get(RdybDatasetMap map, int32_t n, int32_t *idx) {
  RDYB_CHECK(map && (n == 0 || idx), RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(n == map->n, RDYB_ERR_USER, "Wrong size (%d) for the dataset map (should be %d)", n, map->n);
  CU(cudaSetDevice(map->device));
  if (n) CU(cudaMemcpy(idx, map->idx, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost));
  return RDYB_SUCCESS;
}

extern "C" int rdyb_dataset_map_destroy(RdybDatasetMap *map) {
  if (!map || !*map) return RDYB_SUCCESS;
  cudaSetDevice((*map)->device);
  cudaFree((*map)->idx);
  delete *map;
  *map = nullptr;
  return RDYB_SUCCESS;
}

/* the dataset goes up through the operator's staging buffer when it fits, else through a temporary */
static int dataset_upload(RdybOperator op, const double *data, int32_t nvals, double **dev, bool *temp) {
  *temp = false;
  *dev = nullptr;
  if ((size_t)nvals <= 3 * (size_t)op->N && op->stage) *dev = op->stage;
  else { RDYB_CALL(dalloc(dev, (size_t)nvals)); *temp = true; }
  cudaError_t e = cudaMemcpyAsync(*dev, data, sizeof(double) * (size_t)nvals, cudaMemcpyHostToDevice, op->stream);
  if (e != cudaSuccess) { if (*temp) cudaFree(*dev); return rdyb_fail(RDYB_ERR_CUDA, "CUDA error %s while uploading a dataset", cudaGetErrorName(e)); }
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_set_source_from_dataset(RdybOperator op, int32_t component, RdybDatasetMap map, const double *data, int32_t nvals,
                                                     int32_t offset, int32_t stride, double scale) {
  RDYB_CHECK(op && map && data, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(component >= 0 && component < 3, RDYB_ERR_USER, "Invalid source component: %d", component);
  RDYB_CHECK(map->n == op->N, RDYB_ERR_USER, "The dataset map covers %d elements, the operator owns %d cells", map->n, op->N);
  RDYB_CHECK(offset >= 0 && stride >= 1 && (int64_t)offset + (int64_t)(map->ndata - 1) * stride < nvals, RDYB_ERR_USER,
             "The dataset (%d values) is too short for %d data points of stride %d after a header of %d", nvals, map->ndata, stride, offset);
  CU(cudaSetDevice(op->device));
  if (component > 0) op->mom_src = true;
  double *dev;
  bool temp;
  RDYB_CALL(dataset_upload(op, data, nvals, &dev, &temp));
  dataset_gather_source_kernel<<<grid_for(op->N, 256), 256, 0, op->stream>>>(op->N, op->dev_to_owned, map->idx, dev, offset, stride, scale,
                                                                             op->src + (size_t)component * op->stride);
  LAUNCHED();
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(op->stream)); /* the host array may be reused by the caller; the staging buffer by the next setter */
  if (temp) cudaFree(dev);
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_set_boundary_values_from_dataset(RdybOperator op, int32_t b, RdybDatasetMap map, const double *data, int32_t nvals,
                                                              int32_t offset) {
  RDYB_CHECK(op && map && data, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(b >= 0 && b < op->L.num_boundaries, RDYB_ERR_USER, "Invalid boundary index: %d", b);
  const int32_t ne = op->L.b_offsets[b + 1] - op->L.b_offsets[b];
  RDYB_CHECK(map->n == ne, RDYB_ERR_USER, "Wrong number of boundary edges for boundary %d: %d (should be %d)", b, map->n, ne);
  RDYB_CHECK(offset >= 0 && (int64_t)offset + 3 * (int64_t)map->ndata <= nvals, RDYB_ERR_USER,
             "The dataset (%d values) is too short for %d data points of stride 3 after a header of %d", nvals, map->ndata, offset);
  CU(cudaSetDevice(op->device));
  if (ne == 0) return RDYB_SUCCESS;
  double *dev;
  bool temp;
  RDYB_CALL(dataset_upload(op, data, nvals, &dev, &temp));
  dataset_gather_boundary_kernel<<<grid_for(ne, 256), 256, 0, op->stream>>>(ne, map->idx, dev, offset, op->bval + op->L.b_offsets[b], op->nbe);
  LAUNCHED();
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(op->stream));
  if (temp) cudaFree(dev);
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_fill_source(RdybOperator op, int32_t component, double value) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  RDYB_CHECK(component >= 0 && component < 3, RDYB_ERR_USER, "Invalid source component: %d", component);
  CU(cudaSetDevice(op->device));
  if (component > 0 && value != 0.0) op->mom_src = true;
  if (op->N == 0) return RDYB_SUCCESS;
  fill_kernel<<<grid_for(op->N, 256), 256, 0, op->stream>>>(op->N, value, op->src + (size_t)component * op->stride);
  LAUNCHED();
  CU(cudaGetLastError());
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_fill_boundary_values(RdybOperator op, int32_t b, double h, double hu, double hv) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  RDYB_CHECK(b >= 0 && b < op->L.num_boundaries, RDYB_ERR_USER, "Invalid boundary index: %d", b);
  const int32_t ne = op->L.b_offsets[b + 1] - op->L.b_offsets[b];
  CU(cudaSetDevice(op->device));
  if (ne == 0) return RDYB_SUCCESS;
  const double v[3] = {h, hu, hv};
  for (int c = 0; c < 3; ++c) {
    fill_kernel<<<grid_for(ne, 256), 256, 0, op->stream>>>(ne, v[c], op->bval + (size_t)c * op->nbe + op->L.b_offsets[b]);
    LAUNCHED();
  }
  CU(cudaGetLastError());
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_get_boundary_values(RdybOperator op, int32_t b, int32_t num_edges, double *values) {
  RDYB_CHECK(op && values, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(b >= 0 && b < op->L.num_boundaries, RDYB_ERR_USER, "Invalid boundary index: %d", b);
  const int32_t ne = op->L.b_offsets[b + 1] - op->L.b_offsets[b];
  RDYB_CHECK(num_edges == ne, RDYB_ERR_USER, "Wrong number of boundary edges for boundary %d: %d (should be %d)", b, num_edges, ne);
  CU(cudaSetDevice(op->device));
  std::vector<double> tmp(3 * (size_t)ne);
  for (int c = 0; c < 3; ++c)
    CU(cudaMemcpyAsync(tmp.data() + (size_t)c * ne, op->bval + (size_t)c * op->nbe + op->L.b_offsets[b], sizeof(double) * ne, cudaMemcpyDeviceToHost, op->stream));
  CU(cudaStreamSynchronize(op->stream));
  for (int32_t e = 0; e < ne; ++e)
    for (int c = 0; c < 3; ++c) values[3 * (size_t)e + c] = tmp[(size_t)c * ne + e];
  return RDYB_SUCCESS;
}

/* time series (src/time_series.c): the state at n observation sites without downloading the whole solution (WriteObservations,
 * :400-445, a VecScatter of u_global in the reference), and the reset of the boundary-flux accumulators after they have been
 * written (ResetAccumulatedBoundaryFluxes, :508-527) */
extern "C" int rdyb_operator_get_state_at(RdybOperator op, int32_t n, const int32_t *owned_cell_ids, double *values) {
  RDYB_CHECK(op && (n == 0 || (owned_cell_ids && values)), RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(n <= op->N, RDYB_ERR_ARG_SIZ, "more sites (%d) than owned cells (%d)", n, op->N);
  if (n == 0) return RDYB_SUCCESS;
  CU(cudaSetDevice(op->device));
  std::vector<int32_t> idx((size_t)n);
  for (int32_t i = 0; i < n; ++i) {
    RDYB_CHECK(owned_cell_ids[i] >= 0 && owned_cell_ids[i] < op->N, RDYB_ERR_USER, "owned cell id %d out of range", owned_cell_ids[i]);
    idx[i] = op->L.owned_to_dev[owned_cell_ids[i]];
  }
  CU(cudaMemcpyAsync(op->stage_idx, idx.data(), sizeof(int32_t) * n, cudaMemcpyHostToDevice, op->stream));
  sample_state_kernel<<<grid_for(n, 256), 256, 0, op->stream>>>(n, op->stage_idx, op->state[op->cur], op->stride, op->stage);
  LAUNCHED();
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(values, op->stage, sizeof(double) * 3 * (size_t)n, cudaMemcpyDeviceToHost, op->stream));
  CU(cudaStreamSynchronize(op->stream));
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_reset_boundary_fluxes(RdybOperator op) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  CU(cudaSetDevice(op->device));
  if (op->bacc && op->nbe > 0) CU(cudaMemsetAsync(op->bacc, 0, sizeof(double) * 3 * (size_t)op->nbe, op->stream));
  return RDYB_SUCCESS;
}
#endif
