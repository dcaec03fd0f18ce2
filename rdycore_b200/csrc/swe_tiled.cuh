/* swe_tiled.cuh — K2 "tiled" kernel: placeholder until the shared-memory tile kernel lands. */
#ifndef RDYB_SWE_TILED_CUH
#define RDYB_SWE_TILED_CUH
#include <cuda_runtime.h>

#include "device_layout.h"
#include "rdyb200.h"
#include "rdyb_error.h"

struct TiledDev {
  int dummy;
};
static inline void tiled_dev_free(TiledDev *) {}
static inline int tiled_dev_upload(const RdybTiledLayout *, const RdybLayout *, TiledDev *) { return rdyb_fail(RDYB_ERR_SUP, "tiled kernel not built"); }
static inline int tiled_launch(const TiledDev &, const RdybTiledLayout &, const RdybConfig &, const double *, double *, const double *, bool,
                               const double *, const int32_t *, const double *, double *, double *, int32_t, double *, double *, double *,
                               long long *, double, bool, int32_t, int32_t, cudaStream_t) {
  return rdyb_fail(RDYB_ERR_SUP, "tiled kernel not built");
}
#endif
