/* swe_noalloc.cu — rdyb_mass_sum_noalloc: the mass diagnostic (sum of area*h over owned cells) for callers that may not allocate.
 *
 * No counterpart in the reference; added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the allocation-free variant of rdyb_operator_get_mass
 * (swe_device.cu): the same two-stage deterministic reduction - 1024 blocks of 256 threads, a fixed shared-memory tree per block,
 * then the 1024 partial sums added in index order - with two differences:
 *   - the partial sums live in a file-scope `static __device__` array instead of an operator's cudaMalloc'ed scratch, and
 *   - the second stage runs on the device and writes ONE double to a caller-provided DEVICE address, so the call enqueues two
 *     kernels on the caller's stream and returns: no allocation, no copy, no synchronisation.
 *
 * The array (RDYB_NOALLOC_PARTIALS doubles = 8 KiB) is part of the library image: every GPU that loads librdyb200.so holds it until
 * the process exits.  Every call on a device shares it.  Nothing in it has to start at zero - each call overwrites all 1024 partial
 * sums before it reads them - but two calls that OVERLAP on one GPU (different streams, or two host threads without ordering)
 * race on it: order them on one stream or with events.
 *
 * In host code the array's name is a host-side stand-in, not a device address.  The address a kernel may use is obtained once per
 * device, with that device current, through cudaGetSymbolAddress (noalloc_scratch below) and cached; that first call per device
 * may load the module, so make it before a stream capture begins.
 */
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "rdyb200.h"
#include "rdyb_error.h"

#define RDYB_NOALLOC_PARTIALS 1024 /* = the grid of rdyb_operator_get_mass: the partial sums are the same numbers */
#define RDYB_NOALLOC_THREADS 256
#define RDYB_NOALLOC_MAX_DEVICES 64

static_assert(RDYB_NOALLOC_MAX_CELLS <= INT32_MAX - RDYB_NOALLOC_PARTIALS * RDYB_NOALLOC_THREADS, "the grid-stride index is a 32-bit int");

static __device__ __align__(256) double g_noalloc_partial[RDYB_NOALLOC_PARTIALS];

/* stage 1: the reduction of mass_partial_kernel (swe_kernels.cuh) - fixed blocks, fixed tree, no atomics */
__global__ void __launch_bounds__(RDYB_NOALLOC_THREADS) noalloc_mass_partial_kernel(const double *__restrict__ h, const double *__restrict__ area, int32_t n,
                                                                                    double *__restrict__ partial) {
  __shared__ double s[RDYB_NOALLOC_THREADS];
  double acc = 0.0;
  for (int i = blockIdx.x * RDYB_NOALLOC_THREADS + threadIdx.x; i < n; i += gridDim.x * RDYB_NOALLOC_THREADS) acc += area[i] * h[i];
  s[threadIdx.x] = acc;
  __syncthreads();
  for (int off = RDYB_NOALLOC_THREADS / 2; off > 0; off >>= 1) {
    if ((int)threadIdx.x < off) s[threadIdx.x] += s[threadIdx.x + off];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = s[0];
}

/* stage 2: the host loop of rdyb_operator_get_mass (s += partial[i], i ascending) on the device: the partial sums are staged in shared
 * memory by the whole block, one thread adds them in index order */
__global__ void __launch_bounds__(RDYB_NOALLOC_THREADS) noalloc_mass_final_kernel(const double *__restrict__ partial, double *__restrict__ out) {
  __shared__ double s[RDYB_NOALLOC_PARTIALS];
  for (int i = threadIdx.x; i < RDYB_NOALLOC_PARTIALS; i += RDYB_NOALLOC_THREADS) s[i] = partial[i];
  __syncthreads();
  if (threadIdx.x == 0) {
    double acc = 0.0;
    for (int i = 0; i < RDYB_NOALLOC_PARTIALS; ++i) acc = __dadd_rn(acc, s[i]);
    *out = acc;
  }
}

/* device address of g_noalloc_partial on `device`, which must be the current device: resolved once per device and cached */
static int noalloc_scratch(int device, double **ptr) {
  static std::mutex mu;
  static double *cached[RDYB_NOALLOC_MAX_DEVICES];
  std::lock_guard<std::mutex> lock(mu);
  if (!cached[device]) {
    double *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_partial);
    if (e != cudaSuccess || !p)
      return rdyb_fail(RDYB_ERR_CUDA, "rdyb_mass_sum_noalloc: cudaGetSymbolAddress of the static scratch failed on device %d: %s (%s)", device,
                       cudaGetErrorName(e), cudaGetErrorString(e));
    cached[device] = p;
  }
  *ptr = cached[device];
  return RDYB_SUCCESS;
}

extern "C" int rdyb_mass_sum_noalloc(const double *h, const double *area, int64_t num_cells, double *mass, void *cuda_stream) {
  RDYB_CHECK(mass, RDYB_ERR_ARG, "rdyb_mass_sum_noalloc: NULL output");
  RDYB_CHECK(num_cells >= 0, RDYB_ERR_ARG_SIZ, "rdyb_mass_sum_noalloc: negative cell count %lld", (long long)num_cells);
  RDYB_CHECK(num_cells <= (int64_t)RDYB_NOALLOC_MAX_CELLS, RDYB_ERR_ARG_SIZ,
             "rdyb_mass_sum_noalloc: %lld cells exceed the maximum of %lld (RDYB_NOALLOC_MAX_CELLS) of the allocation-free entry point; "
             "use rdyb_operator_get_mass",
             (long long)num_cells, (long long)RDYB_NOALLOC_MAX_CELLS);
  RDYB_CHECK(num_cells == 0 || (h && area), RDYB_ERR_ARG, "rdyb_mass_sum_noalloc: NULL input");
  int device = -1;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess) return rdyb_fail(RDYB_ERR_CUDA, "rdyb_mass_sum_noalloc: no CUDA device: %s (%s)", cudaGetErrorName(e), cudaGetErrorString(e));
  RDYB_CHECK(device >= 0 && device < RDYB_NOALLOC_MAX_DEVICES, RDYB_ERR_SUP, "rdyb_mass_sum_noalloc: device %d out of range", device);
  double *scratch = nullptr;
  RDYB_CALL(noalloc_scratch(device, &scratch));
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  noalloc_mass_partial_kernel<<<RDYB_NOALLOC_PARTIALS, RDYB_NOALLOC_THREADS, 0, stream>>>(h, area, (int32_t)num_cells, scratch);
  noalloc_mass_final_kernel<<<1, RDYB_NOALLOC_THREADS, 0, stream>>>(scratch, mass);
  e = cudaGetLastError();
  if (e != cudaSuccess) return rdyb_fail(RDYB_ERR_CUDA, "rdyb_mass_sum_noalloc: launch failed: %s (%s)", cudaGetErrorName(e), cudaGetErrorString(e));
  return RDYB_SUCCESS;
}
