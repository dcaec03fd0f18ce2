// fp64 pipe microbenchmark on sm_100a: dependent-chain latency and throughput vs. warps per SM and ILP.
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP>
__global__ void chain(double *out, int iters, long long *cycles) {
  double x[ILP];
#pragma unroll
  for (int j = 0; j < ILP; ++j) x[j] = 1.0 + 1e-9 * (threadIdx.x + j);
  const double a = 1.0000001, b = 1e-12;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
#pragma unroll
      for (int j = 0; j < ILP; ++j) x[j] = fma(x[j], a, b);
    }
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int j = 0; j < ILP; ++j) s += x[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
__global__ void mufu_chain(double *out, int iters, long long *cycles) {
  double x = 1.0 + 1e-9 * threadIdx.x;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) { double y; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x)); x = y + 0.5; }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
__global__ void lds_chain(double *out, int iters, long long *cycles) {
  __shared__ int idx[1024];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) idx[i] = (i * 33 + 7) & 1023;
  __syncthreads();
  int p = threadIdx.x;
  long long t0 = clock64();
  for (int i = 0; i < iters * 8; ++i) p = idx[p];
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = p;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
template <int ILP>
void run(int threads, int blocks, double *out, long long *cyc) {
  const int iters = 20000;
  chain<ILP><<<blocks, threads>>>(out, iters, cyc);
  cudaDeviceSynchronize();
  long long c;
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  const double per = (double)c / (iters * 8.0 * ILP);
  printf("DFMA ILP %d, %4d threads/block x %d blocks/SM-ish: %.2f cycles per DFMA per warp-chain; warp-DFMA per cycle per SMSP %.3f\n", ILP, threads, blocks / 148,
         per, (threads / 32.0) * (blocks / 148) / 4.0 / per);
}
int main() {
  double *out; long long *cyc;
  cudaMalloc(&out, 8 * 1024 * 148 * 8); cudaMalloc(&cyc, 8);
  run<1>(32, 148, out, cyc); run<2>(32, 148, out, cyc); run<4>(32, 148, out, cyc); run<8>(32, 148, out, cyc);
  run<1>(128, 148, out, cyc); run<2>(128, 148, out, cyc); run<1>(256, 148, out, cyc); run<1>(512, 148, out, cyc); run<2>(512, 148, out, cyc);
  run<1>(1024, 148, out, cyc); run<4>(512, 148, out, cyc);
  long long c;
  mufu_chain<<<148, 32>>>(out, 20000, cyc); cudaDeviceSynchronize(); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  printf("MUFU.RCP64H + DADD dependent pair: %.2f cycles\n", (double)c / (20000 * 8.0));
  mufu_chain<<<148, 512>>>(out, 20000, cyc); cudaDeviceSynchronize(); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  printf("MUFU.RCP64H + DADD pair, 16 warps/SM: %.2f cycles per pair per warp\n", (double)c / (20000 * 8.0));
  lds_chain<<<148, 32>>>(out, 20000, cyc); cudaDeviceSynchronize(); cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  printf("dependent LDS.32: %.2f cycles\n", (double)c / (20000 * 8.0));
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
