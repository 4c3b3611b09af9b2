// peak_fma.cu — measured FP64 / FP32 FMA-pipe peak (the roofline denominator of this compute-bound path;
// MEASURED_PEAKS.json only carries HBM and bf16 tensor figures).  8 independent FMA chains per thread.
#include <cuda_runtime.h>
#include "rmpc_model.cuh"

template <class T>
__global__ void fma_kernel(T* out, int iters, T a, T b) {
  T x0 = T(threadIdx.x), x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 4
  for (int i = 0; i < iters; ++i) {
    x0 = x0 * a + b; x1 = x1 * a + b; x2 = x2 * a + b; x3 = x3 * a + b;
    x4 = x4 * a + b; x5 = x5 * a + b; x6 = x6 * a + b; x7 = x7 * a + b;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

// the solver's reciprocal / reciprocal square root, exposed for the accuracy test
__global__ void recip_kernel(int n, const double* x, double* rcp, double* rsq) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { rcp[i] = rmpc::trcp(x[i]); rsq[i] = rmpc::trsqrt(x[i]); }
}

extern "C" {
int recip_launch(int n, const void* x, void* rcp, void* rsq, void* stream) {
  recip_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(n, (const double*)x, (double*)rcp, (double*)rsq);
  return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
// returns flops executed; the caller times the call with CUDA events on `stream`.  out must hold blocks*threads elements.
double peak_fma_launch(void* out, int blocks, int threads, int iters, int is_double, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (is_double) fma_kernel<double><<<blocks, threads, 0, st>>>((double*)out, iters, 0.999999, 1e-6);
  else fma_kernel<float><<<blocks, threads, 0, st>>>((float*)out, iters, 0.999999f, 1e-6f);
  if (cudaGetLastError() != cudaSuccess) return -1.0;
  return 2.0 * 8.0 * double(iters) * double(blocks) * double(threads);
}
}
