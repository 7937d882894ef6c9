// fp64 roof of this GPU, measured: dense DMMA (mma.sync.m8n8k4.f64, the instruction of csrc/cholesky.cu, csrc/chol_small.cu
// and csrc/schur.cu) and plain DFMA, register-resident operands, enough independent chains per warp to cover the pipe
// latency.  Prints one JSON line; the numbers are the denominators of the Cholesky's "fraction of the fp64 roof".
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_roof tools/fp64_roof.cu
#include <cuda_runtime.h>

#include <cstdio>

__device__ __forceinline__ void Dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

template <int CHAINS>
__global__ void __launch_bounds__(256) dmma_kernel(int iters, double* out) {
  double c[CHAINS][2];
  const double a = 1.0 + 1e-9 * threadIdx.x, b = 1.0 - 1e-9 * threadIdx.x;
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) { c[i][0] = 0.0; c[i][1] = 0.0; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) Dmma884(c[i][0], c[i][1], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) s += c[i][0] + c[i][1];
  if (s == 12345.678) out[0] = s;
}

template <int CHAINS>
__global__ void __launch_bounds__(256) dfma_kernel(int iters, double* out) {
  double c[CHAINS];
  const double a = 1.0 + 1e-9 * threadIdx.x, b = 1e-9 * threadIdx.x;
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) c[i] = i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) c[i] = fma(c[i], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < CHAINS; ++i) s += c[i];
  if (s == 12345.678) out[0] = s;
}

template <class K>
static double TimeMs(K launch) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  launch();  // warm-up
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  return best;
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out = nullptr;
  cudaMalloc(&out, 8);
  const int iters = 4096, ctas = sms * 8, warps = 8;
  double best_dmma = 0.0, best_dfma = 0.0;
  int best_chains = 0;
  auto dm = [&](auto kernel, int chains) {
    const double ms = TimeMs([&] { kernel<<<ctas, 256>>>(iters, out); });
    const double flops = (double)ctas * warps * iters * chains * (8.0 * 8.0 * 4.0 * 2.0);
    const double tf = flops / (ms * 1e-3) / 1e12;
    if (tf > best_dmma) { best_dmma = tf; best_chains = chains; }
  };
  dm(dmma_kernel<2>, 2); dm(dmma_kernel<4>, 4); dm(dmma_kernel<8>, 8); dm(dmma_kernel<16>, 16);
  auto df = [&](auto kernel, int chains) {
    const double ms = TimeMs([&] { kernel<<<ctas, 256>>>(iters, out); });
    const double flops = (double)ctas * 256 * iters * chains * 2.0;
    const double tf = flops / (ms * 1e-3) / 1e12;
    if (tf > best_dfma) best_dfma = tf;
  };
  df(dfma_kernel<4>, 4); df(dfma_kernel<8>, 8); df(dfma_kernel<16>, 16);
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, 0);
  std::printf("{\"gpu\": \"%s\", \"sms\": %d, \"dmma_m8n8k4_tflops\": %.2f, \"dmma_best_chains_per_warp\": %d, \"dfma_tflops\": %.2f, "
              "\"how\": \"register-resident operands, 8 CTAs x 8 warps per SM, best of 5, CUDA events\"}\n",
              prop.name, sms, best_dmma, best_chains, best_dfma);
  return cudaGetLastError() == cudaSuccess ? 0 : 1;
}
