// FMA microbenchmark for the FP-pipe roofline denominators (MEASURED_PEAKS.json has HBM and bf16 only).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/fp_peak tools/fp_peak.cu ; ./tools/fp_peak
// Prints one JSON object: dependent-issue latency of DFMA / FFMA (clock64 around a chain in one warp),
// and the sustained FMA rate of the whole GPU for several (warps per SM, ILP) points, CUDA-event timed.
#include <cstdio>
#include <cuda_runtime.h>

template <class T, int ILP>
__global__ void fma_chain(T* out, long long* cyc, int iters, T a, T b) {
  T v[ILP];
#pragma unroll
  for (int k = 0; k < ILP; ++k) v[k] = (T)(threadIdx.x + k);
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < ILP; ++k) v[k] = v[k] * a + b;
  }
  const long long t1 = clock64();
  T s = 0;
#pragma unroll
  for (int k = 0; k < ILP; ++k) s += v[k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <class T, int ILP> double rate(int ctas, int threads, int iters, T* out, long long* cyc) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  fma_chain<T, ILP><<<ctas, threads>>>(out, cyc, iters, (T)1.0000001, (T)1e-9);
  cudaEventRecord(e0);
  fma_chain<T, ILP><<<ctas, threads>>>(out, cyc, iters, (T)1.0000001, (T)1e-9);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  return 2.0 * ILP * (double)iters * ctas * threads / (ms * 1e-3) / 1e12;  // TFLOP/s
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  void* out; long long* cyc;
  cudaMalloc(&out, (size_t)sms * 2048 * 8 * 2);
  cudaMallocManaged(&cyc, 8);
  const int it = 1 << 14;
  // latency: one warp, ILP 1
  fma_chain<double, 1><<<1, 32>>>((double*)out, cyc, it, 1.0000001, 1e-9); cudaDeviceSynchronize();
  const double lat64 = (double)*cyc / it;
  fma_chain<float, 1><<<1, 32>>>((float*)out, cyc, it, 1.0000001f, 1e-9f); cudaDeviceSynchronize();
  const double lat32 = (double)*cyc / it;
  // one warp, ILP 8: issue interval of back-to-back independent FMAs
  fma_chain<double, 8><<<1, 32>>>((double*)out, cyc, it, 1.0000001, 1e-9); cudaDeviceSynchronize();
  const double iss64 = (double)*cyc / it / 8;
  fma_chain<float, 8><<<1, 32>>>((float*)out, cyc, it, 1.0000001f, 1e-9f); cudaDeviceSynchronize();
  const double iss32 = (double)*cyc / it / 8;
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_mhz\": %d,\n", p.name, sms, p.clockRate / 1000);
  printf(" \"dfma_latency_cycles\": %.2f, \"ffma_latency_cycles\": %.2f, \"dfma_issue_cycles_one_warp\": %.2f, \"ffma_issue_cycles_one_warp\": %.2f,\n",
         lat64, lat32, iss64, iss32);
  printf(" \"fp64_tflops\": {");
  const int wps[] = {4, 8, 16, 32, 64};
  double best64 = 0, best32 = 0;
  for (int w : wps) {
    const int threads = w * 32 > 1024 ? 1024 : w * 32, per_sm = (w * 32 + threads - 1) / threads;
    const double a = rate<double, 1>(sms * per_sm, threads, it, (double*)out, cyc), b = rate<double, 2>(sms * per_sm, threads, it, (double*)out, cyc),
                 c = rate<double, 8>(sms * per_sm, threads, it, (double*)out, cyc);
    printf("%s\"warps%d\": {\"ilp1\": %.2f, \"ilp2\": %.2f, \"ilp8\": %.2f}", w == 4 ? "" : ", ", w, a, b, c);
    if (c > best64) best64 = c; if (b > best64) best64 = b;
  }
  printf("},\n \"fp32_tflops\": {");
  for (int w : wps) {
    const int threads = w * 32 > 1024 ? 1024 : w * 32, per_sm = (w * 32 + threads - 1) / threads;
    const double a = rate<float, 1>(sms * per_sm, threads, it, (float*)out, cyc), b = rate<float, 2>(sms * per_sm, threads, it, (float*)out, cyc),
                 c = rate<float, 8>(sms * per_sm, threads, it, (float*)out, cyc);
    printf("%s\"warps%d\": {\"ilp1\": %.2f, \"ilp2\": %.2f, \"ilp8\": %.2f}", w == 4 ? "" : ", ", w, a, b, c);
    if (c > best32) best32 = c; if (b > best32) best32 = b;
  }
  printf("},\n \"fp64_peak_tflops\": %.2f, \"fp32_peak_tflops\": %.2f}\n", best64, best32);
  return 0;
}
