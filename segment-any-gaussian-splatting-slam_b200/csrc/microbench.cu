// FP32 pipe microbenchmark for the roofline of the blend kernels ("MEASURED_PEAKS has no FP32 figure --
// add an FMA microbenchmark"). Not part of the product path: built into its own library (lib/libgsr_microbench.so)
// and driven by tests/gpu_fma_peak.py and bench.py's roofline leg.
//
// Every thread runs `iters` rounds over CHAINS independent dependent-free accumulator chains; all SMs are filled with
// resident CTAs. Three instruction mixes:
//   mode 0  FFMA    scalar fp32 fused multiply-add (what the blend kernels issue per channel)
//   mode 1  FFMA2   Blackwell's packed fma.rn.f32x2 (two fp32 FMAs per issued instruction)
//   mode 2  FMUL + FADD pairs (non-fused; the reference's exact alpha / T arithmetic is of this kind)
// Result: elapsed milliseconds measured with CUDA events on the given stream; the caller converts to FLOP/s.
#include <cstdint>
#include <cuda_runtime.h>

namespace {

constexpr int CHAINS = 8;

template <int MODE>
__global__ void __launch_bounds__(256) fma_kernel(float* out, int iters, float a, float b) {
  float acc[CHAINS];
  unsigned long long acc2[CHAINS];
#pragma unroll
  for (int c = 0; c < CHAINS; c++) {
    acc[c] = (float)(threadIdx.x + c);
    float2 t = make_float2(acc[c], acc[c] + 0.5f);
    acc2[c] = *reinterpret_cast<unsigned long long*>(&t);
  }
  float2 ab2 = make_float2(a, a), bb2 = make_float2(b, b);
  const unsigned long long a2 = *reinterpret_cast<unsigned long long*>(&ab2), b2 = *reinterpret_cast<unsigned long long*>(&bb2);
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int r = 0; r < 8; r++) {
#pragma unroll
      for (int c = 0; c < CHAINS; c++) {
        if (MODE == 0) acc[c] = fmaf(acc[c], a, b);
        if (MODE == 1) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc2[c]) : "l"(a2), "l"(b2));
        if (MODE == 2) acc[c] = __fadd_rn(__fmul_rn(acc[c], a), b);
      }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < CHAINS; c++) {
    float2 t = *reinterpret_cast<float2*>(&acc2[c]);
    s += acc[c] + t.x + t.y;
  }
  if (s == 123.456f) out[0] = s;  // never true: keeps the chains alive
}

}  // namespace

extern "C" {

// flops per thread per iteration of the given mode (for the caller's arithmetic)
int gsr_mb_flops_per_thread_iter(int mode) { return 8 * CHAINS * (mode == 1 ? 4 : 2); }

// Launches `blocks` CTAs of 256 threads, `iters` iterations each, `reps` times back to back; returns the average
// elapsed milliseconds of one launch (CUDA events on `stream`), negative on a CUDA error.
float gsr_mb_fma(int mode, int blocks, int iters, int reps, float* scratch, void* stream_) {
  cudaStream_t s = (cudaStream_t)stream_;
  cudaEvent_t e0, e1;
  if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) return -1.f;
  auto launch = [&]() {
    if (mode == 0) fma_kernel<0><<<blocks, 256, 0, s>>>(scratch, iters, 0.999f, 0.001f);
    else if (mode == 1) fma_kernel<1><<<blocks, 256, 0, s>>>(scratch, iters, 0.999f, 0.001f);
    else fma_kernel<2><<<blocks, 256, 0, s>>>(scratch, iters, 0.999f, 0.001f);
  };
  launch();  // warm-up
  cudaEventRecord(e0, s);
  for (int r = 0; r < reps; r++) launch();
  cudaEventRecord(e1, s);
  float ms = -1.f;
  if (cudaEventSynchronize(e1) == cudaSuccess && cudaGetLastError() == cudaSuccess) {
    cudaEventElapsedTime(&ms, e0, e1);
    ms /= (float)reps;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return ms;
}

}  // extern "C"
