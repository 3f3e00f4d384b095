// Throughput probe: ex2.approx f32 vs f16x2 vs bf16x2 (results per clock per SM), and a Horner exp2 on the FMA pipe.
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
template <int MODE>
__global__ void k(float* out, int iters, float seed) {
  float x0 = seed + threadIdx.x * 1e-3f, x1 = x0 + 0.1f, x2 = x0 + 0.2f, x3 = x0 + 0.3f;
  uint32_t h0 = 0x3c003c00u ^ threadIdx.x, h1 = h0 + 1, h2 = h0 + 2, h3 = h0 + 3;
  for (int i = 0; i < iters; ++i) {
    if (MODE == 0) {
      asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x0)); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x1));
      asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x2)); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x3));
    } else if (MODE == 1) {
      asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h0)); asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h1));
      asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h2)); asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h3));
    } else if (MODE == 2) {
      asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h0)); asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h1));
      asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h2)); asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h3));
    } else {
      // exp2 for x <= 0 on the FMA/ALU pipes: split, degree-4 Horner, exponent add
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float& x = u == 0 ? x0 : (u == 1 ? x1 : (u == 2 ? x2 : x3));
        float xc = fmaxf(x, -126.f);
        float fl = floorf(xc);
        float f = xc - fl;
        float p = fmaf(f, 0.0135557f, 0.0520323f);
        p = fmaf(p, f, 0.2413793f); p = fmaf(p, f, 0.6930580f); p = fmaf(p, f, 1.0f);
        x = __int_as_float(__float_as_int(p) + ((int)fl << 23)) - 1.5f;
      }
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + __uint_as_float(h0 ^ h1 ^ h2 ^ h3);
}
template <int MODE>
void run(const char* name, int per_op) {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const int iters = 20000;
  k<MODE><<<148 * 8, 256>>>(out, 100, 0.5f);
  cudaEventRecord(a);
  k<MODE><<<148 * 8, 256>>>(out, iters, 0.5f);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double res = 148.0 * 8 * 256 * iters * 4.0 * per_op;
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("%-28s %8.3f ms  %.3e results/s  %.1f results/clk/SM (at %d MHz)\n", name, ms, res / (ms * 1e-3), res / (ms * 1e-3) / 148 / (clk * 1e3), clk / 1000);
}
int main() {
  run<0>("ex2.approx.ftz.f32", 1);
  run<1>("ex2.approx.f16x2", 2);
  run<2>("ex2.approx.ftz.bf16x2", 2);
  run<3>("poly exp2 (FMA pipe)", 1);
  return 0;
}
