// Throughput probe for the warp-level (legacy) tensor path on sm_100a: mma.sync m16n8k8 / m16n8k16 bf16, alone and
// interleaved with MUFU ex2 -- the mix of the flash-attention kernels (hn_attn_*.cu).  Prints warp-instructions per
// clock per SM sub-partition, i.e. how many clocks one HMMA occupies the pipe.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
template <int MODE>
__global__ void k(float* out, int iters, uint32_t seed) {
  float c[4][4];
  float x[8];
#pragma unroll
  for (int i = 0; i < 4; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = -0.001f * (threadIdx.x + i);
  uint32_t a0 = seed ^ threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (MODE == 0 || MODE == 2)
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+f"(c[u][0]), "+f"(c[u][1]), "+f"(c[u][2]), "+f"(c[u][3]) : "r"(a0), "r"(a1), "r"(b0));
      if (MODE == 1 || MODE == 3)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[u][0]), "+f"(c[u][1]), "+f"(c[u][2]), "+f"(c[u][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
      if (MODE >= 2) {     // two ex2 per mma: the attention ratio is ~1.6 ex2 per mma (d = 4)
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[2 * u]));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[2 * u + 1]));
      }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3] + x[2 * i] + x[2 * i + 1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char* name) {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const int iters = 20000;
  k<MODE><<<148 * 8, 256>>>(out, 100, 1u);
  cudaEventRecord(a);
  k<MODE><<<148 * 8, 256>>>(out, iters, 1u);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  const double warps = 148.0 * 8 * 8, mmas = warps * iters * 4.0;
  const double clocks = ms * 1e-3 * clk * 1e3;
  const double per_smsp = mmas / 148 / 4 / clocks;        // warp-level mma per clock per sub-partition
  const int kdim = (MODE == 0 || MODE == 2) ? 8 : 16;
  printf("%-34s %8.3f ms  %.3f mma/clk/SMSP = %.1f clk per mma;  %.1f dense TFLOP/s", name, ms, per_smsp, 1.0 / per_smsp,
         mmas * 2.0 * 16 * 8 * kdim / (ms * 1e-3) / 1e12);
  if (MODE >= 2) printf(";  %.2f ex2/clk/SM", mmas * 2.0 * 32 / 148 / clocks);
  printf("\n");
}
int main() {
  run<0>("mma.sync m16n8k8 bf16");
  run<1>("mma.sync m16n8k16 bf16");
  run<2>("m16n8k8 + 2 ex2 per mma");
  run<3>("m16n8k16 + 2 ex2 per mma");
  return 0;
}
