// Micro-benchmarks that size the tcgen05 attention design (run on one B200):
//   * MUFU ex2 throughput: f32, f16x2, bf16x2 (results per clock per SM)
//   * issue cost of the softmax helper instructions (fma.rn.f32x2, 3-input max, cvt.rn.bf16x2.f32, fma.rn.bf16x2)
//   * tcgen05.ld / tcgen05.st throughput (32x32b.x32) with 4 and 8 warps per CTA
//   * the softmax inner loop itself on TMEM-resident scores (ld S -> max -> scale/sub -> cvt -> ex2 -> st P)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/build/ubench scripts/ubench.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---------------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void alu_kernel(float* out, int iters, float seed) {
  float x[8];
  uint32_t h[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { x[i] = seed + threadIdx.x * 1e-3f + 0.1f * i; h[i] = (0x3c003c00u ^ threadIdx.x) + i; }
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
    } else if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(h[i]));
    } else if (MODE == 2) {
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h[i]));
    } else if (MODE == 3) {          // packed fp32 FMA: 2 results per instruction
#pragma unroll
      for (int i = 0; i < 8; i += 2) {
        asm volatile("{.reg .b64 a, b, c; mov.b64 a, {%0, %1}; mov.b64 b, {%2, %2}; mov.b64 c, {%3, %3};\n"
                     "fma.rn.f32x2 a, a, b, c; mov.b64 {%0, %1}, a;}" : "+f"(x[i]), "+f"(x[i + 1]) : "f"(0.999f), "f"(1e-3f));
      }
    } else if (MODE == 4) {          // 3-input max
#pragma unroll
      for (int i = 0; i < 8; i += 2) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(x[i + 1]), "f"(seed));
    } else if (MODE == 5) {          // cvt pack
#pragma unroll
      for (int i = 0; i < 8; i += 2) { asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(x[i]), "f"(x[i + 1])); x[i] += __uint_as_float(h[i] << 16); }
    } else if (MODE == 6) {          // packed bf16 FMA
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("fma.rn.bf16x2 %0, %0, %1, %2;" : "+r"(h[i]) : "r"(0x3f803f80u), "r"(0x3c003c00u));
    } else if (MODE == 7) {          // plain FFMA for reference
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(0.999f), "f"(1e-3f));
    } else if (MODE == 8) {          // softmax ALU mix on registers: per 2 elements max3 + ffma2 + cvt + ex2.bf16x2
#pragma unroll
      for (int i = 0; i < 8; i += 2) {
        float m = seed;
        asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(m) : "f"(x[i]), "f"(x[i + 1]));
        asm volatile("{.reg .b64 a, b, c; mov.b64 a, {%0, %1}; mov.b64 b, {%2, %2}; mov.b64 c, {%3, %3};\n"
                     "fma.rn.f32x2 a, a, b, c; mov.b64 {%0, %1}, a;}" : "+f"(x[i]), "+f"(x[i + 1]) : "f"(0.999f), "f"(-m));
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(x[i]), "f"(x[i + 1]));
        asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(h[i]));
        x[i] += 1e-3f;
      }
    } else if (MODE == 9) {          // same with f32 ex2 (2 MUFU per pair) + cvt
#pragma unroll
      for (int i = 0; i < 8; i += 2) {
        float m = seed;
        asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(m) : "f"(x[i]), "f"(x[i + 1]));
        float a = fmaf(x[i], 0.999f, -m), b = fmaf(x[i + 1], 0.999f, -m);
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(b));
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(a), "f"(b));
        x[i] += 1e-3f; x[i + 1] += __uint_as_float(h[i] << 16);
      }
    }
  }
  float s = 0.f;
  uint32_t hx = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) { s += x[i]; hx ^= h[i]; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + __uint_as_float(hx);
}

template <int MODE>
int run_alu(const char* name, double per_iter_results) {
  float* out;
  CK(cudaMalloc(&out, 148 * 8 * 256 * 4));
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  const int iters = 10000;
  alu_kernel<MODE><<<148 * 8, 256>>>(out, 100, 0.5f);
  cudaEventRecord(a);
  alu_kernel<MODE><<<148 * 8, 256>>>(out, iters, 0.5f);
  cudaEventRecord(b);
  CK(cudaEventSynchronize(b));
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  double res = 148.0 * 8 * 256 * iters * per_iter_results;
  int clk;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("%-44s %8.3f ms  %.3e results/s  %6.1f results/clk/SM (nominal %d MHz)\n", name, ms, res / (ms * 1e-3),
         res / (ms * 1e-3) / 148 / (clk * 1e3), clk / 1000);
  cudaFree(out);
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// TMEM load/store throughput and the softmax loop on TMEM-resident scores
#define TMEM_LD32(taddr, r)                                                                                              \
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                 \
               "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];" \
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), \
                 "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),  \
                 "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), \
                 "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])               \
               : "r"(taddr))
#define TMEM_ST16(taddr, r)                                                                                              \
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "                                                           \
               "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"                                               \
               ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),      \
                 "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory")
#define TMEM_ST32(taddr, r)                                                                                              \
  asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "                                                           \
               "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" \
               ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),      \
                 "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), \
                 "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), \
                 "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory")

// MODE 0: ld only; 1: st only; 2: softmax loop with ex2.bf16x2; 3: softmax loop with f32 ex2; 4: mode 2 without the max pass
template <int MODE>
__global__ void __launch_bounds__(256, 1) tmem_kernel(float* out, int iters, long long* clocks) {
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tbase = tmem_slot;
  // warp w owns lanes 32*(w%4); with 8 warps the two warps of a quadrant take different 256-column halves
  const uint32_t trow = tbase + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 256;
  uint32_t r[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(0.01f * (j + lane) - 0.3f);
  for (int c = 0; c < 8; ++c) TMEM_ST32(trow + c * 32, r);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  __syncthreads();
  long long t0 = clock64();
  float acc = 0.f;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        TMEM_LD32(trow + c * 32, r);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        acc += __uint_as_float(r[it & 31]);
      }
    } else if (MODE == 1) {
#pragma unroll
      for (int c = 0; c < 4; ++c) TMEM_ST32(trow + c * 32, r);
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    } else {
      // one 128-column score row per thread: pass 1 max, pass 2 exp + packed store over the first 64 columns
      float m = -1e30f;
      if (MODE != 4) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          TMEM_LD32(trow + c * 32, r);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int j = 0; j < 32; j += 2) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(m) : "r"(r[j]), "r"(r[j + 1]));
        }
      } else {
        m = 0.25f;
      }
      const float sc = 1.4426950408889634f * 0.5f, nm = -m * sc;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        TMEM_LD32(trow + c * 32, r);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        uint32_t p[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          float a = __uint_as_float(r[2 * j]), b = __uint_as_float(r[2 * j + 1]);
          if (MODE == 3) {
            a = fmaf(a, sc, nm); b = fmaf(b, sc, nm);
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a));
            asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(b));
            asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(p[j]) : "f"(b), "f"(a));
          } else {
            asm volatile("{.reg .b64 a, b, c; mov.b64 a, {%0, %1}; mov.b64 b, {%2, %2}; mov.b64 c, {%3, %3};\n"
                         "fma.rn.f32x2 a, a, b, c; mov.b64 {%0, %1}, a;}" : "+f"(a), "+f"(b) : "f"(sc), "f"(nm));
            asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(p[j]) : "f"(b), "f"(a));
            asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(p[j]));
          }
        }
        TMEM_ST16(trow + 128 + c * 16, p);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) clocks[blockIdx.x] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc + __uint_as_float(r[0]);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase) : "memory");
}

template <int MODE>
int run_tmem(const char* name, int threads, double elems_per_iter_per_thread) {
  float* out;
  long long* clocks;
  CK(cudaMalloc(&out, 148 * 256 * 4));
  CK(cudaMalloc(&clocks, 148 * 8));
  const int iters = 2000;
  tmem_kernel<MODE><<<148, threads>>>(out, 10, clocks);
  CK(cudaDeviceSynchronize());
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  tmem_kernel<MODE><<<148, threads>>>(out, iters, clocks);
  cudaEventRecord(b);
  CK(cudaEventSynchronize(b));
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  long long h[148];
  CK(cudaMemcpy(h, clocks, sizeof(h), cudaMemcpyDeviceToHost));
  double cyc = 0;
  for (int i = 0; i < 148; ++i) cyc += h[i];
  cyc /= 148;
  double elems = (double)threads * iters * elems_per_iter_per_thread;
  printf("%-44s %d thr  %8.3f ms  %9.0f clk/CTA  %6.2f elem/clk/SM  (%.1f B/clk/SM as fp32)\n", name, threads, ms, cyc,
         elems / cyc, 4.0 * elems / cyc);
  cudaFree(out); cudaFree(clocks);
  return 0;
}

int main() {
  if (run_alu<0>("ex2.approx.ftz.f32", 8)) return 1;
  run_alu<1>("ex2.approx.f16x2", 16);
  run_alu<2>("ex2.approx.ftz.bf16x2", 16);
  run_alu<7>("fma.rn.f32 (3-reg)", 8);
  run_alu<3>("fma.rn.f32x2", 8);
  run_alu<4>("max.f32 3-input (2 new elems each)", 8);
  run_alu<5>("cvt.rn.bf16x2.f32 (+1 FADD per pair)", 8);
  run_alu<6>("fma.rn.bf16x2", 16);
  run_alu<8>("softmax mix: max3+ffma2+cvt+ex2.bf16x2", 8);
  run_alu<9>("softmax mix: max3+2ffma+2ex2.f32+cvt", 8);
  run_tmem<0>("tcgen05.ld 32x32b.x32 (+wait) x4", 128, 128);
  run_tmem<0>("tcgen05.ld 32x32b.x32 (+wait) x4", 256, 128);
  run_tmem<1>("tcgen05.st 32x32b.x32 x4", 128, 128);
  run_tmem<1>("tcgen05.st 32x32b.x32 x4", 256, 128);
  run_tmem<2>("softmax row: 2x ld + max3 + ffma2/cvt/ex2.bf16x2 + st", 128, 128);
  run_tmem<2>("softmax row: 2x ld + max3 + ffma2/cvt/ex2.bf16x2 + st", 256, 128);
  run_tmem<4>("softmax row: 1x ld (no max) + bf16x2 exp + st", 128, 128);
  run_tmem<4>("softmax row: 1x ld (no max) + bf16x2 exp + st", 256, 128);
  run_tmem<3>("softmax row: 2x ld + max3 + 2ffma/2ex2.f32/cvt + st", 128, 128);
  run_tmem<3>("softmax row: 2x ld + max3 + 2ffma/2ex2.f32/cvt + st", 256, 128);
  return 0;
}
