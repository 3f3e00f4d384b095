// Shared device/host helpers for libhybridnet_sm100.so (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/hybridnet.h"

typedef __nv_bfloat16 bf16;
typedef __nv_bfloat162 bf162;

// ---- error plumbing (thread-local last error string) -----------------------
void hn_set_error(const char* fmt, ...);
#define HN_FAIL(code, ...)        \
  do {                            \
    hn_set_error(__VA_ARGS__);    \
    return (code);                \
  } while (0)
#define HN_REQUIRE(cond, ...)                        \
  do {                                               \
    if (!(cond)) HN_FAIL(HN_EINVAL, __VA_ARGS__);    \
  } while (0)
#define HN_CHECK_LAUNCH(name)                                                   \
  do {                                                                          \
    cudaError_t e__ = cudaGetLastError();                                       \
    if (e__ != cudaSuccess) HN_FAIL(HN_ECUDA, "%s: %s", name, cudaGetErrorString(e__)); \
  } while (0)

static inline int hn_cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

// n / d for n < 2^31 as one multiply-high and a shift (d is fixed per launch; the constants are found on the host).
// A 64-bit `/` or `%` by a runtime divisor is a ~100-instruction subroutine per call on the device.
struct FastDiv { uint32_t d, mul, shr; };
static inline FastDiv hn_fastdiv(uint32_t d) {
  FastDiv f{d, 0u, 0u};
  if (d > 1) {
    uint32_t l = 0;
    while ((1ull << l) < d) ++l;                     // ceil(log2 d)
    f.mul = (uint32_t)(((1ull << (31 + l)) + d - 1) / d);
    f.shr = l - 1;
  }
  return f;
}

// ---- device helpers ----------------------------------------------------------
__device__ __forceinline__ uint32_t fd_div(uint32_t n, const FastDiv f) { return f.d == 1 ? n : __umulhi(n, f.mul) >> f.shr; }
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(sz));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}
__device__ __forceinline__ void ldsm_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x2(uint32_t& r0, uint32_t& r1, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];\n"
               : "=r"(r0), "=r"(r1) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x2_t(uint32_t& r0, uint32_t& r1, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];\n"
               : "=r"(r0), "=r"(r1) : "r"(addr));
}
// D(16x8,f32) += A(16x16,bf16,row) * B(16x8,bf16,col)
__device__ __forceinline__ void mma_16816(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// D(16x8,f32) += A(16x8,bf16,row) * B(8x8,bf16,col)
__device__ __forceinline__ void mma_1688(float* d, uint32_t a0, uint32_t a1, uint32_t b0) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  bf162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float bf16_round(float v) { return __bfloat162float(__float2bfloat16(v)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float gelu_f(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float gelu_grad_f(float x) {
  float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
  float pdf = 0.3989422804014327f * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}
__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + __expf(-x)); }
// Epilogue activation with a RUNTIME selector.  none/ReLU are branch-free (FMNMX + FSEL); GELU and
// sigmoid sit behind a warp-uniform branch to an out-of-line call.  Inlining erff at each of the 64-128 unrolled
// epilogue sites made conv_sc_kernel<8> instruction-fetch bound (ncu: 44 % of samples on BSYNC with stall_no_inst,
// profiles/r1_end_ncu_full_conv_sc_8_64.txt) and put an indirect branch into every conv_tc epilogue chunk.
static __device__ __noinline__ float apply_act_slow(float v, int act) {
  return act == HN_ACT_GELU ? gelu_f(v) : sigmoid_f(v);
}
__device__ __forceinline__ float apply_act(float v, int act) {
  v = act == HN_ACT_RELU ? fmaxf(v, 0.f) : v;          // FMNMX + FSEL, NaN passes through when there is no activation
  if (act > HN_ACT_RELU) v = apply_act_slow(v, act);
  return v;
}

// 8 x bf16 <-> 8 x float through ONE 16-byte access (uint4: a struct of four bf162 is copied
// member-wise by nvcc and degrades to four 32-bit accesses)
__device__ __forceinline__ void load8(const bf16* p, float* f) {
  const uint4 r = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}
__device__ __forceinline__ void store8(bf16* p, const float* f) {
  uint4 r;
  r.x = pack_bf16(f[0], f[1]); r.y = pack_bf16(f[2], f[3]);
  r.z = pack_bf16(f[4], f[5]); r.w = pack_bf16(f[6], f[7]);
  *reinterpret_cast<uint4*>(p) = r;
}

// stateless counter-based RNG for dropout: one 32-bit hash per element index
__device__ __forceinline__ uint32_t hn_hash(uint64_t seed, uint64_t idx) {
  uint64_t z = idx * 0x9E3779B97F4A7C15ull + seed;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  return (uint32_t)(z >> 32);
}
__device__ __forceinline__ uint64_t hn_hash64(uint64_t seed, uint64_t idx) {
  uint64_t z = idx * 0x9E3779B97F4A7C15ull + seed;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// ---- column (per-channel) pass helpers ---------------------------------------------------
// A thread owns channel vector cv = thread % CV (the launch makes the total thread count a
// multiple of CV) and walks pixels pix0, pix0+pstride, ... -- no division in the loop.
struct ColIter { long long pix0, pstride; int cv; };
__device__ __forceinline__ ColIter col_iter(int CV) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long n = gridDim.x * (long long)blockDim.x;
  ColIter it;
  it.pix0 = t / CV; it.pstride = n / CV; it.cv = (int)(t - it.pix0 * CV);
  return it;
}
// Fold 8 per-thread partial sums into s_acc[cv*8 .. +8).  When CV divides 32 the lanes of a
// warp that share cv (lane % CV) are combined by shuffles first, so only CV lanes touch smem.
// Must be called by all threads of the warp.
__device__ __forceinline__ void col_fold8(float* s_acc, float* part, int cv, int CV) {
  const int lane = threadIdx.x & 31;
  if (CV < 32 && (CV & (CV - 1)) == 0) {
    for (int o = CV; o < 32; o <<= 1) {
#pragma unroll
      for (int j = 0; j < 8; ++j) part[j] += __shfl_xor_sync(0xffffffffu, part[j], o);
    }
    if (lane < CV) {
#pragma unroll
      for (int j = 0; j < 8; ++j) atomicAdd(&s_acc[cv * 8 + j], part[j]);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(&s_acc[cv * 8 + j], part[j]);
  }
}
