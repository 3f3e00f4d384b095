// Shared fragment helpers for the flash-attention kernels (hn_attn_fwd.cu,
// hn_attn_bwd.cu).  Head dims D in {4,8,16,32,64}; DP = max(D,8) is the padded
// row length of every shared-memory tile (zero padded), so that a row is at
// least one 16-byte ldmatrix row.  Products whose contraction runs over the head
// dim use m16n8k8 when DP == 8 and m16n8k16 otherwise.
#pragma once
#include "hn_common.cuh"

namespace attn {

constexpr int TILE = 64;      // keys (or queries) per shared-memory tile
constexpr int NWARP = 4;      // 16 rows per warp -> 64 rows per CTA
constexpr float LOG2E = 1.4426950408889634f;

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// packed fp32 pairs (sm_100: one FFMA2 / FADD2 / FMUL2 issue slot for two results)
__device__ __forceinline__ void fma2(float& d0, float& d1, float a0, float a1, float b, float c) {
  asm("{.reg .b64 ra, rb, rc;\n"
      "mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %4}; mov.b64 rc, {%5, %5};\n"
      "fma.rn.ftz.f32x2 ra, ra, rb, rc;\n"
      "mov.b64 {%0, %1}, ra;}" : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b), "f"(c));
}
// (a0, a1) - (c, c)
__device__ __forceinline__ void sub2(float& d0, float& d1, float a0, float a1, float c) {
  asm("{.reg .b64 ra, rc;\n"
      "mov.b64 ra, {%2, %3}; mov.b64 rc, {%4, %4};\n"
      "sub.ftz.f32x2 ra, ra, rc;\n"
      "mov.b64 {%0, %1}, ra;}" : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(c));
}
__device__ __forceinline__ void mul2(float& d0, float& d1, float a0, float a1, float b0, float b1) {
  asm("{.reg .b64 ra, rb;\n"
      "mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5};\n"
      "mul.ftz.f32x2 ra, ra, rb;\n"
      "mov.b64 {%0, %1}, ra;}" : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}
__device__ __forceinline__ void add2(float& d0, float& d1, float a0, float a1) {
  asm("{.reg .b64 ra, rb;\n"
      "mov.b64 ra, {%0, %1}; mov.b64 rb, {%2, %3};\n"
      "add.ftz.f32x2 ra, ra, rb;\n"
      "mov.b64 {%0, %1}, ra;}" : "+f"(d0), "+f"(d1) : "f"(a0), "f"(a1));
}
// (a0, a1) * (b0, b1) + (c, c)
__device__ __forceinline__ void fma2v(float& d0, float& d1, float a0, float a1, float b0, float b1, float c) {
  asm("{.reg .b64 ra, rb, rc;\n"
      "mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5}; mov.b64 rc, {%6, %6};\n"
      "fma.rn.ftz.f32x2 ra, ra, rb, rc;\n"
      "mov.b64 {%0, %1}, ra;}" : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b0), "f"(b1), "f"(c));
}
// exp2 of two arguments <= 0 WITHOUT the exponential unit: the MUFU pipe (16 results/clk/SM) is what binds the
// d = 4 / 8 attention, while the FMA / ALU pipes idle.  Cody-Waite split x = n + f with f in [-0.5, 0.5] (magic-number
// rounding), degree-3 minimax 2^f (max relative error 1.0e-4, far under the bf16 rounding of the probabilities),
// exponent added through the integer ALU.  ~5 issue slots per element in packed form against 8 MUFU clocks.
__device__ __forceinline__ void ex2_poly2(float& p0, float& p1, float x0, float x1) {
  const float MAGIC = 12582912.f;                 // 1.5 * 2^23: x + MAGIC rounds x to the nearest integer
  x0 = fmaxf(x0, -125.f);
  x1 = fmaxf(x1, -125.f);
  float t0, t1, n0, n1, f0, f1;
  fma2(t0, t1, x0, x1, 1.f, MAGIC);               // t = x + MAGIC
  sub2(n0, n1, t0, t1, MAGIC);                    // n = round(x)
  n0 = x0 - n0; n1 = x1 - n1;                     // f = x - n   (kept scalar: two independent FADDs)
  f0 = n0; f1 = n1;
  float q0, q1;
  fma2(q0, q1, f0, f1, 0.05592203512787819f, 0.24264007806777954f);
  fma2v(q0, q1, q0, q1, f0, f1, 0.6931210160255432f);
  fma2v(q0, q1, q0, q1, f0, f1, 0.9999244809150696f);
  p0 = __int_as_float(__float_as_int(q0) + (__float_as_int(t0) << 23));
  p1 = __int_as_float(__float_as_int(q1) + (__float_as_int(t1) << 23));
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

template <int DP>
__device__ __forceinline__ int tswz(int r, int c) {
  constexpr int CPR = DP / 8;
  if (CPR >= 8) return (r * CPR + (c ^ (r & 7))) * 16;
  if (CPR == 4) return (r * 4 + (c ^ ((r >> 1) & 3))) * 16;
  if (CPR == 2) return (r * 2 + (c ^ ((r >> 2) & 1))) * 16;
  return r * 16;
}

// Stage rows [row0, row0+TILE) of a [N, ld] tensor (head offset already applied
// to `base`) into a swizzled [TILE][DP] tile; rows >= N and columns >= D are zero.
// ONES (D == 4 only): column 4 of every in-range row is set to 1.0, so that P V also yields the row sum of P
// (in fp32, from the same bf16-rounded probabilities) in output column 4 -- no FADD per score element.
template <int D, int DP, bool ONES = false>
__device__ __forceinline__ void stage_tile(uint8_t* tile, const bf16* base, int ld, int row0, int N, int tid, int nthreads) {
  if (D >= 8) {
    constexpr int CPR = DP / 8;
    for (int id = tid; id < TILE * CPR; id += nthreads) {
      int r = id / CPR, c = id - r * CPR;
      bool ok = (row0 + r) < N;
      const bf16* src = base + (long long)(ok ? row0 + r : 0) * ld + c * 8;
      cp_async16(smem_u32(tile + tswz<DP>(r, c)), src, ok);
    }
  } else {  // D == 4: 8-byte rows, pad to 16 bytes through registers
    for (int r = tid; r < TILE; r += nthreads) {
      uint2 v = make_uint2(0u, 0u);
      const bool ok = row0 + r < N;
      if (ok) v = *reinterpret_cast<const uint2*>(base + (long long)(row0 + r) * ld);
      *reinterpret_cast<uint4*>(tile + r * 16) = make_uint4(v.x, v.y, (ONES && ok) ? 0x00003F80u : 0u, 0u);
    }
  }
}

// A-operand fragments (16 rows x DP) straight from global memory.
template <int D, int DP>
struct AFrag {
  static constexpr int KS = DP >= 16 ? DP / 16 : 1;
  uint32_t r[KS][4];
  __device__ __forceinline__ void load(const bf16* base, int ld, int row0, int N, int lane) {
    const int g = lane >> 2, t = lane & 3;
    const bool ok0 = (row0 + g) < N, ok1 = (row0 + g + 8) < N;
    const bf16* p0 = base + (long long)(ok0 ? row0 + g : 0) * ld;
    const bf16* p1 = base + (long long)(ok1 ? row0 + g + 8 : 0) * ld;
    if (DP >= 16) {
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        r[ks][0] = ok0 ? *reinterpret_cast<const uint32_t*>(p0 + ks * 16 + 2 * t) : 0u;
        r[ks][1] = ok1 ? *reinterpret_cast<const uint32_t*>(p1 + ks * 16 + 2 * t) : 0u;
        r[ks][2] = ok0 ? *reinterpret_cast<const uint32_t*>(p0 + ks * 16 + 8 + 2 * t) : 0u;
        r[ks][3] = ok1 ? *reinterpret_cast<const uint32_t*>(p1 + ks * 16 + 8 + 2 * t) : 0u;
      }
    } else {
      bool colok = (2 * t) < D;
      r[0][0] = (ok0 && colok) ? *reinterpret_cast<const uint32_t*>(p0 + 2 * t) : 0u;
      r[0][1] = (ok1 && colok) ? *reinterpret_cast<const uint32_t*>(p1 + 2 * t) : 0u;
      r[0][2] = r[0][3] = 0u;
    }
  }
};

// c(16x8) += A(16xDP) * T[ntile*8 .. +8][0..DP)^T   (contraction over the head dim)
template <int D, int DP>
__device__ __forceinline__ void mma_over_d(float* c, const AFrag<D, DP>& a, const uint8_t* tile, int ntile, int lane) {
  if (DP == 8) {
    const int g = lane >> 2, t = lane & 3;
    uint32_t b0 = *reinterpret_cast<const uint32_t*>(tile + (ntile * 8 + g) * 16 + t * 4);
    mma_1688(c, a.r[0][0], a.r[0][1], b0);
  } else {
#pragma unroll
    for (int ks = 0; ks < AFrag<D, DP>::KS; ++ks) {
      uint32_t b0, b1;
      int row = ntile * 8 + (lane & 7);
      int ch = 2 * ks + ((lane >> 3) & 1);
      ldsm_x2(b0, b1, smem_u32(tile + tswz<DP>(row, ch)));
      mma_16816(c, a.r[ks], b0, b1);
    }
  }
}

// o(16 x DP, as DP/8 n-tiles) += P(16 x 16, k16 step kk of the tile rows) * T[kk*16.., :]
template <int DP>
__device__ __forceinline__ void mma_over_rows(float (*o)[4], const uint32_t* pa, const uint8_t* tile, int kk, int lane) {
#pragma unroll
  for (int j = 0; j < DP / 8; ++j) {
    uint32_t b0, b1;
    int row = kk * 16 + (lane & 15);
    ldsm_x2_t(b0, b1, smem_u32(tile + tswz<DP>(row, j)));
    mma_16816(o[j], pa, b0, b1);
  }
}

}  // namespace attn
