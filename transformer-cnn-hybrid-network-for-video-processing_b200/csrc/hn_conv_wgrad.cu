// Weight gradient of the implicit-GEMM convolution (mma.sync path).
//   dW[n][(tap,c)] = sum_p dy[p][n] * x[p@tap][c]
// GEMM with M = Cout (n), N = KH*KW*(C1+C2) in 8-channel groups, reduction over
// the B*OH*OW output pixels, split across CTAs (split-K) and accumulated with
// fp32 atomics straight into the parameter-gradient tensor (OIHW strides).
// For the small-channel layers this is a bandwidth-bound reduction (H4 in
// SURVEY.md 7.3): each CTA streams dy and the gathered x once.
#include "hn_common.cuh"

namespace {

constexpr int PK = 32;        // pixels per pipeline stage (the GEMM K step)
constexpr int TN = 64;        // 8 groups of 8 channels per CTA
constexpr int WSTAGES = 3;
constexpr int NT = 256;

struct WgP {
  HnConv d;
  const bf16* x1;
  const bf16* x2;
  const bf16* dy;
  float* dw;
  int Creal;
  long long s_n, s_c, s_kh, s_kw;
  int M, CG, G, Hs, Ws;
  int chunk;     // pixels per split (multiple of PK)
};

template <int CPR>
__device__ __forceinline__ int swzw(int r, int c) {
  if (CPR >= 8) return (r * CPR + (c ^ (r & 7))) * 16;
  if (CPR == 4) return (r * 4 + (c ^ ((r >> 1) & 3))) * 16;
  if (CPR == 2) return (r * 2 + (c ^ ((r >> 2) & 1))) * 16;
  return (r * CPR + c) * 16;
}

__device__ __forceinline__ const bf16* wg_src(const WgP& p, int b, int oh, int ow, int tap, int c8) {
  const HnConv& d = p.d;
  int kh = tap / d.KW, kw = tap - kh * d.KW;
  int ih = oh * d.stride - d.pad + kh * d.dil;
  int iw = ow * d.stride - d.pad + kw * d.dil;
  if ((unsigned)ih >= (unsigned)d.H || (unsigned)iw >= (unsigned)d.W) return nullptr;
  ih >>= d.ups;
  iw >>= d.ups;
  long long pix = ((long long)b * p.Hs + ih) * p.Ws + iw;
  int c = c8 * 8;
  if (c < d.C1) return p.x1 + pix * d.ld1 + c;
  return p.x2 + pix * d.ld2 + (c - d.C1);
}

// TM = rows (output channels) per CTA: 16, 32 or 64
template <int TM>
__global__ void __launch_bounds__(NT) conv_wgrad_kernel(const WgP p) {
  constexpr int WMW = TM / 16;            // warps along m
  constexpr int WNW = 8 / WMW;            // warps along n
  constexpr int NI = TN / WNW / 8;        // n8 tiles per warp
  constexpr int CPR_A = TM / 8;           // 16B chunks per dy-tile row
  constexpr int A_BYTES = PK * TM * 2;
  constexpr int B_BYTES = PK * TN * 2;
  __shared__ __align__(128) uint8_t sA[WSTAGES * A_BYTES];
  __shared__ __align__(128) uint8_t sB[WSTAGES * B_BYTES];

  const HnConv& d = p.d;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp / WNW, wn = warp % WNW;
  const int n0 = blockIdx.x * TM;          // output-channel tile
  const int g0 = blockIdx.y * 8;           // K-group tile
  const int pbeg = blockIdx.z * p.chunk;
  const int pend = min(p.M, pbeg + p.chunk);
  const int nsteps = (pend - pbeg + PK - 1) / PK;

  // B loader: one 16B chunk per thread per stage: row = pixel, col chunk = group
  const int b_row = tid >> 3, b_g = tid & 7;
  const int gg = g0 + b_g;
  const bool g_ok = gg < p.G;
  const int b_tap = g_ok ? gg / p.CG : 0;
  const int b_c8 = g_ok ? gg - b_tap * p.CG : 0;

  auto load_stage = [&](int stage, int step) {
    int pix0 = pbeg + step * PK;
    {  // x gather
      int m = pix0 + b_row;
      const bf16* src = nullptr;
      if (g_ok && m < pend) {
        int b = m / (d.OH * d.OW);
        int r = m - b * (d.OH * d.OW);
        int oh = r / d.OW, ow = r - oh * d.OW;
        src = wg_src(p, b, oh, ow, b_tap, b_c8);
      }
      cp_async16(smem_u32(sB + stage * B_BYTES + swzw<8>(b_row, b_g)), src ? (const void*)src : (const void*)p.x1,
                 src != nullptr);
    }
    for (int id = tid; id < PK * CPR_A; id += NT) {  // dy tile [PK][TM]
      int row = id / CPR_A, ch = id - row * CPR_A;
      int m = pix0 + row;
      int n = n0 + ch * 8;
      bool ok = (m < pend) && (n < ((d.Cout + 7) & ~7));   // rows n >= Cout are discarded in the epilogue
      const bf16* src = p.dy + (long long)m * d.ldy + n;
      cp_async16(smem_u32(sA + stage * A_BYTES + swzw<CPR_A>(row, ch)), ok ? (const void*)src : (const void*)p.dy, ok);
    }
  };

  float acc[NI][4];
#pragma unroll
  for (int j = 0; j < NI; ++j)
#pragma unroll
    for (int k = 0; k < 4; ++k) acc[j][k] = 0.f;

#pragma unroll
  for (int s = 0; s < WSTAGES - 1; ++s) {
    if (s < nsteps) load_stage(s, s);
    cp_async_commit();
  }
  for (int st = 0; st < nsteps; ++st) {
    cp_async_wait<WSTAGES - 2>();
    __syncthreads();
    {
      int ns = st + WSTAGES - 1;
      if (ns < nsteps) load_stage(ns % WSTAGES, ns);
      cp_async_commit();
    }
    const uint8_t* tA = sA + (st % WSTAGES) * A_BYTES;
    const uint8_t* tB = sB + (st % WSTAGES) * B_BYTES;
#pragma unroll
    for (int ks = 0; ks < PK / 16; ++ks) {
      uint32_t af[4];
      {  // A = dy^T: stored [k=pixel][m=channel] -> transposed ldmatrix
        int i = lane >> 3;
        int krow = ks * 16 + (lane & 7) + ((i >> 1) << 3);
        int mch = wm * 2 + (i & 1);
        ldsm_x4_t(af[0], af[1], af[2], af[3], smem_u32(tA + swzw<CPR_A>(krow, mch)));
      }
#pragma unroll
      for (int j = 0; j < NI; j += 2) {
        uint32_t b0, b1, b2, b3;
        int krow = ks * 16 + (lane & 7) + (((lane >> 3) & 1) << 3);
        int nch = wn * NI + j + (lane >> 4);
        if (NI >= 2) {
          ldsm_x4_t(b0, b1, b2, b3, smem_u32(tB + swzw<8>(krow, nch)));
          mma_16816(acc[j], af, b0, b1);
          mma_16816(acc[j + 1 < NI ? j + 1 : j], af, b2, b3);
        } else {
          nch = wn * NI + j;
          ldsm_x2_t(b0, b1, smem_u32(tB + swzw<8>(krow, nch)));
          mma_16816(acc[j], af, b0, b1);
        }
      }
    }
  }
  cp_async_wait<0>();

  // ---- epilogue: fp32 atomics into dW at the parameter's own strides -------------
  const int g = lane >> 2, t = lane & 3;
#pragma unroll
  for (int j = 0; j < NI; ++j) {
    int col = (wn * NI + j) * 8 + t * 2;          // column within the 64-wide tile
    int grp = g0 + (col >> 3);
    if (grp >= p.G) continue;
    int tap = grp / p.CG;
    int c8 = grp - tap * p.CG;
    int kh = tap / d.KW, kw = tap - kh * d.KW;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      int n = n0 + wm * 16 + g + (e >> 1) * 8;
      int c = c8 * 8 + (col & 7) + (e & 1);
      if (n < d.Cout && c < p.Creal) {
        float v = acc[j][e];
        if (v != 0.f) atomicAdd(p.dw + n * p.s_n + c * p.s_c + kh * p.s_kh + kw * p.s_kw, v);
      }
    }
  }
}

}  // namespace

extern "C" int hn_conv2d_wgrad(const HnConv* d, const void* x1, const void* x2, const void* dy,
                               float* dw, int Creal, long long s_n, long long s_c, long long s_kh,
                               long long s_kw, void* stream) {
  HN_REQUIRE(d && x1 && dy && dw, "wgrad: null pointer");
  HN_REQUIRE(d->mode == 0, "wgrad: only forward-gather geometry (mode 0)");
  HN_REQUIRE(d->C1 > 0 && d->C1 % 8 == 0 && d->C2 % 8 == 0, "wgrad: C1/C2 must be multiples of 8");
  HN_REQUIRE(d->C2 == 0 || x2 != nullptr, "wgrad: C2>0 needs x2");
  HN_REQUIRE(d->ldy % 8 == 0 && d->ldy >= d->Cout, "wgrad: dy ld must be a multiple of 8 >= Cout");
  HN_REQUIRE(d->pshuf == 0, "wgrad: pshuf geometry is expressed as the k2s2 conv of its gradient");
  WgP p;
  p.d = *d;
  p.x1 = (const bf16*)x1; p.x2 = (const bf16*)x2; p.dy = (const bf16*)dy; p.dw = dw;
  p.Creal = Creal; p.s_n = s_n; p.s_c = s_c; p.s_kh = s_kh; p.s_kw = s_kw;
  p.M = d->B * d->OH * d->OW;
  p.CG = (d->C1 + d->C2) / 8;
  p.G = d->KH * d->KW * p.CG;
  p.Hs = d->H >> d->ups; p.Ws = d->W >> d->ups;
  int TM = d->Cout <= 16 ? 16 : (d->Cout <= 32 ? 32 : 64);
  int gx = hn_cdiv(d->Cout, TM), gy = hn_cdiv(p.G, 8);
  // split the pixel reduction so the grid is ~4 waves of 148 SMs (2 CTAs/SM resident)
  int target = 148 * 4;
  int splits = target / (gx * gy);
  if (splits < 1) splits = 1;
  int max_splits = hn_cdiv(p.M, 4 * PK);
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  int chunk = hn_cdiv(hn_cdiv(p.M, splits), PK) * PK;
  splits = hn_cdiv(p.M, chunk);
  p.chunk = chunk;
  dim3 grid(gx, gy, splits);
  cudaStream_t st = (cudaStream_t)stream;
  if (TM == 16) conv_wgrad_kernel<16><<<grid, NT, 0, st>>>(p);
  else if (TM == 32) conv_wgrad_kernel<32><<<grid, NT, 0, st>>>(p);
  else conv_wgrad_kernel<64><<<grid, NT, 0, st>>>(p);
  HN_CHECK_LAUNCH("conv_wgrad_kernel");
  return HN_OK;
}
