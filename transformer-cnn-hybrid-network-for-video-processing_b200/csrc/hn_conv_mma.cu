// Generic implicit-GEMM convolution on NHWC bf16 (fwd / dgrad / transposed) and
// its weight gradient.  Warp-level mma.sync path: this is the kernel for the
// HBM-bound small-channel layers (C <= 32: every FCT / AutoEncoder conv, UNet
// level 1), for strided / dilated / 7x7 / 1x1 geometries, and the correctness
// baseline for the tcgen05 kernel (hn_conv_tc.cu) that takes the C >= 64 3x3s.
//
// GEMM view:  M = B*OH*OW output pixels,  N = Cout,  K = KH*KW*(C1+C2)
// K is ordered (tap, channel) in groups of 8 channels (one 16-byte vector).
#include "hn_common.cuh"

namespace {

constexpr int BK = 32;       // K elements per pipeline stage (4 groups of 8)
constexpr int STAGES = 4;
constexpr int NTHREADS = 256;

struct ConvP {
  HnConv d;
  const bf16* x1;
  const bf16* x2;
  const bf16* w;
  const float* bias;
  const bf16* res;
  bf16* y;
  float* ssum;
  float* ssq;
  int M;        // B*OH*OW
  int CG;       // (C1+C2)/8
  int G;        // KH*KW*CG  (number of 8-channel K groups)
  int KB;       // ceil(G/4)
  int Hs, Ws;   // physical source extent (H>>ups, W>>ups)
};

// swizzled byte offset of 16B chunk `c` in row `r` of a tile with CPR chunks/row
template <int CPR>
__device__ __forceinline__ int swz(int r, int c) {
  if (CPR >= 8) return (r * CPR + (c ^ (r & 7))) * 16;
  if (CPR == 4) return (r * 4 + (c ^ ((r >> 1) & 3))) * 16;
  if (CPR == 2) return (r * 2 + (c ^ ((r >> 2) & 1))) * 16;
  return (r * CPR + c) * 16;
}

// source address of 8 channels for (pixel b,oh,ow ; K-group g) or nullptr if padding
__device__ __forceinline__ const bf16* gather_src(const ConvP& p, int b, int oh, int ow, int tap, int c8) {
  const HnConv& d = p.d;
  int kh = tap / d.KW, kw = tap - kh * d.KW;
  int ih, iw;
  if (d.mode == 0) {
    ih = oh * d.stride - d.pad + kh * d.dil;
    iw = ow * d.stride - d.pad + kw * d.dil;
  } else {
    ih = oh + d.pad - kh * d.dil;
    iw = ow + d.pad - kw * d.dil;
    if (ih < 0 || iw < 0) return nullptr;
    if (d.stride > 1) {
      if ((ih % d.stride) | (iw % d.stride)) return nullptr;
      ih /= d.stride;
      iw /= d.stride;
    }
  }
  if ((unsigned)ih >= (unsigned)d.H || (unsigned)iw >= (unsigned)d.W) return nullptr;
  ih >>= d.ups;
  iw >>= d.ups;
  long long pix = ((long long)b * p.Hs + ih) * p.Ws + iw;
  int c = c8 * 8;
  if (c < d.C1) return p.x1 + pix * d.ld1 + c;
  return p.x2 + pix * d.ld2 + (c - d.C1);
}

template <int BM, int BN, int WM, int WN>
__global__ void __launch_bounds__(NTHREADS) conv_fwd_kernel(const ConvP p) {
  constexpr int MI = BM / WM / 16;
  constexpr int NI = BN / WN / 8;
  constexpr int A_BYTES = BM * BK * 2;
  constexpr int B_BYTES = BN * BK * 2;
  constexpr int A_CH = BM * 4 / NTHREADS;          // 16B chunks per thread (A)
  static_assert(A_CH >= 1, "tile too small");
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ float s_stat[2][BN];
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_BYTES;

  const HnConv& d = p.d;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp / WN, wn = warp % WN;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;

  // per-thread A rows (fixed over K): decode pixel once
  int a_row[A_CH], a_kc[A_CH], a_b[A_CH], a_oh[A_CH], a_ow[A_CH];
  bool a_ok[A_CH];
#pragma unroll
  for (int i = 0; i < A_CH; ++i) {
    int id = tid + i * NTHREADS;
    a_row[i] = id >> 2;
    a_kc[i] = id & 3;
    int m = m0 + a_row[i];
    a_ok[i] = m < p.M;
    int mm = a_ok[i] ? m : 0;
    a_b[i] = mm / (d.OH * d.OW);
    int r = mm - a_b[i] * (d.OH * d.OW);
    a_oh[i] = r / d.OW;
    a_ow[i] = r - a_oh[i] * d.OW;
  }
  // running (tap, c8) of this thread's K group: g = kb*4 + kc
  int a_tap[A_CH], a_c8[A_CH];
#pragma unroll
  for (int i = 0; i < A_CH; ++i) {
    a_tap[i] = a_kc[i] / p.CG;
    a_c8[i] = a_kc[i] - a_tap[i] * p.CG;
  }
  const int KT = d.KH * d.KW;

  auto load_stage = [&](int stage, int kb) {
#pragma unroll
    for (int i = 0; i < A_CH; ++i) {
      const bf16* src = nullptr;
      if (a_ok[i] && a_tap[i] < KT) src = gather_src(p, a_b[i], a_oh[i], a_ow[i], a_tap[i], a_c8[i]);
      uint32_t dst = smem_u32(sA + stage * A_BYTES + swz<4>(a_row[i], a_kc[i]));
      cp_async16(dst, src ? (const void*)src : (const void*)p.x1, src != nullptr);
      a_c8[i] += 4;
      while (a_c8[i] >= p.CG) { a_c8[i] -= p.CG; a_tap[i]++; }
    }
    for (int id = tid; id < BN * 4; id += NTHREADS) {
      int row = id >> 2, kc = id & 3;
      const bf16* src = p.w + (long long)(n0 + row) * d.Kpad + kb * BK + kc * 8;
      cp_async16(smem_u32(sB + stage * B_BYTES + swz<4>(row, kc)), src, true);
    }
  };

  float acc[MI][NI][4];
#pragma unroll
  for (int i = 0; i < MI; ++i)
#pragma unroll
    for (int j = 0; j < NI; ++j)
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[i][j][k] = 0.f;

  // prologue
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (s < p.KB) load_stage(s, s);
    cp_async_commit();
  }
  for (int kb = 0; kb < p.KB; ++kb) {
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    {
      int nk = kb + STAGES - 1;
      if (nk < p.KB) load_stage(nk % STAGES, nk);
      cp_async_commit();
    }
    const uint8_t* tA = sA + (kb % STAGES) * A_BYTES;
    const uint8_t* tB = sB + (kb % STAGES) * B_BYTES;
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
      uint32_t af[MI][4];
#pragma unroll
      for (int i = 0; i < MI; ++i) {
        int row = wm * (MI * 16) + i * 16 + (lane & 15);
        int ch = ks * 2 + (lane >> 4);
        ldsm_x4(af[i][0], af[i][1], af[i][2], af[i][3], smem_u32(tA + swz<4>(row, ch)));
      }
      uint32_t bfr[NI][2];
      if (NI >= 2) {
#pragma unroll
        for (int j = 0; j < NI; j += 2) {
          int row = wn * (NI * 8) + j * 8 + (lane & 7) + ((lane >> 4) << 3);
          int ch = ks * 2 + ((lane >> 3) & 1);
          ldsm_x4(bfr[j][0], bfr[j][1], bfr[j + 1 < NI ? j + 1 : j][0], bfr[j + 1 < NI ? j + 1 : j][1],
                  smem_u32(tB + swz<4>(row, ch)));
        }
      } else {
        int row = wn * 8 + (lane & 7);
        int ch = ks * 2 + ((lane >> 3) & 1);
        ldsm_x2(bfr[0][0], bfr[0][1], smem_u32(tB + swz<4>(row, ch)));
      }
#pragma unroll
      for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NI; ++j) mma_16816(acc[i][j], af[i], bfr[j][0], bfr[j][1]);
    }
  }
  cp_async_wait<0>();

  // ---- epilogue ---------------------------------------------------------------
  const bool want_stats = p.ssum != nullptr;
  if (want_stats) {
    for (int i = tid; i < 2 * BN; i += NTHREADS) (&s_stat[0][0])[i] = 0.f;
    __syncthreads();
  }
  const int g = lane >> 2, t = lane & 3;
  float csum[NI][2], csq[NI][2];
#pragma unroll
  for (int j = 0; j < NI; ++j) { csum[j][0] = csum[j][1] = csq[j][0] = csq[j][1] = 0.f; }

#pragma unroll
  for (int i = 0; i < MI; ++i) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      int m = m0 + wm * (MI * 16) + i * 16 + g + h * 8;
      if (m >= p.M) continue;
      long long opix = m;
      int ob = 0, ooh = 0, oow = 0;
      if (d.pshuf) {
        ob = m / (d.OH * d.OW);
        int r = m - ob * (d.OH * d.OW);
        ooh = r / d.OW;
        oow = r - ooh * d.OW;
      }
#pragma unroll
      for (int j = 0; j < NI; ++j) {
        int n = n0 + wn * (NI * 8) + j * 8 + t * 2;
        float v0 = acc[i][j][h * 2 + 0], v1 = acc[i][j][h * 2 + 1];
        int ntot = d.pshuf ? 4 * d.pshuf : d.Cout;
        if (n >= ntot) continue;
        bool has1 = (n + 1) < ntot;
        int c0 = n, c1 = n + 1;
        long long pix0 = opix, pix1 = opix;
        if (d.pshuf) {
          int tap0 = n / d.pshuf;
          c0 = n - tap0 * d.pshuf;
          pix0 = ((long long)ob * (2 * d.OH) + 2 * ooh + (tap0 >> 1)) * (2 * d.OW) + 2 * oow + (tap0 & 1);
          int tap1 = (n + 1) / d.pshuf;
          c1 = (n + 1) - tap1 * d.pshuf;
          pix1 = ((long long)ob * (2 * d.OH) + 2 * ooh + (tap1 >> 1)) * (2 * d.OW) + 2 * oow + (tap1 & 1);
        }
        if (p.bias) { v0 += p.bias[c0]; if (has1) v1 += p.bias[c1]; }
        if (p.res) {
          v0 += __bfloat162float(p.res[pix0 * d.ldr + c0]);
          if (has1) v1 += __bfloat162float(p.res[pix1 * d.ldr + c1]);
        }
        if (want_stats) {
          float r0 = bf16_round(v0), r1 = has1 ? bf16_round(v1) : 0.f;
          csum[j][0] += r0; csq[j][0] += r0 * r0;
          csum[j][1] += r1; csq[j][1] += r1 * r1;
        }
        v0 = apply_act(v0, d.act);
        v1 = apply_act(v1, d.act);
        bf16* o0 = p.y + pix0 * d.ldy + c0;
        if (has1 && pix1 == pix0 && c1 == c0 + 1 && ((c0 & 1) == 0)) {
          *reinterpret_cast<uint32_t*>(o0) = pack_bf16(v0, v1);
        } else {
          *o0 = __float2bfloat16(v0);
          if (has1) p.y[pix1 * d.ldy + c1] = __float2bfloat16(v1);
        }
      }
    }
  }
  if (want_stats) {
#pragma unroll
    for (int j = 0; j < NI; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        float s = csum[j][e], q = csq[j][e];
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
          s += __shfl_xor_sync(0xffffffffu, s, o);
          q += __shfl_xor_sync(0xffffffffu, q, o);
        }
        if (g == 0) {
          int nl = wn * (NI * 8) + j * 8 + t * 2 + e;
          atomicAdd(&s_stat[0][nl], s);
          atomicAdd(&s_stat[1][nl], q);
        }
      }
    __syncthreads();
    for (int nl = tid; nl < BN; nl += NTHREADS) {
      int n = n0 + nl;
      int ntot = d.pshuf ? 4 * d.pshuf : d.Cout;
      if (n < ntot) {
        int c = d.pshuf ? (n % d.pshuf) : n;
        atomicAdd(p.ssum + c, s_stat[0][nl]);
        atomicAdd(p.ssq + c, s_stat[1][nl]);
      }
    }
  }
}

template <int BM, int BN, int WM, int WN>
int launch_fwd(const ConvP& p, int ntot, cudaStream_t st) {
  constexpr int SMEM = STAGES * (BM + BN) * BK * 2;
  static bool configured = false;
  if (!configured) {
    cudaFuncSetAttribute(conv_fwd_kernel<BM, BN, WM, WN>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    configured = true;
  }
  dim3 grid(hn_cdiv(p.M, BM), hn_cdiv(ntot, BN));
  conv_fwd_kernel<BM, BN, WM, WN><<<grid, NTHREADS, SMEM, st>>>(p);
  HN_CHECK_LAUNCH("conv_fwd_kernel");
  return HN_OK;
}

}  // namespace

extern "C" int hn_conv_kpad(int KH, int KW, int Cin_pad) {
  int k = KH * KW * Cin_pad;
  return (k + BK - 1) / BK * BK;
}
extern "C" int hn_conv_npad(int Cout) { return (Cout + 127) / 128 * 128; }

static int conv_validate(const HnConv* d, const void* x2) {
  HN_REQUIRE(d != nullptr, "conv: null descriptor");
  HN_REQUIRE(d->C1 > 0 && d->C1 % 8 == 0 && d->C2 % 8 == 0 && d->C2 >= 0, "conv: C1/C2 must be multiples of 8 (got %d,%d)", d->C1, d->C2);
  HN_REQUIRE(d->C2 == 0 || x2 != nullptr, "conv: C2>0 needs x2");
  HN_REQUIRE(d->ld1 % 8 == 0 && d->ld1 >= d->C1 && (d->C2 == 0 || (d->ld2 % 8 == 0 && d->ld2 >= d->C2)), "conv: bad input ld");
  HN_REQUIRE(d->B > 0 && d->H > 0 && d->W > 0 && d->OH > 0 && d->OW > 0 && d->Cout > 0, "conv: bad extent");
  HN_REQUIRE(d->KH > 0 && d->KW > 0 && d->stride >= 1 && d->dil >= 1 && d->pad >= 0, "conv: bad kernel geometry");
  HN_REQUIRE(d->ups == 0 || d->ups == 1, "conv: ups must be 0/1");
  HN_REQUIRE(d->ups == 0 || (d->H % 2 == 0 && d->W % 2 == 0), "conv: upsampled extent must be even");
  HN_REQUIRE(d->Kpad == hn_conv_kpad(d->KH, d->KW, d->C1 + d->C2), "conv: Kpad mismatch");
  HN_REQUIRE((long long)d->B * d->OH * d->OW < (1ll << 31), "conv: too many pixels");
  return HN_OK;
}

extern "C" int hn_conv2d_fwd(const HnConv* d, const void* x1, const void* x2, const void* wpacked,
                             const float* bias, const void* res, void* y, float* stat_sum,
                             float* stat_sqsum, void* stream) {
  int rc = conv_validate(d, x2);
  if (rc) return rc;
  HN_REQUIRE(x1 && wpacked && y, "conv: null pointer");
  HN_REQUIRE((stat_sum == nullptr) == (stat_sqsum == nullptr), "conv: stats need both buffers");
  ConvP p;
  p.d = *d;
  p.x1 = (const bf16*)x1; p.x2 = (const bf16*)x2; p.w = (const bf16*)wpacked;
  p.bias = bias; p.res = (const bf16*)res; p.y = (bf16*)y; p.ssum = stat_sum; p.ssq = stat_sqsum;
  p.M = d->B * d->OH * d->OW;
  p.CG = (d->C1 + d->C2) / 8;
  p.G = d->KH * d->KW * p.CG;
  p.KB = (p.G + 3) / 4;
  p.Hs = d->H >> d->ups; p.Ws = d->W >> d->ups;
  int ntot = d->pshuf ? 4 * d->pshuf : d->Cout;
  cudaStream_t st = (cudaStream_t)stream;
  if (ntot <= 8) return launch_fwd<128, 8, 8, 1>(p, ntot, st);
  if (ntot <= 16) return launch_fwd<128, 16, 8, 1>(p, ntot, st);
  if (ntot <= 32) return launch_fwd<128, 32, 8, 1>(p, ntot, st);
  if (ntot <= 64) return launch_fwd<128, 64, 4, 2>(p, ntot, st);
  return launch_fwd<128, 128, 2, 4>(p, ntot, st);
}

// ---- weight packing -------------------------------------------------------------
__global__ void pack_weight_kernel(const float* __restrict__ src, bf16* __restrict__ dst, int N, int Npad,
                                   int Creal, int Cpad, int KH, int KW, int Kpad, long long nmod,
                                   long long s_nhi, long long s_n, long long s_c, long long s_kh, long long s_kw) {
  long long total = (long long)Npad * Kpad;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int n = (int)(i / Kpad), k = (int)(i - (long long)n * Kpad);
    int tap = k / Cpad, c = k - tap * Cpad;
    float v = 0.f;
    if (n < N && tap < KH * KW && c < Creal) {
      int kh = tap / KW, kw = tap - kh * KW;
      v = src[(n / nmod) * s_nhi + (n % nmod) * s_n + c * s_c + kh * s_kh + kw * s_kw];
    }
    dst[i] = __float2bfloat16(v);
  }
}

// Input gradient of a stride-2 conv (k3 p1 or k1 p0) as ONE stride-1 GEMM over dz with a pixel-shuffle store:
// dx[2h+py, 2w+px, ci] = sum_{dh,dw,co} dz[h+dh, w+dw, co] * W[co, ci, kh(py,dh), kw(px,dw)],
// kh(0,0)=1, kh(0,1)=none, kh(1,0)=2, kh(1,1)=0 (k3); k1: only (py,px)=(0,0), no window.
// Packed as B[(py,px,ci)][(dh,dw,co)] -- 9 of the 16 (parity, tap) blocks are non-zero.
__global__ void pack_weight_s2dgrad_kernel(const float* __restrict__ src, bf16* __restrict__ dst, int Cout, int Cin,
                                           int K, int Npad, int Kpad) {
  const long long total = (long long)Npad * Kpad;
  const int win = K == 3 ? 2 : 1;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i / Kpad), k = (int)(i - (long long)n * Kpad);
    float v = 0.f;
    if (n < 4 * Cin && k < win * win * Cout) {
      const int par = n / Cin, ci = n - par * Cin, py = par >> 1, px = par & 1;
      const int tap = k / Cout, co = k - tap * Cout, dh = tap / win, dw = tap - dh * win;
      int kh, kw;
      if (K == 3) {
        kh = py == 0 ? (dh == 0 ? 1 : -1) : (dh == 0 ? 2 : 0);
        kw = px == 0 ? (dw == 0 ? 1 : -1) : (dw == 0 ? 2 : 0);
      } else {
        kh = py == 0 ? 0 : -1;
        kw = px == 0 ? 0 : -1;
      }
      if (kh >= 0 && kw >= 0) v = src[(((long long)co * Cin + ci) * K + kh) * K + kw];
    }
    dst[i] = __float2bfloat16(v);
  }
}

extern "C" int hn_pack_weight_s2dgrad(const float* src, void* dst, int Cout, int Cin, int K, void* stream) {
  HN_REQUIRE(src && dst && Cout > 0 && Cin > 0 && (K == 1 || K == 3) && Cout % 8 == 0, "pack_weight_s2dgrad: bad args");
  const int win = K == 3 ? 2 : 1;
  const int Kpad = hn_conv_kpad(win, win, Cout), Npad = hn_conv_npad(4 * Cin);
  const long long total = (long long)Npad * Kpad;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  pack_weight_s2dgrad_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(src, (bf16*)dst, Cout, Cin, K, Npad, Kpad);
  HN_CHECK_LAUNCH("pack_weight_s2dgrad_kernel");
  return HN_OK;
}

// ---- all weight packs of a model in ONE launch ------------------------------------------------------------------
// jobs: device array of HnPackJob (16 x i64 each, see include/hybridnet.h); blockIdx.y = job.  A training step
// repacks every conv weight after the optimizer update: 45-214 launches of 2-5 us each otherwise.
struct PackJob {
  long long src, dst, N, Npad, Creal, Cpad, KH, KW, Kpad, nmod, s_nhi, s_n, s_c, s_kh, s_kw, s2k;
};
__global__ void __launch_bounds__(256) pack_weight_multi_kernel(const PackJob* __restrict__ jobs) {
  const PackJob j = jobs[blockIdx.y];
  const float* __restrict__ src = reinterpret_cast<const float*>(j.src);
  bf16* __restrict__ dst = reinterpret_cast<bf16*>(j.dst);
  const long long total = j.Npad * j.Kpad;
  const int Kpad = (int)j.Kpad;
  if (j.s2k) {                       // stride-2 input-gradient layout (pack_weight_s2dgrad_kernel): N = Cout, Creal = Cin
    const int Cout = (int)j.N, Cin = (int)j.Creal, K = (int)j.s2k, win = K == 3 ? 2 : 1;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
      const int n = (int)(i / Kpad), k = (int)(i - (long long)n * Kpad);
      float v = 0.f;
      if (n < 4 * Cin && k < win * win * Cout) {
        const int par = n / Cin, ci = n - par * Cin, py = par >> 1, px = par & 1;
        const int tap = k / Cout, co = k - tap * Cout, dh = tap / win, dw = tap - dh * win;
        int kh, kw;
        if (K == 3) {
          kh = py == 0 ? (dh == 0 ? 1 : -1) : (dh == 0 ? 2 : 0);
          kw = px == 0 ? (dw == 0 ? 1 : -1) : (dw == 0 ? 2 : 0);
        } else {
          kh = py == 0 ? 0 : -1;
          kw = px == 0 ? 0 : -1;
        }
        if (kh >= 0 && kw >= 0) v = src[(((long long)co * Cin + ci) * K + kh) * K + kw];
      }
      dst[i] = __float2bfloat16(v);
    }
    return;
  }
  const int N = (int)j.N, Creal = (int)j.Creal, Cpad = (int)j.Cpad, KH = (int)j.KH, KW = (int)j.KW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i / Kpad), k = (int)(i - (long long)n * Kpad);
    const int tap = k / Cpad, c = k - tap * Cpad;
    float v = 0.f;
    if (n < N && tap < KH * KW && c < Creal) {
      const int kh = tap / KW, kw = tap - kh * KW;
      v = src[(n / j.nmod) * j.s_nhi + (n % j.nmod) * j.s_n + c * j.s_c + kh * j.s_kh + kw * j.s_kw];
    }
    dst[i] = __float2bfloat16(v);
  }
}

extern "C" int hn_pack_weight_multi(const void* jobs, int njobs, long long max_elems, void* stream) {
  HN_REQUIRE(jobs && njobs > 0 && njobs <= 65535 && max_elems > 0, "pack_weight_multi: bad args");
  // blocks per job sized for the LARGEST job (2 elements per thread, at most four waves in total); the blocks a
  // small job does not need fall straight through their grid-stride loop
  long long bx = (max_elems + 511) / 512;
  const long long cap = (148LL * 8 * 4 + njobs - 1) / njobs;
  if (bx > cap) bx = cap;
  if (bx < 8) bx = 8;
  pack_weight_multi_kernel<<<dim3((unsigned)bx, njobs), 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const PackJob*>(jobs));
  HN_CHECK_LAUNCH("pack_weight_multi_kernel");
  return HN_OK;
}

extern "C" int hn_pack_weight(const float* src, void* dst, int N, int Creal, int Cpad, int KH, int KW,
                              long long nmod, long long s_nhi, long long s_n, long long s_c,
                              long long s_kh, long long s_kw, void* stream) {
  HN_REQUIRE(src && dst && N > 0 && Creal > 0 && Cpad % 8 == 0 && Creal <= Cpad && nmod > 0, "pack_weight: bad args");
  int Kpad = hn_conv_kpad(KH, KW, Cpad), Npad = hn_conv_npad(N);
  long long total = (long long)Npad * Kpad;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  pack_weight_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(src, (bf16*)dst, N, Npad, Creal, Cpad, KH, KW,
                                                              Kpad, nmod, s_nhi, s_n, s_c, s_kh, s_kw);
  HN_CHECK_LAUNCH("pack_weight_kernel");
  return HN_OK;
}
