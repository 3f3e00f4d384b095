// Fused q/k/v projection of the FCT Attention block (FCT.py:41-57):
//   proj_t = LayerNorm_C(ReLU(depthwise_conv3x3(x; w_t) + b_t)),  t in {q,k,v}
// One pass over x produces all three projections (x is read once, 9 taps from
// L1/L2), LayerNorm statistics are reduced with warp shuffles across the C/8
// lanes that share a pixel.  Backward is two kernels: (1) LN+ReLU backward to
// the dwconv output (recomputing the forward from x), (2) depthwise data- and
// weight-gradient.  All HBM/L2-bound; no tensor-core work here by design.
#include "hn_common.cuh"

namespace {
constexpr int TPB = 256;
#ifndef DWLN_MINB
#define DWLN_MINB 2      // resident CTAs per SM the register allocation is held to
#endif

__device__ __forceinline__ float group_sum(float v, int CV) {
  for (int o = 1; o < CV; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// pre[j] = b + sum_tap x[p+off][c] * w[c][tap]   (zero padding).  xp = x at (b, h, w, c0).  A tap outside the image reads the
// centre pixel instead and is masked to zero: no branch around the loads, so all nine are in flight together (with the
// branch the compiler kept them in program order and the kernel paid nine load latencies per pixel).
__device__ __forceinline__ void dw_pre(const bf16* __restrict__ xp, int h, int w, int H, int W, int ldx, int c0,
                                       const float* __restrict__ sw /*[9][C] for this proj*/, const float* __restrict__ sb,
                                       int C, float* pre) {
  uint4 r[9];
#pragma unroll
  for (int tap = 0; tap < 9; ++tap) {
    const int dh = tap / 3 - 1, dw = tap % 3 - 1;
    const bool in = (unsigned)(h + dh) < (unsigned)H && (unsigned)(w + dw) < (unsigned)W;
    r[tap] = *reinterpret_cast<const uint4*>(xp + (in ? (dh * W + dw) * ldx : 0));
    if (!in) r[tap] = make_uint4(0u, 0u, 0u, 0u);
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) pre[j] = sb[c0 + j];
#pragma unroll
  for (int tap = 0; tap < 9; ++tap) {
    const uint32_t wd[4] = {r[tap].x, r[tap].y, r[tap].z, r[tap].w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      pre[2 * j] += __uint_as_float(wd[j] << 16) * sw[tap * C + c0 + 2 * j];
      pre[2 * j + 1] += __uint_as_float(wd[j] & 0xFFFF0000u) * sw[tap * C + c0 + 2 * j + 1];
    }
  }
}

// pixel index -> (b, h, w) without 64-bit division: the item counts of these kernels stay below 2^31 (checked on the host)
struct PixDiv { FastDiv w, h; int cvs; };      // cvs = log2(channel vectors per pixel)
__device__ __forceinline__ void pix_split(uint32_t pix, const PixDiv d, int W, int H, int& b, int& h, int& w) {
  const uint32_t t2 = fd_div(pix, d.w);
  w = (int)(pix - t2 * (uint32_t)W);
  const uint32_t bq = fd_div(t2, d.h);
  h = (int)(t2 - bq * (uint32_t)H);
  b = (int)bq;
}

// The three projections' parameters are the reference's separate tensors (conv_q/k/v.weight [C,1,3,3], .bias [C],
// layernorm_q/k/v.weight/.bias [C]): three pointers per kind instead of one stacked copy (the host used to torch.stack
// them every forward and autograd unbound the stacked gradients again: ~28 at:: launches per attention layer).
struct C3 { const float* p[3]; };
struct M3 { float* p[3]; };

__global__ void __launch_bounds__(TPB, DWLN_MINB) dwln_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ q, bf16* __restrict__ k, bf16* __restrict__ v,
                                const C3 w, const C3 bias, const C3 gamma, const C3 beta, int B, int H, int W, int C, int ldx,
                                int ldo, float eps, const PixDiv pd) {
  extern __shared__ float sm[];          // w: [3][9][C], b,gamma,beta: [3][C] each
  float* sw = sm;
  float* sb = sm + 27 * C;
  float* sg = sb + 3 * C;
  float* sbe = sg + 3 * C;
  for (int i = threadIdx.x; i < 27 * C; i += blockDim.x) {
    int t = i / (9 * C), r = i % (9 * C), tap = r / C, c = r % C;
    sw[i] = w.p[t][c * 9 + tap];
  }
  for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) {
    const int t = i / C, c = i - t * C;
    sb[i] = bias.p[t] ? bias.p[t][c] : 0.f;
    sg[i] = gamma.p[t][c];
    sbe[i] = beta.p[t][c];
  }
  __syncthreads();
  const int CV = C / 8;
  const uint32_t P = (uint32_t)B * H * W;
  const uint32_t total = ((P << pd.cvs) + 31u) & ~31u;           // whole warps stay alive for the shuffles
  const uint32_t i0 = blockIdx.x * blockDim.x + threadIdx.x;
  const int cv = (int)(i0 & (CV - 1)), c0 = cv * 8;
  const float invC = 1.f / C;
  bf16* outs[3] = {q, k, v};
  for (uint32_t i = i0; i < total; i += gridDim.x * blockDim.x) {
    const uint32_t pix = i >> pd.cvs;
    const bool ok = pix < P;
    const uint32_t pp = ok ? pix : 0u;
    int wq, hq, bq;
    pix_split(pp, pd, W, H, bq, hq, wq);
    const bf16* xp = x + (long long)pp * ldx + c0;
#pragma unroll
    for (int t = 0; t < 3; ++t) {
      float pre[8];
      dw_pre(xp, hq, wq, H, W, ldx, c0, sw + t * 9 * C, sb + t * C, C, pre);
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) { pre[j] = fmaxf(pre[j], 0.f); s += pre[j]; }
      float mean = group_sum(s, CV) * invC;
      float qq = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) { float d = pre[j] - mean; qq += d * d; }
      float rstd = rsqrtf(group_sum(qq, CV) * invC + eps);
#pragma unroll
      for (int j = 0; j < 8; ++j) pre[j] = (pre[j] - mean) * rstd * sg[t * C + c0 + j] + sbe[t * C + c0 + j];
      if (ok) store8(outs[t] + (long long)pix * ldo + c0, pre);
    }
  }
}

// grid.y = projection t
__global__ void __launch_bounds__(TPB, DWLN_MINB) dwln_bwd_pre_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dq, const bf16* __restrict__ dk,
                                    const bf16* __restrict__ dv, bf16* __restrict__ pq, bf16* __restrict__ pk,
                                    bf16* __restrict__ pv, const C3 w, const C3 bias, const C3 gamma, const M3 dgamma,
                                    const M3 dbeta, const M3 db, int B, int H, int W, int C, int ldx, int ldd, float eps,
                                    const PixDiv pd) {
  extern __shared__ float sm[];          // w [9][C], b [C], gamma [C], acc [3][C]
  const int t = blockIdx.y;
  float* sw = sm;
  float* sb = sm + 9 * C;
  float* sg = sb + C;
  float* sacc = sg + C;
  for (int i = threadIdx.x; i < 9 * C; i += blockDim.x) { int tap = i / C, c = i % C; sw[i] = w.p[t][c * 9 + tap]; }
  for (int i = threadIdx.x; i < C; i += blockDim.x) { sb[i] = bias.p[t] ? bias.p[t][i] : 0.f; sg[i] = gamma.p[t][i]; }
  for (int i = threadIdx.x; i < 3 * C; i += blockDim.x) sacc[i] = 0.f;
  __syncthreads();
  const bf16* dys = t == 0 ? dq : (t == 1 ? dk : dv);
  bf16* dps = t == 0 ? pq : (t == 1 ? pk : pv);
  const int CV = C / 8;
  const uint32_t P = (uint32_t)B * H * W;
  const uint32_t total = ((P << pd.cvs) + 31u) & ~31u;
  const uint32_t i0 = blockIdx.x * blockDim.x + threadIdx.x;
  const int cv = (int)(i0 & (CV - 1)), c0 = cv * 8;
  const float invC = 1.f / C;
  float ag[8], ab[8], apb[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) ag[j] = ab[j] = apb[j] = 0.f;
  for (uint32_t i = i0; i < total; i += gridDim.x * blockDim.x) {
    const uint32_t pix = i >> pd.cvs;
    const bool ok = pix < P;
    const uint32_t pp = ok ? pix : 0u;
    int wq, hq, bq;
    pix_split(pp, pd, W, H, bq, hq, wq);
    float pre[8], g[8];
    load8(dys + (long long)pp * ldd + c0, g);
    dw_pre(x + (long long)pp * ldx + c0, hq, wq, H, W, ldx, c0, sw, sb, C, pre);
    if (!ok) {
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] = 0.f;
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) { pre[j] = fmaxf(pre[j], 0.f); s += pre[j]; }
    float mean = group_sum(s, CV) * invC;
    float qq = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) { float d = pre[j] - mean; qq += d * d; }
    float rstd = rsqrtf(group_sum(qq, CV) * invC + eps);
    float a = 0.f, bsum = 0.f, xh[8], gg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      xh[j] = (pre[j] - mean) * rstd;
      gg[j] = g[j] * sg[c0 + j];
      a += gg[j]; bsum += gg[j] * xh[j];
      ag[j] += g[j] * xh[j]; ab[j] += g[j];
    }
    a = group_sum(a, CV) * invC;
    bsum = group_sum(bsum, CV) * invC;
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float dr = rstd * (gg[j] - a - xh[j] * bsum);
      o[j] = pre[j] > 0.f ? dr : 0.f;
      if (ok) apb[j] += bf16_round(o[j]);
    }
    if (ok) store8(dps + (long long)pix * ldd + c0, o);
  }
  // shuffle-fold the lanes that share a channel vector before touching shared memory (C = 8: all 256 threads of
  // the CTA would otherwise contend for the same 24 words with CAS-loop float atomics)
  col_fold8(sacc, ag, cv, CV);
  col_fold8(sacc + C, ab, cv, CV);
  col_fold8(sacc + 2 * C, apb, cv, CV);
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += blockDim.x) {
    atomicAdd(dgamma.p[t] + i, sacc[i]);
    atomicAdd(dbeta.p[t] + i, sacc[C + i]);
    if (db.p[t]) atomicAdd(db.p[t] + i, sacc[2 * C + i]);
  }
}

// thread = (pixel, channel pair): dx and 3x9x2 weight-gradient partials in registers
__global__ void __launch_bounds__(TPB, DWLN_MINB) dw_bwd_dw_kernel(const bf16* __restrict__ x, const bf16* __restrict__ pq,
                                                        const bf16* __restrict__ pk, const bf16* __restrict__ pv,
                                                        bf16* __restrict__ dx, const C3 w, const M3 dw, int B, int H, int W,
                                                        int C, int ldx,
                                                        int ldd, int lddx, const PixDiv pd) {
  extern __shared__ float sm[];          // w [3][9][C], acc [3][9][C]
  float* sw = sm;
  float* sacc = sm + 27 * C;
  for (int i = threadIdx.x; i < 27 * C; i += blockDim.x) {
    int t = i / (9 * C), r = i % (9 * C), tap = r / C, c = r % C;
    sw[i] = w.p[t][c * 9 + tap];
    sacc[i] = 0.f;
  }
  __syncthreads();
  const int C2 = C / 2;
  const uint32_t total = ((uint32_t)B * H * W) << pd.cvs;        // pd.cvs = log2(C / 2) here
  const uint32_t i0 = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = (int)(i0 & (C2 - 1)) * 2;
  float acc[3][9][2];
#pragma unroll
  for (int t = 0; t < 3; ++t)
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) acc[t][tap][0] = acc[t][tap][1] = 0.f;
  const bf16* dps[3] = {pq, pk, pv};
  for (uint32_t i = i0; i < total; i += gridDim.x * blockDim.x) {
    const uint32_t pix = i >> pd.cvs;
    int wq, hq, bq;
    pix_split(pix, pd, W, H, bq, hq, wq);
    const float2 xc = __bfloat1622float2(*reinterpret_cast<const bf162*>(x + (long long)pix * ldx + c0));
    // y[p] = sum_tap x[p+off] w[tap]  =>  dx[q] += dpre[q-off] w[tap];  dw[tap] += dpre[q-off] x[q].
    // All 27 loads first, without branches (a tap outside the image reads the centre pixel and is masked to zero)
    uint32_t raw[3][9];
    const long long base = (long long)pix * ldd + c0;
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      const int dh = 1 - tap / 3, dw_ = 1 - tap % 3;
      const bool in = (unsigned)(hq + dh) < (unsigned)H && (unsigned)(wq + dw_) < (unsigned)W;
      const int off = in ? (dh * W + dw_) * ldd : 0;
#pragma unroll
      for (int t = 0; t < 3; ++t) {
        const uint32_t r = *reinterpret_cast<const uint32_t*>(dps[t] + base + off);
        raw[t][tap] = in ? r : 0u;
      }
    }
    float d0 = 0.f, d1 = 0.f;
#pragma unroll
    for (int tap = 0; tap < 9; ++tap)
#pragma unroll
      for (int t = 0; t < 3; ++t) {
        const float gx = __uint_as_float(raw[t][tap] << 16), gy = __uint_as_float(raw[t][tap] & 0xFFFF0000u);
        const float2 wv = *reinterpret_cast<const float2*>(&sw[(t * 9 + tap) * C + c0]);
        d0 += gx * wv.x;
        d1 += gy * wv.y;
        acc[t][tap][0] += gx * xc.x;
        acc[t][tap][1] += gy * xc.y;
      }
    *reinterpret_cast<bf162*>(dx + (long long)pix * lddx + c0) = __floats2bfloat162_rn(d0, d1);
  }
  // fold lanes that hold the same channel pair, then shared, then global
#pragma unroll
  for (int t = 0; t < 3; ++t)
#pragma unroll
    for (int tap = 0; tap < 9; ++tap)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        float v = acc[t][tap][e];
        for (int o = C2; o < 32; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (C2 >= 32 || (threadIdx.x & 31) < C2) atomicAdd(&sacc[(t * 9 + tap) * C + c0 + e], v);
      }
  __syncthreads();
  for (int i = threadIdx.x; i < 27 * C; i += blockDim.x) {
    int t = i / (9 * C), r = i % (9 * C), tap = r / C, c = r % C;
    atomicAdd(dw.p[t] + c * 9 + tap, sacc[i]);
  }
}
}  // namespace

#define ST ((cudaStream_t)stream)
static int dwln_check(int C) {
  HN_REQUIRE(C >= 8 && C <= 128 && (C & (C - 1)) == 0, "dwconv_relu_ln: C must be a power of two in [8,128], got %d", C);
  return HN_OK;
}
static int dwln_grid(long long items) {
  long long b = (items + TPB - 1) / TPB;
  if (b > 148 * 4) b = 148 * 4;
  if (b < 1) b = 1;
  return (int)b;
}

// division constants of one launch; items = pixels << cvs must stay below 2^31 (32-bit index arithmetic in the kernels)
static int pix_div(int B, int H, int W, int vecs, PixDiv* pd) {
  int cvs = 0;
  while ((1 << cvs) < vecs) ++cvs;
  const long long items = ((long long)B * H * W) << cvs;
  HN_REQUIRE(B > 0 && H > 0 && W > 0 && items + 32 < (1ll << 31), "dwconv_relu_ln: B*H*W*C/%d = %lld items exceed 2^31", 8, items);
  pd->w = hn_fastdiv((uint32_t)W);
  pd->h = hn_fastdiv((uint32_t)H);
  pd->cvs = cvs;
  return HN_OK;
}
static C3 c3(const float* const* a) { C3 r; for (int t = 0; t < 3; ++t) r.p[t] = a ? a[t] : nullptr; return r; }
static M3 m3(float* const* a) { M3 r; for (int t = 0; t < 3; ++t) r.p[t] = a ? a[t] : nullptr; return r; }
static bool all3(const float* const* a) { return a && a[0] && a[1] && a[2]; }

// w / b / gamma / beta (and the gradient targets): HOST arrays of three device pointers, one per projection (q, k, v) --
// the reference's own parameter tensors; b may be null (conv bias=False), as may db
extern "C" int hn_dwconv_relu_ln_fwd(const void* x, void* q, void* k, void* v, const float* const* w, const float* const* b,
                                     const float* const* gamma, const float* const* beta, int B, int H, int W, int C, int ldx,
                                     int ldo, float eps, void* stream) {
  if (int rc = dwln_check(C)) return rc;
  HN_REQUIRE(x && q && k && v && all3(w) && all3(gamma) && all3(beta), "dwconv_relu_ln_fwd: null");
  size_t smem = (27 + 9) * C * sizeof(float);
  PixDiv pd;
  if (int rc = pix_div(B, H, W, C / 8, &pd)) return rc;
  dwln_fwd_kernel<<<dwln_grid((long long)B * H * W * (C / 8)), TPB, smem, ST>>>(
      (const bf16*)x, (bf16*)q, (bf16*)k, (bf16*)v, c3(w), c3(b), c3(gamma), c3(beta), B, H, W, C, ldx, ldo, eps, pd);
  HN_CHECK_LAUNCH("dwln_fwd");
  return HN_OK;
}
extern "C" int hn_dwconv_relu_ln_bwd_pre(const void* x, const void* dq, const void* dk, const void* dv, void* dpre_q,
                                         void* dpre_k, void* dpre_v, const float* const* w, const float* const* b,
                                         const float* const* gamma, float* const* dgamma, float* const* dbeta,
                                         float* const* db, int B, int H, int W, int C, int ldx, int ldd, float eps,
                                         void* stream) {
  if (int rc = dwln_check(C)) return rc;
  HN_REQUIRE(x && dq && dk && dv && dpre_q && dpre_k && dpre_v && all3(w) && all3(gamma) && all3((const float* const*)dgamma) &&
                 all3((const float* const*)dbeta),
             "dwconv_relu_ln_bwd_pre: null");
  size_t smem = (9 + 2 + 3) * C * sizeof(float);
  PixDiv pd;
  if (int rc = pix_div(B, H, W, C / 8, &pd)) return rc;
  dim3 grid(dwln_grid((long long)B * H * W * (C / 8)), 3);
  dwln_bwd_pre_kernel<<<grid, TPB, smem, ST>>>((const bf16*)x, (const bf16*)dq, (const bf16*)dk, (const bf16*)dv,
                                               (bf16*)dpre_q, (bf16*)dpre_k, (bf16*)dpre_v, c3(w), c3(b), c3(gamma), m3(dgamma),
                                               m3(dbeta), m3(db), B, H, W, C, ldx, ldd, eps, pd);
  HN_CHECK_LAUNCH("dwln_bwd_pre");
  return HN_OK;
}
extern "C" int hn_dwconv_bwd_data_weight(const void* x, const void* dpre_q, const void* dpre_k, const void* dpre_v,
                                         void* dx, const float* const* w, float* const* dw, int B, int H, int W, int C, int ldx,
                                         int ldd, int lddx, void* stream) {
  if (int rc = dwln_check(C)) return rc;
  HN_REQUIRE(x && dpre_q && dpre_k && dpre_v && dx && all3(w) && all3((const float* const*)dw), "dwconv_bwd_data_weight: null");
  size_t smem = 54 * C * sizeof(float);
  PixDiv pd;
  if (int rc = pix_div(B, H, W, C / 2, &pd)) return rc;
  long long items = (long long)B * H * W * (C / 2);
  // Every CTA ends with 27 C global atomics onto the same 27 C words (measured on B200: ~1.6 ns x C per CTA, serialised
  // in L2) while a grid-stride iteration costs ~2 us: T(G) = a C G + t items / (256 G) is least at G ~ sqrt(5 items / C)
  // (scripts/dwln_bench.py: 75 -> 44 us at 32x32x32, 80 -> 30 us at 8x8x128 against one CTA per 256 items).
  int grid = dwln_grid(items);
  const int gopt = (int)sqrtf(5.f * (float)items / (float)C);
  if (grid > gopt) grid = gopt < 1 ? 1 : gopt;
  // total threads must be a multiple of C/2 so each thread keeps one channel pair
  dw_bwd_dw_kernel<<<grid, TPB, smem, ST>>>((const bf16*)x, (const bf16*)dpre_q, (const bf16*)dpre_k,
                                            (const bf16*)dpre_v, (bf16*)dx, c3(w), m3(dw), B, H, W, C, ldx, ldd, lddx, pd);
  HN_CHECK_LAUNCH("dw_bwd_dw");
  return HN_OK;
}
