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

__device__ __forceinline__ float group_sum(float v, int CV) {
  for (int o = 1; o < CV; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// pre[j] = b + sum_tap x[p+off][c] * w[c][tap]   (zero padding)
__device__ __forceinline__ void dw_pre(const bf16* __restrict__ x, int b, int h, int w, int H, int W, int ldx, int c0,
                                       const float* __restrict__ sw /*[9][C] for this proj*/, const float* __restrict__ sb,
                                       int C, float* pre) {
#pragma unroll
  for (int j = 0; j < 8; ++j) pre[j] = sb[c0 + j];
#pragma unroll
  for (int tap = 0; tap < 9; ++tap) {
    int ih = h + tap / 3 - 1, iw = w + tap % 3 - 1;
    if ((unsigned)ih < (unsigned)H && (unsigned)iw < (unsigned)W) {
      float f[8];
      load8(x + (((long long)b * H + ih) * W + iw) * ldx + c0, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) pre[j] += f[j] * sw[tap * C + c0 + j];
    }
  }
}

// The three projections' parameters are the reference's separate tensors (conv_q/k/v.weight [C,1,3,3], .bias [C],
// layernorm_q/k/v.weight/.bias [C]): three pointers per kind instead of one stacked copy (the host used to torch.stack
// them every forward and autograd unbound the stacked gradients again: ~28 at:: launches per attention layer).
struct C3 { const float* p[3]; };
struct M3 { float* p[3]; };

__global__ void dwln_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ q, bf16* __restrict__ k, bf16* __restrict__ v,
                                const C3 w, const C3 bias, const C3 gamma, const C3 beta, int B, int H, int W, int C, int ldx,
                                int ldo, float eps) {
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
  const long long P = (long long)B * H * W;
  long long total = (P * CV + 31) / 32 * 32;
  long long i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int cv = (int)(i0 % CV), c0 = cv * 8;
  const float invC = 1.f / C;
  bf16* outs[3] = {q, k, v};
  for (long long i = i0; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long pix = i / CV;
    bool ok = pix < P;
    long long pp = ok ? pix : 0;
    int wq = (int)(pp % W);
    long long t2 = pp / W;
    int hq = (int)(t2 % H), bq = (int)(t2 / H);
#pragma unroll
    for (int t = 0; t < 3; ++t) {
      float pre[8];
      dw_pre(x, bq, hq, wq, H, W, ldx, c0, sw + t * 9 * C, sb + t * C, C, pre);
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
      if (ok) store8(outs[t] + pix * ldo + c0, pre);
    }
  }
}

// grid.y = projection t
__global__ void dwln_bwd_pre_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dq, const bf16* __restrict__ dk,
                                    const bf16* __restrict__ dv, bf16* __restrict__ pq, bf16* __restrict__ pk,
                                    bf16* __restrict__ pv, const C3 w, const C3 bias, const C3 gamma, const M3 dgamma,
                                    const M3 dbeta, const M3 db, int B, int H, int W, int C, int ldx, int ldd, float eps) {
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
  const long long P = (long long)B * H * W;
  long long total = (P * CV + 31) / 32 * 32;
  long long i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int cv = (int)(i0 % CV), c0 = cv * 8;
  const float invC = 1.f / C;
  float ag[8], ab[8], apb[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) ag[j] = ab[j] = apb[j] = 0.f;
  for (long long i = i0; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long pix = i / CV;
    bool ok = pix < P;
    long long pp = ok ? pix : 0;
    int wq = (int)(pp % W);
    long long t2 = pp / W;
    int hq = (int)(t2 % H), bq = (int)(t2 / H);
    float pre[8], g[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    dw_pre(x, bq, hq, wq, H, W, ldx, c0, sw, sb, C, pre);
    if (ok) load8(dys + pix * ldd + c0, g);
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
    if (ok) store8(dps + pix * ldd + c0, o);
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
__global__ void __launch_bounds__(TPB) dw_bwd_dw_kernel(const bf16* __restrict__ x, const bf16* __restrict__ pq,
                                                        const bf16* __restrict__ pk, const bf16* __restrict__ pv,
                                                        bf16* __restrict__ dx, const C3 w, const M3 dw, int B, int H, int W,
                                                        int C, int ldx,
                                                        int ldd, int lddx) {
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
  const long long P = (long long)B * H * W;
  const long long total = P * C2;
  long long i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int c0 = (int)(i0 % C2) * 2;
  float acc[3][9][2];
#pragma unroll
  for (int t = 0; t < 3; ++t)
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) acc[t][tap][0] = acc[t][tap][1] = 0.f;
  const bf16* dps[3] = {pq, pk, pv};
  for (long long i = i0; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long pix = i / C2;
    int wq = (int)(pix % W);
    long long t2 = pix / W;
    int hq = (int)(t2 % H), bq = (int)(t2 / H);
    float2 xc = __bfloat1622float2(*reinterpret_cast<const bf162*>(x + pix * ldx + c0));
    float d0 = 0.f, d1 = 0.f;
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      // y[p] = sum_tap x[p+off] w[tap]  =>  dx[q] += dpre[q-off] w[tap];  dw[tap] += dpre[q-off] x[q]
      int ih = hq - (tap / 3 - 1), iw = wq - (tap % 3 - 1);
      if ((unsigned)ih < (unsigned)H && (unsigned)iw < (unsigned)W) {
        long long np = ((long long)bq * H + ih) * W + iw;
#pragma unroll
        for (int t = 0; t < 3; ++t) {
          float2 g = __bfloat1622float2(*reinterpret_cast<const bf162*>(dps[t] + np * ldd + c0));
          d0 += g.x * sw[(t * 9 + tap) * C + c0];
          d1 += g.y * sw[(t * 9 + tap) * C + c0 + 1];
          acc[t][tap][0] += g.x * xc.x;
          acc[t][tap][1] += g.y * xc.y;
        }
      }
    }
    *reinterpret_cast<bf162*>(dx + pix * lddx + c0) = __floats2bfloat162_rn(d0, d1);
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
  dwln_fwd_kernel<<<dwln_grid((long long)B * H * W * (C / 8)), TPB, smem, ST>>>(
      (const bf16*)x, (bf16*)q, (bf16*)k, (bf16*)v, c3(w), c3(b), c3(gamma), c3(beta), B, H, W, C, ldx, ldo, eps);
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
  dim3 grid(dwln_grid((long long)B * H * W * (C / 8)), 3);
  dwln_bwd_pre_kernel<<<grid, TPB, smem, ST>>>((const bf16*)x, (const bf16*)dq, (const bf16*)dk, (const bf16*)dv,
                                               (bf16*)dpre_q, (bf16*)dpre_k, (bf16*)dpre_v, c3(w), c3(b), c3(gamma), m3(dgamma),
                                               m3(dbeta), m3(db), B, H, W, C, ldx, ldd, eps);
  HN_CHECK_LAUNCH("dwln_bwd_pre");
  return HN_OK;
}
extern "C" int hn_dwconv_bwd_data_weight(const void* x, const void* dpre_q, const void* dpre_k, const void* dpre_v,
                                         void* dx, const float* const* w, float* const* dw, int B, int H, int W, int C, int ldx,
                                         int ldd, int lddx, void* stream) {
  if (int rc = dwln_check(C)) return rc;
  HN_REQUIRE(x && dpre_q && dpre_k && dpre_v && dx && all3(w) && all3((const float* const*)dw), "dwconv_bwd_data_weight: null");
  size_t smem = 54 * C * sizeof(float);
  long long items = (long long)B * H * W * (C / 2);
  int grid = dwln_grid(items);
  // total threads must be a multiple of C/2 so each thread keeps one channel pair
  dw_bwd_dw_kernel<<<grid, TPB, smem, ST>>>((const bf16*)x, (const bf16*)dpre_q, (const bf16*)dpre_k,
                                            (const bf16*)dpre_v, (bf16*)dx, c3(w), m3(dw), B, H, W, C, ldx, ldd, lddx);
  HN_CHECK_LAUNCH("dw_bwd_dw");
  return HN_OK;
}
