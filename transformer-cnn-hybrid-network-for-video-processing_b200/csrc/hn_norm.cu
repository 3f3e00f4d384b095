// LayerNorm over the channel axis and BatchNorm2d (train/eval) on NHWC bf16.
// Warp-shuffle reductions across the lanes that share a pixel (LN) and
// register -> shared -> one-atomic-per-channel column reductions (BN, dgamma).
#include "hn_common.cuh"

namespace {
constexpr int TPB = 256;
inline int grid_for(long long work_items) {
  long long b = (work_items + TPB - 1) / TPB;
  long long cap = 148 * 12;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)((b + 2) / 3 * 3);
}
// BatchNorm column passes: ONE wave of CTAs (3 resident per SM), each thread walking >= 4 pixel pairs.
// Measured (scripts/bn_bench.py, 16x256x256x192): the reduce passes go 5.0 -> 5.7 TB/s and the 8 MB maps
// 2.5x faster than with 4 waves -- the per-CTA tail (2*C global atomics) is paid 444 times instead of 1776.
inline int grid_bn(long long work_items) {
  long long b = (work_items + TPB * 8 - 1) / (TPB * 8);
  const long long cap = 148 * 3;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)((b + 2) / 3 * 3);
}
__device__ __forceinline__ float group_sum(float v, int CV) {
  for (int o = 1; o < CV; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// CV = C/8 lanes cooperate on one pixel (CV is a power of two <= 32)
__global__ void ln_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, const float* __restrict__ gamma,
                              const float* __restrict__ beta, long long P, int C, int ldx, int ldy, float eps) {
  const int CV = C / 8;
  long long total = (P * CV + 31) / 32 * 32;     // keep whole warps alive for the shuffles
  long long i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int cv = (int)(i0 % CV);
  float ga[8], be[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { ga[j] = gamma[cv * 8 + j]; be[j] = beta[cv * 8 + j]; }
  const float invC = 1.f / C;
  for (long long i = i0; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long pix = i / CV;
    bool ok = pix < P;
    float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (ok) load8(x + pix * ldx + cv * 8, f);
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += f[j];
    float mean = group_sum(s, CV) * invC;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) { float d = f[j] - mean; q += d * d; }
    float rstd = rsqrtf(group_sum(q, CV) * invC + eps);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = (f[j] - mean) * rstd * ga[j] + be[j];
    if (ok) store8(y + pix * ldy + cv * 8, f);
  }
}

__global__ void ln_bwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy, bf16* __restrict__ dx,
                              const float* __restrict__ gamma, float* __restrict__ dgamma, float* __restrict__ dbeta,
                              long long P, int C, int ldx, int lddy, int lddx, float eps) {
  extern __shared__ float s_acc[];   // [2][C]
  const int CV = C / 8;
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) s_acc[i] = 0.f;
  __syncthreads();
  long long total = (P * CV + 31) / 32 * 32;
  long long i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int cv = (int)(i0 % CV);
  float ga[8], pg[8], pb[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { ga[j] = gamma[cv * 8 + j]; pg[j] = 0.f; pb[j] = 0.f; }
  const float invC = 1.f / C;
  for (long long i = i0; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long pix = i / CV;
    bool ok = pix < P;
    float f[8] = {0, 0, 0, 0, 0, 0, 0, 0}, g[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (ok) { load8(x + pix * ldx + cv * 8, f); load8(dy + pix * lddy + cv * 8, g); }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += f[j];
    float mean = group_sum(s, CV) * invC;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) { float d = f[j] - mean; q += d * d; }
    float rstd = rsqrtf(group_sum(q, CV) * invC + eps);
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float xh = (f[j] - mean) * rstd;
      float gg = g[j] * ga[j];
      a += gg; b += gg * xh;
      pg[j] += g[j] * xh; pb[j] += g[j];
      f[j] = xh; g[j] = gg;
    }
    a = group_sum(a, CV) * invC;
    b = group_sum(b, CV) * invC;
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = rstd * (g[j] - a - f[j] * b);
    if (ok) store8(dx + pix * lddx + cv * 8, f);
  }
  // lanes that share a channel vector are folded by shuffles first: with C = 8 every thread of the CTA would
  // otherwise hit the same 16 shared-memory words with CAS-loop float atomics (256-way contention)
  col_fold8(s_acc, pg, cv, CV);
  col_fold8(s_acc + C, pb, cv, CV);
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += blockDim.x) {
    atomicAdd(dgamma + i, s_acc[i]);
    atomicAdd(dbeta + i, s_acc[C + i]);
  }
}

// ---- BatchNorm ----------------------------------------------------------------------
// All passes: thread <-> fixed channel vector, two pixels in flight per iteration (loads of both
// issued before any use) so that ~48-96 B per thread are outstanding -- what HBM3e needs.
#define HN_COL_LOOP2(it, P)                                                        \
  for (long long pix = (it).pix0; pix < (P); pix += 2 * (it).pstride)
__global__ void __launch_bounds__(256) bn_stats_kernel(const bf16* __restrict__ x, long long P, int C, int ld,
                                                       float* __restrict__ sum, float* __restrict__ sq) {
  extern __shared__ float s_acc[];   // [2][C]
  const int CV = C / 8;
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) s_acc[i] = 0.f;
  __syncthreads();
  float ps[8] = {0, 0, 0, 0, 0, 0, 0, 0}, pq[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const ColIter it = col_iter(CV);
  HN_COL_LOOP2(it, P) {
    const long long pix2 = pix + it.pstride;
    const bool two = pix2 < P;
    float f[8], h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    load8(x + pix * ld + it.cv * 8, f);
    if (two) load8(x + pix2 * ld + it.cv * 8, h);
#pragma unroll
    for (int j = 0; j < 8; ++j) { ps[j] += f[j] + h[j]; pq[j] += f[j] * f[j] + h[j] * h[j]; }
  }
  col_fold8(s_acc, ps, it.cv, CV);
  col_fold8(s_acc + C, pq, it.cv, CV);
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += blockDim.x) {
    atomicAdd(sum + i, s_acc[i]);
    atomicAdd(sq + i, s_acc[C + i]);
  }
}
__global__ void bn_finalize_kernel(const float* __restrict__ sum, const float* __restrict__ sq, long long count, int C,
                                   float eps, float momentum, float* __restrict__ mean, float* __restrict__ rstd,
                                   float* __restrict__ rmean, float* __restrict__ rvar,
                                   long long* __restrict__ num_batches) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c == 0 && num_batches) *num_batches += 1;      // BatchNorm2d.num_batches_tracked (one launch less per layer)
  if (c >= C) return;
  double m = (double)sum[c] / (double)count;
  double v = (double)sq[c] / (double)count - m * m;
  if (v < 0) v = 0;
  mean[c] = (float)m;
  rstd[c] = (float)(1.0 / sqrt(v + (double)eps));
  if (rmean) {
    double unb = count > 1 ? v * (double)count / (double)(count - 1) : v;
    rmean[c] = (1.f - momentum) * rmean[c] + momentum * (float)m;
    rvar[c] = (1.f - momentum) * rvar[c] + momentum * (float)unb;
  }
}
// U pixels in flight per thread: 2 with a residual stream (64 B outstanding), 4 without -- with one input
// stream two 16-byte loads per thread at 768 threads/SM left HBM3e under-subscribed (3.7 of ~5.5 TB/s).
// FIN: the statistics pass is folded in (training): every CTA derives mean / rstd of all channels from the sums the conv
// epilogue left (one channel per thread, the bn_finalize arithmetic in double) into shared memory; CTA 0 also writes them
// out for the backward pass and moves the running statistics and num_batches_tracked.  One launch less per BatchNorm
// layer (37 per Autoencoder32K step).
struct BnFin {
  const float* sum;
  const float* sq;
  long long count;
  float eps, momentum;
  float* mean_out;
  float* rstd_out;
  float* rmean;
  float* rvar;
  long long* num_batches;
};
template <int U, bool RES, bool FIN>
__global__ void __launch_bounds__(256, 3) bn_apply_kernel(const bf16* __restrict__ x, const bf16* __restrict__ res,
                                                       bf16* __restrict__ y, const float* __restrict__ mean,
                                                       const float* __restrict__ rstd, const float* __restrict__ gamma,
                                                       const float* __restrict__ beta, long long P, int C, int ldx,
                                                       int ldr, int ldy, int act, const BnFin fin) {
  extern __shared__ float s_ms[];      // FIN: [2][C] mean, rstd
  if (FIN) {
    if (blockIdx.x == 0 && threadIdx.x == 0 && fin.num_batches) *fin.num_batches += 1;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      const double m = (double)fin.sum[c] / (double)fin.count;
      double v = (double)fin.sq[c] / (double)fin.count - m * m;
      if (v < 0) v = 0;
      const float mf = (float)m, rf = (float)(1.0 / sqrt(v + (double)fin.eps));
      s_ms[c] = mf;
      s_ms[C + c] = rf;
      if (blockIdx.x == 0) {
        fin.mean_out[c] = mf;
        fin.rstd_out[c] = rf;
        if (fin.rmean) {
          const double unb = fin.count > 1 ? v * (double)fin.count / (double)(fin.count - 1) : v;
          fin.rmean[c] = (1.f - fin.momentum) * fin.rmean[c] + fin.momentum * mf;
          fin.rvar[c] = (1.f - fin.momentum) * fin.rvar[c] + fin.momentum * (float)unb;
        }
      }
    }
    __syncthreads();
    mean = s_ms;
    rstd = s_ms + C;
  }
  const int CV = C / 8;
  const ColIter it = col_iter(CV);
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    int c = it.cv * 8 + j;
    sc[j] = rstd[c] * gamma[c];
    sh[j] = beta[c] - mean[c] * sc[j];
  }
  for (long long pix = it.pix0; pix < P; pix += (long long)U * it.pstride) {
    float f[U][8], r[RES ? U : 1][8];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long pu = pix + (long long)u * it.pstride;
      if (pu < P) load8(x + pu * ldx + it.cv * 8, f[u]);
    }
    if (RES) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long pu = pix + (long long)u * it.pstride;
        if (pu < P) load8(res + pu * ldr + it.cv * 8, r[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long pu = pix + (long long)u * it.pstride;
      if (pu < P) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float v = fmaf(f[u][j], sc[j], sh[j]);
          if (RES) v += r[RES ? u : 0][j];
          f[u][j] = apply_act(v, act);
        }
        store8(y + pu * ldy + it.cv * 8, f[u]);
      }
    }
  }
}
// MASKZ: act is ReLU without residual and the mask is re-derived from x exactly as bn_apply computed
// y (same fp32 expression), so neither y is read nor dz written: 2 passes instead of 4.
template <bool MASKZ, int act>
__global__ void __launch_bounds__(256, 3) bn_bwd_reduce_kernel(
    const bf16* __restrict__ dy, const bf16* __restrict__ y, const bf16* __restrict__ x, bf16* __restrict__ dz,
    const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
    const float* __restrict__ beta, long long P, int C, int lddy, int ldy, int ldx, int lddz,
    float* __restrict__ s1, float* __restrict__ s2) {
  extern __shared__ float s_acc[];   // [2][C]
  const int CV = C / 8;
  for (int i = threadIdx.x; i < 2 * C; i += blockDim.x) s_acc[i] = 0.f;
  __syncthreads();
  const ColIter it = col_iter(CV);
  float mu[8], sc[8], sh[8], p1[8], p2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    int c = it.cv * 8 + j;
    mu[j] = mean[c]; p1[j] = p2[j] = 0.f;
    if (MASKZ) { sc[j] = rstd[c] * gamma[c]; sh[j] = beta[c] - mu[j] * sc[j]; }
  }
  HN_COL_LOOP2(it, P) {
    const long long pix2 = pix + it.pstride;
    const bool two = pix2 < P;
    float g[2][8], a[2][8], f[2][8];
    load8(dy + pix * lddy + it.cv * 8, g[0]);
    load8(x + pix * ldx + it.cv * 8, f[0]);
    if (two) {
      load8(dy + pix2 * lddy + it.cv * 8, g[1]);
      load8(x + pix2 * ldx + it.cv * 8, f[1]);
    }
    if (!MASKZ && act != HN_ACT_NONE) {
      load8(y + pix * ldy + it.cv * 8, a[0]);
      if (two) load8(y + pix2 * ldy + it.cv * 8, a[1]);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !two) break;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v = g[u][j];
        if (MASKZ) v = fmaf(f[u][j], sc[j], sh[j]) > 0.f ? v : 0.f;
        else if (act == HN_ACT_RELU) v = a[u][j] > 0.f ? v : 0.f;
        else if (act == HN_ACT_SIGMOID) v = v * a[u][j] * (1.f - a[u][j]);
        g[u][j] = v;
        float vr = MASKZ ? v : bf16_round(v);
        p1[j] += vr;
        p2[j] += vr * (f[u][j] - mu[j]);
      }
      if (!MASKZ && dz) store8(dz + (u ? pix2 : pix) * lddz + it.cv * 8, g[u]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) p2[j] *= rstd[it.cv * 8 + j];
  col_fold8(s_acc, p1, it.cv, CV);
  col_fold8(s_acc + C, p2, it.cv, CV);
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += blockDim.x) {
    atomicAdd(s1 + i, s_acc[i]);
    atomicAdd(s2 + i, s_acc[C + i]);
  }
}
template <bool MASKZ>
__global__ void __launch_bounds__(256, 3) bn_bwd_apply_kernel(
    const bf16* __restrict__ dz, const bf16* __restrict__ x, bf16* __restrict__ dx, const float* __restrict__ mean,
    const float* __restrict__ rstd, const float* __restrict__ gamma, const float* __restrict__ beta,
    const float* __restrict__ s1, const float* __restrict__ s2, long long P, int C, int lddz, int ldx, int lddx,
    int training) {
  const int CV = C / 8;
  const ColIter it = col_iter(CV);
  // dx = k*(v - m1 - xhat*m2) = k*v + ca + cb*x with per-channel constants
  float k[8], ca[8], cb[8], sh[8];
  const float invP = 1.f / (float)P;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    int c = it.cv * 8 + j;
    const float mu = mean[c], rs = rstd[c];
    k[j] = rs * gamma[c];
    const float m1 = training ? s1[c] * invP : 0.f;
    const float m2 = training ? s2[c] * invP : 0.f;
    cb[j] = -k[j] * rs * m2;
    ca[j] = -k[j] * m1 - cb[j] * mu;
    if (MASKZ) sh[j] = beta[c] - mu * k[j];
  }
  HN_COL_LOOP2(it, P) {
    const long long pix2 = pix + it.pstride;
    const bool two = pix2 < P;
    float g[2][8], f[2][8];
    load8(dz + pix * lddz + it.cv * 8, g[0]);
    load8(x + pix * ldx + it.cv * 8, f[0]);
    if (two) {
      load8(dz + pix2 * lddz + it.cv * 8, g[1]);
      load8(x + pix2 * ldx + it.cv * 8, f[1]);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !two) break;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v = g[u][j];
        if (MASKZ) v = fmaf(f[u][j], k[j], sh[j]) > 0.f ? v : 0.f;
        f[u][j] = fmaf(k[j], v, fmaf(cb[j], f[u][j], ca[j]));
      }
      store8(dx + (u ? pix2 : pix) * lddx + it.cv * 8, f[u]);
    }
  }
}
}  // namespace

#define ST ((cudaStream_t)stream)
#define REQ_CV(C, name) HN_REQUIRE((C) > 0 && (C) % 8 == 0 && (768 % ((C) / 8)) == 0 && (C) <= 2048, name ": unsupported channel count %d", (C))
#define REQ_LN(C, name) HN_REQUIRE((C) >= 8 && (C) <= 256 && ((C) & ((C)-1)) == 0, name ": C must be a power of two in [8,256], got %d", (C))

extern "C" int hn_layernorm_fwd(const void* x, void* y, const float* gamma, const float* beta, long long P, int C, int ldx, int ldy, float eps, void* stream) {
  REQ_LN(C, "layernorm_fwd");
  HN_REQUIRE(x && y && gamma && beta, "layernorm_fwd: null");
  ln_fwd_kernel<<<grid_for(P * (C / 8)), TPB, 0, ST>>>((const bf16*)x, (bf16*)y, gamma, beta, P, C, ldx, ldy, eps);
  HN_CHECK_LAUNCH("ln_fwd");
  return HN_OK;
}
extern "C" int hn_layernorm_bwd(const void* x, const void* dy, void* dx, const float* gamma, float* dgamma, float* dbeta, long long P, int C, int ldx, int lddy, int lddx, float eps, void* stream) {
  REQ_LN(C, "layernorm_bwd");
  HN_REQUIRE(x && dy && dx && gamma && dgamma && dbeta, "layernorm_bwd: null");
  ln_bwd_kernel<<<grid_for(P * (C / 8)), TPB, 2 * C * sizeof(float), ST>>>((const bf16*)x, (const bf16*)dy, (bf16*)dx, gamma, dgamma, dbeta, P, C, ldx, lddy, lddx, eps);
  HN_CHECK_LAUNCH("ln_bwd");
  return HN_OK;
}
extern "C" int hn_bn_stats(const void* x, long long P, int C, int ld, float* sum, float* sqsum, void* stream) {
  REQ_CV(C, "bn_stats");
  HN_REQUIRE(x && sum && sqsum, "bn_stats: null");
  bn_stats_kernel<<<grid_bn(P * (C / 8)), TPB, 2 * C * sizeof(float), ST>>>((const bf16*)x, P, C, ld, sum, sqsum);
  HN_CHECK_LAUNCH("bn_stats");
  return HN_OK;
}
extern "C" int hn_bn_finalize(const float* sum, const float* sqsum, long long count, int C, float eps, float momentum, float* mean, float* rstd, float* running_mean, float* running_var, long long* num_batches_tracked, void* stream) {
  HN_REQUIRE(sum && sqsum && mean && rstd && count > 0 && C > 0, "bn_finalize: bad args");
  HN_REQUIRE((running_mean == nullptr) == (running_var == nullptr), "bn_finalize: running stats come in pairs");
  bn_finalize_kernel<<<hn_cdiv(C, 128), 128, 0, ST>>>(sum, sqsum, count, C, eps, momentum, mean, rstd, running_mean, running_var, num_batches_tracked);
  HN_CHECK_LAUNCH("bn_finalize");
  return HN_OK;
}
extern "C" int hn_bn_apply(const void* x, const void* res, void* y, const float* mean, const float* rstd, const float* gamma, const float* beta, long long P, int C, int ldx, int ldr, int ldy, int act, void* stream) {
  REQ_CV(C, "bn_apply");
  HN_REQUIRE(x && y && mean && rstd && gamma && beta, "bn_apply: null");
  BnFin fin = {};
  if (res) bn_apply_kernel<2, true, false><<<grid_bn(P * (C / 8)), TPB, 0, ST>>>((const bf16*)x, (const bf16*)res, (bf16*)y, mean, rstd, gamma, beta, P, C, ldx, ldr, ldy, act, fin);
  else bn_apply_kernel<4, false, false><<<grid_bn(P * (C / 8)), TPB, 0, ST>>>((const bf16*)x, nullptr, (bf16*)y, mean, rstd, gamma, beta, P, C, ldx, ldr, ldy, act, fin);
  HN_CHECK_LAUNCH("bn_apply");
  return HN_OK;
}
extern "C" int hn_bn_apply_train(const void* x, const void* res, void* y, const float* sum, const float* sqsum, long long count,
                                 float eps, float momentum, float* mean, float* rstd, float* running_mean, float* running_var,
                                 long long* num_batches_tracked, const float* gamma, const float* beta, long long P, int C,
                                 int ldx, int ldr, int ldy, int act, void* stream) {
  REQ_CV(C, "bn_apply_train");
  HN_REQUIRE(x && y && sum && sqsum && mean && rstd && gamma && beta && count > 0, "bn_apply_train: null");
  HN_REQUIRE((running_mean == nullptr) == (running_var == nullptr), "bn_apply_train: running stats come in pairs");
  BnFin fin = {sum, sqsum, count, eps, momentum, mean, rstd, running_mean, running_var, num_batches_tracked};
  const size_t smem = 2 * (size_t)C * sizeof(float);
  if (res) bn_apply_kernel<2, true, true><<<grid_bn(P * (C / 8)), TPB, smem, ST>>>((const bf16*)x, (const bf16*)res, (bf16*)y, nullptr, nullptr, gamma, beta, P, C, ldx, ldr, ldy, act, fin);
  else bn_apply_kernel<4, false, true><<<grid_bn(P * (C / 8)), TPB, smem, ST>>>((const bf16*)x, nullptr, (bf16*)y, nullptr, nullptr, gamma, beta, P, C, ldx, ldr, ldy, act, fin);
  HN_CHECK_LAUNCH("bn_apply_train");
  return HN_OK;
}
extern "C" int hn_bn_bwd_reduce(const void* dy, const void* y, const void* x, void* dz, const float* mean, const float* rstd, long long P, int C, int lddy, int ldy, int ldx, int lddz, int act, float* s1, float* s2, void* stream) {
  REQ_CV(C, "bn_bwd_reduce");
  HN_REQUIRE(dy && x && mean && rstd && s1 && s2 && (act == HN_ACT_NONE || y), "bn_bwd_reduce: null");
#define HN_BN_RED(ACT) bn_bwd_reduce_kernel<false, ACT><<<grid_bn(P * (C / 8)), TPB, 2 * C * sizeof(float), ST>>>((const bf16*)dy, (const bf16*)y, (const bf16*)x, (bf16*)dz, mean, rstd, nullptr, nullptr, P, C, lddy, ldy, ldx, lddz, s1, s2)
  if (act == HN_ACT_NONE) HN_BN_RED(HN_ACT_NONE);
  else if (act == HN_ACT_RELU) HN_BN_RED(HN_ACT_RELU);
  else if (act == HN_ACT_SIGMOID) HN_BN_RED(HN_ACT_SIGMOID);
  else HN_REQUIRE(false, "bn_bwd_reduce: unsupported activation %d", act);
#undef HN_BN_RED
  HN_CHECK_LAUNCH("bn_bwd_reduce");
  return HN_OK;
}
extern "C" int hn_bn_bwd_apply(const void* dz, const void* x, void* dx, const float* mean, const float* rstd, const float* gamma, const float* s1, const float* s2, long long P, int C, int lddz, int ldx, int lddx, int training, void* stream) {
  REQ_CV(C, "bn_bwd_apply");
  HN_REQUIRE(dz && x && dx && mean && rstd && gamma && s1 && s2, "bn_bwd_apply: null");
  bn_bwd_apply_kernel<false><<<grid_bn(P * (C / 8)), TPB, 0, ST>>>((const bf16*)dz, (const bf16*)x, (bf16*)dx, mean, rstd, gamma, nullptr, s1, s2, P, C, lddz, ldx, lddx, training);
  HN_CHECK_LAUNCH("bn_bwd_apply");
  return HN_OK;
}
extern "C" int hn_bn_relu_bwd_reduce(const void* dy, const void* x, const float* mean, const float* rstd, const float* gamma, const float* beta, long long P, int C, int lddy, int ldx, float* s1, float* s2, void* stream) {
  REQ_CV(C, "bn_relu_bwd_reduce");
  HN_REQUIRE(dy && x && mean && rstd && gamma && beta && s1 && s2, "bn_relu_bwd_reduce: null");
  bn_bwd_reduce_kernel<true, HN_ACT_RELU><<<grid_bn(P * (C / 8)), TPB, 2 * C * sizeof(float), ST>>>((const bf16*)dy, nullptr, (const bf16*)x, nullptr, mean, rstd, gamma, beta, P, C, lddy, 0, ldx, 0, s1, s2);
  HN_CHECK_LAUNCH("bn_relu_bwd_reduce");
  return HN_OK;
}
extern "C" int hn_bn_relu_bwd_apply(const void* dy, const void* x, void* dx, const float* mean, const float* rstd, const float* gamma, const float* beta, const float* s1, const float* s2, long long P, int C, int lddy, int ldx, int lddx, int training, void* stream) {
  REQ_CV(C, "bn_relu_bwd_apply");
  HN_REQUIRE(dy && x && dx && mean && rstd && gamma && beta && s1 && s2, "bn_relu_bwd_apply: null");
  bn_bwd_apply_kernel<true><<<grid_bn(P * (C / 8)), TPB, 0, ST>>>((const bf16*)dy, (const bf16*)x, (bf16*)dx, mean, rstd, gamma, beta, s1, s2, P, C, lddy, ldx, lddx, training);
  HN_CHECK_LAUNCH("bn_relu_bwd_apply");
  return HN_OK;
}
