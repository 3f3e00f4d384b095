// Vectorised (16-byte) pointwise / pooling / layout kernels on NHWC bf16.
// All are HBM-bound by construction: one pass, coalesced 16B accesses, grid =
// a multiple of the SM count with a grid-stride loop.
#include <stdarg.h>

#include "hn_common.cuh"

// ---- error string ------------------------------------------------------------
static thread_local char g_err[512] = "";
void hn_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* hn_last_error(void) { return g_err; }
extern "C" int hn_version(void) { return 100; }
extern "C" int hn_check_device(void) {
  int dev = 0;
  cudaDeviceProp prop;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess)
    HN_FAIL(HN_ECUDA, "no CUDA device");
  if (prop.major != 10) HN_FAIL(HN_EUNSUPPORTED, "device is sm_%d%d, library is sm_100a only", prop.major, prop.minor);
  return HN_OK;
}

namespace {
constexpr int TPB = 256;
inline int grid_for(long long work_items) {
  long long b = (work_items + TPB - 1) / TPB;
  long long cap = 148 * 12;                 // multiple of 148 and of 3 (see col_* helpers)
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)((b + 2) / 3 * 3);
}

// ---- column (per-channel) reduction helpers --------------------------------------
// Threads keep a fixed channel-vector cv = global_thread % CV (the launch makes the
// total thread count a multiple of CV), accumulate 8 partial sums in registers,
// then fold into shared and finally one global atomic per channel per CTA.
__device__ __forceinline__ void col_flush(float* s_acc, const float* part, int cv, int C, int nsets) {
  // s_acc: [nsets][C] zeroed by caller; part: [nsets][8]
  for (int s = 0; s < nsets; ++s)
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(&s_acc[s * C + cv * 8 + j], part[s * 8 + j]);
}

// ---- layout ------------------------------------------------------------------------
__global__ void nchw2nhwc_kernel(const float* __restrict__ src, bf16* __restrict__ dst, int B, int C, int HW,
                                 int Cpad, int ld) {
  long long total = (long long)B * HW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int b = (int)(i / HW);
    int r = (int)(i - (long long)b * HW);
    const float* s = src + (long long)b * C * HW + r;
    bf16* o = dst + i * ld;
    for (int c0 = 0; c0 < Cpad; c0 += 8) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = (c0 + j < C) ? s[(long long)(c0 + j) * HW] : 0.f;
      store8(o + c0, f);
    }
  }
}
__global__ void nhwc2nchw_kernel(const bf16* __restrict__ src, float* __restrict__ dst, int B, int C, int HW, int ld) {
  long long total = (long long)B * HW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int b = (int)(i / HW);
    int r = (int)(i - (long long)b * HW);
    const bf16* s = src + i * ld;
    float* o = dst + (long long)b * C * HW + r;
    for (int c0 = 0; c0 < C; c0 += 8) {
      float f[8];
      load8(s + c0, f);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (c0 + j < C) o[(long long)(c0 + j) * HW] = f[j];
    }
  }
}

// ---- pooling -------------------------------------------------------------------------
template <int MODE>  // 0 avg fwd, 1 max fwd
__global__ void pool2_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, int B, int OH, int OW, int CV,
                                 int ldx, int ldy) {
  long long total = (long long)B * OH * OW * CV;
  int W = OW * 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int cv = (int)(i % CV);
    long long op = i / CV;
    int ow = (int)(op % OW);
    long long t = op / OW;
    int oh = (int)(t % OH);
    int b = (int)(t / OH);
    const bf16* s = x + (((long long)b * OH * 2 + oh * 2) * W + ow * 2) * ldx + cv * 8;
    float a[8], c[8], e[8], g[8], o[8];
    load8(s, a); load8(s + ldx, c); load8(s + (long long)W * ldx, e); load8(s + (long long)W * ldx + ldx, g);
#pragma unroll
    for (int j = 0; j < 8; ++j)
      o[j] = MODE == 0 ? 0.25f * (a[j] + c[j] + e[j] + g[j]) : fmaxf(fmaxf(a[j], c[j]), fmaxf(e[j], g[j]));
    store8(y + op * ldy + cv * 8, o);
  }
}
__global__ void maxpool2_bwd_kernel(const bf16* __restrict__ x, const bf16* __restrict__ dy, bf16* __restrict__ dx,
                                    int B, int OH, int OW, int CV, int ldx, int lddy, int lddx) {
  long long total = (long long)B * OH * OW * CV;
  int W = OW * 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int cv = (int)(i % CV);
    long long op = i / CV;
    int ow = (int)(op % OW);
    long long t = op / OW;
    int oh = (int)(t % OH);
    int b = (int)(t / OH);
    long long ip = ((long long)b * OH * 2 + oh * 2) * W + ow * 2;
    const bf16* s = x + ip * ldx + cv * 8;
    float v[4][8], g[8], o[4][8];
    load8(s, v[0]); load8(s + ldx, v[1]); load8(s + (long long)W * ldx, v[2]); load8(s + (long long)W * ldx + ldx, v[3]);
    load8(dy + op * lddy + cv * 8, g);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int best = 0;
      float m = v[0][j];
#pragma unroll
      for (int k = 1; k < 4; ++k)
        if (v[k][j] > m) { m = v[k][j]; best = k; }   // first max wins (PyTorch scan order)
#pragma unroll
      for (int k = 0; k < 4; ++k) o[k][j] = (k == best) ? g[j] : 0.f;
    }
    bf16* d = dx + ip * lddx + cv * 8;
    store8(d, o[0]); store8(d + lddx, o[1]); store8(d + (long long)W * lddx, o[2]); store8(d + (long long)W * lddx + lddx, o[3]);
  }
}
__global__ void upsample2_bwd_kernel(const bf16* __restrict__ dy, bf16* __restrict__ dx, int B, int H, int W, int CV,
                                     int lddy, int lddx) {
  long long total = (long long)B * H * W * CV;
  int W2 = W * 2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    int cv = (int)(i % CV);
    long long op = i / CV;
    int w = (int)(op % W);
    long long t = op / W;
    int h = (int)(t % H);
    int b = (int)(t / H);
    const bf16* s = dy + (((long long)b * H * 2 + h * 2) * W2 + w * 2) * lddy + cv * 8;
    float a[8], c[8], e[8], g[8], o[8];
    load8(s, a); load8(s + lddy, c); load8(s + (long long)W2 * lddy, e); load8(s + (long long)W2 * lddy + lddy, g);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = a[j] + c[j] + e[j] + g[j];
    store8(dx + op * lddx + cv * 8, o);
  }
}

// ---- activation backward (+ bias gradient) -----------------------------------------------
__global__ void __launch_bounds__(256) act_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ yz,
                                                      bf16* __restrict__ dz, long long P, int C, int lddy, int ldy,
                                                      int lddz, int act, float* __restrict__ dbias) {
  extern __shared__ float s_acc[];
  const int CV = C / 8;
  if (dbias) {
    for (int i = threadIdx.x; i < C; i += blockDim.x) s_acc[i] = 0.f;
    __syncthreads();
  }
  float part[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const ColIter it = col_iter(CV);
  for (long long pix = it.pix0; pix < P; pix += 2 * it.pstride) {
    const long long pix2 = pix + it.pstride;
    const bool two = pix2 < P;
    float g[2][8], a[2][8];
    load8(dy + pix * lddy + it.cv * 8, g[0]);
    if (two) load8(dy + pix2 * lddy + it.cv * 8, g[1]);
    if (act != HN_ACT_NONE) {
      load8(yz + pix * ldy + it.cv * 8, a[0]);
      if (two) load8(yz + pix2 * ldy + it.cv * 8, a[1]);
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !two) break;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v = g[u][j];
        if (act == HN_ACT_RELU) v = a[u][j] > 0.f ? v : 0.f;
        else if (act == HN_ACT_SIGMOID) v = v * a[u][j] * (1.f - a[u][j]);
        else if (act == HN_ACT_GELU) v = v * gelu_grad_f(a[u][j]);
        g[u][j] = v;
        part[j] += bf16_round(v);
      }
      store8(dz + (u ? pix2 : pix) * lddz + it.cv * 8, g[u]);
    }
  }
  if (dbias) {
    col_fold8(s_acc, part, it.cv, CV);
    __syncthreads();
    for (int i = threadIdx.x; i < C; i += blockDim.x) atomicAdd(dbias + i, s_acc[i]);
  }
}
__global__ void act_fwd_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, long long P, int CV, int ldx, int ldy, int act) {
  long long total = P * CV;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long pix = i / CV;
    int cv = (int)(i - pix * CV);
    float f[8];
    load8(x + pix * ldx + cv * 8, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] = apply_act(f[j], act);
    store8(y + pix * ldy + cv * 8, f);
  }
}
__global__ void __launch_bounds__(256) colsum_kernel(const bf16* __restrict__ x, long long P, int C, int ld,
                                                     float* __restrict__ out) {
  extern __shared__ float s_acc[];
  const int CV = C / 8;
  for (int i = threadIdx.x; i < C; i += blockDim.x) s_acc[i] = 0.f;
  __syncthreads();
  float part[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const ColIter it = col_iter(CV);
  for (long long pix = it.pix0; pix < P; pix += 4 * it.pstride) {
    float g[4][8];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long q = pix + u * it.pstride;
      if (q < P) load8(x + q * ld + it.cv * 8, g[u]);
      else {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[u][j] = 0.f;
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) part[j] += (g[0][j] + g[1][j]) + (g[2][j] + g[3][j]);
  }
  col_fold8(s_acc, part, it.cv, CV);
  __syncthreads();
  for (int i = threadIdx.x; i < C; i += blockDim.x) atomicAdd(out + i, s_acc[i]);
}

__global__ void add_kernel(const bf16* __restrict__ a, const bf16* __restrict__ b, const bf16* __restrict__ c,
                           bf16* __restrict__ y, long long P, int CV, int lda, int ldb, int ldc, int ldy) {
  long long total = P * CV;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    long long pix = i / CV;
    int cv = (int)(i - pix * CV);
    float fa[8], fb[8], fc[8];
    load8(a + pix * lda + cv * 8, fa);
    load8(b + pix * ldb + cv * 8, fb);
    if (c) {
      load8(c + pix * ldc + cv * 8, fc);
#pragma unroll
      for (int j = 0; j < 8; ++j) fa[j] += fc[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) fa[j] += fb[j];
    store8(y + pix * ldy + cv * 8, fa);
  }
}
__global__ void gelu3_fwd_kernel(const bf16* __restrict__ z1, const bf16* __restrict__ z2, const bf16* __restrict__ z3,
                                 bf16* __restrict__ y, long long nv) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
    float a[8], b[8], c[8];
    load8(z1 + i * 8, a); load8(z2 + i * 8, b); load8(z3 + i * 8, c);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      // each branch is rounded to bf16 like a stored activation would be
      a[j] = bf16_round(gelu_f(a[j])) + bf16_round(gelu_f(b[j])) + bf16_round(gelu_f(c[j]));
    }
    store8(y + i * 8, a);
  }
}
__global__ void gelu3_bwd_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ z1, const bf16* __restrict__ z2,
                                 const bf16* __restrict__ z3, bf16* __restrict__ d1, bf16* __restrict__ d2,
                                 bf16* __restrict__ d3, long long nv) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
    float g[8], a[8], b[8], c[8];
    load8(dy + i * 8, g); load8(z1 + i * 8, a); load8(z2 + i * 8, b); load8(z3 + i * 8, c);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      a[j] = g[j] * gelu_grad_f(a[j]);
      b[j] = g[j] * gelu_grad_f(b[j]);
      c[j] = g[j] * gelu_grad_f(c[j]);
    }
    store8(d1 + i * 8, a); store8(d2 + i * 8, b); store8(d3 + i * 8, c);
  }
}
// keep bit of element e = 16 random bits >= p*65536, four elements per 64-bit hash of (seed, e/4);
// channelwise (Dropout2d): one hash per (b, c) plane index.
__global__ void __launch_bounds__(256) dropout_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, int B,
                                                      long long HW, int C, int ldx, int ldy, float p,
                                                      unsigned long long seed,
                                                      const unsigned long long* __restrict__ step_counter,
                                                      int channelwise) {
  if (step_counter) seed += __ldg(step_counter) * 0x9E3779B97F4A7C15ull;
  const int CV = C / 8;
  const long long P = (long long)B * HW;
  const float scale = 1.f / (1.f - p);
  const uint32_t thr = (uint32_t)(p * 65536.0f);
  const ColIter it = col_iter(CV);
  for (long long pix = it.pix0; pix < P; pix += it.pstride) {
    float f[8];
    load8(x + pix * ldx + it.cv * 8, f);
    uint64_t r0, r1;
    if (channelwise) {
      const uint64_t base = (uint64_t)((pix / HW) * C + it.cv * 8);
      r0 = hn_hash64(seed, base >> 2);
      r1 = hn_hash64(seed, (base >> 2) + 1);
    } else {
      const uint64_t base = (uint64_t)(pix * C + it.cv * 8);
      r0 = hn_hash64(seed, base >> 2);
      r1 = hn_hash64(seed, (base >> 2) + 1);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint32_t bits = (uint32_t)((j < 4 ? r0 : r1) >> (16 * (j & 3))) & 0xFFFFu;
      f[j] = bits >= thr ? f[j] * scale : 0.f;
    }
    store8(y + pix * ldy + it.cv * 8, f);
  }
}
}  // namespace

#define ST ((cudaStream_t)stream)
#define REQ_C8(C, name) HN_REQUIRE((C) > 0 && (C) % 8 == 0, name ": C must be a positive multiple of 8 (padded), got %d", (C))
#define REQ_CV(C, name) HN_REQUIRE((C) > 0 && (C) % 8 == 0 && (768 % ((C) / 8)) == 0 && (C) <= 2048, name ": unsupported channel count %d for a column reduction", (C))

extern "C" int hn_nchw_f32_to_nhwc_bf16(const float* src, void* dst, int B, int C, int H, int W, int Cpad, int ld, void* stream) {
  HN_REQUIRE(src && dst && B > 0 && C > 0 && Cpad % 8 == 0 && Cpad >= C && ld >= Cpad && ld % 8 == 0, "nchw->nhwc: bad args");
  nchw2nhwc_kernel<<<grid_for((long long)B * H * W), TPB, 0, ST>>>(src, (bf16*)dst, B, C, H * W, Cpad, ld);
  HN_CHECK_LAUNCH("nchw2nhwc");
  return HN_OK;
}
extern "C" int hn_nhwc_bf16_to_nchw_f32(const void* src, float* dst, int B, int C, int H, int W, int ld, void* stream) {
  HN_REQUIRE(src && dst && B > 0 && C > 0 && ld % 8 == 0 && ld >= ((C + 7) & ~7), "nhwc->nchw: bad args");
  nhwc2nchw_kernel<<<grid_for((long long)B * H * W), TPB, 0, ST>>>((const bf16*)src, dst, B, C, H * W, ld);
  HN_CHECK_LAUNCH("nhwc2nchw");
  return HN_OK;
}
extern "C" int hn_avgpool2_fwd(const void* x, void* y, int B, int H, int W, int C, int ldx, int ldy, void* stream) {
  REQ_C8(C, "avgpool2");
  HN_REQUIRE(x && y && H % 2 == 0 && W % 2 == 0 && ldx % 8 == 0 && ldy % 8 == 0, "avgpool2: bad args");
  pool2_fwd_kernel<0><<<grid_for((long long)B * (H / 2) * (W / 2) * (C / 8)), TPB, 0, ST>>>((const bf16*)x, (bf16*)y, B, H / 2, W / 2, C / 8, ldx, ldy);
  HN_CHECK_LAUNCH("avgpool2");
  return HN_OK;
}
extern "C" int hn_maxpool2_fwd(const void* x, void* y, int B, int H, int W, int C, int ldx, int ldy, void* stream) {
  REQ_C8(C, "maxpool2");
  HN_REQUIRE(x && y && H % 2 == 0 && W % 2 == 0 && ldx % 8 == 0 && ldy % 8 == 0, "maxpool2: bad args");
  pool2_fwd_kernel<1><<<grid_for((long long)B * (H / 2) * (W / 2) * (C / 8)), TPB, 0, ST>>>((const bf16*)x, (bf16*)y, B, H / 2, W / 2, C / 8, ldx, ldy);
  HN_CHECK_LAUNCH("maxpool2");
  return HN_OK;
}
extern "C" int hn_maxpool2_bwd(const void* x, const void* dy, void* dx, int B, int H, int W, int C, int ldx, int lddy, int lddx, void* stream) {
  REQ_C8(C, "maxpool2_bwd");
  HN_REQUIRE(x && dy && dx && H % 2 == 0 && W % 2 == 0, "maxpool2_bwd: bad args");
  maxpool2_bwd_kernel<<<grid_for((long long)B * (H / 2) * (W / 2) * (C / 8)), TPB, 0, ST>>>((const bf16*)x, (const bf16*)dy, (bf16*)dx, B, H / 2, W / 2, C / 8, ldx, lddy, lddx);
  HN_CHECK_LAUNCH("maxpool2_bwd");
  return HN_OK;
}
extern "C" int hn_upsample2_bwd(const void* dy, void* dx, int B, int H, int W, int C, int lddy, int lddx, void* stream) {
  REQ_C8(C, "upsample2_bwd");
  HN_REQUIRE(dy && dx, "upsample2_bwd: null");
  upsample2_bwd_kernel<<<grid_for((long long)B * H * W * (C / 8)), TPB, 0, ST>>>((const bf16*)dy, (bf16*)dx, B, H, W, C / 8, lddy, lddx);
  HN_CHECK_LAUNCH("upsample2_bwd");
  return HN_OK;
}
extern "C" int hn_act_bwd(const void* dy, const void* y_or_z, void* dz, long long P, int C, int lddy, int ldy, int lddz, int act, float* dbias, void* stream) {
  REQ_CV(C, "act_bwd");
  HN_REQUIRE(dy && dz && (act == HN_ACT_NONE || y_or_z), "act_bwd: null");
  act_bwd_kernel<<<grid_for(P * (C / 8)), TPB, C * sizeof(float), ST>>>((const bf16*)dy, (const bf16*)y_or_z, (bf16*)dz, P, C, lddy, ldy, lddz, act, dbias);
  HN_CHECK_LAUNCH("act_bwd");
  return HN_OK;
}
extern "C" int hn_colsum(const void* x, long long P, int C, int ld, float* out, void* stream) {
  REQ_CV(C, "colsum");
  HN_REQUIRE(x && out, "colsum: null");
  colsum_kernel<<<grid_for(P * (C / 8)), TPB, C * sizeof(float), ST>>>((const bf16*)x, P, C, ld, out);
  HN_CHECK_LAUNCH("colsum");
  return HN_OK;
}
extern "C" int hn_add(const void* a, const void* b, const void* c, void* y, long long P, int C, int lda, int ldb, int ldc, int ldy, void* stream) {
  REQ_C8(C, "add");
  HN_REQUIRE(a && b && y, "add: null");
  add_kernel<<<grid_for(P * (C / 8)), TPB, 0, ST>>>((const bf16*)a, (const bf16*)b, (const bf16*)c, (bf16*)y, P, C / 8, lda, ldb, ldc, ldy);
  HN_CHECK_LAUNCH("add");
  return HN_OK;
}
extern "C" int hn_gelu3_sum_fwd(const void* z1, const void* z2, const void* z3, void* y, long long n, void* stream) {
  HN_REQUIRE(z1 && z2 && z3 && y && n % 8 == 0, "gelu3_sum_fwd: bad args");
  gelu3_fwd_kernel<<<grid_for(n / 8), TPB, 0, ST>>>((const bf16*)z1, (const bf16*)z2, (const bf16*)z3, (bf16*)y, n / 8);
  HN_CHECK_LAUNCH("gelu3_fwd");
  return HN_OK;
}
extern "C" int hn_gelu3_sum_bwd(const void* dy, const void* z1, const void* z2, const void* z3, void* d1, void* d2, void* d3, long long n, void* stream) {
  HN_REQUIRE(dy && z1 && z2 && z3 && d1 && d2 && d3 && n % 8 == 0, "gelu3_sum_bwd: bad args");
  gelu3_bwd_kernel<<<grid_for(n / 8), TPB, 0, ST>>>((const bf16*)dy, (const bf16*)z1, (const bf16*)z2, (const bf16*)z3, (bf16*)d1, (bf16*)d2, (bf16*)d3, n / 8);
  HN_CHECK_LAUNCH("gelu3_bwd");
  return HN_OK;
}
extern "C" int hn_dropout(const void* x, void* y, int B, long long HW, int C, int ldx, int ldy, float p, unsigned long long seed, const unsigned long long* step_counter, int channelwise, void* stream) {
  REQ_CV(C, "dropout");
  HN_REQUIRE(x && y && p >= 0.f && p < 1.f, "dropout: bad args");
  dropout_kernel<<<grid_for((long long)B * HW * (C / 8)), TPB, 0, ST>>>((const bf16*)x, (bf16*)y, B, HW, C, ldx, ldy, p, seed, step_counter, channelwise);
  HN_CHECK_LAUNCH("dropout");
  return HN_OK;
}
extern "C" int hn_act_fwd(const void* x, void* y, long long P, int C, int ldx, int ldy, int act, void* stream) {
  REQ_C8(C, "act_fwd");
  HN_REQUIRE(x && y, "act_fwd: null");
  act_fwd_kernel<<<grid_for(P * (C / 8)), TPB, 0, ST>>>((const bf16*)x, (bf16*)y, P, C / 8, ldx, ldy, act);
  HN_CHECK_LAUNCH("act_fwd");
  return HN_OK;
}
