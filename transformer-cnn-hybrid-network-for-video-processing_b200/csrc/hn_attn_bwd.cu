// Flash attention backward (FCT.py:75 under autograd).  Three kernels:
//   delta[n] = sum_j dO[n,j] O[n,j]
//   dQ  : query-stationary, streams K/V tiles   (dS = P o (dP - delta); dQ += dS K)
//   dKV : key-stationary,   streams Q/dO tiles  (dV += P^T dO; dK += dS^T Q)
// P is recomputed from the saved log-sum-exp; nothing N x N touches HBM and
// there are no atomics (deterministic gradients).
#include "hn_attn.cuh"

using namespace attn;

namespace {

__global__ void attn_delta_kernel(const bf16* __restrict__ o, const bf16* __restrict__ dout, float* __restrict__ delta,
                                  int N, int heads, int D, int ldo, int lddo, long long total) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= total) return;
  int n = (int)(i % N);
  long long r = i / N;
  int head = (int)(r % heads);
  long long b = r / heads;
  const bf16* po = o + (b * N + n) * ldo + head * D;
  const bf16* pd = dout + (b * N + n) * lddo + head * D;
  float s = 0.f;
  for (int j = 0; j < D; j += 2) {
    float2 a = __bfloat1622float2(*reinterpret_cast<const bf162*>(po + j));
    float2 c = __bfloat1622float2(*reinterpret_cast<const bf162*>(pd + j));
    s += a.x * c.x + a.y * c.y;
  }
  delta[i] = s;
}

template <int D>
__global__ void __launch_bounds__(NWARP * 32) attn_bwd_dq_kernel(
    const bf16* __restrict__ q, const bf16* __restrict__ k, const bf16* __restrict__ v, const bf16* __restrict__ dout,
    const float* __restrict__ lse, const float* __restrict__ delta, bf16* __restrict__ dq, int N, int heads, int ldq,
    int ldk, int ldv, int lddo, int lddq, float scale) {
  constexpr int DP = D < 8 ? 8 : D;
  constexpr int TB = TILE * DP * 2;
  __shared__ __align__(128) uint8_t sK[2][TB];
  __shared__ __align__(128) uint8_t sV[2][TB];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int head = blockIdx.y, b = blockIdx.z;
  const int q0 = blockIdx.x * (NWARP * 16) + warp * 16;
  const bf16* qb = q + (long long)b * N * ldq + head * D;
  const bf16* kb = k + (long long)b * N * ldk + head * D;
  const bf16* vb = v + (long long)b * N * ldv + head * D;
  const bf16* dob = dout + (long long)b * N * lddo + head * D;
  const long long stat_base = ((long long)b * heads + head) * N;

  AFrag<D, DP> qa, doa;
  qa.load(qb, ldq, q0, N, lane);
  doa.load(dob, lddo, q0, N, lane);
  float l2[2], dl[2];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    int row = q0 + g + h * 8;
    l2[h] = row < N ? lse[stat_base + row] * LOG2E : 0.f;
    dl[h] = row < N ? delta[stat_base + row] : 0.f;
  }
  float acc[DP / 8][4];
#pragma unroll
  for (int j = 0; j < DP / 8; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
  const float sl2 = scale * LOG2E;
  const int ntiles = (N + TILE - 1) / TILE;

  stage_tile<D, DP>(sK[0], kb, ldk, 0, N, tid, NWARP * 32);
  stage_tile<D, DP>(sV[0], vb, ldv, 0, N, tid, NWARP * 32);
  cp_async_commit();
  for (int kt = 0; kt < ntiles; ++kt) {
    const int cur = kt & 1;
    if (kt + 1 < ntiles) {
      stage_tile<D, DP>(sK[cur ^ 1], kb, ldk, (kt + 1) * TILE, N, tid, NWARP * 32);
      stage_tile<D, DP>(sV[cur ^ 1], vb, ldv, (kt + 1) * TILE, N, tid, NWARP * 32);
    }
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    float ds[TILE / 8][4];
    const bool ragged = (kt + 1) * TILE > N;
#pragma unroll
    for (int j = 0; j < TILE / 8; ++j) {
      float s[4] = {0.f, 0.f, 0.f, 0.f}, dp[4] = {0.f, 0.f, 0.f, 0.f};
      mma_over_d<D, DP>(s, qa, sK[cur], j, lane);
      mma_over_d<D, DP>(dp, doa, sV[cur], j, lane);
      float p0 = ex2(s[0] * sl2 - l2[0]);
      float p1 = ex2(s[1] * sl2 - l2[0]);
      float p2 = ex2(s[2] * sl2 - l2[1]);
      float p3 = ex2(s[3] * sl2 - l2[1]);
      if (ragged) {      // warp-uniform: only the last key tile of an N that is not a multiple of 64
        int key = kt * TILE + j * 8 + 2 * t;
        if (key >= N) p0 = p2 = 0.f;
        if (key + 1 >= N) p1 = p3 = 0.f;
      }
      ds[j][0] = p0 * (dp[0] - dl[0]);
      ds[j][1] = p1 * (dp[1] - dl[0]);
      ds[j][2] = p2 * (dp[2] - dl[1]);
      ds[j][3] = p3 * (dp[3] - dl[1]);
    }
#pragma unroll
    for (int kk = 0; kk < TILE / 16; ++kk) {
      uint32_t pa[4];
      pa[0] = pack_bf16(ds[2 * kk][0], ds[2 * kk][1]);
      pa[1] = pack_bf16(ds[2 * kk][2], ds[2 * kk][3]);
      pa[2] = pack_bf16(ds[2 * kk + 1][0], ds[2 * kk + 1][1]);
      pa[3] = pack_bf16(ds[2 * kk + 1][2], ds[2 * kk + 1][3]);
      mma_over_rows<DP>(acc, pa, sK[cur], kk, lane);
    }
    __syncthreads();
  }
  cp_async_wait<0>();
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    int row = q0 + g + h * 8;
    if (row >= N) continue;
    bf16* op = dq + ((long long)b * N + row) * lddq + head * D;
#pragma unroll
    for (int j = 0; j < DP / 8; ++j) {
      int col = j * 8 + 2 * t;
      if (col < D) *reinterpret_cast<uint32_t*>(op + col) = pack_bf16(acc[j][h * 2] * scale, acc[j][h * 2 + 1] * scale);
    }
  }
}

template <int D>
__global__ void __launch_bounds__(NWARP * 32) attn_bwd_dkv_kernel(
    const bf16* __restrict__ q, const bf16* __restrict__ k, const bf16* __restrict__ v, const bf16* __restrict__ dout,
    const float* __restrict__ lse, const float* __restrict__ delta, bf16* __restrict__ dk, bf16* __restrict__ dv, int N,
    int heads, int ldq, int ldk, int ldv, int lddo, int lddk, int lddv, float scale) {
  constexpr int DP = D < 8 ? 8 : D;
  constexpr int TB = TILE * DP * 2;
  __shared__ __align__(128) uint8_t sQ[2][TB];
  __shared__ __align__(128) uint8_t sDO[2][TB];
  __shared__ __align__(8) float sL[2][TILE], sDl[2][TILE];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int head = blockIdx.y, b = blockIdx.z;
  const int k0 = blockIdx.x * (NWARP * 16) + warp * 16;
  const bf16* qb = q + (long long)b * N * ldq + head * D;
  const bf16* kb = k + (long long)b * N * ldk + head * D;
  const bf16* vb = v + (long long)b * N * ldv + head * D;
  const bf16* dob = dout + (long long)b * N * lddo + head * D;
  const long long stat_base = ((long long)b * heads + head) * N;

  AFrag<D, DP> ka, va;
  ka.load(kb, ldk, k0, N, lane);
  va.load(vb, ldv, k0, N, lane);
  float adk[DP / 8][4], adv[DP / 8][4];
#pragma unroll
  for (int j = 0; j < DP / 8; ++j) {
    adk[j][0] = adk[j][1] = adk[j][2] = adk[j][3] = 0.f;
    adv[j][0] = adv[j][1] = adv[j][2] = adv[j][3] = 0.f;
  }
  const float sl2 = scale * LOG2E;
  const int ntiles = (N + TILE - 1) / TILE;

  auto stage = [&](int buf, int qt) {
    stage_tile<D, DP>(sQ[buf], qb, ldq, qt * TILE, N, tid, NWARP * 32);
    stage_tile<D, DP>(sDO[buf], dob, lddo, qt * TILE, N, tid, NWARP * 32);
    if (tid < TILE) {
      int row = qt * TILE + tid;
      sL[buf][tid] = row < N ? lse[stat_base + row] * LOG2E : INFINITY;   // +inf -> P = 0 for ragged queries
      sDl[buf][tid] = row < N ? delta[stat_base + row] : 0.f;
    }
  };
  stage(0, 0);
  cp_async_commit();
  for (int qt = 0; qt < ntiles; ++qt) {
    const int cur = qt & 1;
    if (qt + 1 < ntiles) stage(cur ^ 1, qt + 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    float pt[TILE / 8][4], dst[TILE / 8][4];
#pragma unroll
    for (int j = 0; j < TILE / 8; ++j) {
      float s[4] = {0.f, 0.f, 0.f, 0.f}, dp[4] = {0.f, 0.f, 0.f, 0.f};
      mma_over_d<D, DP>(s, ka, sQ[cur], j, lane);      // S^T tile: rows = keys, cols = queries
      mma_over_d<D, DP>(dp, va, sDO[cur], j, lane);    // dP^T
      const int qc = j * 8 + 2 * t;
      const float2 lv = *reinterpret_cast<const float2*>(&sL[cur][qc]);
      const float2 dv2 = *reinterpret_cast<const float2*>(&sDl[cur][qc]);
      const float la = lv.x, lb = lv.y, da = dv2.x, db = dv2.y;
      pt[j][0] = ex2(s[0] * sl2 - la);
      pt[j][1] = ex2(s[1] * sl2 - lb);
      pt[j][2] = ex2(s[2] * sl2 - la);
      pt[j][3] = ex2(s[3] * sl2 - lb);
      dst[j][0] = pt[j][0] * (dp[0] - da);
      dst[j][1] = pt[j][1] * (dp[1] - db);
      dst[j][2] = pt[j][2] * (dp[2] - da);
      dst[j][3] = pt[j][3] * (dp[3] - db);
    }
#pragma unroll
    for (int kk = 0; kk < TILE / 16; ++kk) {
      uint32_t pa[4];
      pa[0] = pack_bf16(pt[2 * kk][0], pt[2 * kk][1]);
      pa[1] = pack_bf16(pt[2 * kk][2], pt[2 * kk][3]);
      pa[2] = pack_bf16(pt[2 * kk + 1][0], pt[2 * kk + 1][1]);
      pa[3] = pack_bf16(pt[2 * kk + 1][2], pt[2 * kk + 1][3]);
      mma_over_rows<DP>(adv, pa, sDO[cur], kk, lane);
      pa[0] = pack_bf16(dst[2 * kk][0], dst[2 * kk][1]);
      pa[1] = pack_bf16(dst[2 * kk][2], dst[2 * kk][3]);
      pa[2] = pack_bf16(dst[2 * kk + 1][0], dst[2 * kk + 1][1]);
      pa[3] = pack_bf16(dst[2 * kk + 1][2], dst[2 * kk + 1][3]);
      mma_over_rows<DP>(adk, pa, sQ[cur], kk, lane);
    }
    __syncthreads();
  }
  cp_async_wait<0>();
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    int row = k0 + g + h * 8;
    if (row >= N) continue;
    bf16* pk = dk + ((long long)b * N + row) * lddk + head * D;
    bf16* pv = dv + ((long long)b * N + row) * lddv + head * D;
#pragma unroll
    for (int j = 0; j < DP / 8; ++j) {
      int col = j * 8 + 2 * t;
      if (col < D) {
        *reinterpret_cast<uint32_t*>(pk + col) = pack_bf16(adk[j][h * 2] * scale, adk[j][h * 2 + 1] * scale);
        *reinterpret_cast<uint32_t*>(pv + col) = pack_bf16(adv[j][h * 2], adv[j][h * 2 + 1]);
      }
    }
  }
}

}  // namespace

extern "C" int hn_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout,
                           const float* lse, float* delta_ws, void* dq, void* dk, void* dv, int B, int N, int heads,
                           int D, int ldq, int ldk, int ldv, int ldo, int lddo, int lddq, int lddk, int lddv,
                           float scale, void* stream) {
  HN_REQUIRE(q && k && v && o && dout && lse && delta_ws && dq && dk && dv, "attn_bwd: null pointer");
  HN_REQUIRE(D == 4 || D == 8 || D == 16 || D == 32 || D == 64, "attn_bwd: head dim %d not in {4,8,16,32,64}", D);
  int a = D >= 8 ? 8 : 4;
  HN_REQUIRE(ldq % a == 0 && ldk % a == 0 && ldv % a == 0 && lddo % a == 0 && ldo % 2 == 0 && lddq % 2 == 0 && lddk % 2 == 0 && lddv % 2 == 0,
             "attn_bwd: strides must keep rows 16B (8B for D=4) aligned");
  cudaStream_t st = (cudaStream_t)stream;
  long long total = (long long)B * heads * N;
  attn_delta_kernel<<<hn_cdiv(total, 256), 256, 0, st>>>((const bf16*)o, (const bf16*)dout, delta_ws, N, heads, D, ldo, lddo, total);
  HN_CHECK_LAUNCH("attn_delta");
  dim3 grid((N + NWARP * 16 - 1) / (NWARP * 16), heads, B);
#define LAUNCH(DD)                                                                                                     \
  attn_bwd_dq_kernel<DD><<<grid, NWARP * 32, 0, st>>>((const bf16*)q, (const bf16*)k, (const bf16*)v, (const bf16*)dout, \
                                                      lse, delta_ws, (bf16*)dq, N, heads, ldq, ldk, ldv, lddo, lddq,    \
                                                      scale);                                                          \
  attn_bwd_dkv_kernel<DD><<<grid, NWARP * 32, 0, st>>>((const bf16*)q, (const bf16*)k, (const bf16*)v,                  \
                                                       (const bf16*)dout, lse, delta_ws, (bf16*)dk, (bf16*)dv, N,      \
                                                       heads, ldq, ldk, ldv, lddo, lddk, lddv, scale)
  switch (D) {
    case 4: LAUNCH(4); break;
    case 8: LAUNCH(8); break;
    case 16: LAUNCH(16); break;
    case 32: LAUNCH(32); break;
    default: LAUNCH(64); break;
  }
#undef LAUNCH
  HN_CHECK_LAUNCH("attn_bwd");
  return HN_OK;
}
