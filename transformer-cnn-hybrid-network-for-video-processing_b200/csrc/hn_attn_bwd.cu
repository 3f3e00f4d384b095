// Flash attention backward (FCT.py:75 under autograd).  P is recomputed from the saved log-sum-exp; nothing
// N x N touches HBM.
//   delta[n] = sum_j dO[n,j] O[n,j]                                              (all head dims)
// Head dims 4 and 8 (99 % of FCT's score elements, bound by the exponential unit): ONE pass over the score tiles
// (attn_bwd_fused_kernel below), dQ through an fp32 accumulator.
// Head dims 16..64 (small maps, tensor-bound): two deterministic kernels without atomics
//   dQ  : query-stationary, streams K/V tiles   (dS = P o (dP - delta); dQ += dS K)
//   dKV : key-stationary,   streams Q/dO tiles  (dV += P^T dO; dK += dS^T Q)
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


// ---- one-pass backward for the MUFU-bound head dims (D = 4, 8; 99 % of FCT's score elements) -----------------
// The two-kernel form above recomputes P = exp2(S*c - lse) twice (once for dQ, once for dK/dV): 2*N^2 exponentials
// per head, and the exponential unit is what binds these shapes.  Here every score tile is visited ONCE:
//   CTA  = KB keys of one (batch, head): K, V stationary in shared memory; dK, dV accumulate in registers
//   warp = walks the queries in steps of 16 rows (Q, dO rows and their lse / delta straight from global memory into
//          A fragments, prefetched one step ahead -- no shared-memory staging, no CTA barrier in the loop)
//   per 16-query x 8-key tile:  S = Q K^T, dP = dO V^T (m16n8k8), P = exp2(S*c - lse), dS = P (dP - delta)
//          dQ (16 x D) += dS K           A = packed dS fragments as they are
//          dV (keys x D) += P^T dO       A = P^T  through movmatrix.trans of the same packed fragments
//          dK (keys x D) += dS^T Q       A = dS^T likewise; B = movmatrix.trans of the Q / dO A fragments
//   dQ leaves with vector red.global.add.f32 into an fp32 accumulator (one 16 x D patch per warp per step, the
//   steps of different CTAs rotated so they do not hit the same rows at the same time); dK/dV are reduced over the
//   four warps in shared memory once at the end.
__device__ __forceinline__ uint32_t movm_t(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}
__device__ __forceinline__ void red_add_v2(float* addr, float a, float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

template <int D, int KB, bool RAGGED>
__global__ void __launch_bounds__(NWARP * 32, KB == 128 ? 4 : 6) attn_bwd_fused_kernel(
    const bf16* __restrict__ q, const bf16* __restrict__ k, const bf16* __restrict__ v, const bf16* __restrict__ dout,
    const float* __restrict__ lse, const float* __restrict__ delta, float* __restrict__ dq_acc, bf16* __restrict__ dk,
    bf16* __restrict__ dv, int N, int heads, int ldq, int ldk, int ldv, int lddo, int lddk, int lddv, int C, float scale, int kblock0) {
  constexpr int DP = 8;
  static_assert(D == 4 || D == 8, "fused backward: head dims 4 and 8");
  constexpr int MT = KB / 16;                         // 16-key m-tiles of the dK / dV accumulators
  __shared__ __align__(128) uint8_t sK[KB * 16];
  __shared__ __align__(128) uint8_t sV[KB * 16];
  __shared__ float sAcc[2][KB][DP];                   // cross-warp reduction of dK / dV
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int head = blockIdx.y, b = blockIdx.z;
  const int k0 = (kblock0 + blockIdx.x) * KB;
  const bf16* qb = q + (long long)b * N * ldq + head * D;
  const bf16* kb = k + (long long)b * N * ldk + head * D;
  const bf16* vb = v + (long long)b * N * ldv + head * D;
  const bf16* dob = dout + (long long)b * N * lddo + head * D;
  const long long stat_base = ((long long)b * heads + head) * N;
  float* dqb = dq_acc + (long long)b * N * C + head * D;

  // stationary K / V block: [KB][8] bf16 rows of 16 bytes, rows >= N and channels >= D zero
  for (int r = tid; r < KB; r += NWARP * 32) {
    uint4 kv = make_uint4(0u, 0u, 0u, 0u), vv = kv;
    if (k0 + r < N) {
      if (D == 8) {
        kv = *reinterpret_cast<const uint4*>(kb + (long long)(k0 + r) * ldk);
        vv = *reinterpret_cast<const uint4*>(vb + (long long)(k0 + r) * ldv);
      } else {
        const uint2 a = *reinterpret_cast<const uint2*>(kb + (long long)(k0 + r) * ldk);
        const uint2 c = *reinterpret_cast<const uint2*>(vb + (long long)(k0 + r) * ldv);
        kv = make_uint4(a.x, a.y, 0u, 0u);
        vv = make_uint4(c.x, c.y, 0u, 0u);
      }
    }
    *reinterpret_cast<uint4*>(sK + r * 16) = kv;
    *reinterpret_cast<uint4*>(sV + r * 16) = vv;
  }
  for (int i = tid; i < 2 * KB * DP; i += NWARP * 32) (&sAcc[0][0][0])[i] = 0.f;
  __syncthreads();

  float adk[MT][4], adv[MT][4];
#pragma unroll
  for (int m = 0; m < MT; ++m) {
    adk[m][0] = adk[m][1] = adk[m][2] = adk[m][3] = 0.f;
    adv[m][0] = adv[m][1] = adv[m][2] = adv[m][3] = 0.f;
  }
  const float sl2 = scale * LOG2E;
  const int nsteps = (N + 15) / 16;
  // rotate the query sweep per CTA: CTAs that run at the same time add into different dQ rows
  const int rot = (int)(((long long)(kblock0 + blockIdx.x) * 61 + blockIdx.z * 7) % nsteps);

  AFrag<D, DP> qa, doa, qn, don;
  float l2[2], dl[2], l2n[2], dln[2];
  auto fetch = [&](int step, AFrag<D, DP>& fq, AFrag<D, DP>& fdo, float* fl, float* fd) {
    const int q0 = step * 16;
    fq.load(qb, ldq, q0, N, lane);
    fdo.load(dob, lddo, q0, N, lane);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int row = q0 + g + h * 8;
      fl[h] = row < N ? lse[stat_base + row] * LOG2E : INFINITY;      // +inf -> P = 0 for query rows past the end
      fd[h] = row < N ? delta[stat_base + row] : 0.f;
    }
  };
  int it = warp;
  if (it < nsteps) {
    int st = it + rot; if (st >= nsteps) st -= nsteps;
    fetch(st, qn, don, l2n, dln);
  }
  for (; it < nsteps; it += NWARP) {
    int step = it + rot; if (step >= nsteps) step -= nsteps;
    qa = qn; doa = don;
    l2[0] = l2n[0]; l2[1] = l2n[1]; dl[0] = dln[0]; dl[1] = dln[1];
    if (it + NWARP < nsteps) {
      int st = it + NWARP + rot; if (st >= nsteps) st -= nsteps;
      fetch(st, qn, don, l2n, dln);
    }
    // B operands of the dK / dV products: [16 queries x 8 channels], k = query
    const uint32_t qb0 = movm_t(qa.r[0][0]), qb1 = movm_t(qa.r[0][1]);
    const uint32_t dob0 = movm_t(doa.r[0][0]), dob1 = movm_t(doa.r[0][1]);
    float acc[1][4] = {{0.f, 0.f, 0.f, 0.f}};
#pragma unroll
    for (int jp = 0; jp < MT; ++jp) {
      uint32_t pp[2][2], pd[2][2];
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        const int j = 2 * jp + hf;
        float s[4] = {0.f, 0.f, 0.f, 0.f}, dp[4] = {0.f, 0.f, 0.f, 0.f};
        mma_over_d<D, DP>(s, qa, sK, j, lane);
        mma_over_d<D, DP>(dp, doa, sV, j, lane);
        float x0, x1, x2, x3;
        fma2(x0, x1, s[0], s[1], sl2, -l2[0]);
        fma2(x2, x3, s[2], s[3], sl2, -l2[1]);
        float p0 = ex2(x0), p1 = ex2(x1), p2 = ex2(x2), p3 = ex2(x3);
        if (RAGGED) {               // the last key block of an N that is not a multiple of KB (its own launch)
          const int key = k0 + j * 8 + 2 * t;
          if (key >= N) p0 = p2 = 0.f;
          if (key + 1 >= N) p1 = p3 = 0.f;
        }
        float e0, e1, e2, e3;
        sub2(e0, e1, dp[0], dp[1], dl[0]);
        sub2(e2, e3, dp[2], dp[3], dl[1]);
        mul2(e0, e1, e0, e1, p0, p1);
        mul2(e2, e3, e2, e3, p2, p3);
        pp[hf][0] = pack_bf16(p0, p1);
        pp[hf][1] = pack_bf16(p2, p3);
        pd[hf][0] = pack_bf16(e0, e1);
        pd[hf][1] = pack_bf16(e2, e3);
      }
      // dQ += dS[16 q x 16 keys] K[16 keys x D]
      const uint32_t da[4] = {pd[0][0], pd[0][1], pd[1][0], pd[1][1]};
      mma_over_rows<DP>(acc, da, sK, jp, lane);
      // dV[16 keys x D] += P^T[16 keys x 16 q] dO[16 q x D];  dK likewise with dS^T and Q
      const uint32_t pt[4] = {movm_t(pp[0][0]), movm_t(pp[1][0]), movm_t(pp[0][1]), movm_t(pp[1][1])};
      mma_16816(adv[jp], pt, dob0, dob1);
      const uint32_t dt[4] = {movm_t(pd[0][0]), movm_t(pd[1][0]), movm_t(pd[0][1]), movm_t(pd[1][1])};
      mma_16816(adk[jp], dt, qb0, qb1);
    }
    // this warp's 16 x D patch of dQ (sum over the CTA's KB keys)
    if (2 * t < D) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int row = step * 16 + g + h * 8;
        if (row < N) red_add_v2(dqb + (long long)row * C + 2 * t, acc[0][2 * h], acc[0][2 * h + 1]);
      }
    }
  }
  // dK / dV: sum the four warps' partials
#pragma unroll
  for (int m = 0; m < MT; ++m) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int row = m * 16 + g + h * 8;
      atomicAdd(&sAcc[0][row][2 * t], adk[m][2 * h]);
      atomicAdd(&sAcc[0][row][2 * t + 1], adk[m][2 * h + 1]);
      atomicAdd(&sAcc[1][row][2 * t], adv[m][2 * h]);
      atomicAdd(&sAcc[1][row][2 * t + 1], adv[m][2 * h + 1]);
    }
  }
  __syncthreads();
  for (int i = tid; i < KB * (D / 2); i += NWARP * 32) {
    const int row = i / (D / 2), c2 = (i - row * (D / 2)) * 2;
    if (k0 + row >= N) continue;
    *reinterpret_cast<uint32_t*>(dk + ((long long)b * N + k0 + row) * lddk + head * D + c2) =
        pack_bf16(sAcc[0][row][c2] * scale, sAcc[0][row][c2 + 1] * scale);
    *reinterpret_cast<uint32_t*>(dv + ((long long)b * N + k0 + row) * lddv + head * D + c2) =
        pack_bf16(sAcc[1][row][c2], sAcc[1][row][c2 + 1]);
  }
}

// dq = bf16(scale * dq_acc): the fp32 accumulator [B*N, C] of the fused kernel -> the NHWC bf16 gradient
__global__ void attn_dq_cast_kernel(const float* __restrict__ acc, bf16* __restrict__ dq, long long rows, int C, int lddq, float scale) {
  const int cv = C / 2;
  const long long total = rows * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cv;
    const int c = (int)(i - r * cv) * 2;
    const float2 a = *reinterpret_cast<const float2*>(acc + r * C + c);
    *reinterpret_cast<uint32_t*>(dq + r * lddq + c) = pack_bf16(a.x * scale, a.y * scale);
  }
}

}  // namespace

// keys per CTA of the one-pass kernel: 128 (fewer dQ reductions) or 64 (fewer registers, more resident warps)
static int g_fused_kb = 128;
extern "C" int hn_attn_tune(int fused_bwd_keys_per_cta) {
  HN_REQUIRE(fused_bwd_keys_per_cta == 64 || fused_bwd_keys_per_cta == 128, "attn_tune: keys per CTA must be 64 or 128");
  g_fused_kb = fused_bwd_keys_per_cta;
  return HN_OK;
}

extern "C" long long hn_attn_bwd_workspace(int B, int N, int heads, int D) {
  // delta [B, heads, N] (+ the fp32 dQ accumulator [B, N, heads*D] of the one-pass kernel for D <= 8)
  long long n = ((long long)B * heads * N + 3) / 4 * 4;
  if (D <= 8) n += (long long)B * N * heads * D;
  return (n + 3) / 4 * 4;
}

extern "C" int hn_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout,
                           const float* lse, float* ws, long long ws_floats, void* dq, void* dk, void* dv, int B, int N,
                           int heads, int D, int ldq, int ldk, int ldv, int ldo, int lddo, int lddq, int lddk, int lddv,
                           float scale, void* stream) {
  HN_REQUIRE(q && k && v && o && dout && lse && ws && dq && dk && dv, "attn_bwd: null pointer");
  HN_REQUIRE(D == 4 || D == 8 || D == 16 || D == 32 || D == 64, "attn_bwd: head dim %d not in {4,8,16,32,64}", D);
  HN_REQUIRE(ws_floats >= hn_attn_bwd_workspace(B, N, heads, D), "attn_bwd: workspace too small (hn_attn_bwd_workspace)");
  int a = D >= 8 ? 8 : 4;
  HN_REQUIRE(ldq % a == 0 && ldk % a == 0 && ldv % a == 0 && lddo % a == 0 && ldo % 2 == 0 && lddq % 2 == 0 && lddk % 2 == 0 && lddv % 2 == 0,
             "attn_bwd: strides must keep rows 16B (8B for D=4) aligned");
  cudaStream_t st = (cudaStream_t)stream;
  float* delta_ws = ws;
  long long total = (long long)B * heads * N;
  attn_delta_kernel<<<hn_cdiv(total, 256), 256, 0, st>>>((const bf16*)o, (const bf16*)dout, delta_ws, N, heads, D, ldo, lddo, total);
  HN_CHECK_LAUNCH("attn_delta");
  if (D <= 8) {
    // one-pass kernel: P recomputed once; dQ through an fp32 accumulator
    const int C = heads * D;
    float* dq_acc = ws + (total + 3) / 4 * 4;
    cudaError_t e = cudaMemsetAsync(dq_acc, 0, sizeof(float) * (size_t)B * N * C, st);
    if (e != cudaSuccess) HN_FAIL(HN_ECUDA, "attn_bwd: memset: %s", cudaGetErrorString(e));
    const int KB = g_fused_kb;
    const int full = N / KB;                        // key blocks without a ragged edge
#define LAUNCHF(DD, KBB, RG, GX, K0)                                                                                    \
  attn_bwd_fused_kernel<DD, KBB, RG><<<dim3(GX, heads, B), NWARP * 32, 0, st>>>(                                        \
      (const bf16*)q, (const bf16*)k, (const bf16*)v, (const bf16*)dout, lse, delta_ws, dq_acc, (bf16*)dk, (bf16*)dv, N, \
      heads, ldq, ldk, ldv, lddo, lddk, lddv, C, scale, K0)
#define LAUNCHD(KBB, RG, GX, K0)                                                  \
  do {                                                                            \
    if (D == 4) LAUNCHF(4, KBB, RG, GX, K0); else LAUNCHF(8, KBB, RG, GX, K0);    \
  } while (0)
    if (full > 0) {
      if (KB == 128) LAUNCHD(128, false, full, 0); else LAUNCHD(64, false, full, 0);
    }
    if (N % KB) {
      if (KB == 128) LAUNCHD(128, true, 1, full); else LAUNCHD(64, true, 1, full);
    }
#undef LAUNCHD
#undef LAUNCHF
    HN_CHECK_LAUNCH("attn_bwd_fused");
    const long long rows = (long long)B * N;
    long long blocks = (rows * (C / 2) + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    attn_dq_cast_kernel<<<(int)blocks, 256, 0, st>>>(dq_acc, (bf16*)dq, rows, C, lddq, scale);
    HN_CHECK_LAUNCH("attn_dq_cast");
    return HN_OK;
  }
  dim3 grid((N + NWARP * 16 - 1) / (NWARP * 16), heads, B);
#define LAUNCH(DD)                                                                                                     \
  attn_bwd_dq_kernel<DD><<<grid, NWARP * 32, 0, st>>>((const bf16*)q, (const bf16*)k, (const bf16*)v, (const bf16*)dout, \
                                                      lse, delta_ws, (bf16*)dq, N, heads, ldq, ldk, ldv, lddo, lddq,    \
                                                      scale);                                                          \
  attn_bwd_dkv_kernel<DD><<<grid, NWARP * 32, 0, st>>>((const bf16*)q, (const bf16*)k, (const bf16*)v,                  \
                                                       (const bf16*)dout, lse, delta_ws, (bf16*)dk, (bf16*)dv, N,      \
                                                       heads, ldq, ldk, ldv, lddo, lddk, lddv, scale)
  switch (D) {
    case 16: LAUNCH(16); break;
    case 32: LAUNCH(32); break;
    default: LAUNCH(64); break;
  }
#undef LAUNCH
  HN_CHECK_LAUNCH("attn_bwd");
  return HN_OK;
}
