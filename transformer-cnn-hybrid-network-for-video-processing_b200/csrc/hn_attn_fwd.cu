// Flash attention forward for the FCT conv-attention block (FCT.py:37,75):
//   O = softmax(Q K^T * scale) V   per (batch, head), never materialising N x N.
// The softmax (MUFU ex2) is the binding unit for head dims 4..16;
// QK^T / PV run on the warp-level tensor path with the head dim padded to 8/16.
#include "hn_attn.cuh"

using namespace attn;

namespace {

// RAGGED: N is not a multiple of the 64-key tile (never the case for the FCT maps): keys past the end are masked
// POLY: every POLY-th 8-key tile takes its exponentials from the FMA-pipe polynomial instead of MUFU (0 = never)
template <int D, bool RAGGED, int POLY>
__global__ void __launch_bounds__(NWARP * 32) attn_fwd_kernel(const bf16* __restrict__ q, const bf16* __restrict__ k,
                                                              const bf16* __restrict__ v, bf16* __restrict__ o,
                                                              float* __restrict__ lse, int N, int heads, int ldq, int ldk,
                                                              int ldv, int ldo, float scale) {
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

  AFrag<D, DP> qa;
  qa.load(qb, ldq, q0, N, lane);

  float oacc[DP / 8][4];
#pragma unroll
  for (int j = 0; j < DP / 8; ++j) oacc[j][0] = oacc[j][1] = oacc[j][2] = oacc[j][3] = 0.f;
  float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
  const float sl2 = scale * LOG2E;
  const int ntiles = (N + TILE - 1) / TILE;

  // D == 4: V rows are padded to 8 channels; channel 4 carries 1.0 so that output column 4 accumulates the row sum
  constexpr bool ONES = D == 4;
  stage_tile<D, DP>(sK[0], kb, ldk, 0, N, tid, NWARP * 32);
  stage_tile<D, DP, ONES>(sV[0], vb, ldv, 0, N, tid, NWARP * 32);
  cp_async_commit();

  for (int kt = 0; kt < ntiles; ++kt) {
    const int cur = kt & 1;
    if (kt + 1 < ntiles) {
      stage_tile<D, DP>(sK[cur ^ 1], kb, ldk, (kt + 1) * TILE, N, tid, NWARP * 32);
      stage_tile<D, DP, ONES>(sV[cur ^ 1], vb, ldv, (kt + 1) * TILE, N, tid, NWARP * 32);
    }
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();

    float s[TILE / 8][4];
#pragma unroll
    for (int j = 0; j < TILE / 8; ++j) {
      s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
      mma_over_d<D, DP>(s[j], qa, sK[cur], j, lane);
    }
    if (RAGGED && (kt + 1) * TILE > N) {   // ragged last tile: keys >= N never win
#pragma unroll
      for (int j = 0; j < TILE / 8; ++j) {
        int key = kt * TILE + j * 8 + 2 * t;
        if (key >= N) s[j][0] = s[j][2] = -INFINITY;
        if (key + 1 >= N) s[j][1] = s[j][3] = -INFINITY;
      }
    }
    float mx[2] = {mrow[0], mrow[1]};
#pragma unroll
    for (int j = 0; j < TILE / 8; ++j) {
      mx[0] = max3(mx[0], s[j][0], s[j][1]);
      mx[1] = max3(mx[1], s[j][2], s[j][3]);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
      mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
    }
    float corr[2], moff[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      corr[h] = ex2((mrow[h] - mx[h]) * sl2);
      mrow[h] = mx[h];
      moff[h] = -mx[h] * sl2;
      lrow[h] *= corr[h];
    }
#pragma unroll
    for (int j = 0; j < DP / 8; ++j) {
      oacc[j][0] *= corr[0]; oacc[j][1] *= corr[0];
      oacc[j][2] *= corr[1]; oacc[j][3] *= corr[1];
    }
#pragma unroll
    for (int j = 0; j < TILE / 8; ++j) {
      fma2(s[j][0], s[j][1], s[j][0], s[j][1], sl2, moff[0]);
      fma2(s[j][2], s[j][3], s[j][2], s[j][3], sl2, moff[1]);
      if (POLY > 0 && (j % (POLY > 0 ? POLY : 1)) == (POLY > 0 ? POLY : 1) - 1) {
        ex2_poly2(s[j][0], s[j][1], s[j][0], s[j][1]);
        ex2_poly2(s[j][2], s[j][3], s[j][2], s[j][3]);
      } else {
        s[j][0] = ex2(s[j][0]); s[j][1] = ex2(s[j][1]);
        s[j][2] = ex2(s[j][2]); s[j][3] = ex2(s[j][3]);
      }
      if (!ONES) {
        add2(lrow[0], lrow[1], s[j][0], s[j][2]);
        add2(lrow[0], lrow[1], s[j][1], s[j][3]);
      }
    }
#pragma unroll
    for (int kk = 0; kk < TILE / 16; ++kk) {
      uint32_t pa[4];
      pa[0] = pack_bf16(s[2 * kk][0], s[2 * kk][1]);
      pa[1] = pack_bf16(s[2 * kk][2], s[2 * kk][3]);
      pa[2] = pack_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa[3] = pack_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
      mma_over_rows<DP>(oacc, pa, sV[cur], kk, lane);
    }
    __syncthreads();
  }
  cp_async_wait<0>();

  if (ONES) {       // the row sums sit in output column 4: lane t == 2 of each row group holds them
    lrow[0] = __shfl_sync(0xffffffffu, oacc[0][0], (lane & ~3) | 2);
    lrow[1] = __shfl_sync(0xffffffffu, oacc[0][2], (lane & ~3) | 2);
  } else {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      lrow[h] += __shfl_xor_sync(0xffffffffu, lrow[h], 1);
      lrow[h] += __shfl_xor_sync(0xffffffffu, lrow[h], 2);
    }
  }
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    int row = q0 + g + h * 8;
    if (row >= N) continue;
    float inv = 1.f / lrow[h];
    bf16* op = o + ((long long)b * N + row) * ldo + head * D;
#pragma unroll
    for (int j = 0; j < DP / 8; ++j) {
      int col = j * 8 + 2 * t;
      if (col < D) *reinterpret_cast<uint32_t*>(op + col) = pack_bf16(oacc[j][h * 2] * inv, oacc[j][h * 2 + 1] * inv);
    }
    if (t == 0) lse[((long long)b * heads + head) * N + row] = mrow[h] * scale + logf(lrow[h]);
  }
}

}  // namespace

static int g_fwd_poly = 4;
extern "C" int hn_attn_tune_fwd(int poly_every) {
  HN_REQUIRE(poly_every == 0 || poly_every == 2 || poly_every == 3 || poly_every == 4 || poly_every == 8, "attn_tune_fwd: 0, 2, 3, 4 or 8");
  g_fwd_poly = poly_every;
  return HN_OK;
}

extern "C" int hn_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int N, int heads,
                           int D, int ldq, int ldk, int ldv, int ldo, float scale, void* stream) {
  HN_REQUIRE(q && k && v && o && lse && B > 0 && N > 0 && heads > 0, "attn_fwd: bad args");
  HN_REQUIRE(ldq % 2 == 0 && ldk % 4 == 0 && ldv % 4 == 0 && ldo % 2 == 0, "attn_fwd: bad strides");
  dim3 grid((N + NWARP * 16 - 1) / (NWARP * 16), heads, B);
  cudaStream_t st = (cudaStream_t)stream;
#define LAUNCH_P(DD, RG, PL)                                                                                        \
  attn_fwd_kernel<DD, RG, PL><<<grid, NWARP * 32, 0, st>>>((const bf16*)q, (const bf16*)k, (const bf16*)v, (bf16*)o, lse, \
                                                         N, heads, ldq, ldk, ldv, ldo, scale)
// head dims 4 / 8 are bound by the exponential unit: a share of the tiles goes to the polynomial (g_fwd_poly, default 4
// = every 4th tile); the larger head dims are tensor-bound and keep MUFU
#define LAUNCH(DD)                                                              \
  do {                                                                          \
    if (N % TILE) LAUNCH_P(DD, true, 0);                                        \
    else if (DD > 8 || g_fwd_poly == 0) LAUNCH_P(DD, false, 0);                 \
    else if (g_fwd_poly == 2) LAUNCH_P(DD, false, 2);                           \
    else if (g_fwd_poly == 3) LAUNCH_P(DD, false, 3);                           \
    else if (g_fwd_poly == 8) LAUNCH_P(DD, false, 8);                           \
    else LAUNCH_P(DD, false, 4);                                                \
  } while (0)
  switch (D) {
    case 4: LAUNCH(4); break;
    case 8: HN_REQUIRE(ldk % 8 == 0 && ldv % 8 == 0, "attn_fwd: D>=8 needs 16B-aligned rows"); LAUNCH(8); break;
    case 16: HN_REQUIRE(ldk % 8 == 0 && ldv % 8 == 0, "attn_fwd: D>=8 needs 16B-aligned rows"); LAUNCH(16); break;
    case 32: HN_REQUIRE(ldk % 8 == 0 && ldv % 8 == 0, "attn_fwd: D>=8 needs 16B-aligned rows"); LAUNCH(32); break;
    case 64: HN_REQUIRE(ldk % 8 == 0 && ldv % 8 == 0, "attn_fwd: D>=8 needs 16B-aligned rows"); LAUNCH(64); break;
    default: HN_FAIL(HN_EUNSUPPORTED, "attn_fwd: head dim %d not in {4,8,16,32,64}", D);
  }
#undef LAUNCH
  HN_CHECK_LAUNCH("attn_fwd");
  return HN_OK;
}
