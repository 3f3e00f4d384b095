// Small-channel weight gradient (same layers as hn_conv_sc.cu: Cin, Cout <= 64, stride 1):
//   dW[n][c][kh][kw] += sum_px dz[px][n] * x[px @ (kh,kw)][c]
// HBM-bound, so again built around traffic: a persistent CTA stages, per 16x32-pixel tile, the input
// halo and the dz tile ONCE in shared memory; the contraction over pixels runs on mma.sync m16n8k16
// with both operands read through ldmatrix.trans (pixels are the K dimension).  The 8 warps form a
// CGRP x KSPLIT grid: the (tap, 8-channel) column chunks of dW are dealt round-robin to CGRP warp
// columns, the 32 pixel k-steps of a tile to KSPLIT warp rows; every warp keeps its slice of dW in
// registers across all tiles of the CTA; at the end the slices are summed in shared memory and
// flushed with one fp32 global atomic per element per CTA.
#include "hn_common.cuh"

namespace {

constexpr int TH = 16, TW = 32;
constexpr int WS_THREADS = 256;

struct WsP {
  const bf16* x1;
  const bf16* x2;
  const bf16* dz;
  float* dw;
  int B, H, W, C1, C2, ld1, ld2, lddz, Cout, Creal;
  int KH, KW, pad, dil, ups;             // ups: x1 is the H/2 x W/2 map read through nearest-neighbour 2x upsampling
  int CG, CG1, lgCG, sh;                 // input 8-channel chunks
  int NC, lgNC, shn;                     // dz 8-channel chunks
  int Kc;                                // taps * CG
  int cgrp, lgcgrp, ksplit;              // warp grid: chunk columns x pixel rows (cgrp * ksplit = 8)
  int HR, WR;
  int tiles_w, tiles_h, tiles;
  long long s_n, s_c, s_kh, s_kw;
};

template <int MTN, int CPW>   // 16-row blocks of Cout, column chunks per warp
__global__ void __launch_bounds__(WS_THREADS, 2) wgrad_sc_kernel(const WsP p) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int s_tap[9];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int CG = p.CG, cgm = CG - 1, NC = p.NC, ncm = NC - 1;
  const int halo_chunks = p.HR * p.WR * CG;
  uint8_t* sH = smem;
  uint8_t* sD = smem + halo_chunks * 16;

  if (tid < p.KH * p.KW) {
    const int kh = tid / p.KW, kw = tid - kh * p.KW;
    s_tap[tid] = ((kh * p.dil) << 16) | (kw * p.dil);
  }
  float acc[MTN][CPW][4];
#pragma unroll
  for (int i = 0; i < MTN; ++i)
#pragma unroll
    for (int j = 0; j < CPW; ++j) acc[i][j][0] = acc[i][j][1] = acc[i][j][2] = acc[i][j][3] = 0.f;

  // this warp's column chunks: wc, wc+cgrp, ...  (chunk -> tap, c8) and pixel k-steps wk, wk+ksplit, ...
  const int wc = warp & (p.cgrp - 1), wk = warp >> p.lgcgrp;
  int c_dr[CPW], c_dc[CPW], c_c8[CPW];
  bool c_ok[CPW];
  __syncthreads();
#pragma unroll
  for (int j = 0; j < CPW; ++j) {
    const int ch = wc + p.cgrp * j;
    c_ok[j] = ch < p.Kc;
    const int tap = c_ok[j] ? (ch >> p.lgCG) : 0;
    c_c8[j] = ch & cgm;
    c_dr[j] = s_tap[tap] >> 16;
    c_dc[j] = s_tap[tap] & 0xFFFF;
  }

  // Shared-memory byte offsets of this lane's ldmatrix rows, without the tile-row term: the swizzle key of a halo pixel
  // depends only on its column, the key of a dz pixel only on the low bits of its index, so everything but
  // (row * pitch) is a per-warp constant.  Computed per k-step (with run-time shifts) this address arithmetic was 17
  // instructions per HMMA: 25 % IMAD, 12 % SHF, 12 % LOP3 against 6 % HMMA in the ncu instruction mix of the C = 32 layers.
  const uint32_t sH_addr = smem_u32(sH), sD_addr = smem_u32(sD);
  // (the key uses column bits below 4 only, so the second 16-pixel half of a row is a constant 16 pixels further on)
  uint32_t b_off[CPW], a_off[MTN];
#pragma unroll
  for (int j = 0; j < CPW; ++j) {
    const int cc0 = (lane & 15) + c_dc[j];
    b_off[j] = (uint32_t)((((c_dr[j] * p.WR + cc0) << p.lgCG) + (c_c8[j] ^ ((cc0 >> p.sh) & cgm))) * 16);
  }
  if (NC == 1) {
    a_off[0] = (uint32_t)((lane & 15) * 16);
  } else {
    const int ql = (lane & 7) + ((lane >> 4) << 3);       // pixel of this lane's row inside the 16-pixel k-step
#pragma unroll
    for (int i = 0; i < MTN; ++i) {
      const int nc = i * 2 + ((lane >> 3) & 1);
      a_off[i] = (uint32_t)(((ql << p.lgNC) + (nc ^ ((ql >> p.shn) & ncm))) * 16);
    }
  }

  for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) {
    const int tw = tile % p.tiles_w, th = (tile / p.tiles_w) % p.tiles_h, b = tile / (p.tiles_w * p.tiles_h);
    const int h0 = th * TH, w0 = tw * TW;
    __syncthreads();                                       // previous tile fully consumed
    // one warp per halo / tile row, lanes over the row's 16-byte chunks: no per-chunk division, the 64-bit row base once per
    // row (the flat-index form spent ~80 instructions per chunk: half of the kernel's instructions on the C = 32 layers)
    for (int r = warp; r < p.HR; r += WS_THREADS / 32) {
      const int ih = h0 - p.pad + r;
      const bool rok = (unsigned)ih < (unsigned)p.H;
      const int srow = rok ? ih : 0;
      const long long rowpix = p.ups ? ((long long)b * (p.H >> 1) + (srow >> 1)) * (p.W >> 1) : ((long long)b * p.H + srow) * p.W;
      const uint32_t rowaddr = sH_addr + (uint32_t)(((r * p.WR) << p.lgCG) * 16);
      for (int idx = lane; idx < (p.WR << p.lgCG); idx += 32) {
        const int c = idx >> p.lgCG, c8 = idx & cgm;
        const int iw = w0 - p.pad + c;
        const bool ok = rok && (unsigned)iw < (unsigned)p.W;
        const int sw_ = ok ? iw : 0;
        const long long pix = rowpix + (p.ups ? (sw_ >> 1) : sw_);
        const bf16* src = c8 < p.CG1 ? p.x1 + pix * p.ld1 + c8 * 8 : p.x2 + pix * p.ld2 + (c8 - p.CG1) * 8;
        cp_async16(rowaddr + (uint32_t)(((c << p.lgCG) + (c8 ^ ((c >> p.sh) & cgm))) * 16), src, ok);
      }
    }
    for (int r = warp; r < TH; r += WS_THREADS / 32) {
      const long long rowpix = ((long long)b * p.H + h0 + r) * p.W + w0;
      for (int idx = lane; idx < (TW << p.lgNC); idx += 32) {
        const int cpx = idx >> p.lgNC, nc = idx & ncm;
        const int q = r * TW + cpx;
        cp_async16(sD_addr + (uint32_t)(((q << p.lgNC) + (nc ^ ((q >> p.shn) & ncm))) * 16), p.dz + (rowpix + cpx) * p.lddz + nc * 8, true);
      }
    }
    cp_async_commit();
    cp_async_wait<0>();
    __syncthreads();

    for (int ks = wk; ks < TH * 2; ks += p.ksplit) {       // 16 pixels of one tile row per k-step
      const int r = ks >> 1, hsel = ks & 1;
      // A fragments: dz^T  (rows = output channels, k = pixels)
      uint32_t a[MTN][4];
      const uint32_t arow = sD_addr + (uint32_t)(((r * 32 + hsel * 16) << p.lgNC) * 16);
      if (NC == 1) {
        ldsm_x2_t(a[0][0], a[0][2], arow + a_off[0]);
        a[0][1] = a[0][3] = 0u;
      } else {
#pragma unroll
        for (int i = 0; i < MTN; ++i) ldsm_x4_t(a[i][0], a[i][1], a[i][2], a[i][3], arow + a_off[i]);
      }
      const uint32_t brow = sH_addr + (uint32_t)(((r * p.WR + hsel * 16) << p.lgCG) * 16);
#pragma unroll
      for (int j = 0; j < CPW; ++j) {
        if (!c_ok[j]) continue;                            // warp-uniform
        uint32_t b0, b1;
        ldsm_x2_t(b0, b1, brow + b_off[j]);
#pragma unroll
        for (int i = 0; i < MTN; ++i) mma_16816(acc[i][j], a[i], b0, b1);
      }
    }
  }
  // ---- reduce the KSPLIT partial slices in shared memory, then one global atomic per element ------
  // acc[i][j] = dW rows n = i*16 + g (+8), columns chunk*8 + 2t (+1) with chunk = (tap, c8)
  const int ncols = p.Kc * 8;
  float* sR = reinterpret_cast<float*>(smem);              // [MTN*16][Kc*8]
  __syncthreads();
  for (int i = tid; i < MTN * 16 * ncols; i += WS_THREADS) sR[i] = 0.f;
  __syncthreads();
#pragma unroll
  for (int j = 0; j < CPW; ++j) {
    if (!c_ok[j]) continue;
    const int col = (wc + p.cgrp * j) * 8 + 2 * t;
#pragma unroll
    for (int i = 0; i < MTN; ++i)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float* dst = sR + (i * 16 + g + h * 8) * ncols + col;
        atomicAdd(dst, acc[i][j][h * 2]);
        atomicAdd(dst + 1, acc[i][j][h * 2 + 1]);
      }
  }
  __syncthreads();
  for (int i = tid; i < MTN * 16 * ncols; i += WS_THREADS) {
    const int n = i / ncols, col = i - n * ncols;
    const int ch = col >> 3, tap = ch >> p.lgCG;
    const int c = (ch & cgm) * 8 + (col & 7);
    const float v = sR[i];
    if (n < p.Cout && c < p.Creal && v != 0.f) {
      const int kh = tap / p.KW, kw = tap - kh * p.KW;
      atomicAdd(p.dw + n * p.s_n + c * p.s_c + kh * p.s_kh + kw * p.s_kw, v);
    }
  }
}

int ws_smem_bytes(const HnConv* d) {
  const int cg = (d->C1 + d->C2) / 8;
  const int n8 = (d->Cout + 7) / 8 * 8;
  return (TH + (d->KH - 1) * d->dil) * (TW + (d->KW - 1) * d->dil) * cg * 16 + TH * TW * n8 * 2;
}
bool ws_ok(const HnConv* d) {
  if (!d || d->mode != 0 || d->stride != 1 || d->pshuf != 0) return false;
  if (d->ups && (d->C2 != 0 || (d->H & 1) || (d->W & 1))) return false;
  if (d->KH != d->KW || (d->KH != 1 && d->KH != 3) || d->dil < 1 || d->dil > 8) return false;
  if (d->OH != d->H || d->OW != d->W || d->pad != d->dil * (d->KH - 1) / 2) return false;
  if (d->H % TH || d->W % TW) return false;
  const int cin = d->C1 + d->C2;
  if (d->C1 <= 0 || d->C1 % 8 || d->C2 % 8) return false;
  if (cin != 8 && cin != 16 && cin != 32 && cin != 64) return false;
  const int n8 = (d->Cout + 7) / 8 * 8;
  if (d->Cout < 1 || (n8 != 8 && n8 != 16 && n8 != 32 && n8 != 64) || cin * n8 > 2048) return false;
  if (d->ldy % 8 || d->ldy < n8 || d->ld1 % 8 || (d->C2 && d->ld2 % 8)) return false;
  return ws_smem_bytes(d) <= 160 * 1024;
}

template <int MTN, int CPW>
int launch_ws(const WsP& p, int smem, cudaStream_t st) {
  static bool configured = false;
  static int occ[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_sc_kernel<MTN, CPW>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    if (e != cudaSuccess) HN_FAIL(HN_ECUDA, "conv_wgrad_sc: smem attribute: %s", cudaGetErrorString(e));
    configured = true;
  }
  const int bucket = smem / (16 * 1024);
  if (occ[bucket] == 0) {
    int nb = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, wgrad_sc_kernel<MTN, CPW>, WS_THREADS, (bucket + 1) * 16 * 1024);
    occ[bucket] = nb > 0 ? nb : 1;
  }
  int per_sm = occ[bucket] > 3 ? 3 : occ[bucket];          // more CTAs only multiply the final atomics
  int grid = 148 * per_sm;
  if (grid > (p.tiles + 1) / 2) grid = (p.tiles + 1) / 2;   // >= 2 tiles per CTA (measured: 4 starves the small FCT levels, 1 doubles the flush atomics)
  wgrad_sc_kernel<MTN, CPW><<<grid, WS_THREADS, smem, st>>>(p);
  HN_CHECK_LAUNCH("wgrad_sc_kernel");
  return HN_OK;
}

template <int MTN>
int dispatch_cpw(const WsP& p, int cpw, int smem, cudaStream_t st) {
  if (cpw <= 1) return launch_ws<MTN, 1>(p, smem, st);
  if (cpw <= 2) return launch_ws<MTN, 2>(p, smem, st);
  if (cpw <= 3) return launch_ws<MTN, 3>(p, smem, st);
  if (cpw <= 5) return launch_ws<MTN, 5>(p, smem, st);
  if (MTN <= 2) return launch_ws<(MTN <= 2 ? MTN : 2), 9>(p, smem, st);
  HN_FAIL(HN_EUNSUPPORTED, "conv_wgrad_sc: Cin*Cout too large");
}

}  // namespace

extern "C" int hn_conv_wgrad_sc_supported(const HnConv* d) { return ws_ok(d) ? 1 : 0; }

extern "C" int hn_conv_wgrad_sc(const HnConv* d, const void* x1, const void* x2, const void* dy, float* dw, int Creal,
                                long long s_n, long long s_c, long long s_kh, long long s_kw, void* stream) {
  HN_REQUIRE(ws_ok(d), "conv_wgrad_sc: geometry not supported by the small-channel path (use hn_conv2d_wgrad)");
  HN_REQUIRE(x1 && dy && dw && (d->C2 == 0 || x2), "conv_wgrad_sc: null pointer");
  HN_REQUIRE(((uintptr_t)x1 % 16) == 0 && ((uintptr_t)dy % 16) == 0 && ((uintptr_t)x2 % 16) == 0, "conv_wgrad_sc: alignment");
  WsP p;
  p.x1 = (const bf16*)x1; p.x2 = (const bf16*)x2; p.dz = (const bf16*)dy; p.dw = dw;
  p.B = d->B; p.H = d->H; p.W = d->W; p.C1 = d->C1; p.C2 = d->C2; p.ld1 = d->ld1; p.ld2 = d->ld2; p.lddz = d->ldy;
  p.Cout = d->Cout; p.Creal = Creal; p.KH = d->KH; p.KW = d->KW; p.pad = d->pad; p.dil = d->dil; p.ups = d->ups;
  const int cin = d->C1 + d->C2;
  p.CG = cin / 8; p.CG1 = d->C1 / 8;
  p.lgCG = p.CG == 1 ? 0 : (p.CG == 2 ? 1 : (p.CG == 4 ? 2 : 3));
  p.sh = p.CG == 2 ? 2 : (p.CG == 4 ? 1 : 0);
  const int n8 = (d->Cout + 7) / 8 * 8;
  p.NC = n8 / 8;
  p.lgNC = p.NC == 1 ? 0 : (p.NC == 2 ? 1 : (p.NC == 4 ? 2 : 3));
  p.shn = p.NC == 2 ? 2 : (p.NC == 4 ? 1 : 0);
  p.Kc = d->KH * d->KW * p.CG;
  p.HR = TH + (d->KH - 1) * d->dil; p.WR = TW + (d->KW - 1) * d->dil;
  p.tiles_w = d->W / TW; p.tiles_h = d->H / TH; p.tiles = d->B * p.tiles_w * p.tiles_h;
  p.s_n = s_n; p.s_c = s_c; p.s_kh = s_kh; p.s_kw = s_kw;
  int smem = ws_smem_bytes(d);
  const int red = (n8 < 16 ? 16 : n8) * p.Kc * 8 * 4;       // the final reduction reuses the tile buffers
  if (smem < red) smem = red;
  // warp grid: as few chunk columns as keep the per-warp dW slice in registers (9 chunks; 5 when Cout > 32)
  const int cpw_max = n8 > 32 ? 5 : 9;
  p.cgrp = 1; p.lgcgrp = 0;
  while (p.cgrp < 8 && (p.Kc + p.cgrp - 1) / p.cgrp > cpw_max) { p.cgrp *= 2; p.lgcgrp++; }
  p.ksplit = 8 / p.cgrp;
  const int cpw = (p.Kc + p.cgrp - 1) / p.cgrp;
  cudaStream_t st = (cudaStream_t)stream;
  if (n8 <= 16) return dispatch_cpw<1>(p, cpw, smem, st);
  if (n8 <= 32) return dispatch_cpw<2>(p, cpw, smem, st);
  return dispatch_cpw<4>(p, cpw, smem, st);
}
