// Small-channel weight gradient (same layers as hn_conv_sc.cu: Cin, Cout <= 64, stride 1):
//   dW[n][c][kh][kw] += sum_px dz[px][n] * x[px @ (kh,kw)][c]
// HBM-bound, so again built around traffic: a persistent CTA stages, per 16x32-pixel tile, the input
// halo and the dz tile ONCE in shared memory; the contraction over pixels runs on mma.sync m16n8k16
// with both operands read through ldmatrix.trans (pixels are the K dimension).  The 8 warps form a
// CGRP x KSPLIT grid: the (tap, 8-channel) column chunks of dW are dealt round-robin to CGRP warp
// columns, the 32 pixel k-steps of a tile to KSPLIT warp rows; every warp keeps its slice of dW in
// registers across all tiles of the CTA; at the end every warp parks its slice in shared memory (fragment order, its own
// region: no shared atomics), the KSPLIT slices of an element are summed by one thread and flushed with one fp32 global
// atomic per element per CTA -- or, launched as thread-block clusters of 8 CTAs, per element per CLUSTER: the CTAs first sum
// their dW through distributed shared memory (see the end of the kernel).  Two tile buffers when they fit beside a second CTA: the next tile streams in while this
// one is contracted (single-buffered, the loads of every tile were fully exposed: 3.5 tiles per CTA, ~6 us each at C = 8).
#include "hn_common.cuh"

namespace {

constexpr int TH = 16, TW = 32;
constexpr int WS_THREADS = 256;
#ifndef WS_CLUSTER
#define WS_CLUSTER 8       // CTAs per thread-block cluster (final reduction through distributed shared memory)
#endif
#ifndef WS_CLUSTER_MIN_FLUSH
#define WS_CLUSTER_MIN_FLUSH 100000   // cluster launches only when grid x dW elements reaches this many global atomics
#endif
#ifndef WS_MINB
#define WS_MINB(MTN) 2   // resident CTAs per SM the register allocation is held to (3 for the 16-row kernels measured no gain: 109 -> 124 us at 64 -> 8 channels)
#endif
#ifndef WS_MIN_TILES
#define WS_MIN_TILES 2     // tiles per CTA the grid is sized for
#endif

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
  int nbuf, tile_bytes;                  // tile buffers (halo + dz each) and their pitch
  int red_bytes;                         // parked accumulator slices in the final reduction (the CTA's sums follow them)
  FastDiv fd_tw, fd_th;
  long long s_n, s_c, s_kh, s_kw;
};

// thread-block cluster plumbing for the final reduction (a launch without the cluster attribute is a cluster of one)
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t cluster_size() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t dsmem_addr(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ float ld_dsmem(uint32_t addr) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}

template <int MTN, int CPW>   // 16-row blocks of Cout, column chunks per warp
__global__ void __launch_bounds__(WS_THREADS, WS_MINB(MTN)) wgrad_sc_kernel(const WsP p) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int s_tap[9];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
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

  const bool one_src = !p.ups && p.CG1 == CG;
  // halo and dz tile of `tile` -> buffer `buf`: one warp per halo / tile row, lanes over the row's 16-byte chunks (no
  // per-chunk division, the 64-bit row base once per row)
  auto stage = [&](int tile, int buf) {
    const uint32_t q0 = fd_div((uint32_t)tile, p.fd_tw);
    const int tw = tile - (int)q0 * p.tiles_w;
    const int b = (int)fd_div(q0, p.fd_th);
    const int th = (int)q0 - b * p.tiles_h;
    const int h0 = th * TH, w0 = tw * TW;
    const uint32_t hb = sH_addr + (uint32_t)(buf * p.tile_bytes), db = sD_addr + (uint32_t)(buf * p.tile_bytes);
    for (int r = warp; r < p.HR; r += WS_THREADS / 32) {
      const int ih = h0 - p.pad + r;
      const bool rok = (unsigned)ih < (unsigned)p.H;
      const int srow = rok ? ih : 0;
      const long long rowpix = p.ups ? ((long long)b * (p.H >> 1) + (srow >> 1)) * (p.W >> 1) : ((long long)b * p.H + srow) * p.W;
      const uint32_t rowaddr = hb + (uint32_t)(((r * p.WR) << p.lgCG) * 16);
      if (one_src) {
        const bf16* rowp = p.x1 + rowpix * p.ld1;
        for (int idx = lane; idx < (p.WR << p.lgCG); idx += 32) {
          const int c = idx >> p.lgCG, c8 = idx & cgm;
          const int iw = w0 - p.pad + c;
          const bool ok = rok && (unsigned)iw < (unsigned)p.W;
          cp_async16(rowaddr + (uint32_t)(((c << p.lgCG) + (c8 ^ ((c >> p.sh) & cgm))) * 16), rowp + (ok ? iw * p.ld1 + c8 * 8 : 0), ok);
        }
      } else {
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
    }
    for (int r = warp; r < TH; r += WS_THREADS / 32) {
      const bf16* rowp = p.dz + (((long long)b * p.H + h0 + r) * p.W + w0) * p.lddz;
      for (int idx = lane; idx < (TW << p.lgNC); idx += 32) {
        const int cpx = idx >> p.lgNC, nc = idx & ncm;
        const int q = r * TW + cpx;
        cp_async16(db + (uint32_t)(((q << p.lgNC) + (nc ^ ((q >> p.shn) & ncm))) * 16), rowp + cpx * p.lddz + nc * 8, true);
      }
    }
  };

  stage(blockIdx.x, 0);
  cp_async_commit();
  int it = 0;
  for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
    const int cur = p.nbuf == 2 ? (it & 1) : 0;
    if (p.nbuf == 2) {
      if (tile + (int)gridDim.x < p.tiles) stage(tile + gridDim.x, cur ^ 1);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const uint32_t sHc = sH_addr + (uint32_t)(cur * p.tile_bytes), sDc = sD_addr + (uint32_t)(cur * p.tile_bytes);

    for (int ks = wk; ks < TH * 2; ks += p.ksplit) {       // 16 pixels of one tile row per k-step
      const int r = ks >> 1, hsel = ks & 1;
      // A fragments: dz^T  (rows = output channels, k = pixels)
      uint32_t a[MTN][4];
      const uint32_t arow = sDc + (uint32_t)(((r * 32 + hsel * 16) << p.lgNC) * 16);
      if (NC == 1) {
        ldsm_x2_t(a[0][0], a[0][2], arow + a_off[0]);
        a[0][1] = a[0][3] = 0u;
      } else {
#pragma unroll
        for (int i = 0; i < MTN; ++i) ldsm_x4_t(a[i][0], a[i][1], a[i][2], a[i][3], arow + a_off[i]);
      }
      const uint32_t brow = sHc + (uint32_t)(((r * p.WR + hsel * 16) << p.lgCG) * 16);
#pragma unroll
      for (int j = 0; j < CPW; ++j) {
        if (!c_ok[j]) continue;                            // warp-uniform
        uint32_t b0, b1;
        ldsm_x2_t(b0, b1, brow + b_off[j]);
#pragma unroll
        for (int i = 0; i < MTN; ++i) mma_16816(acc[i][j], a[i], b0, b1);
      }
    }
    __syncthreads();                                       // this buffer is free for the tile after next
    if (p.nbuf == 1 && tile + (int)gridDim.x < p.tiles) {
      stage(tile + gridDim.x, 0);
      cp_async_commit();
    }
  }
  cp_async_wait<0>();
  // ---- park every warp's slice in shared memory (fragment order: one conflict-free 16-byte store per lane and
  // fragment); then the CTAs of a thread-block cluster share the dW elements out among themselves and each sums ITS
  // elements over all KSPLIT slices of all CTAs through distributed shared memory before the global atomic.  The flush
  // was the bound of this kernel: every CTA used to add its whole dW (up to 9216 words at 32 -> 32 channels) onto the
  // same addresses, ~20 G contended atomics / s in L2 (43 of 49 us at 64x64x32); a cluster of 8 issues an eighth of them.
  // acc[i][j] = dW rows n = i*16 + g (+8), columns chunk*8 + 2t (+1) with chunk = wc + cgrp*j = (tap, c8)
  const int ncols = p.Kc * 8;
  float4* sR = reinterpret_cast<float4*>(smem);            // [warp][MTN][CPW][32 lanes] x {h0e0, h0e1, h1e0, h1e1}
  __syncthreads();
#pragma unroll
  for (int i = 0; i < MTN; ++i)
#pragma unroll
    for (int j = 0; j < CPW; ++j)
      sR[((warp * MTN + i) * CPW + j) * 32 + lane] = make_float4(acc[i][j][0], acc[i][j][1], acc[i][j][2], acc[i][j][3]);
  __syncthreads();
  const float* sRf = reinterpret_cast<const float*>(smem);
  const int E = MTN * 16 * ncols;
  const int wstride = p.cgrp * MTN * CPW * 32 * 4;         // floats between the slices of warps cw + cgrp * k
  const uint32_t cs = cluster_size(), cr = cluster_rank();
  if (cs == 1) {
    // no cluster: one thread per dW element sums its KSPLIT slices and issues the global atomic
    for (int i = tid; i < E; i += WS_THREADS) {
      const int n = i / ncols, col = i - n * ncols;
      const int ch = col >> 3, tap = ch >> p.lgCG;
      const int c = (ch & cgm) * 8 + (col & 7);
      if (n < p.Cout && c < p.Creal) {
        const int mi = n >> 4, gg = n & 7, hh = (n >> 3) & 1;
        const int cw = ch & (p.cgrp - 1), jj = ch >> p.lgcgrp;
        const int ln = gg * 4 + ((col & 7) >> 1);
        const float* src = sRf + ((((cw * MTN + mi) * CPW + jj) * 32 + ln) << 2) + hh * 2 + (col & 1);
        float v = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (k < p.ksplit) v += src[k * wstride];
        if (v != 0.f) {
          const int kh = tap / p.KW, kw = tap - kh * p.KW;
          atomicAdd(p.dw + n * p.s_n + c * p.s_c + kh * p.s_kh + kw * p.s_kw, v);
        }
      }
    }
    return;
  }
  // this CTA's sums over its KSPLIT slices -> sOut[element] (behind the parked slices)
  float* sOut = reinterpret_cast<float*>(smem + p.red_bytes);
  for (int i = tid; i < E; i += WS_THREADS) {
    const int n = i / ncols, col = i - n * ncols;
    const int ch = col >> 3;
    const int mi = n >> 4, gg = n & 7, hh = (n >> 3) & 1;
    const int cw = ch & (p.cgrp - 1), jj = ch >> p.lgcgrp;
    const int ln = gg * 4 + ((col & 7) >> 1);
    const float* src = sRf + ((((cw * MTN + mi) * CPW + jj) * 32 + ln) << 2) + hh * 2 + (col & 1);
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (k < p.ksplit) v += src[k * wstride];
    sOut[i] = v;
  }
  cluster_barrier();
  // the cluster's CTAs take every cs-th block of 256 elements each and sum them over all CTAs (8 independent remote loads)
  const uint32_t so = smem_u32(sOut);
  for (int i = tid + (int)cr * WS_THREADS; i < E; i += WS_THREADS * (int)cs) {
    const int n = i / ncols, col = i - n * ncols;
    const int ch = col >> 3, tap = ch >> p.lgCG;
    const int c = (ch & cgm) * 8 + (col & 7);
    if (n < p.Cout && c < p.Creal) {
      float part[8];
#pragma unroll
      for (int rk = 0; rk < 8; ++rk) part[rk] = (uint32_t)rk < cs ? ld_dsmem(dsmem_addr(so + (uint32_t)i * 4u, (uint32_t)rk)) : 0.f;
      const float v = ((part[0] + part[1]) + (part[2] + part[3])) + ((part[4] + part[5]) + (part[6] + part[7]));
      if (v != 0.f) {
        const int kh = tap / p.KW, kw = tap - kh * p.KW;
        atomicAdd(p.dw + n * p.s_n + c * p.s_c + kh * p.s_kh + kw * p.s_kw, v);
      }
    }
  }
  cluster_barrier();                                       // nobody leaves while a neighbour still reads its sums
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

// smem: plain launch; smem_cl: cluster launch (the CTA's own sums need room behind the parked slices) or 0 when that would
// cost a resident CTA per SM
template <int MTN, int CPW>
int launch_ws(const WsP& p, int smem, int smem_cl, cudaStream_t st) {
  static bool configured = false;
  static int occ[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  static int occ_cl[11] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};   // resident clusters of WS_CLUSTER CTAs on the device (-1: none)
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
  const int grid_cl = grid > p.tiles ? p.tiles : grid;      // a cluster launch affords one tile per CTA
  const int min_tiles = WS_MIN_TILES;                       // plain: >= 2 tiles per CTA (1 doubles the flush atomics, 4 starves the small FCT levels)
  if (grid > (p.tiles + min_tiles - 1) / min_tiles) grid = (p.tiles + min_tiles - 1) / min_tiles;
  // Clusters of WS_CLUSTER CTAs reduce their dW among themselves first.  Only for the layers whose flush dominates: few
  // tiles per CTA and a large dW (measured on B200, scripts/conv_bench.py --wgrad --graph).  The grid of a cluster launch
  // is sized by cudaOccupancyMaxActiveClusters: clusters that do not fit beside the first wave would run after it.
  const long long flush = (long long)grid_cl * p.Cout * p.Creal * p.KH * p.KW;
  if (WS_CLUSTER > 1 && smem_cl > 0 && grid_cl >= WS_CLUSTER && flush >= WS_CLUSTER_MIN_FLUSH) {
    const int bcl = smem_cl / (16 * 1024);
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(WS_THREADS);
    cfg.dynamicSmemBytes = (size_t)((bcl + 1) * 16 * 1024 > 160 * 1024 ? 160 * 1024 : (bcl + 1) * 16 * 1024);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = WS_CLUSTER; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (occ_cl[bcl] == 0) {
      int nc = 0;
      cfg.gridDim = dim3(148 * 3 / WS_CLUSTER * WS_CLUSTER);
      cudaError_t e = cudaOccupancyMaxActiveClusters(&nc, wgrad_sc_kernel<MTN, CPW>, &cfg);
      if (e != cudaSuccess) (void)cudaGetLastError();
      occ_cl[bcl] = (e == cudaSuccess && nc > 0) ? nc : -1;
    }
    if (occ_cl[bcl] > 0) {
      int gcl = grid_cl - grid_cl % WS_CLUSTER;
      if (gcl > occ_cl[bcl] * WS_CLUSTER) gcl = occ_cl[bcl] * WS_CLUSTER;
      cfg.gridDim = dim3((unsigned)gcl);
      cfg.dynamicSmemBytes = (size_t)smem_cl;
      cudaError_t e = cudaLaunchKernelEx(&cfg, wgrad_sc_kernel<MTN, CPW>, p);
      if (e == cudaSuccess) return HN_OK;
      (void)cudaGetLastError();                            // fall through to the plain launch
      occ_cl[bcl] = -1;
    }
  }
  wgrad_sc_kernel<MTN, CPW><<<grid, WS_THREADS, smem, st>>>(p);
  HN_CHECK_LAUNCH("wgrad_sc_kernel");
  return HN_OK;
}

template <int MTN>
int dispatch_cpw(const WsP& p, int cpw, int smem, int smem_cl, cudaStream_t st) {
  if (cpw <= 1) return launch_ws<MTN, 1>(p, smem, smem_cl, st);
  if (cpw <= 2) return launch_ws<MTN, 2>(p, smem, smem_cl, st);
  if (cpw <= 3) return launch_ws<MTN, 3>(p, smem, smem_cl, st);
  if (cpw <= 5) return launch_ws<MTN, 5>(p, smem, smem_cl, st);
  if (MTN <= 2) return launch_ws<(MTN <= 2 ? MTN : 2), 9>(p, smem, smem_cl, st);
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
  p.tile_bytes = ws_smem_bytes(d);
  p.nbuf = 2 * p.tile_bytes <= 100 * 1024 ? 2 : 1;           // double-buffer while two CTAs per SM still fit
  p.fd_tw = hn_fastdiv((uint32_t)p.tiles_w); p.fd_th = hn_fastdiv((uint32_t)p.tiles_h);
  int smem = p.nbuf * p.tile_bytes;
  // warp grid: as few chunk columns as keep the per-warp dW slice in registers (9 chunks; 5 when Cout > 32)
  const int cpw_max = n8 > 32 ? 5 : 9;
  p.cgrp = 1; p.lgcgrp = 0;
  while (p.cgrp < 8 && (p.Kc + p.cgrp - 1) / p.cgrp > cpw_max) { p.cgrp *= 2; p.lgcgrp++; }
  p.ksplit = 8 / p.cgrp;
  const int cpw = (p.Kc + p.cgrp - 1) / p.cgrp;
  // the final reduction parks 8 warps x MTN x CPW fragments of 512 bytes in the tile buffers (CPW as instantiated) and
  // puts the CTA's own sums (one float per dW element of the padded MTN*16 x Kc*8 matrix) behind them
  const int cpw_inst = cpw <= 3 ? cpw : (cpw <= 5 ? 5 : 9);
  const int mtn = n8 <= 16 ? 1 : (n8 <= 32 ? 2 : 4);
  p.red_bytes = 8 * mtn * cpw_inst * 512;
  if (smem < p.red_bytes) smem = p.red_bytes;
  int smem_cl = p.red_bytes + mtn * 16 * p.Kc * 8 * 4;
  if (smem_cl < smem) smem_cl = smem;
  const int half = 113 * 1024;                               // two CTAs per SM fit below this
  if ((smem <= half && smem_cl > half) || smem_cl > 160 * 1024) smem_cl = 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (n8 <= 16) return dispatch_cpw<1>(p, cpw, smem, smem_cl, st);
  if (n8 <= 32) return dispatch_cpw<2>(p, cpw, smem, smem_cl, st);
  return dispatch_cpw<4>(p, cpw, smem, smem_cl, st);
}
