// Small-channel direct convolution (Cin <= 64, Cout <= 64, Cin*Cout <= 2048): every FCT / AutoEncoder
// conv, UNet level 1, the stems and heads.  These layers are HBM-bound (AI = 4.5*C FLOP/B), so the
// kernel is built around memory traffic, not the tensor pipe:
//   * a persistent CTA loads the whole (tiny) packed weight matrix into shared memory once;
//   * per 16x32-pixel output tile it stages the input halo ((16+2p) x (32+2p) pixels, all channels)
//     ONCE with 16-byte cp.async (zero fill = padding) -- every input byte crosses L2->SM once
//     instead of once per filter tap;
//   * the 9 taps are shifted ldmatrix reads of that halo feeding mma.sync m16n8k16 (K ordered
//     (tap, 8-channel chunk), two chunks per k-step, possibly from two different taps);
//   * bias / activation / BatchNorm statistics in the epilogue, 16-byte-per-pixel-quad stores,
//     one global atomic per channel per CTA for the statistics.
// Input gradients of stride-1 convs use the same kernel with tap-flipped packed weights.
#include "hn_common.cuh"

namespace {

constexpr int TH = 16, TW = 32;          // output tile (pixels)
constexpr int SC_THREADS = 256;          // 8 warps, 2 tile rows each

struct ScP {
  const bf16* x1;
  const bf16* x2;
  const bf16* w;
  const float* bias;
  bf16* y;
  float* ssum;
  float* ssq;
  const bf16* res;                       // optional residual added before the activation (pixel stride ldr)
  int B, H, W, C1, C2, ld1, ld2, ldy, Cout, Kpad;
  int KH, KW, pad, dil, act, ups, ldr;   // ups: x1 is the H/2 x W/2 map, read through nearest-neighbour 2x upsampling
  int CG, CG1, lgCG, sh;                 // 8-channel chunks (total, from x1), log2, swizzle shift
  int Kc, KS;                            // K chunks = taps*CG, k16 steps
  int wrow;                              // weight row pitch in shared memory (bytes, odd multiple of 16)
  int HR, WR;                            // halo rows / cols
  int tiles_w, tiles_h, tiles;
  int nbuf;                              // halo buffers (2 unless they do not fit)
  FastDiv fd_tw, fd_th;                  // tile index -> (b, th, tw) without runtime division
};

// tile -> image, tile row, tile column
__device__ __forceinline__ void sc_tile(const ScP& p, int tile, int& b, int& th, int& tw) {
  const uint32_t q = fd_div((uint32_t)tile, p.fd_tw);
  tw = tile - (int)q * p.tiles_w;
  const uint32_t bb = fd_div(q, p.fd_th);
  th = (int)q - (int)bb * p.tiles_h;
  b = (int)bb;
}

// Epilogue of one pass: bias, statistics, residual, activation, 4-byte stores (a warp store = 8 pixels x 16 bytes).
// FAST (no residual, activation none / ReLU) and STATS are compile-time here and chosen by ONE warp-uniform branch per
// pass: as run-time tests inside the unrolled (m-tile, row half, channel group) loops they cost a predicated-off copy of
// the slow activation and the residual load per element (ncu: ~240 warp instructions per 8-pixel store at C = 8).
// ytile / rtile: output / residual at the tile's first pixel; the per-store offset is 32-bit.
template <int NT, int MT, bool FAST, bool STATS>
__device__ __forceinline__ void sc_epilogue(const ScP& p, const float (&acc)[MT][NT][4], const float (&bia)[NT][2],
                                            float (&csum)[NT][2], float (&csq)[NT][2], bf16* __restrict__ ytile,
                                            const bf16* __restrict__ rtile, int pass, int warp, int g, int t) {
  const bool relu = p.act == HN_ACT_RELU;
  const int nlim = p.Cout - 2 * t;       // channel n = j * 8 + 2 t (+ 1) is real while j * 8 (+ 1) < nlim
#pragma unroll
  for (int i = 0; i < MT; ++i) {
    const int mt = pass * MT + i;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int off = (warp * 2 + (mt >> 1)) * p.W + (mt & 1) * 16 + g + h * 8;
      bf16* dst = ytile + (long long)off * p.ldy + 2 * t;
      const bf16* rsd = (!FAST && rtile) ? rtile + (long long)off * p.ldr + 2 * t : nullptr;
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        float v0 = acc[i][j][h * 2] + bia[j][0], v1 = acc[i][j][h * 2 + 1] + bia[j][1];
        if (STATS) {
          const float r0 = bf16_round(v0), r1 = bf16_round(v1);
          csum[j][0] += r0; csq[j][0] += r0 * r0;
          csum[j][1] += r1; csq[j][1] += r1 * r1;
        }
        if (FAST) {
          v0 = relu ? fmaxf(v0, 0.f) : v0;
          v1 = relu ? fmaxf(v1, 0.f) : v1;
        } else {
          if (rsd) {
            const float2 rr = __bfloat1622float2(*reinterpret_cast<const bf162*>(rsd + j * 8));
            v0 += rr.x; v1 += rr.y;
          }
          v0 = apply_act(v0, p.act);
          v1 = apply_act(v1, p.act);
        }
        v0 = j * 8 < nlim ? v0 : 0.f;
        v1 = j * 8 + 1 < nlim ? v1 : 0.f;
        *reinterpret_cast<uint32_t*>(dst + j * 8) = pack_bf16(v0, v1);
      }
    }
  }
}

template <int NT>   // N/8 output-channel groups
__global__ void __launch_bounds__(SC_THREADS, NT == 1 ? 4 : (NT == 8 ? 1 : 2)) conv_sc_kernel(const ScP p) {
  constexpr int MT = NT == 8 ? 2 : 4;    // m-tiles (16 px) per pass, 4 per warp in total
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ float s_stat[2][NT * 8];
  __shared__ int s_tap[9];               // (kh*dil) << 16 | (kw*dil)
  uint8_t* sW = smem;
  uint8_t* sZ = smem + NT * 8 * p.wrow;                 // 16 zero bytes (K padding chunk)
  uint8_t* sH = sZ + 128;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int CG = p.CG, cgm = CG - 1;

  // ---- one-time: weights, tap table, zero chunk -------------------------------------------
  {
    const int wch = p.wrow >> 4;
    for (int id = tid; id < NT * 8 * wch; id += SC_THREADS) {
      const int n = id / wch, j = id - n * wch;
      const bool ok = j < p.Kc;
      cp_async16(smem_u32(sW + n * p.wrow + j * 16), ok ? (const void*)(p.w + (long long)n * p.Kpad + j * 8) : (const void*)p.w, ok);
    }
    if (tid < 8) reinterpret_cast<uint32_t*>(sZ)[tid] = 0u;
    if (tid < p.KH * p.KW) {
      const int kh = tid / p.KW, kw = tid - kh * p.KW;
      s_tap[tid] = ((kh * p.dil) << 16) | (kw * p.dil);
    }
    if (tid < 2 * NT * 8) (&s_stat[0][0])[tid] = 0.f;
  }
  float csum[NT][2], csq[NT][2];
#pragma unroll
  for (int j = 0; j < NT; ++j) csum[j][0] = csum[j][1] = csq[j][0] = csq[j][1] = 0.f;
  float bia[NT][2];
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const int n = j * 8 + 2 * t;
    bia[j][0] = (p.bias && n < p.Cout) ? p.bias[n] : 0.f;
    bia[j][1] = (p.bias && n + 1 < p.Cout) ? p.bias[n + 1] : 0.f;
  }
  const int halo_chunks = p.HR * p.WR * CG;

  // halo of one tile -> buffer `buf` (cp.async, zero fill outside the image)
  const int halo_bytes = halo_chunks * 16;
  const uint32_t sHa = smem_u32(sH), sZa = smem_u32(sZ), sWa = smem_u32(sW);
  const bool one_src = !p.ups && p.CG1 == CG;      // one plain input: the common case gets the short address path
  auto stage_halo = [&](int tile, int buf) {
    int tw, th, b;
    sc_tile(p, tile, b, th, tw);
    const int h0 = th * TH, w0 = tw * TW;
    // one warp per halo row, lanes over the row's 16-byte chunks (no per-chunk division; the 64-bit row base once per row)
    const uint32_t dsta = sHa + (uint32_t)(buf * halo_bytes);
    for (int r = warp; r < p.HR; r += SC_THREADS / 32) {
      const int ih = h0 - p.pad + r;
      const bool rok = (unsigned)ih < (unsigned)p.H;
      const int srow = rok ? ih : 0;
      const long long rowpix = p.ups ? ((long long)b * (p.H >> 1) + (srow >> 1)) * (p.W >> 1) : ((long long)b * p.H + srow) * p.W;
      const uint32_t rowaddr = dsta + (uint32_t)(((r * p.WR) << p.lgCG) * 16);
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
  };
  // CG == 1 (8 input channels): one chunk per pixel, no swizzle, a k-step = two taps (lanes 0-15 / 16-31).  The tap offset
  // of each k-step is a per-lane constant of the kernel: found once here instead of once per (k-step, m-tile, tile)
  constexpr uint32_t NOKEY = 0xFFFFFFFFu;
  uint32_t koff[5] = {NOKEY, NOKEY, NOKEY, NOKEY, NOKEY};     // Kc <= 9 chunks -> at most 5 k16 steps
  if (CG == 1 && NT <= 2) {      // wider outputs keep the generic k-step loop below (register budget)
#pragma unroll
    for (int ks = 0; ks < 5; ++ks) {
      const int gi = ks * 2 + (lane >> 4);
      if (gi < p.Kc) {
        const int kh = p.KW == 3 ? (gi * 11) >> 5 : 0;       // gi / 3 for gi < 10 (KW is 1 or 3 here: sc_ok); no run-time division
        const int kw = gi - kh * p.KW;
        koff[ks] = (uint32_t)(((kh * p.dil) * p.WR + kw * p.dil) * 16);
      }
    }
  }
  // two halo buffers: the next tile streams in while this one is computed (keeps loads in flight)
  stage_halo(blockIdx.x, 0);
  cp_async_commit();
  int it = 0;
  for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
    int tw, th, b;
    sc_tile(p, tile, b, th, tw);
    const int h0 = th * TH, w0 = tw * TW;
    const int cur = p.nbuf == 2 ? (it & 1) : 0;
    if (p.nbuf == 2) {
      if (tile + (int)gridDim.x < p.tiles) stage_halo(tile + gridDim.x, cur ^ 1);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const uint32_t sHc = sHa + (uint32_t)(cur * halo_bytes);

#pragma unroll
    for (int pass = 0; pass < 4 / MT; ++pass) {
      float acc[MT][NT][4];
#pragma unroll
      for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j) acc[i][j][0] = acc[i][j][1] = acc[i][j][2] = acc[i][j][3] = 0.f;

      if (CG >= 2) {
        // K order (tap, c8) with an even chunk count per tap: tap outer, so the halo address of a pixel
        // is computed once per (tap, m-tile) and a k-step only XORs the chunk index into it
        const int kpt = CG >> 1;                              // k16 steps per tap
        const int taps = p.KH * p.KW;
        int ks = 0;
        for (int tap = 0; tap < taps; ++tap) {
          const int tp = s_tap[tap];
          const int dr = tp >> 16, dc = tp & 0xFFFF;
          uint32_t abase[MT];
          int akey[MT];
#pragma unroll
          for (int i = 0; i < MT; ++i) {
            const int mt = pass * MT + i;
            const int r = warp * 2 + (mt >> 1) + dr;
            const int c = (mt & 1) * 16 + (lane & 15) + dc;
            abase[i] = sHc + (uint32_t)(((r * p.WR + c) << p.lgCG) * 16);
            akey[i] = (c >> p.sh) & cgm;
          }
          for (int kk = 0; kk < kpt; ++kk, ++ks) {
            uint32_t bfr[NT][2];
            if (NT >= 2) {
#pragma unroll
              for (int j = 0; j < NT; j += 2) {
                const int row = j * 8 + (lane & 7) + ((lane >> 4) << 3);
                const int ch = ks * 2 + ((lane >> 3) & 1);
                ldsm_x4(bfr[j][0], bfr[j][1], bfr[j + 1 < NT ? j + 1 : j][0], bfr[j + 1 < NT ? j + 1 : j][1],
                        sWa + (uint32_t)(row * p.wrow + ch * 16));
              }
            } else {
              ldsm_x2(bfr[0][0], bfr[0][1], sWa + (uint32_t)((lane & 7) * p.wrow + (ks * 2 + ((lane >> 3) & 1)) * 16));
            }
            const int c8 = kk * 2 + (lane >> 4);
            uint32_t af[MT][4];                               // all A fragments first: the MMAs then issue back to back
#pragma unroll
            for (int i = 0; i < MT; ++i)
              ldsm_x4(af[i][0], af[i][1], af[i][2], af[i][3], abase[i] + (uint32_t)((c8 ^ akey[i]) << 4));
#pragma unroll
            for (int i = 0; i < MT; ++i)
#pragma unroll
              for (int j = 0; j < NT; ++j) mma_16816(acc[i][j], af[i], bfr[j][0], bfr[j][1]);
          }
        }
      } else if (NT <= 2) {
        uint32_t abase[MT];
#pragma unroll
        for (int i = 0; i < MT; ++i) {
          const int mt = pass * MT + i;
          abase[i] = sHc + (uint32_t)(((warp * 2 + (mt >> 1)) * p.WR + (mt & 1) * 16 + (lane & 15)) * 16);
        }
#pragma unroll
        for (int ks = 0; ks < 5; ++ks) {
          if (ks < p.KS) {
            uint32_t bfr[NT][2];
            if (NT >= 2) {
#pragma unroll
              for (int j = 0; j < NT; j += 2) {
                const int row = j * 8 + (lane & 7) + ((lane >> 4) << 3);
                const int ch = ks * 2 + ((lane >> 3) & 1);
                ldsm_x4(bfr[j][0], bfr[j][1], bfr[j + 1 < NT ? j + 1 : j][0], bfr[j + 1 < NT ? j + 1 : j][1],
                        sWa + (uint32_t)(row * p.wrow + ch * 16));
              }
            } else {
              ldsm_x2(bfr[0][0], bfr[0][1], sWa + (uint32_t)((lane & 7) * p.wrow + (ks * 2 + ((lane >> 3) & 1)) * 16));
            }
            uint32_t af[MT][4];
#pragma unroll
            for (int i = 0; i < MT; ++i)
              ldsm_x4(af[i][0], af[i][1], af[i][2], af[i][3], koff[ks] != NOKEY ? abase[i] + koff[ks] : sZa);
#pragma unroll
            for (int i = 0; i < MT; ++i)
#pragma unroll
              for (int j = 0; j < NT; ++j) mma_16816(acc[i][j], af[i], bfr[j][0], bfr[j][1]);
          }
        }
      } else
      for (int ks = 0; ks < p.KS; ++ks) {
        // B fragments (weights) for this k-step
        uint32_t bfr[NT][2];
        if (NT >= 2) {
#pragma unroll
          for (int j = 0; j < NT; j += 2) {
            const int row = j * 8 + (lane & 7) + ((lane >> 4) << 3);
            const int ch = ks * 2 + ((lane >> 3) & 1);
            ldsm_x4(bfr[j][0], bfr[j][1], bfr[j + 1 < NT ? j + 1 : j][0], bfr[j + 1 < NT ? j + 1 : j][1],
                    sWa + (uint32_t)(row * p.wrow + ch * 16));
          }
        } else {
          ldsm_x2(bfr[0][0], bfr[0][1], sWa + (uint32_t)((lane & 7) * p.wrow + (ks * 2 + ((lane >> 3) & 1)) * 16));
        }
        // A fragments: chunk gi of the K order = (tap, c8); lanes 0-15 first chunk, 16-31 second
        const int gi = ks * 2 + (lane >> 4);
        const bool kok = gi < p.Kc;
        const int tap = kok ? (gi >> p.lgCG) : 0, c8 = gi & cgm;
        const int tp = s_tap[tap];
        const int dr = tp >> 16, dc = tp & 0xFFFF;
#pragma unroll
        for (int i = 0; i < MT; ++i) {
          const int mt = pass * MT + i;
          const int r = warp * 2 + (mt >> 1) + dr;
          const int c = (mt & 1) * 16 + (lane & 15) + dc;
          const uint32_t addr = kok ? sHc + (uint32_t)((((r * p.WR + c) << p.lgCG) + (c8 ^ ((c >> p.sh) & cgm))) * 16) : sZa;
          uint32_t af[4];
          ldsm_x4(af[0], af[1], af[2], af[3], addr);
#pragma unroll
          for (int j = 0; j < NT; ++j) mma_16816(acc[i][j], af, bfr[j][0], bfr[j][1]);
        }
      }
      // ---- epilogue of this pass: one warp-uniform choice of the compiled variant ----------------
      const long long tile_pix = ((long long)b * p.H + h0) * p.W + w0;
      bf16* ytile = p.y + tile_pix * p.ldy;
      const bf16* rtile = p.res ? p.res + tile_pix * p.ldr : nullptr;
      const bool fast = p.act <= HN_ACT_RELU && !p.res;
      if (fast) {
        if (p.ssum) sc_epilogue<NT, MT, true, true>(p, acc, bia, csum, csq, ytile, rtile, pass, warp, g, t);
        else sc_epilogue<NT, MT, true, false>(p, acc, bia, csum, csq, ytile, rtile, pass, warp, g, t);
      } else {
        if (p.ssum) sc_epilogue<NT, MT, false, true>(p, acc, bia, csum, csq, ytile, rtile, pass, warp, g, t);
        else sc_epilogue<NT, MT, false, false>(p, acc, bia, csum, csq, ytile, rtile, pass, warp, g, t);
      }
    }
    __syncthreads();                                       // everyone is done reading this halo buffer
    if (p.nbuf == 1 && tile + (int)gridDim.x < p.tiles) {
      stage_halo(tile + gridDim.x, 0);
      cp_async_commit();
    }
  }
  cp_async_wait<0>();
  if (p.ssum) {
#pragma unroll
    for (int j = 0; j < NT; ++j)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        float s = csum[j][e], q = csq[j][e];
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
          s += __shfl_xor_sync(0xffffffffu, s, o);
          q += __shfl_xor_sync(0xffffffffu, q, o);
        }
        if (g == 0) {
          atomicAdd(&s_stat[0][j * 8 + 2 * t + e], s);
          atomicAdd(&s_stat[1][j * 8 + 2 * t + e], q);
        }
      }
    __syncthreads();
    for (int n = tid; n < NT * 8; n += SC_THREADS)
      if (n < p.Cout) {
        atomicAdd(p.ssum + n, s_stat[0][n]);
        atomicAdd(p.ssq + n, s_stat[1][n]);
      }
  }
}

constexpr int SC_SMEM_MAX = 200 * 1024;
int sc_smem_bytes(const HnConv* d, int nbuf) {
  const int cg = (d->C1 + d->C2) / 8, kc = d->KH * d->KW * cg;
  const int wch = (kc & 1) ? kc + 2 : kc + 1;
  const int n8 = (d->Cout + 7) / 8 * 8;
  return n8 * wch * 16 + 128 + nbuf * (TH + (d->KH - 1) * d->dil) * (TW + (d->KW - 1) * d->dil) * cg * 16;
}
bool sc_ok(const HnConv* d) {
  if (!d || d->mode != 0 || d->stride != 1 || d->pshuf != 0) return false;
  if (d->ups && (d->C2 != 0 || (d->H & 1) || (d->W & 1))) return false;
  if (d->ldr && (d->ldr % 2 || d->ldr < d->Cout)) return false;
  if (d->KH != d->KW || (d->KH != 1 && d->KH != 3) || d->dil < 1 || d->dil > 8) return false;
  if (d->OH != d->H || d->OW != d->W || d->pad != d->dil * (d->KH - 1) / 2) return false;
  if (d->H % TH || d->W % TW) return false;
  const int cin = d->C1 + d->C2;
  if (d->C1 <= 0 || d->C1 % 8 || d->C2 % 8) return false;
  if (cin != 8 && cin != 16 && cin != 32 && cin != 64) return false;
  const int n8 = (d->Cout + 7) / 8 * 8;
  if (d->Cout < 1 || (n8 != 8 && n8 != 16 && n8 != 32 && n8 != 64) || cin * n8 > 2048) return false;
  if (d->ldy % 8 || d->ldy < n8 || d->ld1 % 8 || (d->C2 && d->ld2 % 8)) return false;
  if (d->Kpad < d->KH * d->KW * cin) return false;
  return sc_smem_bytes(d, 1) <= SC_SMEM_MAX;
}

template <int NT>
int launch_sc(const ScP& p, int smem, cudaStream_t st) {
  static int max_smem = 0;
  static int blocks_per_sm[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (smem > max_smem) {
    cudaError_t e = cudaFuncSetAttribute(conv_sc_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, SC_SMEM_MAX);
    if (e != cudaSuccess) HN_FAIL(HN_ECUDA, "conv_sc: smem attribute: %s", cudaGetErrorString(e));
    max_smem = SC_SMEM_MAX;
  }
  // resident CTAs per SM for this shared-memory footprint (bucketed by 16 KB)
  const int bucket = smem / (16 * 1024) > 7 ? 7 : smem / (16 * 1024);
  if (blocks_per_sm[bucket] == 0) {
    int nb = 0;
    const int q = bucket == 7 ? SC_SMEM_MAX : (bucket + 1) * 16 * 1024;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, conv_sc_kernel<NT>, SC_THREADS, q);
    blocks_per_sm[bucket] = nb > 0 ? nb : 1;
  }
  int grid = 148 * blocks_per_sm[bucket];
  if (grid > p.tiles) grid = p.tiles;
  conv_sc_kernel<NT><<<grid, SC_THREADS, smem, st>>>(p);
  HN_CHECK_LAUNCH("conv_sc_kernel");
  return HN_OK;
}

}  // namespace

extern "C" int hn_conv_sc_supported(const HnConv* d) { return sc_ok(d) ? 1 : 0; }

extern "C" int hn_conv_sc_fwd(const HnConv* d, const void* x1, const void* x2, const void* wpacked, const float* bias,
                              void* y, float* stat_sum, float* stat_sqsum, void* stream) {
  HN_REQUIRE(d && d->ldr == 0, "conv_sc_fwd: a residual (ldr != 0) needs hn_conv_sc_fwd_res");
  return hn_conv_sc_fwd_res(d, x1, x2, wpacked, bias, nullptr, y, stat_sum, stat_sqsum, stream);
}

extern "C" int hn_conv_sc_fwd_res(const HnConv* d, const void* x1, const void* x2, const void* wpacked, const float* bias,
                                  const void* res, void* y, float* stat_sum, float* stat_sqsum, void* stream) {
  HN_REQUIRE(sc_ok(d), "conv_sc_fwd: geometry not supported by the small-channel path (use hn_conv2d_fwd)");
  HN_REQUIRE((res != nullptr) == (d->ldr != 0) && ((uintptr_t)res % 4) == 0, "conv_sc_fwd: residual pointer and ldr must come together");
  HN_REQUIRE(x1 && wpacked && y && (d->C2 == 0 || x2), "conv_sc_fwd: null pointer");
  HN_REQUIRE((stat_sum == nullptr) == (stat_sqsum == nullptr), "conv_sc_fwd: stats need both buffers");
  HN_REQUIRE(((uintptr_t)x1 % 16) == 0 && ((uintptr_t)wpacked % 16) == 0 && ((uintptr_t)y % 4) == 0 && ((uintptr_t)x2 % 16) == 0,
             "conv_sc_fwd: alignment");
  ScP p;
  p.x1 = (const bf16*)x1; p.x2 = (const bf16*)x2; p.w = (const bf16*)wpacked; p.bias = bias; p.y = (bf16*)y;
  p.ssum = stat_sum; p.ssq = stat_sqsum;
  p.res = (const bf16*)res; p.ups = d->ups; p.ldr = d->ldr;
  p.B = d->B; p.H = d->H; p.W = d->W; p.C1 = d->C1; p.C2 = d->C2; p.ld1 = d->ld1; p.ld2 = d->ld2; p.ldy = d->ldy;
  p.Cout = d->Cout; p.Kpad = d->Kpad; p.KH = d->KH; p.KW = d->KW; p.pad = d->pad; p.dil = d->dil; p.act = d->act;
  const int cin = d->C1 + d->C2;
  p.CG = cin / 8; p.CG1 = d->C1 / 8;
  p.lgCG = p.CG == 1 ? 0 : (p.CG == 2 ? 1 : (p.CG == 4 ? 2 : 3));
  p.sh = p.CG == 2 ? 2 : (p.CG == 4 ? 1 : 0);
  p.Kc = d->KH * d->KW * p.CG;
  p.KS = (p.Kc + 1) / 2;
  const int wch = (p.Kc & 1) ? p.Kc + 2 : p.Kc + 1;        // odd chunk count: conflict-free ldmatrix, zero K padding
  p.wrow = wch * 16;
  p.HR = TH + (d->KH - 1) * d->dil; p.WR = TW + (d->KW - 1) * d->dil;
  p.tiles_w = d->W / TW; p.tiles_h = d->H / TH; p.tiles = d->B * p.tiles_w * p.tiles_h;
  p.fd_tw = hn_fastdiv((uint32_t)p.tiles_w); p.fd_th = hn_fastdiv((uint32_t)p.tiles_h);
  const int n8 = (d->Cout + 7) / 8 * 8;
  const int NT = n8 <= 8 ? 1 : (n8 <= 16 ? 2 : (n8 <= 32 ? 4 : 8));
  p.nbuf = sc_smem_bytes(d, 2) <= 112 * 1024 ? 2 : 1;      // double-buffer when two CTAs per SM still fit
  const int smem = sc_smem_bytes(d, p.nbuf);
  HN_REQUIRE(smem <= SC_SMEM_MAX, "conv_sc_fwd: tile does not fit shared memory (%d B)", smem);
  cudaStream_t st = (cudaStream_t)stream;
  switch (NT) {
    case 1: return launch_sc<1>(p, smem, st);
    case 2: return launch_sc<2>(p, smem, st);
    case 4: return launch_sc<4>(p, smem, st);
    default: return launch_sc<8>(p, smem, st);
  }
}
