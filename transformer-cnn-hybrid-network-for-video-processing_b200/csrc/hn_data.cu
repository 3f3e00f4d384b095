// Input side of the hot path: the step right before `model(x)`.
//   Dataloader.py:16-24   PIL RGB -> ToTensor: uint8 HWC -> float CHW / 255
//   Model.py:44-53        batch-wide hflip, image + N(0, 0.05^2) noise, up to three 16x16 blackout patches
// One pass reads the decoded frames (uint8 HWC as the loader produces them, or fp32 NCHW as the reference's
// tensors are) and writes the augmented batch DIRECTLY as the first conv's operand (NHWC bf16, 3 -> 8 channels
// zero padded) and, when asked, as the fp32 NCHW tensor the reference's trainer would hold.  The random
// DECISIONS (flip? noise? which patches?) stay on the host exactly as in the reference (python `random`), so a
// seeded run takes the same branches; only the Gaussian noise values come from a counter-based generator
// (Box-Muller over the dropout hash), which cannot match torch.randn bit for bit.
// Also here: the AvgPool2d(2,2) backward (FCT.py:238-240 under autograd).
#include "hn_common.cuh"

namespace {
constexpr int TPB = 256;
inline int grid_for(long long items) {
  long long b = (items + TPB - 1) / TPB;
  if (b > 148 * 12) b = 148 * 12;
  return b < 1 ? 1 : (int)b;
}

struct Patches { int n; int x[4]; int y[4]; int size; };

__device__ __forceinline__ float gauss(unsigned long long seed, unsigned long long idx) {
  const uint64_t h = hn_hash64(seed, idx);
  const float u1 = ((uint32_t)(h >> 40) + 1u) * (1.0f / 16777217.0f);      // (0, 1]
  const float u2 = (uint32_t)(h & 0xFFFFFFu) * (1.0f / 16777216.0f);
  return sqrtf(-2.f * __logf(u1)) * __cosf(6.283185307179586f * u2);
}

// one thread per output pixel
template <bool U8>
__global__ void __launch_bounds__(TPB) augment_pack_kernel(const void* __restrict__ src, bf16* __restrict__ dst_nhwc,
                                                          float* __restrict__ dst_nchw, int B, int H, int W, int ld,
                                                          int flip, float noise_std, unsigned long long seed,
                                                          const unsigned long long* __restrict__ step_counter, Patches pt) {
  const long long HW = (long long)H * W, total = (long long)B * HW;
  const unsigned long long sd = seed + (step_counter ? step_counter[0] * 0x9E3779B97F4A7C15ull : 0ull);
  for (long long i = blockIdx.x * (long long)TPB + threadIdx.x; i < total; i += (long long)gridDim.x * TPB) {
    const int b = (int)(i / HW);
    const int r = (int)(i - (long long)b * HW);
    const int h = r / W, w = r - h * W;
    // Model.py:44-47: the noise branch starts again from the UNFLIPPED image (`aug_image = image + noise`)
    const bool noisy = noise_std > 0.f;
    const int ws = (flip && !noisy) ? (W - 1 - w) : w;
    float v[3];
    if (U8) {
      const uint8_t* s = reinterpret_cast<const uint8_t*>(src) + ((long long)b * HW + (long long)h * W + ws) * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c] = __fdiv_rn((float)s[c], 255.0f);     // ToTensor(): a true division, bit-identical to torch's .div(255)
    } else {
      const float* s = reinterpret_cast<const float*>(src) + (long long)b * 3 * HW + (long long)h * W + ws;
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c] = s[c * HW];
    }
    if (noisy) {
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c] += noise_std * gauss(sd, (unsigned long long)(i * 3 + c));
    }
    bool black = false;
    for (int k = 0; k < pt.n; ++k)       // aug_image[:, :, x:x+16, y:y+16] = 0 (x indexes H, y indexes W)
      black |= (h >= pt.x[k] && h < pt.x[k] + pt.size && w >= pt.y[k] && w < pt.y[k] + pt.size);
    if (black) v[0] = v[1] = v[2] = 0.f;
    if (dst_nhwc) {
      float f[8] = {v[0], v[1], v[2], 0.f, 0.f, 0.f, 0.f, 0.f};
      store8(dst_nhwc + i * ld, f);
    }
    if (dst_nchw) {
      float* o = dst_nchw + (long long)b * 3 * HW + r;
#pragma unroll
      for (int c = 0; c < 3; ++c) o[c * HW] = v[c];
    }
  }
}

__global__ void __launch_bounds__(TPB) avgpool2_bwd_kernel(const bf16* __restrict__ dy, bf16* __restrict__ dx, int B, int H,
                                                          int W, int CV, int lddy, int lddx) {
  // H, W: extent of dy (the pooled map); dx is 2H x 2W
  const long long total = (long long)B * H * W * CV;
  const int W2 = 2 * W;
  for (long long i = blockIdx.x * (long long)TPB + threadIdx.x; i < total; i += (long long)gridDim.x * TPB) {
    const int cv = (int)(i % CV);
    const long long op = i / CV;
    const int w = (int)(op % W);
    const long long t = op / W;
    const int h = (int)(t % H), b = (int)(t / H);
    float f[8];
    load8(dy + op * lddy + cv * 8, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= 0.25f;
    bf16* o = dx + (((long long)b * 2 * H + 2 * h) * W2 + 2 * w) * lddx + cv * 8;
    store8(o, f); store8(o + lddx, f); store8(o + (long long)W2 * lddx, f); store8(o + (long long)W2 * lddx + lddx, f);
  }
}
}  // namespace

extern "C" int hn_augment_pack(const void* src, int src_is_u8, void* dst_nhwc, float* dst_nchw, int B, int H, int W, int ld,
                               int flip, float noise_std, unsigned long long seed, const unsigned long long* step_counter,
                               int npatch, const int* patch_x, const int* patch_y, int patch_size, void* stream) {
  HN_REQUIRE(src && (dst_nhwc || dst_nchw) && B > 0 && H > 0 && W > 0, "augment_pack: bad args");
  HN_REQUIRE(!dst_nhwc || (ld % 8 == 0 && ld >= 8 && (uintptr_t)dst_nhwc % 16 == 0), "augment_pack: NHWC destination needs ld %% 8 == 0");
  HN_REQUIRE(npatch >= 0 && npatch <= 4 && (npatch == 0 || (patch_x && patch_y && patch_size > 0)), "augment_pack: at most 4 patches");
  Patches pt;
  pt.n = npatch; pt.size = patch_size;
  for (int k = 0; k < 4; ++k) { pt.x[k] = k < npatch ? patch_x[k] : 0; pt.y[k] = k < npatch ? patch_y[k] : 0; }
  const int grid = grid_for((long long)B * H * W);
  cudaStream_t st = (cudaStream_t)stream;
  if (src_is_u8)
    augment_pack_kernel<true><<<grid, TPB, 0, st>>>(src, (bf16*)dst_nhwc, dst_nchw, B, H, W, ld, flip, noise_std, seed, step_counter, pt);
  else
    augment_pack_kernel<false><<<grid, TPB, 0, st>>>(src, (bf16*)dst_nhwc, dst_nchw, B, H, W, ld, flip, noise_std, seed, step_counter, pt);
  HN_CHECK_LAUNCH("augment_pack");
  return HN_OK;
}

extern "C" int hn_avgpool2_bwd(const void* dy, void* dx, int B, int H, int W, int C, int lddy, int lddx, void* stream) {
  HN_REQUIRE(dy && dx && B > 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0 && lddy % 8 == 0 && lddx % 8 == 0, "avgpool2_bwd: bad args");
  avgpool2_bwd_kernel<<<grid_for((long long)B * H * W * (C / 8)), TPB, 0, (cudaStream_t)stream>>>(
      (const bf16*)dy, (bf16*)dx, B, H, W, C / 8, lddy, lddx);
  HN_CHECK_LAUNCH("avgpool2_bwd");
  return HN_OK;
}
