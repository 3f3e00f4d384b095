// Loss / metric kernels of the step right after the hot path:
//   DiceLoss      Metrics.py:14-22   1 - (2*sum(p*t) + 1) / (sum(p) + sum(t) + 1) on channel 0
//   nn.MSELoss    Model.py:57 (the L2 term of MixedLoss), PSNR of Model.py:30 from the same pass
//   JaccardScore  Metrics.py:31-39   sum(p!=0 & t!=0) / sum(p!=0 | t!=0) on channel 0
// Each forward is ONE launch: grid-wide sums land in a small caller-owned workspace through per-CTA
// atomics and the last CTA to finish turns them into the scalar, keeps what backward needs and re-zeroes
// the accumulators -- no host read, no second launch, CUDA-graph capturable.  The backward kernels write
// dL/dy (the gradient seed of the whole backward pass) in one pass.
//
// Workspace layout (float[16], zero before the first call, self-cleaning afterwards):
//   [0..3] accumulators   [4] CTA-done counter (as uint)
// hn_dice_fwd writes loss3 = {loss, sum(p t), sum p + sum t + smooth}; hn_dice_bwd reads it back.
#include "hn_common.cuh"

namespace {

constexpr int LOSS_TPB = 256;
constexpr int LOSS_GRID = 148 * 4;

__device__ __forceinline__ float block_sum(float v, float* sh) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  float r = 0.f;
  if (warp == 0) {
    r = lane < (LOSS_TPB / 32) ? sh[lane] : 0.f;
    r = warp_sum(r);
  }
  __syncthreads();
  return r;          // valid in warp 0
}

// returns true in exactly one thread of the grid (thread 0 of the last CTA), after every CTA's atomics are visible
__device__ __forceinline__ bool last_cta_done(float* ws) {
  __shared__ bool last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int prev = atomicAdd(reinterpret_cast<unsigned int*>(ws + 4), 1u);
    last = prev == gridDim.x - 1;
  }
  __syncthreads();
  if (last) __threadfence();
  return last && threadIdx.x == 0;
}

// channel 0 of an NCHW fp32 tensor: frame b at base + b*C*HW, HW contiguous values
__global__ void __launch_bounds__(LOSS_TPB) dice_fwd_kernel(const float* __restrict__ pred, const float* __restrict__ tgt,
                                                           int B, long long pstride, long long tstride, int HW,
                                                           float smooth, float* __restrict__ ws, float* __restrict__ loss) {
  __shared__ float sh[LOSS_TPB / 32];
  float si = 0.f, sp = 0.f, st = 0.f;
  const int hw4 = HW >> 2;
  const long long total4 = (long long)B * hw4;
  for (long long i = blockIdx.x * (long long)LOSS_TPB + threadIdx.x; i < total4; i += (long long)gridDim.x * LOSS_TPB) {
    const int b = (int)(i / hw4), r = (int)(i - (long long)b * hw4);
    const float4 p = *reinterpret_cast<const float4*>(pred + b * pstride + 4 * r);
    const float4 t = *reinterpret_cast<const float4*>(tgt + b * tstride + 4 * r);
    si += p.x * t.x + p.y * t.y + p.z * t.z + p.w * t.w;
    sp += p.x + p.y + p.z + p.w;
    st += t.x + t.y + t.z + t.w;
  }
  const float a = block_sum(si, sh), b2 = block_sum(sp, sh), c = block_sum(st, sh);
  if (threadIdx.x == 0) { atomicAdd(ws + 0, a); atomicAdd(ws + 1, b2); atomicAdd(ws + 2, c); }
  if (last_cta_done(ws)) {
    const float I = atomicAdd(ws + 0, 0.f), Sp = atomicAdd(ws + 1, 0.f), St = atomicAdd(ws + 2, 0.f);
    const float S = Sp + St + smooth;
    loss[0] = 1.f - (2.f * I + smooth) / S;
    loss[1] = I; loss[2] = S;            // what dice_bwd needs
    ws[0] = ws[1] = ws[2] = ws[3] = 0.f;
    *reinterpret_cast<unsigned int*>(ws + 4) = 0u;
  }
}

// dL/dp = -g * (2 t S - (2 I + smooth)) / S^2 on channel 0, zero on the other channels
__global__ void __launch_bounds__(LOSS_TPB) dice_bwd_kernel(const float* __restrict__ tgt, const float* __restrict__ fwd,
                                                           const float* __restrict__ gout, float* __restrict__ dpred,
                                                           int B, int Cp, long long tstride, int HW, float smooth) {
  const float I = fwd[1], S = fwd[2], g = gout ? gout[0] : 1.f;
  const float k1 = -g * 2.f / S, k0 = g * (2.f * I + smooth) / (S * S);
  const int hw4 = HW >> 2;
  const long long total4 = (long long)B * Cp * hw4;
  for (long long i = blockIdx.x * (long long)LOSS_TPB + threadIdx.x; i < total4; i += (long long)gridDim.x * LOSS_TPB) {
    const long long bc = i / hw4;
    const int r = (int)(i - bc * hw4);
    const int b = (int)(bc / Cp), c = (int)(bc - (long long)b * Cp);
    float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c == 0) {
      const float4 t = *reinterpret_cast<const float4*>(tgt + b * tstride + 4 * r);
      d = make_float4(fmaf(k1, t.x, k0), fmaf(k1, t.y, k0), fmaf(k1, t.z, k0), fmaf(k1, t.w, k0));
    }
    *reinterpret_cast<float4*>(dpred + bc * HW + 4 * r) = d;
  }
}

__global__ void __launch_bounds__(LOSS_TPB) mse_fwd_kernel(const float* __restrict__ a, const float* __restrict__ b, long long n,
                                                          float* __restrict__ ws, float* __restrict__ out) {
  __shared__ float sh[LOSS_TPB / 32];
  float s = 0.f;
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * (long long)LOSS_TPB + threadIdx.x; i < n4; i += (long long)gridDim.x * LOSS_TPB) {
    const float4 x = reinterpret_cast<const float4*>(a)[i], y = reinterpret_cast<const float4*>(b)[i];
    const float d0 = x.x - y.x, d1 = x.y - y.y, d2 = x.z - y.z, d3 = x.w - y.w;
    s += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
  }
  const long long tail = (n4 << 2) + blockIdx.x * (long long)LOSS_TPB + threadIdx.x;
  if (tail < n) { const float d = a[tail] - b[tail]; s += d * d; }
  const float r = block_sum(s, sh);
  if (threadIdx.x == 0) atomicAdd(ws + 0, r);
  if (last_cta_done(ws)) {
    const float mse = atomicAdd(ws + 0, 0.f) / (float)n;
    out[0] = mse;
    out[1] = 20.f * log10f(1.f / sqrtf(mse));        // PSNR with max_val = 1 (Model.py:30)
    ws[0] = 0.f;
    *reinterpret_cast<unsigned int*>(ws + 4) = 0u;
  }
}

__global__ void __launch_bounds__(LOSS_TPB) mse_bwd_kernel(const float* __restrict__ a, const float* __restrict__ b, long long n,
                                                          const float* __restrict__ gout, float* __restrict__ da) {
  const float k = 2.f * (gout ? gout[0] : 1.f) / (float)n;
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * (long long)LOSS_TPB + threadIdx.x; i < n4; i += (long long)gridDim.x * LOSS_TPB) {
    const float4 x = reinterpret_cast<const float4*>(a)[i], y = reinterpret_cast<const float4*>(b)[i];
    reinterpret_cast<float4*>(da)[i] = make_float4(k * (x.x - y.x), k * (x.y - y.y), k * (x.z - y.z), k * (x.w - y.w));
  }
  const long long tail = (n4 << 2) + blockIdx.x * (long long)LOSS_TPB + threadIdx.x;
  if (tail < n) da[tail] = k * (a[tail] - b[tail]);
}

__global__ void __launch_bounds__(LOSS_TPB) jaccard_kernel(const float* __restrict__ pred, const float* __restrict__ tgt, int B,
                                                          long long pstride, long long tstride, int HW, float* __restrict__ ws,
                                                          float* __restrict__ out) {
  __shared__ float sh[LOSS_TPB / 32];
  float inter = 0.f, uni = 0.f;
  const long long total = (long long)B * HW;
  for (long long i = blockIdx.x * (long long)LOSS_TPB + threadIdx.x; i < total; i += (long long)gridDim.x * LOSS_TPB) {
    const int b = (int)(i / HW), r = (int)(i - (long long)b * HW);
    const bool p = pred[b * pstride + r] != 0.f, t = tgt[b * tstride + r] != 0.f;     // torch.logical_and/or: nonzero is True
    inter += (p && t) ? 1.f : 0.f;
    uni += (p || t) ? 1.f : 0.f;
  }
  const float a = block_sum(inter, sh), c = block_sum(uni, sh);
  if (threadIdx.x == 0) { atomicAdd(ws + 0, a); atomicAdd(ws + 1, c); }
  if (last_cta_done(ws)) {
    out[0] = atomicAdd(ws + 0, 0.f) / atomicAdd(ws + 1, 0.f);        // 0/0 = NaN like the reference
    ws[0] = ws[1] = 0.f;
    *reinterpret_cast<unsigned int*>(ws + 4) = 0u;
  }
}

int loss_grid(long long items) {
  long long b = (items + LOSS_TPB - 1) / LOSS_TPB;
  if (b > LOSS_GRID) b = LOSS_GRID;
  return b < 1 ? 1 : (int)b;
}

}  // namespace

extern "C" int hn_dice_fwd(const float* pred, const float* target, int B, int Cpred, int Ctarget, int HW, float smooth,
                           float* ws, float* loss3, void* stream) {
  HN_REQUIRE(pred && target && ws && loss3 && B > 0 && Cpred > 0 && Ctarget > 0 && HW > 0, "dice_fwd: bad args");
  HN_REQUIRE(HW % 4 == 0 && ((uintptr_t)pred | (uintptr_t)target) % 16 == 0, "dice_fwd: H*W must be a multiple of 4, tensors 16B aligned");
  dice_fwd_kernel<<<loss_grid((long long)B * HW / 4), LOSS_TPB, 0, (cudaStream_t)stream>>>(
      pred, target, B, (long long)Cpred * HW, (long long)Ctarget * HW, HW, smooth, ws, loss3);
  HN_CHECK_LAUNCH("dice_fwd");
  return HN_OK;
}

extern "C" int hn_dice_bwd(const float* target, const float* fwd_out3, const float* grad_out, float* dpred, int B, int Cpred,
                           int Ctarget, int HW, float smooth, void* stream) {
  HN_REQUIRE(target && fwd_out3 && dpred && B > 0 && Cpred > 0 && Ctarget > 0 && HW > 0 && HW % 4 == 0, "dice_bwd: bad args");
  dice_bwd_kernel<<<loss_grid((long long)B * Cpred * HW / 4), LOSS_TPB, 0, (cudaStream_t)stream>>>(
      target, fwd_out3, grad_out, dpred, B, Cpred, (long long)Ctarget * HW, HW, smooth);
  HN_CHECK_LAUNCH("dice_bwd");
  return HN_OK;
}

extern "C" int hn_mse_fwd(const float* a, const float* b, long long n, float* ws, float* out2, void* stream) {
  HN_REQUIRE(a && b && ws && out2 && n > 0, "mse_fwd: bad args");
  HN_REQUIRE(((uintptr_t)a | (uintptr_t)b) % 16 == 0, "mse_fwd: tensors must be 16B aligned");
  mse_fwd_kernel<<<loss_grid(n / 4 + 1), LOSS_TPB, 0, (cudaStream_t)stream>>>(a, b, n, ws, out2);
  HN_CHECK_LAUNCH("mse_fwd");
  return HN_OK;
}

extern "C" int hn_mse_bwd(const float* a, const float* b, long long n, const float* grad_out, float* da, void* stream) {
  HN_REQUIRE(a && b && da && n > 0, "mse_bwd: bad args");
  HN_REQUIRE(((uintptr_t)a | (uintptr_t)b | (uintptr_t)da) % 16 == 0, "mse_bwd: tensors must be 16B aligned");
  mse_bwd_kernel<<<loss_grid(n / 4 + 1), LOSS_TPB, 0, (cudaStream_t)stream>>>(a, b, n, grad_out, da);
  HN_CHECK_LAUNCH("mse_bwd");
  return HN_OK;
}

extern "C" int hn_jaccard(const float* pred, const float* target, int B, int Cpred, int Ctarget, int HW, float* ws,
                          float* out, void* stream) {
  HN_REQUIRE(pred && target && ws && out && B > 0 && Cpred > 0 && Ctarget > 0 && HW > 0, "jaccard: bad args");
  jaccard_kernel<<<loss_grid((long long)B * HW), LOSS_TPB, 0, (cudaStream_t)stream>>>(
      pred, target, B, (long long)Cpred * HW, (long long)Ctarget * HW, HW, ws, out);
  HN_CHECK_LAUNCH("jaccard");
  return HN_OK;
}
