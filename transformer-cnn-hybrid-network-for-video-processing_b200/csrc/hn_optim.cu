// Fused AdamW over a flat fp32 arena (torch.optim.AdamW semantics, FCT.py:305,338):
//   p *= 1 - lr*wd;  m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;
//   p -= lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps)
// One launch for the whole model instead of ~300 per-tensor launches; grad_scale
// folds the 1/world_size of the data-parallel all-reduce into the same pass.
#include "hn_common.cuh"

namespace {
// step_dev != nullptr: the step count lives on the device ([0] = steps done so far, [1] = CTA-done counter); the
// bias corrections are derived from it in-kernel and the last CTA to finish bumps it, so a CUDA graph that
// contains the optimizer step replays with the right corrections (a host-side step count would be frozen at
// capture time).
__global__ void adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                             float* __restrict__ v, long long n, float lr, float b1, float b2, float eps, float wd,
                             float bc1, float bc2s, float gscale, long long* __restrict__ step_dev) {
  if (step_dev) {
    const float t = (float)(step_dev[0] + 1);
    bc1 = 1.f - powf(b1, t);
    bc2s = sqrtf(1.f - powf(b2, t));
  }
  long long nv = n / 4;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i], gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
    float* P = &pp.x; float* G = &gg.x; float* M = &mm.x; float* V = &vv.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float gr = G[j] * gscale;
      P[j] *= 1.f - lr * wd;
      M[j] = b1 * M[j] + (1.f - b1) * gr;
      V[j] = b2 * V[j] + (1.f - b2) * gr * gr;
      P[j] -= (lr / bc1) * M[j] / (sqrtf(V[j]) / bc2s + eps);
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  long long tail = nv * 4 + blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (tail < n) {
    float gr = g[tail] * gscale;
    float pv = p[tail] * (1.f - lr * wd);
    float mv = b1 * m[tail] + (1.f - b1) * gr;
    float vv2 = b2 * v[tail] + (1.f - b2) * gr * gr;
    p[tail] = pv - (lr / bc1) * mv / (sqrtf(vv2) / bc2s + eps);
    m[tail] = mv;
    v[tail] = vv2;
  }
  if (step_dev) {
    __syncthreads();                      // every thread of this CTA has read step_dev[0]
    if (threadIdx.x == 0) {
      const unsigned long long prev = atomicAdd(reinterpret_cast<unsigned long long*>(step_dev + 1), 1ull);
      if (prev == gridDim.x - 1) {        // all CTAs have read the step: bump it, reset the counter
        step_dev[1] = 0;
        step_dev[0] += 1;
      }
    }
  }
}
}  // namespace

extern "C" int hn_adamw_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                             float beta1, float beta2, float eps, float weight_decay, int step, float grad_scale,
                             void* stream) {
  HN_REQUIRE(param && grad && exp_avg && exp_avg_sq && n > 0 && step >= 1, "adamw_step: bad args");
  HN_REQUIRE(((uintptr_t)param | (uintptr_t)grad | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) % 16 == 0, "adamw_step: arenas must be 16B aligned");
  float bc1 = 1.f - powf(beta1, (float)step);
  float bc2s = sqrtf(1.f - powf(beta2, (float)step));
  long long b = (n / 4 + 255) / 256;
  if (b > 148 * 8) b = 148 * 8;
  if (b < 1) b = 1;
  adamw_kernel<<<(int)b, 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps,
                                                       weight_decay, bc1, bc2s, grad_scale, nullptr);
  HN_CHECK_LAUNCH("adamw");
  return HN_OK;
}

extern "C" int hn_adamw_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                                 float beta1, float beta2, float eps, float weight_decay, long long* step_dev,
                                 float grad_scale, void* stream) {
  HN_REQUIRE(param && grad && exp_avg && exp_avg_sq && step_dev && n > 0, "adamw_step_dev: bad args");
  HN_REQUIRE(((uintptr_t)param | (uintptr_t)grad | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) % 16 == 0, "adamw_step_dev: arenas must be 16B aligned");
  long long b = (n / 4 + 255) / 256;
  if (b > 148 * 8) b = 148 * 8;
  if (b < 1) b = 1;
  adamw_kernel<<<(int)b, 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps,
                                                       weight_decay, 1.f, 1.f, grad_scale, step_dev);
  HN_CHECK_LAUNCH("adamw_dev");
  return HN_OK;
}
