// Output head of the mask nets: the final activation and the NHWC bf16 -> NCHW fp32 boundary in ONE pass each way.
//   forward   out[b,c,h,w] (fp32) = act(z[b,h,w,c])          z = the last conv's (or BatchNorm's) pre-activation, bf16
//   backward  dz[b,h,w,c] (bf16)  = dout[b,c,h,w] * act'(out)  with act' taken from the fp32 output
// Replaces "activation in the conv epilogue, mask stored in bf16, separate layout pass" (FCT.py:205 sigmoid(conv3),
// UNet.py:51 sigmoid(conv), pyc:AE_32K L168 Sigmoid(BatchNorm)).  Why it matters beyond the saved pass: a TRAINED mask
// saturates, and sigma' = y (1 - y) computed from a bf16-stored y has only 2^-8 absolute resolution in (1 - y) --
// every pixel with y > 0.998 lost its gradient entirely, a systematic error of several per cent on every gradient of
// the net (measured on the trained-weights fixture); from the fp32 output sigma' is exact to fp32.
#include "hn_common.cuh"

namespace {
constexpr int TPB = 256;
inline int grid_for(long long items) {
  long long b = (items + TPB - 1) / TPB;
  if (b > 148 * 12) b = 148 * 12;
  return b < 1 ? 1 : (int)b;
}

__global__ void __launch_bounds__(TPB) head_fwd_kernel(const bf16* __restrict__ z, float* __restrict__ out, int B, int C,
                                                      long long HW, int ldz, int act) {
  const long long total = (long long)B * HW;
  for (long long i = blockIdx.x * (long long)TPB + threadIdx.x; i < total; i += (long long)gridDim.x * TPB) {
    const int b = (int)(i / HW);
    const long long r = i - (long long)b * HW;
    for (int c0 = 0; c0 < C; c0 += 8) {
      float f[8];
      load8(z + i * ldz + c0, f);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (c0 + j < C) out[((long long)b * C + c0 + j) * HW + r] = apply_act(f[j], act);
    }
  }
}

__global__ void __launch_bounds__(TPB) head_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ out,
                                                      bf16* __restrict__ dz, int B, int C, long long HW, int lddz, int Cpad,
                                                      int act) {
  const long long total = (long long)B * HW;
  for (long long i = blockIdx.x * (long long)TPB + threadIdx.x; i < total; i += (long long)gridDim.x * TPB) {
    const int b = (int)(i / HW);
    const long long r = i - (long long)b * HW;
    for (int c0 = 0; c0 < Cpad; c0 += 8) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float g = 0.f;
        if (c0 + j < C) {
          const long long idx = ((long long)b * C + c0 + j) * HW + r;
          g = dout[idx];
          const float y = out[idx];
          if (act == HN_ACT_SIGMOID) g *= y * (1.f - y);
          else if (act == HN_ACT_RELU) g = y > 0.f ? g : 0.f;
        }
        f[j] = g;
      }
      store8(dz + i * lddz + c0, f);
    }
  }
}
}  // namespace

extern "C" int hn_head_fwd(const void* z, float* out, int B, int C, int H, int W, int ldz, int act, void* stream) {
  HN_REQUIRE(z && out && B > 0 && C > 0 && H > 0 && W > 0 && ldz % 8 == 0 && ldz >= ((C + 7) & ~7), "head_fwd: bad args");
  HN_REQUIRE(act == HN_ACT_NONE || act == HN_ACT_RELU || act == HN_ACT_SIGMOID, "head_fwd: none / relu / sigmoid");
  head_fwd_kernel<<<grid_for((long long)B * H * W), TPB, 0, (cudaStream_t)stream>>>((const bf16*)z, out, B, C, (long long)H * W, ldz, act);
  HN_CHECK_LAUNCH("head_fwd");
  return HN_OK;
}

extern "C" int hn_head_bwd(const float* dout, const float* out, void* dz, int B, int C, int H, int W, int lddz, int act,
                           void* stream) {
  HN_REQUIRE(dout && out && dz && B > 0 && C > 0 && H > 0 && W > 0 && lddz % 8 == 0 && lddz >= ((C + 7) & ~7), "head_bwd: bad args");
  HN_REQUIRE(act == HN_ACT_NONE || act == HN_ACT_RELU || act == HN_ACT_SIGMOID, "head_bwd: none / relu / sigmoid");
  head_bwd_kernel<<<grid_for((long long)B * H * W), TPB, 0, (cudaStream_t)stream>>>(dout, out, (bf16*)dz, B, C, (long long)H * W,
                                                                                lddz, (C + 7) & ~7, act);
  HN_CHECK_LAUNCH("head_bwd");
  return HN_OK;
}
