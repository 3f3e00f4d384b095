// tcgen05 / TMEM / TMA implicit-GEMM 3x3 convolution -- placeholder until the
// kernel lands: reports "unsupported" so the host routes to the mma.sync path.
#include "hn_common.cuh"
extern "C" int hn_conv3x3_tc_supported(const HnConv* d) { (void)d; return 0; }
extern "C" int hn_conv3x3_tc_fwd(const HnConv* d, const void* x1, const void* x2, const void* wpacked, const float* bias,
                                 void* y, float* stat_sum, float* stat_sqsum, void* stream) {
  (void)d; (void)x1; (void)x2; (void)wpacked; (void)bias; (void)y; (void)stat_sum; (void)stat_sqsum; (void)stream;
  HN_FAIL(HN_EUNSUPPORTED, "conv3x3_tc_fwd: not built");
}
