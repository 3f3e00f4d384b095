/* libhybridnet_sm100.so -- C ABI of the B200 (sm_100a) hot path of the hybrid
 * segmentation nets (FCT / UNet / AutoEncoder / Autoencoder32K).
 *
 * The reference (spygaurad/Transformer-CNN-Hybrid-Network-for-Video-Processing)
 * has NO FFI: its hot path is `model(x)` / `loss.backward()` over torch.nn
 * layers (FCT.py:330,337; Model.py:56,58).  Each entry point below replaces the
 * ATen/cuDNN call behind one of those layers; the reference line it stands in
 * for is cited beside it.  The host side (package `hybridnet`, Python, because
 * the reference is Python) binds these with ctypes -- see INTEGRATION.md.
 *
 * Conventions
 *  - every activation is NHWC bf16 in device memory; a tensor is (ptr, ld) where
 *    ld = elements between consecutive pixels (>= C, multiple of 8, so a channel
 *    slice of a wider buffer is a tensor too -- that is how concat is free);
 *  - parameters / parameter gradients are fp32 in the reference's own layout
 *    (OIHW etc.), i.e. exactly the state_dict tensors;
 *  - the caller owns all memory; no entry point allocates, synchronises or
 *    keeps state, so every call is CUDA-graph capturable;
 *  - `stream` is a cudaStream_t passed as void*;
 *  - return 0 on success, negative HN_E* otherwise; hn_last_error() gives the
 *    thread-local message.
 */
#ifndef HYBRIDNET_H_
#define HYBRIDNET_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define HN_OK 0
#define HN_EINVAL (-1)
#define HN_ECUDA (-2)
#define HN_EUNSUPPORTED (-3)

enum { HN_ACT_NONE = 0, HN_ACT_RELU = 1, HN_ACT_GELU = 2, HN_ACT_SIGMOID = 3 };

const char* hn_last_error(void);
int hn_version(void);
/* device sanity: returns 0 iff the current device is sm_100 */
int hn_check_device(void);

/* ---- layout (module boundary: x.to(device) FCT.py:326; final mask) ---------- */
/* [B,C,H,W] fp32 -> [B,H,W,ld] bf16, channels >= C zero-filled up to Cpad */
int hn_nchw_f32_to_nhwc_bf16(const float* src, void* dst, int B, int C, int H, int W,
                             int Cpad, int ld, void* stream);
int hn_nhwc_bf16_to_nchw_f32(const void* src, float* dst, int B, int C, int H, int W,
                             int ld, void* stream);
/* Output head: final activation (none / ReLU / sigmoid) + NHWC bf16 -> NCHW fp32 in one pass (FCT.py:205,
 * UNet.py:51, pyc:AE_32K L168), and its backward dz (NHWC bf16, padded channels zero) = dout * act'(out) with act'
 * from the fp32 output (a bf16-stored saturated mask loses sigma' = y(1-y) for y > 0.998).              */
int hn_head_fwd(const void* z, float* out, int B, int C, int H, int W, int ldz, int act, void* stream);
int hn_head_bwd(const float* dout, const float* out, void* dz, int B, int C, int H, int W, int lddz, int act,
                void* stream);
/* AvgPool2d(2,2) on NHWC bf16 (FCT.py:238-240 pyramid) */
int hn_avgpool2_fwd(const void* x, void* y, int B, int H, int W, int C, int ldx, int ldy,
                    void* stream);

/* ---- convolution as implicit GEMM ----------------------------------------- */
/* Replaces nn.Conv2d / nn.ConvTranspose2d forward and input-gradient
 * (FCT.py:88,110-113,140-143,170-172,194-196; UNet.py:19-26,58-61;
 *  AutoEncoder.py:9-12,36-38,57-59; pyc:AE_32K L24-31,L65-81,L123-150).       */
typedef struct HnConv {
  int B, H, W;            /* logical input extent (after the optional x2 upsample) */
  int C1, C2;             /* channels taken from x1 then x2 (virtual concat); %8==0 */
  int ld1, ld2;           /* pixel strides of x1 / x2                              */
  int OH, OW, Cout;       /* output extent and channels (any Cout >= 1)           */
  int ldy;                /* pixel stride of y                                     */
  int KH, KW, stride, pad, dil;
  int mode;               /* 0: ih = oh*stride - pad + kh*dil  (forward conv)
                             1: ih = (oh + pad - kh*dil)/stride when divisible
                                (input gradient / transposed conv)                */
  int ups;                /* 1: x is read through nearest Upsample(x2): x[ih>>1]  */
  int act;                /* HN_ACT_* applied after bias (+ residual)             */
  int pshuf;              /* >0: GEMM column n = tap*pshuf + co is stored to pixel
                             (2*oh + tap/2, 2*ow + tap%2), channel co (ConvT k2s2) */
  int ldr;                /* pixel stride of residual (if res != NULL)            */
  int Kpad;               /* row length of packed weights (from hn_conv_kpad)     */
} HnConv;

int hn_conv_kpad(int KH, int KW, int Cin_pad);   /* packed row length in elements   */
int hn_conv_npad(int Cout);                      /* packed row count                 */
/* fp32 parameter -> packed bf16 [Npad][Kpad], k = (kh*KW+kw)*Cpad + c.
 * element (n,c,kh,kw) is read from src[(n/nmod)*s_nhi + (n%nmod)*s_n + c*s_c + kh*s_kh + kw*s_kw];
 * c >= Creal or n >= N is written as zero. */
int hn_pack_weight(const float* src, void* dst, int N, int Creal, int Cpad, int KH, int KW,
                   long long nmod, long long s_nhi, long long s_n, long long s_c,
                   long long s_kh, long long s_kw, void* stream);
/* Every weight pack of a model in one launch.  jobs = DEVICE array of njobs records of 16 x int64:
 *   { src, dst, N, Npad, Creal, Cpad, KH, KW, Kpad, nmod, s_nhi, s_n, s_c, s_kh, s_kw, s2k }
 * = the arguments of hn_pack_weight (Npad = hn_conv_npad(N), Kpad = hn_conv_kpad(KH,KW,Cpad)); s2k != 0 selects
 * the hn_pack_weight_s2dgrad layout instead (N = Cout, Creal = Cin, s2k = K, Npad/Kpad as that call defines). */
int hn_pack_weight_multi(const void* jobs, int njobs, long long max_elems /* largest Npad*Kpad */, void* stream);
/* Packed B operand for the input gradient of a stride-2 conv (k in {1,3}, pad k/2; the reference's
 * Bottleneck conv2 / downsample, pyc:AE_32K:L20-53) run as one stride-1 tcgen05 GEMM over dz with a 2x2
 * (k3) or 1x1 (k1) window and a pixel-shuffle store: rows (py,px,ci), columns (dh,dw,co).
 * src = Conv2d weight [Cout][Cin][k][k] fp32; dst = bf16 [hn_conv_npad(4*Cin)][hn_conv_kpad(w,w,Cout)]. */
int hn_pack_weight_s2dgrad(const float* src, void* dst, int Cout, int Cin, int K, void* stream);
/* y = act(conv(x1|x2, w) + bias [+ res]); optional per-channel sum / sum-of-squares
 * of the stored (bf16-rounded, pre-activation) values for BatchNorm.            */
int hn_conv2d_fwd(const HnConv* d, const void* x1, const void* x2, const void* wpacked,
                  const float* bias, const void* res, void* y, float* stat_sum,
                  float* stat_sqsum, void* stream);
/* dW (fp32, += into a zeroed buffer) for the same HnConv geometry (mode 0 only):
 * dW[n,c,kh,kw] += sum_p dy[p,n] * x[p@(kh,kw), c], written at
 * dw[n*s_n + c*s_c + kh*s_kh + kw*s_kw]; c >= Creal skipped.                     */
int hn_conv2d_wgrad(const HnConv* d, const void* x1, const void* x2, const void* dy,
                    float* dw, int Creal, long long s_n, long long s_c, long long s_kh,
                    long long s_kw, void* stream);

/* ---- pointwise / reductions ------------------------------------------------ */
/* dz = dy * act'(.) (ReLU/sigmoid from the saved output y, GELU from the saved
 * pre-activation); optional fp32 column sums of dz (bias gradient, += )         */
int hn_act_bwd(const void* dy, const void* y_or_z, void* dz, long long P, int C, int lddy,
               int ldy, int lddz, int act, float* dbias, void* stream);
/* y = act(x) (GELU after a conv whose pre-activation must be kept for backward) */
int hn_act_fwd(const void* x, void* y, long long P, int C, int ldx, int ldy, int act, void* stream);
int hn_colsum(const void* x, long long P, int C, int ld, float* out, void* stream);
/* y = a + b (+ c) elementwise on [P,C] tensors */
int hn_add(const void* a, const void* b, const void* c, void* y, long long P, int C, int lda,
           int ldb, int ldc, int ldy, void* stream);
/* y = gelu(z1)+gelu(z2)+gelu(z3) (Wide_Focus, FCT.py:118-129) and its backward */
int hn_gelu3_sum_fwd(const void* z1, const void* z2, const void* z3, void* y, long long n,
                     void* stream);
int hn_gelu3_sum_bwd(const void* dy, const void* z1, const void* z2, const void* z3, void* d1,
                     void* d2, void* d3, long long n, void* stream);
/* Wide_Focus glue fused (FCT.py:117-132, dropout 0.1 as hn_dropout's hash with the same step-counter mixing):
 *   combine: y = sum_i dropout_i(gelu(z_i + b_i)), i=1..3 (z_i = the dilated convs run WITHOUT bias; b_i fp32 [C] or NULL)
 *   out:     out = res + dropout(gelu(z + bias))   (res may be NULL)
 * backward: dz_i = dy * keep_i/(1-p) * gelu'(z_i + b_i); dbias_i (fp32 [C], accumulated, may be NULL) = column sums. */
int hn_wf_combine_fwd(const void* z1, const void* z2, const void* z3, const float* b1, const float* b2,
                      const float* b3, void* y, long long P, int C, int ldz, int ldy, float p,
                      unsigned long long seed1, unsigned long long seed2, unsigned long long seed3,
                      const unsigned long long* step_counter, void* stream);
int hn_wf_combine_bwd(const void* dy, const void* z1, const void* z2, const void* z3, const float* b1,
                      const float* b2, const float* b3, void* d1, void* d2, void* d3, long long P, int C,
                      int lddy, int ldz, int ldd, float p, unsigned long long seed1, unsigned long long seed2,
                      unsigned long long seed3, const unsigned long long* step_counter, float* dbias1,
                      float* dbias2, float* dbias3, void* stream);
int hn_wf_out_fwd(const void* z, const float* bias, const void* res, void* out, long long P, int C, int ldz, int ldr,
                  int ldo, float p, unsigned long long seed, const unsigned long long* step_counter, void* stream);
int hn_wf_out_bwd(const void* dout, const void* z, const float* bias, void* dz, long long P, int C, int lddo, int ldz,
                  int lddz, float p, unsigned long long seed, const unsigned long long* step_counter, float* dbias,
                  void* stream);
/* MaxPool2d(2) (FCT.py:147; UNet.py:12) forward / backward (first max wins) */
int hn_maxpool2_fwd(const void* x, void* y, int B, int H, int W, int C, int ldx, int ldy,
                    void* stream);
int hn_maxpool2_bwd(const void* x, const void* dy, void* dx, int B, int H, int W, int C, int ldx,
                    int lddy, int lddx, void* stream);
/* gradient of nearest Upsample(x2): dx[h,w] = sum of the 2x2 block of dy */
int hn_upsample2_bwd(const void* dy, void* dx, int B, int H, int W, int C, int lddy, int lddx,
                     void* stream);
/* Dropout / Dropout2d: y = x * keep/(1-p); keep bit = hash(seed', idx) (idx = element,
 * or (b,c) when channelwise).  The same call with dy gives the backward.
 * seed' = seed when step_counter is NULL, else seed + *step_counter * 0x9E3779B97F4A7C15:
 * step_counter is a DEVICE u64 read by the kernel, so that a training step captured once in a
 * CUDA graph still draws a fresh mask on every replay (the graph bumps the counter itself).   */
int hn_dropout(const void* x, void* y, int B, long long HW, int C, int ldx, int ldy, float p,
               unsigned long long seed, const unsigned long long* step_counter, int channelwise,
               void* stream);

/* ---- LayerNorm over C (FCT.py:32-36,89) ------------------------------------- */
int hn_layernorm_fwd(const void* x, void* y, const float* gamma, const float* beta, long long P,
                     int C, int ldx, int ldy, float eps, void* stream);
/* dx, and += dgamma/dbeta (fp32[C]); recomputes mean/rstd from x */
int hn_layernorm_bwd(const void* x, const void* dy, void* dx, const float* gamma, float* dgamma,
                     float* dbeta, long long P, int C, int ldx, int lddy, int lddx, float eps,
                     void* stream);

/* ---- BatchNorm2d, training and eval (UNet.py:59,62; pyc:AE_32K) ------------- */
int hn_bn_stats(const void* x, long long P, int C, int ld, float* sum, float* sqsum, void* stream);
/* mean/rstd from sums; running stats update with momentum and unbiased var      */
int hn_bn_finalize(const float* sum, const float* sqsum, long long count, int C, float eps,
                   float momentum, float* mean, float* rstd, float* running_mean,
                   float* running_var, long long* num_batches_tracked /* device i64 or NULL: += 1 */,
                   void* stream);
/* y = act((x-mean)*rstd*gamma+beta [+ res]) */
int hn_bn_apply(const void* x, const void* res, void* y, const float* mean, const float* rstd,
                const float* gamma, const float* beta, long long P, int C, int ldx, int ldr,
                int ldy, int act, void* stream);
/* hn_bn_finalize + hn_bn_apply in ONE launch (training): every CTA derives mean / rstd from the sums (count = pixels per
 * channel), CTA 0 writes them to mean / rstd for the backward pass and updates running_mean / running_var (momentum,
 * unbiased variance) and num_batches_tracked exactly as hn_bn_finalize does.  UNet.py:58-63, pyc:AE_32K Bottleneck. */
int hn_bn_apply_train(const void* x, const void* res, void* y, const float* sum, const float* sqsum, long long count,
                      float eps, float momentum, float* mean, float* rstd, float* running_mean, float* running_var,
                      long long* num_batches_tracked, const float* gamma, const float* beta, long long P, int C,
                      int ldx, int ldr, int ldy, int act, void* stream);
/* pass 1: dz = dy*act'(y) (written to dz), s1 += sum dz, s2 += sum dz*xhat      */
int hn_bn_bwd_reduce(const void* dy, const void* y, const void* x, void* dz, const float* mean,
                     const float* rstd, long long P, int C, int lddy, int ldy, int ldx, int lddz,
                     int act, float* s1, float* s2, void* stream);
/* pass 2: dx = gamma*rstd*(dz - s1/P - xhat*s2/P) (training) or gamma*rstd*dz (eval) */
int hn_bn_bwd_apply(const void* dz, const void* x, void* dx, const float* mean, const float* rstd,
                    const float* gamma, const float* s1, const float* s2, long long P, int C,
                    int lddz, int ldx, int lddx, int training, void* stream);

/* BatchNorm + ReLU without residual (every UNet.py:58-63 block, most of pyc:AE_32K): the ReLU mask is
 * re-derived from x with the exact fp32 expression hn_bn_apply used, so backward reads dy and x only
 * (reduce: 2 passes, apply: 3) instead of dy,y,x(+dz). */
int hn_bn_relu_bwd_reduce(const void* dy, const void* x, const float* mean, const float* rstd,
                          const float* gamma, const float* beta, long long P, int C, int lddy,
                          int ldx, float* s1, float* s2, void* stream);
int hn_bn_relu_bwd_apply(const void* dy, const void* x, void* dx, const float* mean,
                         const float* rstd, const float* gamma, const float* beta, const float* s1,
                         const float* s2, long long P, int C, int lddy, int ldx, int lddx,
                         int training, void* stream);

/* ---- conv-projected attention (FCT.py:24-79) -------------------------------- */
/* q,k,v = LayerNorm_C(ReLU(dwconv3x3(x)+b)) for the three projections in one pass.
 * w, b, gamma, beta: HOST arrays of three device pointers (q, k, v) to the reference's own parameter tensors
 * (conv_q/k/v.weight [C,1,3,3] fp32, .bias [C] -- the array or its entries may be null --, layernorm_q/k/v.weight /
 * .bias [C]); out: [3] tensors of [P,C]                                                            */
int hn_dwconv_relu_ln_fwd(const void* x, void* q, void* k, void* v, const float* const* w, const float* const* b,
                          const float* const* gamma, const float* const* beta, int B, int H, int W, int C,
                          int ldx, int ldo, float eps, void* stream);
/* backward: dpre[3] (grad at the dwconv output, after LN+ReLU backward), += dgamma/dbeta/db (arrays of three fp32 [C]
 * accumulation targets, e.g. slices of a gradient arena; db may be null)                            */
int hn_dwconv_relu_ln_bwd_pre(const void* x, const void* dq, const void* dk, const void* dv,
                              void* dpre_q, void* dpre_k, void* dpre_v, const float* const* w,
                              const float* const* b, const float* const* gamma, float* const* dgamma,
                              float* const* dbeta, float* const* db, int B, int H, int W, int C, int ldx, int ldd,
                              float eps, void* stream);
/* dx = sum_t dwconv^T(dpre_t, w_t);  dw[t][C][9] += sum_p dpre_t[p,c] * x[p@tap,c]  (three targets) */
int hn_dwconv_bwd_data_weight(const void* x, const void* dpre_q, const void* dpre_k,
                              const void* dpre_v, void* dx, const float* const* w, float* const* dw, int B, int H,
                              int W, int C, int ldx, int ldd, int lddx, void* stream);
/* softmax(Q K^T * scale) V per head, flash style.  q,k,v,o: [B,N,heads*D] with
 * pixel strides ld*; lse: fp32 [B,heads,N] (log-sum-exp, natural log).
 * D in {4,8,16,32,64}.  Replaces nn.MultiheadAttention's core (FCT.py:37,75).   */
int hn_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int N,
                int heads, int D, int ldq, int ldk, int ldv, int ldo, float scale, void* stream);
/* tcgen05 / TMEM / TMA form of hn_attn_fwd for the exponential-bound shapes: D in {4, 8}, two heads, N %% 128 == 0,
 * 16-byte aligned rows.  K / V tiles arrive as TMA boxes, S = Q K^T and O += P [V | 1] run on tcgen05.mma (S, P and O
 * live in TMEM; the ones panel makes the tensor core deliver the softmax row sum), the softmax warps read S with
 * tcgen05.ld.  poly_every in {0, 2, 3, 4}: share of the exponentials taken on the FMA pipe.             */
int hn_attn_fwd_tc_supported(int N, int D, int ldq, int ldk, int ldv, int ldo);
int hn_attn_fwd_tc(const void* q, const void* k, const void* v, void* o, float* lse, int B, int N, int heads, int D,
                   int ldq, int ldk, int ldv, int ldo, float scale, int poly_every, void* stream);
/* tcgen05 / TMEM / TMA one-pass backward for the same shapes (csrc/hn_attn_tc.cu): CTA = 128 keys, dK / dV accumulate in
 * TMEM, the per-query lse / delta terms ride the idle upper half of the K = 16 contraction as three-term bf16 splits, dQ
 * leaves through an fp32 accumulator.  workspace (fp32 elements) = hn_attn_bwd_tc_workspace().             */
long long hn_attn_bwd_tc_workspace(int B, int N, int heads, int D);
int hn_attn_bwd_tc_supported(int N, int D, int ldq, int ldk, int ldv, int ldo, int lddo, int lddq, int lddk, int lddv);
int hn_attn_bwd_tc(const void* q, const void* k, const void* v, const void* o, const void* dout, const float* lse,
                   float* workspace, long long workspace_floats, void* dq, void* dk, void* dv, int B, int N, int heads,
                   int D, int ldq, int ldk, int ldv, int ldo, int lddo, int lddq, int lddk, int lddv, float scale,
                   int poly_every, void* stream);
/* hn_attn_bwd_tc without a caller workspace: the scratch (auxiliary panels + fp32 dQ accumulator) is a file-scope device
 * array of HN_ATTN_BWD_TC_STATIC_MAX_FLOATS fp32 (8 MiB, resident on every GPU that loads the library), so the call
 * allocates nothing: for code under CUDA graph capture, serving loops and harnesses that forbid cudaMalloc.  No
 * counterpart in the reference.  Shapes as hn_attn_bwd_tc, with
 * B * N * (16 + 2 D) <= HN_ATTN_BWD_TC_STATIC_MAX_FLOATS (e.g. B = 4, N = 16384, D = 8); larger sizes are rejected with
 * HN_EINVAL.  Every call shares the array (the entry point re-initialises what it needs): calls that overlap on the GPU race. */
#define HN_ATTN_BWD_TC_STATIC_MAX_FLOATS (2 * 1024 * 1024)
int hn_attn_bwd_tc_static(const void* q, const void* k, const void* v, const void* o, const void* dout, const float* lse,
                          void* dq, void* dk, void* dv, int B, int N, int heads, int D, int ldq, int ldk, int ldv,
                          int ldo, int lddo, int lddq, int lddk, int lddv, float scale, int poly_every, void* stream);
/* forward, D <= 8: every poly_every-th 8-key tile computes exp2 on the FMA pipe (polynomial) instead of MUFU;
 * 0 = never, default 4; process-wide */
int hn_attn_tune_fwd(int poly_every);
/* tuning knob of the one-pass backward (D <= 8): keys per CTA, 64 or 128 (default 128); process-wide */
int hn_attn_tune(int fused_bwd_keys_per_cta);
/* workspace (fp32 elements, caller-owned, uninitialised): delta [B,heads,N] and, for D <= 8, the fp32 dQ
 * accumulator of the one-pass kernel */
long long hn_attn_bwd_workspace(int B, int N, int heads, int D);
int hn_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout,
                const float* lse, float* workspace, long long workspace_floats, void* dq, void* dk, void* dv, int B, int N,
                int heads, int D, int ldq, int ldk, int ldv, int ldo, int lddo, int lddq, int lddk,
                int lddv, float scale, void* stream);

/* ---- tcgen05/TMEM implicit-GEMM convolution (C >= 64 layers) ---------------- */
/* Same contract as hn_conv2d_fwd (no residual) for the geometries the tensor-core path takes:
 * stride 1, "same" output extent, C1%64==0, C2%64==0, GEMM N a multiple of 32 (>= 64), pixel
 * tiles of 128 (W a power of two <= 128 or a multiple of 128); 1x1 / 3x3 / dilated, ConvT k2s2
 * through pshuf.  A: 4-D TMA boxes shifted per tap (OOB zero fill = padding), B: packed weights,
 * accumulators in TMEM, persistent CTAs.  hn_conv_tc_supported() tells the host which path to use. */
int hn_conv_tc_supported(const HnConv* d);
/* 1 when hn_conv_tc_fwd takes its halo variant for d (3x3 stride 1, W % 128 == 0, weights resident in smem:
 * one TMA box of 128+2*pad pixels per filter row instead of one per tap) -- the host prefers it over
 * hn_conv_sc_fwd for the C = 32 / 64 layers. */
int hn_conv_tc_halo_supported(const HnConv* d);
int hn_conv_tc_fwd(const HnConv* d, const void* x1, const void* x2, const void* wpacked,
                      const float* bias, void* y, float* stat_sum, float* stat_sqsum,
                      void* stream);

/* ---- small-channel direct convolution (Cin <= 64, Cout <= 64, Cin*Cout <= 2048) -------------- */
/* Same contract as hn_conv2d_fwd (no residual) for stride-1 "same" 1x1 / 3x3 (dilated) convs on maps
 * with H % 16 == 0, W % 32 == 0: all of FCT.py:88,110-113,140-143,170-172,194-196,205 and
 * AutoEncoder.py:9-12,36-38,57-59, UNet.py level 1.  HBM-bound layers: the input halo of a 16x32
 * tile is staged once in shared memory and the taps are shifted ldmatrix reads of it. */
int hn_conv_sc_supported(const HnConv* d);
int hn_conv_sc_fwd(const HnConv* d, const void* x1, const void* x2, const void* wpacked,
                   const float* bias, void* y, float* stat_sum, float* stat_sqsum, void* stream);
/* the same with the residual epilogue of hn_conv2d_fwd (y = act(conv + bias + res), d->ldr = pixel stride of res): the
 * Transformer's `conv1(x1) + x` (FCT.py:93-95).  d->ups = 1 (x1 is the H/2 x W/2 map read through nearest-neighbour 2x
 * upsampling, C2 = 0: Block_decoder's Upsample -> conv, FCT.py:167-172) is accepted by both entry points and by
 * hn_conv_wgrad_sc. */
int hn_conv_sc_fwd_res(const HnConv* d, const void* x1, const void* x2, const void* wpacked,
                       const float* bias, const void* res, void* y, float* stat_sum,
                       float* stat_sqsum, void* stream);

/* small-channel weight gradient: same contract as hn_conv2d_wgrad for the hn_conv_sc_fwd geometries (dw is accumulated
 * into with fp32 atomics: zero it first).  Large flushes are launched as thread-block clusters of 8 CTAs that sum dW
 * through distributed shared memory before the atomics; graph capturable, allocates nothing. */
int hn_conv_wgrad_sc_supported(const HnConv* d);
int hn_conv_wgrad_sc(const HnConv* d, const void* x1, const void* x2, const void* dy, float* dw,
                     int Creal, long long s_n, long long s_c, long long s_kh, long long s_kw,
                     void* stream);

/* tcgen05 weight gradient: same contract as hn_conv2d_wgrad for stride 1/2 convs with C1%64==0,
 * C2%64==0, Cout>=64 and 64-pixel tileable output maps.  Per tap: D[Cout,Cin] = dz^T x_tap with both
 * operands MN-major straight from TMA boxes, accumulators in TMEM, split over pixels + fp32 atomics. */
int hn_conv_wgrad_tc_supported(const HnConv* d);
/* workspace (fp32 elements, caller-owned, uninitialised) = hn_conv_wgrad_tc_workspace(d, s_c): with it the
 * per-split partial sums are stored and reduced by a second pass instead of scalar atomics (3x3 weights);
 * NULL / too small / s_c == 1 (1x1 weights, vector reductions): atomics straight into dw.           */
long long hn_conv_wgrad_tc_workspace(const HnConv* d, long long s_c);
int hn_conv_wgrad_tc(const HnConv* d, const void* x1, const void* x2, const void* dy, float* dw,
                     int Creal, long long s_n, long long s_c, long long s_kh, long long s_kw,
                     float* workspace, long long workspace_floats, void* stream);

/* ---- fused multi-tensor AdamW on a flat fp32 arena (FCT.py:305,338) --------- */
int hn_adamw_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq,
                  long long n, float lr, float beta1, float beta2, float eps, float weight_decay,
                  int step, float grad_scale, void* stream);

/* same update with the step count kept on the device: step_dev[0] = steps done so far (the kernel derives the
 * bias corrections from it and the last CTA bumps it), step_dev[1] = scratch counter (zero).  This is the form a
 * CUDA graph can replay (torch.optim.AdamW.step, FCT.py:338, under parallel.CapturedStep).           */
int hn_adamw_step_dev(float* param, const float* grad, float* exp_avg, float* exp_avg_sq,
                      long long n, float lr, float beta1, float beta2, float eps, float weight_decay,
                      long long* step_dev, float grad_scale, void* stream);

/* ---- losses / metrics: the step right after forward -------- */
/* All tensors NCHW fp32 as the reference's modules return them.  ws = float[16], zero before the first call
 * (self-cleaning afterwards).  One launch each: the last CTA reduces, writes the scalar, keeps what backward needs. */
/* DiceLoss.forward, Metrics.py:14-22: 1 - (2 sum(p t) + smooth) / (sum p + sum t + smooth) on channel 0 */
int hn_dice_fwd(const float* pred, const float* target, int B, int Cpred, int Ctarget, int HW, float smooth,
                float* ws, float* loss3 /* {loss, sum(p t), sum p + sum t + smooth} */, void* stream);
/* dL/dpred [B,Cpred,H,W] (zero outside channel 0); fwd_out3 = loss3 of the forward; grad_out = device scalar or NULL (= 1) */
int hn_dice_bwd(const float* target, const float* fwd_out3, const float* grad_out, float* dpred, int B, int Cpred,
                int Ctarget, int HW, float smooth, void* stream);
/* nn.MSELoss (Model.py:57 L2 term): out2[0] = mean((a-b)^2), out2[1] = PSNR = 20 log10(1/sqrt(mse)) (Model.py:30) */
int hn_mse_fwd(const float* a, const float* b, long long n, float* ws, float* out2, void* stream);
int hn_mse_bwd(const float* a, const float* b, long long n, const float* grad_out, float* da, void* stream);
/* JaccardScore.forward, Metrics.py:31-39: sum(p!=0 & t!=0) / sum(p!=0 | t!=0) on channel 0 (NaN for an empty union) */
int hn_jaccard(const float* pred, const float* target, int B, int Cpred, int Ctarget, int HW, float* ws,
               float* out, void* stream);

/* ---- input pipeline: the step right before forward --------- */
/* Dataloader.py:16-24 (ToTensor: uint8 HWC -> float/255) + Model.py:44-53 (batch hflip | image + N(0,std^2) noise
 * and up to four patch_size x patch_size blackout patches at rows patch_x / columns patch_y) in one pass.
 * src: uint8 [B,H,W,3] (src_is_u8) or fp32 [B,3,H,W].  Writes dst_nhwc (bf16 [B,H,W,ld], channels 3..7 zero: the
 * first conv's operand) and/or dst_nchw (fp32 [B,3,H,W]).  noise_std > 0 selects the noise branch, which -- as in
 * the reference -- starts from the UNFLIPPED image.  step_counter (device, may be NULL) is mixed into the seed. */
int hn_augment_pack(const void* src, int src_is_u8, void* dst_nhwc, float* dst_nchw, int B, int H, int W, int ld,
                    int flip, float noise_std, unsigned long long seed, const unsigned long long* step_counter,
                    int npatch, const int* patch_x, const int* patch_y, int patch_size, void* stream);
/* AvgPool2d(2,2) backward: dx[2h+i, 2w+j] = dy[h, w] / 4; H, W = extent of dy */
int hn_avgpool2_bwd(const void* dy, void* dx, int B, int H, int W, int C, int lddy, int lddx, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HYBRIDNET_H_ */
