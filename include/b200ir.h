/* b200ir.h — C ABI of libb200ir.so: hand-written sm_100a kernels for the Restormer-family
 * transformer block (LayerNorm -> MDTA -> LayerNorm -> GDFN) and its U-Net glue.
 *
 * The reference (1601339895/Image-Restoration) has no FFI layer: its boundary for this path is
 * the nn.Module surface of Restormer.py.  Each entry point below replaces the chain of stock
 * ATen calls of one reference forward() and cites it (file:line in /root/reference).  The
 * Python mirror of the reference modules (image-restoration_b200/modules.py) binds these with
 * ctypes; INTEGRATION.md shows the stub.
 *
 * Conventions
 *  - All activations are NCHW, channel stride = H*W elements, explicit batch stride (elements),
 *    so channel slices of a wider buffer (q|k|v chunks, concat halves) are passed without copies.
 *  - dtype: B200IR_F32 or B200IR_BF16 for activations; parameters are always fp32 device arrays
 *    in the reference's own state_dict layout (packed/bf16 copies are made by b200ir_pack_*).
 *  - The library never allocates, frees or synchronises: caller owns every buffer (including
 *    workspaces) and the stream; every call only enqueues work on `stream`.
 *  - Return 0 on success; >0 = cudaError_t; <0 = B200IR_E_* (bad argument / unsupported shape).
 *    b200ir_last_error_string() gives the thread-local message.  No C++ exception crosses.
 */
#ifndef B200IR_H
#define B200IR_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef void* b200ir_stream_t; /* cudaStream_t */

enum { B200IR_F32 = 0, B200IR_BF16 = 1 };
enum { B200IR_LN_NONE = 0, B200IR_LN_WITHBIAS = 1, B200IR_LN_BIASFREE = 2 };
enum { B200IR_SHUFFLE_NONE = 0, B200IR_UNSHUFFLE2 = 1, B200IR_SHUFFLE2 = 2 };
enum { B200IR_IMPL_AUTO = 0, B200IR_IMPL_SIMT = 1, B200IR_IMPL_TC = 2 };
enum { B200IR_E_BADARG = -1, B200IR_E_UNSUPPORTED = -2, B200IR_E_WORKSPACE = -3, B200IR_E_DRIVER = -4 };

int b200ir_version(void);
const char* b200ir_last_error_string(void);

/* y[b,o,p] = (res[b,o,p]) + bias[o] + sum_c w[b][o,c] * LN(x)[b,c,p]
 * Replaces LayerNorm.forward + the following 1x1 nn.Conv2d (+ the residual add of
 * TransformerBlock.forward): Restormer.py:37-39,54-57,68-70 + :114 (qkv), :89 (project_in),
 * :92 (project_out), :131 (project_out of attention, with the per-sample folded weight),
 * :147-148 (x = x + ...), :261,:266 (reduce_chan_level{3,2}).
 * w: fp32 [Cout,Cin] (w_bstride = 0) or per-sample [B,Cout,Cin] (w_bstride = Cout*Cin).
 * ln_type NONE: ln_w/ln_b ignored.  res may be NULL.  SIMT fp32-FMA implementation. */
int b200ir_ln_conv1x1_fwd(const void* x, int64_t x_bstride, int x_dtype,
                          const float* w, int64_t w_bstride, const float* bias,
                          int ln_type, const float* ln_w, const float* ln_b,
                          const void* res, int64_t res_bstride, int res_dtype,
                          void* y, int64_t y_bstride, int y_dtype,
                          int B, int Cin, int Cout, int64_t HW, b200ir_stream_t stream);

/* Depthwise 3x3, stride 1, zero pad 1 (+bias): Attention.qkv_dwconv, Restormer.py:106,114.
 * w fp32 [C,3,3]. */
int b200ir_dwconv3x3_fwd(const void* x, int64_t x_bstride, const float* w, const float* bias,
                         void* y, int64_t y_bstride, int dtype,
                         int B, int C, int H, int W, b200ir_stream_t stream);

/* FeedForward.dwconv + chunk(2) + gelu(x1)*x2: Restormer.py:84,90-91.
 * x has 2*Hd channels, y has Hd: y[c] = gelu_erf(dw(x)[c]) * dw(x)[c+Hd].  w fp32 [2Hd,3,3]. */
int b200ir_dwconv3x3_gelu_gate_fwd(const void* x, int64_t x_bstride, const float* w, const float* bias,
                                   void* y, int64_t y_bstride, int dtype,
                                   int B, int Hd, int H, int W, b200ir_stream_t stream);

/* MDTA stage 1 — split-K Gram over HW slabs: for every (b, head, slab) the partial
 * q.k^T [d,d] and partial sum q^2 [d], sum k^2 [d] (d = C/heads), fp32, fixed order
 * (deterministic).  Replaces F.normalize + q@k^T, Restormer.py:117-124.
 * qkv: [B,3C,HW] (q|k|v chunks, Restormer.py:115).  partials: [B,heads,nslab,d*d+2d] fp32.
 * b200ir_mdta_gram_nslab gives nslab for a shape (so the caller can size `partials`). */
int b200ir_mdta_gram_nslab(int B, int C, int heads, int64_t HW);
int b200ir_mdta_gram_fwd(const void* qkv, int64_t bstride, int dtype, int B, int C, int heads, int64_t HW,
                         float* partials, int nslab, int impl, b200ir_stream_t stream);

/* MDTA stage 2 — (a) reduce slabs, attn = softmax(q^.k^^T * temperature) (Restormer.py:121-125), then (b) fold the
 * output projection: m[b][o, h*d+j] = sum_i w_out[o, h*d+i] * attn[b,h][i,j], so that project_out(attn @ v)
 * (Restormer.py:127-131) becomes ONE 1x1 GEMM with a per-sample weight.
 * attn: fp32 [B,heads,d,d] (always written; kept for backward).  m: fp32 [B,C,C] or NULL.
 * mp: bf16 [B,nrows,kpad] operand of b200ir_gemm_tc_fwd (geometry from b200ir_gemm_tc_geometry(C,C)) or NULL. */
int b200ir_mdta_softmax_fold_fwd(const float* partials, int nslab, const float* temperature,
                                 const float* w_out, float* m, float* attn, void* mp, int kpad, int nrows,
                                 int B, int C, int heads, b200ir_stream_t stream);
/* Same with a per-sample temperature [B][heads] when temp_per_sample != 0: the dynamic temperature of the reference's
 * RectiFormer experiment (temp = temperature * dynamic_temp[B,1,1,1], experiment/RectiFormer.py:309-311). */
int b200ir_mdta_softmax_fold_fwd_ex(const float* partials, int nslab, const float* temperature, int temp_per_sample,
                                    const float* w_out, float* m, float* attn, void* mp, int kpad, int nrows,
                                    int B, int C, int heads, b200ir_stream_t stream);
/* Split-gate FFN (experiment/RectiFormer.py:231-266): out[c] = gelu(dw(t)[c]) * sigmoid(dw(t)[c + Hd]); w [2*Hd][9] =
 * dwconv_content | dwconv_gate. */
int b200ir_dwconv3x3_gelu_sigmoid_gate_fwd(const void* x, int64_t x_bstride, const float* w, const float* bias, void* y,
                                           int64_t y_bstride, int dtype, int B, int Hd, int H, int W,
                                           b200ir_stream_t stream);

/* Dense 3x3 conv, pad 1 (+bias) (+res) with PixelUnshuffle(2)/PixelShuffle(2) folded into the
 * store: OverlapPatchEmbed (Restormer.py:162-165), Downsample (:171-179), Upsample (:181-189),
 * output conv + inp_img (:281).  w fp32 [Cout,Cin,3,3].  res (optional) has the OUTPUT shape.
 * y shape: NONE [B,Cout,H,W]; UNSHUFFLE2 [B,4Cout,H/2,W/2]; SHUFFLE2 [B,Cout/4,2H,2W]. */
int b200ir_conv3x3_fwd(const void* x, int64_t x_bstride, int x_dtype, const float* w, const float* bias,
                       const void* res, int64_t res_bstride, int res_dtype,
                       void* y, int64_t y_bstride, int y_dtype,
                       int B, int Cin, int Cout, int H, int W, int shuffle_mode, b200ir_stream_t stream);

/* Strided NCHW channel-slice copy / dtype cast (used to place a skip tensor into a concat
 * buffer, Restormer.py:260,265,270, and for fp32<->bf16 I/O casts at the module boundary). */
int b200ir_copy_cast(const void* x, int64_t x_bstride, int x_dtype, void* y, int64_t y_bstride, int y_dtype,
                     int B, int64_t CHW, b200ir_stream_t stream);

/* Depthwise K x K conv, stride 1, zero padding K/2, K in {3,5,7}: the kv_dwconv (7x7) of MoCE-IR's CrossAttention
 * (moce_ir.py:341-342).  w fp32 [C,1,K,K], bias fp32 [C] or NULL; x/y share `dtype`. */
int b200ir_dwconv_kxk_fwd(const void* x, int64_t x_bstride, const float* w, const float* bias, void* y,
                          int64_t y_bstride, int dtype, int B, int C, int H, int W, int K, b200ir_stream_t stream);
/* Its backward: input gradient (the same conv with the taps read in reverse order) and the weight/bias gradient as
 * fixed-order partial sums parts fp32 [nparts][C][K*K+1] (last entry = bias gradient), folded by
 * b200ir_reduce_partials; nparts from b200ir_dwconv_kxk_wgrad_nparts. */
int b200ir_dwconv_kxk_bwd_input(const void* dy, int64_t dy_bstride, const float* w, void* dx, int64_t dx_bstride,
                                int dtype, int B, int C, int H, int W, int K, b200ir_stream_t stream);
int b200ir_dwconv_kxk_wgrad_nparts(int B, int H, int W);
int b200ir_dwconv_kxk_wgrad(const void* dy, int64_t dy_bstride, const void* x, int64_t x_bstride, int dtype,
                            float* parts, int nparts, int B, int C, int H, int W, int K, b200ir_stream_t stream);

/* Multiplicative sigmoid gate of the gated MDTA variants (GatedMDTA.forward of experiment/
 * baseline_restormer_new_train/20251116_use_headwise_gate_model/.../moce_ir.py:155-166): y[b,c,p] = u[b,c,p] *
 * sigmoid(g[b, c / (C/G), p]).  G = C: element-wise gate, G = heads: head-wise gate.  u, g, y (and dy, du, dg) share
 * `dtype`.  Backward: du = dy * sigmoid(g), dg = sigmoid'(g) * sum over the gated channels of dy * u. */
int b200ir_sigmoid_gate_fwd(const void* u, int64_t u_bstride, const void* g, int64_t g_bstride, int dtype, void* y,
                            int64_t y_bstride, int B, int C, int G, int64_t HW, b200ir_stream_t stream);
int b200ir_sigmoid_gate_bwd(const void* u, int64_t u_bstride, const void* g, int64_t g_bstride, const void* dy,
                            int64_t dy_bstride, int dtype, void* du, int64_t du_bstride, void* dg, int64_t dg_bstride,
                            int B, int C, int G, int64_t HW, b200ir_stream_t stream);

/* ---- tcgen05 / TMEM / TMA path (bf16 activations) -------------------------------------------------
 * Same math as b200ir_ln_conv1x1_fwd, as a persistent tensor-core GEMM (csrc/gemm_tc.cu).  Weights are
 * first packed once per parameter update:
 *   wp[b][nrows][kpad] bf16 = w[b][o][c] * ln_w[c] (zero padded),
 *   st[0][o] = sum_c bf16(w*ln_w)[o,c],  st[1][o] = bias[o] + sum_c w[o,c]*ln_b[c]     (fp32 [2,nrows])
 * so that W.LN(x) = rstd*(wp.x - mu*st0) + st1 is applied in the GEMM epilogue.
 * b200ir_gemm_tc_geometry returns 0 and the padded sizes if (Cin,Cout) is supported.
 * x must be bf16 with HW%8==0, HW>=64; Bw = 1 (shared weight) or B (per-sample weight, e.g. the folded
 * MDTA matrix of b200ir_mdta_softmax_fold_fwd). */
int b200ir_gemm_tc_geometry(int Cin, int Cout, int* kpad, int* nrows);
int b200ir_pack_conv1x1(const float* w, int64_t w_bstride, int Bw, const float* bias, int ln_type,
                        const float* ln_w, const float* ln_b, int Cin, int Cout, void* wp, float* st,
                        b200ir_stream_t stream);
/* Operand of the INPUT-gradient GEMM dx = W^T.dy of a 1x1 conv (autograd of nn.Conv2d(Cin,Cout,1), Restormer.py:80,86,
 * 103,107): packs w^T straight from w fp32 [Cout][Cin] (tiled shared-memory transpose), i.e. the result equals
 * b200ir_pack_conv1x1(w^T, ln NONE, no bias) for a GEMM with Cin' = Cout, Cout' = Cin; sizes from
 * b200ir_gemm_tc_geometry(Cout, Cin). */
int b200ir_pack_conv1x1_t(const float* w, int Cin, int Cout, void* wp, float* st, b200ir_stream_t stream);
/* All weight packs of a model in ONE launch (a training step repacks every 1x1 weight after the optimizer update: ~260
 * tiny launches otherwise).  descs: DEVICE array of ndesc descriptors; desc i owns CTAs [first_block, next first_block)
 * of the launch (total_blocks CTAs of 256 threads): transposed = 0 -> ceil(Nrows/8) CTAs, same result as
 * b200ir_pack_conv1x1 (Bw = 1); transposed = 1 -> ceil(Kpad/32)*ceil(Nrows/32) CTAs, same result as
 * b200ir_pack_conv1x1_t.  Kpad / Nrows from b200ir_gemm_tc_geometry of the GEMM the pack feeds. */
typedef struct {
  const float* w;      /* [Cout][Cin] */
  const float* bias;   /* [Cout] or NULL (transposed: ignored) */
  const float* ln_w;   /* [Cin] or NULL */
  const float* ln_b;   /* [Cin] or NULL */
  void* wp;            /* bf16 [Nrows][Kpad] */
  float* st;           /* fp32 [2][Nrows] */
  int Cin, Cout, Kpad, Nrows, ln_type, transposed, first_block, reserved;
} b200ir_pack_desc_t;
int b200ir_pack_conv1x1_multi(const b200ir_pack_desc_t* descs, int ndesc, int total_blocks, b200ir_stream_t stream);
int b200ir_gemm_tc_fwd(const void* x, int64_t x_bstride, const void* wp, int Bw, const float* st, int ln_type,
                       const void* res, int64_t res_bstride, int res_dtype, void* y, int64_t y_bstride, int y_dtype,
                       int B, int Cin, int Cout, int64_t HW, b200ir_stream_t stream);
/* Same call with the kernel variant chosen by the caller instead of by shape (tests, A/B timing): AUTO = what
 * b200ir_gemm_tc_fwd picks; RESIDENT = weights resident in shared memory (gemm_tc.cu); STREAM = both operands streamed
 * (gemm_stream_tc.cu, Cin >= 64); RESIDENT_CH = resident weights with the channels-on-TMEM-lanes epilogue.  The
 * library reads no environment variable: every switch is an argument. */
enum { B200IR_GEMM_AUTO = 0, B200IR_GEMM_RESIDENT = 1, B200IR_GEMM_STREAM = 2, B200IR_GEMM_RESIDENT_CH = 3 };
int b200ir_gemm_tc_fwd_ex(const void* x, int64_t x_bstride, const void* wp, int Bw, const float* st, int ln_type,
                          const void* res, int64_t res_bstride, int res_dtype, void* y, int64_t y_bstride, int y_dtype,
                          int B, int Cin, int Cout, int64_t HW, int variant, b200ir_stream_t stream);

/* LayerNorm + 1x1 conv + depthwise 3x3 (+ GELU gate) as ONE kernel (fused_ln_proj_dw.cu): the widest tensor of the
 * block never reaches HBM.  gate = 1: GDFN, out[c] = gelu(dw(t)[c]) * dw(t)[c+Cout] with t = project_in(norm2(x)), Cout =
 * hidden_features, conv weights w [2*Cout][C], dw_w [2*Cout][9] (Restormer.py:88-91).  gate = 0: MDTA front end,
 * out = qkv_dwconv(qkv(norm1(x))), Cout = 3*C (Restormer.py:114).  x, y bf16 NCHW.  wf = b200ir_pack_fused_proj_dw(...)
 * (bf16 [nrows][kf], sizes from b200ir_fused_proj_dw_geometry): LayerNorm gamma folded into the weights, bias + W.beta
 * carried by two constant-one K rows.  b200ir_ln_proj_dw_supported: 1 if the shape fits (C <= 240, shared memory). */
int b200ir_fused_proj_dw_geometry(int C, int Cout, int gate, int* kf, int* nrows);
int b200ir_pack_fused_proj_dw(const float* w, const float* bias, int ln_type, const float* ln_w, const float* ln_b,
                              int C, int Cout, int gate, void* wf, b200ir_stream_t stream);
int b200ir_ln_proj_dw_supported(int C, int Cout, int H, int W, int gate);
int b200ir_ln_proj_dw_fwd(const void* x, int64_t x_bstride, const void* wf, const float* dw_w, const float* dw_b,
                          int ln_type, void* y, int64_t y_bstride, int B, int C, int Cout, int H, int W, int gate,
                          b200ir_stream_t stream);

/* Dense 3x3 conv (pad 1, +bias) on the same tcgen05 kernel: 9 tap-shifted GEMMs accumulated in TMEM, zero padding
 * from TMA out-of-bounds fill, PixelUnshuffle(2)/PixelShuffle(2) folded into the store (Downsample / Upsample,
 * Restormer.py:171-189).  x bf16 NHWC with Cin%8==0; wp/st from b200ir_pack_conv3x3 (w fp32 [Cout,Cin,3,3]). */
int b200ir_conv3x3_tc_geometry(int Cin, int Cout, int* kpad, int* nrows);
/* bf16 NCHW -> NHWC staging copy: b200ir_conv3x3_tc_fwd takes its activations as [B][H][W][Cin] (channels
 * contiguous, x_bstride = H*W*Cin) so that tap shifts only touch non-innermost TMA coordinates. */
int b200ir_nchw_to_nhwc(const void* x, int64_t x_bstride, void* y, int B, int C, int64_t HW, b200ir_stream_t stream);
int b200ir_pack_conv3x3(const float* w, const float* bias, int Cin, int Cout, void* wp, float* st,
                        b200ir_stream_t stream);
int b200ir_conv3x3_tc_fwd(const void* x, int64_t x_bstride, const void* wp, const float* st, void* y,
                          int64_t y_bstride, int y_dtype, int B, int Cin, int Cout, int H, int W, int shuffle_mode,
                          b200ir_stream_t stream);

/* ---- backward / training step (autograd backward of the ops above, nn.L1Loss, AdamW) -------------
 * Input gradients of the 1x1 / depthwise / dense 3x3 convs reuse the forward entry points with transposed / flipped
 * weights.  The entry points below are the reductions over pixels and the non-linear pieces.  All reductions write
 * fixed-order partial sums that b200ir_reduce_partials folds: results are bitwise reproducible. */

/* Input gradient of the depthwise 3x3 conv (autograd of nn.Conv2d(C,C,3,padding=1,groups=C), Restormer.py:84,106):
 * dx = depthwise conv of dy with the taps of w [C,1,3,3] read in reverse order (no flipped copy of w needed). */
int b200ir_dwconv3x3_bwd_input(const void* dy, int64_t dy_bstride, const float* w, void* dx, int64_t dx_bstride,
                               int dtype, int B, int C, int H, int W, b200ir_stream_t stream);
/* y = LN(x) materialised (operand of the weight gradient of the fused LN + 1x1 conv; Restormer.py:37-39,54-57). */
int b200ir_ln_fwd(const void* x, int64_t x_bstride, int x_dtype, int ln_type, const float* ln_w, const float* ln_b,
                  void* y, int64_t y_bstride, int y_dtype, int B, int C, int64_t HW, b200ir_stream_t stream);
/* LayerNorm backward: dx = res + dLN(x)^T g;  parts fp32 [nparts][2][C] = partial (dweight | dbias), nparts from
 * b200ir_ln_bwd_nparts.  res (optional) = gradient arriving through the residual add (Restormer.py:147-148).
 * xn (optional, dtype/layout of dx, needs ln_b for WithBias) receives LN(x) as a by-product: the operand of the
 * following weight-gradient GEMM, so that no separate b200ir_ln_fwd pass over x is needed. */
int b200ir_ln_bwd_nparts(int B, int64_t HW);
int b200ir_ln_bwd(const void* x, int64_t x_bstride, int x_dtype, const void* g, int64_t g_bstride, int g_dtype,
                  int ln_type, const float* ln_w, const float* ln_b, const void* res, int64_t res_bstride,
                  int res_dtype, void* dx, int64_t dx_bstride, int dx_dtype, void* xn, int64_t xn_bstride,
                  float* parts, int nparts, int B, int C, int64_t HW, b200ir_stream_t stream);
/* out[k] = sum_{i<n} in[i*stride+k], fixed order. */
int b200ir_reduce_partials(const float* in, float* out, int n, int64_t K, int64_t stride, b200ir_stream_t stream);
/* The fold of depthwise-conv partial sums in[n][C][nv] (nv - 1 taps and the bias sum per channel) into the two parameter
 * gradients as dense tensors: dw fp32 [C][nv - 1], db fp32 [C] (NULL: no bias).  Same order of additions. */
int b200ir_reduce_partials_taps(const float* in, float* dw, float* db, int n, int C, int nv, b200ir_stream_t stream);
/* Weight (+bias) gradient of a 1x1 (taps=1) or dense 3x3 pad-1 (taps=9) conv = reduction over pixels:
 *   parts_w[split][g][o][c*taps+t] = partial sum_p dy[b,o,p]*x[b,c,p+shift(t)],  parts_b[split][g][o] = partial sum_p dy
 * g = sample if per_sample (MDTA: dM_b = dout_b.v_b^T, gradient of the folded per-sample weight) else the whole batch
 * (G=1).  nsplit from b200ir_conv_wgrad_nsplit.  parts_b may be NULL.  impl: B200IR_IMPL_* (TC: bf16 tcgen05). */
int b200ir_conv_wgrad_nsplit(int B, int Cout, int Cin, int taps, int64_t HW, int per_sample);
int b200ir_conv_wgrad(const void* dy, int64_t dy_bstride, int dy_dtype, const void* x, int64_t x_bstride, int x_dtype,
                      float* parts_w, float* parts_b, int nsplit, int B, int Cout, int Cin, int H, int W, int taps,
                      int per_sample, int impl, b200ir_stream_t stream);
/* y bf16 [B,Cin*9,H*W]: y[b][c*9+t][p] = x[b][c][p+shift(t)] (zero padded) — turns the dense 3x3 weight gradient into
 * b200ir_conv_wgrad(taps=1) over Cin*9 channels so that it runs on the tcgen05 kernel. */
int b200ir_im2col3x3(const void* x, int64_t x_bstride, int x_dtype, void* y, int B, int Cin, int H, int W,
                     b200ir_stream_t stream);
/* Depthwise 3x3 weight/bias gradient (csrc/dw_backward.cu): parts fp32 [nparts][C][10] (9 taps + bias) with
 * nparts = b200ir_dwconv3x3_bwd_nparts(B, C, H, W); fold with b200ir_reduce_partials. */
int b200ir_dwconv3x3_bwd_nparts(int B, int C, int H, int W);
int b200ir_dwconv3x3_wgrad(const void* dy, int64_t dy_bstride, const void* x, int64_t x_bstride, int dtype,
                           float* parts, int nparts, int B, int C, int H, int W, b200ir_stream_t stream);
/* Backward of b200ir_dwconv3x3_gelu_gate_fwd w.r.t. the depthwise-conv OUTPUT (recomputed from t) fused with the
 * depthwise weight/bias gradient: da [B,2Hd,H,W] from t [B,2Hd,H,W] and dg [B,Hd,H,W]; parts fp32 [nparts][2Hd][10]
 * with nparts = b200ir_dwconv3x3_bwd_nparts(B, Hd, H, W).  dt then follows from b200ir_dwconv3x3_fwd (flipped w). */
int b200ir_dwconv3x3_gelu_gate_bwd(const void* t, int64_t t_bstride, const float* w, const float* bias, const void* dg,
                                   int64_t dg_bstride, void* da, int64_t da_bstride, int dtype, float* parts,
                                   int nparts, int B, int Hd, int H, int W, b200ir_stream_t stream);
/* Same for the split-gate FFN of experiment/RectiFormer.py:231-266, y = gelu(a1) * sigmoid(a2):
 * da[c] = dg*sigmoid(a2)*gelu'(a1), da[c+Hd] = dg*gelu(a1)*sigmoid'(a2). */
int b200ir_dwconv3x3_gelu_sigmoid_gate_bwd(const void* t, int64_t t_bstride, const float* w, const float* bias, const void* dg,
                                   int64_t dg_bstride, void* da, int64_t da_bstride, int dtype, float* parts,
                                   int nparts, int B, int Hd, int H, int W, b200ir_stream_t stream);
/* MDTA backward, small per-(sample, head) part (softmax, temperature, F.normalize; Restormer.py:121-131).
 * In: Gram partials + attn of the forward, dM [B,C,C] (= conv_wgrad(dout, v, per_sample)).
 * Out: mt [B,C,C] = M_b^T (dv = mt_b.dout), mqk [B,2C,2C] (d[q;k] = mqk_b.[q;k]), dw_out_part [B,C,C] and
 * dtemp_part [B,heads] (sum over B = dW_out, dtemperature). */
int b200ir_mdta_bwd(const float* partials, int nslab, const float* attn, const float* temperature, const float* w_out,
                    const float* dM, float* mt, float* mqk, float* dw_out_part, float* dtemp_part, int B, int C,
                    int heads, b200ir_stream_t stream);
/* Same with a per-sample temperature [B][heads] (temp_per_sample = 1): the dynamic temperature of
 * experiment/RectiFormer.py:309-311; dtemp_part [B][heads] is then the gradient of that effective temperature. */
int b200ir_mdta_bwd_ex(const float* partials, int nslab, const float* attn, const float* temperature, int temp_per_sample,
                       const float* w_out, const float* dM, float* mt, float* mqk, float* dw_out_part, float* dtemp_part,
                       int B, int C, int heads, b200ir_stream_t stream);
/* The two halves of b200ir_mdta_bwd_ex separately (they only share inputs, so a training step runs them on two streams):
 * CHAIN = softmax / temperature / normalize backward -> mqk, dtemp_part; FOLD = mt (the dv operand) and dw_out_part.
 * Outputs of a half that is not selected may be NULL. */
enum { B200IR_MDTA_BWD_CHAIN = 1, B200IR_MDTA_BWD_FOLD = 2 };
int b200ir_mdta_bwd_parts(const float* partials, int nslab, const float* attn, const float* temperature,
                          int temp_per_sample, const float* w_out, const float* dM, float* mt, float* mqk,
                          float* dw_out_part, float* dtemp_part, int B, int C, int heads, int parts,
                          b200ir_stream_t stream);
/* nn.PixelShuffle(2) / nn.PixelUnshuffle(2) as a copy (x [B,C,H,W]; gradients of Upsample / Downsample). */
int b200ir_pixel_shuffle(const void* x, int64_t x_bstride, void* y, int64_t y_bstride, int dtype, int B, int C, int H,
                         int W, int mode, b200ir_stream_t stream);
/* nn.L1Loss (mean) forward / backward (MoCE-IR-main/src/train.py:51,65).  parts: fp32 [nparts] scratch. */
int b200ir_l1_loss_fwd(const void* pred, int pred_dtype, const void* target, int target_dtype, int64_t n, float* parts,
                       int nparts, float* loss, b200ir_stream_t stream);
/* Allocation-free variant of b200ir_l1_loss_fwd.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops).  The per-block partial
 * sums go to a 16 KB file-scope __device__ array inside the library (held by every GPU that loads it, until exit) instead
 * of a caller buffer; the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).  Limit:
 * n <= B200IR_L1_NOALLOC_MAX_N elements (16 Mi: e.g. 16 x 3 x 512 x 512), else B200IR_E_WORKSPACE with a message in
 * b200ir_last_error_string().  It uses nparts = ceil(n / 4096) blocks, and is bitwise equal to b200ir_l1_loss_fwd called with that
 * nparts.  All calls on a device share the array: calls that overlap on the GPU (different streams) race; calls on one
 * stream are safe, and nothing needs clearing between them (each call rewrites what it reads). */
#define B200IR_L1_NOALLOC_MAX_PARTS 4096
#define B200IR_L1_NOALLOC_ELEMS_PER_PART 4096
#define B200IR_L1_NOALLOC_MAX_N ((int64_t)B200IR_L1_NOALLOC_MAX_PARTS * B200IR_L1_NOALLOC_ELEMS_PER_PART)
int b200ir_l1_loss_fwd_noalloc(const void* pred, int pred_dtype, const void* target, int target_dtype, int64_t n,
                               float* loss, b200ir_stream_t stream);
int b200ir_l1_loss_bwd(const void* pred, int pred_dtype, const void* target, int target_dtype, int64_t n,
                       const float* gscale, void* grad, int grad_dtype, b200ir_stream_t stream);
/* torch.optim.AdamW step on one flat fp32 buffer (MoCE-IR-main/src/train.py:83); g is scaled by grad_scale first
 * (1/world_size after a sum all-reduce).  The step count (bias correction) is either the host value `step` (>= 1,
 * step_dev NULL) or a device counter `step_dev` that this call increments on the stream before the update — the form
 * that can be captured in a CUDA graph and replayed. */
int b200ir_adamw_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                      float eps, float weight_decay, int step, int* step_dev, float grad_scale,
                      b200ir_stream_t stream);
/* Same update over a sub-range of the flat buffers.  tick = 0 leaves step_dev as it is (several ranges share one
 * optimizer step: parameters whose gradient is None are skipped the way torch.optim.AdamW skips them, so the buffer is
 * updated as a few contiguous runs).  hyper_dev (optional, device float[2] = {lr, weight_decay}) overrides the host
 * scalars: an LR scheduler (LinearWarmupCosineAnnealingLR in MoCE-IR-main/src/train.py:84-86) changes the learning rate
 * between replays of a captured CUDA graph by writing that buffer. */
int b200ir_adamw_step_ex(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                         float eps, float weight_decay, int step, int* step_dev, int tick, const float* hyper_dev,
                         float grad_scale, b200ir_stream_t stream);

/* ---- MoCE-IR decoder pieces (moce_ops.cu) ----
 * Per p x p patch and channel: mode 0 y = circular convolution of a and b = irfft2(rfft2(a) * rfft2(b)) of FFTAttention
 * (moce_ir.py:406-412; planes are zero-padded to a multiple of p like F.pad at :391 and cropped back); mode 1
 * y[s] = sum_i a[i] b[(i - s) mod p] (the two gradients: dq = corr(dout, k), dk = corr(dout, q)).  p in {2,4,8,16,32}. */
int b200ir_circconv_patch(const void* a, int64_t a_bstride, const void* b, int64_t b_bstride, void* y, int64_t y_bstride,
                          int dtype, int B, int C, int H, int W, int patch, int mode, b200ir_stream_t stream);
/* Elementwise over [B][per_sample] views: op 0 a*b, 1 a*silu(b) (ModExpert.process, moce_ir.py:556), 2 a+b, 3 gelu(a)
 * (b NULL), 5 b * dsilu(a). */
int b200ir_ew(const void* a, int64_t a_bstride, const void* b, int64_t b_bstride, void* y, int64_t y_bstride, int dtype,
              int B, int64_t per_sample, int op, b200ir_stream_t stream);
/* y[b] (+)= g[b] * x[b]: per-sample gate of an expert output (AdapterLayer.forward, moce_ir.py:672-676). */
int b200ir_scale_rows(const void* x, int64_t x_bstride, const float* g, void* y, int64_t y_bstride, int dtype, int B,
                      int64_t per_sample, int accumulate, b200ir_stream_t stream);
/* out[b][c] = mean_hw x[b,c,:] (gelu = 1: mean of gelu(x)): AdaptiveAvgPool2d(1) of the router (moce_ir.py:703-707) and
 * x.mean(dim=(-2,-1)) of FrequencyEmbedding (:1072).  out fp32 [B][C]. */
int b200ir_spatial_mean(const void* x, int64_t x_bstride, int dtype, float* out, int B, int C, int64_t HW, int gelu,
                        b200ir_stream_t stream);
/* Backward of b200ir_spatial_mean: dx[b,c,:] = g[b,c] / HW (times gelu'(x) with gelu = 1).  g fp32 [B][C]. */
int b200ir_spatial_mean_bwd(const float* g, const void* x, int64_t x_bstride, void* dx, int64_t dx_bstride, int dtype,
                            int B, int C, int64_t HW, int gelu, b200ir_stream_t stream);
/* Training mode of the MoCE-IR adapter, SparseDispatcher.dispatch / combine (moce_ir.py:94-123): rows = samples of
 * `per` elements.  y[dst(j)] (+)= scale[j] * x[src(j)], j < n; src_idx / dst_idx int32 device arrays or NULL (identity),
 * scale fp32 device array or NULL (1).  dispatch: `inp[self._batch_index]` = gather by src_idx; combine:
 * `zeros.index_add(0, batch_index, stitched * gates)` = scatter-accumulate by dst_idx with the gates as scale.  The indices
 * of one call must be distinct. */
int b200ir_rows_copy(const void* x, int64_t x_rowstride, void* y, int64_t y_rowstride, const int* src_idx,
                     const int* dst_idx, const float* scale, int n, int64_t per, int accumulate, int dtype,
                     b200ir_stream_t stream);
/* out[j] = sum_i a[a_idx(j), i] * b[b_idx(j), i] (fp32, fixed order): gradient of the gates of the weighted combine
 * (autograd of `stitched.mul(self._nonzero_gates)`, moce_ir.py:117). */
int b200ir_rowdot(const void* a, int64_t a_rowstride, const int* a_idx, const void* b, int64_t b_rowstride,
                  const int* b_idx, float* out, int n, int64_t per, int dtype, b200ir_stream_t stream);

/* ---- AdaIR frequency mining / modulation pieces (adair_ops.cu; AdaIR-main/net/model.py:234-372) ----
 * lowfreq_split: FreModule.fft.  hw_counts int32 [B][2] = (h_, w_) of the centred spectrum mask per sample; low =
 * |ifft2(fft * mask)|, high = |ifft2(fft * (1 - mask))|, the <= max_modes kept bins synthesised directly (no FFT).
 * coef_ws: fp32 scratch [B][C][64][2]. */
int b200ir_lowfreq_split(const void* x, int64_t x_bstride, const int* hw_counts, float* coef_ws, void* high,
                         int64_t h_bstride, void* low, int64_t l_bstride, int dtype, int B, int C, int H, int W,
                         int max_modes, b200ir_stream_t stream);
/* F.interpolate(mode='bilinear', align_corners=False) (model.py:313). */
int b200ir_bilinear_resize(const void* x, int64_t x_bstride, void* y, int64_t y_bstride, int dtype, int B, int C, int Hin,
                           int Win, int Hout, int Wout, b200ir_stream_t stream);
/* SpatialGate (model.py:234-246): gate fp32 [B][H*W] = sigmoid(conv7x7([max_c x, mean_c x])); mm_ws fp32 [B][2][H*W]. */
int b200ir_spatial_gate(const void* x, int64_t x_bstride, int dtype, const float* w7, float* mm_ws, float* gate, int B,
                        int C, int H, int W, b200ir_stream_t stream);
/* AdaptiveMaxPool2d(1) (ChannelGate, model.py:253): out fp32 [B][C]. */
int b200ir_spatial_max(const void* x, int64_t x_bstride, int dtype, float* out, int B, int C, int64_t HW,
                       b200ir_stream_t stream);
/* FreRefine gating (model.py:279-284): y = low * sw[b][p] + high * cw[b][c]. */
int b200ir_fre_combine(const void* low, int64_t l_bstride, const void* high, int64_t h_bstride, const float* sw,
                       const float* cw, void* y, int64_t y_bstride, int dtype, int B, int C, int64_t HW,
                       b200ir_stream_t stream);
/* y = a * p1[c] + b * p2[c] (model.py:324). */
int b200ir_affine2(const void* a, int64_t a_bstride, const void* b, int64_t b_bstride, const float* p1, const float* p2,
                   void* y, int64_t y_bstride, int dtype, int B, int C, int64_t HW, b200ir_stream_t stream);

/* ---- backward of the AdaIR pieces (adair_bwd.cu): what torch autograd does on AdaIR-main/net/model.py:234-372 in
 * AdaIR-main/train.py.  Every reduction is in a fixed order (bitwise deterministic). ----
 * lowfreq_split_bwd: low = |P x|, high = |x - P x| (P = ifft2 . mask . fft2, Hermitian), so
 * dx = Re(u_h) + Re(P (u_l - u_h)), u = g * z / |z| (0 where |z| = 0, torch's sgn).  coef = the forward's coef_ws;
 * coefw_ws fp32 scratch [B][C][64][2].  Nothing flows to the thresholds (they pass through .int(), model.py:345-346). */
int b200ir_lowfreq_split_bwd(const void* x, int64_t x_bstride, const int* hw_counts, const float* coef, const void* g_high,
                             int64_t gh_bstride, const void* g_low, int64_t gl_bstride, float* coefw_ws, void* dx,
                             int64_t dx_bstride, int dtype, int B, int C, int H, int W, int max_modes,
                             b200ir_stream_t stream);
/* SpatialGate backward: mm / gate as written by b200ir_spatial_gate, dgate fp32 [B][H*W], ds_ws fp32 scratch [B][H*W];
 * dw7 fp32 [2][7][7]; dx (+)= dmean / C + [c == first argmax_c x] * dmax (accumulate != 0 adds to dx). */
int b200ir_spatial_gate_bwd(const void* x, int64_t x_bstride, int dtype, const float* w7, const float* mm,
                            const float* gate, const float* dgate, float* ds_ws, void* dx, int64_t dx_bstride, float* dw7,
                            int B, int C, int H, int W, int accumulate, b200ir_stream_t stream);
/* AdaptiveMaxPool2d(1) backward: dx[b][c][:] = 0 except g[b][c] at the first maximum of the plane. */
int b200ir_spatial_max_bwd(const void* x, int64_t x_bstride, const float* g, void* dx, int64_t dx_bstride, int dtype, int B,
                           int C, int64_t HW, b200ir_stream_t stream);
/* fre_combine backward: dlow = g sw, dhigh = g cw, dsw[b][p] = sum_c g low, dcw[b][c] = sum_p g high (fp32). */
int b200ir_fre_combine_bwd(const void* g, int64_t g_bstride, const void* low, int64_t l_bstride, const void* high,
                           int64_t h_bstride, const float* sw, const float* cw, void* dlow, int64_t dl_bstride, void* dhigh,
                           int64_t dh_bstride, float* dsw, float* dcw, int dtype, int B, int C, int64_t HW,
                           b200ir_stream_t stream);
/* affine2 backward: da = g p1[c], db = g p2[c], parts fp32 [B][C][2] = per-sample (sum_p g a, sum_p g b): the caller
 * sums them over B for dp1 / dp2. */
int b200ir_affine2_bwd(const void* g, int64_t g_bstride, const void* a, int64_t a_bstride, const void* b, int64_t b_bstride,
                       const float* p1, const float* p2, void* da, int64_t da_bstride, void* db, int64_t db_bstride,
                       float* parts, int dtype, int B, int C, int64_t HW, b200ir_stream_t stream);

/* ---- input pipeline and evaluation metrics on the device (pipeline_ops.cu) ----
 * Random-crop + flip/rot90 augmentation (mode 0..7 of MoCE-IR-main/src/utils/image_utils.py:136-166) + Gaussian-noise
 * synthesis clip(clean + noise*sigma, 0, 255).astype(uint8) (degradation_utils.py:21-24) + ToTensor.  src uint8
 * [B][Hs][Ws][3]; offsets int32 [B][2] (row, column of the crop); modes int32 [B]; sigma fp32 [B] (on the 0..255 scale,
 * NULL / 0 = no noise); noise fp32 [B][P][P][3] standard-normal samples (caller's generator); clean / degraded fp32
 * [B][3][P][P] in [0,1] (degraded may be NULL). */
int b200ir_crop_augment_degrade(const uint8_t* src, int Hs, int Ws, const int* offsets, const int* modes,
                                const float* sigma, const float* noise, float* clean, float* degraded, int B, int P,
                                b200ir_stream_t stream);
/* Full-frame inference in tiles (the E / W accumulate of the reference's tile helper, 20260104_CG_IR/src/data/
 * dataset_utils.py:513-534, MEASNet-main/MEASNet/utils/dataset_utils.py:226-247).  tile_gather: patches fp32
 * [ntiles][B][C][tile][tile] cut from src fp32 [B][C][H][W] at origins int32 [ntiles][2] (row, column).  tile_blend: out =
 * (sum of the tiles covering a pixel, in (h_idx, w_idx) order) / (their number), optionally clamped to [0, 1]; patches are
 * ordered h_idx-major ([nh * nw][B][C][tile][tile]).  No atomics: bitwise the reference's accumulate order. */
int b200ir_tile_gather(const float* src, const int* origins, float* patches, int ntiles, int B, int C, int H, int W, int tile,
                       b200ir_stream_t stream);
int b200ir_tile_blend(const float* patches, const int* h_idx, int nh, const int* w_idx, int nw, float* out, int B, int C,
                      int H, int W, int tile, int clamp01, b200ir_stream_t stream);
/* out fp32 [B][2] = (PSNR, SSIM) of clamp(restored,0,1) vs clean, fp32 NCHW, skimage semantics (test.py:46-52):
 * data_range 1, SSIM with an 11x11 Gaussian window (sigma 1.5), sample covariance, mean without a 5-pixel border, then
 * over channels.  ws: fp32 scratch [B * (1 + C) * b200ir_psnr_ssim_nblocks(H, W)]. */
int b200ir_psnr_ssim_nblocks(int H, int W);
int b200ir_psnr_ssim(const float* restored, const float* clean, float* ws, float* out, int B, int C, int H, int W,
                     b200ir_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif
