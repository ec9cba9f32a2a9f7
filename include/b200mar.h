/* b200mar.h — C ABI of the B200-native NGswin hot path (libb200mar.so).
 *
 * Drop-in boundary for the N-gram window-attention transformer blocks of
 * mpat247/TransMAR-GAN's generator (models/generator/ngswin.py::NGswin).  The reference is
 * pure Python/PyTorch and has no FFI of its own; each entry point below replaces the
 * reference *module method* cited beside it, and is what a ctypes / torch.library binding in
 * the reference tree would call (see INTEGRATION.md for the stubs).
 *
 * Conventions (all entry points):
 *   - every pointer is a DEVICE pointer to a contiguous, 16-byte aligned buffer owned by the
 *     caller (no ownership transfer, nothing is allocated or freed here);
 *   - `stream` is a cudaStream_t passed as void*; kernels are enqueued on it and the call
 *     returns without synchronising (safe under CUDA-graph capture);
 *   - return value: 0 on success, otherwise a cudaError_t (or -1 for an argument the path
 *     does not support, e.g. a head count other than 4 or 6);
 *   - token tensors are fp32 `[B, ph*pw, 64]`, row-major over (row, col), channels last
 *     (models/generator/ngswin.py:172); parameters are in the reference state_dict layout
 *     (nn.Linear weight `[out, in]`, conv weight `[out, in/groups, kh, kw]`);
 *   - embed dim 64, window 8x8, MLP hidden 128, n-gram 2 are compile-time constants of the
 *     reference's only configuration (models/generator/ngswin.py:43-49);
 *   - `*_bwd` entry points ACCUMULATE (+=) into parameter-gradient buffers, so one zeroed
 *     flat gradient bucket can be shared by all blocks; activation gradients are overwritten
 *     unless stated otherwise.
 *   - no CPU implementation exists behind this ABI: on a machine without an sm_100 GPU every
 *     compute entry point returns cudaErrorNoKernelImageForDevice / cudaErrorNoDevice.
 */
#ifndef B200MAR_H
#define B200MAR_H

#ifdef __cplusplus
extern "C" {
#endif

/* ABI version, bumped whenever a signature changes. */
int b200mar_abi_version(void);   /* 7 */
/* Name of the last failing CUDA call of this thread (static string; never NULL). */
const char* b200mar_last_error(void);
/* 1 iff device `dev` is compute capability 10.x (B200); no kernels are launched. */
int b200mar_device_ok(int dev);

/* ---- a3: NGramContext.unigram_embed  (components/window_utils.py:36-38,78) --------------
 * u[b,h,w,g] = bias[g] + sum_{k<2,i<8,j<8} weight[g,k,i,j] * z[b, 8h+i, 8w+j, 2g+k]
 * z [B,ph,pw,64] -> u [B,wh,ww,32] (channels last).                                        */
int b200mar_unigram_fwd(const float* z, const float* weight /*[32,2,8,8]*/,
                        const float* bias /*[32]*/, float* u,
                        int B, int ph, int pw, void* stream);
/* dz += scatter(du * weight)  (dz must already hold the other contributions);
 * dweight += , dbias += .                                                                  */
int b200mar_unigram_bwd(const float* z, const float* weight, const float* du,
                        float* dz, float* dweight, float* dbias,
                        int B, int ph, int pw, void* stream);
/* Same, with one more addend: dz = dz + dz_add + scatter(du * weight).  dz_add (may be null) is the
 * gradient of the within-stage sum z_{i+1} = z_i + out_i (components/encoder_decoder.py:138-141) that
 * autograd would otherwise add with a separate elementwise pass.                              */
int b200mar_unigram_bwd_add(const float* z, const float* weight, const float* du, float* dz,
                            const float* dz_add, float* dweight, float* dbias, int B, int ph, int pw,
                            void* stream);

/* ---- a3: NGramContext.forward after the unigram conv (components/window_utils.py:43-90) --
 * seq_refl_win_pad (fwd/bwd), 2x2 sliding WindowAttention(dim 32, nh heads, window 2),
 * AvgPool2d(2), concat, 1x1 merge conv.  u [B,wh,ww,32] -> ctx [B,wh,ww,64].
 * nh in {4,6}; head_dim = 32/nh (8 or 5).                                                  */
int b200mar_ngram_ctx_fwd(const float* u,
                          const float* qkv_w /*[3*na,32]*/, const float* qkv_b /*[3*na]*/,
                          const float* proj_w /*[32,na]*/, const float* proj_b /*[32]*/,
                          const float* logit_scale /*[nh]*/, const float* bias_table /*[9,nh]*/,
                          const float* merge_w /*[64,64]*/, const float* merge_b /*[64]*/,
                          float* ctx, int B, int wh, int ww, int nh, void* stream);
/* Inference pair: the unigram pass also applies the n-gram attention's qkv Linear once per window
 * (nq [B*wh*ww, 96], rows >= 3*NA zero); b200mar_ngram_ctx_pre_fwd gathers those rows instead of
 * re-deriving them for every neighbourhood.  Same results as unigram_fwd + ngram_ctx_fwd.       */
int b200mar_unigram_qkv_fwd(const float* z, const float* weight, const float* bias,
                            const float* qkv_w, const float* qkv_b, float* nq,
                            int B, int ph, int pw, int nh, void* stream);
/* Training form of b200mar_unigram_qkv_fwd (components/window_utils.py:36-38,75-78): the same pass also writes the
 * unigram embedding u [B, ph/8, pw/8, 32] that b200mar_ngram_ctx_bwd / b200mar_unigram_bwd consume.            */
int b200mar_unigram_qkv_u_fwd(const float* z, const float* weight, const float* bias,
                              const float* qkv_w, const float* qkv_b, float* nq, float* u,
                              int B, int ph, int pw, int nh, void* stream);
int b200mar_ngram_ctx_pre_fwd(const float* nq, const float* proj_w, const float* proj_b,
                              const float* logit_scale, const float* bias_table,
                              const float* merge_w, const float* merge_b, float* ctx,
                              int B, int wh, int ww, int nh, void* stream);
/* NO-ALLOCATION variant of the inference pair above (b200mar_unigram_qkv_fwd + b200mar_ngram_ctx_pre_fwd in one
 * call).  It has NO COUNTERPART IN THE REFERENCE and was added for callers
 * that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it): the nq scratch
 * between the two kernels lives in a file-scope `static __device__` array of the library (6 MiB per GPU, held
 * until the process exits) instead of a caller-allocated buffer; the call performs no cudaMalloc / cudaFree and
 * calls no allocating routine.  Limit: B*(ph/8)*(pw/8) <= b200mar_ngram_ctx_infer_noalloc_max_windows() = 16384
 * windows (one 1024x1024 slice, or six 416x416 slices, at stage 1); larger calls return -5 with a message in
 * b200mar_last_error().  All calls share the array, which keeps the last call's contents (nothing needs to start
 * at zero: every element read is written first); calls that OVERLAP on the GPU (different streams) race on it -
 * serialise them.  ctx [B, ph/8, pw/8, 64] is bit-identical to the two-call form.                              */
int b200mar_ngram_ctx_infer_noalloc(const float* z, const float* uni_w, const float* uni_b,
                                    const float* qkv_w, const float* qkv_b,
                                    const float* proj_w, const float* proj_b,
                                    const float* logit_scale, const float* bias_table,
                                    const float* merge_w, const float* merge_b, float* ctx,
                                    int B, int ph, int pw, int nh, void* stream);
int b200mar_ngram_ctx_infer_noalloc_max_windows(void);
int b200mar_ngram_ctx_bwd(const float* u, const float* dctx,
                          const float* qkv_w, const float* qkv_b,
                          const float* proj_w, const float* proj_b,
                          const float* logit_scale, const float* bias_table,
                          const float* merge_w, const float* merge_b,
                          float* du,
                          float* d_qkv_w, float* d_qkv_b, float* d_proj_w, float* d_proj_b,
                          float* d_logit_scale, float* d_bias_table,
                          float* d_merge_w, float* d_merge_b,
                          int B, int wh, int ww, int nh, void* stream);

/* ---- a4+a5+a6+a7+a8(first half): NGramWindowPartition + WindowAttention + unpartition +
 * reverse roll + `x + norm1(.)`  (components/window_utils.py:127-147,
 * components/attention.py:70-101, components/transformer_blocks.py:63-112,117) ------------
 * For every (shifted) window: gather z+ctx with the cyclic shift folded into the address
 * math, QKV, scaled-cosine logits * exp(min(logit_scale, ln 100)) + relative-position bias
 * + analytic 9-region shift mask (-100), softmax, PV, proj, LayerNorm(norm1), residual,
 * scatter back to the source token positions:  a = z + LN1(attn(partition(z + ctx))).
 * shift in {0,4}; nh in {4,6} (head_dim 16 / 10).                                          */
int b200mar_attn_fwd(const float* z, const float* ctx,
                     const float* qkv_w /*[3A,64]*/, const float* qkv_b /*[3A]*/,
                     const float* proj_w /*[64,A]*/, const float* proj_b /*[64]*/,
                     const float* logit_scale /*[nh]*/, const float* bias_table /*[225,nh]*/,
                     const float* ln_w /*[64]*/, const float* ln_b /*[64]*/,
                     float* a, int B, int ph, int pw, int nh, int shift, void* stream);
/* Same contract as b200mar_attn_fwd, computed on the tcgen05 tensor cores (kind::tf32 operands,
 * fp32 accumulate in TMEM): two windows (128 tokens) per tile, two workers per persistent CTA;
 * q/k are normalised in fp32 before the down-conversion, softmax / LayerNorm / residual in fp32. */
int b200mar_attn_tc_fwd(const float* z, const float* ctx,
                        const float* qkv_w, const float* qkv_b,
                        const float* proj_w, const float* proj_b,
                        const float* logit_scale, const float* bias_table,
                        const float* ln_w, const float* ln_b,
                        float* a, int B, int ph, int pw, int nh, int shift, int residual,
                        void* stream);
/* Training forward: as b200mar_attn_tc_fwd(residual = 1) and, in addition, the pre-projection attention
 * output O as bf16 [ceil(B*(ph/8)*(pw/8)/2)*128][64] in window order (o16) for b200mar_attn_bwd_tc.  */
int b200mar_attn_tc_fwd_save(const float* z, const float* ctx,
                        const float* qkv_w, const float* qkv_b,
                        const float* proj_w, const float* proj_b,
                        const float* logit_scale, const float* bias_table,
                        const float* ln_w, const float* ln_b,
                        float* a, void* o16, int B, int ph, int pw, int nh, int shift,
                        void* stream);
/* Precision level "tf32c" of the per-unit plan: as b200mar_attn_tc_fwd(residual = 1), or as
 * b200mar_attn_tc_fwd_save when o16 != NULL, with the qkv projection (models/generator/components/attention.py:75)
 * on compensated FP16 hi + lo operand pairs: X = Xh + Xl, W = Wh + Wl, Xh Wh + Xl Wh + Xh Wl accumulated in fp32,
 * i.e. ~22 significant bits instead of 11; two workers per persistent CTA.                                    */
int b200mar_attn_tc_fwd_comp(const float* z, const float* ctx,
                        const float* qkv_w, const float* qkv_b,
                        const float* proj_w, const float* proj_b,
                        const float* logit_scale, const float* bias_table,
                        const float* ln_w, const float* ln_b,
                        float* a, void* o16 /* may be NULL */, int B, int ph, int pw, int nh, int shift,
                        void* stream);
/* residual = 1: a = z + LN1(attn) as above.  residual = 0: a = LN1(attn) only - the consumer
 * (b200mar_mlp_tc_fwd with a_is_delta = 1) adds z, which it reads anyway.                   */
/* dz = da + scatter(dX) (overwrites dz); dctx += per-source-window sums of dX (dctx must be
 * zeroed by the caller); parameter gradients are accumulated.                              */
int b200mar_attn_bwd(const float* z, const float* ctx, const float* da,
                     const float* qkv_w, const float* qkv_b,
                     const float* proj_w, const float* proj_b,
                     const float* logit_scale, const float* bias_table,
                     const float* ln_w, const float* ln_b,
                     float* dz, float* dctx,
                     float* d_qkv_w, float* d_qkv_b, float* d_proj_w, float* d_proj_b,
                     float* d_logit_scale, float* d_bias_table, float* d_ln_w, float* d_ln_b,
                     int B, int ph, int pw, int nh, int shift, void* stream);
/* attn_bwd on the tcgen05 tensor cores (fp16 / bf16 operands, fp32 accumulate): same contract; `scratch`
 * is a caller-owned device buffer of ceil(B*(ph/8)*(pw/8)/2) * 128 * 768 bytes (window-order bf16
 * operands of the weight gradients: X, O, dY, per-head packed d[q|k|v]).  o16_saved: the O tensor written
 * by b200mar_attn_tc_fwd_save for the same inputs (skips the recomputation pass), or NULL.      */
int b200mar_attn_bwd_tc(const float* z, const float* ctx, const float* da,
                     const float* qkv_w, const float* qkv_b,
                     const float* proj_w, const float* proj_b,
                     const float* logit_scale, const float* bias_table,
                     const float* ln_w, const float* ln_b,
                     float* dz, float* dctx,
                     float* d_qkv_w, float* d_qkv_b, float* d_proj_w, float* d_proj_b,
                     float* d_logit_scale, float* d_bias_table, float* d_ln_w, float* d_ln_b, void* scratch, const void* o16_saved,
                     int B, int ph, int pw, int nh, int shift, void* stream);

/* ---- a8(second half)+a9: `x + norm2(ffn(x))` and the within-stage residual
 * (components/transformer_blocks.py:119, components/encoder_decoder.py:138-141) -----------
 * out = a + LN2(fc2(GELU_erf(fc1(a))));  if zsum != NULL: zsum = out + z (the next block's
 * input).  out may be NULL when only zsum is wanted.  n_tok = B*ph*pw.                     */
int b200mar_mlp_fwd(const float* a, const float* z,
                    const float* fc1_w /*[128,64]*/, const float* fc1_b /*[128]*/,
                    const float* fc2_w /*[64,128]*/, const float* fc2_b /*[64]*/,
                    const float* ln_w, const float* ln_b,
                    float* out, float* zsum, long long n_tok, void* stream);
/* Same contract as b200mar_mlp_fwd, computed on the tcgen05 tensor cores (kind::f16 operands with
 * power-of-two range scaling, fp32 accumulate in TMEM): 128-token tiles, four workers per CTA. */
int b200mar_mlp_tc_fwd(const float* a, const float* z,
                       const float* fc1_w, const float* fc1_b,
                       const float* fc2_w, const float* fc2_b,
                       const float* ln_w, const float* ln_b,
                       float* out, float* zsum, float* a_out, int a_is_delta,
                       long long n_tok, void* stream);
/* a_is_delta = 1: `a` holds LN1(attn) only (b200mar_attn_tc_fwd with residual = 0) and the kernel
 * forms a = z + a itself (z required); a_out (optional) receives that sum for the backward.   */
/* In-place form for the blocks inside a stage (components/encoder_decoder.py:138-141 consumes
 * only the sum): out = a + LN2(mlp(a)) (optional, may be NULL) and z_inout += out (optional, may be
 * NULL; at least one of the two).  Same arithmetic as b200mar_mlp_tc_fwd; the stream moves through
 * TMA bulk tensor loads/stores and the sum is a TMA reduce-add, so z is never read by the SMs.
 * Returns -2/-3 if the driver's tensor-map encoder is unavailable / rejects the buffers.          */
int b200mar_mlp_tma_fwd(const float* a, float* z_inout,
                        const float* fc1_w, const float* fc1_b,
                        const float* fc2_w, const float* fc2_b,
                        const float* ln_w, const float* ln_b,
                        float* out, long long n_tok, void* stream);
/* Precision level "tf32c" of the MLP half: the same contract with fc1 / fc2
 * (models/generator/components/transformer_blocks.py:119) on compensated FP16 hi + lo operand pairs (three MMA
 * passes, ~22 significant bits) and the 1.5e-7 erf GELU; two workers per persistent CTA.                    */
int b200mar_mlp_tma_fwd_comp(const float* a, float* z_inout,
                        const float* fc1_w, const float* fc1_b,
                        const float* fc2_w, const float* fc2_b,
                        const float* ln_w, const float* ln_b,
                        float* out, long long n_tok, void* stream);
/* ---- backward twins of the fused glue (training path).  Parameter gradients are ACCUMULATED,
 * activation gradients written.
 * final_recon_bwd: backward of b200mar_final_recon_fwd (models/generator/components/reconstruction.py:14-19,
 *   models/generator/ngswin.py:190,204,208): dy, y [B,1,H0,W0]; taps [B*H*W, 9] as written by the forward;
 *   g1_scratch [B*H*W] floats; ddec, dc0 [B*H*W, 64] written; d_w1 [64,9], d_b1 [1], d_w2 [9], d_b2 [1], d_ln_* [64].
 * shallow_tok_bwd: backward of b200mar_shallow_tok_fwd (components/encoder_decoder.py:23-27): dc0 [B*H*W, 64] ->
 *   dw [64,9], db [64] accumulated; dx [B,1,H,W] written if not NULL.                                              */
int b200mar_final_recon_bwd(const float* dy, const float* y, const float* taps, const float* dec,
                            const float* c0, const float* ln_w, const float* ln_b, const float* w1,
                            const float* b1, const float* w2, float* g1_scratch, float* ddec, float* dc0,
                            float* d_ln_w, float* d_ln_b, float* d_w1, float* d_b1, float* d_w2,
                            float* d_b2, int B, int H, int W, int H0, int W0, void* stream);
/* cascade_proj_bwd: backward of the cascade_proj forward, fp32 or tcgen05 (components/pooling.py:13-17 InterPool with
 *   nn.MaxPool2d's first-maximum routing, models/generator/ngswin.py:175-178, encoder_decoder.py:117,136): dout
 *   [B*h*wd, 64]; dpool (may be NULL) = gradient arriving on the returned poolA; d_srcA / d_srcB [B*4*h*wd, 64] and
 *   d_direct [B*h*wd, 64] written; dw [64, K], db [64] accumulated (K = 128 without srcB, 192 with).                */
int b200mar_cascade_proj_bwd(const float* srcA, const float* srcB, const float* direct, const float* w,
                             const float* dout, const float* dpool, float* d_srcA, float* d_srcB,
                             float* d_direct, float* dw, float* db, int B, int h, int wd, void* stream);
/* bottleneck_bwd: backward of the bottleneck_cat + bottleneck_tail (fp32 or tcgen05) forward pair (SCDPBottleneck,
 *   components/encoder_decoder.py:159-164,188-200; BottleneckPool pooling.py:32-37; the "+ e1" of ngswin.py:189):
 *   cat [B*H*W, 84] as written by the forward, dout [B*H*W, 64]; ddwc_scratch [B*H*W, 84] floats;
 *   dc0, de1 [B*H*W, 64], dp0, de2 [B*H*W/4, 64], dp00, de3 [B*H*W/16, 64] written (de1 includes dout, the tail's
 *   residual); d_dw_w [84,9], d_dw_b [84], d_pw_w [64,84], d_pw_b [64], d_ln_w, d_ln_b [64] accumulated.           */
int b200mar_bottleneck_bwd(const float* cat, const float* dout, const float* c0, const float* p0,
                           const float* p00, const float* dw_w, const float* dw_b, const float* pw_w,
                           const float* pw_b, const float* ln_w, float* ddwc_scratch, float* dc0,
                           float* dp0, float* dp00, float* de1, float* de2, float* de3, float* d_dw_w,
                           float* d_dw_b, float* d_pw_w, float* d_pw_b, float* d_ln_w, float* d_ln_b,
                           int B, int H, int W, void* stream);
int b200mar_shallow_tok_bwd(const float* x, const float* w, const float* dc0, float* dw, float* db,
                            float* dx, int B, int H, int W, void* stream);
/* ---- precision mode "bf16" (BASELINE.json configs[1]: "bf16 on 1xB200"): the NSTB residual
 * stream inside a stage is BF16 in HBM, which halves every pass of the three per-block kernels.  Arithmetic is
 * unchanged (fp32 q/k normalisation, softmax, LayerNorm, residual adds; FP16 / TF32 tensor-core operands): rows are
 * widened to fp32 on read and rounded to bf16 once, on write; z16 += out is a bf16 TMA reduce-add.  Same reference
 * contracts as their fp32-stream siblings: components/window_utils.py:75-90 (unigram + n-gram qkv),
 * components/attention.py:70-101 + transformer_blocks.py:83-117 (attention + LN1 + residual),
 * transformer_blocks.py:119 + encoder_decoder.py:138-141 (MLP + LN2 + residual + within-stage sum).
 * z16 / a16 / out16: bf16 [B*ph*pw, 64]; ctx, nq and every parameter: fp32.                                      */
/* stage entry / exit of the mode: dst = bf16(src) (to_bf16 = 1) or dst = fp32(src) (to_bf16 = 0); n elements,
 * n % 8 == 0 (the reference has no counterpart: its stream is fp32 throughout).                                  */
int b200mar_cast_stream(const void* src, void* dst, long long n, int to_bf16, void* stream);
int b200mar_unigram_qkv_fwd_s16(const void* z16, const float* weight, const float* bias,
                                const float* qkv_w, const float* qkv_b, float* nq,
                                int B, int ph, int pw, int nh, void* stream);
int b200mar_attn_tc_fwd_s16(const void* z16, const float* ctx,
                            const float* qkv_w, const float* qkv_b,
                            const float* proj_w, const float* proj_b,
                            const float* logit_scale, const float* bias_table,
                            const float* ln_w, const float* ln_b,
                            void* a16, int B, int ph, int pw, int nh, int shift, void* stream);
int b200mar_mlp_tma_fwd_s16(const void* a16, void* z16_inout,
                            const float* fc1_w, const float* fc1_b,
                            const float* fc2_w, const float* fc2_b,
                            const float* ln_w, const float* ln_b,
                            void* out16, long long n_tok, void* stream);
/* 1 if a tensor-core kernel of this process ever timed out on an mbarrier (every wait is
 * bounded so a descriptor bug cannot hang the device); 0 otherwise; -1 on a CUDA error.     */
int b200mar_tc_error(void);
/* da = dout + d(MLP branch) (overwrites da); parameter gradients accumulated.             */
int b200mar_mlp_bwd(const float* a, const float* dout,
                    const float* fc1_w, const float* fc1_b,
                    const float* fc2_w, const float* fc2_b,
                    const float* ln_w, const float* ln_b,
                    float* da,
                    float* d_fc1_w, float* d_fc1_b, float* d_fc2_w, float* d_fc2_b,
                    float* d_ln_w, float* d_ln_b, long long n_tok, void* stream);
/* mlp_bwd on the tcgen05 tensor cores (BF16 operands, fp32 accumulate; timm Mlp + norm2 backward,
 * components/transformer_blocks.py:119).  `scratch` is a caller-owned device buffer of
 * n_tok * 768 bytes (the bf16 operands a, g, dh, dy of the weight-gradient MMAs).            */
int b200mar_mlp_bwd_tc(const float* a, const float* dout,
                       const float* fc1_w, const float* fc1_b,
                       const float* fc2_w, const float* fc2_b,
                       const float* ln_w, const float* ln_b,
                       float* da,
                       float* d_fc1_w, float* d_fc1_b, float* d_fc2_w, float* d_fc2_b,
                       float* d_ln_w, float* d_ln_b, void* scratch, long long n_tok, void* stream);
/* Weight gradient of one nn.Linear on the tensor cores: dw[cout][cin] += sum_t g16[t][cout] *
 * in16[t][cin] and (optionally, cin == 64) dbias[cout] += sum_t g16[t][cout]; in16 / g16 are
 * row-major bf16 device buffers; cin in {64,128}, cout in {64,...,256} a multiple of 32.       */
int b200mar_wgrad_tc(const void* in16, const void* g16, float* dw, float* dbias, long long n_tok,
                     int cin, int cout, void* stream);

/* ---- a10: PatchMerging.forward (components/encoder_decoder.py:51-74) -------------------
 * 2x2 gather in channel order (even,even),(even,odd),(odd,even),(odd,odd) -> LN(256) ->
 * Linear(256->64, no bias).  x [B,ph,pw,64] -> y [B,ph/2,pw/2,64].                          */
int b200mar_patch_merge_fwd(const float* x, const float* ln_w /*[256]*/, const float* ln_b,
                            const float* red_w /*[64,256]*/, float* y,
                            int B, int ph, int pw, void* stream);
/* Same contract on TMA + tcgen05 (kind::tf32): stride-2 tensor-map boxes are the MMA A operands,
 * LayerNorm folded around the GEMM (per-quadrant centring), TMA store.  ph, pw even.          */
int b200mar_patch_merge_tc_fwd(const float* x, const float* ln_w, const float* ln_b,
                               const float* red_w, float* y, int B, int ph, int pw, void* stream);
int b200mar_patch_merge_bwd(const float* x, const float* dy,
                            const float* ln_w, const float* ln_b, const float* red_w,
                            float* dx, float* d_ln_w, float* d_ln_b, float* d_red_w,
                            int B, int ph, int pw, void* stream);

/* ---- a12 / a13: LayerNorm(64) of the stage glue on the TRAINING path - SCDPBottleneck.norm
 * (components/encoder_decoder.py:199-200) and NGswin.norm (ngswin.py:204); eps 1e-5, affine.
 * x, y, dy, dx [n_tok,64].  bwd recomputes the row statistics from x; dweight += , dbias += .  */
int b200mar_ln64_fwd(const float* x, const float* weight /*[64]*/, const float* bias /*[64]*/,
                     float* y, long long n_tok, void* stream);
int b200mar_ln64_bwd(const float* x, const float* dy, const float* weight, float* dx,
                     float* dweight, float* dbias, long long n_tok, void* stream);

/* ---- the image-space terms of the generator loss, values and dL/dfake in one
 * pass over (fake, real, ct), each [nimg, H, W] fp32 (losses/gan_losses.py:48-219 as combined in
 * training/train_combined.py:456-485):
 *   M = ct > threshold;  w = min(1 + beta * maxpool_{2r+1}(M), w_max);
 *   rec = mean|w (fake-real)|;  edge = mean w (|gx(fake)-gx(real)| + |gy(fake)-gy(real)|) with forward
 *   differences whose last column / row is replicated;  metal = mean|M (fake-real)|.
 * sums4 += {rec, edge, metal, lam_rec*rec + lam_edge*edge + lam_metal*metal} (zeroed by the caller);
 * dfake = d(sums4[3]) / d(fake), written (not accumulated).  H, W >= 2, 0 <= radius <= 32.   */
int b200mar_metal_losses(const float* fake, const float* real, const float* ct, float* dfake,
                         float* sums4, int nimg, int H, int W, int radius, float threshold,
                         float beta, float w_max, float lam_rec, float lam_edge, float lam_metal,
                         void* stream);

/* ---- parallel-beam Radon projector of the physics-consistency loss
 * (training/train_combined.py:269-272 `Radon(PATCH_SIZE, angles)`; losses/gan_losses.py:222-253
 * `projector.forward`).  torch-radon is an empty submodule of the reference: restated from its
 * published algorithm, parity unpinned.
 * fwd: img [nimg,H,W] -> sino [nimg,n_angles,det_count]; ray-driven, detector k on the line
 *      x cos th + y sin th = (k - det_count/2 + 1/2) * det_spacing (image-centred coordinates),
 *      <= 1-pixel steps sampled at their midpoints with bilinear interpolation, zero outside.
 * bwd: sino -> img, pixel-driven backprojection with linear detector interpolation (what
 *      torch-radon's autograd backward runs).  angles: device float[n_angles], n_angles <= 1024. */
int b200mar_radon_fwd(const float* img, const float* angles, float* sino, int nimg, int H, int W,
                      int n_angles, int det_count, float det_spacing, void* stream);
int b200mar_radon_bwd(const float* sino, const float* angles, float* img, int nimg, int H, int W,
                      int n_angles, int det_count, float det_spacing, void* stream);

/* ---- a12: pixel_shuffle_permute of the three encoder skips + channel concat
 * (components/encoder_decoder.py:159-164,191-192) ----------------------------------------
 * out[b, (h*s+a)*W + (w*s+c'), off_i + c] = x_i[b, h*pw_i + w, (c*s + a)*s + c'],  s = 2^i,
 * channel blocks 64 | 16 | 4 -> out [B, H*W, 84].  inverse=1 runs the same map backwards
 * (the gradient): reads `out`-layout data from `cat` and writes x0/x1/x2-layout buffers.   */
int b200mar_pixel_shuffle_cat(float* x0 /*[B,H*W,64]*/, float* x1 /*[B,H*W/4,64]*/,
                              float* x2 /*[B,H*W/16,64]*/, float* cat /*[B,H*W,84]*/,
                              int B, int H, int W, int inverse, void* stream);

/* ---- a2, a9, a11, a12, a13, a14: the generator glue between the transformer stages, fused and
 * kept in the channels-last token layout (inference path; training uses the stock-PyTorch
 * composition of the same modules so that autograd covers it) ------------------------------
 * a2  ShallowExtractor.conv1 (components/encoder_decoder.py:23-27) + the token rearrange of
 *     ngswin.py:172:  x [B,1,H,W] -> c0 [B,H*W,64].                                          */
int b200mar_shallow_tok_fwd(const float* x, const float* w /*[64,1,3,3]*/, const float* b /*[64]*/,
                            float* c0, int B, int H, int W, void* stream);
/* a11+a9 InterPool (components/pooling.py:13-17) of the cascade, torch.cat (ngswin.py:175,178)
 *     and EncoderLayer.across_cascade_proj (components/encoder_decoder.py:117,136):
 *     out[b,t,:] = w * cat(maxpool2x2(srcA)[t], maxpool2x2(srcB)[t] (if srcB), direct[t]) + bias.
 *     srcA/srcB [B,2h,2wd,64], direct/out [B,h*wd,64], w [64,128] (srcB NULL) or [64,192];
 *     poolA (optional) receives maxpool2x2(srcA).                                            */
int b200mar_cascade_proj_fwd(const float* srcA, const float* srcB, const float* direct,
                             const float* w, const float* bias, float* out, float* poolA,
                             int B, int h, int wd, void* stream);
/* Same contract on TMA + tcgen05 (kind::tf32): stride-2 boxes + running max build the pooled
 * operands in place, the direct tile is an MMA operand as delivered, TMA stores.              */
int b200mar_cascade_proj_tc_fwd(const float* srcA, const float* srcB, const float* direct,
                                const float* w, const float* bias, float* out, float* poolA,
                                int B, int h, int wd, void* stream);
/* a12 SCDPBottleneck front (components/encoder_decoder.py:159-164,191-192; pooling.py:32-37):
 *     cat[b,t,:] = [ e1 + lrelu(c0) | shuffle2(e2 + lrelu(p0)) | shuffle4(e3 + lrelu(p00)) ],
 *     p0 = maxpool2(c0), p00 = maxpool4(c0) (token-major), cat [B,H*W,84].                   */
int b200mar_bottleneck_cat_fwd(const float* c0, const float* p0, const float* p00,
                               const float* e1, const float* e2, const float* e3,
                               float* cat, int B, int H, int W, void* stream);
/* a12+a13 SCDPBottleneck tail (components/encoder_decoder.py:194-200) + the decoder input add
 *     (ngswin.py:189): out = LN(pointwise(GELU(depthwise3x3(cat)))) + e1.                    */
int b200mar_bottleneck_tail_fwd(const float* cat, const float* dw_w /*[84,1,3,3]*/,
                                const float* dw_b, const float* pw_w /*[64,84]*/,
                                const float* pw_b, const float* ln_w, const float* ln_b,
                                const float* e1, float* out, int B, int H, int W, void* stream);
/* Same contract on TMA + tcgen05: one bulk tensor load per 18x10x84 halo (zero padding = the map's
 * out-of-bounds fill), depthwise + GELU per token thread, pointwise as fp16 MMAs, TMA store.    */
int b200mar_bottleneck_tail_tc_fwd(const float* cat, const float* dw_w, const float* dw_b,
                                   const float* pw_w, const float* pw_b,
                                   const float* ln_w, const float* ln_b,
                                   const float* e1, float* out, int B, int H, int W, void* stream);
/* a13+a14 decoder tail: LN(norm) + global skip (ngswin.py:190,204), Reconstruction
 *     (components/reconstruction.py:14-19) and the crop of ngswin.py:208:
 *     out [B,1,H0,W0] = tanh(conv3x3(conv3x3(LN(dec) + c0; w1,b1); w2,b2))[:H0,:W0].
 *     taps [B,H*W,9] is caller-provided scratch.                                             */
int b200mar_final_recon_fwd(const float* dec, const float* c0, const float* ln_w,
                            const float* ln_b, const float* w1 /*[1,64,3,3]*/, const float* b1,
                            const float* w2 /*[1,1,3,3]*/, const float* b2, float* taps,
                            float* out, int B, int H, int W, int H0, int W0, void* stream);

/* ---- integer index maps (bit-exact contract of a4/a6/a7) -------------------------------
 * tok_idx/ctx_idx [B*wh*ww, 64] int64: flat source token / source window of every window
 * slot, exactly what the attention kernels use (same device function).                     */
int b200mar_partition_index(long long* tok_idx, long long* ctx_idx,
                            int B, int ph, int pw, int shift, void* stream);
/* mask [wh*ww, 64, 64] fp32 (0 / -100) from the kernels' analytic region rule, and
 * rel_index [64,64] int64 from the kernels' bias addressing.                               */
int b200mar_shift_mask(float* mask, int ph, int pw, int shift, void* stream);
int b200mar_rel_index(long long* rel_index, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200MAR_H */
