/*
 * iit_b200 -- C-ABI of the B200 (sm_100a) kernels behind the interchange-intervention step.
 *
 * Plain `extern "C"`: raw device pointers, POD descriptors, a cudaStream_t passed as void*.
 * No torch types. Every entry point returns 0 on success; on failure it returns non-zero and
 * iit_last_error() holds a message (thread-local). Unless a comment says otherwise, no entry
 * point allocates, frees, retains pointers or synchronises: the caller owns every buffer
 * (inputs, outputs, workspaces) and the kernels are enqueued on the given stream.
 *
 * The reference (tkwa/iit) is pure Python over PyTorch eager ops and has no FFI of its own; each
 * entry point below names the reference lines whose arithmetic it replaces (paths relative to
 * the reference root). INTEGRATION.md shows the ctypes binding a maintainer would add.
 */
#ifndef IIT_B200_H
#define IIT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IIT_MAX_RANK 6

enum { IIT_F32 = 0, IIT_BF16 = 1 };

/* ------------------------------------------------------------------------------------------ */
/* library                                                                                     */
/* ------------------------------------------------------------------------------------------ */
const char* iit_last_error(void);
int iit_version(void);
/* number of kernels this library has enqueued since load (bench.py's gpu_launches) */
unsigned long long iit_launch_count(void);
/* Which tensor-core kernel instantiation the convolution dispatchers picked, counted per launch since load: slot ids
 * 0 .. iit_kernel_variant_slots()-1, each with the kernel's name. Diagnostics for tests and profiles (they assert that a
 * configuration really ran on, e.g., the CTA-pair kernel); no reference counterpart. */
int iit_kernel_variant_slots(void);
const char* iit_kernel_variant_name(int slot);
unsigned long long iit_kernel_variant_count(int slot);

/* ------------------------------------------------------------------------------------------ */
/* (a) patch: TorchIndex box gather/scatter                                                     */
/* A box is what `TorchIndex.as_index` (iit/utils/index.py:36-38) selects when every entry is   */
/* None, an int or a step-less slice: per dim a [start, start+extent) range. Strides are in     */
/* elements, so NCHW-shaped views of channels-last memory patch with contiguous C runs.          */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
  int32_t rank;
  int32_t elem_bytes;               /* 2, 4 or 8 */
  int64_t start[IIT_MAX_RANK];
  int64_t extent[IIT_MAX_RANK];
  int64_t dst_stride[IIT_MAX_RANK];
  int64_t src_stride[IIT_MAX_RANK];
} iit_box_t;

/* dst[box] = src[box], in place on dst.
 * Replaces `out = hook_point_out.clone(); out[idx] = cache[name][idx]`
 * (iit/model_pairs/base_model_pair.py:157-161, and the HL twin at :125-137). */
int iit_patch_fwd(void* dst, const void* src, const iit_box_t* box, void* stream);

/* grad[box] = 0, in place: the autograd CopySlices backward of the assignment above
 * (the patched values are constants; base_model_pair.py:160 under autograd). */
int iit_patch_bwd(void* grad, const iit_box_t* box, void* stream);

/* dst_compact = src[box] packed densely in box order (source-run capture of only the
 * Correspondence's slice; replaces caching every hook point, base_model_pair.py:86). */
int iit_capture_box(void* dst_compact, const void* src, const iit_box_t* box, void* stream);

/* ------------------------------------------------------------------------------------------ */
/* (c) high-level models and losses                                                             */
/* ------------------------------------------------------------------------------------------ */
/* MNIST-PVR HL with one variable interchanged: v = base_vars; v[:, var] = src_vars[:, var]
 * (var < 0: no interchange); hl_out[b] = (tr,bl,br)[class_map[tl]-1].
 * Replaces MNIST_PVR_HL.forward + hl_ablation_hook (iit/tasks/mnist_pvr/pvr_hl.py:44-56,
 * base_model_pair.py:80,89-91,120-142) and the leaky variant (pvr_check_leaky_hl.py:45-66).
 * A top-left variable outside [-10, 9] (where the reference's indexing raises IndexError) gives hl_out[b] = -1 and
 * sets *err_flag (device int32, may be NULL) to 1. */
int iit_hl_pvr(const int64_t* base_vars, const int64_t* src_vars, int var, int64_t batch,
               int64_t* hl_out, int32_t* err_flag, void* stream);

enum { IIT_IOI_NONE = 0, IIT_IOI_ALL_NODES = 1, IIT_IOI_DUPLICATE = 2, IIT_IOI_S_INHIBITION = 3,
       IIT_IOI_NAME_MOVER = 4 };

/* Which high-level model produces the class index a row of LL logits is scored against (iit_hl_ce_fwd). */
typedef struct {
  int32_t mode;          /* 0: targets are given; 1: MNIST-PVR HL (pvr_hl.py:44-56); 2: IOI HL, last position (ioi_hl.py:102-130) */
  int32_t node;          /* mode 1: index of the interchanged variable (-1: none); mode 2: IIT_IOI_* */
  const int64_t* base;   /* mode 1: base run's intermediate variables [rows,4]; mode 2: base tokens [rows,seq] */
  const int64_t* src;    /* the same from the source run (NULL: no interchange) */
  const int64_t* names;  /* mode 2: sorted name token ids */
  int32_t n_names;
  int32_t seq;           /* mode 2: tokens per prompt (<= 64) */
  int64_t d_vocab;       /* mode 2 */
} iit_hl_spec_t;

/* IOI HL with one hook interchanged, evaluated sparsely at the LAST position:
 * out_argmax[b] = argmax_v logits[b, S-1, v] with first-index tie-breaking (torch.argmax).
 * Replaces IOI_HL.forward + heads (iit/tasks/ioi/ioi_hl.py:16-24,37-46,54-68,102-130) and the
 * argmax at ioi_model_pair.py:64,82 without materialising [B,S,V]. names must be sorted. */
int iit_hl_ioi_last_argmax(const int64_t* base_tok, const int64_t* src_tok, int node,
                           const int64_t* names, int n_names, int64_t batch, int seq, int64_t d_vocab,
                           int64_t* out_argmax, void* stream);

/* Dense IOI HL output [B,S,V] fp32 (API compatibility for do_intervention's hl_output). */
int iit_hl_ioi_dense(const int64_t* base_tok, const int64_t* src_tok, int node,
                     const int64_t* names, int n_names, int64_t batch, int seq, int64_t d_vocab,
                     float* out, void* stream);

/* Cross-entropy over rows with a class-index target, mean over rows:
 *   row_lse[r] = logsumexp(logits[r,:]);  loss += (row_lse[r] - logits[r,target[r]]) / rows
 * loss must be zeroed by the caller (one float). Replaces t.nn.CrossEntropyLoss()(ll, hl)
 * (iit_behavior_model_pair.py:20-30, base_model_pair.py:173) and, with target = argmax of the
 * one-hot, CE with a probability target (ioi_model_pair.py:27-35,64-68). */
int iit_ce_index_fwd(const void* logits, int dtype, int64_t rows, int64_t classes, int64_t row_stride,
                     const int64_t* target, float* row_lse, float* loss, void* stream);

/* The high-level side of a loss term and the loss in ONE launch (north_star (c): "the HL model's intervened evaluation fused
 * with the losses"): per row the class index comes from `hl` -- given (mode 0), MNIST_PVR_HL with one variable interchanged
 * (mode 1; replaces pvr_hl.py:44-56 + base_model_pair.py:80,89-91), or the argmax of IOI_HL's last-position logits with one
 * hook interchanged (mode 2; ioi_hl.py:102-130 + ioi_model_pair.py:64-66) -- and is written to target[] (modes 1, 2); then
 * row_lse[r] = logsumexp(logits[r, :classes]), and *loss = mean_r (row_lse[r] - logits[r, target[r]]) (written, not added;
 * fixed summation order). row_nll: fp32 [rows] scratch. ticket: one device uint32 that is zero before the first call (the
 * kernel leaves it zero). A PVR pointer variable outside [-10, 9] or a target outside [0, classes) sets *err_flag (may be
 * NULL) and gives NaN, where torch raises. Replaces CrossEntropyLoss (iit_behavior_model_pair.py:20-30, base_model_pair.py:173). */
int iit_hl_ce_fwd(const void* logits, int dtype, int64_t rows, int64_t classes, int64_t row_stride, const iit_hl_spec_t* hl,
                  int64_t* target, float* row_lse, float* row_nll, float* loss, unsigned int* ticket, int32_t* err_flag,
                  void* stream);

/* dlogits[r,c] = (exp(logits[r,c]-row_lse[r]) - [c==target[r]]) * (*grad_out) / rows for c < classes; when d_row_stride >
 * classes the columns classes .. d_row_stride-1 of every row are written too, as zeros (padded logits get a padded gradient). */
int iit_ce_index_bwd(const void* logits, int dtype, int64_t rows, int64_t classes, int64_t row_stride,
                     const int64_t* target, const float* row_lse, const float* grad_out,
                     void* dlogits, int64_t d_row_stride, void* stream);

/* Same with a dense probability target prob[r,c] (fp32), as F.cross_entropy does for float
 * targets (ioi_model_pair.py:29,35): loss_r = lse_r * sum_c p - sum_c p*l. */
int iit_ce_prob_fwd(const void* logits, int dtype, int64_t rows, int64_t classes, int64_t row_stride,
                    const float* prob, int64_t p_row_stride, float* row_lse, float* row_psum,
                    float* loss, void* stream);
int iit_ce_prob_bwd(const void* logits, int dtype, int64_t rows, int64_t classes, int64_t row_stride,
                    const float* prob, int64_t p_row_stride, const float* row_lse, const float* row_psum,
                    const float* grad_out, void* dlogits, int64_t d_row_stride, void* stream);

/* ------------------------------------------------------------------------------------------ */
/* (d) interchange-intervention accuracy                                                        */
/* ------------------------------------------------------------------------------------------ */
/* top1[r] = argmax_c logits[r,c] (first maximal index; NaN counts as maximal, as torch.argmax);
 * counts[0] += #{r : top1[r]==hl[r] and (changed==NULL or changed[r])}
 * counts[1] += #{r : changed==NULL or changed[r]}
 * top1 may be NULL; hl may be NULL (then only top1 is written). counts is accumulated, not
 * cleared. Replaces argmax/eq/float().mean() (iit_behavior_model_pair.py:118-120,
 * ioi_model_pair.py:90-92) and the masked reduction of eval_causality.py:74-83. */
int iit_iia_count(const void* logits, int dtype, int64_t rows, int64_t classes, int64_t row_stride,
                  const int64_t* hl, const uint8_t* changed, int64_t* top1,
                  unsigned long long* counts, void* stream);

/* ------------------------------------------------------------------------------------------ */
/* (b) low-level ResNet-18 layers, NHWC ("channels-last") activations                           */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
  int32_t n, h, w, c;      /* input  [n,h,w,c]            */
  int32_t k, r, s;         /* filter [k,r,s,c] (KRSC)     */
  int32_t stride, pad;
  int32_t p, q;            /* output [n,p,q,k]            */
} iit_conv_t;

/* y = conv(x, w). Replaces torch's cuDNN conv for every Conv2d of torchvision resnet18
 * (iit/tasks/mnist_pvr/get_alignment.py:11-15). If bn_sum/bn_sqsum are non-NULL (fp64 [k],
 * zeroed by the caller) the epilogue also accumulates per-channel sum and sum of squares of y
 * for the BatchNorm that follows. */
int iit_conv2d_fprop(const void* x, const void* w, void* y, const iit_conv_t* d, int dtype,
                     double* bn_sum, double* bn_sqsum, void* stream);
/* dx = conv_transpose(dy, w);  w_crsk is the filter re-laid out as [c,r,s,k]. */
int iit_conv2d_dgrad(const void* dy, const void* w_crsk, void* dx, const iit_conv_t* d, int dtype,
                     void* stream);
/* dw[k,r,s,c] (fp32) = sum over pixels dy * x. scratch: iit_conv2d_wgrad_scratch_bytes(). */
size_t iit_conv2d_wgrad_scratch_bytes(const iit_conv_t* d);
int iit_conv2d_wgrad(const void* dy, const void* x, float* dw, const iit_conv_t* d, int dtype,
                     void* scratch, size_t scratch_bytes, void* stream);
/* Same result as iit_conv2d_wgrad, for callers that cannot give the kernels a scratch buffer because
 * they must not (or cannot) allocate device memory: FFI plug-ins that only forward the caller's tensors,
 * processes that reserved the whole device at start-up and may not call cudaMalloc afterwards, loops that
 * cannot afford cudaMalloc's device-wide synchronisation. This entry point has no counterpart in the
 * reference. The split-K partial sums go to an 8 MiB array built into the library (each GPU that loads
 * the library holds those 8 MiB until the process exits); a problem that needs more is rejected with
 * an error naming both sizes. It calls no cudaMalloc / cudaFree / allocating routine. The first call on
 * a device looks up the array's device address (once per device, with that device current): make that
 * first call before starting any stream capture. All calls share the one array and it keeps the last
 * call's partial sums (nothing needs clearing: a launch overwrites what it reads), so two calls whose
 * kernels may overlap on the GPU - different streams - race with each other: serialise them. */
int iit_conv2d_wgrad_noalloc(const void* dy, const void* x, float* dw, const iit_conv_t* d, int dtype,
                             void* stream);

/* Tensor-core path (tcgen05.mma + TMA + TMEM accumulator) for bf16 3x3 stride-1 pad-1 convolutions whose
 * channel counts are multiples of 64 -- 80 % of ResNet-18's FLOPs. Same results contract as
 * iit_conv2d_fprop / iit_conv2d_dgrad (fp32 accumulation, bf16 output); w_krsc / w_crsk are bf16.
 * iit_conv2d_tc_supported() returns 1 when a descriptor qualifies. */
int iit_conv2d_tc_supported(const iit_conv_t* d, int dtype);
int iit_conv2d_fprop_tc(const void* x, const void* w_krsc, void* y, const iit_conv_t* d, double* bn_sum,
                        double* bn_sqsum, void* stream);
/* Evaluation-mode convolution + BatchNorm (+ residual) (+ ReLU) in one kernel: y = act(conv(x, w) * scale[k] + shift[k]
 * (+ residual)), scale = gamma / sqrt(running_var + eps), shift = beta - running_mean * scale (iit_bn_eval_affine), applied to
 * the fp32 accumulator in the epilogue. residual (may be NULL) is laid out like y. The evaluation passes of the causality
 * sweep and of run_eval_step use it wherever no hook sits between the convolution and its BatchNorm. */
int iit_conv2d_fprop_tc_affine(const void* x, const void* w_krsc, void* y, const iit_conv_t* d, const float* scale,
                               const float* shift, const void* residual, int relu, void* stream);
int iit_bn_eval_affine(const float* gamma, const float* beta, const float* running_mean, const float* running_var, int c,
                       float eps, float* scale, float* shift, void* stream);
int iit_conv2d_dgrad_tc(const void* dy, const void* w_crsk, void* dx, const iit_conv_t* d, void* stream);
/* dx += dgrad(dy): the same kernels with their epilogue's stores replaced by 16-byte bf16 reductions (the addition is done
 * in L2; every element receives exactly one addend, so the result is order-independent and equals a bf16 tensor addition).
 * Where a tensor feeds two consumers -- a residual block's input goes to conv1 and to the skip path -- the second gradient
 * is added straight onto the first: the sum autograd would otherwise make with a separate element-wise kernel
 * (torchvision BasicBlock.forward `out += identity`, its backward). */
int iit_conv2d_dgrad_tc_accumulate(const void* dy, const void* w_crsk, void* dx, const iit_conv_t* d, void* stream);
/* The interchange patch INSIDE the hooked layer (north_star (a); replaces the Python forward hook of
 * base_model_pair.py:151-163 for a convolution's hook point): iit_conv2d_fprop_tc whose epilogue, for output rows p in
 * [box6[0], box6[1]), columns q in [box6[2], box6[3]) and channels in [box6[4], box6[5]) of every image, stores src's value
 * (the cached source-run activation, laid out like y) instead of the convolution's; the BatchNorm sums are taken from the
 * patched values, as a BatchNorm after the hook would see them. iit_conv2d_patch_fusable: 1 when the descriptor runs on the
 * kernel that carries this epilogue. */
int iit_conv2d_patch_fusable(const iit_conv_t* d);
int iit_conv2d_fprop_tc_patch(const void* x, const void* w_krsc, void* y, const iit_conv_t* d, double* bn_sum, double* bn_sqsum,
                              const void* src, const int32_t* box6, void* stream);
/* The transformer's linear maps (transformer_lens HookedTransformer: attention projections, MLP, unembedding; reference
 * train_ioi.py:30-34) on the same tensor-core GEMM with their element-wise neighbours applied to the fp32 accumulator in the
 * epilogue, before the one rounding to bf16:
 *   iit_gemm_bias_act_tc:   y[rows, n_out] = act(x[rows, n_in] @ w_oi[n_out, n_in]^T + bias (+ resid)), act 0 identity /
 *                           1 gelu_new. With act 1, pre_out (may be NULL) receives the pre-activation for the backward
 *                           (mlp.hook_pre's value); resid must then be NULL. bias fp32 [n_out] (may be NULL).
 *   iit_gemm_dgrad_gelu_tc: dpre[rows, n_in] = (dy[rows, n_out] @ w_io[n_in, n_out]^T) * gelu_new'(pre[rows, n_in]) -- the
 *                           input gradient of the MLP's output map already taken through the activation; col_sum / col_sq
 *                           (fp64 [n_in], both or neither; ADDED to) receive the column sums of dpre (= the gradient of
 *                           the MLP's input bias) and of its squares (a by-product of the shared epilogue).
 * Widths are multiples of 64, tensors bf16, 16-byte aligned. */
int iit_gemm_bias_act_tc(const void* x, const void* w_oi, const float* bias, const void* resid, void* y, void* pre_out,
                         int64_t rows, int n_in, int n_out, int act, void* stream);
int iit_gemm_dgrad_gelu_tc(const void* dy, const void* w_io, const void* pre, void* dpre, double* col_sum, double* col_sq,
                           int64_t rows, int n_in, int n_out, void* stream);
/* iit_conv2d_wgrad_tc with the split-K reduction written straight INTO parameter gradients whose layout is not the
 * [k][r][s][c] matrix (the transformer's per-head [H, d_model, d_head] tensors and [in][out] matrices): for every map m and
 * destination element idx (decomposed by size[] in row-major order), dst[idx] += sum over partials of the matrix element at
 * offset + a*stride[0] + b*stride[1] + c*stride[2]. maps_dev: DEVICE array of n_maps (<= 16) descriptors; max_map_numel: the
 * largest destination. Replaces the strided additions autograd would otherwise make per parameter. */
typedef struct {
  float* dst;
  int32_t size[3];
  int32_t reserved;
  int64_t stride[3];
  int64_t offset;
} iit_wgrad_map_t;
int iit_conv2d_wgrad_tc_scatter(const void* dy, const void* x, const iit_conv_t* d, void* scratch, size_t scratch_bytes,
                                const iit_wgrad_map_t* maps_dev, int n_maps, int64_t max_map_numel, void* stream);
/* 3x3 / stride-1 / pad-1 convolutions with 64 -> 64 channels on images whose padded form fits shared memory
 * (layer1 of the ResNet) run on a "shifted-window" kernel inside the two entry points above: the padded image
 * is loaded once and every filter tap addresses it at a row offset. Same results bit for bit; this switch
 * (returns the previous setting) exists so that tests and profiles can compare the two kernels. */
int iit_conv_win_enable(int on);
/* Layers with a multiple of 256 filters run on a CTA-pair kernel (tcgen05 cta_group::2: two SMs share one
 * 256 pixel x 256 filter tile, each loading half of the filter block). Same results bit for bit as the single-CTA
 * kernel; this switch (returns the previous setting) lets tests and profiles compare the two. */
int iit_conv_pair_enable(int on);
/* The shifted-window kernels store by TMA where an M tile is a rectangle of the output (default on; 0 = always the
 * register-store epilogue). Same results either way; for A/B measurements. Returns the previous setting. */
int iit_conv_win_tma_store(int on);
/* The 7x7 / stride-2 / pad-3 stem on 3 input channels (torchvision resnet18.conv1 under
 * iit/utils/wrapper.py:33-66) as a 4x4 stride-1 convolution over the 2x2 space-to-depth image:
 *   iit_stem_s2d:      x [n,3,h,w] (fp32 or bf16, any strides; h, w even) -> s2d [n,(h+6)/2,(w+6)/2,16] bf16,
 *                      channel (a*2+b)*3+c of s2d[y,x] = padded image [c, 2y+a, 2x+b]; channels 12..15 zero;
 *   iit_stem_filter_s2d: w [64,3,7,7] fp32 (any strides) -> w2 [64,4,4,16] bf16 regrouped the same way;
 *   iit_stem_conv_s2d: y [n,h/2,w/2,64] bf16 = conv(s2d, w2) on the shifted-window tcgen05 kernel, optionally
 *                      adding the per-channel sum / sum of squares of y into bn_sum / bn_sqsum (fp64). */
int iit_stem_s2d(const void* x, int dtype, const int64_t* strides_nchw, void* s2d, int n, int h, int w, void* stream);
int iit_stem_filter_s2d(const float* w, const int64_t* strides_kcrs, void* w2, void* stream);
int iit_stem_conv_s2d(const void* s2d, const void* w2, void* y, int n, int sh, int sw, double* bn_sum,
                      double* bn_sqsum, void* stream);
/*   iit_stem_wgrad_s2d: dw [64,3,7,7] fp32 (strides given, e.g. the parameter's channels-last strides) from
 *                      dy [n,sh-3,sw-3,64] bf16 and the s2d image, on the shifted-window tcgen05 kernel; scratch
 *                      (iit_stem_wgrad_s2d_scratch_bytes, 0 = this image size is not supported: s2d width > 32)
 *                      holds one partial sum per CTA, reduced in fixed order. */
size_t iit_stem_wgrad_s2d_scratch_bytes(int n, int sh, int sw);
int iit_stem_wgrad_s2d(const void* dy, const void* s2d, float* dw, const int64_t* dw_strides_kcrs, int n, int sh,
                       int sw, void* scratch, size_t scratch_bytes, void* stream);
/* dw[k,r,s,c] fp32: pixels are the GEMM K dimension (MN-major operands straight from NHWC), split across
 * CTAs; scratch holds the per-split partial sums (iit_conv2d_wgrad_tc_scratch_bytes), reduced in fixed order. */
size_t iit_conv2d_wgrad_tc_scratch_bytes(const iit_conv_t* d);
int iit_conv2d_wgrad_tc(const void* dy, const void* x, float* dw, const iit_conv_t* d, void* scratch,
                        size_t scratch_bytes, void* stream);
/* The same, adding the result into dw (e.g. the optimizer's gradient buffer) instead of overwriting it. */
int iit_conv2d_wgrad_tc_accumulate(const void* dy, const void* x, float* dw, const iit_conv_t* d, void* scratch,
                        size_t scratch_bytes, void* stream);

/* Stem (7x7 stride-2 pad-3, 3 input channels): im2col matrix col[n*p*q][192] in bf16 (147 columns in
 * (r,s,c) order + zero padding) gathered from an image tensor addressed as [n,c,h,w] through the given
 * element strides (fp32 or bf16, any memory format); the stem then runs as a 1x1 convolution with 192
 * input channels on iit_conv2d_fprop_tc / iit_conv2d_wgrad_tc. */
int iit_stem_im2col(const void* x, int dtype, const int64_t* strides_nchw, void* col, int n, int h, int w,
                    void* stream);

/* filter re-layout: fp32 master filter addressed as [k,c,r,s] through the 4 element strides given
 * (any memory format) -> dense KRSC (to_crsk=0) or CRSK (to_crsk=1) in `dtype`. */
int iit_filter_relayout(const float* w, const int64_t* strides_kcrs, void* w_out, int k, int c, int r,
                        int s, int to_crsk, int dtype, void* stream);
/* The same for many filters in one launch: `table_dev` is an array of `entries` descriptors in DEVICE memory
 * (max_elements = the largest k*c*r*s among them, for the launch geometry). Used to refresh every bf16 filter copy of
 * the ResNet right after an optimizer update; replayable from a CUDA graph because nothing is baked into the launch. */
typedef struct {
  const void* w;      /* fp32 master filter [k,c,r,s], element strides below */
  void* out;          /* dense KRSC (to_crsk = 0) or CRSK (to_crsk = 1) buffer in `dtype` */
  int64_t stride[4];
  int32_t k, c, r, s;
  int32_t to_crsk, dtype;
  int64_t out_pitch;   /* elements between consecutive OUTER indices of the output (k for KRSC, c for CRSK); 0 = dense.
                          Lets several entries fill column blocks of one wider matrix, or leave zero padding per row. */
} iit_relayout_t;
int iit_filter_relayout_batched(const iit_relayout_t* table_dev, int entries, int64_t max_elements, void* stream);

/* BatchNorm2d (training): per-channel statistics over n*h*w rows of an NHWC tensor.
 * sum/sqsum fp64 [c] zeroed by caller. Replaces F.batch_norm's statistics pass. */
int iit_bn_stats(const void* x, int dtype, int64_t rows, int c, double* sum, double* sqsum, void* stream);
/* From sums: mean, invstd (biased var + eps), and running-stat update with momentum and the
 * unbiased variance (torch.nn.BatchNorm2d semantics). running_* may be NULL (no update).
 * num_batches_tracked (may be NULL) is incremented; clear_sums != 0 zeroes sum / sqsum after reading them, so a
 * persistent accumulator is ready for its next use without a fill launch. */
int iit_bn_finalize(double* sum, double* sqsum, int64_t rows, int c, float eps, float momentum,
                    float* mean, float* invstd, float* running_mean, float* running_var,
                    int64_t* num_batches_tracked, int clear_sums, void* stream);
/* y = act((x-mean)*invstd*gamma+beta (+ residual)); act = relu if relu!=0. In eval mode pass
 * mean=running_mean and invstd=1/sqrt(running_var+eps) (iit_bn_eval_invstd).
 * relu_mask (may be NULL): one byte per pack of 8 (bf16) or 4 (fp32) consecutive elements, bit j = element j of y is
 * > 0. The backward reads it instead of y (1/16 of the bytes); rows*c/8 (rows*c/4) bytes. */
int iit_bn_eval_invstd(const float* running_var, int c, float eps, float* invstd, void* stream);
int iit_bn_act_fwd(const void* x, const void* residual, void* y, int dtype, int64_t rows, int c,
                   const float* mean, const float* invstd, const float* gamma, const float* beta,
                   int relu, uint8_t* relu_mask, void* stream);
/* Backward of the above in training mode. Pass 1 reduces dgamma = sum(g*xhat), dbeta = sum(g)
 * with g = dy * [y>0 if relu]; pass 2 writes dx (and dresidual = g when dres != NULL).
 * dgamma/dbeta fp64 [c] zeroed by the caller. relu_mask: the mask iit_bn_act_fwd wrote, or NULL (y is read). */
int iit_bn_act_bwd_reduce(const void* dy, const void* y, const void* x, int dtype, int64_t rows, int c,
                          const float* mean, const float* invstd, int relu,
                          double* dgamma, double* dbeta, const uint8_t* relu_mask, void* stream);
int iit_bn_act_bwd_apply(const void* dy, const void* y, const void* x, void* dx, void* dres, int dtype,
                         int64_t rows, int c, const float* mean, const float* invstd, const float* gamma,
                         const double* dgamma, const double* dbeta, int relu, const uint8_t* relu_mask,
                         void* stream);
/* y = x / scale, element-wise (n a multiple of 4; y may alias x). The forward hook of the stop-gradient pair
 * (`hook_point_out / scale`, iit/model_pairs/stop_grad_pair.py:37-44); its backward is the same call on the gradient. */
int iit_div_scalar(const void* x, void* y, int dtype, int64_t n, float scale, void* stream);
/* iit_bn_act_bwd_apply for a BatchNorm whose input was interchange-patched inside the producing convolution
 * (iit_conv2d_fprop_tc_patch): dx is written as ZERO for rows p in [box6[0], box6[1]), columns q in [box6[2], box6[3]) of every
 * h x w image and channels in [box6[4], box6[5]) -- the CopySlices backward of `out[idx] = cache[idx]`
 * (base_model_pair.py:157-161) without a separate zeroing kernel. Vector path only (16-byte aligned, bf16 c % 8 / fp32 c % 4). */
int iit_bn_act_bwd_apply_box(const void* dy, const void* y, const void* x, void* dx, void* dres, int dtype, int64_t rows, int c,
                             const float* mean, const float* invstd, const float* gamma, const double* dgamma,
                             const double* dbeta, int relu, const uint8_t* relu_mask, int h, int w, const int32_t* box6,
                             void* stream);
int iit_relu_fwd(const void* x, void* y, int dtype, int64_t n, void* stream);
int iit_relu_bwd(const void* dy, const void* y, void* dx, int dtype, int64_t n, void* stream);
int iit_add_relu_fwd(const void* a, const void* b, void* y, int dtype, int64_t n, int relu, void* stream);

/* MaxPool2d(3, stride 2, pad 1), NHWC. idx (int8, window slot) is saved for the backward. */
int iit_maxpool3x3s2_fwd(const void* x, void* y, int8_t* idx, int dtype, int n, int h, int w, int c,
                         void* stream);
int iit_maxpool3x3s2_bwd(const void* dy, const int8_t* idx, void* dx, int dtype, int n, int h, int w,
                         int c, void* stream);
/* The stem tail bn1 -> relu -> maxpool (torchvision resnet.py forward; hook points
 * mod.bn1 / mod.relu / mod.maxpool of iit/utils/wrapper.py:33-66) as one kernel: x is the stem conv
 * output [n,h,w,c], y the pooled [n,(h+1)/2,(w+1)/2,c] tensor, idx the winning window slot per pooled
 * element. Values and window winners are identical to iit_bn_act_fwd(relu) followed by
 * iit_maxpool3x3s2_fwd. xwin (same shape as y; may be NULL when no backward follows) receives the raw
 * value of x at each winner. The backward ADDS the per-channel sums into dgamma / dbeta (the caller zeroes
 * them) from the pooled tensors dy, y, xwin, then writes dx, recomputing the normalised activation from x
 * (one thread per 2 x 2 input block, which only four pooling windows cover); two launches. */
int iit_bn_relu_maxpool_fwd(const void* x, void* y, int8_t* idx, void* xwin, int dtype, int n, int h, int w,
                            int c, const float* mean, const float* invstd, const float* gamma,
                            const float* beta, void* stream);
int iit_bn_relu_maxpool_bwd(const void* dy, const int8_t* idx, const void* x, const void* y, const void* xwin,
                            void* dx, int dtype, int n, int h, int w, int c, const float* mean,
                            const float* invstd, const float* gamma, const float* beta, double* dgamma,
                            double* dbeta, void* stream);
/* BatchNorm backward, last step: sums = fp64 [2][c] (dgamma sums, then dbeta sums) as accumulated by
 * iit_bn_act_bwd_reduce / iit_bn_relu_maxpool_bwd. dgamma[i] (+)= (float)sums[i], dbeta[i] (+)= (float)sums[c+i]
 * (accumulate != 0 adds into the existing values, e.g. the optimizer's gradient buffer); clear_sums != 0 zeroes
 * sums afterwards so that a persistent accumulator is ready for the next backward. */
int iit_bn_param_grads(double* sums, int c, float* dgamma, float* dbeta, int accumulate, int clear_sums,
                       void* stream);
/* AdaptiveAvgPool2d(1) + flatten: [n,hw,c] -> fp32 [n,c]; and its backward. */
int iit_avgpool_fwd(const void* x, float* y, int dtype, int n, int hw, int c, void* stream);
int iit_avgpool_bwd(const float* dy, void* dx, int dtype, int n, int hw, int c, void* stream);
/* Linear: y[n,o] = x[n,i] @ w[o,i]^T + b (fp32), and its three gradients. */
int iit_linear_fwd(const float* x, const float* w, const float* b, float* y, int n, int in, int out,
                   void* stream);
int iit_linear_bwd(const float* x, const float* w, const float* dy, float* dx, float* dw, float* db,
                   int n, int in, int out, void* stream);

/* kl[r] = sum_c B[r,c] * (log B[r,c] - log A[r,c]) with A / B = the row itself (x_is_pmf != 0) or softmax(row);
 * fp32 rows with the given strides. The metric of the resample-ablation evaluation
 * (iit/utils/eval_metrics.py:4-20, iit/utils/eval_ablations.py:59-61). */
int iit_kl_div_rows(const float* a, const float* b, int64_t rows, int64_t classes, int64_t a_row_stride,
                    int64_t b_row_stride, int a_is_pmf, int b_is_pmf, float* kl, void* stream);

/* ------------------------------------------------------------------------------------------ */
/* Either side of the step: batches of pairs and counterfactual inputs.  */
/* err_flag: device int the kernels set to non-zero on an out-of-range index (the wrapper checks it). */
/* ------------------------------------------------------------------------------------------ */
/* dst[b, :] = src[idx[b], :], rows of row_bytes bytes: the collate step of IITDataset
 * (iit/utils/iit_dataset.py:24-59) over a device-resident dataset. */
int iit_gather_rows(void* dst, const void* src, const int64_t* idx, int64_t rows, int64_t row_bytes,
                    int64_t src_rows, int* err_flag, void* stream);
/* dst[b, :, r0:r0+h, c0:c0+w] = bank[idx[b]] (bank: [items, c, h, w] dense; dst: given NCHW strides): the
 * quadrant swap of MNIST_PVR_Dataset.patch_at_hl_idx (iit/tasks/mnist_pvr/dataset.py:125-155). */
int iit_gather_box(void* dst, const int64_t* dst_strides_nchw, const void* bank, const int64_t* idx, int dtype,
                   int n, int c, int h, int w, int r0, int c0, int64_t bank_items, int* err_flag, void* stream);
/* For each sample, a uniform draw among the bank items whose label differs from cur_class[b] (the bank is sorted
 * by label; class_start has n_classes + 1 entries): the rejection loop of dataset.py:137-149 in one pass.
 * Counter-based generator (splitmix64 of seed and b): reproducible, not the reference's numpy stream. */
int iit_pick_other_class(const int64_t* cur_class, const int64_t* class_start, int n_classes, int64_t bank_items,
                         uint64_t seed, int64_t n, int64_t* out_idx, int64_t* out_label, int* err_flag,
                         void* stream);

/* ------------------------------------------------------------------------------------------ */
/* (b) low-level HookedTransformer (train_ioi.py:30-34; transformer_lens 1.12.1 semantics):        */
/* everything but the dense contractions, which run on the convolution GEMMs above (a linear layer  */
/* over tokens is a 1x1 convolution with n = tokens, h = w = 1).                                    */
/* ------------------------------------------------------------------------------------------ */
/* resid[t,:] = W_E[tokens[t],:] + W_pos[t % seq,:]  (embed + pos_embed hooks' sum) and its scatter-add
 * backward into fp32 dW_E / dW_pos (accumulated; caller zeroes). */
int iit_embed_fwd(const int64_t* tokens, const float* W_E, const float* W_pos, void* out, int dtype,
                  int64_t ntok, int seq, int d_model, int64_t d_vocab, void* stream);
int iit_embed_bwd(const int64_t* tokens, const void* dresid, int dtype, float* dW_E, float* dW_pos,
                  int64_t ntok, int seq, int d_model, int64_t d_vocab, void* stream);
/* LayerNormPre (folded layer norm): y = (x - mean) / sqrt(var + eps); rstd saved for the backward. */
int iit_lnpre_fwd(const void* x, void* y, float* rstd, int dtype, int64_t rows, int d, float eps, void* stream);
int iit_lnpre_bwd(const void* dy, const void* y, const float* rstd, void* dx, int dtype, int64_t rows, int d,
                  void* stream);
/* dx += the same: the residual stream feeds the LayerNorm and the skip connection, and the LayerNorm's input gradient is added
 * onto the skip path's (which is already there) instead of being summed by a separate element-wise kernel. */
int iit_lnpre_bwd_accumulate(const void* dy, const void* y, const float* rstd, void* dx, int dtype, int64_t rows, int d,
                             void* stream);
/* pre = x + bias (+ resid); out = act(pre), act 0: identity, 1: gelu_new. pre_out optional (hook_pre). */
int iit_bias_act_fwd(const void* x, const float* bias, const void* resid, void* pre_out, void* out, int dtype,
                     int64_t rows, int n, int act, void* stream);
int iit_gelu_bwd(const void* dpost, const void* pre, void* dpre, int dtype, int64_t n, void* stream);
/* The MLP activation's backward with the input bias folded in and its gradient taken in the same pass (bf16, n % 8 == 0):
 * dpre[r,c] = dpost[r,c] * gelu_new'(h[r,c] + bias[c]); colsum[c] += sum_r dpre[r,c] (fp32, caller zeroes; may be NULL).
 * `h` is the raw result of the MLP's input map: the forward keeps it instead of the pre-activation h + bias. */
int iit_gelu_bwd_bias_colsum(const void* dpost, const void* h, const float* bias, void* dpre, float* colsum, int64_t rows,
                             int64_t n, void* stream);
/* out[c] += sum_r x[r,c] (bias gradients; fp32, caller zeroes) */
int iit_colsum(const void* x, int dtype, float* out, int64_t rows, int64_t n, void* stream);
/* Causal attention core on the fused projection qkv [B,S,3,H,DH] -> z [B,S,H,DH] (blocks.i.attn.hook_z),
 * scores / sqrt(DH), softmax over keys <= query. Backward recomputes the pattern. */
int iit_attn_fwd(const void* qkv, void* z, int dtype, int64_t batch, int seq, int heads, int d_head, void* stream);
int iit_attn_bwd(const void* qkv, const void* dz, void* dqkv, int dtype, int64_t batch, int seq, int heads,
                 int d_head, void* stream);

/* ------------------------------------------------------------------------------------------ */
/* optimiser step pieces (iit_behavior_model_pair.py:60-64, base_model_pair.py:176-180,229)     */
/* ------------------------------------------------------------------------------------------ */
/* sqnorm[0] += sum(g^2) over a flat fp32 buffer (clip_grad_norm_'s total norm) */
int iit_sqnorm(const float* g, int64_t n, double* sqnorm, void* stream);
/* Adam (torch.optim.Adam defaults: no weight decay, no amsgrad) on a flat fp32 buffer with the
 * clip factor min(1, max_norm/(sqrt(*sqnorm)+1e-6)) applied to g on the fly (max_norm<=0: none).
 * step is the 1-based step count (host value). */
int iit_adam_step(float* p, const float* g, float* m, float* v, int64_t n, double lr, double beta1,
                  double beta2, double eps, int64_t step, const double* sqnorm, double max_norm,
                  void* stream);

/* Same update with the 1-based step count and the learning rate read from device memory (the call
 * first increments *step_dev): no schedule value is baked into the launch, so it can be replayed from a
 * captured CUDA graph. */
int iit_adam_step_dev(float* p, const float* g, float* m, float* v, int64_t n, const float* lr_dev,
                      double beta1, double beta2, double eps, long long* step_dev, const double* sqnorm,
                      double max_norm, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* IIT_B200_H */
