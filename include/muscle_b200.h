/* muscle_b200 — C ABI of the B200-native contrastive-loss hot path of SCoulY/MuSCLe.
 *
 * The reference is pure Python/PyTorch and has no FFI layer: its boundary for this path is the
 * set of Python call expressions in train_mcl.py / train_muscle.py / src/MuSCLe.py (SURVEY.md
 * §8b).  These entry points are what a binding for that boundary calls; INTEGRATION.md shows the
 * ctypes stubs.  Conventions (all functions):
 *   - plain device pointers + explicit sizes; the caller owns every buffer (inputs, outputs,
 *     workspace); nothing is allocated, freed or synchronised inside; all work is enqueued on
 *     `stream` (a cudaStream_t passed as void*); re-entrant, no global state;
 *   - return value 0 = MUSCLE_OK, otherwise a MUSCLE_ERR_* code; muscle_last_error() returns a
 *     thread-local description of the most recent failure;
 *   - float tensors are fp32, row-major/contiguous in the stated layout; nothing is mutated
 *     except the documented outputs and the workspace.
 */
#ifndef MUSCLE_B200_H_
#define MUSCLE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MUSCLE_OK 0
#define MUSCLE_ERR_BAD_ARG 1       /* unsupported shape / null pointer / misaligned buffer */
#define MUSCLE_ERR_WORKSPACE 2     /* workspace too small */
#define MUSCLE_ERR_CUDA 3          /* a CUDA runtime / driver call failed */
#define MUSCLE_ERR_UNSUPPORTED 4   /* device is not sm_100 */

/* Layout of a [pixels, dim] feature operand. */
#define MUSCLE_LAYOUT_PIXEL_MAJOR 0 /* x[n*dim + d]            (N, D)            */
#define MUSCLE_LAYOUT_NCHW 1        /* x[(b*dim + d)*P + p]    (B, D, h, w), P=h*w */

#define MUSCLE_MAX_CLASSES 31

const char* muscle_last_error(void);
int muscle_version(void);

/* ------------------------------------------------------------------------------------------
 * Dense label-masked pixel InfoNCE, forward + input gradient in one call.
 * Replaces: the composition of reference `info_nce` (src/loss_multilabel.py:9-21) over a whole
 * batch fixed in SURVEY.md §8c — for every valid pixel q of class c:
 *     pos_q = e_q . mean_{k: cls_k = c} e_k ,  negs_q = { e_q . e_k : cls_k != c } ,
 *     loss  = mean_q [ logsumexp([pos_q, negs_q] / T) - pos_q / T ],  e = x / max(||x||, eps)
 * x        [B*P, dim] in `layout`, raw (un-normalised) embeddings
 * cls      per-pixel class id in [0, num_classes), < 0 = pixel takes no part; int32 or int64
 * loss_out device float[1];  dx_out device float[B*P*dim] in `layout`, or NULL for forward only
 * dim in {64,128,256}; num_classes <= MUSCLE_MAX_CLASSES.
 */
size_t muscle_dense_infonce_workspace_bytes(int64_t n_pixels, int dim, int num_classes);
int muscle_dense_infonce_fwd_bwd(const float* x, int layout, int64_t batch,
                                 int64_t pixels_per_image, int dim, const void* cls,
                                 int cls_is_int64, int num_classes, float temperature, float eps,
                                 float* loss_out, float* dx_out, void* workspace,
                                 size_t workspace_bytes, void* stream);

/* The same op as an autograd pair (what torch.autograd.Function.forward / .backward call): _fwd runs
 * everything up to the loss and leaves its state in `workspace`; _bwd, given the SAME untouched
 * workspace and the same shape arguments, runs the gradient pass and writes
 * dx_out = grad_scale[0] * d(loss)/dx  (grad_scale: device float[1], NULL = 1), so the upstream
 * gradient costs no extra pass over dx. */
int muscle_dense_infonce_fwd(const float* x, int layout, int64_t batch, int64_t pixels_per_image,
                             int dim, const void* cls, int cls_is_int64, int num_classes,
                             float temperature, float eps, float* loss_out, void* workspace,
                             size_t workspace_bytes, void* stream);
int muscle_dense_infonce_bwd(int layout, int64_t batch, int64_t pixels_per_image, int dim,
                             int num_classes, float temperature, const float* grad_scale,
                             float* dx_out, void* workspace, size_t workspace_bytes, void* stream);

/* The same pair with the class prototypes HANDED IN (proto_in device float [num_classes, dim], or NULL for the
 * class means): CAM-weighted prototypes from muscle_pool_prototypes_fwd take the place of the hard class mean
 * as every pixel's positive.  The backward then returns d loss / d prototype in dproto_out (times grad_scale)
 * for the pooling op's backward; with proto_external != 0 no gradient flows to the pixels through the
 * prototypes inside this op (the pooling op owns that path). */
int muscle_dense_infonce_proto_fwd(const float* x, int layout, int64_t batch, int64_t pixels_per_image,
                                   int dim, const void* cls, int cls_is_int64, int num_classes,
                                   float temperature, float eps, const float* proto_in, float* loss_out,
                                   void* workspace, size_t workspace_bytes, void* stream);
int muscle_dense_infonce_proto_bwd(int layout, int64_t batch, int64_t pixels_per_image, int dim,
                                   int num_classes, float temperature, const float* grad_scale,
                                   int proto_external, float* dx_out, float* dproto_out, void* workspace,
                                   size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * Staged form of the same op, for the batch-sharded multi-GPU path: the host
 * runs  prepare -> [all-gather ehat/bits/tor, all-reduce class_sum/class_mass] -> plan ->
 * rowsum -> mid -> [all-gather w, all-reduce dproto/loss_acc] -> grad -> final.
 * "rows" are this rank's pixels, "cols" all pixels of all ranks; *_pad = round_up(n, 256).
 * Inside the library the pixels are kept SORTED BY CLASS (stable counting sort; invalid pixels
 * and zero pad rows last): arrays marked [sorted] are indexed by sorted row, `perm` maps a
 * source pixel to its sorted row.  `tor` has one word per 128 sorted rows: the OR of the class
 * bits in the tile, bit 31 set when the tile holds an invalid or pad row; the pair passes skip
 * every (row tile, column tile) whose pairs are all masked and drop the mask where none is.
 */
int64_t muscle_pad_rows(int64_t n);
size_t muscle_infonce_prepare_scratch_bytes(int64_t n_pixels);

/* Class sort + L2-normalise + bf16 cast + per-class sums (accumulated with atomics: the caller
 * zeroes class_sum [C*dim] and class_mass [C] first).  Pad rows are zeroed.
 * ehat: bf16 [rows_pad, dim] [sorted]; bits: uint32 [rows_pad] [sorted] (1<<class, 0 = invalid);
 * inv_norm: float [rows_pad] [sorted]; perm: int32 [n]; tor: uint32 [rows_pad/128];
 * seg: uint32 [num_classes+2] first sorted row of every class, of the invalid pixels, and n. */
int muscle_infonce_prepare(const float* x, int layout, int64_t batch, int64_t pixels_per_image,
                           int dim, const void* cls, int cls_is_int64, int num_classes, float eps,
                           void* ehat, uint32_t* bits, float* inv_norm, int32_t* perm,
                           uint32_t* tor, uint32_t* seg, float* class_sum, float* class_mass,
                           void* scratch, size_t scratch_bytes, void* stream);

/* Work partition of the two pair passes over rows x cols: unit_off [rows_pad/128 + 1] = running count of the LIVE (row block, column tile) units, i.e. those
 * with at least one unmasked pair.  Every CTA of a pair pass takes an equal contiguous share of
 * that range, so the passes stay balanced whatever the class mix. */
int muscle_infonce_plan(const uint32_t* tor_rows, int64_t rows_pad, const uint32_t* tor_cols,
                        int64_t cols_pad, int dim, uint32_t* unit_off, void* stream);

/* rowsum[q] += sum_k [cls_k != cls_q] exp((e_q.e_k - 1)/T)   — tcgen05 pass 1 (caller zeroes
 * rowsum first). */
int muscle_infonce_rowsum(const void* ehat_rows, const uint32_t* bits_rows,
                          const uint32_t* tor_rows, int64_t rows_pad, const void* ehat_cols,
                          const uint32_t* bits_cols, const uint32_t* tor_cols, int64_t cols_pad,
                          const uint32_t* unit_off, int dim, float temperature, float* rowsum,
                          void* stream);

/* The row-sum pass when rows and columns are THE SAME pixels (single GPU; dim 64 or 128): the
 * similarity matrix and the label mask are symmetric, so only the live units with column tile >= row
 * tile are visited and every off-diagonal tile feeds the row sums of both its row pixels and its
 * column pixels (half the products and half the exponentials of muscle_infonce_rowsum; same result up
 * to fp32 summation order).  muscle_infonce_plan_sym numbers those units: unit_off_sym [rows_pad/128 + 1].
 * No reference counterpart: the reference materialises the logits (src/loss_multilabel.py:15-21). */
int muscle_infonce_plan_sym(const uint32_t* tor, int64_t rows_pad, uint32_t* unit_off_sym, void* stream);
int muscle_infonce_rowsum_sym(const void* ehat, const uint32_t* bits, const uint32_t* tor,
                              int64_t rows_pad, const uint32_t* unit_off_sym, int dim,
                              float temperature, float* rowsum, void* stream);
/* The symmetric pass over `world` ranks: the global similarity matrix is symmetric too, so every
 * unordered tile pair is computed by ONE rank.  The ranks' pixel blocks form a ring: a rank takes its own block's
 * pairs with column tile >= row tile, all pairs with the next ceil((world-1)/2) blocks and (even world) half of
 * the pairs with the opposite block.  rowsum_all [cols_pad] (zeroed by the caller) receives the rank's own row
 * sums at rank * rows_pad AND its column-sum contributions to the other ranks' pixels: ONE all-reduce (sum) of
 * rowsum_all over the ranks completes every pixel's row sum (cols_pad floats; 1 MiB at BASELINE config 4).
 * cols_pad == world * rows_pad; scratch: muscle_infonce_plan_sym_ring_scratch_bytes. */
size_t muscle_infonce_plan_sym_ring_scratch_bytes(int64_t rows_pad);
int muscle_infonce_plan_sym_ring(const uint32_t* tor_rows, int64_t rows_pad, const uint32_t* tor_cols,
                                 int64_t cols_pad, int rank, int world, void* scratch,
                                 uint32_t* unit_off_sym, void* stream);
/* muscle_infonce_plan_sym_ring without caller scratch and WITHOUT ANY ALLOCATION.  No counterpart in the reference:
 * added for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  The per-row-tile counts live in a file-scope
 * `static __device__` array of the library (64 KB, held by every GPU that loads the library until exit), so
 * rows_pad is limited to muscle_infonce_plan_sym_ring_static_max_rows() = 2 097 152 rows; larger sizes are refused
 * with MUSCLE_ERR_BAD_ARG (use the variant above).  Same inputs, same unit_off_sym (bit-identical).  The entry
 * point calls no cudaMalloc / cudaFree / allocating library routine.  ALL calls on one GPU share the one array: it
 * keeps the last call's contents (every entry a call reads is overwritten by that call first, nothing needs
 * clearing), and two calls that overlap on the GPU (different streams, no ordering between them) RACE on it --
 * order them, or give one of them caller-owned scratch through muscle_infonce_plan_sym_ring. */
int muscle_infonce_plan_sym_ring_static_max_rows(void);
int muscle_infonce_plan_sym_ring_static(const uint32_t* tor_rows, int64_t rows_pad, const uint32_t* tor_cols,
                                        int64_t cols_pad, int rank, int world, uint32_t* unit_off_sym,
                                        void* stream);
int muscle_infonce_rowsum_sym_ring(const void* ehat_rows, const uint32_t* bits_rows,
                                   const uint32_t* tor_rows, int64_t rows_pad, const void* ehat_cols,
                                   const uint32_t* bits_cols, const uint32_t* tor_cols, int64_t cols_pad,
                                   const uint32_t* unit_off_sym, int rank, int world, int dim,
                                   float temperature, float* rowsum_all, void* stream);
/* The same exchange with the work split by COUNT instead of by block: every rank numbers the live unordered tile
 * pairs of the whole gathered array identically (muscle_infonce_plan_sym_ring over the gathered arrays as rows AND
 * columns, rank 0 of world 1: the gathered tiles are class-sorted per rank block only, which that plan does not
 * mind) and computes share `share` of `n_shares` equal contiguous parts of that list.  Whatever the class mix of a
 * rank's own images, all ranks finish the pass together (at 8 ranks the ring split left the pass 1.17-1.38 ms
 * apart).  rowsum_all as above: zeroed by the caller, completed by one all-reduce. */
int muscle_infonce_rowsum_sym_share(const void* ehat_all, const uint32_t* bits_all, const uint32_t* tor_all,
                                    int64_t cols_pad, const uint32_t* unit_off_sym, int share, int n_shares,
                                    int dim, float temperature, float* rowsum_all, void* stream);

/* pos[q] = e_q . proto_{c_q} / T; rowsum[q] += exp(pos - 1/T); w[q] = 1/rowsum[q];
 * gpos[q] = (exp(pos - 1/T) w - 1)/T; loss_acc (double[1]) += sum_q (1/T + ln rowsum[q] - pos);
 * dproto [C*dim] += gpos[q] e_q  (accumulated: caller zeroes loss_acc and dproto).
 * Rows outside every class get w = gpos = 0.  class_sum / class_mass are the GLOBAL sums. */
int muscle_infonce_mid(const void* ehat, const uint32_t* seg, int64_t rows_pad, int dim,
                       int num_classes, const float* class_sum, const float* class_mass,
                       float* rowsum, float temperature, float* w, float* gpos, double* loss_acc,
                       float* dproto, void* stream);

/* part[s][q][:] = sum_{k in share s} [cls_k != cls_q] exp((e_q.e_k - 1)/T) (w_rows[q] +
 * w_cols[k]) e_k   — tcgen05 pass 2.  part: float [muscle_infonce_part_slots(), rows_pad, dim];
 * slot s of a row holds the share of the s-th CTA that worked on the row's block (see plan);
 * slots no CTA wrote are never read. */
int muscle_infonce_part_slots(void);
int muscle_infonce_grad(const void* ehat_rows, const uint32_t* bits_rows,
                        const uint32_t* tor_rows, const float* w_rows, int64_t rows_pad,
                        const void* ehat_cols, const uint32_t* bits_cols,
                        const uint32_t* tor_cols, const float* w_cols, int64_t cols_pad,
                        const uint32_t* unit_off, int dim, float temperature, float* part,
                        void* stream);

/* dx = d(loss)/dx in the caller's pixel order and `layout`: sums the slots, adds the prototype
 * terms, back-propagates through the L2 normalisation.  loss_out = loss_acc / n_valid
 * (n_valid = sum of class_mass).  dx may be NULL (loss only); loss_out may be NULL (gradient only).
 * grad_scale: device pointer to the upstream gradient of the (scalar) loss, multiplied into dx
 * inside the kernel — what autograd hands to backward(); NULL = 1. */
int muscle_infonce_final(const void* ehat, const uint32_t* bits, const float* inv_norm,
                         const int32_t* perm, const float* part, const uint32_t* unit_off,
                         int64_t rows_pad, int64_t cols_pad, int64_t batch, int64_t pixels_per_image, int dim, int num_classes,
                         const float* class_sum, const float* class_mass, const float* gpos,
                         const float* dproto, const double* loss_acc, float temperature,
                         int layout, float* dx, float* loss_out, const float* grad_scale, void* stream);

/* ------------------------------------------------------------------------------------------
 * info_nce — replaces src/loss_multilabel.py:9-21 with distinct operand sets: query [Q, dim],
 * positive_keys [P, dim], negative_keys [K, dim], fp32 row-major, unit-norm rows (like the reference
 * the function does not normalise; the tensor-pipe passes assume |q.k| <= 1).  loss_out float[1]
 * = mean (reduction_mean != 0) or sum over queries; row_loss_out [Q] (reduction 'none') or NULL;
 * dquery_out [Q,dim], dpositive_out [P,dim], dnegative_out [K,dim]: gradients of loss_out, each
 * may be NULL.  dim in {64,128,256}. */
size_t muscle_info_nce_workspace_bytes(int64_t n_query, int64_t n_negative, int dim);
int muscle_info_nce_fwd_bwd(const float* query, const float* positive_keys,
                            const float* negative_keys, int64_t n_query, int64_t n_positive,
                            int64_t n_negative, int dim, float temperature, int reduction_mean,
                            float* loss_out, float* row_loss_out, float* dquery_out,
                            float* dpositive_out, float* dnegative_out, void* workspace,
                            size_t workspace_bytes, void* stream);
/* The same for rows of ANY norm (the reference, src/loss_multilabel.py:9-21, never normalises) and with an upstream
 * gradient per row (the backward of reduction='none').  query_norm_bound >= max_q |q|, key_norm_bound >= max(max_k |k|,
 * |mean_p positive_keys[p]|): the passes then run on q / query_norm_bound and k / key_norm_bound at temperature / (the
 * product of the bounds), which must stay >= MUSCLE_MIN_TEMPERATURE; results are those of the unscaled problem.
 * muscle_info_nce_fwd_bwd is the case of both bounds = 1 (rows shorter than 1 keep their length).
 * row_weight [Q] or NULL: the three gradients become those of sum_q row_weight[q] * loss_q (times 1/Q when
 * reduction_mean); loss_out stays the unweighted value and row_loss_out must be NULL. */
#define MUSCLE_MIN_TEMPERATURE 0.0125f
int muscle_info_nce_fwd_bwd_ex(const float* query, const float* positive_keys,
                               const float* negative_keys, int64_t n_query, int64_t n_positive,
                               int64_t n_negative, int dim, float temperature, int reduction_mean,
                               float query_norm_bound, float key_norm_bound, const float* row_weight,
                               float* loss_out, float* row_loss_out, float* dquery_out,
                               float* dpositive_out, float* dnegative_out, void* workspace,
                               size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * CAM-weighted (soft) prototype pooling and its gradient scatter (north_star: "CAM-weighted pooling of
 * backbone features into class and region prototypes ... and gradient-scatter kernels").  Instances in the
 * reference: the positive-key mean of src/loss_multilabel.py:12 (0/1 weights), the region mean of
 * src/loss_multilabel.py:255, the CAM aggregation `cam @ aff` of src/MuSCLe.py:222 (w = cam^T, `sum` output).
 *   fwd: sum[c][:] = sum_p w[p][c] e[p][:];  mass[c] = sum_p w[p][c];  proto = sum / max(mass, eps) (proto may be NULL)
 *   bwd: given dproto [C, dim]:  de[p][:] = sum_c w[p][c] dproto[c][:] / mass[c]
 *                                dw[p][c] = (e[p] . dproto[c] - proto[c] . dproto[c]) / mass[c]     (either may be NULL)
 * e [n, dim] pixel-major, bf16 (e_is_bf16 != 0: e.g. the class-sorted `ehat` of the InfoNCE path) or fp32;
 * w [n, num_classes] pixel-major fp32; dim % 4 == 0, dim <= 256; num_classes <= 24.  de is fp32 [n, dim]. */
int muscle_pool_prototypes_fwd(const void* e, int e_is_bf16, const float* w, int64_t n, int dim,
                               int num_classes, float eps, float* sum, float* mass, float* proto,
                               void* stream);
/* The forward with caller scratch (muscle_pool_prototypes_fwd_scratch_bytes(dim, num_classes) bytes, 16-byte aligned):
 * for bf16 e with dim in {64,128,256} (128-byte aligned) the product runs on the tensor pipe (csrc/pool_tc.cuh: weights
 * split into bf16 hi + lo parts, fp32 accumulation in TMEM), every CTA stores its part into the scratch and a second
 * kernel adds the parts in a fixed order: no atomics, bit-reproducible, two launches.  Other inputs (fp32 e, other
 * dims) and scratch == NULL take muscle_pool_prototypes_fwd's path; results agree to fp32 rounding. */
size_t muscle_pool_prototypes_fwd_scratch_bytes(int dim, int num_classes);
int muscle_pool_prototypes_fwd_ws(const void* e, int e_is_bf16, const float* w, int64_t n, int dim,
                                  int num_classes, float eps, float* sum, float* mass, float* proto,
                                  void* scratch, size_t scratch_bytes, void* stream);
int muscle_pool_prototypes_bwd(const void* e, int e_is_bf16, const float* w, int64_t n, int dim,
                               int num_classes, float eps, const float* proto, const float* mass,
                               const float* dproto, float* de, float* dw, void* stream);

/* ==========================================================================================
 * The other members of the path.  Same conventions as above.
 * ========================================================================================== */

/* ------------------------------------------------------------------------------------------
 * image_level_contrast — replaces src/loss_multilabel.py:36-66 (called at train_mcl.py:193).
 * emb [batch, dim] fp32, label [batch, n_labels] fp32 multi-hot.  loss_out float[1];
 * n_used_out int32[1] = number of images that contributed (0: the reference returns python 0);
 * demb_out [batch, dim] = d loss / d emb, or NULL. */
size_t muscle_imc_workspace_bytes(int batch, int dim);
int muscle_imc_fwd_bwd(const float* emb, const float* label, int batch, int dim, int n_labels,
                       float temperature, float* loss_out, int32_t* n_used_out, float* demb_out,
                       void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * Classification losses of the MCL stage — replace FocalLoss (src/loss_multilabel.py:68-91),
 * Log_Sum_Exp_Pairwise_Loss (:24-33) and nn.MultiLabelSoftMarginLoss as composed at train_mcl.py:180-183:
 *   loss_cls = focal(sigmoid(l)) + softmargin(l) + mean_b LSEP(sigmoid(l)),  l = logits[:, offset : offset+num_fg]
 * in ONE launch.  losses_out float[4] = {loss_cls, focal, soft margin, pair}; dlogits_out [batch, logit_stride] =
 * grad_scale[0] * d loss_cls / d logits (columns outside the class range zeroed), or NULL.  label [batch, num_fg].
 * muscle_focal_loss_fwd_bwd / muscle_lsep_fwd_bwd are the two reference callables on their own inputs
 * (probabilities): loss_out4[1] = focal; rows_out [batch] = the LSEP values, dpred = sum_b row_grad[b] d row_b. */
int muscle_cls_losses_fwd_bwd(const float* logits, const float* label, int batch, int num_fg,
                              int logit_stride, int logit_offset, float gamma, float alpha,
                              const float* grad_scale, float* losses_out, float* dlogits_out, void* stream);
int muscle_focal_loss_fwd_bwd(const float* prob, const float* target, int batch, int num_classes,
                              float gamma, float alpha, const float* grad_scale, float* loss_out4,
                              float* dprob_out, void* stream);
int muscle_lsep_fwd_bwd(const float* pred, const float* labels, int batch, int num_classes,
                        const float* row_grad, float* rows_out, float* dpred_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * PixPro — replaces src/loss_multilabel.py:93-105 (train_mcl.py:209).  fm1, fm2 [batch, 21,
 * height, width] NCHW; box1, box2 device int32 [batch,4] = (top, left, h, w), equal h,w per image,
 * inside the map; max_box_pixels = max_b h*w (grid sizing).  loss_out float[1]; dfm1_out (same
 * shape as fm1, fully written) or NULL; view 2 receives no gradient (:103).  scratch: double[1]. */
int muscle_pixpro_fwd_bwd(const float* fm1, const float* fm2, const int32_t* box1,
                          const int32_t* box2, int batch, int channels, int height, int width,
                          int max_box_pixels, float* loss_out, float* dfm1_out, double* scratch,
                          void* stream);

/* ------------------------------------------------------------------------------------------
 * CAM head — replaces src/MuSCLe.py:242-247 (== :263-268): cam = relu(sum_ch p7 * W), W =
 * fc.weight.data (no gradient to W).  p7 [batch, channels, pixels]; fc_weight [num_classes,
 * channels]; cam [batch, num_classes, pixels].  bwd: dp7 = W^T (dcam [cam > 0]) (+ demb/pixels when
 * demb != NULL — the GAP backward of :240, so dp7 is written once; dcam may be NULL). */
int muscle_cam_head_fwd(const float* p7, const float* fc_weight, int batch, int channels,
                        int pixels, int num_classes, float* cam_out, void* stream);
int muscle_cam_head_bwd(const float* dcam, const float* cam, const float* fc_weight,
                        const float* demb, int batch, int channels, int pixels, int num_classes,
                        float* dp7_out, void* stream);
/* Image embedding + logits — replaces src/MuSCLe.py:182-183,240-241: emb = mean over pixels,
 * logits = emb W^T (gradient reaches W here). */
int muscle_gap_fwd(const float* p7, int batch, int channels, int pixels, float* emb_out,
                   void* stream);
int muscle_linear_fwd(const float* emb, const float* weight, int batch, int channels,
                      int num_classes, float* logits_out, void* stream);
int muscle_linear_bwd(const float* dlogits, const float* emb, const float* weight, int batch,
                      int channels, int num_classes, float* demb_out, float* dweight_out,
                      void* stream);

/* ------------------------------------------------------------------------------------------
 * Projection head + PCM — replaces src/MuSCLe.py:181 (`fuse`, 1x1 conv Cin -> 128 + bias) and
 * src/MuSCLe.py:213-223 (`PCM`): ghat = g/(||g||+1e-5); S = relu(ghat^T ghat); out = cam S /
 * (colsum S + 1e-5).  cam [batch, num_classes, pixels], f [batch, cin, pixels] (detached in the
 * reference, :254), fuse_w [128, cin], fuse_b [128], out [batch, num_classes, pixels]; fp32.
 * `state` (muscle_pcm_state_bytes) is written by the forward and read by the backward: bf16 copies
 * of f and ghat, ||g||, the column sums.  Backward: dcam (may be NULL), dfuse_w [128,cin] and
 * dfuse_b [128] (both or neither); `scratch` (muscle_pcm_bwd_scratch_bytes) is free afterwards.
 * num_classes <= 24, cin <= 383.  The P x P affinity never reaches HBM. */
size_t muscle_pcm_state_bytes(int batch, int cin, int64_t pixels);
size_t muscle_pcm_bwd_scratch_bytes(int batch, int cin, int64_t pixels);
int muscle_pcm_fwd(const float* cam, const float* f, const float* fuse_w, const float* fuse_b,
                   int batch, int num_classes, int cin, int64_t pixels, float* out, void* state,
                   size_t state_bytes, void* stream);
int muscle_pcm_bwd(const float* dout, const float* cam, const float* out, int batch,
                   int num_classes, int cin, int64_t pixels, const void* state,
                   size_t state_bytes, float* dcam, float* dfuse_w, float* dfuse_b,
                   void* scratch, size_t scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * The trainers' CAM normalisers — replace train_mcl.py:30-36 (cam_softmaxnorm: softmax over the
 * fg channels, bg = 1 - max fg) and train_mcl.py:21-28 (cam_maxnorm: relu, per-plane min-max with
 * -1e-6 / +1e-6, relu).  cam [batch, num_classes, pixels]; planes = batch*num_classes;
 * lohi float[planes*2], argmm int64[planes*2] are saved by the forward for the backward. */
int muscle_cam_softmaxnorm_fwd(const float* cam, int batch, int num_classes, int64_t pixels,
                               float* out, void* stream);
int muscle_cam_softmaxnorm_bwd(const float* out, const float* dout, int batch, int num_classes,
                               int64_t pixels, float* din, void* stream);
int muscle_cam_maxnorm_fwd(const float* cam, int64_t planes, int64_t pixels, float* out,
                           float* lohi, int64_t* argmm, void* stream);
int muscle_cam_maxnorm_bwd(const float* cam, const float* dout, const float* lohi,
                           const int64_t* argmm, int64_t planes, int64_t pixels, float* din,
                           void* stream);

/* ------------------------------------------------------------------------------------------
 * SURVEY.md 8f rank 1 — bilinear re-sampling with align_corners=True (F.interpolate as called at
 * src/MuSCLe.py:256-257 and :285) and the ER loss of train_mcl.py:175-188.
 * x [planes, height, width] -> y [planes, out_height, out_width]; the backward is the exact adjoint (separable, two
 * launches; scratch: muscle_upsample_bilinear_bwd_scratch_bytes).
 * ER loss: raw_cams, raw_sgcs [batch, 21, height, width] (the model's outputs at full resolution), label [batch, 20]:
 *   cams = cam_softmaxnorm(raw_cams) (no gradient), sgcs = cam_softmaxnorm(raw_sgcs), both times the label mask
 *   (background on); per image the k = int(0.2 * int(label.sum()) * H * W) largest |cams - sgcs|; loss = their mean over
 *   all batch * k values.  Forward: six launches, no host synchronisation (the reference reads valid_channel back and
 *   sorts 21 H W values per image); `state` (muscle_er_loss_state_bytes) carries the per-image thresholds to the
 *   backward, which writes dsgcs_out [batch, 21, H, W] = grad_scale[0] * d loss / d raw_sgcs. */
int muscle_upsample_bilinear_fwd(const float* x, int64_t planes, int height, int width, int out_height,
                                 int out_width, float* y, void* stream);
size_t muscle_upsample_bilinear_bwd_scratch_bytes(int64_t planes, int out_height, int width);
int muscle_upsample_bilinear_bwd(const float* dy, int64_t planes, int height, int width, int out_height,
                                 int out_width, float* dx, void* scratch, size_t scratch_bytes, void* stream);
size_t muscle_er_loss_state_bytes(int batch);
int muscle_er_loss_fwd(const float* raw_cams, const float* raw_sgcs, const float* label, int batch,
                       int num_classes, int height, int width, float* loss_out, void* state,
                       size_t state_bytes, void* stream);
int muscle_er_loss_bwd(const float* raw_cams, const float* raw_sgcs, const float* label, int batch,
                       int num_classes, int height, int width, const void* state, size_t state_bytes,
                       const float* grad_scale, float* dsgcs_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * FieldLoss — replaces src/edge.py:175-384 (constructed train_muscle.py:163, called :196).
 * Stage 1 (muscle_field_edges; replaces Mix_fg.forward :423-440, OrientQuantize :65-89 and the
 * selection of FieldLoss.forward :370-374): seg_map [batch, C, H, W], label_with_bg [batch, C]
 * float -> orient u8 [batch, C-1, H, W] (bins 0..7), pos u8 (boundary flags), mag_sum
 * [batch, H, W] (the second return value of the reference), plane_max u32 [batch, C-1] (float
 * bits) and counts int32 [batch, C-1, 3] = {boundary pixels, kept outside indices, kept inside
 * indices} (in_out_div :195-227 with `step`).  prob_scratch, mag_scratch: float [batch,C-1,H,W].
 * The host reads `counts`, applies the gates (:378, :297) and draws the samples (random.sample).
 * Stage 2 (muscle_field_compact): raster-ordered outside / inside index lists of the n_active
 * planes (planes[i] = b*(C-1)+c; offsets[2i], offsets[2i+1] = start of the plane's outside /
 * inside list in `lists`).
 * Stage 3 (muscle_field_loss; replaces bifilter :261-361 + loss_constructor :229-258):
 * sel int32 [n_active, 2, k] = positions in the plane's lists; channel softmax of dense_ft
 * [batch, channels, H, W] and mask [batch, C, H, W] at the sampled pixels only; loss_out float[1]
 * (already / batch); ddense_out = d loss / d dense_ft (fully written) or NULL. */
int muscle_field_edges(const float* seg_map, const float* label_with_bg, int batch, int num_classes,
                       int height, int width, float beta, int step, float* prob_scratch,
                       float* mag_scratch, uint8_t* orient_out, uint8_t* pos_out,
                       float* mag_sum_out, uint32_t* plane_max_out, int32_t* counts_out,
                       void* stream);
/* The same with the Sobel pair chosen like the reference class's constructor argument (src/edge.py:29-43:
 * kernel_size 3 or 5, each with its 1e-6 stand-ins for zero); muscle_field_edges is sobel_size = 5, the value
 * train_muscle.py:163 passes. */
int muscle_field_edges_sobel(const float* seg_map, const float* label_with_bg, int batch, int num_classes,
                             int height, int width, float beta, int step, int sobel_size, float* prob_scratch,
                             float* mag_scratch, uint8_t* orient_out, uint8_t* pos_out,
                             float* mag_sum_out, uint32_t* plane_max_out, int32_t* counts_out,
                             void* stream);
int muscle_field_compact(const uint8_t* orient, const uint8_t* pos, int height, int width, int step,
                         const int32_t* planes, const int64_t* offsets, int n_active,
                         int32_t* lists, void* stream);
size_t muscle_field_loss_scratch_bytes(int n_active, int k, int channels, int num_classes);
int muscle_field_loss(const float* dense_ft, const float* mask, int batch, int channels,
                      int num_classes, int height, int width, const int32_t* planes,
                      const int64_t* offsets, const int32_t* sel, const int32_t* lists, int n_active,
                      int k, float* loss_out, float* ddense_out, void* scratch, size_t scratch_bytes,
                      void* stream);

/* ------------------------------------------------------------------------------------------
 * Region crops — replaces the tensor work of torchutils.get_dynamic_crops
 * (src/torchutils.py:217-291; called train_mcl.py:220): every window of a batch in one launch.
 * x [batch, channels, height, width]; desc int32 [n_crops, 10] = {image, top, left, lh, lw, rh, rw,
 * pool, oh, ow}: window (top,left,lh,lw) of the map (absolute coordinates), resized bilinearly with
 * align_corners=True to (rh,rw) (:270), then 4x4 average-pooled when pool != 0 (:271-272, :279;
 * view-2 windows use rh=lh, rw=lw, pool=1); output [channels, oh, ow] at element offsets[i] of
 * `flat`; offsets int64 [n_crops+1] ascending, offsets[n_crops] = total.  The window plan itself
 * (numpy RNG, :240-254) is host integer work.  bwd: dx (fully written) from dflat. */
int muscle_crops_fwd(const float* x, int channels, int height, int width, const int32_t* desc,
                     const int64_t* offsets, int n_crops, int64_t total, float* flat_out,
                     void* stream);
int muscle_crops_bwd(const float* dflat, int batch, int channels, int height, int width,
                     const int32_t* desc, const int64_t* offsets, int n_crops, int64_t total,
                     float* dx_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * EMD()(crops1, crops2, mode='dynamic') — replaces src/loss_multilabel.py:201-257,287-326
 * (train_mcl.py:221).  Regions live in two flat buffers as [21, pixels] blocks; pair_desc int64
 * [n_pairs, 6] = {offset in flat1, N, offset in flat2, M, image, 0}, pairs of one image listed in
 * the reference's order (crop1 outer, crop2 inner).  Per pair: cost 1 - x y^T, weights
 * x.mean(y) / y.mean(x), `maxiter` log-domain Sinkhorn iterations with `reg`, value
 * sum(exp(M) C)/(N M); per image the FIRST pair with the smallest value is kept; loss_out =
 * mean over images; dflat1_out (size total1, fully written) = d loss / d flat1 through the
 * unrolled iterations of the kept pairs, or NULL.  N, M <= 1024. */
size_t muscle_emd_scratch_bytes(int n_pairs, int n_images, int64_t max_n_plus_m, int maxiter);
int muscle_emd_dynamic_fwd_bwd(const float* flat1, const float* flat2, const int64_t* pair_desc,
                               int n_pairs, int n_images, int channels, int64_t max_n,
                               int64_t max_m, int64_t max_n_plus_m, float reg, int maxiter,
                               float* loss_out, float* dflat1_out, int64_t total1, void* scratch,
                               size_t scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * EMD()(crops1, crops2, mode='static') — replaces src/loss_multilabel.py:259-285 (the class's
 * default mode; fed by get_static_crops, src/torchutils.py:177-214).  Regions are [n, C, h, w]
 * blocks in two flat buffers; pair_desc int64 [n_pairs, 12] = {offset in flat1, n*C, h1, w1, oh1,
 * ow1, offset in flat2, h2, w2, oh2, ow2, n} with (oh, ow) the resize target of :265-269 /
 * :275-278 (host integer arithmetic), h, w <= 1024.  Per pair: mean over the resized position
 * pairs and the batch dim of 1 - x.y; loss_out = sum of the two smallest (stable order);
 * dists_out [n_pairs] (optional) = every pair's value; dflat1_out (size total1, fully written) =
 * d loss / d flat1, or NULL.  n_pairs >= 2 (the reference indexes dists[1]). */
size_t muscle_emd_static_scratch_bytes(int n_pairs, int max_planes);
int muscle_emd_static_fwd_bwd(const float* flat1, const float* flat2, const int64_t* pair_desc,
                              int n_pairs, int max_planes, float* loss_out, float* dists_out,
                              float* dflat1_out, int64_t total1, void* scratch,
                              size_t scratch_bytes, void* stream);

/* SURVEY.md 8f rank 2 — the decoder tail (src/MuSCLe.py:285-286: dense_ft = bilinear up-sampling of p3_dec, seg_map =
 * fuse_dec(dense_ft)) feeding FieldLoss without ever materialising dense_ft: the 1x1 fuse_dec commutes with the
 * up-sampling, so seg_map = upsample(muscle_conv1x1_fwd(p3_dec)) (21 channels instead of 256 at full resolution), and
 * stage 3 interpolates only the <= 2 k sampled columns per plane from p3_dec [batch, channels, small_height, small_width]
 * (align_corners=True); dp3_out has p3_dec's shape.  Everything else as muscle_field_loss. */
int muscle_field_loss_decoder(const float* p3_dec, int small_height, int small_width, const float* mask,
                              int batch, int channels, int num_classes, int height, int width,
                              const int32_t* planes, const int64_t* offsets, const int32_t* sel,
                              const int32_t* lists, int n_active, int k, float* loss_out, float* dp3_out,
                              void* scratch, size_t scratch_bytes, void* stream);
/* y[b][k][p] = bias[k] + sum_c weight[k][c] x[b][c][p]  (NCHW, out_channels <= 32; bias may be NULL) */
int muscle_conv1x1_fwd(const float* x, const float* weight, const float* bias, int batch, int channels,
                       int out_channels, int64_t pixels, float* y, void* stream);

/* ------------------------------------------------------------------------------------------
 * SURVEY.md 8f rank 3 — IRN random walk (src/indexing.py:116-150; inference-time refinement, no gradient).
 * muscle_irn_transition_init: trans_out[i][j] = affinity[i][j]^beta / sum_i affinity[i][j]^beta  ([n, n] row-major;
 * colsum_scratch float[n]) — the element-wise half of to_transition_matrix (:117-119) in two passes.  The `times`
 * squarings (:120-121) and the final product (:147) are plain dense fp32 GEMMs and are left to the GEMM library by
 * the Python mirror (muscle_b200.indexing).  muscle_irn_mask_edge: y = x * (1 - edge) (:145), x [planes, hw], edge [hw]. */
int muscle_irn_transition_init(const float* affinity, int n, float beta, float* colsum_scratch,
                               float* trans_out, void* stream);
int muscle_irn_mask_edge(const float* x, const float* edge, int planes, int hw, float* y, void* stream);
/* The squarings T <- T T (src/indexing.py:120-121) and the final product (x * (1 - edge)) T (:147) as fp32-ACCURATE products
 * on the tensor pipe: every fp32 value is split into three bf16 parts (x = h + m + l, 24 mantissa bits) and
 * A B ~= Ah Bh + Ah Bm + Am Bh + Ah Bl + Al Bh + Am Bm is accumulated in fp32 (TMEM); a transition matrix has no
 * negative entries, so an output entry is as accurate as its terms (~2^-23 relative).
 *   muscle_irn_split3: parts [3][rows_pad][cols_pad] bf16 (hi, mid, lo), zero outside rows x cols; pads multiples of 128.
 *   muscle_irn_matmul_split3: C = A B from the parts of A [m_pad, k_pad] and B [k_pad, n_pad]; c_out fp32 [m, n] (row-major,
 *   unpadded) and / or c_parts_out [3][m_pad][n_pad] (the next squaring's operand: no separate split pass); either may be
 *   NULL, not both; the output must not alias an input. */
size_t muscle_irn_split3_bytes(int rows_pad, int cols_pad);
int muscle_irn_split3(const float* x, int rows, int cols, int rows_pad, int cols_pad, void* parts, void* stream);
int muscle_irn_matmul_split3(const void* a_parts, const void* b_parts, int m_pad, int k_pad, int n_pad, int m, int n,
                             float* c_out, void* c_parts_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MUSCLE_B200_H_ */
