/* cmf_b200 -- C ABI of the B200 (sm_100a) kernels behind the non-square log-det hot path of
 * Canonical Manifold Flows (k-flouris/cmf).
 *
 * This header is the drop-in boundary.  Every entry point takes plain device pointers and
 * sizes (no torch types), launches on the caller's stream, allocates nothing, keeps no state
 * between calls and never synchronises.  All tensors are dense row-major fp32 unless noted.
 * Return value: 0 = ok, <0 = bad argument (see cmf_last_error_string), >0 = cudaError_t.
 *
 * Which reference code each entry point replaces is cited as path:line relative to the
 * reference checkout (cmf/... = cmf/models/components/...).
 *
 * Layout conventions
 *   B   samples in the (per-GPU) batch shard
 *   d   tangent columns carried next to the primal (latent dimension, or Hutchinson probes)
 *   M   elements of the half a coupling layer transforms, stored COMPACT: z_m (B,M), V_m (B,d,M)
 *   Q   elements of one conditioner output ("shift" or "log-scale") per sample; the
 *       conditioner emits (B, 2Q): shift = [0,Q), log-scale = [Q,2Q)  (couplers.py:52-59)
 *   map (M) int32: position in [0,Q) of compact element j (NULL = identity, requires M == Q)
 *   inv (Q) int32: compact index of conditioner position q, or -1 (NULL = identity)
 */
#ifndef CMF_B200_H
#define CMF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* cmf_stream_t; /* a cudaStream_t */

#define CMF_B200_ABI_VERSION 1

#define CMF_PENALTY_NONE    0
#define CMF_PENALTY_OFFDIAG 1 /* sum_{i!=j} |G_ij|  non_square.py:95-100 (g_ij_loss) */
#define CMF_PENALTY_DIAG    2 /* sum_k |G_kk|       non_square.py:87-92  (g_kk_loss) */

int cmf_version(void);
const char* cmf_last_error_string(void);
/* Number of kernel launches issued by this library since it was loaded (all threads); bench.py
 * reports the difference over its timed region as "gpu_launches". */
unsigned long long cmf_launch_count(void);

/* (a) Affine coupling, primal + all d tangent columns in one pass.
 * Replaces acl.py:53-66 (checkerboard _jvp), acl.py:125-146 (channelwise _jvp) called d times
 * from non_square.py:298-311, and with d == 0 the primal-only acl.py:48-51,113-123 (_z_to_x)
 * used by flow_forward (non_square.py:313-320).
 *   e      = exp(-log_scale)
 *   x_m    = z_m * e - shift
 *   V_out  = e * (V_m - z_m * log_scale_dot) - shift_dot
 */
int cmf_coupling_jvp_fwd(int B, int d, int M, int Q,
                         const float* z_m, const float* V_m,
                         const float* st, const float* ST, const int32_t* map,
                         float* x_m, float* V_out, cmf_stream_t stream);

/* Adjoint of cmf_coupling_jvp_fwd (what autograd derives from the ~10 ATen ops per layer per
 * column of acl.py:53-66 / :125-146).  g_x may be NULL (treated as zero).  g_st (B,2Q) and
 * g_ST (B,d,2Q) are written densely (zeros where inv[q] < 0).  workspace must hold
 * cmf_coupling_jvp_bwd_workspace(B,d,M,Q) bytes (may be 0).  The query is conservative: the mapped
 * (32-bit) path keeps its tangent-chunk partials there for a second, reducing launch; the identity-map
 * 128-bit path reduces the chunks inside the CTA (one launch) and leaves the workspace untouched
 * except for tiny grids with d > 256. */
size_t cmf_coupling_jvp_bwd_workspace(int B, int d, int M, int Q);
int cmf_coupling_jvp_bwd(int B, int d, int M, int Q,
                         const float* z_m, const float* V_m,
                         const float* st, const float* ST, const int32_t* map, const int32_t* inv,
                         const float* g_x, const float* g_V,
                         float* g_z, float* g_Vin, float* g_st, float* g_ST,
                         void* workspace, size_t workspace_bytes, cmf_stream_t stream);

/* Allocation-free variant of cmf_coupling_jvp_bwd.  It has NO counterpart in the reference: it was
 * added for callers that cannot call cudaMalloc (code under
 * CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same arguments and results
 * as cmf_coupling_jvp_bwd minus the workspace: the tangent-chunk partials live in a file-scope
 * `static __device__` array of CMF_COUPLING_BWD_STATIC_MAX_BYTES that every GPU loading the library
 * holds until exit.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library
 * routine).  Sizes whose cmf_coupling_jvp_bwd_workspace(B,d,M,Q) exceeds the array are rejected
 * (return < 0, cmf_last_error_string says how many bytes were needed).  ALL calls share the one
 * array per GPU and it keeps the previous call's contents (nothing in it needs to start at zero:
 * every partial read is written first by the same call): calls that overlap on the GPU -- two
 * streams, two host threads -- race.  The MNIST-shape layer (B=50, d=64, M=Q=392) needs 2.4 MiB. */
#define CMF_COUPLING_BWD_STATIC_MAX_BYTES (4u << 20)
size_t cmf_coupling_jvp_bwd_static_max_bytes(void);
int cmf_coupling_jvp_bwd_static(int B, int d, int M, int Q,
                                const float* z_m, const float* V_m,
                                const float* st, const float* ST, const int32_t* map, const int32_t* inv,
                                const float* g_x, const float* g_V,
                                float* g_z, float* g_Vin, float* g_st, float* g_ST,
                                cmf_stream_t stream);

/* Index maps folded into addressing: out[r, i] = idx[i] < 0 ? 0 : cat(in0, in1)[r, idx[i]].
 * Replaces the zero-pad / permute / squeeze / split copies of non_square.py:397-410,
 * split.py:26-30,50-52, reshaping.py:28-29,102-114 and the mask*z products of acl.py:54.
 * in1 may be NULL (n1 = 0).  idx must be injective on its non-negative entries. */
int cmf_gather_rows(long long rows, int n0, int n1, int n_out,
                    const float* in0, const float* in1, const int32_t* idx,
                    float* out, cmf_stream_t stream);

/* (b) Per-sample metric G = J^T J with tangents stored as V (B,d,n) (V = J^T, row k = column k
 * of J restricted to this block): G[b] (+)= V[b] V[b]^T.  Replaces torch.stack + torch.bmm of
 * non_square.py:307-308.  accumulate != 0 adds into G (several blocks of one sample). */
int cmf_gram_fwd(int B, int d, int n, const float* V, float* G, int accumulate, cmf_stream_t stream);

/* Same, with the kernel chosen explicitly (parity tests pin both):
 *   CMF_GRAM_TC   tcgen05.mma kind::tf32 with TMEM accumulators, FP32-accurate 3xTF32 split, TMA
 *                 operand staging; needs d >= 16, n >= 32, n % 4 == 0, V 16-byte aligned
 *   CMF_GRAM_FMA  FP32 FMA SYRK (any shape; the tabular / 2-D configs with d <= 10)
 *   CMF_GRAM_AUTO TC when its shape constraints hold, else FMA (what cmf_gram_fwd does). */
#define CMF_GRAM_AUTO 0
#define CMF_GRAM_FMA  1
#define CMF_GRAM_TC   2
int cmf_gram_fwd_path(int B, int d, int n, const float* V, float* G, int accumulate, int path,
                      cmf_stream_t stream);

/* Two blocks of the same samples in one pass: G[b] (+)= V1[b] V1[b]^T + V2[b] V2[b]^T, the sum over
 * the layer blocks of non_square.py:307-308 for the usual two-block split of J^T.  One launch with
 * one accumulation over the n1 + n2 columns when both blocks qualify for the tensor-core kernel
 * (each as in CMF_GRAM_TC); otherwise the same result from two cmf_gram_fwd calls. */
int cmf_gram_fwd2(int B, int d, int n1, const float* V1, int n2, const float* V2, float* G, int accumulate,
                  cmf_stream_t stream);

/* Adjoint: g_V[b] = (g_G[b] + g_G[b]^T) V[b]   (B,d,n). */
int cmf_gram_bwd(int B, int d, int n, const float* V, const float* g_G, float* g_V, cmf_stream_t stream);
/* CMF_GRAM_TC here: D[n,j] = sum_k V[k,n] S[k,j] with V as the MN-major tcgen05 operand (TMA,
 * 32-byte-atom 128B swizzle) and S = g_G + g_G^T staged by TMA; additionally needs d % 4 == 0 and
 * g_G 16-byte aligned.  AUTO picks it for d > 64. */
int cmf_gram_bwd_path(int B, int d, int n, const float* V, const float* g_G, float* g_V, int path,
                      cmf_stream_t stream);

/* (c) Batched Cholesky log-det + metric penalties.  Replaces torch.linalg.cholesky + diagonal +
 * log + sum (non_square.py:282,293-294) and masked_select + vector_norm (non_square.py:95-100,
 * :87-92).  Outputs: L (B,d,d) lower factor (upper triangle zero), logdet (B) = 2 sum log L_kk,
 * l1_off (B), l1_diag (B).  info (1 int32, device, MUST be zeroed by the caller): set to
 * 1 + index of a sample whose leading minor is not positive (the RuntimeError of
 * non_square.py:283-288); the host then adds eps*I and retries like the reference.
 * d <= 128: factor held in shared memory (several samples per CTA for small d); 128 < d <= 1024
 * (CIFAR-shape latent sweep): panel-blocked out of L2, one CTA per sample. */
int cmf_chol_logdet_fwd(int B, int d, const float* G, float* L, float* logdet,
                        float* l1_off, float* l1_diag, int32_t* info, cmf_stream_t stream);

/* Adjoint: g_G = g_logdet * G^{-1} + g_l1 * sign(G) (.) mask(penalty)   (B,d,d), with G^{-1}
 * rebuilt from L in shared memory.  g_l1 may be NULL when penalty == CMF_PENALTY_NONE. */
int cmf_chol_logdet_bwd(int B, int d, const float* L, const float* G,
                        const float* g_logdet, const float* g_l1, int penalty,
                        float* g_G, cmf_stream_t stream);

/* Conditioner convolution: 3x3, 64 -> 64 channels, stride 1, zero padding 1, NCHW
 * fp32 -- the two convolutions of every residual block of the coupler ResNet
 * (cmf/models/components/networks.py:35-94,116-161), applied to the primal and to all d tangent
 * columns (jvp_layers.py:49-64).  Replaces the cuDNN FP32 calls F.conv2d issues there.  tcgen05
 * kind::tf32 with the 3xTF32 split (pixel operand staged in tensor memory, weights pre-split by
 * cmf_conv3x3_c64_prep), FP32 round-to-nearest accumulation across K segments: results agree with
 * an FP32 convolution to ~1e-6 relative.
 *   prep:  w (64,64,3,3) -> wsplit (2*9*64*64 floats, 16-byte aligned).  transpose = 0 for the
 *          forward / tangent push, 1 for the data gradient (g_x = conv(g_y, prep(w, 1))).
 *   fwd:   x (N,64,H,W) -> y (N,64,H,W) (+ bias (64), may be NULL).  Needs H*W % 4 == 0 and
 *          W <= 63 (cmf_conv3x3_c64_usable); other shapes stay on the library path. */
int cmf_conv3x3_c64_usable(int N, int H, int W);
int cmf_conv3x3_c64_prep(const float* w, int transpose, float* wsplit, cmf_stream_t stream);
int cmf_conv3x3_c64_fwd(int N, int H, int W, const float* x, const float* wsplit, const float* bias,
                        float* y, cmf_stream_t stream);
/* Same convolution with the elementwise neighbours of the residual block fused into the epilogue
 * (networks.py:62-79: relu JVP = mask by the primal's sign, skip connection):
 *     t = conv(x) + bias;  if (mask_first) t *= m;  if (res) t += res;  y = t;  if (y2) y2 = t * m
 * with m = (mask_src[n / mask_div, co, p] > 0), mask_src = pre-ReLU PRIMAL activations
 * (N / mask_div, 64, H, W): mask_div = d tangent columns per sample.  mask_src, res, y2 may be NULL. */
int cmf_conv3x3_c64_fused(int N, int H, int W, const float* x, const float* wsplit, const float* bias,
                          const float* mask_src, int mask_div, int mask_first, const float* res,
                          float* y, float* y2, cmf_stream_t stream);
/* The same with the mask given as packed sign bits: cmf_relu_mask_pack turns the primal activations
 * x (N, 64, HW) into bits (N, HW, 2) uint32 -- bit e of word 0 / word 1 = (x[n, e, p] > 0) /
 * (x[n, 32 + e, p] > 0), 8-byte aligned -- once per residual block; the d tangent images of a
 * sample then read ONE 8-byte word per output pixel instead of 64 strided floats, in the forward
 * and in the backward launches. */
int cmf_relu_mask_pack(int N, int HW, const float* x, uint32_t* bits, cmf_stream_t stream);
int cmf_conv3x3_c64_fused_bits(int N, int H, int W, const float* x, const float* wsplit, const float* bias,
                               const uint32_t* mask_bits, int mask_div, int mask_first, const float* res,
                               float* y, float* y2, int drain_stages, cmf_stream_t stream);
/* Joint batch (one launch for the tangents AND the primal of a residual-block convolution in the
 * backward pass, where nothing orders them): images [0, Nt) are tangent images -- image n belongs to
 * sample n / mask_div and gets no bias -- and images [Nt, N) are the primal images of samples
 * 0 .. N - Nt - 1 (Nt / mask_div == N - Nt), which read mask row n - Nt and are the only ones that get
 * `bias`.  Everything else as cmf_conv3x3_c64_fused_bits.  Replaces the separate aten::convolution_backward
 * calls autograd issues for the primal and the tangent use of one weight (networks.py:62-79). */
int cmf_conv3x3_c64_joint(int N, int Nt, int H, int W, const float* x, const float* wsplit, const float* bias,
                          const uint32_t* mask_bits, int mask_div, int mask_first, const float* res,
                          float* y, float* y2, int drain_stages, cmf_stream_t stream);
/* drain_stages = K = 32 stages accumulated on the tensor core (truncating FP32 adds) between two
 * round-to-nearest drains into registers: 1 for everything that feeds the metric G (its inverse
 * amplifies forward errors by cond(G) ~ 5e3), 2 for the data-gradient launches of the backward
 * pass (a 4e-7 instead of 2e-7 relative error on a gradient; 7 % faster).  mask_bits may be NULL. */

/* Weight gradient of the same convolution: g_w (64,64,3,3) (+)= sum_{n,p} g_y[n,co,p] x[n,ci,p+tap]
 * (what autograd derives for F.conv2d at cmf/trainer.py:216, aten::convolution_backward).
 * tcgen05 3xTF32 with pixels as the K dimension; CTAs write partial sums into `workspace`
 * (cmf_conv3x3_c64_wgrad_workspace bytes, 16-byte aligned) and a second kernel adds them in a fixed
 * order.  Needs H*W % 4 == 0 and W <= 47 (cmf_conv3x3_c64_wgrad_usable). */
int cmf_conv3x3_c64_wgrad_usable(int N, int H, int W);
size_t cmf_conv3x3_c64_wgrad_workspace(int N, int H, int W);
int cmf_conv3x3_c64_wgrad(int N, int H, int W, const float* x, const float* g_y, float* workspace,
                          int accumulate, float* g_w, cmf_stream_t stream);
/* Same with the reproducibility of the result chosen by the caller (what
 * torch.use_deterministic_algorithms selects for aten::convolution_backward):
 *   deterministic = 0  two MMA issuers share each accumulator; the order of their truncating FP32
 *                      accumulations depends on timing: same inputs reproduce to 1-3 ulp (the
 *                      default above; accuracy against FP64 is the same in both modes)
 *   deterministic = 1  one in-order MMA stream per accumulator: bitwise reproducible, ~9 % slower */
int cmf_conv3x3_c64_wgrad_det(int N, int H, int W, const float* x, const float* g_y, float* workspace,
                              int accumulate, float* g_w, int deterministic, cmf_stream_t stream);

/* Rational-quadratic spline coupling, inverse direction (z -> x), primal + d tangent columns.
 * Arithmetic of gitmodules/nsf/nde/transforms/splines/rational_quadratic.py:13-52,96-112 (linear
 * tails: identity outside [-tail_bound, tail_bound]) and the bin search of
 * gitmodules/nsf/utils/torchutils.py:133-138, done with __ballot_sync/__popc over the K+1 <= 32
 * height knots of an element.  The reference's cmf wrapper (bijections/nsf.py:19-42) has no _jvp
 * (SURVEY D4): the tangent push is new; its oracle is autograd over the nsf arithmetic.
 *   cw, ch, dv (B,M,K+1): cumulative widths / heights and knot derivatives, built by the host with
 *   the reference's torch ops (bit-identical knots => bit-identical bins); CW, CH, DV (B,d,M,K+1)
 *   their tangents.  bin (B,M) int32 (may be NULL): selected bin, -1 outside the interval. */
int cmf_rqs_coupling_jvp_fwd(int B, int d, int M, int K, float tail_bound,
                             const float* z_m, const float* V_m,
                             const float* cw, const float* ch, const float* dv,
                             const float* CW, const float* CH, const float* DV,
                             float* x_m, float* V_out, int32_t* bin, cmf_stream_t stream);
/* Adjoint w.r.t. z, its tangents, the three knot tables and their tangents (dense, zero outside
 * the two selected knots).  g_x may be NULL. */
int cmf_rqs_coupling_jvp_bwd(int B, int d, int M, int K, float tail_bound,
                             const float* z_m, const float* V_m,
                             const float* cw, const float* ch, const float* dv,
                             const float* CW, const float* CH, const float* DV,
                             const float* g_x, const float* g_V,
                             float* g_z, float* g_Vin, float* g_cw, float* g_ch, float* g_dv,
                             float* g_CW, float* g_CH, float* g_DV, cmf_stream_t stream);

/* Encode pass and loss epilogue: primal-sized elementwise work around the head, one CTA
 * per sample.
 *   preproc_logit: DequantizationDensity + scalar-mult + scalar-add + logit (wrapper.py:27-30,
 *     math.py:41-53,70-105; config/schemas.py:30-50) in one pass.  x (B,n) gets `noise` (B,n, may be NULL)
 *     added IN PLACE like the reference; y = logit(mult * x + add); logjac (B) = n log|mult| + sum(-log u
 *     - log(1 - u)) with u clamped to [1e-7, 1 - 1e-7] in the log-Jacobian only. */
int cmf_preproc_logit_fwd(int B, int n, float* x, const float* noise, float mult, float add,
                          float* y, float* logjac, cmf_stream_t stream);
/*   coupling_x_to_z: affine coupling in the encode direction (acl.py:43-46,101-111), whole layer:
 *     z[b,i] = x[b,i] for passthrough positions, (x[b,i] + t) * exp(s) for modified ones, logjac (B) = sum s.
 *     pos2j (D) int32: compact index j of the modified element at flat position i, or -1; qmap (M) int32
 *     or NULL: conditioner position of compact element j (NULL: q = j); st (B,2Q) = [shift | log-scale].
 *     Adjoint: q2pos (Q) int32 = flat position read through conditioner position q, or -1; g_st is
 *     written densely.  g_z, g_logjac may be NULL (zero). */
int cmf_coupling_x_to_z_fwd(int B, int D, int Q, const float* x, const float* st, const int32_t* pos2j,
                            const int32_t* qmap, float* z, float* logjac, cmf_stream_t stream);
int cmf_coupling_x_to_z_bwd(int B, int D, int Q, const float* z, const float* st, const int32_t* pos2j,
                            const int32_t* qmap, const int32_t* q2pos, const float* g_z, const float* g_logjac,
                            float* g_x, float* g_st, cmf_stream_t stream);
/*   sumsq_diff: out[b] = sum_i (a[b,i] - c[b,i])^2, c may be NULL: the reconstruction error
 *     (non_square.py:111-114) and the quadratic form of the standard-normal log-prob (gaussian.py:9-22).
 *     Adjoint w.r.t. a: g_a = 2 g[b] (a - c). */
int cmf_sumsq_diff_fwd(int B, int n, const float* a, const float* c, float* out, cmf_stream_t stream);
int cmf_sumsq_diff_bwd(int B, int n, const float* a, const float* c, const float* g, float* g_a, cmf_stream_t stream);

/* Whole-stack kernel for the small flat configs (the 2-D group of config/two_d.py:
 * `flat-realnvp`, D <= 8, tanh MLP conditioners of width 10): ONE launch pushes the primal and all d
 * tangent columns through the tail pad (non_square.py:397-410), every affine coupling layer INCLUDING
 * its conditioner MLP with the reference's JVP rules (acl.py:125-146, networks.py:206-224,
 * jvp_layers.py:38-64) and forms G = J^T J (non_square.py:307-308); one more launch is the adjoint of all
 * of it.  The model's parameters live in shared memory, a sample's state in registers / local memory;
 * per sample the kernel reads z (B,d) and writes G (B,d,d) and the primal output x (B,D).
 *   desc (int32, host memory): D, d, L, nh, P, pos[d] (logical position of latent k after the tail pad),
 *     then per layer in z -> x order: n_keep, n_mod, keep[n_keep], mod[n_mod] (logical positions in
 *     conditioner-input / -output order), width[nh + 2] (MLP widths, first = n_keep, last = 2 n_mod),
 *     w_off[nh + 1], b_off[nh + 1] (offsets of each Linear's weight (out,in) / bias in `params`).
 *   params (P) fp32: the packed parameters.  bwd: g_G, g_x may be NULL (zero); g_params_partial holds
 *     cmf_flat_stack_bwd_partials(B) x P floats, one partial per CTA (the host adds them). */
#define CMF_FLAT_MAX_D       8
#define CMF_FLAT_MAX_LATENT  8
#define CMF_FLAT_MAX_WIDTH   16
#define CMF_FLAT_MAX_HIDDEN  4
#define CMF_FLAT_MAX_LAYERS  16
#define CMF_FLAT_MAX_PARAMS  4096
int cmf_flat_stack_usable(int D, int d, int L, int nh, int max_width, int P);
int cmf_flat_stack_fwd(const int32_t* desc, int desc_len, int B, const float* z, const float* params,
                       float* G, float* x, cmf_stream_t stream);
int cmf_flat_stack_bwd_partials(int B);
int cmf_flat_stack_bwd(const int32_t* desc, int desc_len, int B, const float* z, const float* params,
                       const float* g_G, const float* g_x, float* g_z, float* g_params_partial,
                       cmf_stream_t stream);

/* ---- tangent-batched tanh-MLP conditioner of the flat couplers --------------------
 * The whole MLP of a coupler -- Linear, tanh, ..., Linear; primal + d tangent columns -- in one launch,
 * and its adjoint (data, weight and bias gradients) in one launch plus a fixed-order reduction.
 * Replaces cmf/models/components/networks.py:206-224 (get_mlp), :24-32 (NN_Sequential_JVP.jvp),
 * jvp_layers.py:38-47 (bias-free Linear on the tangent, tanh' of the primal as the tangent's mask) and
 * what autograd derives from them at cmf/trainer.py:216.  Exact FP32 FMA (the reference runs these
 * Linears in FP32), bitwise reproducible.
 *   width[0 .. nlin]: input width, hidden widths, output width; every width in 1 .. 128, nlin <= 8;
 *   d in 0 .. 127 (d = 0: primal only, V / ST / g_ST / g_V / save_U may be null).
 *   x (B, width[0]), V (B, d, width[0]) -> st (B, width[nlin]) (with bias), ST (B, d, width[nlin]).
 *   cmf_mlp_pack: W[l] (width[l+1], width[l]) row-major and b[l] (width[l+1]) (host arrays of device
 *     pointers) -> wpack (cmf_mlp_pack_floats(nlin) floats): both operand layouts, zero padded.
 *   save_T, save_U: cmf_mlp_save_floats(B, d, nlin) floats each, written by the forward launch for the
 *     adjoint (masked / unmasked hidden vectors, tile by tile); both null = inference.
 *   partial: cmf_mlp_partial_floats(B, d, nlin) floats of scratch (one partial per CTA, no zeroing needed);
 *   gpack (nlin x (128*128 + 128) floats): g_W[l][j][k] at l*16512 + j*128 + k, g_b[l][j] at l*16512 + 16384 + j.
 *   16-byte alignment of wpack, save_T, save_U.  No allocation, no synchronisation. */
#define CMF_MLP_MAX_LINEARS 8
int cmf_mlp_usable(int nlin, const int* width, int d);
long long cmf_mlp_pack_floats(int nlin);
int cmf_mlp_tiles(int B, int d);
long long cmf_mlp_save_floats(int B, int d, int nlin);
int cmf_mlp_partials(int B, int d);
long long cmf_mlp_partial_floats(int B, int d, int nlin);
int cmf_mlp_pack(int nlin, const int* width, const float* const* W, const float* const* b, float* wpack,
                 cmf_stream_t stream);
int cmf_mlp_jvp_fwd(int B, int d, int nlin, const int* width, const float* x, const float* V,
                    const float* wpack, float* st, float* ST, float* save_T, float* save_U, cmf_stream_t stream);
int cmf_mlp_jvp_bwd(int B, int d, int nlin, const int* width, const float* x, const float* V,
                    const float* wpack, const float* save_T, const float* save_U, const float* g_st,
                    const float* g_ST, float* g_x, float* g_V, float* partial, float* gpack, cmf_stream_t stream);

/* ---- tail of the image couplers' conditioner ------------------------------------------
 * ReLU -> 1x1 convolution (C -> O) -> scaled tanh on the primal hidden activations h (B, C, HW) and, by the
 * JVP rules, on the d tangent images Hv (B d, C, HW), evaluated at the P spatial positions the coupling
 * layer uses (inv: position -> compact index or < 0; null = all positions, P == HW):
 *   y (B, O, P) = sw * tanh(W2 relu(h) + b2) + sb,   Y (B, d, O, P) = (W2 (Hv * [h > 0])) * (1 - tanh^2) * sw.
 * Replaces cmf/models/components/networks.py:156-161 (ReLU, 1x1 conv of get_resnet), :96-113
 * (ScaledTanh2dModule and its jvp), jvp_layers.py:38-64, the `(1 - mask)` selection of acl.py:53-66, and
 * their autograd.  The adjoint writes DENSE g_h (B, C, HW) / g_Hv (B d, C, HW) (zeros at unused positions) and
 * g_params = [g_W2 (O, C) | g_b2 (O) | g_sw (O) | g_sb (O)]; partial: cmf_resnet_tail_partial_floats floats of
 * scratch (no zeroing needed).  g_y may be null.  C <= 64, O <= 32.  Bitwise reproducible. */
#define CMF_TAIL_MAX_C 64
#define CMF_TAIL_MAX_OUT 32
int cmf_resnet_tail_usable(int C, int O);
long long cmf_resnet_tail_partial_floats(int B, int d, int HW, int P, int C, int O);
int cmf_resnet_tail_fwd(int B, int d, int C, int HW, int P, int O, const float* h, const float* Hv,
                        const int32_t* inv, const float* w2, const float* b2, const float* sw, const float* sb,
                        float* y, float* Y, cmf_stream_t stream);
int cmf_resnet_tail_bwd(int B, int d, int C, int HW, int P, int O, const float* h, const float* Hv,
                        const int32_t* inv, const float* w2, const float* b2, const float* sw, const float* sb,
                        const float* g_y, const float* g_Y, float* g_h, float* g_Hv, float* partial,
                        float* g_params, cmf_stream_t stream);

/* ---- first convolution of the image couplers' conditioner ------------------------------
 * C_in -> 64 channels, 3 x 3, padding 1, no bias; x (N, C_in, H, W), w (64, C_in, 3, 3), y (N, 64, H, W).
 * Replaces cmf/models/components/networks.py:131-139 (first get_conv of get_resnet) and its JVP rule
 * jvp_layers.py:49-64 (the same convolution on every tangent image), forward / data gradient / weight
 * gradient (cuDNN in the reference).  C_in <= 16 (the gradients run in passes of <= 4 input channels),
 * W <= 64, H <= 256, N <= 65535 per call.
 * cmf_convin_wgrad: partial = cmf_convin_wgrad_partials(N, H, W) x 64 * min(C_in, 4) * 9 floats of scratch (no zeroing
 * needed); g_w (64, C_in, 3, 3) = the partials summed in block order (bitwise reproducible). */
#define CMF_CONVIN_MAX_CIN 16
int cmf_convin_usable(int Cin, int Cout, int H, int W);
int cmf_convin_wgrad_partials(int N, int H, int W);
int cmf_convin_fwd(int N, int Cin, int H, int W, const float* x, const float* w, float* y, cmf_stream_t stream);
int cmf_convin_dgrad(int N, int Cin, int H, int W, const float* g_y, const float* w, float* g_x, cmf_stream_t stream);
int cmf_convin_wgrad(int N, int Cin, int H, int W, const float* x, const float* g_y, float* partial, float* g_w,
                     cmf_stream_t stream);

/* ---- bias gradient of a convolution ----------------------------------------------------------------------
 * out[c] = sum over n, p of x[n, c, p]; x (N, C, HW) contiguous (a batch slice of a larger tensor is fine).
 * Replaces autograd's g_y.sum(dim = (0, 2, 3)) behind every biased Conv2d of the coupler ResNet
 * (cmf/models/components/networks.py:35-94 at cmf/trainer.py:216).  partial: C x cmf_channel_sum_splits(N, C)
 * floats of scratch (no zeroing needed).  Fixed summation order. */
int cmf_channel_sum_splits(int N, int C);
int cmf_channel_sum(int N, int C, int HW, const float* x, float* partial, float* out, cmf_stream_t stream);

/* ---- ReLU + packed sign bits in one pass ----------------------------------------------------------------
 * relu_out (N, 64, HW) = max(x, 0) and bits as cmf_relu_mask_pack writes them, from one read of x (N, 64, HW):
 * the primal side of every residual block (cmf/models/components/networks.py:62-79) and the mask its tangents
 * take (jvp_layers.py:38-46).  relu_out may be a batch slice of a larger tensor; it must not alias x. */
int cmf_relu_and_pack(int N, int HW, const float* x, uint32_t* bits, float* relu_out, cmf_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* CMF_B200_H */
