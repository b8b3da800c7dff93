/* fm_b200.h -- C ABI of the B200-native coarse-to-fine matching head.
 *
 * Drop-in boundary for the three hot-path modules of Iheckzza/FeatureMatching
 * (paths relative to the reference tree):
 *
 *   fm_coarse_match  <->  CoarseMatching.forward + get_coarse_match
 *                         network/utils/coarse_matching_new.py:43-143   (called at network/net.py:75)
 *   fm_fine_gather   <->  FinePreprocess.forward
 *                         network/module/fine_preprocess.py:32-62       (called at network/net.py:78)
 *   fm_fine_match    <->  FineMatching.forward
 *                         network/utils/fine_matching_new.py:22-79      (called at network/net.py:83)
 *   fm_fine_transformer <-> LocalFeatureTransformer.forward on the fine windows
 *   fm_coarse_transformer <-> LocalFeatureTransformer.forward on the coarse tokens
 *                         network/module/transformer.py:78-97           (called at network/net.py:80)
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless stated;
 *   - the caller owns all buffers (inputs, outputs, workspace); the library allocates nothing
 *     on the device and keeps no reference to caller memory after the call's work completes;
 *   - every call only ENQUEUES work on `stream` (a cudaStream_t passed as void*); results,
 *     including the match count and the status word, are valid after the stream is synchronised;
 *   - return value: FM_OK or a negative FM_ERR_* for argument / launch errors detected on the
 *     host (fm_last_error_string() has the detail).  Conditions only the device can detect
 *     (capacity overflow, lost dynamic range) are reported in the status word, see below;
 *   - no CPU fallback exists: without a CUDA device every compute entry point fails.
 *
 * The host mirror of the reference interface (same class names, constructor dicts and `data`
 * keys) is featurematching_b200/modules.py; INTEGRATION.md shows the binding a reference
 * maintainer adds.
 */
#ifndef FM_B200_H_
#define FM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FM_B200_VERSION 100 /* 0.1.0 */

#if defined(__GNUC__)
#define FM_API __attribute__((visibility("default")))
#else
#define FM_API
#endif

/* return codes */
#define FM_OK 0
#define FM_ERR_BAD_ARG (-1)
#define FM_ERR_WORKSPACE (-2) /* workspace too small or misaligned */
#define FM_ERR_CUDA (-3)      /* a CUDA runtime / driver call failed */
#define FM_ERR_UNSUPPORTED (-4)

/* element types of feature tensors */
#define FM_DTYPE_BF16 0
#define FM_DTYPE_F32 1

/* fine feature-map layouts (logical shape is always [B, C, H, W]) */
#define FM_LAYOUT_NHWC 0 /* channels-last strides: the layout the gather reads directly */
#define FM_LAYOUT_NCHW 1 /* what the reference backbone emits; transposed once per call */

/* status word: int32[FM_STATUS_WORDS] on the device, written by fm_coarse_match */
#define FM_STATUS_WORDS 8
#define FM_ST_COUNT 0     /* number of matches M (clamped to `capacity`)            */
#define FM_ST_FLAGS 1     /* bitmask of FM_FLAG_*                                   */
#define FM_ST_MAX_CAND 2  /* largest per-unit candidate count seen (sizing hint)    */
#define FM_ST_LOST 3      /* tokens whose soft-max sums left fp32 range             */
#define FM_ST_CAND_TOTAL 4
#define FM_ST_REDO 5      /* 1 = the exact (second) candidate-extraction pass ran           */

#define FM_FLAG_CAND_OVERFLOW 1 /* candidate buffer too small even for the exact extraction pass
                                   (entries whose row soft-max exceeds 0.9 thr): retry with a
                                   larger cand_per_row                                        */
#define FM_FLAG_OUT_OVERFLOW 2  /* more matches than `capacity`                               */
#define FM_FLAG_TIE_OVERFLOW 4  /* exactly tied matches beyond 4 per row exceed the shared list
                                   of 16384 entries (duplicate descriptors en masse); up to
                                   that, all ties are emitted like the reference's == test    */
#define FM_FLAG_RANGE 8         /* FM_ST_LOST > 0: scores span > 157 nats below their bound   */

FM_API int fm_version(void);
FM_API const char* fm_last_error_string(void);

/* Number of SMs of the current device (used for grid sizing); <0 on error. */
FM_API int fm_device_sm_count(void);

/* Kernel launches this library has issued since it was loaded (all threads, all devices).  Take the
 * difference around a region to count the library's own launches in it. */
FM_API long long fm_launch_count(void);

/* ---------------------------------------------------------------- coarse stage
 * Bytes of workspace fm_coarse_match needs.  `cand_per_row` sizes the candidate buffers
 * (records per image-0 token per 256-column block; 4 is ample for thr >= 0.1, the Python shim
 * grows it when FM_FLAG_CAND_OVERFLOW comes back).  Returns 0 on invalid arguments.
 * Candidate extraction is single-pass (a chunk is kept when its best entry exceeds 0.9 thr of the
 * row's 128-column partial soft-max sum).  When that weak test overflows the buffers -- diffuse rows:
 * random-init networks, thr << 1/128 -- the call re-runs the statistics kernel ON ITS OWN (decided on
 * the device, no host round trip; FM_ST_REDO reports it) with the full-row sums as the reference,
 * which keeps only entries whose row soft-max really exceeds 0.9 thr. */
FM_API size_t fm_coarse_workspace_bytes(int B, int N0, int N1, int C, int feat_dtype, int cand_per_row);

/* Dual-softmax mutual-nearest-neighbour coarse matching of B frame pairs.
 *
 *   conf[b,i,j] = softmax_i(s)[b,i,j] * softmax_j(s)[b,i,j],
 *   s[b,i,j]    = <feat_c0[b,i], feat_c1[b,j]> / (C * temperature)      (coarse_matching_new.py:64-68)
 *   match (b,i,j) iff conf > thr, both cells >= border_rm from their grid borders, and conf is
 *   the maximum of its row and of its column                            (coarse_matching_new.py:99-106)
 *
 * feat_c0 [B,N0,C], feat_c1 [B,N1,C] contiguous, bf16 or fp32 (fp32 is rounded to bf16 once,
 * on the device: the contraction is bf16 x bf16 -> fp32).  C must be a multiple of 64, <= 256.
 * N0 = h0c*w0c, N1 = h1c*w1c.  `scale` = image height / coarse height (coarse_matching_new.py:126);
 * scale0/scale1: optional [B,2] per-pair multipliers (data['scale0'] / data['scale1']), or NULL.
 *
 * Outputs (capacity entries each; sorted lexicographically by (b,i,j) like torch.where):
 *   b_ids,i_ids,j_ids int64[capacity]; mconf float[capacity]; mkpts0_c,mkpts1_c float[capacity,2]
 *   (x,y) in pixels; status int32[FM_STATUS_WORDS].
 * The N0 x N1 score / confidence matrix is never written to memory (eval path only;
 * data['conf_matrix'] is not produced). */
FM_API int fm_coarse_match(const void* feat_c0, const void* feat_c1, int feat_dtype, int B, int N0, int N1,
                    int C, int h0c, int w0c, int h1c, int w1c, float temperature, float thr,
                    int border_rm, float scale, const float* scale0, const float* scale1,
                    int64_t* b_ids, int64_t* i_ids, int64_t* j_ids, float* mconf, float* mkpts0_c,
                    float* mkpts1_c, int32_t* status, int capacity, void* workspace,
                    size_t workspace_bytes, int cand_per_row, void* stream);

/* ---------------------------------------------------------------- fine stage: windows
 * Transposes a fine map [B,C,H,W] (NCHW, bf16 or fp32) to channels-last bf16 [B,H,W,C]. */
FM_API int fm_fine_to_nhwc_bf16(const void* src, int src_dtype, int B, int C, int H, int W, void* dst,
                         void* stream);

/* FinePreprocess: W x W windows (stride `stride`, zero padding `pad`) of both fine maps at the
 * matched coarse cells, optionally merged with the down-projected coarse tokens:
 *   win_k[m, ww, :] = merge_w[:, :Cf] . window_k[m, ww, :] + merge_w[:, Cf:] . ctx_k[m] + merge_b
 *   ctx_0[m] = down_w . feat_c0[b, i] + down_b,  ctx_1[m] = down_w . feat_c1[b, j] + down_b
 *                                                               (fine_preprocess.py:43-60)
 * feat_f0/feat_f1: channels-last bf16 [B,Hf,Wf,Cf] (FM_LAYOUT_NHWC; use fm_fine_to_nhwc_bf16
 * first for NCHW input).  feat_c0/feat_c1: [B,N0,Cc]/[B,N1,Cc] bf16 or fp32.
 * down_w [Cf,Cc], down_b [Cf], merge_w [Cf,2Cf], merge_b [Cf] fp32, or all NULL when
 * fine_concat_coarse_feat is off.  `count` points at the match count on the device (status word
 * of fm_coarse_match) so no host round-trip is needed; entries >= *count are not touched.
 * win0/win1: [capacity, W*W, Cf] in out_dtype.  Cf must be 64, W*W <= 64.
 * With merge parameters the work runs on the tensor cores (fp32 weights as 3-term bf16 splits, so
 * products are fp32-grade) and needs `workspace` of fm_fine_gather_workspace_bytes(...) bytes
 * (16-byte aligned); without them it is a pure copy and workspace may be NULL. */
FM_API size_t fm_fine_gather_workspace_bytes(int B, int N0, int N1, int Cc, int Cf, int featc_dtype);
FM_API int fm_fine_gather(const void* feat_f0, const void* feat_f1, const void* feat_c0,
                   const void* feat_c1, int featc_dtype, const int64_t* b_ids, const int64_t* i_ids,
                   const int64_t* j_ids, const int32_t* count, int capacity, int B, int Hf, int Wf,
                   int Cf, int N0, int N1, int Cc, int w0c, int w1c, int stride, int pad, int W,
                   const float* down_w, const float* down_b, const float* merge_w,
                   const float* merge_b, void* win0, void* win1, int out_dtype, void* workspace,
                   size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------- fine stage: matching
 * FineMatching (dual correlate): mixed centre descriptor of each window against the OTHER
 * image's window, softmax(./sqrt(Cf)), spatial expectation on the [-1,1]^2 window grid, std:
 *   mkpts0_f[m] = (mkpts0_c[m] + coords0[m] * (W/2) * scale + W/2, std0[m])   (fine_matching_new.py:50-79)
 * win0/win1 [capacity, WW, Cf] bf16 or fp32; mix_w* [WW], mix_b* [1] fp32;
 * mkpts*_c float[capacity,2]; outputs float[capacity,3].  `scale` = image height / fine height. */
FM_API int fm_fine_match(const void* win0, const void* win1, int win_dtype, const float* mkpts0_c,
                  const float* mkpts1_c, const float* mix_w0, const float* mix_b0,
                  const float* mix_w1, const float* mix_b1, const int32_t* count, int capacity,
                  int WW, int Cf, float scale, float* mkpts0_f, float* mkpts1_f, void* stream);

/* ---------------------------------------------------------------- fine stage: window transformer
 * LocalFeatureTransformer on the gathered windows -- the step between fm_fine_gather and
 * fm_fine_match (network/net.py:79-80 -> network/module/transformer.py:78-97; EncoderLayer
 * transformer.py:34-57 with LinearAttention attentions.py:22-46, no masks):
 *   'self' : feat0 = layer(feat0, feat0); feat1 = layer(feat1, feat1)
 *   'cross': feat0 = layer(feat0, feat1); feat1 = layer(feat1, feat0)      (the UPDATED feat0)
 * win0/win1, out0/out1: float [capacity, WW, C]; out may alias win (in place).  C must be 64,
 * nhead 8, WW <= 49 (the shipped fine config, demo/net_config.py:40-44).
 * layer_kinds: HOST array [n_layers], 0 = 'self', 1 = 'cross'.
 * layer_params: HOST array [n_layers * 10] of DEVICE pointers to the fp32 parameters of each
 *   layer in state-dict order q_proj.weight [C,C], k_proj.weight, v_proj.weight, merge.weight,
 *   mlp.0.weight [2C,2C], mlp.2.weight [C,2C], norm1.weight [C], norm1.bias, norm2.weight, norm2.bias.
 * workspace: fm_fine_transformer_workspace_bytes(n_layers) bytes (packed weights); windows, outputs and
 * workspace must be 32-byte aligned (256-bit global accesses).
 * All GEMMs run on the tensor cores with fp32-grade 3-term bf16 splits; entries >= *count are
 * not touched. */
/* Both transformers: the first bytes of `workspace` receive the layers' weights in operand form (bf16 hi + lo planes,
 * pre-swizzled).  `layer_params` == NULL skips that packing: the workspace must then be the one (same pointer, same
 * n_layers) an earlier call with the same parameters packed -- what a module does in eval mode, where the weights do
 * not change between calls (network/module/transformer.py:59-97 holds them as nn.Parameters). */
FM_API size_t fm_fine_transformer_workspace_bytes(int n_layers);
FM_API int fm_fine_transformer(const float* win0, const float* win1, const int32_t* count, int capacity,
                        int WW, int C, int nhead, int n_layers, const int32_t* layer_kinds,
                        const float* const* layer_params, float* out0, float* out1, void* workspace,
                        size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------- coarse stage: token transformer
 * LocalFeatureTransformer on the 1/8-resolution tokens -- the producer of the coarse matcher's
 * inputs (network/net.py:74 -> network/module/transformer.py:59-97; EncoderLayer transformer.py:34-57
 * with LinearAttention attentions.py:22-46, no masks); semantics of 'self' / 'cross' as above.
 * feat0 [B, L0, C], feat1 [B, L1, C] fp32; out0 / out1 the updated tokens (fp32; may alias feat).
 * out0_bf16 / out1_bf16 (may be NULL): the same tokens rounded to bf16 -- exactly the tensors
 * fm_coarse_match reads, written without a separate cast pass.  C must be 256, nhead 8 (the shipped
 * coarse config, demo/net_config.py:20-25).  layer_kinds / layer_params as for fm_fine_transformer.
 * workspace: fm_coarse_transformer_workspace_bytes(B, L0, L1, n_layers) bytes, 256-byte aligned;
 * token pointers 32-byte aligned.  All GEMMs run on the tensor cores with fp32-grade 3-term bf16
 * splits.  B == 0 or L0 == L1 == 0 is a no-op; exactly one empty side is FM_ERR_BAD_ARG (the reference
 * divides by the source length, attentions.py:40-41). */
FM_API size_t fm_coarse_transformer_workspace_bytes(int B, int L0, int L1, int n_layers);
FM_API int fm_coarse_transformer(const float* feat0, const float* feat1, int B, int L0, int L1, int C,
                          int nhead, int n_layers, const int32_t* layer_kinds,
                          const float* const* layer_params, float* out0, float* out1, void* out0_bf16,
                          void* out1_bf16, void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------- match records
 * Packs the match list into 40-byte records for the cross-GPU gather (the analogue of the
 * reference's pickled all_gather, utils/comm.py:141-176, used at lightning_new.py:291-295):
 *   record = { int32 pair, int32 i, int32 j, float conf, float x0, y0, std0, x1, y1, std1 }
 * records[0] is a header { int32 count, int32 pair_offset, 0... }; match m goes to records[1 + m]
 * with pair = b_ids[m] + pair_offset.  records: (capacity + 1) * 40 bytes. */
#define FM_RECORD_BYTES 40
FM_API int fm_pack_matches(const int64_t* b_ids, const int64_t* i_ids, const int64_t* j_ids,
                           const float* mconf, const float* mkpts0_f, const float* mkpts1_f,
                           const int32_t* count, int capacity, int pair_offset, void* records,
                           void* stream);

/* Append form for a batch processed in chunks of pairs on one stream: chunk `chunk_index`
 * (0 .. FM_MAX_CHUNKS-1, in call order) writes its records behind those of the earlier chunks and
 * leaves its count in chunk_counts[chunk_index] (device int32[FM_MAX_CHUNKS]); the header record
 * carries the running total and the pair_offset of chunk 0.  records: (total_capacity + 1) * 40
 * bytes.  One device->host copy of `records` then returns the whole batch. */
#define FM_MAX_CHUNKS 64
FM_API int fm_pack_matches_append(const int64_t* b_ids, const int64_t* i_ids, const int64_t* j_ids,
                                  const float* mconf, const float* mkpts0_f, const float* mkpts1_f,
                                  const int32_t* count, int capacity, int pair_offset,
                                  int32_t* chunk_counts, int chunk_index, int total_capacity,
                                  void* records, void* stream);

/* ---------------------------------------------------------------- backbone output heads
 * The hand-over from the Swin-UNet to the matcher, network/backbone/swin_backbone_modify_unetarchecture.py:689-702:
 *   x = norm_up[j](x); out = liner_middle(x) (128 -> 256, coarse tokens) | liner_end(x) (64 -> 64, fine map);
 *   out.view(-1, Wh, Ww, C).permute(0, 3, 1, 2).contiguous()
 * as one pass per head: LayerNorm(Cin, eps) + Linear(Cin -> Cout, no bias) on `rows` = B * Wh * Ww rows of
 * x [rows, Cin] fp32, written row-major [rows, Cout] -- i.e. CHANNELS-LAST maps / token sequences, the layouts the
 * matcher's kernels read -- as fp32 (out_f32, the coarse transformer's input) and / or bf16 (out_bf16, the window
 * gather's input); either may be NULL.  linear_weight [Cout, Cin] row-major (nn.Linear.weight).
 * Supported (Cin, Cout): (64, 64), (128, 256). */
FM_API int fm_backbone_emit(const float* x, long long rows, int Cin, int Cout, const float* ln_weight,
                            const float* ln_bias, float ln_eps, const float* linear_weight, float* out_f32,
                            void* out_bf16, void* stream);

/* ---------------------------------------------------------------- pose from matches
 * Relative pose of every frame pair of a batch from its fine matches, on the device -- the consumer of
 * mkpts*_f that the reference runs through OpenCV on the host after a device->host copy:
 *   network/utils/pose_pred.py:130-151  estimate_pose: cv2.findEssentialMat(kpts0, kpts1, focal, pp, RANSAC,
 *                                       0.999, 1.0) + cv2.recoverPose(E, kpts0, kpts1, np.eye(3), 1e9, mask)
 *   network/utils/pose_pred.py:170-179  get_opencv_pose (R = I, t = 0 when no pose comes out)
 *   utils/metrics.py:79-109,124-159     the same per batch element (mask = m_bids == bs)
 * b_ids int64[capacity] sorted ascending (as fm_coarse_match emits them); mkpts0_f / mkpts1_f float rows of
 * `pts_stride` floats (3 for the head's [M, 3] outputs, 2 for plain [M, 2]); count = device match count or
 * NULL (= capacity); K0 float[B, 9] row-major intrinsics of image 0 per pair -- like the reference, focal =
 * (fx + fy) / 2 and the principal point of K0 normalise BOTH images; threshold_px = RANSAC threshold in pixels
 * (the reference: 1.0).  A match is an inlier iff its Sampson error <= (threshold_px / focal)^2 (OpenCV's
 * definition).  num_hypotheses 8-point samples per pair (4096 recommended), deterministic in `seed`.
 * flags:  FM_POSE_POLISH       least-squares local optimisation of the best hypotheses (recommended; OpenCV
 *                              returns its minimal-sample model unrefined)
 *         FM_POSE_IDENTITY_K   rank the (R, +-t) candidates on PIXEL coordinates with an identity camera matrix,
 *                              exactly as the reference's recoverPose call does; without it the normalised
 *                              coordinates are used (what cv2.recoverPose(E, p0, p1, K) computes)
 * Outputs per pair: R float[B, 9] row-major, t float[B, 3] unit vector, E float[B, 9] (may be NULL), n_inliers,
 * pose_status (0 = ok, 1 = fewer than 8 matches, 2 = no sample gave a model, 3 = a model was found but no inlier
 * triangulates in front of both cameras -- recoverPose's n = 0, the reference's None; R = I, t = 0 for 1, 2 and 3)
 * and, if not NULL, inlier_mask uint8[capacity] aligned with the match list.
 * Known difference: the reference's 5-point solver also answers pairs with 5, 6 or 7 matches; this 8-point solver
 * reports them as status 1 (identity pose), like pairs with fewer than 5.
 * workspace: fm_pose_workspace_bytes(B, capacity, num_hypotheses) bytes, 16-byte aligned. */
#define FM_POSE_POLISH 1
#define FM_POSE_IDENTITY_K 2
FM_API size_t fm_pose_workspace_bytes(int B, int capacity, int num_hypotheses);
FM_API int fm_pose_from_matches(const int64_t* b_ids, const float* mkpts0_f, const float* mkpts1_f, int pts_stride,
                                const int32_t* count, int capacity, int B, const float* K0, float threshold_px,
                                int num_hypotheses, int flags, uint32_t seed, float* R, float* t, float* E,
                                int32_t* n_inliers, int32_t* pose_status, uint8_t* inlier_mask, void* workspace,
                                size_t workspace_bytes, void* stream);

/* Pose from matches without a caller-provided workspace.  No counterpart in the reference (its pose step is a host-side
 * OpenCV call, pose_pred.py:130-151, which has no device scratch to speak of): this entry point exists for hosts of
 * THIS library that cannot allocate device memory where they need a pose -- a plug-in in a serving process whose
 * memory arena belongs to the framework, a per-frame callback of a live tracker where allocation latency is not
 * allowed, a bare cgo / JNI / C binding that has no allocator at all.  Arguments, outputs, status codes and results are
 * those of fm_pose_from_matches (same kernels, bit-identical output for the same inputs and seed); the intermediate
 * buffers live in an array of FM_POSE_STATIC_SCRATCH_BYTES that is part of the library image on every GPU that loads
 * it, for the life of the process.
 *   - Limit: fm_pose_workspace_bytes(B, capacity, num_hypotheses) <= FM_POSE_STATIC_SCRATCH_BYTES (e.g. 8 pairs x 4096
 *     hypotheses x 32768 matches); anything larger returns FM_ERR_WORKSPACE and names the size it would have needed.
 *   - The call allocates nothing: no cudaMalloc / cudaFree / allocating library routine, ever.
 *   - The first call on each device resolves the array's device address (not stream work; it may load a module):
 *     make it at start-up, before any stream capture begins.  Later calls only enqueue kernels on `stream`.
 *   - One array per device, shared by all threads and streams: calls that can overlap on the GPU race on it and
 *     return garbage -- keep them on one stream or order them with events.  Contents left by an earlier call are
 *     harmless (every word is rewritten before it is read). */
#define FM_POSE_STATIC_SCRATCH_BYTES (2 * 1024 * 1024)
FM_API int fm_pose_from_matches_static(const int64_t* b_ids, const float* mkpts0_f, const float* mkpts1_f,
                                       int pts_stride, const int32_t* count, int capacity, int B, const float* K0,
                                       float threshold_px, int num_hypotheses, int flags, uint32_t seed, float* R,
                                       float* t, float* E, int32_t* n_inliers, int32_t* pose_status,
                                       uint8_t* inlier_mask, void* stream);

/* ---------------------------------------------------------------- test / profiling hooks
 * Same contract as fm_coarse_match, but the score statistics come from a plain SIMT kernel
 * instead of the tcgen05 kernel.  Exists so tests can tell a tensor-core plumbing bug from a
 * resolve bug; never used by the product path. */
FM_API int fm_debug_coarse_match_simt(const void* feat_c0, const void* feat_c1, int feat_dtype, int B,
                               int N0, int N1, int C, int h0c, int w0c, int h1c, int w1c,
                               float temperature, float thr, int border_rm, float scale,
                               const float* scale0, const float* scale1, int64_t* b_ids,
                               int64_t* i_ids, int64_t* j_ids, float* mconf, float* mkpts0_c,
                               float* mkpts1_c, int32_t* status, int capacity, void* workspace,
                               size_t workspace_bytes, int cand_per_row, void* stream);

/* Copies the per-token log2 soft-max sums of the last fm_coarse_match call on this workspace:
 * l2row float[B*N0], l2col float[B*N1] (device pointers). */
FM_API int fm_debug_coarse_lse(const void* workspace, int B, int N0, int N1, int C, int feat_dtype,
                        int cand_per_row, float* l2row, float* l2col, void* stream);

/* Number of hypotheses per pair that produced a model in the fm_pose_from_matches call that last used `workspace`
 * (same B, capacity, num_hypotheses): valid int32[B], 0 for pairs with fewer than 8 matches.  A healthy build turns
 * nearly every sample of a non-degenerate pair into a model; tests assert that (a miscompiled null-space routine once
 * rejected every sample and the API answered with identity poses, status 2). */
FM_API int fm_debug_pose_valid_hypotheses(const void* workspace, int B, int capacity, int num_hypotheses,
                                          int32_t* valid, void* stream);

/* Times only the coarse statistics kernel (the tcgen05 GEMM + fused dual-softmax sums):
 * runs it `iters` times on `stream` and returns the mean milliseconds per launch in *ms. */
FM_API int fm_profile_coarse_stats(const void* feat_c0, const void* feat_c1, int B, int N0, int N1, int C,
                            float temperature, float thr, void* workspace, size_t workspace_bytes,
                            int cand_per_row, int iters, float* ms, void* stream);

/* Stage timers.  While enabled, every fm_coarse_match call brackets its statistics kernel (the
 * tcgen05 GEMM + fused dual-softmax sums) with a pair of CUDA events on the call's stream (up to
 * 4096 calls are kept).  fm_profile_collect synchronises on the recorded events, returns the
 * summed milliseconds and the number of launches, and clears the list. */
FM_API int fm_profile_enable(int on);
FM_API int fm_profile_collect(float* ms_sum, int* launches);

#ifdef __cplusplus
}
#endif
#endif /* FM_B200_H_ */
