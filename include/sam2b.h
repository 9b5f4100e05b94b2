/* sam2b — C ABI of the B200-native (sm_100a) SAM2.1 Hiera-L hot path.
 *
 * This is the drop-in boundary for the sub-graphs FasterProcess/sam2_opt hands to a backend executor
 * (reference: sam2/ytools/executor/model_executor.py:133-169 `ModelExectuor.Inference`), i.e. what its
 * TensorRT engines compute:
 *   E1/E2  image encoder   sam2/sam2/sam2_image_predictor.py:252-276 (set_image_e2e_*),
 *                          sam2/sam2/modeling/sam2_base_official.py:566-598 (inference_image_*)
 *   E3     memory attention sam2/sam2/modeling/memory_attention.py:283-381 (inference_memory_attention_*)
 *   E4     memory encoder   sam2/sam2/modeling/memory_encoder.py:228-251 (inference_memory_*)
 *
 * Conventions: every pointer marked `dev` is a CUDA device pointer on the engine's device; tensors are
 * contiguous fp32 in the reference's own layouts; `stream` is a cudaStream_t (work is enqueued on it and the call
 * returns without synchronising, like TensorRTExecutor's event hand-off, tensorrt_executor.py:251-289);
 * inputs are borrowed for the duration of the enqueued work, outputs are caller-allocated.
 * Every function returns 0 on success; otherwise sam2b_last_error() describes the failure (thread-local).
 * There is no CPU fallback: without a CUDA device sam2b_create fails.
 */
#ifndef SAM2B_H_
#define SAM2B_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sam2b_engine* sam2b_handle;

const char* sam2b_last_error(void);
const char* sam2b_version(void);
/* number of kernels this library has launched so far in this process (all engines) */
unsigned long long sam2b_launch_count(void);

/* ---- engine life cycle (replaces TensorRTExecutor.__init__/Release, tensorrt_executor.py:16-73,311-334) ---- */
int sam2b_create(int device, sam2b_handle* out);
int sam2b_destroy(sam2b_handle h);
/* Register one tensor of the reference state_dict (reference names, e.g. "image_encoder.trunk.blocks.0.attn.qkv.weight")
 * or a derived constant ("derived.*", see sam2_opt_b200/weights.py).  The data is copied (device to device). */
int sam2b_set_tensor(sam2b_handle h, const char* name, const float* dev_data, const int64_t* shape, int ndim);
/* Repack weights into kernel layouts (bf16 K-major) and check that everything the path needs is present. */
int sam2b_finalize(sam2b_handle h);

/* ---- E1 / E2: Hiera-L trunk + FpnNeck + conv_s0/conv_s1 ----
 * img      dev [n,3,1024,1024]; normalised=1: already (x-mean)/std (E2, inference_image);
 *                               normalised=0: raw [0,1] pixels, normalisation is fused (E1, set_image_e2e)
 * add_no_mem_embed: 1 adds no_mem_embed to the 64x64 level (E1, sam2_image_predictor.py:259-260)
 * fpn0 dev [n,32,256,256], fpn1 dev [n,64,128,128], fpn2 dev [n,256,64,64]  (NCHW fp32)
 * (vision_features == fpn2; the vision_pos_enc outputs are constants, see sam2b_position_encoding) */
int sam2b_image_encode(sam2b_handle h, const float* img, int n, int normalised, int add_no_mem_embed,
                       float* fpn0, float* fpn1, float* fpn2, void* stream);
/* Same, with HOST buffers (pinned or pageable): H2D of the images and D2H of the three outputs are issued on `stream`
 * in micro-batches overlapped with compute; returns after the stream has been synchronised. */
int sam2b_image_encode_host(sam2b_handle h, const float* img_host, int n, int normalised, int add_no_mem_embed,
                            float* fpn0_host, float* fpn1_host, float* fpn2_host, void* stream);
/* The same call for a serving loop: returns as soon as everything is enqueued, so that the uploads of call k + 1 overlap the
 * compute of call k (the reference's executor is synchronous per call, tensorrt_executor.py:240-309; this is what a
 * maintainer would use from a double-buffered loader).  The host buffers of a call must stay valid and untouched until
 * sam2b_host_sync(h, stream) has returned, which waits for all uploads, compute and downloads issued so far. */
int sam2b_image_encode_host_async(sam2b_handle h, const float* img_host, int n, int normalised, int add_no_mem_embed,
                                  float* fpn0_host, float* fpn1_host, float* fpn2_host, void* stream);
int sam2b_host_sync(sam2b_handle h, void* stream);
/* Same, from raw frames: hwc_u8_host is HOST uint8 [n][1024][1024][3] (what set_image / set_image_batch receive,
 * sam2_image_predictor.py:173-196, already at the model resolution).  The frames cross PCIe as bytes (a quarter of the fp32
 * traffic) and ToTensor -> Normalize (sam2/utils/transforms.py:27-46) runs on the device, bit-identical to the reference;
 * the outputs equal sam2b_image_encode_host(normalised=1) on the reference-preprocessed frames bit for bit. */
int sam2b_image_encode_host_u8(sam2b_handle h, const void* hwc_u8_host, int n, int add_no_mem_embed,
                               float* fpn0_host, float* fpn1_host, float* fpn2_host, void* stream);

/* ---- constant outputs of E2 / E4: the sine position encodings ----
 * E2's vision_pos_enc0..2 ([n,256,256,256], [n,256,128,128], [n,256,64,64]) and E4's mask_memory_pos_enc ([n,64,64,64]) do
 * not depend on the input (PositionEmbeddingSine, sam2/sam2/modeling/position_encoding.py:90-124, which the reference
 * re-materialises with .repeat(n, ...) on every call).  out_chw dev [channels][h][w] fp32: one level, computed on `stream`;
 * the caller broadcasts it over the batch.  channels must be a multiple of 4 (the reference uses 256 and 64). */
int sam2b_position_encoding(int channels, int h, int w, float* out_chw, void* stream);

/* ---- E3: 4-layer MemoryAttention ----
 * curr, curr_pos dev [4096,n,256]; memory, memory_pos dev [L,4096,n,64]; memory_exclude, memory_pos_exclude dev [P,n,64]
 * (P may be 0, then both may be NULL: the `none` variant); out dev [4096,n,256]. L in [1,16] (the reference's
 * engines: [1,7]; more than one conditioning frame needs more), P in [0,64]. */
int sam2b_memory_attention(sam2b_handle h, const float* curr, const float* curr_pos, const float* memory,
                           const float* memory_pos, int L, const float* memory_exclude,
                           const float* memory_pos_exclude, int P, int n, float* out, void* stream);

/* The same for objects whose memory banks differ (added on different frames, re-prompted ...; SURVEY.md section 8a "dynamic
 * shapes"): the tensors are padded to L_max frames / P_max pointer tokens and object b uses memory[0 .. L_rows[b]) and
 * memory_exclude[0 .. P_rows[b]) only (the padding is never read).  L_rows / P_rows are HOST arrays of n entries (either may be
 * NULL: that count is the same for every object); 1 <= L_rows[b] <= L_max, 0 <= P_rows[b] <= P_max, multiples of 4. */
int sam2b_memory_attention_varlen(sam2b_handle h, const float* curr, const float* curr_pos, const float* memory,
                                  const float* memory_pos, int L_max, const float* memory_exclude,
                                  const float* memory_pos_exclude, int P_max, int n, const int32_t* L_rows,
                                  const int32_t* P_rows, float* out, void* stream);

/* The same with the per-object counts on the DEVICE: counts_dev is a device int32 array [2][n] = (frames L_b, pointer tokens
 * P_b).  Nothing crosses from the host, so the call can sit inside a captured CUDA graph whose bank fills up from frame to frame
 * (sam2_opt_b200.predictor._GraphedTrack).  The values are clamped to [1, L_max] / [0, P_max] (P_b rounded down to a multiple
 * of 4) on the device, since the host cannot validate them. */
int sam2b_memory_attention_varlen_dev(sam2b_handle h, const float* curr, const float* curr_pos, const float* memory,
                                      const float* memory_pos, int L_max, const float* memory_exclude,
                                      const float* memory_pos_exclude, int P_max, int n, const int32_t* counts_dev, float* out,
                                      void* stream);

/* ---- E4: MemoryEncoder ----
 * pix_feat dev [n,256,64,64], masks dev [n,1,1024,1024] (already sigmoid-scaled) -> out_x dev [n,64,64,64].
 * (the second reference output, the sine position encoding, is a constant: sam2b_position_encoding) */
int sam2b_memory_encode(sam2b_handle h, const float* pix_feat, const float* masks, int n, float* out_x, void* stream);

/* The same from the LOW-RESOLUTION mask logits the SAM heads produce: what SAM2Base._encode_new_memory does around the encoder
 * (sam2_base_official.py:978-1026) is folded in.  low_res_masks dev [n,1,256,256]; the 1024^2 input of the down-sampler,
 *   mask_for_mem = sigmoid(bilinear_x4(low)) * 20 - 10     (binarize = 1: (bilinear_x4(low) > 0) * 20 - 10, prompted frames),
 * is evaluated inside the first convolution and never reaches memory (the reference writes the high-res mask, its sigmoid and
 * the scaled copy: 4 x 4 MB per object and frame).  pix_batch_stride: 1048576, or 0 when all n objects share one frame's
 * pix_feat [1,256,64,64].  object_score_logits dev [n] or NULL: objects with a score <= 0 get no_obj_embed_spatial added to
 * their output (sam2_base_official.py:1019-1024). */
int sam2b_memory_encode_low_res(sam2b_handle h, const float* pix_feat, int64_t pix_batch_stride, const float* low_res_masks,
                                int binarize, const float* object_score_logits, int n, float* out_x, void* stream);

/* ---- SAM heads: prompt encoder (points) + two-way mask decoder + object-pointer head ----
 * Replaces SAM2Base._forward_sam_heads (sam2/sam2/modeling/sam2_base_official.py:338-494), i.e. PromptEncoder._embed_points
 * (sam/prompt_encoder.py:124-150), MaskDecoder.forward / predict_masks (sam/mask_decoder.py:116-316, incl. the dynamic
 * multimask fallback :337-382) and TwoWayTransformer (sam/transformer.py:97-219), for n prompts / objects in one call.
 * backbone_features dev: the 256-channel 64x64 embedding of every prompt, element (b, c, p = y*64+x) at
 *                  backbone_features[b*feat_stride_n + c*feat_stride_c + p*feat_stride_p]; one of the two inner strides must be 1
 *                  (NCHW [n,256,64,64]: 1048576, 4096, 1; the E3 output [4096,n,256]: 256, 1, n*256; feat_stride_n = 0 repeats
 *                  one image for every prompt, mask_decoder.py:252-259 repeat_image)
 * feat_s0 dev [.,32,256,256], feat_s1 dev [.,64,128,128]: the conv_s0 / conv_s1 levels (fpn0 / fpn1 of sam2b_image_encode), batch
 *                  strides in elements (0: one image for every prompt)
 * point_coords dev [n,num_points,2] (x, y in the 1024 frame, before the +0.5 shift), point_labels dev int32 [n,num_points]
 *                  (-1 padding, 0/1 negative/positive click, 2/3 box corners); num_points = 0: no prompt (a tracked frame).
 *                  The padding point the reference appends is added internally; num_points <= 25.
 * sparse_embeddings dev [n,num_sparse,256] or NULL: the sparse prompt embeddings themselves (MaskDecoder.forward's
 *                  `sparse_prompt_embeddings`, e.g. from the reference's own PromptEncoder); when given they follow the 6 output
 *                  tokens as they are (no padding point is added, num_sparse in [0,26]) and the point arguments are ignored
 * dense_embedding  dev [n,256,64,64] (the prompt encoder's mask_downscaling output) or NULL: no_mask_embed
 * multimask_output 1: the 3 multimask outputs and the best of them; 0: mask 0 unless it is unstable (dynamic fallback)
 * decoder_only     1: plain MaskDecoder.forward semantics (no NO_OBJ_SCORE fill: what SAM2ImagePredictor.predict uses)
 * outputs (any may be NULL), M = multimask_output ? 3 : 1:
 *   low_res_multimasks dev [n,M,256,256], high_res_multimasks dev [n,M,1024,1024], ious dev [n,M],
 *   low_res_masks dev [n,1,256,256] / high_res_masks dev [n,1,1024,1024] (the best mask), obj_ptr dev [n,256],
 *   object_score_logits dev [n,1], mask_tokens_out dev [n,4,256] (the four mask tokens after the transformer, from which
 *   MaskDecoder.forward slices its `sam_tokens_out`). */
int sam2b_sam_heads(sam2b_handle h, const float* backbone_features, int64_t feat_stride_n, int64_t feat_stride_c,
                    int64_t feat_stride_p, const float* feat_s0, int64_t s0_batch_stride, const float* feat_s1,
                    int64_t s1_batch_stride, const float* point_coords, const int32_t* point_labels, int num_points,
                    const float* sparse_embeddings, int num_sparse, const float* dense_embedding, int n, int multimask_output,
                    int decoder_only, float* low_res_multimasks, float* high_res_multimasks, float* ious, float* low_res_masks,
                    float* high_res_masks, float* obj_ptr, float* object_score_logits, float* mask_tokens_out, void* stream);

/* ---- CUDA graphs over engine calls ----
 * The reference replays nothing: every frame re-enqueues its TensorRT contexts (tensorrt_executor.py:232-309).  Here the
 * engine calls are capturable (cudaStreamBeginCapture on `stream`, then sam2b_memory_attention / sam2b_sam_heads /
 * sam2b_memory_encode ... with sizes the workspaces already hold: a capture must follow one eager run of the same calls):
 * under capture the calls leave out the cross-stream hand-off events that serialise callers on the shared workspaces, so a
 * REPLAY must be bracketed by sam2b_stream_acquire(h, stream) / sam2b_stream_release(h, stream) on the replaying stream.
 * sam2b_workspace_generation changes whenever a workspace buffer was re-allocated (a larger problem arrived): graphs captured
 * under an older generation hold stale addresses and must be captured again. */
int sam2b_stream_acquire(sam2b_handle h, void* stream);
int sam2b_stream_release(sam2b_handle h, void* stream);
unsigned long long sam2b_workspace_generation(sam2b_handle h);

/* The dense embedding of a mask prompt: PromptEncoder._embed_masks (sam2/sam2/modeling/sam/prompt_encoder.py:97-109,152-155),
 * i.e. conv 2x2 s2 + LayerNorm2d + GELU twice and a 1x1 conv.  mask_prompt dev [n,1,256,256] (logits at the prompt encoder's
 * mask_input_size) -> dense_embedding dev [n,256,64,64], the `dense_embedding` argument of sam2b_sam_heads.  Touches no
 * workspace of the engine (no serialisation against other calls needed). */
int sam2b_mask_downscale(sam2b_handle h, const float* mask_prompt, int n, float* dense_embedding, void* stream);

/* ---- peer memory for the embedding gather (the only collective of the path) ----
 * The reference has no multi-GPU inference path (one process, one GPU: sam2/ytools/tensorrt/tensorrt_executor.py:26-29);
 * this is the B200-box replacement: one process per GPU, every rank owns a gather buffer that its peers map through CUDA
 * IPC, the neck writes its results straight into the local buffer and sam2b_peer_copy pushes them to the peers over
 * NVLink on the copy engines (no SM work, so the next step's persistent kernels keep all 148 SMs).
 * sam2b_peer_alloc: cudaMalloc of `bytes` on `device` + its 64-byte IPC handle.  sam2b_peer_open: map a peer's handle into
 * this process for `device` (peer access is enabled lazily).  sam2b_peer_copy: asynchronous copy on `stream`; either side
 * may be a mapped peer pointer. */
int sam2b_peer_alloc(int device, int64_t bytes, void** dev_ptr, unsigned char* ipc_handle_64);
int sam2b_peer_free(int device, void* dev_ptr);
int sam2b_peer_open(int device, const unsigned char* ipc_handle_64, void** mapped_ptr);
int sam2b_peer_close(int device, void* mapped_ptr);
int sam2b_peer_copy(void* dst, const void* src, int64_t bytes, void* stream);

/* ---- per-launch profiling (CUDA event pair around every kernel this library launches, on the launching stream) ----
 * begin: clear + enable.  end: disable, synchronise the device and write a CSV "name,ms,flops,bytes" (one line per
 * launch; flops/bytes are the ALGORITHMIC work of that launch) into buf; *needed receives the full length. */
int sam2b_profile_begin(void);
int sam2b_profile_end(char* buf, int64_t buflen, int64_t* needed);

/* ---- building blocks (used by the parity tests and available to other callers) ---- */
/* C[b][m][n] = act(alpha * A[b] B[b]^T + bias) * gamma + residual ; A [batch,M,K], B [batch or 1,N,K] bf16 K-major.
 * act: 0 none, 1 exact GELU, 2 ReLU.  Exactly one of out_f32 / out_bf16 may be NULL. */
int sam2b_op_gemm(const void* a_bf16, int64_t lda, int64_t a_batch_stride, const void* b_bf16, int64_t ldb,
                  int64_t b_batch_stride, int m, int n, int k, int batch, float alpha, const float* bias,
                  int bias_along_m, int act, const float* gamma, const float* residual, int64_t ld_res,
                  int64_t res_batch_stride, int res_row_mod, float* out_f32, void* out_bf16, int64_t ldc,
                  int64_t c_batch_stride, int force_block_n, void* stream);
/* LayerNorm fold (replaces nn.LayerNorm + nn.Linear pairs of MultiScaleBlock, reference hieradet.py:134-166):
 * producer: out_f32 = A B^T + bias + residual, plus a bf16 copy (out2_bf16) and per-row partial sums
 *   stats [m][*parts][2] = (sum, sum of squares) of every output row, *parts <= stats_capacity_parts;
 * consumer: out = act(LN(x) W^T + b) computed from the RAW rows a_bf16 = bf16(x) as rstd * (a W''^T) + (b + W beta):
 *   b_bf16 = W'' = W diag(gamma) with every row centred to zero mean (the mean subtraction of LN is absorbed by the
 *   weights: sum_k x_k W''[n][k] = sum_k (x_k - mu) (W diag(gamma))[n][k]), bias = b + W beta, rstd from stats over k columns. */
int sam2b_op_gemm_ln_producer(const void* a_bf16, const void* b_bf16, int m, int n, int k, const float* bias,
                              const float* residual, float* out_f32, void* out2_bf16, float* stats,
                              int stats_capacity_parts, int* parts, int force_block_n, void* stream);
int sam2b_op_gemm_ln_consumer(const void* a_bf16, const void* b_bf16, int m, int n, int k, const float* bias,
                              const float* stats, int parts, float eps, int act, float* out_f32,
                              void* out_bf16, int force_block_n, void* stream);
int sam2b_op_layernorm(const float* x, const float* gamma, const float* beta, float eps, void* y_bf16, float* y_f32,
                       int64_t rows, int c, void* stream);
int sam2b_op_cast_bf16(const float* x, void* y_bf16, int64_t n, void* stream);
int sam2b_op_maxpool2x2_nhwc(const float* x, float* y, int b, int hgt, int wid, int c, void* stream);
int sam2b_op_window_attention(const void* qkv_bf16, void* out_bf16, int b, int hgt, int wid, int heads, int ws,
                              int q_pool, void* stream);
/* Connected components with areas, batched (replaces the reference's only native kernel, sam2/csrc/connected_components.cu:
 * get_connected_componnets, called by fill_holes_in_mask_scores, sam2/utils/misc.py:311-338).  mask_u8 dev [n][h][w]
 * (non-zero = foreground), labels / counts dev int32 [n][h][w]: label 0 and count 0 for background, otherwise a positive
 * label shared by the 8-connected component and the component's area in pixels. */
int sam2b_op_connected_components(const void* mask_u8, int32_t* labels, int32_t* counts, int n, int h, int w, void* stream);
/* Pre-processing of camera / video frames (replaces SAM2Transforms.__call__, sam2/utils/transforms.py:38-46, at the model
 * resolution where its Resize is the identity): hwc_u8 dev uint8 [n][h][w][3] -> out_nchw dev fp32 [n][3][h][w] = x / 255 and,
 * with normalise != 0, (x / 255 - mean) / std with the ImageNet constants of transforms.py:27-28.  Same IEEE operations in
 * the same order as the reference: bit-exact.  w must be a multiple of 4. */
int sam2b_op_preprocess_u8(const void* hwc_u8, float* out_nchw, int n, int h, int w, int normalise, void* stream);
/* Frames that are NOT at the model resolution (replaces the Resize of SAM2Transforms, sam2/utils/transforms.py:30, applied to
 * `images / 255.0` in SAM2ImagePredictor._set_image_, sam2_image_predictor.py:193): antialiased bilinear interpolation
 * (F.interpolate(mode="bilinear", antialias=True, align_corners=False): a separable triangle filter whose support grows with
 * the down-scaling factor).  src dev: uint8 [n][h][w][3] when src_is_u8_hwc != 0 (x / 255 first), else fp32 [n][3][h][w];
 * out_nchw dev fp32 [n][3][out_h][out_w], with normalise != 0 followed by (x - mean) / std.  out_w must be a multiple of 4.
 * workspace: sam2b_op_resize_frames_workspace(...) bytes of device memory.  Bit-exact against oracle/preprocess_oracle.py:
 * resize_bilinear_aa, which reproduces the reference framework's CPU operator bit for bit. */
int64_t sam2b_op_resize_frames_workspace(int n, int h, int w, int out_h, int out_w);
int sam2b_op_resize_frames(const void* src, int src_is_u8_hwc, int n, int h, int w, float* out_nchw, int out_h, int out_w,
                           int normalise, void* workspace, void* stream);
/* Allocation-free variant of sam2b_op_resize_frames.  No counterpart in the reference: added
 * for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same arguments and results, but the workspace is a static device array of SAM2B_RESIZE_STATIC_MAX_BYTES inside the library
 * (held by every GPU that loads it, until exit); the call allocates nothing.  Problems whose
 * sam2b_op_resize_frames_workspace(...) exceeds that size are rejected with an error (16 MiB covers one 1920 x 1080 frame
 * -> 1024 x 1024).  All calls on a device share the array: calls that overlap on the GPU (different streams) race; calls
 * on one stream are ordered.  The array keeps the previous call's contents, every call rewrites what it reads. */
#define SAM2B_RESIZE_STATIC_MAX_BYTES (16u << 20)
int sam2b_op_resize_frames_static(const void* src, int src_is_u8_hwc, int n, int h, int w, float* out_nchw, int out_h,
                                  int out_w, int normalise, void* stream);
/* Post-processing of predicted masks (replaces the F.interpolate of SAM2Transforms.postprocess_masks, sam2/utils/transforms.py:119,
 * and the `masks > mask_threshold` that follows it, sam2_image_predictor.py:300-302 / the video predictor's `video_res_masks > 0`):
 * low dev fp32 [n][h][w] mask logits -> out dev fp32 [n][out_h][out_w] (bilinear, align_corners = False; may be NULL) and
 * mask_u8 dev uint8 [n][out_h][out_w] = (out > threshold) (may be NULL; at least one of the two).  fp32 arithmetic of the reference framework's
 * bilinear up-sampling operator (coordinate = one fma, blend = separately rounded operations): bit-exact against oracle/postprocess_oracle.py:upsample_bilinear. */
int sam2b_op_upsample_masks(const float* low, float* out, void* mask_u8, int64_t n, int h, int w, int out_h, int out_w,
                            float threshold, void* stream);
int sam2b_op_softmax_rows(const float* s, int64_t lds, void* p_bf16, int64_t ldp, int64_t rows, int cols, void* stream);
/* Single-head flash attention, head_dim 256 (the MemoryAttention core; reference sam/transformer.py:265-296):
 * out[b] = softmax(q[b] k[b]^T / 16) v[b];  q [n][tq][256], k, v [n][lk][256], out [n][tq][256], all bf16 device memory,
 * tq a multiple of 128.  workspace: at least sam2b_op_attention_hd256_workspace(n, tq, lk) bytes of device memory. */
int64_t sam2b_op_attention_hd256_workspace(int n, int tq, int64_t lk);
int sam2b_op_attention_hd256(const void* q_bf16, const void* k_bf16, const void* v_bf16, int n, int tq, int64_t lk,
                             void* out_bf16, void* workspace, void* stream);

/* The cross-attention form of the same core (reference sam/transformer.py:345-424 with kv_in_dim = 64): the values are the
 * UN-PROJECTED 64-wide memory rows, out[b] = softmax(q[b] k[b]^T / 16) v64[b] with v64 [n][lk][64], out [n][tq][64] (v_proj is
 * linear and the attention weights sum to one, so it is applied afterwards together with out_proj).  lk_rows_dev: optional DEVICE
 * int32 array of n per-object key counts (<= lk); k / v64 rows at or past an object's count are ignored but must be finite. */
int64_t sam2b_op_attention_hd256_v64_workspace(int n, int tq, int64_t lk);
int sam2b_op_attention_hd256_v64(const void* q_bf16, const void* k_bf16, const void* v64_bf16, int n, int tq, int64_t lk,
                                 const int32_t* lk_rows_dev, void* out_bf16, void* workspace, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SAM2B_H_ */
