// Internal C++ interface between the kernels (*.cu) and the engine (engine.cu).  Every function
// enqueues work on `stream` and returns immediately; errors are reported through set_error().
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

namespace sam2b {

typedef __nv_bfloat16 bf16;

void set_error(const char* fmt, ...);
const char* get_error();
int num_sms();
// "This kernel's launch attributes have been set on the current device."  cudaFuncSetAttribute applies to the current device
// only, so a process-wide `static bool` would leave the second GPU of a multi-device process without its opt-in shared
// memory.  Two host threads may both find the flag clear and both set the (idempotent) attribute; neither launches before
// its own call has returned.
class PerDeviceFlag {
 public:
  bool is_set() const { return (bits_.load(std::memory_order_acquire) >> cur()) & 1ull; }
  void set() { bits_.fetch_or(1ull << cur(), std::memory_order_release); }
 private:
  static int cur() { int d = 0; cudaGetDevice(&d); return d & 63; }
  std::atomic<unsigned long long> bits_{0};
};
// counts kernel launches made by this library (bench.py's gpu_launches)
extern std::atomic<unsigned long long> g_launch_count;   // several host threads may drive one library (video_multi_thread.py)

#define SAM2B_CHECK_LAUNCH(name)                                              \
  do {                                                                        \
    ::sam2b::g_launch_count.fetch_add(1, std::memory_order_relaxed);         \
    cudaError_t e__ = cudaGetLastError();                                     \
    if (e__ != cudaSuccess) {                                                 \
      ::sam2b::set_error("%s launch failed: %s", name, cudaGetErrorString(e__)); \
      return 1;                                                               \
    }                                                                         \
  } while (0)

// ---- optional per-launch CUDA-event profiling (bench.py's roofline leg; off by default) ----
// A ProfRange placed around a launch records an event pair on the launching stream when profiling is enabled.
void prof_push(const char* name, double flops, double bytes, cudaStream_t st);
void prof_pop(cudaStream_t st);
extern bool g_prof_on;
struct ProfRange {
  cudaStream_t st; bool on;
  ProfRange(cudaStream_t s, const char* name, double flops, double bytes) : st(s), on(g_prof_on) { if (on) prof_push(name, flops, bytes, st); }
  ~ProfRange() { if (on) prof_pop(st); }
};

enum Act { ACT_NONE = 0, ACT_GELU = 1, ACT_RELU = 2 };

// C[b][m][n] = epilogue( alpha * sum_k A[b][m][k] * B[b][n][k] )      (both operands K-major, bf16)
// epilogue(v) = (act(v + bias)) * gamma[n] + residual[b][m % res_row_mod][n]
struct GemmArgs {
  const bf16* A = nullptr; long lda = 0; long a_batch_stride = 0;
  const bf16* B = nullptr; long ldb = 0; long b_batch_stride = 0;  // b_batch_stride == 0: B shared by all batches
  int M = 0, N = 0, K = 0, batch = 1;
  float alpha = 1.0f;
  const float* bias = nullptr; int bias_along_m = 0;
  int act = ACT_NONE;
  const float* gamma = nullptr;
  const float* residual = nullptr; long ld_res = 0; long res_batch_stride = 0; int res_row_mod = 0;
  // RoPE in the epilogue (bf16 output, N % 256 == 0): rope_table [rope_positions][256] holds (cos, sin) of column pair i at
  // columns (2i, 2i + 1); row m of a batch is rotated by position m % rope_positions if m < rope_rows (rope_rows_dev[b] when given)
  const float* rope_table = nullptr; int rope_positions = 0; int rope_rows = 0; const int* rope_rows_dev = nullptr;
  float* out_f32 = nullptr; bf16* out_bf16 = nullptr; long ldc = 0; long c_batch_stride = 0;
  int force_bn = 0;  // 0 = pick automatically; 1000 + BLOCK_N forces CTA pairs
  // ---- LayerNorm fold (see GemmMode in gemm_kernel.cuh) ----
  // producer: also write a bf16 copy of the fp32 output and per-row partial (sum, sum x^2) into stats_out
  // [batch*M][stats_parts][2]; the number of parts (2 * N tiles) is returned through *stats_parts_out.
  bf16* out2_bf16 = nullptr; long ldc2 = 0; long c2_batch_stride = 0;
  float* stats_out = nullptr; int stats_capacity_parts = 0; int* stats_parts_out = nullptr;
  // consumer: A holds raw rows x; B holds W'' = W diag(gamma) with every row centred to zero mean; bias holds b + W beta;
  // the epilogue computes rstd * acc + bias with rstd from ln_stats [M][ln_parts][2] over the K columns of A.
  const float* ln_stats = nullptr; int ln_parts = 0; float ln_eps = 0.f;
};
int gemm_bf16(const GemmArgs& a, cudaStream_t stream);

// rows x C fp32 -> bf16 (LayerNorm with affine, or plain cast when gamma == nullptr)
int layernorm_rows(const float* x, const float* gamma, const float* beta, float eps, bf16* y_bf16, float* y_f32,
                   long rows, int C, cudaStream_t stream);
int cast_f32_bf16(const float* x, bf16* y, long n, cudaStream_t stream);
// 2x2 max-pool over NHWC fp32: [B,H,W,C] -> [B,H/2,W/2,C]
int maxpool2x2_nhwc(const float* x, float* y, int B, int H, int W, int C, cudaStream_t stream);
// im2col for the 7x7 stride-4 pad-3 patch embedding: img [B,3,S,S] fp32 (optionally (x-mean)/std first)
// -> cols [B*(S/4)^2, 160] bf16 (147 taps in (c,ky,kx) order, zero padded to 160)
int patch_im2col(const float* img, bf16* cols, int B, int S, int normalise, cudaStream_t stream);
// dst[b,c,y,x] += src[b,c,y/2,x/2]   (nearest x2 top-down add of the FPN), NCHW fp32
int upsample2x_add_nchw(float* dst, const float* src, int B, int C, int H, int W, cudaStream_t stream);

// Hiera window attention on a packed qkv tensor [B, H, W, 3, heads, 72] bf16 (token-major).
// ws = window side (ws == H: global).  q_pool: queries are the 2x2 max of q inside the window and the
// output lives on the (H/2, W/2) grid.  out: [B, Ho, Wo, heads*72] bf16.
int window_attention_hd72(const bf16* qkv, bf16* out, int B, int H, int W, int heads, int ws, int q_pool,
                          cudaStream_t stream);

// tcgen05 flash kernel for the 16x16-window and global (ws == W) blocks; same tensor contract as above, no q-pool.
bool flash_attention_hd72_supported(int H, int W, int ws, int q_pool);
int flash_attention_hd72(const bf16* qkv, bf16* out, int B, int H, int W, int heads, int ws, cudaStream_t stream);

// K4 (attn_hd256.cu): softmax(q k^T / 16) v, single head, head_dim 256.  q [n][Tq][256], k / v [n][Lk][256] bf16 (batch
// strides in elements), out [n][Tq][256] bf16.  o_part / ml_part: workspaces of splits*n*Tq*256 and splits*n*Tq*2 floats
// with splits = flash_attention_hd256_splits(n, Tq, Lk).
int flash_attention_hd256_splits(int n, int Tq, long Lk);
int flash_attention_hd256(const bf16* q, long q_batch_stride, const bf16* k, long k_batch_stride, const bf16* v,
                          long v_batch_stride, int n, int Tq, long Lk, float* o_part, float* ml_part, bf16* out,
                          cudaStream_t stream);

// Sine position encoding of one feature level (reference PositionEmbeddingSine): out [C][H][W] fp32
int position_sine(float* out, int C, int H, int W, cudaStream_t stream);

// K4b (attn_hd256_v64.cu): cross-attention core with the value projection folded behind P V: q [n][Tq][256], k [n][Lk][256],
// v [n][Lk][64] (the un-projected memory) bf16 -> out [n][Tq][64] bf16 = softmax(q k^T / 16) v.  lk_rows: optional DEVICE array of
// per-object key counts (<= Lk); v rows in [lk_rows[b], Lk) must hold finite values.  o_part / ml_part: splits*n*Tq*64 and
// splits*n*Tq*2 floats with splits = flash_attention_hd256_v64_splits(n, Tq, Lk).
int flash_attention_hd256_v64_splits(int n, int Tq, long Lk);
int flash_attention_hd256_v64(const bf16* q, long q_batch_stride, const bf16* k, long k_batch_stride, const bf16* v,
                              long v_batch_stride, int n, int Tq, long Lk, const int* lk_rows, float* o_part, float* ml_part,
                              bf16* out, cudaStream_t stream);

// Row softmax over fp32 scores -> bf16 probabilities (row stride lds / ldp elements), `cols` valid columns;
// columns in [cols, ldp) are written as zero.
int softmax_rows(const float* s, long lds, bf16* p, long ldp, long rows, int cols, cudaStream_t stream);
// Connected components of `mask != 0` under 8-connectivity (connected_components.cu): labels int32 [n][H][W] (0 = background,
// otherwise 1 + smallest pixel index of the component), counts = area of the pixel's component (0 for background).
int connected_components(const uint8_t* mask, int* labels, int* counts, int n, int H, int W, cudaStream_t stream);
// uint8 HWC [n][H][W][3] -> fp32 NCHW [n][3][H][W] = x / 255, optionally (x - mean) / std (SAM2Transforms, bit-exact)
int preprocess_u8_hwc(const uint8_t* in, float* out, int n, int H, int W, int normalise, cudaStream_t stream);
// low fp32 [n][h][w] -> out fp32 [n][H][W] (nullable) and mask uint8 [n][H][W] = out > threshold (nullable); bilinear, align_corners = False
int upsample_bilinear_mask(const float* low, float* out, uint8_t* mask, long n, int h, int w, int H, int W, float threshold, cudaStream_t stream);
// Antialiased bilinear resize of frames (resize.cu): src = uint8 [n][h][w][3] (src_is_u8_hwc) or fp32 [n][3][h][w] ->
// out fp32 [n][3][out_h][out_w] = resize(x / 255 | x) and, with normalise, (x - mean) / std.  workspace: resize_frames_workspace bytes.
long resize_frames_workspace(int n, int h, int w, int out_h, int out_w);
int resize_frames(const void* src, int src_is_u8_hwc, int n, int h, int w, float* out, int out_h, int out_w, int normalise,
                  void* workspace, cudaStream_t stream);
// Fused trunk MLP for C = 144 / 288 (mlp_fused.cu): x += fc2(GELU(rstd * (xb W1''^T) + b1)) + b2 in place, bf16 copy and
// per-row (sum, sum of squares) in two parts for the next folded LayerNorm; the hidden layer never reaches HBM.
bool mlp_fused_supported(int C);
int mlp_fused(const bf16* xb, const float* stats_in, int parts_in, float eps, const bf16* w1, const float* b1, const bf16* w2,
              const float* b2, float* x, bf16* xb_out, float* stats_out, long M, int C, cudaStream_t stream);

}  // namespace sam2b

namespace sam2b {
// ---- memory attention helpers (memory_attention.cu) ----
// rows [3][n] (spatial keys, pointer tokens, total) from device counts [2][n] (frames, pointer tokens), clamped to L_max / P_max
int memattn_rows_from_device(const int* counts, int* rows, int n, int L_max, int P_max, cudaStream_t stream);
// x[b][t][c] = curr[t][b][c] + 0.1 * curr_pos[t][b][c]      (seq-first [T,n,C] -> batch-major [n,T,C], fp32)
int memattn_prep_x(const float* curr, const float* curr_pos, float* x, int T, int n, int C, cudaStream_t stream);
// out[t][b][c] = x[b][t][c]
int batch_to_seq(const float* x, float* out, int T, int n, int C, cudaStream_t stream);
// kin[b][j] = bf16(mem + pos), vin[b][j] = bf16(mem);  j = l*4096+t for the L spatial frames then L*4096+p for the P
// pointer tokens; rows [Lk, ldk) of each batch are zero-filled.
// rows: optional DEVICE array [2][n] = (spatial keys L_b * 4096, pointer tokens P_b) of every object (nullptr: L * 4096 and
// P for all); object b then uses memory[0 .. L_b) and excl[0 .. P_b), packed back to back, zero rows behind.
int memattn_prep_memory(const float* memory, const float* memory_pos, int L, const float* excl, const float* pos_excl,
                        int P, int n, const int* rows, bf16* kin, bf16* vin, long ldk, cudaStream_t stream);
}  // namespace sam2b

namespace sam2b {
// ---- memory encoder helpers (memory_encoder.cu); reference modeling/memory_encoder.py:19-137 ----
// LayerNorm rows followed by exact GELU, bf16 out (LayerNorm2d + GELU of the mask down-sampler in channels-last form)
int layernorm_gelu_rows(const float* x, const float* gamma, const float* beta, float eps, bf16* y, long rows, int C,
                        cudaStream_t stream);
// MaskDownSampler stages 1-3: conv3x3 s2 p1 + LayerNorm2d(eps 1e-6) + GELU, fp32, channels-last outputs.
//   stage 1: mask [n,1,S,S]        -> [n,S/2,S/2,4]
//   stage 2: [n,H,H,4]  NHWC       -> [n,H/2,H/2,16]
//   stage 3: [n,H,H,16] NHWC       -> [n,H/2,H/2,64]
// w: conv weight in the reference layout [Cout,Cin,3,3]; b, ln_w, ln_b: [Cout]
int mask_ds_stage(int stage, const float* in, float* out, int n, int H_in, const float* w, const float* b,
                  const float* ln_w, const float* ln_b, cudaStream_t stream);
// stage 1 from the low-resolution logits [n,256,256]: the 1024^2 input sigmoid(bilinear_x4(low)) * 20 - 10 (or its binarised
// form) is evaluated per tap, never materialised -> [n,512,512,4]
int mask_ds1_from_low(const float* low, float* out, int n, int binarize, const float* w, const float* b, const float* ln_w,
                      const float* ln_b, cudaStream_t stream);
// x [n,64,4096] += no_obj_embed_spatial[c] where object_score_logits[b] <= 0
int add_no_obj_embed(float* x, const float* obj, const float* no_obj, int n, cudaStream_t stream);
// im2col for conv3x3 s2 p1 on NHWC fp32 [n,H,H,C] -> bf16 [n*(H/2)^2, 9*C], k = (ky*3+kx)*C + c
int im2col3x3s2_nhwc(const float* in, bf16* cols, int n, int H, int C, cudaStream_t stream);
int im2col3x3s2_nhwc_bf16(const bf16* in, bf16* cols, int n, int H, int C, cudaStream_t stream);
// [n,C,HW] fp32 -> [n,HW,C] bf16
// in_batch_stride (elements): < 0 = C * HW (contiguous); 0 = one input image repeated for every output batch entry
int nchw_to_nhwc_bf16(const float* in, bf16* out, int n, int C, int HW, cudaStream_t stream, long in_batch_stride = -1);
// CXBlock front half: depthwise 7x7 (pad 3) + LayerNorm2d(eps 1e-6) on NHWC fp32 [n,H,H,256] -> bf16 [n*H*H,256].
// dw_w: [49][256] (tap-major), dw_b/ln_w/ln_b: [256]
int dwconv7_ln_nhwc(const float* x, bf16* y, int n, int H, const float* dw_w, const float* dw_b, const float* ln_w,
                    const float* ln_b, cudaStream_t stream);
}  // namespace sam2b

namespace sam2b {
// ---- SAM heads (sam_heads.cu); reference modeling/sam/{prompt_encoder,transformer,mask_decoder}.py, sam2_base_official.py:338-494 ----
// One small fp32 linear problem of the token side: Y[r][j] = act(sum_k (X[r][k] + X2[r][k]) W[j][k] + b[j]); act: 0 none, 1 ReLU, 2 sigmoid
struct LinJob {
  const float* X = nullptr; long ldx = 0; const float* X2 = nullptr; long ldx2 = 0;
  const float* W = nullptr; long ldw = 0; const float* b = nullptr;
  float* Y = nullptr; long ldy = 0;
  int rows = 0, N = 0, K = 0, act = 0;
};
struct LinJobs { static const int kMax = 16; LinJob j[kMax]; };
int tok_linear(const LinJobs& jobs, int njobs, cudaStream_t st);
// y[r][0..256) = LN(sum_s part[s][r] + bias + res[r]) (g == nullptr: no LayerNorm)
int tok_sum_res_ln(const float* part, long part_stride, int S, const float* bias, const float* res, const float* g,
                   const float* be, float eps, float* y, int rows, cudaStream_t st);
int tok_self_attn(const float* q, const float* k, const float* v, int n, int T, float* o, cudaStream_t st);
// src = image embedding + dense embedding -> keys fp32 [n*4096][256] + split bf16 operand [n*4096][768] = [hi | lo | hi]
int heads_prep_src(const float* pix, long sn, long sc, long sp, const float* dense, const float* nomask, float* keys, bf16* ks,
                   int n, cudaStream_t st);
int ln256_split(const float* x, const float* g, const float* b, float eps, float* y, bf16* ys, long rows, cudaStream_t st);
// PromptEncoder.mask_downscaling: mask [n,1,256,256] -> dense embedding [n,256,64,64]; w = {conv0 w, b, ln1 w, b, conv3 w, b, ln4 w, b,
// conv6 w, b} in the reference's layouts
int mask_downscale(const float* mask, const float* const w[11], float* out, int n, cudaStream_t st);
int heads_tokens(const float* coords, const int* labels, int P, int n, const float* gauss, const float* not_a_point,
                 const float* const pt_emb[4], const float* obj_tok, const float* iou_tok, const float* mask_toks,
                 const float* sparse, int S, float* tok, float* qpe, cudaStream_t st);
// token -> image attention: kv rows of ld floats (k at column 0, v at column 128), q [n*T][128]; part: n*128*T*8*18 floats; out [n*T][128]
int t2i_attention(const float* kv, int ld, const float* q, int n, int T, float* part, float* out, cudaStream_t st);
// image -> token attention: q = columns [q_col, q_col+128) of kvq; tk / tv [n*T][128]; os [n*4096][384] split operand
int i2t_attention(const float* kvq, int ld, int q_col, const float* tk, const float* tv, int n, int T, bf16* os, cudaStream_t st);
int heads_upscale(const float* u1, const float* b1, const float* s1, long s1_bs, const float* ln_g, const float* ln_b,
                  const float* w2, const float* b2, const float* s0, long s0_bs, const float* hyper, float* masks, int* stab,
                  float delta, int n, cudaStream_t st);
int heads_select(const float* iou, const float* obj, const float* ptr4, const int* stab, const float* no_obj_ptr, int multimask,
                 int dynamic, float thresh, float* iou_out, float* ptr_out, float* obj_out, int* sel, int decoder_only, int n,
                 cudaStream_t st);
int heads_write_masks(const float* masks, const int* sel, int multimask, float no_obj_score, float* low_multi, float* low_best,
                      int n, cudaStream_t st);
}  // namespace sam2b
