// Engine: weight store + workspace + the kernel schedules of the three sub-graphs, behind the C ABI of
// include/sam2b.h.  Reference call sites it replaces: ImageEncoder.forward (modeling/backbones/image_encoder.py:29-42)
// = Hiera.forward (hieradet.py:283-299) + FpnNeck.forward (image_encoder.py:102-134) + conv_s0/s1
// (sam2_base_official.py:571-577); MemoryAttention (memory_attention.py:299-349); MemoryEncoder (memory_encoder.py:233-241).
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/sam2b.h"
#include "ops.h"

namespace sam2b {

static thread_local char g_err[1024] = "";
std::atomic<unsigned long long> g_launch_count{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }

// ---- profiling ----
bool g_prof_on = false;
struct ProfRec { char name[96]; double flops, bytes; cudaEvent_t a, b; };
static std::vector<ProfRec> g_prof;
void prof_push(const char* name, double flops, double bytes, cudaStream_t st) {
  ProfRec r;
  snprintf(r.name, sizeof(r.name), "%s", name);
  r.flops = flops; r.bytes = bytes;
  cudaEventCreate(&r.a); cudaEventCreate(&r.b);
  cudaEventRecord(r.a, st);
  g_prof.push_back(r);
}
void prof_pop(cudaStream_t st) { cudaEventRecord(g_prof.back().b, st); }

int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

#define CUDA_OK(expr)                                                              \
  do {                                                                             \
    cudaError_t e__ = (expr);                                                      \
    if (e__ != cudaSuccess) {                                                      \
      set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return 1;                                                                    \
    }                                                                              \
  } while (0)
#define TRY(expr) do { if (expr) return 1; } while (0)

// Makes `dev` the calling thread's current device for the lifetime of the guard and restores the previous one afterwards,
// so that a call on an engine of another GPU does not silently change the caller's (and torch's) current device.
struct DeviceGuard {
  int prev = -1;
  bool switched = false;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int dev) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != dev) { err = cudaSetDevice(dev); switched = (err == cudaSuccess); }
  }
  ~DeviceGuard() { if (switched) cudaSetDevice(prev); }
  DeviceGuard(const DeviceGuard&) = delete;
  DeviceGuard& operator=(const DeviceGuard&) = delete;
};
#define ENTER_DEVICE(dev)          \
  DeviceGuard device_guard__(dev); \
  CUDA_OK(device_guard__.err)

struct Tensor {
  float* f32 = nullptr;
  bf16* b16 = nullptr;
  std::vector<int64_t> shape;
  int64_t numel = 0;
};

struct Buf {
  void* p = nullptr;
  size_t bytes = 0;
};

struct BlockCfg { int dim, dim_out, heads, window, q_pool; };

}  // namespace sam2b

using namespace sam2b;

struct sam2b_engine {
  int device = 0;
  bool finalized = false;
  bool has_heads = false;                 // the SAM-heads tensors were registered (sam2b_sam_heads is available)
  std::map<std::string, Tensor> w;
  std::map<std::string, Buf> ws;   // named workspace buffers, grow-only
  std::vector<BlockCfg> blocks;
  cudaStream_t copy_stream = nullptr;      // host -> device
  cudaStream_t copy_stream2 = nullptr;     // device -> host (PCIe is full duplex: the two directions overlap)
  cudaEvent_t ev[12] = {};
  // Calls on one engine share the named workspaces, so they must not overlap on the DEVICE either (two host threads with
  // private streams drive one predictor in the reference's video_multi_thread.py:36-60).  `mu` serialises the host enqueue;
  // `done` is recorded on the caller's stream at the end of every call and the next call's stream waits for it first: the
  // same hand-off the reference executor gets from its private work stream (tensorrt_executor.py:248-252,288-291).
  std::mutex mu;
  cudaEvent_t done = nullptr;
  bool done_valid = false;
  unsigned long long ws_generation = 0;   // bumped whenever a workspace buffer is re-allocated
  unsigned long long host_it = 0;         // micro-batches pushed through the host-buffer path so far (staging slot = host_it & 1)

  Tensor* find(const std::string& n) {
    auto it = w.find(n);
    return it == w.end() ? nullptr : &it->second;
  }
};

namespace sam2b {

// Serialises one engine call against every other call on the same engine, on the host (mutex) and on the device (event
// chain across the callers' streams).  Constructed after ENTER_DEVICE; `begin()` reports CUDA errors.
struct EngineCall {
  sam2b_engine* e;
  cudaStream_t st;
  std::unique_lock<std::mutex> lk;
  bool capturing = false;
  EngineCall(sam2b_engine* eng, cudaStream_t s) : e(eng), st(s), lk(eng->mu) {}
  int begin() {
    // Under CUDA graph capture the hand-off events must stay out of the graph (waiting on an event recorded outside the
    // capture is an error, and an event recorded inside would become a captured one): a caller that replays a captured
    // sequence brackets the replay with sam2b_stream_acquire / sam2b_stream_release instead.
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    CUDA_OK(cudaStreamIsCapturing(st, &cs));
    capturing = (cs != cudaStreamCaptureStatusNone);
    if (!capturing && e->done_valid) CUDA_OK(cudaStreamWaitEvent(st, e->done, 0));
    return 0;
  }
  ~EngineCall() {
    if (!capturing && cudaEventRecord(e->done, st) == cudaSuccess) e->done_valid = true;
  }
};
#define ENGINE_CALL(h, st)        \
  EngineCall engine_call__(h, st); \
  TRY(engine_call__.begin())

static int ws_get(sam2b_engine* e, const char* name, size_t bytes, void** out) {
  Buf& b = e->ws[name];
  if (b.bytes < bytes) {
    ++e->ws_generation;               // captured graphs that baked the old address must be re-captured (sam2b_workspace_generation)
    if (b.p) CUDA_OK(cudaFree(b.p));
    b.p = nullptr; b.bytes = 0;
    CUDA_OK(cudaMalloc(&b.p, bytes));
    b.bytes = bytes;
  }
  *out = b.p;
  return 0;
}
template <typename T>
static int ws_get_t(sam2b_engine* e, const char* name, size_t count, T** out) {
  void* p = nullptr;
  TRY(ws_get(e, name, count * sizeof(T), &p));
  *out = reinterpret_cast<T*>(p);
  return 0;
}

static void build_block_cfgs(std::vector<BlockCfg>& out) {
  // Hiera-L: configs/sam2.1/sam2.1_hiera_l.yaml:10-16, hieradet.py:233-260 (window lags one block at stage changes)
  const int stages[4] = {2, 6, 36, 4};
  const int window_spec[4] = {8, 4, 16, 8};
  const int stage_ends[4] = {1, 7, 43, 47};
  int dim = 144, heads = 2, cur_stage = 1;
  for (int i = 0; i < stages[0] + stages[1] + stages[2] + stages[3]; ++i) {
    int dim_out = dim;
    int window = window_spec[cur_stage - 1];
    if (i == 23 || i == 33 || i == 43) window = 0;
    bool after_end = false;
    for (int s : stage_ends) after_end |= (i - 1 == s);
    if (after_end) { dim_out = dim * 2; heads *= 2; cur_stage += 1; }
    BlockCfg c{dim, dim_out, heads, window, (i == 2 || i == 8 || i == 44) ? 1 : 0};
    out.push_back(c);
    dim = dim_out;
  }
}

static int need(sam2b_engine* e, const std::string& name, Tensor** t, bool want_bf16) {
  Tensor* x = e->find(name);
  if (!x) { set_error("missing tensor '%s' (call sam2b_set_tensor for every state_dict / derived entry)", name.c_str()); return 1; }
  if (want_bf16 && !x->b16) {
    CUDA_OK(cudaMalloc(&x->b16, (size_t)x->numel * sizeof(bf16)));
    TRY(cast_f32_bf16(x->f32, x->b16, x->numel, 0));
  }
  if (t) *t = x;
  return 0;
}

static const float* W32(sam2b_engine* e, const std::string& n) { return e->w.at(n).f32; }
static const bf16* W16(sam2b_engine* e, const std::string& n) { return e->w.at(n).b16; }

static int finalize(sam2b_engine* e) {
  ENTER_DEVICE(e->device);
  const std::string t = "image_encoder.trunk.";
  for (size_t i = 0; i < e->blocks.size(); ++i) {
    const std::string b = t + "blocks." + std::to_string(i) + ".";
    for (const char* ln : {"norm1.", "norm2."}) {
      TRY(need(e, b + ln + "weight", nullptr, false));
      TRY(need(e, b + ln + "bias", nullptr, false));
    }
    for (const char* lin : {"attn.qkv.", "attn.proj.", "mlp.layers.0.", "mlp.layers.1."}) {
      TRY(need(e, b + lin + "weight", nullptr, true));
      TRY(need(e, b + lin + "bias", nullptr, false));
    }
    if (e->blocks[i].dim != e->blocks[i].dim_out) {
      TRY(need(e, b + "proj.weight", nullptr, true));
      TRY(need(e, b + "proj.bias", nullptr, false));
    }
    // LayerNorm-folded forms of the norm -> Linear pairs (weights.py derive())
    const std::string d = "derived.blocks." + std::to_string(i) + ".";
    for (const char* lin : {"qkv.", "fc1.", "skip."}) {
      if (std::string(lin) == "skip." && e->blocks[i].dim == e->blocks[i].dim_out) continue;
      TRY(need(e, d + lin + "weight", nullptr, true));
      TRY(need(e, d + lin + "bias", nullptr, false));
    }
  }
  TRY(need(e, "derived.patch_embed.weight", nullptr, true));   // [144,160] (147 taps zero padded)
  TRY(need(e, t + "patch_embed.proj.bias", nullptr, false));
  TRY(need(e, "derived.pos_embed", nullptr, false));           // [256*256,144] NHWC
  TRY(need(e, "derived.neck.lat3.weight", nullptr, true));     // [256,1152]
  TRY(need(e, "derived.neck.lat3.bias", nullptr, false));
  TRY(need(e, "derived.neck.lat2.weight", nullptr, true));     // [256,576]
  TRY(need(e, "derived.neck.lat2.bias", nullptr, false));
  TRY(need(e, "derived.neck.lat2.bias_nomem", nullptr, false));
  TRY(need(e, "derived.neck.s1.weight", nullptr, true));       // conv_s1 o lateral: [64,288]
  TRY(need(e, "derived.neck.s1.bias", nullptr, false));
  TRY(need(e, "derived.neck.s0.weight", nullptr, true));       // conv_s0 o lateral: [32,144]
  TRY(need(e, "derived.neck.s0.bias", nullptr, false));
  // ---- memory attention (memory_attention.py:18-109, transformer.py:297-424) ----
  for (int l = 0; l < 4; ++l) {
    const std::string p = "memory_attention.layers." + std::to_string(l) + ".";
    const std::string d = "derived.memattn." + std::to_string(l) + ".";
    TRY(need(e, d + "sa_qk.weight", nullptr, true));          // [512,256] = cat(q_proj, k_proj)
    TRY(need(e, d + "sa_qk.bias", nullptr, false));
    TRY(need(e, d + "ca_ov.weight", nullptr, true));          // [256,64] = out_proj o v_proj of the cross-attention
    TRY(need(e, d + "ca_ov.bias", nullptr, false));
    for (const char* lin : {"self_attn.v_proj.", "self_attn.out_proj.", "cross_attn_image.q_proj.", "cross_attn_image.k_proj.",
                            "cross_attn_image.v_proj.", "cross_attn_image.out_proj.", "linear1.", "linear2."}) {
      TRY(need(e, p + lin + "weight", nullptr, true));
      TRY(need(e, p + lin + "bias", nullptr, false));
    }
    for (const char* ln : {"norm1.", "norm2.", "norm3."}) {
      TRY(need(e, p + ln + "weight", nullptr, false));
      TRY(need(e, p + ln + "bias", nullptr, false));
    }
  }
  TRY(need(e, "memory_attention.norm.weight", nullptr, false));
  TRY(need(e, "memory_attention.norm.bias", nullptr, false));
  // ---- memory encoder (memory_encoder.py) ----
  {
    const std::string m = "memory_encoder.", en = m + "mask_downsampler.encoder.";
    for (int j = 0; j < 4; ++j) {
      if (j < 3) TRY(need(e, en + std::to_string(3 * j) + ".weight", nullptr, false));
      TRY(need(e, en + std::to_string(3 * j) + ".bias", nullptr, false));
      TRY(need(e, en + std::to_string(3 * j + 1) + ".weight", nullptr, false));
      TRY(need(e, en + std::to_string(3 * j + 1) + ".bias", nullptr, false));
    }
    TRY(need(e, "derived.memenc.ds4.weight", nullptr, true));   // [256, (ky,kx,cin)=576]
    TRY(need(e, "derived.memenc.ds3.weight", nullptr, true));   // [64, (ky,kx,cin)=144]
    for (const char* c : {"mask_downsampler.encoder.12.", "pix_feat_proj.", "out_proj."}) {
      TRY(need(e, m + c + "weight", nullptr, true));
      TRY(need(e, m + c + "bias", nullptr, false));
    }
    for (int l = 0; l < 2; ++l) {
      const std::string f = m + "fuser.layers." + std::to_string(l) + ".";
      TRY(need(e, "derived.memenc.fuser." + std::to_string(l) + ".dw.weight", nullptr, false));   // [49,256]
      TRY(need(e, f + "dwconv.bias", nullptr, false));
      TRY(need(e, f + "norm.weight", nullptr, false));
      TRY(need(e, f + "norm.bias", nullptr, false));
      TRY(need(e, f + "gamma", nullptr, false));
      for (const char* c : {"pwconv1.", "pwconv2."}) {
        TRY(need(e, f + c + "weight", nullptr, true));
        TRY(need(e, f + c + "bias", nullptr, false));
      }
    }
  }
  TRY(need(e, "derived.rope.cs", nullptr, false));             // [4096,256]: (cos, sin) interleaved per column pair
  // ---- SAM heads (optional: an engine built from encoder-only weights has no sam2b_sam_heads) ----
  if (e->find("sam_mask_decoder.iou_token.weight") && e->find("derived.heads.up1.weight")) {
    const std::string md = "sam_mask_decoder.", tr = md + "transformer.", pe = "sam_prompt_encoder.";
    for (const char* nm : {"pe_layer.positional_encoding_gaussian_matrix", "not_a_point_embed.weight", "no_mask_embed.weight",
                           "point_embeddings.0.weight", "point_embeddings.1.weight", "point_embeddings.2.weight",
                           "point_embeddings.3.weight"})
      TRY(need(e, pe + nm, nullptr, false));
    for (const char* nm : {"obj_score_token.weight", "iou_token.weight", "mask_tokens.weight", "output_upscaling.0.bias",
                           "output_upscaling.1.weight", "output_upscaling.1.bias", "output_upscaling.3.bias"})
      TRY(need(e, md + nm, nullptr, false));
    auto attn = [&](const std::string& p) -> int {
      for (const char* nm : {"q_proj.", "k_proj.", "v_proj.", "out_proj."}) {
        TRY(need(e, p + nm + "weight", nullptr, false));
        TRY(need(e, p + nm + "bias", nullptr, false));
      }
      return 0;
    };
    for (int l = 0; l < 2; ++l) {
      const std::string p = tr + "layers." + std::to_string(l) + ".", d = "derived.heads." + std::to_string(l) + ".";
      TRY(attn(p + "self_attn."));
      TRY(attn(p + "cross_attn_token_to_image."));
      TRY(attn(p + "cross_attn_image_to_token."));
      for (const char* nm : {"norm1.", "norm2.", "norm3.", "norm4.", "mlp.layers.0.", "mlp.layers.1."}) {
        TRY(need(e, p + nm + "weight", nullptr, false));
        TRY(need(e, p + nm + "bias", nullptr, false));
      }
      TRY(need(e, d + "kvq.weight", nullptr, true));       // [384, 768] split form of (t2i k | t2i v | i2t q)
      TRY(need(e, d + "kvq.table", nullptr, false));       // [4096, 384] = dense_pe W^T + bias of the same three projections
      TRY(need(e, d + "i2t_out.weight", nullptr, true));   // [256, 384] split form of the i2t out_proj
    }
    TRY(attn(tr + "final_attn_token_to_image."));
    TRY(need(e, tr + "norm_final_attn.weight", nullptr, false));
    TRY(need(e, tr + "norm_final_attn.bias", nullptr, false));
    TRY(need(e, "derived.heads.final.kv.weight", nullptr, true));   // [256, 768]
    TRY(need(e, "derived.heads.final.kv.table", nullptr, false));   // [4096, 256]
    TRY(need(e, "derived.heads.up1.weight", nullptr, true));        // [256 = (dy,dx,c), 768] ConvTranspose2d(256 -> 64) as a GEMM
    TRY(need(e, "derived.heads.up2.weight", nullptr, false));       // [64][4 = (ey,ex)][32]
    for (const char* head : {"sam_mask_decoder.iou_prediction_head.", "sam_mask_decoder.pred_obj_score_head.", "obj_ptr_proj.",
                             "sam_mask_decoder.output_hypernetworks_mlps.0.", "sam_mask_decoder.output_hypernetworks_mlps.1.",
                             "sam_mask_decoder.output_hypernetworks_mlps.2.", "sam_mask_decoder.output_hypernetworks_mlps.3."})
      for (int i = 0; i < 3; ++i) {
        TRY(need(e, std::string(head) + "layers." + std::to_string(i) + ".weight", nullptr, false));
        TRY(need(e, std::string(head) + "layers." + std::to_string(i) + ".bias", nullptr, false));
      }
    TRY(need(e, "no_obj_ptr", nullptr, false));
    for (const char* nm : {"0.weight", "0.bias", "1.weight", "1.bias", "3.weight", "3.bias", "4.weight", "4.bias", "6.weight", "6.bias"})
      TRY(need(e, pe + "mask_downscaling." + nm, nullptr, false));
    e->has_heads = true;
  }
  CUDA_OK(cudaStreamSynchronize(0));
  e->finalized = true;
  return 0;
}

// One micro-batch (n images resident on the device) through trunk + neck.
// early_ev (optional, 2 events): recorded on `st` as soon as fpn0 / fpn1 are complete -- the level-0/1 neck outputs depend
// only on the stage-1/2 taps, so they are produced right after blocks 1 and 7 and can leave the device while the rest of
// the trunk is still running.
static int image_encode_mb(sam2b_engine* e, const float* img, int n, int normalised, int add_no_mem, float* fpn0,
                           float* fpn1, float* fpn2, cudaStream_t st, cudaEvent_t* early_ev = nullptr) {
  const long T0 = 65536L * n;
  float *X, *X2, *SC, *LAT3;
  bf16 *XN, *QKV, *ATT, *HID, *COLS, *TAP[4];
  TRY(ws_get_t(e, "x", (size_t)T0 * 144, &X));
  TRY(ws_get_t(e, "x2", (size_t)T0 * 144, &X2));
  TRY(ws_get_t(e, "sc", (size_t)T0 * 288, &SC));
  TRY(ws_get_t(e, "xn", (size_t)T0 * 144, &XN));
  TRY(ws_get_t(e, "qkv", (size_t)T0 * 864, &QKV));
  TRY(ws_get_t(e, "att", (size_t)T0 * 144, &ATT));
  TRY(ws_get_t(e, "hid", (size_t)T0 * 576, &HID));
  TRY(ws_get_t(e, "cols", (size_t)T0 * 160, &COLS));
  TRY(ws_get_t(e, "tap0", (size_t)T0 * 144, &TAP[0]));
  TRY(ws_get_t(e, "tap1", (size_t)T0 / 4 * 288, &TAP[1]));
  TRY(ws_get_t(e, "tap2", (size_t)T0 / 16 * 576, &TAP[2]));
  TRY(ws_get_t(e, "tap3", (size_t)T0 / 64 * 1152, &TAP[3]));
  TRY(ws_get_t(e, "lat3", (size_t)n * 256 * 1024, &LAT3));

  // ---- FpnNeck (+ conv_s0 / conv_s1 folded into the level-0/1 laterals): out[b][co][hw] = W[co][:] . x[b][hw][:] ----
  struct NeckJob { const char* w; const char* bias; const bf16* x; int cin, cout, hw; float* out; };
  const NeckJob jobs[4] = {
      {"derived.neck.lat3.weight", "derived.neck.lat3.bias", TAP[3], 1152, 256, 1024, LAT3},
      {"derived.neck.lat2.weight", add_no_mem ? "derived.neck.lat2.bias_nomem" : "derived.neck.lat2.bias", TAP[2], 576, 256, 4096, fpn2},
      {"derived.neck.s1.weight", "derived.neck.s1.bias", TAP[1], 288, 64, 16384, fpn1},
      {"derived.neck.s0.weight", "derived.neck.s0.bias", TAP[0], 144, 32, 65536, fpn0},
  };
  auto neck = [&](int k) -> int {
    const NeckJob& j = jobs[k];
    GemmArgs g;
    g.A = W16(e, j.w); g.lda = j.cin; g.M = j.cout; g.K = j.cin;   // weights shared by the batch
    g.B = j.x; g.ldb = j.cin; g.N = j.hw; g.b_batch_stride = (long)j.hw * j.cin;
    g.batch = n;
    g.bias = W32(e, j.bias); g.bias_along_m = 1;
    g.out_f32 = j.out; g.ldc = j.hw; g.c_batch_stride = (long)j.cout * j.hw;
    return gemm_bf16(g, st);
  };
  bool early_done = false;
  const std::string t = "image_encoder.trunk.";
  static int ln_fold = -1;
  if (ln_fold < 0) { const char* ev = getenv("SAM2B_LN_FOLD"); ln_fold = ev ? atoi(ev) : 1; }
  if (ln_fold) {
    // ============ LayerNorm never runs as a kernel: every GEMM that writes the fp32 residual stream also emits its bf16
    // copy and per-row (sum, sum x^2) partials; the norm -> Linear consumers fold the normalisation into their epilogue.
    bf16 *XB, *XB2;
    float *ST1, *ST2;
    // partial-sum slots: 2 per N tile of the producing GEMM (BLOCK_N >= 128): <= 4 / 6 / 10 / 18 per row in stages 1..4,
    // whose row counts shrink 4x per stage -> T0 * 8 floats cover every stage
    const size_t kStatFloats = (size_t)T0 * 8;
    auto stat_cap = [&](long rows) { return (int)(kStatFloats / ((size_t)rows * 2)); };
    TRY(ws_get_t(e, "xb", (size_t)T0 * 144, &XB));
    TRY(ws_get_t(e, "xb2", (size_t)T0 * 144, &XB2));
    TRY(ws_get_t(e, "st1", kStatFloats, &ST1));
    TRY(ws_get_t(e, "st2", kStatFloats, &ST2));
    TRY(patch_im2col(img, COLS, n, 1024, normalised ? 0 : 1, st));
    int parts1 = 0, parts2 = 0;         // partial-sum slots per row in ST1 (block input) / ST2 (after attention)
    // stage-1 MLPs (C = 144) run as ONE kernel (mlp_fused.cu: the hidden layer never leaves the SM; 1.1 - 1.3 ms against 2.0 ms
    // for the fc1 / fc2 GEMM pair per block per 32 images).  SAM2B_FUSED_MLP=0 selects the GEMM pair (read per call: the
    // equivalence test switches it)
    const char* fused_env = getenv("SAM2B_FUSED_MLP");
    const bool fused_mlp = !fused_env || atoi(fused_env) != 0;
    {
      GemmArgs g;
      g.A = COLS; g.lda = 160; g.M = (int)T0; g.K = 160;
      g.B = W16(e, "derived.patch_embed.weight"); g.ldb = 160; g.N = 144;
      g.bias = W32(e, t + "patch_embed.proj.bias");
      g.residual = W32(e, "derived.pos_embed"); g.ld_res = 144; g.res_row_mod = 65536;
      g.out_f32 = X; g.ldc = 144;
      g.out2_bf16 = XB; g.ldc2 = 144; g.stats_out = ST1; g.stats_capacity_parts = stat_cap(T0); g.stats_parts_out = &parts1;
      TRY(gemm_bf16(g, st));
    }
    const bf16* xb_in = XB;            // bf16 copy of the current residual stream (block input)
    int H = 256, tap = 0;
    for (size_t i = 0; i < e->blocks.size(); ++i) {
      const BlockCfg& c = e->blocks[i];
      const std::string b = t + "blocks." + std::to_string(i) + ".";
      const std::string d = "derived.blocks." + std::to_string(i) + ".";
      const long T = (long)n * H * H;
      auto ln_linear = [&](const bf16* A, int K, const std::string& w, int N, const float* stats, int parts, int act,
                           float* o32, bf16* o16, long M) -> int {
        GemmArgs g;
        g.A = A; g.lda = K; g.M = (int)M; g.K = K;
        g.B = W16(e, w + "weight"); g.ldb = K; g.N = N;
        g.bias = W32(e, w + "bias"); g.act = act;
        g.ln_stats = stats; g.ln_parts = parts; g.ln_eps = 1e-6f;
        g.out_f32 = o32; g.out_bf16 = o16; g.ldc = N;
        return gemm_bf16(g, st);
      };
      float* shortcut = X;
      const bool trans = c.dim != c.dim_out;
      if (trans) {
        TRY(ln_linear(xb_in, c.dim, d + "skip.", c.dim_out, ST1, parts1, ACT_NONE, SC, nullptr, T));
        TRY(maxpool2x2_nhwc(SC, X2, n, H, H, c.dim_out, st));
        shortcut = X2;
      }
      TRY(ln_linear(xb_in, c.dim, d + "qkv.", 3 * c.dim_out, ST1, parts1, ACT_NONE, nullptr, QKV, T));
      TRY(window_attention_hd72(QKV, ATT, n, H, H, c.heads, c.window ? c.window : H, c.q_pool, st));
      if (c.q_pool) H /= 2;
      const long Tq = (long)n * H * H;
      {
        GemmArgs g;
        g.A = ATT; g.lda = c.dim_out; g.M = (int)Tq; g.K = c.dim_out;
        g.B = W16(e, b + "attn.proj.weight"); g.ldb = c.dim_out; g.N = c.dim_out;
        g.bias = W32(e, b + "attn.proj.bias");
        g.residual = shortcut; g.ld_res = c.dim_out;
        g.out_f32 = shortcut; g.ldc = c.dim_out;
        g.out2_bf16 = XB2; g.ldc2 = c.dim_out; g.stats_out = ST2; g.stats_capacity_parts = stat_cap(Tq); g.stats_parts_out = &parts2;
        TRY(gemm_bf16(g, st));
      }
      if (trans) { float* tmp = X; X = X2; X2 = tmp; }
      const bool is_tap = (i == 1 || i == 7 || i == 43 || i == 47);
      bf16* xb_out = is_tap ? TAP[tap] : XB;      // stage taps ARE the bf16 copy of the stream (no cast kernel)
      if (fused_mlp && mlp_fused_supported(c.dim_out)) {
        // stages 1 and 2: fc1 -> GELU -> fc2 in one kernel, the hidden layer stays in tensor memory
        TRY(mlp_fused(XB2, ST2, parts2, 1e-6f, W16(e, d + "fc1.weight"), W32(e, d + "fc1.bias"), W16(e, b + "mlp.layers.1.weight"),
                      W32(e, b + "mlp.layers.1.bias"), X, xb_out, ST1, Tq, c.dim_out, st));
        parts1 = 1;              // the read-out warp that owns a row sums the whole row
      } else {
      TRY(ln_linear(XB2, c.dim_out, d + "fc1.", 4 * c.dim_out, ST2, parts2, ACT_GELU, nullptr, HID, Tq));
      {
        GemmArgs g;
        g.A = HID; g.lda = 4 * c.dim_out; g.M = (int)Tq; g.K = 4 * c.dim_out;
        g.B = W16(e, b + "mlp.layers.1.weight"); g.ldb = 4 * c.dim_out; g.N = c.dim_out;
        g.bias = W32(e, b + "mlp.layers.1.bias");
        g.residual = X; g.ld_res = c.dim_out;
        g.out_f32 = X; g.ldc = c.dim_out;
        g.out2_bf16 = xb_out; g.ldc2 = c.dim_out; g.stats_out = ST1; g.stats_capacity_parts = stat_cap(Tq); g.stats_parts_out = &parts1;
        TRY(gemm_bf16(g, st));
      }
      }
      xb_in = xb_out;
      if (is_tap) ++tap;
      if (i == 1) {          // level 0 (conv_s0 o lateral) straight from the stage-1 tap
        TRY(neck(3));
        if (early_ev) CUDA_OK(cudaEventRecord(early_ev[0], st));
      } else if (i == 7) {   // level 1 from the stage-2 tap
        TRY(neck(2));
        if (early_ev) CUDA_OK(cudaEventRecord(early_ev[1], st));
        early_done = true;
      }
    }
  } else {
  // ---- patch embed (im2col + GEMM) with the constant position embedding added in the epilogue ----
  TRY(patch_im2col(img, COLS, n, 1024, normalised ? 0 : 1, st));
  {
    GemmArgs g;
    g.A = COLS; g.lda = 160; g.M = (int)T0; g.K = 160;
    g.B = W16(e, "derived.patch_embed.weight"); g.ldb = 160; g.N = 144;
    g.bias = W32(e, t + "patch_embed.proj.bias");
    g.residual = W32(e, "derived.pos_embed"); g.ld_res = 144; g.res_row_mod = 65536;
    g.out_f32 = X; g.ldc = 144;
    TRY(gemm_bf16(g, st));
  }
  int H = 256, tap = 0;
  for (size_t i = 0; i < e->blocks.size(); ++i) {
    const BlockCfg& c = e->blocks[i];
    const std::string b = t + "blocks." + std::to_string(i) + ".";
    const long T = (long)n * H * H;
    TRY(layernorm_rows(X, W32(e, b + "norm1.weight"), W32(e, b + "norm1.bias"), 1e-6f, XN, nullptr, T, c.dim, st));
    float* shortcut = X;
    const bool trans = c.dim != c.dim_out;
    if (trans) {
      GemmArgs g;
      g.A = XN; g.lda = c.dim; g.M = (int)T; g.K = c.dim;
      g.B = W16(e, b + "proj.weight"); g.ldb = c.dim; g.N = c.dim_out;
      g.bias = W32(e, b + "proj.bias");
      g.out_f32 = SC; g.ldc = c.dim_out;
      TRY(gemm_bf16(g, st));
      TRY(maxpool2x2_nhwc(SC, X2, n, H, H, c.dim_out, st));
      shortcut = X2;
    }
    {
      GemmArgs g;
      g.A = XN; g.lda = c.dim; g.M = (int)T; g.K = c.dim;
      g.B = W16(e, b + "attn.qkv.weight"); g.ldb = c.dim; g.N = 3 * c.dim_out;
      g.bias = W32(e, b + "attn.qkv.bias");
      g.out_bf16 = QKV; g.ldc = 3 * c.dim_out;
      TRY(gemm_bf16(g, st));
    }
    TRY(window_attention_hd72(QKV, ATT, n, H, H, c.heads, c.window ? c.window : H, c.q_pool, st));
    if (c.q_pool) H /= 2;
    const long Tq = (long)n * H * H;
    {
      GemmArgs g;
      g.A = ATT; g.lda = c.dim_out; g.M = (int)Tq; g.K = c.dim_out;
      g.B = W16(e, b + "attn.proj.weight"); g.ldb = c.dim_out; g.N = c.dim_out;
      g.bias = W32(e, b + "attn.proj.bias");
      g.residual = shortcut; g.ld_res = c.dim_out;
      g.out_f32 = shortcut; g.ldc = c.dim_out;
      TRY(gemm_bf16(g, st));
    }
    if (trans) { float* tmp = X; X = X2; X2 = tmp; }
    TRY(layernorm_rows(X, W32(e, b + "norm2.weight"), W32(e, b + "norm2.bias"), 1e-6f, XN, nullptr, Tq, c.dim_out, st));
    {
      GemmArgs g;
      g.A = XN; g.lda = c.dim_out; g.M = (int)Tq; g.K = c.dim_out;
      g.B = W16(e, b + "mlp.layers.0.weight"); g.ldb = c.dim_out; g.N = 4 * c.dim_out;
      g.bias = W32(e, b + "mlp.layers.0.bias"); g.act = ACT_GELU;
      g.out_bf16 = HID; g.ldc = 4 * c.dim_out;
      TRY(gemm_bf16(g, st));
    }
    {
      GemmArgs g;
      g.A = HID; g.lda = 4 * c.dim_out; g.M = (int)Tq; g.K = 4 * c.dim_out;
      g.B = W16(e, b + "mlp.layers.1.weight"); g.ldb = 4 * c.dim_out; g.N = c.dim_out;
      g.bias = W32(e, b + "mlp.layers.1.bias");
      g.residual = X; g.ld_res = c.dim_out;
      g.out_f32 = X; g.ldc = c.dim_out;
      TRY(gemm_bf16(g, st));
    }
    if (i == 1 || i == 7 || i == 43 || i == 47) {
      TRY(cast_f32_bf16(X, TAP[tap], Tq * c.dim_out, st));
      ++tap;
    }
  }
  }
  TRY(neck(0));
  TRY(neck(1));
  if (!early_done) {
    TRY(neck(2));
    TRY(neck(3));
    if (early_ev) { CUDA_OK(cudaEventRecord(early_ev[0], st)); CUDA_OK(cudaEventRecord(early_ev[1], st)); }
  }
  TRY(upsample2x_add_nchw(fpn2, LAT3, n, 256, 64, 64, st));
  return 0;
}

// ------------------------------------------------------------------------------------------------------------
// MemoryAttention (boundary E3) for a micro-batch of n objects.  All contractions run on the tcgen05 GEMM:
// projections and FFN run on the tcgen05 GEMM (K1), softmax(q k^T / 16) v on the hd-256 flash kernel (K4): the score and
// probability matrices never reach HBM.
// ------------------------------------------------------------------------------------------------------------
// rows_host (optional, [2][n]): frames L_b and pointer tokens P_b of every object (objects added on different frames have
// different banks); nullptr = L and P for all.
static int memory_attention_mb(sam2b_engine* e, const float* curr, const float* curr_pos, const float* memory,
                               const float* memory_pos, int L, const float* excl, const float* pos_excl, int P, int n,
                               int n_total, const int* rows_host, float* out, cudaStream_t st, const int* counts_dev = nullptr) {
  // curr/memory pointers address object 0 of this micro-batch inside tensors whose batch dimension is n_total
  const int T = 4096, C = 256;
  const long Lk = (long)L * 4096 + P;
  const long Lkp = (Lk + 127) / 128 * 128;
  const long TT = (long)n * T;
  float *X, *OUT;
  bf16 *XN, *Qb, *Kb, *Vb, *KIN, *VIN, *Kc, *O, *O64, *HID;
  // per-object row counts on the device: [0][b] spatial keys, [1][b] pointer tokens, [2][b] keys in total
  int* ROWS = nullptr;
  if (rows_host) {
    TRY(ws_get_t(e, "ma_rows", (size_t)3 * n, &ROWS));
    std::vector<int> r(3 * (size_t)n);
    for (int b = 0; b < n; ++b) {
      r[b] = rows_host[b] * 4096;
      r[n + b] = rows_host[n + b];
      r[2 * n + b] = r[b] + r[n + b];
    }
    // pageable source: staged by the runtime before the call returns, stream-ordered afterwards
    CUDA_OK(cudaMemcpyAsync(ROWS, r.data(), r.size() * sizeof(int), cudaMemcpyHostToDevice, st));
  } else if (counts_dev) {      // the counts live on the device (a captured graph reads them from its control vector)
    TRY(ws_get_t(e, "ma_rows", (size_t)3 * n, &ROWS));
    TRY(memattn_rows_from_device(counts_dev, ROWS, n, L, P, st));
  }
  TRY(ws_get_t(e, "ma_x", (size_t)TT * C, &X));
  TRY(ws_get_t(e, "ma_out", (size_t)TT * C, &OUT));
  TRY(ws_get_t(e, "ma_xn", (size_t)TT * C, &XN));
  TRY(ws_get_t(e, "ma_qb", (size_t)TT * C, &Qb));
  TRY(ws_get_t(e, "ma_kb", (size_t)TT * C, &Kb));
  TRY(ws_get_t(e, "ma_vb", (size_t)TT * C, &Vb));
  TRY(ws_get_t(e, "ma_kin", (size_t)n * Lkp * 64, &KIN));
  TRY(ws_get_t(e, "ma_vin", (size_t)n * Lkp * 64, &VIN));
  TRY(ws_get_t(e, "ma_kc", (size_t)n * Lkp * C, &Kc));
  TRY(ws_get_t(e, "ma_o", (size_t)TT * C, &O));
  TRY(ws_get_t(e, "ma_o64", (size_t)TT * 64, &O64));
  TRY(ws_get_t(e, "ma_hid", (size_t)TT * 2048, &HID));
  const float* rope_cs = W32(e, "derived.rope.cs");

  if (n == n_total) {
    TRY(memattn_prep_x(curr, curr_pos, X, T, n, C, st));
    TRY(memattn_prep_memory(memory, memory_pos, L, excl, pos_excl, P, n, ROWS, KIN, VIN, Lkp, st));
  } else {
    set_error("memory_attention: object micro-batching of a strided batch is not supported (n=%d of %d)", n, n_total);
    return 1;
  }

  auto linear = [&](const bf16* A, long lda, int M, const std::string& wname, int N, int K, int act, const float* residual,
                    float* o32, bf16* o16, long ldc) -> int {
    GemmArgs g;
    g.A = A; g.lda = lda; g.M = M; g.K = K;
    g.B = W16(e, wname + "weight"); g.ldb = K; g.N = N;
    g.bias = W32(e, wname + "bias"); g.act = act;
    g.residual = residual; g.ld_res = ldc;
    g.out_f32 = o32; g.out_bf16 = o16; g.ldc = ldc;
    return gemm_bf16(g, st);
  };
  // attention core (K4, attn_hd256.cu) on prepared operands: q [n][T][256], k / v [n][rows_alloc][256], `keys` valid rows
  // projection + RoPE + bf16 cast in one GEMM (rotation in the epilogue, table rows = token position m % 4096):
  // rows [w0, w0 + N) of the weight `wname`, A [batch][M][K], out [batch][M][N]; rows at or past rope_rows stay un-rotated
  auto linear_rope = [&](const bf16* A, long a_batch_stride, int M, int batch, const std::string& wname, int w0, int K, bf16* o16,
                         long c_batch_stride, int rope_rows, const int* rope_rows_dev) -> int {
    GemmArgs g;
    g.A = A; g.lda = K; g.a_batch_stride = a_batch_stride; g.M = M; g.K = K; g.batch = batch;
    g.B = W16(e, wname + "weight") + (size_t)w0 * K; g.ldb = K; g.N = C;
    g.bias = W32(e, wname + "bias") + w0;
    g.out_bf16 = o16; g.ldc = C; g.c_batch_stride = c_batch_stride;
    g.rope_table = rope_cs; g.rope_positions = 4096; g.rope_rows = rope_rows; g.rope_rows_dev = rope_rows_dev;
    return gemm_bf16(g, st);
  };
  auto attend = [&](const bf16* q, const bf16* k, const bf16* v, long rows_alloc, long keys) -> int {
    const int splits = flash_attention_hd256_splits(n, T, keys);
    float* part = nullptr;
    TRY(ws_get_t(e, "ma_part", (size_t)splits * n * T * (C + 2), &part));
    return flash_attention_hd256(q, (long)T * C, k, rows_alloc * C, v, rows_alloc * C, n, T, keys, part,
                                 part + (size_t)splits * n * T * C, O, st);
  };
  // cross-attention core (K4b, attn_hd256_v64.cu): the values are the un-projected 64-wide memory rows (VIN), v_proj is applied
  // together with out_proj afterwards ("ca_ov"); per-object key counts when the objects' banks differ
  auto attend_v64 = [&](const bf16* q, const bf16* k, const bf16* v) -> int {
    const int splits = flash_attention_hd256_v64_splits(n, T, Lk);
    float* part = nullptr;
    TRY(ws_get_t(e, "ma_part", (size_t)splits * n * T * (64 + 2), &part));
    return flash_attention_hd256_v64(q, (long)T * C, k, Lkp * C, v, Lkp * 64, n, T, Lk, ROWS ? ROWS + 2 * n : nullptr, part,
                                     part + (size_t)splits * n * T * 64, O64, st);
  };
  for (int l = 0; l < 4; ++l) {
    const std::string p = "memory_attention.layers." + std::to_string(l) + ".";
    const std::string d = "derived.memattn." + std::to_string(l) + ".";
    // ---- self attention (RoPE on q and k) ----
    TRY(layernorm_rows(X, W32(e, p + "norm1.weight"), W32(e, p + "norm1.bias"), 1e-5f, XN, nullptr, TT, C, st));
    TRY(linear_rope(XN, (long)T * C, T, n, d + "sa_qk.", 0, C, Qb, (long)T * C, T, nullptr));       // q, rotated
    TRY(linear_rope(XN, (long)T * C, T, n, d + "sa_qk.", 256, C, Kb, (long)T * C, T, nullptr));     // k, rotated
    TRY(linear(XN, C, (int)TT, p + "self_attn.v_proj.", C, C, ACT_NONE, nullptr, nullptr, Vb, C));
    TRY(attend(Qb, Kb, Vb, T, T));
    TRY(linear(O, C, (int)TT, p + "self_attn.out_proj.", C, C, ACT_NONE, X, X, nullptr, C));
    // ---- cross attention over the memory bank (RoPE on q and on the L*4096 spatial keys only) ----
    TRY(layernorm_rows(X, W32(e, p + "norm2.weight"), W32(e, p + "norm2.bias"), 1e-5f, XN, nullptr, TT, C, st));
    TRY(linear_rope(XN, (long)T * C, T, n, p + "cross_attn_image.q_proj.", 0, C, Qb, (long)T * C, T, nullptr));
    // k = k_proj(memory + pos) over all Lkp rows of every object (the padding rows of KIN are zero: finite keys the attention
    // kernel masks); only the spatial keys are rotated, the pointer tokens pass through
    TRY(linear_rope(KIN, Lkp * 64, (int)Lkp, n, p + "cross_attn_image.k_proj.", 0, 64, Kc, Lkp * C, L * 4096, ROWS));
    TRY(attend_v64(Qb, Kc, VIN));
    // x += (softmax(q k^T / 16) memory) (Wo Wv)^T + (Wo bv + bo): v_proj and out_proj in one 64 -> 256 linear map
    TRY(linear(O64, 64, (int)TT, d + "ca_ov.", C, 64, ACT_NONE, X, X, nullptr, C));
    // ---- FFN ----
    TRY(layernorm_rows(X, W32(e, p + "norm3.weight"), W32(e, p + "norm3.bias"), 1e-5f, XN, nullptr, TT, C, st));
    TRY(linear(XN, C, (int)TT, p + "linear1.", 2048, C, ACT_RELU, nullptr, nullptr, HID, 2048));
    TRY(linear(HID, 2048, (int)TT, p + "linear2.", C, 2048, ACT_NONE, X, X, nullptr, C));
  }
  TRY(layernorm_rows(X, W32(e, "memory_attention.norm.weight"), W32(e, "memory_attention.norm.bias"), 1e-5f, nullptr,
                     n == 1 ? out : OUT, TT, C, st));
  if (n > 1) TRY(batch_to_seq(OUT, out, T, n, C, st));
  return 0;
}

// ------------------------------------------------------------------------------------------------------------
// MemoryEncoder (boundary E4): MaskDownSampler -> + pix_feat_proj -> 2 x CXBlock -> out_proj, channels-last inside.
// ------------------------------------------------------------------------------------------------------------
// low (optional, instead of masks): the low-resolution logits [n,256,256]; the 1024^2 mask_for_mem is then evaluated inside the
// first down-sampler stage (binarize: the (mask > 0) form of prompted frames).  obj (optional): object_score_logits [n].
static int memory_encode_mb(sam2b_engine* e, const float* pix_feat, const float* masks, int n, float* out_x, cudaStream_t st,
                            const float* low = nullptr, int binarize = 0, const float* obj = nullptr, long pix_batch_stride = -1) {
  const long T = 4096L * n;
  float *D1, *D2, *D3, *D4, *MDS, *X;
  bf16 *COLS, *D3B, *D4N, *PIXT, *XN, *HID, *XB;
  TRY(ws_get_t(e, "me_d1", (size_t)n * 512 * 512 * 4, &D1));
  TRY(ws_get_t(e, "me_d2", (size_t)n * 256 * 256 * 16, &D2));
  TRY(ws_get_t(e, "me_d3", (size_t)n * 128 * 128 * 64, &D3));
  TRY(ws_get_t(e, "me_d4", (size_t)T * 256, &D4));
  TRY(ws_get_t(e, "me_mds", (size_t)T * 256, &MDS));
  TRY(ws_get_t(e, "me_x", (size_t)T * 256, &X));
  TRY(ws_get_t(e, "me_cols", (size_t)T * 576, &COLS));
  TRY(ws_get_t(e, "me_d3b", (size_t)n * 128 * 128 * 64, &D3B));
  TRY(ws_get_t(e, "me_d4n", (size_t)T * 256, &D4N));
  TRY(ws_get_t(e, "me_pixt", (size_t)T * 256, &PIXT));
  TRY(ws_get_t(e, "me_xn", (size_t)T * 256, &XN));
  TRY(ws_get_t(e, "me_hid", (size_t)T * 1024, &HID));
  TRY(ws_get_t(e, "me_xb", (size_t)T * 256, &XB));
  const std::string m = "memory_encoder.", en = m + "mask_downsampler.encoder.";
  auto P = [&](const std::string& name) { return W32(e, name); };
  float* stage_out[2] = {D1, D2};
  const float* stage_in = masks;
  int H = 1024;
  for (int j = 0; j < 2; ++j) {       // 1 -> 4 -> 16 channels: a few taps per output, fp32 FMA kernels (HBM bound)
    const std::string c = en + std::to_string(3 * j) + ".", l = en + std::to_string(3 * j + 1) + ".";
    if (j == 0 && low) TRY(mask_ds1_from_low(low, D1, n, binarize, P(c + "weight"), P(c + "bias"), P(l + "weight"), P(l + "bias"), st));
    else TRY(mask_ds_stage(j + 1, stage_in, stage_out[j], n, H, P(c + "weight"), P(c + "bias"), P(l + "weight"), P(l + "bias"), st));
    stage_in = stage_out[j];
    H /= 2;
  }
  auto linear = [&](const bf16* A, long M, int K, const bf16* Wt, const float* bias, int N, int act, const float* gamma,
                    const float* residual, float* o32, bf16* o16) -> int {
    GemmArgs g;
    g.A = A; g.lda = K; g.M = (int)M; g.K = K;
    g.B = Wt; g.ldb = K; g.N = N;
    g.bias = bias; g.act = act; g.gamma = gamma;
    g.residual = residual; g.ld_res = N;
    g.out_f32 = o32; g.out_bf16 = o16; g.ldc = N;
    return gemm_bf16(g, st);
  };
  // 16 -> 64 and 64 -> 256 channels: dense contractions over (3 x 3 x C_in) = 144 / 576 -> im2col + tcgen05 GEMM + LN2d + GELU
  // (the 16 -> 64 stage as a direct fp32 FMA kernel ran at 8 TFLOP/s: 0.59 ms per frame at 16 objects)
  TRY(im2col3x3s2_nhwc(D2, COLS, n, 256, 16, st));
  TRY(linear(COLS, 16384L * n, 144, W16(e, "derived.memenc.ds3.weight"), P(en + "6.bias"), 64, ACT_NONE, nullptr, nullptr, D3, nullptr));
  TRY(layernorm_gelu_rows(D3, P(en + "7.weight"), P(en + "7.bias"), 1e-6f, D3B, 16384L * n, 64, st));
  TRY(im2col3x3s2_nhwc_bf16(D3B, COLS, n, 128, 64, st));
  TRY(linear(COLS, T, 576, W16(e, "derived.memenc.ds4.weight"), P(en + "9.bias"), 256, ACT_NONE, nullptr, nullptr, D4, nullptr));
  TRY(layernorm_gelu_rows(D4, P(en + "10.weight"), P(en + "10.bias"), 1e-6f, D4N, T, 256, st));
  TRY(linear(D4N, T, 256, W16(e, en + "12.weight"), P(en + "12.bias"), 256, ACT_NONE, nullptr, nullptr, MDS, nullptr));
  TRY(nchw_to_nhwc_bf16(pix_feat, PIXT, n, 256, 4096, st, pix_batch_stride));
  TRY(linear(PIXT, T, 256, W16(e, m + "pix_feat_proj.weight"), P(m + "pix_feat_proj.bias"), 256, ACT_NONE, nullptr, MDS, X, nullptr));
  for (int l = 0; l < 2; ++l) {
    const std::string f = m + "fuser.layers." + std::to_string(l) + ".";
    TRY(dwconv7_ln_nhwc(X, XN, n, 64, P("derived.memenc.fuser." + std::to_string(l) + ".dw.weight"), P(f + "dwconv.bias"),
                        P(f + "norm.weight"), P(f + "norm.bias"), st));
    TRY(linear(XN, T, 256, W16(e, f + "pwconv1.weight"), P(f + "pwconv1.bias"), 1024, ACT_GELU, nullptr, nullptr, nullptr, HID));
    TRY(linear(HID, T, 1024, W16(e, f + "pwconv2.weight"), P(f + "pwconv2.bias"), 256, ACT_NONE, P(f + "gamma"), X, X, nullptr));
  }
  TRY(cast_f32_bf16(X, XB, T * 256, st));
  {
    GemmArgs g;   // out_proj with swapped operands -> NCHW [n][64][4096]
    g.A = W16(e, m + "out_proj.weight"); g.lda = 256; g.M = 64; g.K = 256;
    g.B = XB; g.ldb = 256; g.b_batch_stride = 4096L * 256; g.N = 4096;
    g.batch = n;
    g.bias = P(m + "out_proj.bias"); g.bias_along_m = 1;
    g.out_f32 = out_x; g.ldc = 4096; g.c_batch_stride = 64L * 4096;
    TRY(gemm_bf16(g, st));
  }
  if (obj) {
    if (!e->find("no_obj_embed_spatial")) { set_error("memory_encode: object scores given but the engine has no no_obj_embed_spatial tensor"); return 1; }
    TRY(add_no_obj_embed(out_x, obj, W32(e, "no_obj_embed_spatial"), n, st));
  }
  return 0;
}


// ------------------------------------------------------------------------------------------------------------
// SAM heads for n prompts / objects: prompt encoder (points), two-way transformer, up-scaling,
// hyper-network masks, IoU / object-score heads, best-mask selection, object pointer.  Reference:
// modeling/sam2_base_official.py:338-494 (_forward_sam_heads), modeling/sam/mask_decoder.py:116-316,
// modeling/sam/transformer.py:97-219.  Image side (4096 rows per object): split-bf16 tcgen05 GEMMs (sam_heads.cu header);
// token side: fp32 kernels.
// ------------------------------------------------------------------------------------------------------------
struct HeadsIO {
  const float* pix; long sn, sc, sp;             // image embedding, addressed pix[b*sn + c*sc + p*sp]
  const float* s0; long s0_bs;                   // conv_s0 level [.,32,256,256], batch stride (0: one image for all)
  const float* s1; long s1_bs;                   // conv_s1 level [.,64,128,128]
  const float* coords; const int* labels; int P; // device, [n,P,2] / [n,P]
  const float* sparse; int S;                    // optional: the sparse prompt embeddings themselves [n,S,256] (then no points)
  const float* dense;                            // optional dense prompt embedding [n,256,64,64]
  int multimask, decoder_only;
  float *low_multi, *high_multi, *ious, *low_best, *high_best, *obj_ptr, *obj_logits, *mask_tokens;
};

static int sam_heads_mb(sam2b_engine* e, const HeadsIO& io, int n, cudaStream_t st) {
  const int P = io.P > 0 ? io.P : 1, T = io.sparse ? 6 + io.S : 6 + P + 1, nT = n * T;
  const long R = 4096L * n;
  float *KEYS, *KVQ, *XR, *TOK, *QPE, *TQ, *TK, *TV, *TA, *PART, *HID, *TPART, *H1, *H2, *IOU, *OBJ, *HYP, *PTR4, *MASKS, *IOUT, *OBJT, *PTRT;
  bf16 *KS, *OS;
  int *STAB, *SEL, *LABP;
  TRY(ws_get_t(e, "sh_keys", (size_t)R * 256, &KEYS));
  TRY(ws_get_t(e, "sh_ks", (size_t)R * 768, &KS));
  TRY(ws_get_t(e, "sh_kvq", (size_t)R * 384, &KVQ));
  TRY(ws_get_t(e, "sh_os", (size_t)R * 384, &OS));
  TRY(ws_get_t(e, "sh_xr", (size_t)R * 256, &XR));
  TRY(ws_get_t(e, "sh_tok", (size_t)nT * 256, &TOK));
  TRY(ws_get_t(e, "sh_qpe", (size_t)nT * 256, &QPE));
  TRY(ws_get_t(e, "sh_tq", (size_t)nT * 256, &TQ));
  TRY(ws_get_t(e, "sh_tk", (size_t)nT * 256, &TK));
  TRY(ws_get_t(e, "sh_tv", (size_t)nT * 256, &TV));
  TRY(ws_get_t(e, "sh_ta", (size_t)nT * 256, &TA));
  TRY(ws_get_t(e, "sh_part", (size_t)8 * nT * 256, &PART));
  TRY(ws_get_t(e, "sh_hid", (size_t)nT * 2048, &HID));
  TRY(ws_get_t(e, "sh_tpart", (size_t)n * 128 * T * 8 * 18, &TPART));      // 128 key tiles x T tokens x 8 heads x (max, sum, 16 outputs)
  TRY(ws_get_t(e, "sh_h1", (size_t)10 * n * 256, &H1));
  TRY(ws_get_t(e, "sh_h2", (size_t)10 * n * 256, &H2));
  TRY(ws_get_t(e, "sh_iou", (size_t)n * 4, &IOU));
  TRY(ws_get_t(e, "sh_obj", (size_t)n, &OBJ));
  TRY(ws_get_t(e, "sh_hyp", (size_t)n * 128, &HYP));
  TRY(ws_get_t(e, "sh_ptr4", (size_t)n * 4 * 256, &PTR4));
  TRY(ws_get_t(e, "sh_masks", (size_t)n * 4 * 65536, &MASKS));
  TRY(ws_get_t(e, "sh_iout", (size_t)n * 4, &IOUT));
  TRY(ws_get_t(e, "sh_objt", (size_t)n, &OBJT));
  TRY(ws_get_t(e, "sh_ptrt", (size_t)n * 256, &PTRT));
  TRY(ws_get_t(e, "sh_stab", (size_t)n * 2, &STAB));
  TRY(ws_get_t(e, "sh_sel", (size_t)n * 4, &SEL));
  TRY(ws_get_t(e, "sh_labp", (size_t)n * 4, &LABP));
  auto P32 = [&](const std::string& name) { return W32(e, name); };
  const std::string md = "sam_mask_decoder.", tr = md + "transformer.", pe = "sam_prompt_encoder.";

  // prompt encoder: no prompt at all = one point with label -1 (sam2_base_official.py:391-395) + the padding point
  const float* coords = io.coords;
  const int* labels = io.labels;
  if (io.P <= 0 && !io.sparse) {
    CUDA_OK(cudaMemsetAsync(LABP, 0xFF, (size_t)n * 4 * sizeof(int), st));     // labels = -1; the coordinates are never read
    labels = LABP;
    coords = reinterpret_cast<const float*>(LABP);
  }
  const float* pt_emb[4] = {P32(pe + "point_embeddings.0.weight"), P32(pe + "point_embeddings.1.weight"),
                            P32(pe + "point_embeddings.2.weight"), P32(pe + "point_embeddings.3.weight")};
  TRY(heads_tokens(coords, labels, P, n, P32(pe + "pe_layer.positional_encoding_gaussian_matrix"), P32(pe + "not_a_point_embed.weight"),
                   pt_emb, P32(md + "obj_score_token.weight"), P32(md + "iou_token.weight"), P32(md + "mask_tokens.weight"),
                   io.sparse, io.S, TOK, QPE, st));
  TRY(heads_prep_src(io.pix, io.sn, io.sc, io.sp, io.dense, P32(pe + "no_mask_embed.weight"), KEYS, KS, n, st));

  auto job = [&](const float* X, long ldx, const float* X2, const float* Wt, long ldw, const float* b, float* Y, long ldy, int rows,
                 int N, int K, int act) {
    LinJob j;
    j.X = X; j.ldx = ldx; j.X2 = X2; j.ldx2 = ldx; j.W = Wt; j.ldw = ldw; j.b = b; j.Y = Y; j.ldy = ldy;
    j.rows = rows; j.N = N; j.K = K; j.act = act;
    return j;
  };
  // linear (no bias) into PART, then bias + residual + LayerNorm -> TOK
  auto proj_res_ln = [&](const float* X, int K, const std::string& lin, const float* res, const std::string& norm) -> int {
    LinJobs J;
    const int S = K > 256 ? K / 256 : 1;
    for (int s = 0; s < S; ++s)
      J.j[s] = job(X + s * 256, K, nullptr, P32(lin + "weight") + s * 256, K, nullptr, PART + (size_t)s * nT * 256, 256, nT, 256,
                   S > 1 ? 256 : K, 0);
    TRY(tok_linear(J, S, st));
    return tok_sum_res_ln(PART, (long)nT * 256, S, P32(lin + "bias"), res, P32(norm + "weight"), P32(norm + "bias"), 1e-5f, TOK, nT, st);
  };
  auto image_gemm = [&](const bf16* A, int K, const std::string& wname, int N, const float* bias, const float* residual, long ld_res,
                        int res_row_mod, float* out, long ldc) -> int {
    GemmArgs g;
    g.A = A; g.lda = K; g.M = (int)R; g.K = K;
    g.B = W16(e, wname); g.ldb = K; g.N = N;
    g.bias = bias; g.residual = residual; g.ld_res = ld_res; g.res_row_mod = res_row_mod;
    g.out_f32 = out; g.ldc = ldc;
    return gemm_bf16(g, st);
  };
  // token -> image attention block shared by the two layers and the final attention
  auto t2i_block = [&](const std::string& p, int ld, const std::string& norm) -> int {
    LinJobs J;
    J.j[0] = job(TOK, 256, QPE, P32(p + "q_proj.weight"), 256, P32(p + "q_proj.bias"), TQ, 128, nT, 128, 256, 0);
    TRY(tok_linear(J, 1, st));
    TRY(t2i_attention(KVQ, ld, TQ, n, T, TPART, TA, st));
    return proj_res_ln(TA, 128, p + "out_proj.", TOK, norm);
  };

  for (int l = 0; l < 2; ++l) {
    const std::string p = tr + "layers." + std::to_string(l) + ".", d = "derived.heads." + std::to_string(l) + ".";
    {   // self-attention of the tokens (transformer.py:164-172): layer 0 without the positional term and without residual
      const std::string sa = p + "self_attn.";
      const float* qpe = l ? QPE : nullptr;
      LinJobs J;
      J.j[0] = job(TOK, 256, qpe, P32(sa + "q_proj.weight"), 256, P32(sa + "q_proj.bias"), TQ, 256, nT, 256, 256, 0);
      J.j[1] = job(TOK, 256, qpe, P32(sa + "k_proj.weight"), 256, P32(sa + "k_proj.bias"), TK, 256, nT, 256, 256, 0);
      J.j[2] = job(TOK, 256, nullptr, P32(sa + "v_proj.weight"), 256, P32(sa + "v_proj.bias"), TV, 256, nT, 256, 256, 0);
      TRY(tok_linear(J, 3, st));
      TRY(tok_self_attn(TQ, TK, TV, n, T, TA, st));
      TRY(proj_res_ln(TA, 256, sa + "out_proj.", l ? TOK : nullptr, p + "norm1."));
    }
    // k, v of the token -> image attention and q of the image -> token attention: one GEMM over the keys of this layer,
    // the dense positional encoding enters as the constant table (dense_pe W^T + b) added per position
    TRY(image_gemm(KS, 768, d + "kvq.weight", 384, nullptr, P32(d + "kvq.table"), 384, 4096, KVQ, 384));
    TRY(t2i_block(p + "cross_attn_token_to_image.", 384, p + "norm2."));
    {   // MLP 256 -> 2048 -> 256 (ReLU), transformer.py:184-187
      LinJobs J;
      J.j[0] = job(TOK, 256, nullptr, P32(p + "mlp.layers.0.weight"), 256, P32(p + "mlp.layers.0.bias"), HID, 2048, nT, 2048, 256, 1);
      TRY(tok_linear(J, 1, st));
      TRY(proj_res_ln(HID, 2048, p + "mlp.layers.1.", TOK, p + "norm3."));
    }
    {   // image -> token attention (transformer.py:189-196): keys = norm4(keys + attn(q = keys + pe, k = queries + pe, v = queries))
      const std::string ca = p + "cross_attn_image_to_token.";
      LinJobs J;
      J.j[0] = job(TOK, 256, QPE, P32(ca + "k_proj.weight"), 256, P32(ca + "k_proj.bias"), TK, 128, nT, 128, 256, 0);
      J.j[1] = job(TOK, 256, nullptr, P32(ca + "v_proj.weight"), 256, P32(ca + "v_proj.bias"), TV, 128, nT, 128, 256, 0);
      TRY(tok_linear(J, 2, st));
      TRY(i2t_attention(KVQ, 384, 256, TK, TV, n, T, OS, st));
      TRY(image_gemm(OS, 384, d + "i2t_out.weight", 256, P32(ca + "out_proj.bias"), KEYS, 256, 0, XR, 256));
      TRY(ln256_split(XR, P32(p + "norm4.weight"), P32(p + "norm4.bias"), 1e-5f, KEYS, KS, R, st));
    }
  }
  TRY(image_gemm(KS, 768, "derived.heads.final.kv.weight", 256, nullptr, P32("derived.heads.final.kv.table"), 256, 4096, KVQ, 256));
  TRY(t2i_block(tr + "final_attn_token_to_image.", 256, tr + "norm_final_attn."));

  if (io.mask_tokens)     // MaskDecoder.forward's third output comes from these: the four mask tokens after the transformer
    CUDA_OK(cudaMemcpy2DAsync(io.mask_tokens, 4 * 256 * sizeof(float), TOK + 2 * 256, (size_t)T * 256 * sizeof(float),
                              4 * 256 * sizeof(float), n, cudaMemcpyDeviceToDevice, st));
  // the small heads on the output tokens (mask_decoder.py:302-316, sam2_base_official.py:474-494): IoU (token 1), object score
  // (token 0), the four hyper-network MLPs and obj_ptr_proj on the four mask tokens (tokens 2..5); three launches of 10 jobs
  const std::string heads[10] = {md + "iou_prediction_head.", md + "pred_obj_score_head.", md + "output_hypernetworks_mlps.0.",
                                 md + "output_hypernetworks_mlps.1.", md + "output_hypernetworks_mlps.2.",
                                 md + "output_hypernetworks_mlps.3.", "obj_ptr_proj.", "obj_ptr_proj.", "obj_ptr_proj.", "obj_ptr_proj."};
  const int head_tok[10] = {1, 0, 2, 3, 4, 5, 2, 3, 4, 5};
  {
    LinJobs J0, J1, J2;
    for (int i = 0; i < 10; ++i) {
      const std::string L = heads[i] + "layers.";
      J0.j[i] = job(TOK + head_tok[i] * 256, (long)T * 256, nullptr, P32(L + "0.weight"), 256, P32(L + "0.bias"), H1 + (size_t)i * n * 256,
                    256, n, 256, 256, 1);
      J1.j[i] = job(H1 + (size_t)i * n * 256, 256, nullptr, P32(L + "1.weight"), 256, P32(L + "1.bias"), H2 + (size_t)i * n * 256, 256, n,
                    256, 256, 1);
      float* Y; long ldy; int N, act = 0;
      if (i == 0) { Y = IOU; ldy = 4; N = 4; act = 2; }
      else if (i == 1) { Y = OBJ; ldy = 1; N = 1; }
      else if (i < 6) { Y = HYP + (i - 2) * 32; ldy = 128; N = 32; }
      else { Y = PTR4 + (i - 6) * 256; ldy = 1024; N = 256; }
      J2.j[i] = job(H2 + (size_t)i * n * 256, 256, nullptr, P32(L + "2.weight"), 256, P32(L + "2.bias"), Y, ldy, n, N, 256, act);
    }
    TRY(tok_linear(J0, 10, st));
    TRY(tok_linear(J1, 10, st));
    TRY(tok_linear(J2, 10, st));
  }
  // up-scaling: ConvTranspose2d(256 -> 64, k2 s2) of the final keys as a GEMM, everything after it in one kernel
  TRY(image_gemm(KS, 768, "derived.heads.up1.weight", 256, nullptr, nullptr, 0, 0, XR, 256));
  const int dynamic = !io.multimask;        // dynamic_multimask_via_stability (mask_decoder.py:215-219; on in every SAM2.1 config)
  if (dynamic) CUDA_OK(cudaMemsetAsync(STAB, 0, (size_t)n * 2 * sizeof(int), st));
  TRY(heads_upscale(XR, P32(md + "output_upscaling.0.bias"), io.s1, io.s1_bs, P32(md + "output_upscaling.1.weight"),
                    P32(md + "output_upscaling.1.bias"), P32("derived.heads.up2.weight"), P32(md + "output_upscaling.3.bias"), io.s0,
                    io.s0_bs, HYP, MASKS, dynamic ? STAB : nullptr, 0.05f, n, st));
  TRY(heads_select(IOU, OBJ, PTR4, STAB, P32("no_obj_ptr"), io.multimask, dynamic, 0.98f, io.ious ? io.ious : IOUT,
                   io.obj_ptr ? io.obj_ptr : PTRT, io.obj_logits ? io.obj_logits : OBJT, SEL, io.decoder_only, n, st));
  float* low_best = io.low_best;
  if (!low_best && io.high_best) { TRY(ws_get_t(e, "sh_lowbest", (size_t)n * 65536, &low_best)); }
  TRY(heads_write_masks(MASKS, SEL, io.multimask, -1024.0f, io.low_multi, low_best, n, st));
  if (io.high_best) TRY(upsample_bilinear_mask(low_best, io.high_best, nullptr, n, 256, 256, 1024, 1024, 0.f, st));
  if (io.high_multi && io.low_multi)
    TRY(upsample_bilinear_mask(io.low_multi, io.high_multi, nullptr, (long)n * (io.multimask ? 3 : 1), 256, 256, 1024, 1024, 0.f, st));
  return 0;
}

static const int kMaxMicroBatchCap = 32;
static const int kHostMicroBatch = 16;     // host path: two or more micro-batches so that copies overlap compute
// images per pass through the trunk (SAM2B_MICROBATCH overrides).  Every GEMM is a persistent kernel over 148 SMs (74 CTA
// pairs): at 16 images the stage-3 tile counts leave the last wave 40-60 % empty (fc2: 768 pair tiles = 10.4 waves), at 32
// images the loss is ~1 % (20.8 waves) - 2.7 % of the step.  Results are bit-identical for any micro-batch (tested).
static int max_micro_batch() {
  static int v = 0;
  if (!v) {
    const char* e = getenv("SAM2B_MICROBATCH");
    v = e ? atoi(e) : 32;
    if (v < 1) v = 1;
    if (v > kMaxMicroBatchCap) v = kMaxMicroBatchCap;
  }
  return v;
}

}  // namespace sam2b

// ============================================================================================
// C ABI
// ============================================================================================
extern "C" {

const char* sam2b_last_error(void) { return get_error(); }
const char* sam2b_version(void) { return "sam2b 0.1 (sm_100a)"; }
unsigned long long sam2b_launch_count(void) { return g_launch_count.load(std::memory_order_relaxed); }

int sam2b_create(int device, sam2b_handle* out) {
  if (!out) { set_error("sam2b_create: null out"); return 1; }
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) {
    set_error("sam2b_create: no CUDA device available (this library has no CPU path)");
    return 1;
  }
  if (device < 0 || device >= count) { set_error("sam2b_create: device %d out of range (%d devices)", device, count); return 1; }
  cudaDeviceProp prop;
  CUDA_OK(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) { set_error("sam2b_create: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor); return 1; }
  ENTER_DEVICE(device);
  sam2b_engine* e = new sam2b_engine();
  e->device = device;
  build_block_cfgs(e->blocks);
  CUDA_OK(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
  CUDA_OK(cudaStreamCreateWithFlags(&e->copy_stream2, cudaStreamNonBlocking));
  for (auto& ev : e->ev) CUDA_OK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  CUDA_OK(cudaEventCreateWithFlags(&e->done, cudaEventDisableTiming));
  *out = e;
  return 0;
}

int sam2b_destroy(sam2b_handle h) {
  if (!h) return 0;
  DeviceGuard guard(h->device);
  cudaDeviceSynchronize();
  for (auto& kv : h->w) { if (kv.second.f32) cudaFree(kv.second.f32); if (kv.second.b16) cudaFree(kv.second.b16); }
  for (auto& kv : h->ws) if (kv.second.p) cudaFree(kv.second.p);
  for (auto& ev : h->ev) if (ev) cudaEventDestroy(ev);
  if (h->done) cudaEventDestroy(h->done);
  if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
  if (h->copy_stream2) cudaStreamDestroy(h->copy_stream2);
  delete h;
  return 0;
}

int sam2b_set_tensor(sam2b_handle h, const char* name, const float* dev_data, const int64_t* shape, int ndim) {
  if (!h || !name || !dev_data) { set_error("sam2b_set_tensor: null argument"); return 1; }
  ENTER_DEVICE(h->device);
  std::lock_guard<std::mutex> lk(h->mu);
  Tensor& t = h->w[name];
  int64_t numel = 1;
  std::vector<int64_t> shp;
  for (int i = 0; i < ndim; ++i) { numel *= shape[i]; shp.push_back(shape[i]); }
  if (t.f32 && t.numel != numel) { cudaFree(t.f32); t.f32 = nullptr; }
  if (t.b16) { cudaFree(t.b16); t.b16 = nullptr; }
  // round the allocation up so vectorised (16 B) reads of the tail stay in bounds
  if (!t.f32) CUDA_OK(cudaMalloc(&t.f32, ((size_t)numel * 4 + 63) & ~(size_t)63));
  CUDA_OK(cudaMemcpy(t.f32, dev_data, (size_t)numel * 4, cudaMemcpyDeviceToDevice));
  t.numel = numel; t.shape = shp;
  h->finalized = false;
  return 0;
}

int sam2b_finalize(sam2b_handle h) {
  if (!h) { set_error("sam2b_finalize: null handle"); return 1; }
  std::lock_guard<std::mutex> lk(h->mu);
  return finalize(h);
}

int sam2b_image_encode(sam2b_handle h, const float* img, int n, int normalised, int add_no_mem_embed, float* fpn0,
                       float* fpn1, float* fpn2, void* stream) {
  if (!h || !h->finalized) { set_error("sam2b_image_encode: engine not finalized"); return 1; }
  if (n <= 0 || !img || !fpn0 || !fpn1 || !fpn2) { set_error("sam2b_image_encode: bad arguments (n=%d)", n); return 1; }
  ENTER_DEVICE(h->device);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  ENGINE_CALL(h, st);
  const int kMaxMicroBatch = max_micro_batch();
  for (int i0 = 0; i0 < n; i0 += kMaxMicroBatch) {
    const int mb = (n - i0) < kMaxMicroBatch ? (n - i0) : kMaxMicroBatch;
    TRY(image_encode_mb(h, img + (size_t)i0 * 3 * 1024 * 1024, mb, normalised, add_no_mem_embed,
                        fpn0 + (size_t)i0 * 32 * 65536, fpn1 + (size_t)i0 * 64 * 16384, fpn2 + (size_t)i0 * 256 * 4096, st));
  }
  return 0;
}

// Host-buffer encode.  `u8` = 0: img_host is fp32 [n][3][1024][1024]; `u8` = 1: img_host is uint8 [n][1024][1024][3] (raw frames),
// uploaded as bytes (a quarter of the PCIe traffic) and converted on the device by preprocess_u8_hwc (/255, then mean/std).
// async_mode: return without synchronising and without fencing the copy streams behind the caller's earlier work, so that
// the uploads of this call overlap the compute of the previous one (sam2b_host_sync waits for everything).  The two staging
// slots and their events live across calls (h->host_it), which is what orders slot re-use between calls.
static int image_encode_host_impl(sam2b_handle h, const void* img_host_v, int u8, int n, int normalised, int add_no_mem_embed,
                                  float* fpn0_host, float* fpn1_host, float* fpn2_host, void* stream, int async_mode = 0) {
  if (!h || !h->finalized) { set_error("sam2b_image_encode_host: engine not finalized"); return 1; }
  if (n <= 0 || !img_host_v || !fpn0_host || !fpn1_host || !fpn2_host) { set_error("sam2b_image_encode_host: bad arguments"); return 1; }
  ENTER_DEVICE(h->device);
  const float* img_host = reinterpret_cast<const float*>(img_host_v);
  const uint8_t* img_host_u8 = reinterpret_cast<const uint8_t*>(img_host_v);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  ENGINE_CALL(h, st);
  // double-buffered staging: uploads run on copy_stream, downloads on copy_stream2, compute on `st`
  const size_t in_sz = (size_t)3 * 1024 * 1024, o0 = (size_t)32 * 65536, o1 = (size_t)64 * 16384, o2 = (size_t)256 * 4096;
  const int kMaxMicroBatch = max_micro_batch() < kHostMicroBatch ? max_micro_batch() : kHostMicroBatch;
  float *in[2], *f0[2], *f1[2], *f2[2];
  const char* names[2][4] = {{"h_in0", "h_f00", "h_f10", "h_f20"}, {"h_in1", "h_f01", "h_f11", "h_f21"}};
  for (int s = 0; s < 2; ++s) {
    TRY(ws_get_t(h, names[s][0], in_sz * kMaxMicroBatch, &in[s]));
    TRY(ws_get_t(h, names[s][1], o0 * kMaxMicroBatch, &f0[s]));
    TRY(ws_get_t(h, names[s][2], o1 * kMaxMicroBatch, &f1[s]));
    TRY(ws_get_t(h, names[s][3], o2 * kMaxMicroBatch, &f2[s]));
  }
  uint8_t* raw[2] = {nullptr, nullptr};
  if (u8) {
    TRY(ws_get_t(h, "h_raw0", in_sz * kMaxMicroBatch, &raw[0]));
    TRY(ws_get_t(h, "h_raw1", in_sz * kMaxMicroBatch, &raw[1]));
  }
  cudaStream_t cs = h->copy_stream, ds = h->copy_stream2;
  // ev[0..1]: input slot s uploaded; ev[2..3]: compute on slot s done; ev[4..5]: outputs of slot s downloaded;
  // ev[8..9], ev[10..11]: fpn0 / fpn1 of slot s complete (early, see image_encode_mb)
  if (!async_mode || h->host_it == 0) {
    CUDA_OK(cudaEventRecord(h->ev[6], st));
    CUDA_OK(cudaStreamWaitEvent(cs, h->ev[6], 0));
    CUDA_OK(cudaStreamWaitEvent(ds, h->ev[6], 0));
  }
  for (int i0 = 0; i0 < n; i0 += kMaxMicroBatch) {
    const unsigned long long it = h->host_it++;
    const int s = (int)(it & 1);
    const int mb = (n - i0) < kMaxMicroBatch ? (n - i0) : kMaxMicroBatch;
    if (it >= 2) CUDA_OK(cudaStreamWaitEvent(cs, h->ev[2 + s], 0));       // slot's previous compute has consumed the input
    if (u8) CUDA_OK(cudaMemcpyAsync(raw[s], img_host_u8 + (size_t)i0 * in_sz, in_sz * mb, cudaMemcpyHostToDevice, cs));
    else CUDA_OK(cudaMemcpyAsync(in[s], img_host + (size_t)i0 * in_sz, in_sz * mb * 4, cudaMemcpyHostToDevice, cs));
    CUDA_OK(cudaEventRecord(h->ev[s], cs));
    CUDA_OK(cudaStreamWaitEvent(st, h->ev[s], 0));
    if (u8) TRY(preprocess_u8_hwc(raw[s], in[s], mb, 1024, 1024, 1, st));
    if (it >= 2) CUDA_OK(cudaStreamWaitEvent(st, h->ev[4 + s], 0));       // slot's previous outputs have left the device
    cudaEvent_t early[2] = {h->ev[8 + s], h->ev[10 + s]};
    TRY(image_encode_mb(h, in[s], mb, normalised, add_no_mem_embed, f0[s], f1[s], f2[s], st, early));
    CUDA_OK(cudaEventRecord(h->ev[2 + s], st));
    CUDA_OK(cudaStreamWaitEvent(ds, early[0], 0));
    CUDA_OK(cudaMemcpyAsync(fpn0_host + (size_t)i0 * o0, f0[s], o0 * mb * 4, cudaMemcpyDeviceToHost, ds));
    CUDA_OK(cudaStreamWaitEvent(ds, early[1], 0));
    CUDA_OK(cudaMemcpyAsync(fpn1_host + (size_t)i0 * o1, f1[s], o1 * mb * 4, cudaMemcpyDeviceToHost, ds));
    CUDA_OK(cudaStreamWaitEvent(ds, h->ev[2 + s], 0));
    CUDA_OK(cudaMemcpyAsync(fpn2_host + (size_t)i0 * o2, f2[s], o2 * mb * 4, cudaMemcpyDeviceToHost, ds));
    CUDA_OK(cudaEventRecord(h->ev[4 + s], ds));
  }
  if (async_mode) return 0;
  CUDA_OK(cudaStreamSynchronize(ds));
  CUDA_OK(cudaStreamSynchronize(cs));
  CUDA_OK(cudaStreamSynchronize(st));
  return 0;
}

int sam2b_image_encode_host(sam2b_handle h, const float* img_host, int n, int normalised, int add_no_mem_embed,
                            float* fpn0_host, float* fpn1_host, float* fpn2_host, void* stream) {
  return image_encode_host_impl(h, img_host, 0, n, normalised, add_no_mem_embed, fpn0_host, fpn1_host, fpn2_host, stream);
}

int sam2b_image_encode_host_async(sam2b_handle h, const float* img_host, int n, int normalised, int add_no_mem_embed,
                                  float* fpn0_host, float* fpn1_host, float* fpn2_host, void* stream) {
  return image_encode_host_impl(h, img_host, 0, n, normalised, add_no_mem_embed, fpn0_host, fpn1_host, fpn2_host, stream, 1);
}
int sam2b_host_sync(sam2b_handle h, void* stream) {
  if (!h) { set_error("sam2b_host_sync: null engine"); return 1; }
  ENTER_DEVICE(h->device);
  CUDA_OK(cudaStreamSynchronize(h->copy_stream2));
  CUDA_OK(cudaStreamSynchronize(h->copy_stream));
  CUDA_OK(cudaStreamSynchronize(reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

int sam2b_image_encode_host_u8(sam2b_handle h, const void* hwc_u8_host, int n, int add_no_mem_embed,
                               float* fpn0_host, float* fpn1_host, float* fpn2_host, void* stream) {
  return image_encode_host_impl(h, hwc_u8_host, 1, n, 1, add_no_mem_embed, fpn0_host, fpn1_host, fpn2_host, stream);
}

int sam2b_memory_attention(sam2b_handle h, const float* curr, const float* curr_pos, const float* memory,
                           const float* memory_pos, int L, const float* memory_exclude, const float* memory_pos_exclude,
                           int P, int n, float* out, void* stream) {
  if (!h || !h->finalized) { set_error("sam2b_memory_attention: engine not finalized"); return 1; }
  if (!curr || !curr_pos || !memory || !memory_pos || !out || n <= 0) { set_error("sam2b_memory_attention: null/empty argument"); return 1; }
  // the reference's exported engines carry L in [1,7] (one conditioning frame + 6 recent ones); every kernel here is general
  // in L, so several conditioning frames (L up to 16) are accepted as the reference's torch path accepts them
  if (L < 1 || L > 16 || P < 0 || P > 64 || (P % 4)) { set_error("sam2b_memory_attention: L=%d must be in [1,16], P=%d in [0,64] and a multiple of 4", L, P); return 1; }
  if (P > 0 && (!memory_exclude || !memory_pos_exclude)) { set_error("sam2b_memory_attention: P=%d but memory_exclude is NULL", P); return 1; }
  ENTER_DEVICE(h->device);
  ENGINE_CALL(h, reinterpret_cast<cudaStream_t>(stream));
  return memory_attention_mb(h, curr, curr_pos, memory, memory_pos, L, memory_exclude, memory_pos_exclude, P, n, n, nullptr, out,
                             reinterpret_cast<cudaStream_t>(stream));
}

int sam2b_memory_attention_varlen(sam2b_handle h, const float* curr, const float* curr_pos, const float* memory,
                                  const float* memory_pos, int L_max, const float* memory_exclude, const float* memory_pos_exclude,
                                  int P_max, int n, const int32_t* L_rows, const int32_t* P_rows, float* out, void* stream) {
  if (!L_rows && !P_rows)
    return sam2b_memory_attention(h, curr, curr_pos, memory, memory_pos, L_max, memory_exclude, memory_pos_exclude, P_max, n, out, stream);
  if (!h || !h->finalized) { set_error("sam2b_memory_attention_varlen: engine not finalized"); return 1; }
  if (!curr || !curr_pos || !memory || !memory_pos || !out || n <= 0) { set_error("sam2b_memory_attention_varlen: null/empty argument"); return 1; }
  if (L_max < 1 || L_max > 16 || P_max < 0 || P_max > 64 || (P_max % 4)) { set_error("sam2b_memory_attention_varlen: L_max=%d must be in [1,16], P_max=%d in [0,64] and a multiple of 4", L_max, P_max); return 1; }
  if (P_max > 0 && (!memory_exclude || !memory_pos_exclude)) { set_error("sam2b_memory_attention_varlen: P_max=%d but memory_exclude is NULL", P_max); return 1; }
  std::vector<int> rows(2 * (size_t)n);
  for (int b = 0; b < n; ++b) {
    const int Lb = L_rows ? L_rows[b] : L_max, Pb = P_rows ? P_rows[b] : P_max;
    if (Lb < 1 || Lb > L_max || Pb < 0 || Pb > P_max || (Pb % 4)) {
      set_error("sam2b_memory_attention_varlen: object %d has L=%d (must be in [1,%d]) / P=%d (in [0,%d], multiple of 4)", b, Lb, L_max, Pb, P_max);
      return 1;
    }
    rows[b] = Lb; rows[n + b] = Pb;
  }
  ENTER_DEVICE(h->device);
  ENGINE_CALL(h, reinterpret_cast<cudaStream_t>(stream));
  return memory_attention_mb(h, curr, curr_pos, memory, memory_pos, L_max, memory_exclude, memory_pos_exclude, P_max, n, n, rows.data(),
                             out, reinterpret_cast<cudaStream_t>(stream));
}

int sam2b_memory_attention_varlen_dev(sam2b_handle h, const float* curr, const float* curr_pos, const float* memory,
                                      const float* memory_pos, int L_max, const float* memory_exclude, const float* memory_pos_exclude,
                                      int P_max, int n, const int32_t* counts_dev, float* out, void* stream) {
  if (!h || !h->finalized) { set_error("sam2b_memory_attention_varlen_dev: engine not finalized"); return 1; }
  if (!curr || !curr_pos || !memory || !memory_pos || !out || !counts_dev || n <= 0) { set_error("sam2b_memory_attention_varlen_dev: null/empty argument"); return 1; }
  if (L_max < 1 || L_max > 16 || P_max < 0 || P_max > 64 || (P_max % 4)) { set_error("sam2b_memory_attention_varlen_dev: L_max=%d must be in [1,16], P_max=%d in [0,64] and a multiple of 4", L_max, P_max); return 1; }
  if (P_max > 0 && (!memory_exclude || !memory_pos_exclude)) { set_error("sam2b_memory_attention_varlen_dev: P_max=%d but memory_exclude is NULL", P_max); return 1; }
  ENTER_DEVICE(h->device);
  ENGINE_CALL(h, reinterpret_cast<cudaStream_t>(stream));
  return memory_attention_mb(h, curr, curr_pos, memory, memory_pos, L_max, memory_exclude, memory_pos_exclude, P_max, n, n, nullptr, out,
                             reinterpret_cast<cudaStream_t>(stream), counts_dev);
}

int sam2b_memory_encode(sam2b_handle h, const float* pix_feat, const float* masks, int n, float* out_x, void* stream) {
  if (!h || !h->finalized) { set_error("sam2b_memory_encode: engine not finalized"); return 1; }
  if (!pix_feat || !masks || !out_x || n <= 0) { set_error("sam2b_memory_encode: null/empty argument"); return 1; }
  ENTER_DEVICE(h->device);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  ENGINE_CALL(h, st);
  const int kMaxMicroBatch = 16;
  for (int i0 = 0; i0 < n; i0 += kMaxMicroBatch) {
    const int mb = (n - i0) < kMaxMicroBatch ? (n - i0) : kMaxMicroBatch;
    TRY(memory_encode_mb(h, pix_feat + (size_t)i0 * 256 * 4096, masks + (size_t)i0 * 1024 * 1024, mb, out_x + (size_t)i0 * 64 * 4096, st));
  }
  return 0;
}

int sam2b_sam_heads(sam2b_handle h, const float* backbone_features, int64_t feat_stride_n, int64_t feat_stride_c,
                    int64_t feat_stride_p, const float* feat_s0, int64_t s0_batch_stride, const float* feat_s1,
                    int64_t s1_batch_stride, const float* point_coords, const int32_t* point_labels, int num_points,
                    const float* sparse_embeddings, int num_sparse, const float* dense_embedding, int n, int multimask_output,
                    int decoder_only, float* low_res_multimasks, float* high_res_multimasks, float* ious, float* low_res_masks,
                    float* high_res_masks, float* obj_ptr, float* object_score_logits, float* mask_tokens_out, void* stream) {
  if (!h || !h->finalized) { set_error("sam2b_sam_heads: engine not finalized"); return 1; }
  if (!h->has_heads) { set_error("sam2b_sam_heads: the engine was built without the sam_mask_decoder / sam_prompt_encoder tensors"); return 1; }
  if (!backbone_features || !feat_s0 || !feat_s1 || n <= 0) { set_error("sam2b_sam_heads: null/empty argument"); return 1; }
  if (num_points < 0 || num_points > 25) { set_error("sam2b_sam_heads: %d points per prompt (0..25)", num_points); return 1; }
  if (sparse_embeddings && (num_sparse < 0 || num_sparse > 26)) { set_error("sam2b_sam_heads: %d sparse embeddings per prompt (0..26)", num_sparse); return 1; }
  if (!sparse_embeddings && num_points > 0 && (!point_coords || !point_labels)) { set_error("sam2b_sam_heads: num_points=%d but the point arrays are NULL", num_points); return 1; }
  if (high_res_multimasks && !low_res_multimasks) { set_error("sam2b_sam_heads: high_res_multimasks needs low_res_multimasks"); return 1; }
  ENTER_DEVICE(h->device);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  ENGINE_CALL(h, st);
  const int kMaxMicroBatch = 64, M = multimask_output ? 3 : 1;
  for (int i0 = 0; i0 < n; i0 += kMaxMicroBatch) {
    const int mb = (n - i0) < kMaxMicroBatch ? (n - i0) : kMaxMicroBatch;
    HeadsIO io;
    io.pix = backbone_features + (size_t)i0 * feat_stride_n; io.sn = feat_stride_n; io.sc = feat_stride_c; io.sp = feat_stride_p;
    io.s0 = feat_s0 + (size_t)i0 * s0_batch_stride; io.s0_bs = s0_batch_stride;
    io.s1 = feat_s1 + (size_t)i0 * s1_batch_stride; io.s1_bs = s1_batch_stride;
    io.coords = point_coords ? point_coords + (size_t)i0 * num_points * 2 : nullptr;
    io.labels = point_labels ? point_labels + (size_t)i0 * num_points : nullptr;
    io.P = num_points;
    io.sparse = sparse_embeddings ? sparse_embeddings + (size_t)i0 * num_sparse * 256 : nullptr; io.S = num_sparse;
    io.mask_tokens = mask_tokens_out ? mask_tokens_out + (size_t)i0 * 4 * 256 : nullptr;
    io.dense = dense_embedding ? dense_embedding + (size_t)i0 * 256 * 4096 : nullptr;
    io.multimask = multimask_output ? 1 : 0; io.decoder_only = decoder_only ? 1 : 0;
    io.low_multi = low_res_multimasks ? low_res_multimasks + (size_t)i0 * M * 65536 : nullptr;
    io.high_multi = high_res_multimasks ? high_res_multimasks + (size_t)i0 * M * 1048576 : nullptr;
    io.ious = ious ? ious + (size_t)i0 * M : nullptr;
    io.low_best = low_res_masks ? low_res_masks + (size_t)i0 * 65536 : nullptr;
    io.high_best = high_res_masks ? high_res_masks + (size_t)i0 * 1048576 : nullptr;
    io.obj_ptr = obj_ptr ? obj_ptr + (size_t)i0 * 256 : nullptr;
    io.obj_logits = object_score_logits ? object_score_logits + i0 : nullptr;
    TRY(sam_heads_mb(h, io, mb, st));
  }
  return 0;
}

int sam2b_mask_downscale(sam2b_handle h, const float* mask_prompt, int n, float* dense_embedding, void* stream) {
  if (!h || !h->finalized || !h->has_heads) { set_error("sam2b_mask_downscale: engine not finalized or built without the SAM heads"); return 1; }
  if (!mask_prompt || !dense_embedding || n <= 0) { set_error("sam2b_mask_downscale: null/empty argument"); return 1; }
  ENTER_DEVICE(h->device);
  const std::string d = "sam_prompt_encoder.mask_downscaling.";
  const float* w[11] = {W32(h, d + "0.weight"), W32(h, d + "0.bias"), W32(h, d + "1.weight"), W32(h, d + "1.bias"),
                        W32(h, d + "3.weight"), W32(h, d + "3.bias"), W32(h, d + "4.weight"), W32(h, d + "4.bias"),
                        W32(h, d + "6.weight"), W32(h, d + "6.bias"), nullptr};
  return mask_downscale(mask_prompt, w, dense_embedding, n, reinterpret_cast<cudaStream_t>(stream));
}

int sam2b_memory_encode_low_res(sam2b_handle h, const float* pix_feat, int64_t pix_batch_stride, const float* low_res_masks,
                                int binarize, const float* object_score_logits, int n, float* out_x, void* stream) {
  if (!h || !h->finalized) { set_error("sam2b_memory_encode_low_res: engine not finalized"); return 1; }
  if (!pix_feat || !low_res_masks || !out_x || n <= 0) { set_error("sam2b_memory_encode_low_res: null/empty argument"); return 1; }
  if (pix_batch_stride != 0 && pix_batch_stride != 256 * 4096) { set_error("sam2b_memory_encode_low_res: pix_batch_stride must be 0 (one frame for all objects) or 1048576"); return 1; }
  ENTER_DEVICE(h->device);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  ENGINE_CALL(h, st);
  const int kMaxMicroBatch = 16;
  for (int i0 = 0; i0 < n; i0 += kMaxMicroBatch) {
    const int mb = (n - i0) < kMaxMicroBatch ? (n - i0) : kMaxMicroBatch;
    TRY(memory_encode_mb(h, pix_feat + (size_t)i0 * pix_batch_stride, nullptr, mb, out_x + (size_t)i0 * 64 * 4096, st,
                         low_res_masks + (size_t)i0 * 65536, binarize, object_score_logits ? object_score_logits + i0 : nullptr,
                         pix_batch_stride));
  }
  return 0;
}

// ---- CUDA-graph support: replaying a captured sequence of engine calls ----
int sam2b_stream_acquire(sam2b_handle h, void* stream) {
  if (!h) { set_error("sam2b_stream_acquire: null engine"); return 1; }
  ENTER_DEVICE(h->device);
  std::lock_guard<std::mutex> lk(h->mu);
  if (h->done_valid) CUDA_OK(cudaStreamWaitEvent(reinterpret_cast<cudaStream_t>(stream), h->done, 0));
  return 0;
}
int sam2b_stream_release(sam2b_handle h, void* stream) {
  if (!h) { set_error("sam2b_stream_release: null engine"); return 1; }
  ENTER_DEVICE(h->device);
  std::lock_guard<std::mutex> lk(h->mu);
  CUDA_OK(cudaEventRecord(h->done, reinterpret_cast<cudaStream_t>(stream)));
  h->done_valid = true;
  return 0;
}
unsigned long long sam2b_workspace_generation(sam2b_handle h) { return h ? h->ws_generation : 0; }

// ---- profiling ----
int sam2b_profile_begin(void) {
  for (auto& r : g_prof) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  g_prof.clear();
  g_prof_on = true;
  return 0;
}
int sam2b_profile_end(char* buf, int64_t buflen, int64_t* needed) {
  g_prof_on = false;
  cudaDeviceSynchronize();
  std::string out = "name,ms,flops,bytes\n";
  char line[256];
  for (auto& r : g_prof) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, r.a, r.b);
    snprintf(line, sizeof(line), "%s,%.6f,%.6g,%.6g\n", r.name, ms, r.flops, r.bytes);
    out += line;
    cudaEventDestroy(r.a); cudaEventDestroy(r.b);
  }
  g_prof.clear();
  if (needed) *needed = (int64_t)out.size() + 1;
  if (buf && buflen > 0) { snprintf(buf, (size_t)buflen, "%s", out.c_str()); }
  return 0;
}

// ---- building blocks ----
int sam2b_op_gemm(const void* a, int64_t lda, int64_t a_bs, const void* b, int64_t ldb, int64_t b_bs, int m, int n, int k,
                  int batch, float alpha, const float* bias, int bias_along_m, int act, const float* gamma,
                  const float* residual, int64_t ld_res, int64_t res_bs, int res_row_mod, float* out_f32, void* out_bf16,
                  int64_t ldc, int64_t c_bs, int force_block_n, void* stream) {
  GemmArgs g;
  g.A = reinterpret_cast<const bf16*>(a); g.lda = lda; g.a_batch_stride = a_bs;
  g.B = reinterpret_cast<const bf16*>(b); g.ldb = ldb; g.b_batch_stride = b_bs;
  g.M = m; g.N = n; g.K = k; g.batch = batch; g.alpha = alpha;
  g.bias = bias; g.bias_along_m = bias_along_m; g.act = act; g.gamma = gamma;
  g.residual = residual; g.ld_res = ld_res; g.res_batch_stride = res_bs; g.res_row_mod = res_row_mod;
  g.out_f32 = out_f32; g.out_bf16 = reinterpret_cast<bf16*>(out_bf16); g.ldc = ldc; g.c_batch_stride = c_bs;
  g.force_bn = force_block_n;
  return gemm_bf16(g, reinterpret_cast<cudaStream_t>(stream));
}
int sam2b_op_gemm_ln_producer(const void* a, const void* b, int m, int n, int k, const float* bias, const float* residual,
                              float* out_f32, void* out2_bf16, float* stats, int stats_capacity_parts, int* parts,
                              int force_block_n, void* stream) {
  GemmArgs g;
  g.A = reinterpret_cast<const bf16*>(a); g.lda = k; g.B = reinterpret_cast<const bf16*>(b); g.ldb = k;
  g.M = m; g.N = n; g.K = k; g.bias = bias; g.residual = residual; g.ld_res = n;
  g.out_f32 = out_f32; g.ldc = n; g.out2_bf16 = reinterpret_cast<bf16*>(out2_bf16); g.ldc2 = n;
  g.stats_out = stats; g.stats_capacity_parts = stats_capacity_parts; g.stats_parts_out = parts;
  g.force_bn = force_block_n;
  return gemm_bf16(g, reinterpret_cast<cudaStream_t>(stream));
}
int sam2b_op_gemm_ln_consumer(const void* a, const void* b, int m, int n, int k, const float* bias,
                              const float* stats, int parts, float eps, int act, float* out_f32, void* out_bf16,
                              int force_block_n, void* stream) {
  GemmArgs g;
  g.A = reinterpret_cast<const bf16*>(a); g.lda = k; g.B = reinterpret_cast<const bf16*>(b); g.ldb = k;
  g.M = m; g.N = n; g.K = k; g.bias = bias; g.act = act;
  g.out_f32 = out_f32; g.out_bf16 = reinterpret_cast<bf16*>(out_bf16); g.ldc = n;
  g.ln_stats = stats; g.ln_parts = parts; g.ln_eps = eps;
  g.force_bn = force_block_n;
  return gemm_bf16(g, reinterpret_cast<cudaStream_t>(stream));
}
int sam2b_op_layernorm(const float* x, const float* gamma, const float* beta, float eps, void* y_bf16, float* y_f32,
                       int64_t rows, int c, void* stream) {
  return layernorm_rows(x, gamma, beta, eps, reinterpret_cast<bf16*>(y_bf16), y_f32, rows, c, reinterpret_cast<cudaStream_t>(stream));
}
int sam2b_op_cast_bf16(const float* x, void* y, int64_t n, void* stream) {
  return cast_f32_bf16(x, reinterpret_cast<bf16*>(y), n, reinterpret_cast<cudaStream_t>(stream));
}
int sam2b_op_maxpool2x2_nhwc(const float* x, float* y, int b, int hgt, int wid, int c, void* stream) {
  return maxpool2x2_nhwc(x, y, b, hgt, wid, c, reinterpret_cast<cudaStream_t>(stream));
}
int sam2b_op_window_attention(const void* qkv, void* out, int b, int hgt, int wid, int heads, int ws, int q_pool, void* stream) {
  return window_attention_hd72(reinterpret_cast<const bf16*>(qkv), reinterpret_cast<bf16*>(out), b, hgt, wid, heads, ws, q_pool,
                               reinterpret_cast<cudaStream_t>(stream));
}
int sam2b_op_connected_components(const void* mask_u8, int32_t* labels, int32_t* counts, int n, int h, int w, void* stream) {
  if (!mask_u8 || !labels || !counts) { set_error("sam2b_op_connected_components: null argument"); return 1; }
  return connected_components(reinterpret_cast<const uint8_t*>(mask_u8), labels, counts, n, h, w, reinterpret_cast<cudaStream_t>(stream));
}
int sam2b_op_preprocess_u8(const void* hwc_u8, float* out_nchw, int n, int h, int w, int normalise, void* stream) {
  if (!hwc_u8 || !out_nchw) { set_error("sam2b_op_preprocess_u8: null argument"); return 1; }
  return preprocess_u8_hwc(reinterpret_cast<const uint8_t*>(hwc_u8), out_nchw, n, h, w, normalise, reinterpret_cast<cudaStream_t>(stream));
}
int64_t sam2b_op_resize_frames_workspace(int n, int h, int w, int out_h, int out_w) {
  return (int64_t)resize_frames_workspace(n, h, w, out_h, out_w);
}
int sam2b_op_resize_frames(const void* src, int src_is_u8_hwc, int n, int h, int w, float* out_nchw, int out_h, int out_w,
                           int normalise, void* workspace, void* stream) {
  return resize_frames(src, src_is_u8_hwc, n, h, w, out_nchw, out_h, out_w, normalise, workspace, reinterpret_cast<cudaStream_t>(stream));
}
int sam2b_op_upsample_masks(const float* low, float* out, void* mask_u8, int64_t n, int h, int w, int out_h, int out_w, float threshold,
                            void* stream) {
  if (!low || (!out && !mask_u8)) { set_error("sam2b_op_upsample_masks: null argument"); return 1; }
  return upsample_bilinear_mask(low, out, reinterpret_cast<uint8_t*>(mask_u8), (long)n, h, w, out_h, out_w, threshold,
                                reinterpret_cast<cudaStream_t>(stream));
}
int sam2b_position_encoding(int channels, int h, int w, float* out_chw, void* stream) {
  return position_sine(out_chw, channels, h, w, reinterpret_cast<cudaStream_t>(stream));
}
int sam2b_op_softmax_rows(const float* s, int64_t lds, void* p, int64_t ldp, int64_t rows, int cols, void* stream) {
  return softmax_rows(s, lds, reinterpret_cast<bf16*>(p), ldp, rows, cols, reinterpret_cast<cudaStream_t>(stream));
}

int64_t sam2b_op_attention_hd256_workspace(int n, int tq, int64_t lk) {
  if (n <= 0 || tq <= 0 || lk <= 0) return 0;
  return (int64_t)flash_attention_hd256_splits(n, tq, lk) * n * tq * (256 + 2) * 4;
}
int sam2b_op_attention_hd256(const void* q, const void* k, const void* v, int n, int tq, int64_t lk, void* out, void* workspace,
                             void* stream) {
  if (!q || !k || !v || !out || !workspace) { set_error("sam2b_op_attention_hd256: null argument"); return 1; }
  const int splits = flash_attention_hd256_splits(n, tq, lk);
  float* o_part = reinterpret_cast<float*>(workspace);
  float* ml_part = o_part + (size_t)splits * n * tq * 256;
  return flash_attention_hd256(reinterpret_cast<const bf16*>(q), (long)tq * 256, reinterpret_cast<const bf16*>(k), lk * 256,
                               reinterpret_cast<const bf16*>(v), lk * 256, n, tq, lk, o_part, ml_part,
                               reinterpret_cast<bf16*>(out), reinterpret_cast<cudaStream_t>(stream));
}

int64_t sam2b_op_attention_hd256_v64_workspace(int n, int tq, int64_t lk) {
  if (n <= 0 || tq <= 0 || lk <= 0) return 0;
  return (int64_t)flash_attention_hd256_v64_splits(n, tq, lk) * n * tq * (64 + 2) * 4;
}
int sam2b_op_attention_hd256_v64(const void* q, const void* k, const void* v64, int n, int tq, int64_t lk, const int32_t* lk_rows_dev,
                                 void* out, void* workspace, void* stream) {
  if (!q || !k || !v64 || !out || !workspace) { set_error("sam2b_op_attention_hd256_v64: null argument"); return 1; }
  const int splits = flash_attention_hd256_v64_splits(n, tq, lk);
  float* o_part = reinterpret_cast<float*>(workspace);
  float* ml_part = o_part + (size_t)splits * n * tq * 64;
  return flash_attention_hd256_v64(reinterpret_cast<const bf16*>(q), (long)tq * 256, reinterpret_cast<const bf16*>(k), lk * 256,
                                   reinterpret_cast<const bf16*>(v64), lk * 64, n, tq, lk, lk_rows_dev, o_part, ml_part,
                                   reinterpret_cast<bf16*>(out), reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"
