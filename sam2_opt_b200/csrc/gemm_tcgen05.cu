// K1 host side: tensor maps, BLOCK_N choice and dispatch to the per-BLOCK_N translation units (gemm_bn*.cu), which
// instantiate the kernel of gemm_kernel.cuh.
#include <mutex>

#include "gemm_kernel.cuh"

namespace sam2b {

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(f);
  });
  return fn;
}

// 3-D bf16 tensor map over [batch][rows][K] (K contiguous), box = {64, box_rows, 1}, 128B swizzle, zero OOB fill.
int make_tmap(CUtensorMap* m, const bf16* ptr, long rows, long K, long ld, long batch, long batch_stride,
                     int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return 1; }
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (ld * 2) % 16 || (batch > 1 && (batch_stride * 2) % 16)) {
    set_error("gemm operand not 16B aligned (ptr=%p ld=%ld batch_stride=%ld)", (const void*)ptr, ld, batch_stride);
    return 1;
  }
  cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)rows, (cuuint64_t)(batch > 0 ? batch : 1)};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)(batch > 1 ? batch_stride : ld * rows) * 2};
  cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<bf16*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) rows=%ld K=%ld ld=%ld batch=%ld bs=%ld box_rows=%d", (int)r, rows, K,
              ld, batch, batch_stride, box_rows);
    return 1;
  }
  return 0;
}

// 3-D tensor map over the output / residual [batch][rows][cols] (cols contiguous): box = 32 rows x 32 columns,
// 64B swizzle for bf16 (64 B rows), 128B swizzle for fp32 (128 B rows) -- the layouts the epilogue warps write.
int make_tmap_out(CUtensorMap* m, const void* ptr, int is_bf16, long rows, long cols, long ld, long batch,
                         long batch_stride) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return 1; }
  const long es = is_bf16 ? 2 : 4;
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (ld * es) % 16 || (batch > 1 && (batch_stride * es) % 16)) {
    set_error("gemm output/residual not 16B aligned (ptr=%p ld=%ld batch_stride=%ld)", ptr, ld, batch_stride);
    return 1;
  }
  cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)(batch > 0 ? batch : 1)};
  cuuint64_t strides[2] = {(cuuint64_t)(ld * es), (cuuint64_t)((batch > 1 ? batch_stride : ld * rows) * es)};
  cuuint32_t box[3] = {32, 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(ptr),
                   dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   is_bf16 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(out) failed (%d) rows=%ld cols=%ld ld=%ld batch=%ld bs=%ld", (int)r, rows, cols, ld, batch, batch_stride);
    return 1;
  }
  return 0;
}

// Tile choice: BLOCK_N and whether CTA pairs (256-row tiles) are used.  Cost model per tile ~ MMA time (prop. to BN) plus a
// fixed overhead (epilogue drain, barriers); a pair tile moves 1/3 less operand data per flop and has a deeper smem ring.
static void pick_tile(const GemmArgs& a, int* bn_out, int* cta2_out) {
  static const int cand[] = {256, 224, 192, 160, 128, 96, 64, 32};
  static int allow2 = -1;
  if (allow2 < 0) { const char* e = getenv("SAM2B_GEMM_2CTA"); allow2 = e ? atoi(e) : 1; }
  double best = 1e30; int best_bn = 128, best_c2 = 0;
  // A LayerNorm producer writes one partial (sum, sum x^2) per N tile and chunk parity; fixing BLOCK_N by N keeps that
  // partition -- and with it every rounding of the statistics -- independent of M, so an image's result is bit-identical
  // whatever batch it is encoded in.  Least padding wins, ties go to the larger tile: 144, 288 -> 160; 576, 1152 -> 192.
  int ln_bn = 0;
  if (a.out2_bf16) {
    long waste_best = 1L << 40;
    for (int bn : {128, 160, 192}) {
      const long waste = (long)((a.N + bn - 1) / bn) * bn - a.N;
      if (waste <= waste_best) { waste_best = waste; ln_bn = bn; }
    }
  }
  for (int c2 = 0; c2 <= (allow2 ? 1 : 0); ++c2) {
    const int tm = c2 ? 2 * BM : BM;
    const long units = c2 ? num_sms() / 2 : num_sms();
    const long tiles_m = (a.M + tm - 1) / tm;
    for (int bn : cand) {
      if (c2 && bn < 128) continue;                      // pair kernels are instantiated for the large tiles only
      if (a.out2_bf16 && bn != ln_bn) continue;          // LayerNorm producer: BLOCK_N is a function of N alone (see above)
      const long tiles = tiles_m * ((a.N + bn - 1) / bn) * a.batch;
      // measured on B200 (tests/_gemm_bench.py): pairs win 3-7 % on the K >= 1152 stage-4 shapes (long main loop), and lose
      // a few % on the K = 576 / 288 / 144 shapes whose time is set by the epilogue, not by operand traffic
      // and, combined with 16 epilogue warps, 7-11 % on the bf16-output forms from K = 576 up
      const bool pair_ok = a.K >= 1024 || (a.out_bf16 && a.K >= 512);
      if (c2 && (tiles < units || !pair_ok || (a.res_row_mod && a.res_row_mod % tm))) continue;
      const long waves = (tiles + units - 1) / units;
      const double cost = (double)waves * (c2 ? 0.90 * bn + 64.0 : bn + 64.0);
      if (cost < best - 1e-9) { best = cost; best_bn = bn; best_c2 = c2; }
    }
  }
  *bn_out = best_bn; *cta2_out = best_c2;
}

// 16 epilogue warps pay off where the epilogue, not the tensor pipe, sets the tile time.  The choice depends on (N, K, form)
// only -- never on M -- so that the statistics partition of a LayerNorm producer stays batch-independent.
static int pick_ew16(const GemmArgs& a, int bn) {
  static int mode = -1;     // SAM2B_GEMM_EW16: 0 never, 1 heuristic (default), 2 always
  if (mode < 0) { const char* e = getenv("SAM2B_GEMM_EW16"); mode = e ? atoi(e) : 1; }
  if (bn < 128 || mode == 0) return 0;
  if (mode == 2) return 1;
  // measured (tests/_gemm_bench.py): +14..42 % on the K <= 288 bf16-output GEMMs (GELU most), +7..11 % with CTA pairs on the
  // K >= 576 ones; nothing on the fp32 + residual forms, which are HBM-bound -- those keep 8 warps and double-buffered staging
  return a.out_bf16 != nullptr && !a.residual;
}

int gemm_bf16(const GemmArgs& a, cudaStream_t stream) {
  if (a.M <= 0 || a.N <= 0 || a.K <= 0 || a.batch <= 0) { set_error("gemm: empty problem M=%d N=%d K=%d batch=%d", a.M, a.N, a.K, a.batch); return 1; }
  if (!a.out_f32 == !a.out_bf16) { set_error("gemm: exactly one of out_f32 / out_bf16 must be given"); return 1; }
  if (a.residual && a.out_bf16) { set_error("gemm: a residual needs an fp32 output"); return 1; }
  if (a.residual && a.res_row_mod && (a.res_row_mod % BM)) {
    set_error("gemm: res_row_mod=%d must be a multiple of %d", a.res_row_mod, BM);
    return 1;
  }
  if (a.rope_table && (a.residual || !a.out_bf16 || a.N % 256 || a.rope_positions <= 0 || a.rope_positions % (2 * BM) || a.ln_stats || a.out2_bf16)) {
    set_error("gemm: the RoPE epilogue needs a bf16 output, N %% 256 == 0, no residual, a table of a multiple of %d positions", 2 * BM);
    return 1;
  }
  int bn = 0, cta2 = 0, ew16 = 0;
  if (a.force_bn) {     // tests / tuning: BLOCK_N + 1000 (CTA pairs) + 2000 (16 epilogue warps)
    bn = a.force_bn % 1000; cta2 = (a.force_bn / 1000) & 1; ew16 = (a.force_bn / 2000) & 1;
  } else {
    pick_tile(a, &bn, &cta2);
    ew16 = pick_ew16(a, bn);
  }
  if (ew16 && bn < 128) ew16 = 0;
  if (cta2 && (bn < 128 || (a.res_row_mod && a.res_row_mod % (2 * BM)))) { set_error("gemm: CTA pairs need BLOCK_N >= 128 and res_row_mod %% 256 == 0"); return 1; }
  // compile-time specialised epilogues for the trunk's three dominant GEMM forms (fewer instructions per element)
  int mode = MODE_GENERIC;
  const bool plain = a.alpha == 1.0f && a.bias && !a.bias_along_m && !a.gamma && (a.N % 4) == 0 && !a.rope_table && !getenv("SAM2B_GEMM_GENERIC");
  if (plain && a.out_bf16 && !a.residual && a.act == ACT_NONE) mode = MODE_BF16_BIAS;
  else if (plain && a.out_bf16 && !a.residual && a.act == ACT_GELU) mode = MODE_BF16_BIAS_GELU;
  else if (plain && a.out_f32 && a.residual && a.act == ACT_NONE) mode = a.out2_bf16 ? MODE_F32_BIAS_RES_LN : MODE_F32_BIAS_RES;
  if (a.out2_bf16) {
    if (mode != MODE_F32_BIAS_RES_LN || bn < 128) { set_error("gemm: the LayerNorm-producer epilogue needs bias + fp32 residual, no activation, BLOCK_N >= 128"); return 1; }
    const int parts = (ew16 ? 4 : 2) * ((a.N + bn - 1) / bn);
    if (a.stats_out && parts > a.stats_capacity_parts) { set_error("gemm: stats buffer holds %d parts per row, %d needed", a.stats_capacity_parts, parts); return 1; }
    if (a.stats_parts_out) *a.stats_parts_out = parts;
  }
  if (a.ln_stats && (a.residual || a.ln_parts <= 0 || !a.bias || a.bias_along_m)) { set_error("gemm: LayerNorm-fold consumer needs bias along N, parts > 0 and no residual"); return 1; }
  switch (bn) {
    case 256: return launch_gemm_bn256(a, mode, cta2, ew16, stream);
    case 224: return launch_gemm_bn224(a, mode, cta2, ew16, stream);
    case 192: return launch_gemm_bn192(a, mode, cta2, ew16, stream);
    case 160: return launch_gemm_bn160(a, mode, cta2, ew16, stream);
    case 128: return launch_gemm_bn128(a, mode, cta2, ew16, stream);
    case 96: return launch_gemm_bn96(a, mode, cta2, ew16, stream);
    case 64: return launch_gemm_bn64(a, mode, cta2, ew16, stream);
    case 32: return launch_gemm_bn32(a, mode, cta2, ew16, stream);
  }
  set_error("gemm: unsupported BLOCK_N %d (256, 224, 192, 160, 128, 96, 64, 32)", bn);
  return 1;
}

}  // namespace sam2b
