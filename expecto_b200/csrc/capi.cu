// C ABI implementation (include/beluga_b200.h): context, weight staging, layer launches.
// Reference path replaced: chromatin.py:102-106 (model load), :262-279 (encode + batched forward),
// :281-286 (diff/ref/alt), compute_expecto_features.py:113-124, predict.py:87-136.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <atomic>
#include <algorithm>
#include <vector>

#include "../../include/beluga_b200.h"
#include "gemm_tc.cuh"
#include "kernels_misc.cuh"
#include "genome_pack.cuh"
#include "variants.cuh"
#include "score_tc.cuh"
#include <cub/device/device_scan.cuh>

using namespace bx;

namespace {

constexpr int kMaxDevices = 64;
thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define CK(expr)                                                                                      \
  do {                                                                                                \
    cudaError_t _e = (expr);                                                                          \
    if (_e != cudaSuccess)                                                                            \
      return fail(BX_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

// Entry points run on their context's device and put the caller's current device back on return.
struct DeviceGuard {
  int prev = -1;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int device) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    err = cudaSetDevice(device);
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};
#define ON_DEVICE(dev)          \
  DeviceGuard _guard(dev);      \
  CK(_guard.err)

PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;

int load_driver_entry() {
  if (g_encode) return 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
    return fail(BX_E_CUDA, "cuTensorMapEncodeTiled not available from the driver (%s)", cudaGetErrorString(e));
  g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  return 0;
}

// 2D K-major 16-bit matrix [rows, cols] (row pitch ld elements), box [box_rows, 64 cols], SWIZZLE_128B.
int make_tmap(CUtensorMap* tm, int dtype, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
              uint32_t box_rows) {
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {64, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(tm, dtype == BX_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
                        2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(BX_E_CUDA, "cuTensorMapEncodeTiled failed: %d (rows=%llu cols=%llu ld=%llu box_rows=%u)", (int)r,
                (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, box_rows);
  return 0;
}

}  // namespace

// Layer table (SURVEY §3.2 / Beluga.py:23-45).  Row pitches per window are chosen so that pooled
// layers see a pitch that is a multiple of 4 and everything else carries as few junk rows as possible.
struct LayerDef {
  int taps, cin, cout, block_n, epi;
  int pitch_in, pitch_out, l_store;
};
static const LayerDef kLayers[6] = {
    {1, 64, 320, 160, EPI_BIAS_RELU, 1996, 1996, 1996},       // conv1 on the im2col one-hot (K = 32 valid)
    {8, 320, 320, 160, EPI_POOL4_BIAS_RELU, 1996, 496, 496},  // conv2 + ReLU + MaxPool4
    {8, 320, 480, 240, EPI_BIAS_RELU, 496, 492, 492},         // conv3 + ReLU (489 valid)
    {8, 480, 480, 240, EPI_POOL4_BIAS_RELU, 492, 120, 120},   // conv4 + ReLU + MaxPool4
    {8, 480, 640, 160, EPI_BIAS_RELU, 120, 113, 113},         // conv5 + ReLU
    {8, 640, 640, 160, EPI_BIAS_RELU, 113, 106, 106},         // conv6 + ReLU
};

struct bx_ctx {
  int device = 0, dtype = 0, cg = 2, cap = 0, num_sms = 0;
  // Split-operand mode (BX_DTYPE_F16X3): layer i (0..5 = conv1..6, 6 = fc1, 7 = fc2) reads its activation and its
  // weights as fp16 (hi, lo) pairs and issues Ah*Wh + Ah*Wl + Al*Wh; entry 8 (the probabilities) is never split.
  // conv1's activation is the exact one-hot, so only its weights are split (K = 32 -> 64).
  bool split_in[9] = {};
  bool split_any = false;
  bool weights_ready = false;
  void* wconv[6] = {};   // 16-bit [taps*cout, cin]
  float* bconv[6] = {};
  void* wfc1 = nullptr;  // 16-bit [2048, 67840]
  float* bfc1 = nullptr; // [2048]
  void* wfc2 = nullptr;  // 16-bit [2048, 2048]
  float* bfc2 = nullptr; // [2048]
  void* act[7] = {};     // act[0] = im2col one-hot [cap*1996,64]; act[i] = output of conv i
  float* acc1 = nullptr; // fp32 [cap, 2048]
  float* acc1_view = nullptr;  // partial sums the next finaliser reads (acc1, or a slab of acc_ref in delta mode)
  void* hid = nullptr;   // 16-bit [cap, 2048]
  const uint32_t* bases2 = nullptr;
  const uint32_t* nmask = nullptr;
  int64_t n_bases = 0;
  // per-stage CUDA-event timing (bx_ctx_profile): stage 0 = encode/im2col, 1..6 = conv1..6, 7 = fc1, 8 = fc2
  bool prof_on = false;
  struct ProfRec { int stage; cudaEvent_t e0, e1; };
  std::vector<ProfRec> prof_recs;
  std::vector<cudaEvent_t> prof_pool;
  double prof_ms[BX_NSTAGES] = {};
  int64_t prof_cnt[BX_NSTAGES] = {};
  // host-buffer pipeline (bx_*_host): copy streams, double-buffered staging
  cudaStream_t s_comp = nullptr, s_in = nullptr, s_out = nullptr;
  cudaEvent_t ev_in[2] = {}, ev_comp[2] = {}, ev_out[2] = {};
  void* stage_in[2] = {};
  size_t stage_in_bytes = 0;
  float* stage_out[2] = {};
  size_t stage_out_bytes = 0;
  float* probs_scratch = nullptr;
  size_t probs_scratch_bytes = 0;
  uint8_t* alleles_stage = nullptr;  // bx_chromatin_shift_host: the call's allele codes (grown on demand, kept)
  size_t alleles_stage_bytes = 0;
  // segment mode: phase-B conv6 output and the dense FC1 operand gathered from the segment activations
  void* act6b = nullptr;
  void* fc_in = nullptr;
  void* act4h = nullptr;  // conv4 half-pooled (row pairs), both pooling phases are finished from it
  void* act4b = nullptr;  // phase-B pooled conv4 output
  float* acc_ref = nullptr;     // delta mode: FC1 partial sums of a chunk's ref windows [8][n_ref][2048]
  size_t acc_ref_bytes = 0;
  float* probs_dense = nullptr; // delta mode: FC2 output of one piece in dense order, before the scatter
  void* patch6a = nullptr;  // delta mode: conv6 outputs of the alt patches (phase A / B)
  void* patch6b = nullptr;
  size_t patch6_bytes = 0;
  int* si_of_k_dev = nullptr;   // [200] shift index of each window offset (EPI_CHROMATIN, delta mode's alt rows)
  float* ref_scratch = nullptr; // averaged outputs in delta mode: per-strand ref probabilities of a chunk
  size_t ref_scratch_bytes = 0;
  struct ChromPlan* plan = nullptr;  // bx_chromatin_open .. bx_chromatin_close
};

// Where the variant path's FC2 epilogue puts ref / alt / diff (see ChromEpi): device arrays, rows ld floats apart.
struct ChromSink {
  float* ref; float* alt; float* diff;
  const int* dst;                                  // device [group size]: output row of each group member (or null)
  long long shift_stride, strand_stride, row0;
  int average;
  long long ld = BX_FEATURES;
};

// RAII-ish stage timer: records an event pair around a stage when profiling is on.
struct StageTimer {
  bx_ctx* c; cudaStream_t st; cudaEvent_t e1 = nullptr;
  StageTimer(bx_ctx* ctx, int stage, cudaStream_t s) : c(ctx), st(s) {
    if (!c->prof_on) return;
    cudaEvent_t ev[2];
    for (int i = 0; i < 2; ++i) {
      if (!c->prof_pool.empty()) { ev[i] = c->prof_pool.back(); c->prof_pool.pop_back(); }
      else cudaEventCreate(&ev[i]);
    }
    cudaEventRecord(ev[0], st);
    e1 = ev[1];
    c->prof_recs.push_back({stage, ev[0], ev[1]});
  }
  ~StageTimer() { if (e1) cudaEventRecord(e1, st); }
};

static size_t act_rows(int i) { return i == 0 ? kLayers[0].pitch_in : kLayers[i - 1].pitch_out; }
static size_t act_cols(int i) { return i == 0 ? 64 : kLayers[i - 1].cout; }
// physical row width of activation i (the input of layer i): [hi | lo] when layer i runs on split operands
static int act_mul(const bx_ctx* c, int i) { return (i >= 1 && c->split_in[i]) ? 2 : 1; }
static size_t act_width(const bx_ctx* c, int i) { return act_cols(i) * act_mul(c, i); }

// ------------------------------------------------------------------ GEMM launch
// a_cols / b_cols: physical row widths (elements) of the A and B matrices ([hi | lo] pairs in split-operand mode).
template <typename T, int TAPS, int BN, int CG, int EPI, int A_MODE = 0, int SPLIT_OUT = 0, int PARTS = 1>
static int launch_gemm_t(const bx_ctx* c, const void* a, uint64_t a_rows, uint64_t a_cols, const void* b,
                         uint64_t b_rows, uint64_t b_cols, GemmParams p, cudaStream_t st,
                         const EncParams& enc = EncParams()) {
  using Cfg = GemmCfg<TAPS, BN, CG>;
  CUtensorMap tmA, tmB;
  if (int r = make_tmap(&tmB, c->dtype, b, b_rows, b_cols, b_cols, Cfg::B_ROWS)) return r;
  if (A_MODE != 0) tmA = tmB;  // unused: the producer warp builds the A tile in shared memory
  else if (int r = make_tmap(&tmA, c->dtype, a, a_rows, a_cols, a_cols, Cfg::A_ROWS)) return r;
  if (p.parts < 1) p.parts = 1;
  if (p.parts != PARTS) return fail(BX_E_STATE, "kernel instance with %d operand parts launched for %d", PARTS, p.parts);
  if (p.a_grp < 1) { p.a_grp = 1 << 30; p.a_grp_stride = 0; }
  auto kern = gemm_tc_kernel<T, TAPS, BN, CG, EPI, A_MODE, SPLIT_OUT, PARTS>;
  // function attributes are per device: one flag per device, so contexts on several GPUs of one process all get it
  static std::atomic<bool> attr_set[kMaxDevices];
  if (c->device < 0 || c->device >= kMaxDevices) return fail(BX_E_ARG, "device index %d out of range", c->device);
  if (!attr_set[c->device].load(std::memory_order_acquire)) {
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_set[c->device].store(true, std::memory_order_release);
  }
  p.m_tiles = (p.m_rows + 128 * CG - 1) / (128 * CG);
  p.n_chunks = p.n_total / BN;
  static const int dbg = getenv("BX_DEBUG_FLAGS") ? atoi(getenv("BX_DEBUG_FLAGS")) : 0;  // timing diagnostics only
  p.debug_flags = dbg;
  int units = (A_MODE != 0) ? p.m_tiles : p.m_tiles * p.n_chunks * p.k_splits;
  if (p.n_tile_ranges > 0) {
    if (p.n_tile_ranges != p.m_tiles) return fail(BX_E_ARG, "tile ranges (%d) != M tiles (%d)", p.n_tile_ranges, p.m_tiles);
    units = 0;
    for (int t = 0; t < p.n_tile_ranges; ++t) units += p.tile_n[t] * p.n_chunks;
  }
  int clusters = std::min(units, c->num_sms / CG);
  if (clusters < 1) clusters = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(clusters * CG);
  cfg.blockDim = dim3(192);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  CK(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, p, enc));
  g_launches++;
  return 0;
}

template <typename T, int CG>
static int launch_gemm_tc(const bx_ctx* c, int taps, int bn, int epi, const void* a, uint64_t a_rows,
                          uint64_t a_cols, const void* b, uint64_t b_rows, uint64_t b_cols, const GemmParams& p,
                          bool split_out, cudaStream_t st) {
#define BX_CASE(TP, N, E) \
  if (taps == TP && bn == N && epi == E && !split_out && p.parts <= 1) \
    return launch_gemm_t<T, TP, N, CG, E>(c, a, a_rows, a_cols, b, b_rows, b_cols, p, st)
  BX_CASE(1, 160, EPI_BIAS_RELU);
  BX_CASE(8, 160, EPI_POOL4_BIAS_RELU);
  BX_CASE(8, 240, EPI_BIAS_RELU);
  BX_CASE(8, 240, EPI_POOL4_BIAS_RELU);
  BX_CASE(8, 240, EPI_POOL2_BIAS_RELU);
  BX_CASE(8, 160, EPI_BIAS_RELU);
  BX_CASE(1, 256, EPI_PARTIAL_F32);
  BX_CASE(1, 256, EPI_BIAS_SIGMOID);
  BX_CASE(1, 256, EPI_CHROMATIN);
#undef BX_CASE
  if constexpr (std::is_same<T, __half>::value) {  // split-operand mode exists for fp16 operands only
    // split-operand instances: conv layers read and write (hi, lo) pairs, the dense layers read them
#define BX_CASE(TP, N, E) \
  if (taps == TP && bn == N && epi == E && split_out && p.parts == 3) \
    return launch_gemm_t<T, TP, N, CG, E, 0, 1, 3>(c, a, a_rows, a_cols, b, b_rows, b_cols, p, st)
#define BX_CASE_FC(E) \
  if (taps == 1 && bn == 256 && epi == E && !split_out && p.parts == 3) \
    return launch_gemm_t<T, 1, 256, CG, E, 0, 0, 3>(c, a, a_rows, a_cols, b, b_rows, b_cols, p, st)
    BX_CASE_FC(EPI_PARTIAL_F32);
    BX_CASE_FC(EPI_BIAS_SIGMOID);
    BX_CASE_FC(EPI_CHROMATIN);
#undef BX_CASE_FC
    BX_CASE(8, 160, EPI_POOL4_BIAS_RELU);
    BX_CASE(8, 240, EPI_BIAS_RELU);
    BX_CASE(8, 240, EPI_POOL4_BIAS_RELU);
    BX_CASE(8, 240, EPI_POOL2_BIAS_RELU);
    BX_CASE(8, 160, EPI_BIAS_RELU);
#undef BX_CASE
  }
  return fail(BX_E_ARG, "no kernel instance for taps=%d block_n=%d epi=%d split_out=%d parts=%d", taps, bn, epi,
              (int)split_out, p.parts);
}

static int launch_gemm(const bx_ctx* c, int taps, int bn, int epi, const void* a, uint64_t a_rows, uint64_t a_cols,
                       const void* b, uint64_t b_rows, uint64_t b_cols, const GemmParams& p, bool split_out,
                       cudaStream_t st) {
  if (c->dtype == BX_DTYPE_F16) {
    if (c->cg == 1)
      return launch_gemm_tc<__half, 1>(c, taps, bn, epi, a, a_rows, a_cols, b, b_rows, b_cols, p, split_out, st);
    return launch_gemm_tc<__half, 2>(c, taps, bn, epi, a, a_rows, a_cols, b, b_rows, b_cols, p, split_out, st);
  }
  if (c->cg == 1)
    return launch_gemm_tc<__nv_bfloat16, 1>(c, taps, bn, epi, a, a_rows, a_cols, b, b_rows, b_cols, p, split_out, st);
  return launch_gemm_tc<__nv_bfloat16, 2>(c, taps, bn, epi, a, a_rows, a_cols, b, b_rows, b_cols, p, split_out, st);
}

// Split-operand fields of conv layer i (1..5): A = act[i] [rows, cin * mul], B = wconv[i] [8*cout, cin * mul].
static void conv_split_params(const bx_ctx* c, int i, GemmParams& p) {
  const LayerDef& L = kLayers[i];
  p.parts = (i >= 1 && c->split_in[i]) ? 3 : 1;
  p.a_grp = 1 << 30;
  p.a_grp_stride = 0;
  p.a_lo_off = L.cin;
  p.b_lo_off = L.cin;
  p.ld_out = L.cout * (c->split_in[i + 1] ? 2 : 1);
  p.out_lo_off = L.cout;
}

// conv1 with the operand generated inside the kernel (north-star item 1: extraction + one-hot + reverse
// complement fused into conv1's staged input): amode 1 = descriptors over the packed genome, 2 = one-hot bytes.
static int launch_conv1_gen(const bx_ctx* c, int amode, const EncParams& enc, const GemmParams& p, cudaStream_t st) {
  const LayerDef& L = kLayers[0];
  if (p.pitch_in < 136) return fail(BX_E_ARG, "conv1 row pitch %d too small for in-kernel encoding", p.pitch_in);
#define BX_GEN(T, CG, M) \
  return launch_gemm_t<T, 1, 160, CG, EPI_BIAS_RELU, M>(c, nullptr, 0, L.cin, c->wconv[0], L.cout, L.cin, p, st, enc)
#define BX_GEN_S(CG, M) \
  return launch_gemm_t<__half, 1, 160, CG, EPI_BIAS_RELU, M, 1>(c, nullptr, 0, L.cin, c->wconv[0], L.cout, L.cin, p, st, enc)
  const bool f16 = c->dtype == BX_DTYPE_F16;
  if (c->split_in[1]) {  // conv2 reads conv1's output as (hi, lo) pairs
    if (amode == 1) { if (c->cg == 1) BX_GEN_S(1, 1); BX_GEN_S(2, 1); }
    if (c->cg == 1) BX_GEN_S(1, 2);
    BX_GEN_S(2, 2);
  }
  if (amode == 1) {
    if (f16) { if (c->cg == 1) BX_GEN(__half, 1, 1); BX_GEN(__half, 2, 1); }
    if (c->cg == 1) BX_GEN(__nv_bfloat16, 1, 1);
    BX_GEN(__nv_bfloat16, 2, 1);
  }
  if (f16) { if (c->cg == 1) BX_GEN(__half, 1, 2); BX_GEN(__half, 2, 2); }
  if (c->cg == 1) BX_GEN(__nv_bfloat16, 1, 2);
  BX_GEN(__nv_bfloat16, 2, 2);
#undef BX_GEN
#undef BX_GEN_S
}

// BX_CONV1_IM2COL=1 keeps the two-kernel form (encode -> im2col matrix in HBM -> TMA) for A/B comparisons.
static bool conv1_gen_enabled(const bx_ctx* c) {
  static const bool on = !(getenv("BX_CONV1_IM2COL") && atoi(getenv("BX_CONV1_IM2COL")) != 0);
  return on || c->split_any;  // the two-kernel form has no split-operand variant
}

// conv layer i (0..5) over n windows: act[i] -> act[i+1]; `enc` (conv1 only) selects in-kernel operand generation
static int run_conv(bx_ctx* c, int i, int n, cudaStream_t st, const EncParams* enc = nullptr, int amode = 0) {
  const LayerDef& L = kLayers[i];
  GemmParams p = {};
  p.m_rows = n * L.pitch_in;
  p.k_valid = (i == 0) ? (c->split_in[0] ? 64 : 32) : L.cin;
  p.n_total = L.cout;
  p.k_splits = 1;
  p.k_blocks = (p.k_valid + 63) / 64;
  p.kb_per_split = p.k_blocks;
  p.pitch_in = L.pitch_in;
  p.pitch_out = L.pitch_out;
  p.l_store = L.l_store;
  p.n_store = L.cout;
  p.bias = c->bconv[i];
  p.out = c->act[i + 1];
  conv_split_params(c, i, p);
  if (i == 0 && enc) return launch_conv1_gen(c, amode, *enc, p, st);
  const uint64_t width = (uint64_t)L.cin * act_mul(c, i);
  return launch_gemm(c, L.taps, L.block_n, L.epi, c->act[i], (uint64_t)n * L.pitch_in, width, c->wconv[i],
                     (uint64_t)L.taps * L.cout, width, p, c->split_in[i + 1], st);
}

template <typename T>
static int run_fc1_finalize(bx_ctx* c, int n, int k_splits, size_t split_stride, cudaStream_t st) {
  const size_t total = (size_t)n * 2048;
  fc1_finalize_kernel<T><<<(unsigned)((total / 4 + 255) / 256), 256, 0, st>>>(c->acc1_view, k_splits, split_stride, c->bfc1,
                                                                             (T*)c->hid, total, c->split_in[7] ? 1 : 0);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

// fc1 always splits K into the same kFc1Splits slices, whatever the batch: the fp32 summation order of every
// output is then independent of batch size and position (bit-identical results across batchings).
constexpr int kFc1Splits = 8;
static size_t fc1_scratch_rows(int cap) { return (size_t)cap * kFc1Splits; }

// fc1 GEMM: [n, 67840] x [2048, 67840]^T; each K slice writes its own fp32 partial into out[ks][row][2048]
// (slices split_stride elements apart).  ranges: optional per-M-tile slice ranges (GemmParams::tile_lo/tile_n).
static int launch_fc1_gemm(bx_ctx* c, int n, float* out, size_t split_stride, const GemmParams* ranges, cudaStream_t st) {
  GemmParams p = {};
  p.m_rows = n;
  p.k_valid = 67840;
  p.n_total = 2048;
  p.k_blocks = 67840 / 64;
  p.kb_per_split = (p.k_blocks + kFc1Splits - 1) / kFc1Splits;
  p.k_splits = (p.k_blocks + p.kb_per_split - 1) / p.kb_per_split;
  p.split_stride = (long long)split_stride;
  p.pitch_in = p.pitch_out = p.l_store = 1 << 30;
  p.ld_out = 2048;
  p.n_store = 2048;
  p.bias = nullptr;
  p.out = out;
  if (ranges) {
    p.n_tile_ranges = ranges->n_tile_ranges;
    memcpy(p.tile_lo, ranges->tile_lo, sizeof(p.tile_lo));
    memcpy(p.tile_n, ranges->tile_n, sizeof(p.tile_n));
  }
  // operand rows are 106 conv6 rows of [hi(640) | lo(640)] in split mode: ten K-blocks per conv6 row
  const int mul = c->split_in[6] ? 2 : 1;
  p.parts = c->split_in[6] ? 3 : 1;
  p.a_grp = 10;
  p.a_grp_stride = 640 * mul;
  p.a_lo_off = 640;
  p.b_lo_off = 67840;
  return launch_gemm(c, 1, 256, EPI_PARTIAL_F32, c->act[6], n, (uint64_t)67840 * mul, c->wfc1, 2048,
                     (uint64_t)67840 * mul, p, false, st);
}
constexpr int kFc1KbPerSplit = (67840 / 64 + kFc1Splits - 1) / kFc1Splits;

static int run_fc1(bx_ctx* c, int n, cudaStream_t st, float* part = nullptr, size_t split_stride = 0) {
  // the finaliser adds the slices in index order (deterministic, no atomics)
  if (!part) { part = c->acc1; split_stride = (size_t)n * 2048; }
  if (int r = launch_fc1_gemm(c, n, part, split_stride, nullptr, st)) return r;
  if (c->dtype == BX_DTYPE_F16) {
    c->acc1_view = part;
    return run_fc1_finalize<__half>(c, n, kFc1Splits, split_stride, st);
  }
  c->acc1_view = part;
  return run_fc1_finalize<__nv_bfloat16>(c, n, kFc1Splits, split_stride, st);
}

static int run_fc2(bx_ctx* c, int n, float* probs, cudaStream_t st, const ChromEpi* chrom = nullptr) {
  GemmParams p = {};
  p.m_rows = n;
  p.k_valid = 2048;
  p.n_total = 2048;
  p.k_blocks = 32;
  p.kb_per_split = 32;
  p.k_splits = 1;
  p.pitch_in = p.pitch_out = p.l_store = 1 << 30;
  p.ld_out = BX_FEATURES;
  p.n_store = BX_FEATURES;
  p.bias = c->bfc2;
  p.out = probs;
  const int mul = c->split_in[7] ? 2 : 1;
  p.parts = c->split_in[7] ? 3 : 1;
  p.a_grp = 1 << 30;
  p.a_lo_off = 2048;
  p.b_lo_off = 2048;
  if (chrom) p.chrom = *chrom;
  return launch_gemm(c, 1, 256, chrom ? EPI_CHROMATIN : EPI_BIAS_SIGMOID, c->hid, n, (uint64_t)2048 * mul, c->wfc2, 2048,
                     (uint64_t)2048 * mul, p, false, st);
}

// FC2 epilogue parameters of a launch whose rows start at index `first` of row order `mode` (ChromEpi).
static ChromEpi make_chrom(const bx_ctx* c, const ChromSink& s, int mode, int ng, int n_shifts, long long first,
                           int si_fixed = 0) {
  ChromEpi e = {};
  e.ref = s.ref; e.alt = s.alt; e.diff = s.diff;
  e.ref_scratch = c->ref_scratch;
  e.dst = s.dst;
  e.si_of_k = c->si_of_k_dev;
  e.shift_stride = s.shift_stride; e.strand_stride = s.strand_stride; e.row0 = s.row0;
  e.ld = s.ld;
  e.first = first;
  e.mode = mode; e.ng = ng; e.S = n_shifts; e.si_fixed = si_fixed; e.average = s.average;
  return e;
}

// conv1..conv6, fc1, fc2 over n windows whose conv1 operand is already in act[0] or is generated from `enc`.
static int run_network(bx_ctx* c, int n, float* probs, cudaStream_t st, const EncParams* enc = nullptr,
                       int amode = 0, const ChromEpi* chrom = nullptr) {
  for (int i = 0; i < 6; ++i) {
    StageTimer t(c, 1 + i, st);
    if (int r = run_conv(c, i, n, st, enc, amode)) return r;
  }
  {
    StageTimer t(c, 7, st);
    if (int r = run_fc1(c, n, st)) return r;
  }
  StageTimer t(c, 8, st);
  return run_fc2(c, n, probs, st, chrom);
}

// ------------------------------------------------------------------ segment mode
// Row pitches of every activation for "windows" of L positions (L = 2000 reproduces kLayers).
struct Geom {
  int L, p1, p2, p3, p4, p5, p6, l4b, ph;
};
static Geom make_geom(int L) {
  Geom g;
  g.L = L;
  g.p1 = (L - 7 + 3) / 4 * 4;       // conv1 rows (L-7 valid), pooled by conv2's epilogue -> multiple of 4
  g.p2 = (L - 14) / 4;               // conv2 + MaxPool4
  g.p3 = (g.p2 - 7 + 3) / 4 * 4;     // conv3 rows, pooled by conv4's epilogue
  g.p4 = (g.p2 - 14) / 4;            // conv4 + MaxPool4 (phase A: pooling groups start at row 0)
  g.l4b = (g.p2 - 16) / 4;           // phase B: pooling groups start at row 2
  g.ph = (g.p2 - 14) / 2;            // conv4 rows pooled in pairs (segment mode: both phases come from these)
  g.p5 = g.p4 - 7;
  g.p6 = g.p5 - 7;
  return g;
}

// conv layer i over n_seg segments with explicit geometry; phase_b shifts conv4's input by two rows.
static int run_conv_geom(bx_ctx* c, int i, int n_seg, const Geom& g, const void* in, void* out, bool phase_b,
                         cudaStream_t st, const EncParams* enc = nullptr, bool half_pool = false) {
  const LayerDef& L = kLayers[i];
  const int pin[6] = {g.p1, g.p1, g.p2, g.p3, g.p4, g.p5};
  const int pout[6] = {g.p1, g.p2, g.p3, g.p4, g.p5, g.p6};
  GemmParams p = {};
  const int skip = phase_b ? 2 : 0;
  p.m_rows = n_seg * pin[i] - skip;
  p.k_valid = (i == 0) ? (c->split_in[0] ? 64 : 32) : L.cin;
  p.n_total = L.cout;
  p.k_splits = 1;
  p.k_blocks = (p.k_valid + 63) / 64;
  p.kb_per_split = p.k_blocks;
  p.pitch_in = pin[i];
  p.pitch_out = pout[i];
  p.l_store = (i == 3 && phase_b) ? g.l4b : pout[i];
  p.n_store = L.cout;
  p.bias = c->bconv[i];
  p.out = out;
  conv_split_params(c, i, p);
  if (i == 0 && enc) return launch_conv1_gen(c, 1, *enc, p, st);
  const uint64_t width = (uint64_t)L.cin * act_mul(c, i);
  const char* a = reinterpret_cast<const char*>(in) + (size_t)skip * width * 2;
  int epi = L.epi;
  if (half_pool) {  // conv4 in segment mode: MaxPool over row pairs only, [n_seg, ph, 480]
    epi = EPI_POOL2_BIAS_RELU;
    p.pitch_out = p.l_store = g.ph;
  }
  return launch_gemm(c, L.taps, L.block_n, epi, a, (uint64_t)p.m_rows, width, c->wconv[i], (uint64_t)L.taps * L.cout,
                     width, p, c->split_in[i + 1], st);
}

// Largest number of L-position segments whose activations fit the scratch sized for `cap` 2000-bp windows.
static int max_segments(const bx_ctx* c, const Geom& g) {
  const int need[7] = {g.p1, g.p1, g.p2, g.p3, g.p4, g.p5, g.p6};
  long long best = 1LL << 40;
  for (int i = 0; i < 7; ++i) best = std::min(best, (long long)c->cap * (long long)act_rows(i) / need[i]);
  return (int)std::max(0LL, best);
}

// ------------------------------------------------------------------ C ABI
extern "C" {

int bx_version(void) { return 100; }
const char* bx_last_error(void) { return g_err; }
int64_t bx_launch_count(void) { return g_launches.load(); }

int bx_ctx_create(int device, int dtype, int cta_group, int capacity_windows, bx_ctx** out) {
  if (!out) return fail(BX_E_ARG, "out is null");
  *out = nullptr;
  if (dtype != BX_DTYPE_F16 && dtype != BX_DTYPE_BF16 && dtype != BX_DTYPE_F16X3)
    return fail(BX_E_ARG, "dtype must be 0 (f16), 1 (bf16) or 2 (f16x3: split fp16 operands)");
  if (cta_group != 1 && cta_group != 2) return fail(BX_E_ARG, "cta_group must be 1 or 2");
  if (capacity_windows < 1) return fail(BX_E_ARG, "capacity_windows must be >= 1");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(BX_E_CUDA, "no CUDA device (%s): this library has no CPU fallback", cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return fail(BX_E_ARG, "device %d out of range (%d devices)", device, ndev);
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(BX_E_CUDA, "device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major,
                prop.minor);
  ON_DEVICE(device);
  if (int r = load_driver_entry()) return r;
  bx_ctx* c = new bx_ctx();
  c->device = device;
  c->dtype = dtype == BX_DTYPE_F16X3 ? BX_DTYPE_F16 : dtype;
  if (dtype == BX_DTYPE_F16X3) {
    for (int i = 0; i < 8; ++i) c->split_in[i] = true;   // every layer (the kernel instances exist for all-or-nothing)
    c->split_any = true;
  }
  c->cg = cta_group;
  c->cap = capacity_windows;
  c->num_sms = prop.multiProcessorCount;
  const size_t cap = capacity_windows;
  for (int i = 0; i < 7; ++i) {
    // +256 rows of slack is not needed: TMA zero-fills rows past the tensor end.
    const size_t bytes = cap * act_rows(i) * act_width(c, i) * 2;
    if (cudaMalloc(&c->act[i], bytes) != cudaSuccess) {
      bx_ctx_destroy(c);
      return fail(BX_E_CUDA, "cudaMalloc of activation %d (%zu bytes) failed", i, bytes);
    }
    cudaMemset(c->act[i], 0, bytes);
  }
  if (cudaMalloc(&c->acc1, fc1_scratch_rows(capacity_windows) * 2048 * 4) != cudaSuccess || cudaMalloc(&c->hid, cap * 2048 * 2 * (c->split_in[7] ? 2 : 1)) != cudaSuccess) {
    bx_ctx_destroy(c);
    return fail(BX_E_CUDA, "cudaMalloc of fc scratch failed");
  }
  *out = c;
  return 0;
}

int bx_ctx_destroy(bx_ctx* c) {
  if (!c) return 0;
  DeviceGuard guard(c->device);
  for (int i = 0; i < 6; ++i) { cudaFree(c->wconv[i]); cudaFree(c->bconv[i]); }
  cudaFree(c->wfc1); cudaFree(c->bfc1); cudaFree(c->wfc2); cudaFree(c->bfc2);
  for (int i = 0; i < 7; ++i) cudaFree(c->act[i]);
  cudaFree(c->acc1); cudaFree(c->hid);
  for (int i = 0; i < 2; ++i) {
    cudaFree(c->stage_in[i]); cudaFree(c->stage_out[i]);
    if (c->ev_in[i]) cudaEventDestroy(c->ev_in[i]);
    if (c->ev_comp[i]) cudaEventDestroy(c->ev_comp[i]);
    if (c->ev_out[i]) cudaEventDestroy(c->ev_out[i]);
  }
  cudaFree(c->probs_scratch);
  cudaFree(c->alleles_stage);
  cudaFree(c->act6b); cudaFree(c->fc_in); cudaFree(c->act4h); cudaFree(c->act4b);
  cudaFree(c->patch6a); cudaFree(c->patch6b); cudaFree(c->acc_ref); cudaFree(c->probs_dense);
  cudaFree(c->si_of_k_dev); cudaFree(c->ref_scratch);
  bx_chromatin_close(c);
  if (c->s_comp) cudaStreamDestroy(c->s_comp);
  if (c->s_in) cudaStreamDestroy(c->s_in);
  if (c->s_out) cudaStreamDestroy(c->s_out);
  for (auto& r : c->prof_recs) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  for (auto e : c->prof_pool) cudaEventDestroy(e);
  delete c;
  return 0;
}

}  // extern "C"

template <typename T>
static int load_weights_t(bx_ctx* c, const float* const w[8], const float* const b[8]) {
  cudaStream_t st = 0;
  float* tmp = nullptr;
  const size_t fc1_elems = (size_t)2003 * 67840;
  CK(cudaMalloc(&tmp, fc1_elems * sizeof(float)));
  auto cleanup = [&]() { cudaFree(tmp); };
  for (int i = 0; i < 6; ++i) {
    const LayerDef& L = kLayers[i];
    const int cin_real = (i == 0) ? 4 : L.cin;
    const size_t src_elems = (size_t)L.cout * cin_real * 8;
    const size_t dst_elems = (size_t)L.taps * L.cout * L.cin;
    const int split = c->split_in[i] ? 1 : 0;
    if (!c->wconv[i]) CK(cudaMalloc(&c->wconv[i], dst_elems * 2 * ((split && i > 0) ? 2 : 1)));
    if (!c->bconv[i]) CK(cudaMalloc(&c->bconv[i], L.cout * sizeof(float)));
    CK(cudaMemcpy(tmp, w[i], src_elems * sizeof(float), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->bconv[i], b[i], L.cout * sizeof(float), cudaMemcpyHostToDevice));
    if (i == 0)
      prep_conv1_kernel<T><<<(L.cout * 64 + 255) / 256, 256, 0, st>>>(tmp, (T*)c->wconv[i], L.cout, split);
    else
      prep_conv_kernel<T><<<(unsigned)((dst_elems + 255) / 256), 256, 0, st>>>(tmp, (T*)c->wconv[i], L.cout, L.cin, split);
    g_launches++;
    CK(cudaStreamSynchronize(st));
  }
  if (!c->wfc1) CK(cudaMalloc(&c->wfc1, (size_t)2048 * 67840 * 2 * (c->split_in[6] ? 2 : 1)));
  if (!c->bfc1) CK(cudaMalloc(&c->bfc1, 2048 * sizeof(float)));
  if (!c->wfc2) CK(cudaMalloc(&c->wfc2, (size_t)2048 * 2048 * 2 * (c->split_in[7] ? 2 : 1)));
  if (!c->bfc2) CK(cudaMalloc(&c->bfc2, 2048 * sizeof(float)));
  CK(cudaMemcpy(tmp, w[6], fc1_elems * sizeof(float), cudaMemcpyHostToDevice));
  prep_fc1_kernel<T><<<(unsigned)(((size_t)2048 * 67840 + 255) / 256), 256, 0, st>>>(tmp, (T*)c->wfc1,
                                                                                     c->split_in[6] ? 1 : 0);
  g_launches++;
  CK(cudaStreamSynchronize(st));
  CK(cudaMemcpy(tmp, w[7], (size_t)2002 * 2003 * sizeof(float), cudaMemcpyHostToDevice));
  prep_fc2_kernel<T><<<(2048 * 2048 + 255) / 256, 256, 0, st>>>(tmp, (T*)c->wfc2, c->split_in[7] ? 1 : 0);
  g_launches++;
  CK(cudaMemset(c->bfc1, 0, 2048 * sizeof(float)));
  CK(cudaMemset(c->bfc2, 0, 2048 * sizeof(float)));
  CK(cudaMemcpy(c->bfc1, b[6], 2003 * sizeof(float), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(c->bfc2, b[7], 2002 * sizeof(float), cudaMemcpyHostToDevice));
  CK(cudaStreamSynchronize(st));
  CK(cudaGetLastError());
  cleanup();
  c->weights_ready = true;
  return 0;
}

extern "C" {

int bx_ctx_load_weights(bx_ctx* c, const float* const w[8], const float* const b[8]) {
  if (!c || !w || !b) return fail(BX_E_ARG, "null argument");
  for (int i = 0; i < 8; ++i)
    if (!w[i] || !b[i]) return fail(BX_E_ARG, "weight/bias %d is null", i);
  ON_DEVICE(c->device);
  return c->dtype == BX_DTYPE_F16 ? load_weights_t<__half>(c, w, b) : load_weights_t<__nv_bfloat16>(c, w, b);
}

int bx_ctx_set_genome(bx_ctx* c, const uint32_t* bases2_dev, const uint32_t* nmask_dev, int64_t n_bases) {
  if (!c || !bases2_dev || !nmask_dev || n_bases <= 0) return fail(BX_E_ARG, "invalid genome arguments");
  c->bases2 = bases2_dev;
  c->nmask = nmask_dev;
  c->n_bases = n_bases;
  return 0;
}

int bx_encode_windows(bx_ctx* c, const bx_window_desc* descs_dev, const uint8_t* alleles_dev, int n,
                      uint8_t* out_dev, void* stream) {
  if (!c || n < 0) return fail(BX_E_ARG, "invalid arguments");
  if (n == 0) return 0;  // empty input: nothing to do, buffers may be null
  if (!descs_dev || !alleles_dev || !out_dev) return fail(BX_E_ARG, "invalid arguments");
  if (!c->bases2) return fail(BX_E_STATE, "genome not set (bx_ctx_set_genome)");
  ON_DEVICE(c->device);
  encode_onehot_kernel<<<n, 256, 0, (cudaStream_t)stream>>>(descs_dev, alleles_dev, c->bases2, c->nmask, c->n_bases,
                                                             out_dev);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

}  // extern "C"

// One chunk (m <= cap windows): kernel 1 fused into the conv1 operand, then the network.
static int forward_chunk_windows(bx_ctx* c, const bx_window_desc* descs_dev, const uint8_t* alleles_dev, int m,
                                 float* probs_dev, cudaStream_t st, const ChromEpi* chrom = nullptr) {
  if (conv1_gen_enabled(c)) {
    EncParams e = {};
    e.descs = descs_dev; e.alleles = alleles_dev; e.bases2 = c->bases2; e.nmask = c->nmask; e.n_bases = c->n_bases;
    e.L = BX_WINDOW; e.n_seg = m;
    return run_network(c, m, probs_dev, st, &e, 1, chrom);
  }
  {
    StageTimer t(c, 0, st);
    if (c->dtype == BX_DTYPE_F16)
      encode_im2col_kernel<__half><<<m, 256, 0, st>>>(descs_dev, alleles_dev, c->bases2, c->nmask, c->n_bases,
                                                      (__half*)c->act[0]);
    else
      encode_im2col_kernel<__nv_bfloat16><<<m, 256, 0, st>>>(descs_dev, alleles_dev, c->bases2, c->nmask, c->n_bases,
                                                             (__nv_bfloat16*)c->act[0]);
    g_launches++;
    CK(cudaGetLastError());
  }
  return run_network(c, m, probs_dev, st, nullptr, 0, chrom);
}

static int forward_chunk_onehot(bx_ctx* c, const uint8_t* x, int m, float* probs_dev, cudaStream_t st) {
  if (conv1_gen_enabled(c)) {
    EncParams e = {};
    e.onehot = x; e.L = BX_WINDOW; e.n_seg = m;
    return run_network(c, m, probs_dev, st, &e, 2);
  }
  {
    StageTimer t(c, 0, st);
    if (c->dtype == BX_DTYPE_F16)
      onehot_im2col_kernel<__half><<<m, 256, 0, st>>>(x, (__half*)c->act[0]);
    else
      onehot_im2col_kernel<__nv_bfloat16><<<m, 256, 0, st>>>(x, (__nv_bfloat16*)c->act[0]);
    g_launches++;
    CK(cudaGetLastError());
  }
  return run_network(c, m, probs_dev, st);
}

// Lazily created streams / events / staging buffers of the host-buffer entry points.
static int ensure_pipe(bx_ctx* c, size_t in_bytes, size_t out_bytes, size_t probs_bytes) {
  if (!c->s_comp) {
    CK(cudaStreamCreateWithFlags(&c->s_comp, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&c->s_in, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&c->s_out, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      CK(cudaEventCreateWithFlags(&c->ev_in[i], cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&c->ev_comp[i], cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&c->ev_out[i], cudaEventDisableTiming));
    }
  }
  if (in_bytes > c->stage_in_bytes) {
    for (int i = 0; i < 2; ++i) { cudaFree(c->stage_in[i]); c->stage_in[i] = nullptr; CK(cudaMalloc(&c->stage_in[i], in_bytes)); }
    c->stage_in_bytes = in_bytes;
  }
  if (out_bytes > c->stage_out_bytes) {
    for (int i = 0; i < 2; ++i) { cudaFree(c->stage_out[i]); c->stage_out[i] = nullptr; CK(cudaMalloc(&c->stage_out[i], out_bytes)); }
    c->stage_out_bytes = out_bytes;
  }
  if (probs_bytes > c->probs_scratch_bytes) {
    cudaFree(c->probs_scratch); c->probs_scratch = nullptr;
    CK(cudaMalloc(&c->probs_scratch, probs_bytes));
    c->probs_scratch_bytes = probs_bytes;
  }
  return 0;
}

extern "C" {

int bx_forward_windows(bx_ctx* c, const bx_window_desc* descs_dev, const uint8_t* alleles_dev, int n,
                       float* probs_dev, void* stream) {
  if (!c || n < 0) return fail(BX_E_ARG, "invalid arguments");
  if (n == 0) return 0;
  if (!descs_dev || !alleles_dev || !probs_dev) return fail(BX_E_ARG, "invalid arguments");
  if (!c->bases2) return fail(BX_E_STATE, "genome not set (bx_ctx_set_genome)");
  if (!c->weights_ready) return fail(BX_E_STATE, "weights not loaded (bx_ctx_load_weights)");
  ON_DEVICE(c->device);
  cudaStream_t st = (cudaStream_t)stream;
  for (int off = 0; off < n; off += c->cap) {
    const int m = std::min(c->cap, n - off);
    if (int r = forward_chunk_windows(c, descs_dev + off, alleles_dev, m, probs_dev + (size_t)off * BX_FEATURES, st))
      return r;
  }
  return 0;
}

int bx_forward_onehot(bx_ctx* c, const uint8_t* onehot_dev, int n, float* probs_dev, void* stream) {
  if (!c || n < 0) return fail(BX_E_ARG, "invalid arguments");
  if (n == 0) return 0;
  if (!onehot_dev || !probs_dev) return fail(BX_E_ARG, "invalid arguments");
  if (!c->weights_ready) return fail(BX_E_STATE, "weights not loaded (bx_ctx_load_weights)");
  ON_DEVICE(c->device);
  cudaStream_t st = (cudaStream_t)stream;
  for (int off = 0; off < n; off += c->cap) {
    const int m = std::min(c->cap, n - off);
    if (int r = forward_chunk_onehot(c, onehot_dev + (size_t)off * 4 * BX_WINDOW, m,
                                     probs_dev + (size_t)off * BX_FEATURES, st))
      return r;
  }
  return 0;
}

int bx_forward_onehot_host(bx_ctx* c, const uint8_t* onehot_host, int n, float* probs_host) {
  if (!c || n < 0) return fail(BX_E_ARG, "invalid arguments");
  if (n == 0) return 0;
  if (!onehot_host || !probs_host) return fail(BX_E_ARG, "invalid arguments");
  if (!c->weights_ready) return fail(BX_E_STATE, "weights not loaded (bx_ctx_load_weights)");
  ON_DEVICE(c->device);
  const size_t in_chunk = (size_t)c->cap * 4 * BX_WINDOW, out_chunk = (size_t)c->cap * BX_FEATURES * sizeof(float);
  if (int r = ensure_pipe(c, in_chunk, out_chunk, 0)) return r;
  int i = 0;
  for (int off = 0; off < n; off += c->cap, ++i) {
    const int m = std::min(c->cap, n - off), b = i & 1;
    // H2D of chunk i overlaps the compute of chunk i-1; its staging buffer was last read by chunk i-2.
    CK(cudaStreamWaitEvent(c->s_in, c->ev_comp[b], 0));
    CK(cudaMemcpyAsync(c->stage_in[b], onehot_host + (size_t)off * 4 * BX_WINDOW, (size_t)m * 4 * BX_WINDOW,
                       cudaMemcpyHostToDevice, c->s_in));
    CK(cudaEventRecord(c->ev_in[b], c->s_in));
    CK(cudaStreamWaitEvent(c->s_comp, c->ev_in[b], 0));
    CK(cudaStreamWaitEvent(c->s_comp, c->ev_out[b], 0));
    if (int r = forward_chunk_onehot(c, (const uint8_t*)c->stage_in[b], m, c->stage_out[b], c->s_comp)) return r;
    CK(cudaEventRecord(c->ev_comp[b], c->s_comp));
    CK(cudaStreamWaitEvent(c->s_out, c->ev_comp[b], 0));
    CK(cudaMemcpyAsync(probs_host + (size_t)off * BX_FEATURES, c->stage_out[b], (size_t)m * BX_FEATURES * sizeof(float),
                       cudaMemcpyDeviceToHost, c->s_out));
    CK(cudaEventRecord(c->ev_out[b], c->s_out));
  }
  CK(cudaStreamSynchronize(c->s_out));
  CK(cudaStreamSynchronize(c->s_comp));
  return 0;
}

int bx_chromatin_shift_host(bx_ctx* c, const bx_window_desc* descs_host, const uint8_t* alleles_host,
                            int64_t n_alleles, int n_var, int average, float* ref_host, float* alt_host,
                            float* diff_host) {
  if (!c || n_var < 0 || n_alleles < 0) return fail(BX_E_ARG, "invalid arguments");
  if (n_var == 0) return 0;
  if (!descs_host || !ref_host || !alt_host || !diff_host || (n_alleles > 0 && !alleles_host))
    return fail(BX_E_ARG, "invalid arguments");
  if (!c->bases2) return fail(BX_E_STATE, "genome not set (bx_ctx_set_genome)");
  if (!c->weights_ready) return fail(BX_E_STATE, "weights not loaded (bx_ctx_load_weights)");
  if (c->cap < 4) return fail(BX_E_ARG, "context capacity must be >= 4 windows");
  ON_DEVICE(c->device);
  const int vcap = c->cap / 4;  // variants per chunk (4 windows each)
  const size_t rows_chunk = (size_t)(average ? 1 : 2) * vcap;
  const size_t out_chunk = 3 * rows_chunk * BX_FEATURES * sizeof(float);
  const size_t in_chunk = (size_t)4 * vcap * sizeof(bx_window_desc);
  if (int r = ensure_pipe(c, in_chunk, out_chunk, (size_t)4 * vcap * BX_FEATURES * sizeof(float))) return r;
  // allele codes: one upload for the whole call (descriptor aoff values index this table)
  if ((size_t)std::max<int64_t>(n_alleles, 1) > c->alleles_stage_bytes) {  // first call / a longer table only
    CK(cudaStreamSynchronize(c->s_comp));
    cudaFree(c->alleles_stage);
    c->alleles_stage = nullptr; c->alleles_stage_bytes = 0;
    const size_t want = std::max<size_t>((size_t)n_alleles * 2, 1 << 16);
    CK(cudaMalloc(&c->alleles_stage, want));
    c->alleles_stage_bytes = want;
  }
  uint8_t* alleles_dev = c->alleles_stage;
  // the previous call's kernels have finished (every call ends with a synchronise), so the table can be overwritten;
  // the first chunk's compute waits for this copy through ev_in (same stream, recorded after it)
  if (n_alleles > 0) CK(cudaMemcpyAsync(alleles_dev, alleles_host, (size_t)n_alleles, cudaMemcpyHostToDevice, c->s_in));
  const size_t F = BX_FEATURES;
  int i = 0, rc = 0;
  for (int v0 = 0; v0 < n_var && rc == 0; v0 += vcap, ++i) {
    const int m = std::min(vcap, n_var - v0), b = i & 1;
    const size_t rows = (size_t)(average ? 1 : 2) * m;
    float* o_ref = c->stage_out[b];
    float* o_alt = o_ref + rows * F;
    float* o_diff = o_alt + rows * F;
    auto step = [&]() -> int {
      CK(cudaStreamWaitEvent(c->s_in, c->ev_comp[b], 0));
      CK(cudaMemcpyAsync(c->stage_in[b], descs_host + (size_t)4 * v0, (size_t)4 * m * sizeof(bx_window_desc),
                         cudaMemcpyHostToDevice, c->s_in));
      CK(cudaEventRecord(c->ev_in[b], c->s_in));
      CK(cudaStreamWaitEvent(c->s_comp, c->ev_in[b], 0));
      CK(cudaStreamWaitEvent(c->s_comp, c->ev_out[b], 0));
      if (int r = forward_chunk_windows(c, (const bx_window_desc*)c->stage_in[b], alleles_dev, 4 * m, c->probs_scratch,
                                        c->s_comp))
        return r;
      {
        StageTimer t(c, 9, c->s_comp);
        if (average)
          chromatin_outputs_avg_kernel<<<m, 256, 0, c->s_comp>>>(c->probs_scratch, m, o_ref, o_alt, o_diff);
        else
          chromatin_outputs_kernel<<<m, 256, 0, c->s_comp>>>(c->probs_scratch, m, o_ref, o_alt, o_diff);
        g_launches++;
        CK(cudaGetLastError());
      }
      CK(cudaEventRecord(c->ev_comp[b], c->s_comp));
      CK(cudaStreamWaitEvent(c->s_out, c->ev_comp[b], 0));
      float* dst[3] = {ref_host, alt_host, diff_host};
      const float* src[3] = {o_ref, o_alt, o_diff};
      for (int k = 0; k < 3; ++k) {
        // forward-strand rows land at v, reverse-complement rows at n_var + v (chromatin.py:170-171 row order)
        CK(cudaMemcpyAsync(dst[k] + (size_t)v0 * F, src[k], (size_t)m * F * sizeof(float), cudaMemcpyDeviceToHost,
                           c->s_out));
        if (!average)
          CK(cudaMemcpyAsync(dst[k] + ((size_t)n_var + v0) * F, src[k] + (size_t)m * F, (size_t)m * F * sizeof(float),
                             cudaMemcpyDeviceToHost, c->s_out));
      }
      CK(cudaEventRecord(c->ev_out[b], c->s_out));
      return 0;
    };
    rc = step();
  }
  cudaStreamSynchronize(c->s_out);
  cudaError_t e = cudaStreamSynchronize(c->s_comp);
  if (rc) return rc;
  if (e != cudaSuccess) return fail(BX_E_CUDA, "pipeline failed: %s", cudaGetErrorString(e));
  return 0;
}

int bx_ctx_profile(bx_ctx* c, int enable) {
  if (!c) return fail(BX_E_ARG, "ctx is null");
  c->prof_on = enable != 0;
  return 0;
}

int bx_ctx_profile_read(bx_ctx* c, double* ms_out, int64_t* count_out, int reset) {
  if (!c || !ms_out || !count_out) return fail(BX_E_ARG, "null argument");
  ON_DEVICE(c->device);
  for (auto& r : c->prof_recs) {
    CK(cudaEventSynchronize(r.e1));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, r.e0, r.e1));
    c->prof_ms[r.stage] += ms;
    c->prof_cnt[r.stage] += 1;
    c->prof_pool.push_back(r.e0);
    c->prof_pool.push_back(r.e1);
  }
  c->prof_recs.clear();
  for (int i = 0; i < BX_NSTAGES; ++i) { ms_out[i] = c->prof_ms[i]; count_out[i] = c->prof_cnt[i]; }
  if (reset)
    for (int i = 0; i < BX_NSTAGES; ++i) { c->prof_ms[i] = 0; c->prof_cnt[i] = 0; }
  return 0;
}

int bx_debug_activation(bx_ctx* c, int layer, int n, void* dst_dev, int64_t* count, void* stream) {
  if (!c || layer < 0 || layer > 7 || n < 1 || n > c->cap || !dst_dev) return fail(BX_E_ARG, "invalid arguments");
  if (c->split_any) return fail(BX_E_STATE, "activation dumps are not available in split-operand mode");
  const void* src = layer == 7 ? c->hid : c->act[layer];
  const size_t elems = layer == 7 ? (size_t)n * 2048 : (size_t)n * act_rows(layer) * act_cols(layer);
  CK(cudaMemcpyAsync(dst_dev, src, elems * 2, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  if (count) *count = (int64_t)elems;
  return 0;
}

int bx_debug_run_layer(bx_ctx* c, int layer, int n, float* probs_dev, void* stream) {
  if (!c || layer < 0 || layer > 7 || n < 1 || n > c->cap) return fail(BX_E_ARG, "invalid arguments");
  if (!c->weights_ready) return fail(BX_E_STATE, "weights not loaded");
  ON_DEVICE(c->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (layer < 6) return run_conv(c, layer, n, st);
  if (layer == 6) return run_fc1(c, n, st);
  if (!probs_dev) return fail(BX_E_ARG, "probs_dev is null");
  return run_fc2(c, n, probs_dev, st);
}

int bx_chromatin_outputs(const float* probs_dev, int n_var, float* ref_dev, float* alt_dev, float* diff_dev,
                         void* stream) {
  if (n_var < 0) return fail(BX_E_ARG, "invalid arguments");
  if (n_var == 0) return 0;
  if (!probs_dev || !ref_dev || !alt_dev || !diff_dev) return fail(BX_E_ARG, "invalid arguments");
  chromatin_outputs_kernel<<<n_var, 256, 0, (cudaStream_t)stream>>>(probs_dev, n_var, ref_dev, alt_dev, diff_dev);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

int bx_chromatin_outputs_avg(const float* probs_dev, int n_var, float* ref_dev, float* alt_dev, float* diff_dev,
                             void* stream) {
  if (n_var < 0) return fail(BX_E_ARG, "invalid arguments");
  if (n_var == 0) return 0;
  if (!probs_dev || !ref_dev || !alt_dev || !diff_dev) return fail(BX_E_ARG, "invalid arguments");
  chromatin_outputs_avg_kernel<<<n_var, 256, 0, (cudaStream_t)stream>>>(probs_dev, n_var, ref_dev, alt_dev,
                                                                       diff_dev);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

int bx_strand_average(const float* x_dev, int n, float* out_dev, void* stream) {
  if (n < 0) return fail(BX_E_ARG, "invalid arguments");
  if (n == 0) return 0;
  if (!x_dev || !out_dev) return fail(BX_E_ARG, "invalid arguments");
  const size_t total2 = (size_t)n * (kFeat / 2);
  const unsigned blocks = (unsigned)std::min<size_t>((total2 + 255) / 256, 148 * 16);
  strand_average_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(x_dev, n, out_dev);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

int bx_decay_transform_fixed(const float* probs_dev, const float* weights_dev, int n_genes, int n_shifts,
                             float* out_dev, void* stream) {
  if (n_genes < 0 || n_shifts < 1 || n_shifts > 1024) return fail(BX_E_ARG, "invalid arguments");
  if (n_genes == 0) return 0;
  if (!probs_dev || !weights_dev || !out_dev) return fail(BX_E_ARG, "invalid arguments");
  const size_t smem = (size_t)(10 * n_shifts + (kTfSlices - 1) * 20 * kTfF) * sizeof(float);
  dim3 grid((kFeat2 + kTfF - 1) / kTfF, n_genes), block(kTfF, kTfSlices);
  int dev = 0;
  CK(cudaGetDevice(&dev));
  static std::atomic<bool> attr_set[kMaxDevices];  // per device, like the GEMM launcher's
  if (dev < 0 || dev >= kMaxDevices) return fail(BX_E_ARG, "device index %d out of range", dev);
  if (!attr_set[dev].load(std::memory_order_acquire)) {
    CK(cudaFuncSetAttribute(decay_fixed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    attr_set[dev].store(true, std::memory_order_release);
  }
  decay_fixed_kernel<<<grid, block, smem, (cudaStream_t)stream>>>(probs_dev, weights_dev, n_shifts, out_dev);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

int bx_decay_transform_variant(const float* x_dev, const double* dist_dev, const int8_t* strand_dev,
                               const int32_t* shifts_dev, int n, int n_shifts, float* out_dev, void* stream) {
  if (n < 0 || n_shifts < 1 || n_shifts > 25) return fail(BX_E_ARG, "invalid arguments (n_shifts must be 1..25)");
  if (n == 0) return 0;
  if (!x_dev || !dist_dev || !strand_dev || !shifts_dev || !out_dev) return fail(BX_E_ARG, "invalid arguments");
  if (n_shifts == 9)  // chromatin.py's 9 shifts: all loads of a column unrolled and in flight together
    decay_variant_kernel<9><<<n, 256, 0, (cudaStream_t)stream>>>(x_dev, dist_dev, strand_dev, shifts_dev, n, n_shifts,
                                                                 out_dev);
  else
    decay_variant_kernel<0><<<n, 256, 0, (cudaStream_t)stream>>>(x_dev, dist_dev, strand_dev, shifts_dev, n, n_shifts,
                                                                 out_dev);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

int bx_decay_score_variant(const float* x_dev, const double* dist_dev, const int8_t* strand_dev,
                            const int32_t* shifts_dev, const float* w_dev, const float* offset_dev, int n,
                            int n_shifts, int n_models, float* out_dev, void* stream) {
  if (n < 0 || n_models < 1 || n_shifts < 1 || n_shifts > 25)
    return fail(BX_E_ARG, "invalid arguments (n_shifts must be 1..25, n_models >= 1)");
  if (n == 0) return 0;
  if (!x_dev || !dist_dev || !strand_dev || !shifts_dev || !w_dev || !offset_dev || !out_dev)
    return fail(BX_E_ARG, "invalid arguments");
  dim3 grid(n, (n_models + kScoreModels - 1) / kScoreModels);
  const size_t smem = (size_t)(n_shifts * 10 + 8 * kScoreModels) * sizeof(float);
  decay_score_variant_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(x_dev, dist_dev, strand_dev, shifts_dev, w_dev,
                                                                        offset_dev, n, n_shifts, n_models, out_dev);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

// ---- transform + scoring as one tcgen05 GEMM (score_tc.cuh)
static int check_device();
static int score_group_pad(int m) { return (std::min(m, 256) + 15) / 16 * 16; }

int64_t bx_score_weights_bytes(int n_models) {
  if (n_models < 1) return 0;
  int64_t rows = 0;
  for (int m0 = 0; m0 < n_models; m0 += 256) rows += score_group_pad(n_models - m0);
  return rows * 2 * (int64_t)kScK * (int64_t)sizeof(__half);
}

int bx_score_prepare_weights(const float* w_dev, int n_models, void* wp_dev, void* stream) {
  if (n_models < 1 || !w_dev || !wp_dev) return fail(BX_E_ARG, "invalid arguments");
  if (int r = check_device()) return r;
  __half* out = reinterpret_cast<__half*>(wp_dev);
  for (int m0 = 0; m0 < n_models; m0 += 256) {
    const int m = std::min(256, n_models - m0), n_pad = score_group_pad(m);
    const size_t total = (size_t)n_pad * kScK;
    score_prep_weights_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        w_dev + (size_t)m0 * 20020, m, n_pad, out);
    g_launches++;
    CK(cudaGetLastError());
    out += (size_t)n_pad * 2 * kScK;
  }
  return 0;
}

int bx_decay_score_variant_tc(const float* x_dev, int64_t x_row_pitch, const double* dist_dev, const int8_t* strand_dev,
                              const int32_t* shifts_dev, const void* wp_dev, const float* offset_dev, int n,
                              int n_shifts, int n_models, float* out_dev, void* stream) {
  if (n < 0 || n_models < 1) return fail(BX_E_ARG, "invalid arguments");
  if (x_row_pitch < BX_FEATURES || (x_row_pitch & 1))
    return fail(BX_E_ARG, "x_row_pitch must be an even number of floats >= %d", BX_FEATURES);
  if (n_shifts != kScShifts)
    return fail(BX_E_ARG, "the tensor-core scoring kernel is built for %d shifts (--maxshift 800); got %d", kScShifts, n_shifts);
  if (n == 0) return 0;
  if (!x_dev || !dist_dev || !strand_dev || !shifts_dev || !wp_dev || !offset_dev || !out_dev)
    return fail(BX_E_ARG, "invalid arguments");
  if (int r = check_device()) return r;
  if (int r = load_driver_entry()) return r;
  int dev = 0;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(BX_E_ARG, "device index %d out of range", dev);
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const __half* wp = reinterpret_cast<const __half*>(wp_dev);
  // Rows on 16-byte boundaries: x is staged by TMA (score_tma_kernel).  Otherwise (e.g. dense 8008-byte rows) the
  // kernel that loads x through registers (score_tc_kernel); same operand layout, same results.
#ifndef BX_SC_L2PROMO
#define BX_SC_L2PROMO CU_TENSOR_MAP_L2_PROMOTION_L2_128B
#endif
  // measurement switches (tools/score_probe.py), read once: force the register-load kernel, cap the weight ring / grid
  static const bool no_tma = getenv("BX_SCORE_NO_TMA") != nullptr;
  static const int sb_cap = getenv("BX_SCORE_SB") ? atoi(getenv("BX_SCORE_SB")) : 0;
  static const int cluster_cap = getenv("BX_SCORE_CLUSTERS") ? atoi(getenv("BX_SCORE_CLUSTERS")) : 0;
  const bool x_tma = (x_row_pitch % 4 == 0) && (reinterpret_cast<uintptr_t>(x_dev) % 16 == 0) && !no_tma;
  CUtensorMap tmX;
  if (x_tma) {
    cuuint64_t dims[3] = {(cuuint64_t)BX_FEATURES, (cuuint64_t)n, (cuuint64_t)kScShifts};
    cuuint64_t strides[2] = {(cuuint64_t)x_row_pitch * 4, (cuuint64_t)n * (cuuint64_t)x_row_pitch * 4};
    cuuint32_t box[3] = {8, 128, (cuuint32_t)kScShifts};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = g_encode(&tmX, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(x_dev), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, BX_SC_L2PROMO,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
      return fail(BX_E_CUDA, "cuTensorMapEncodeTiled failed for x: %d (n=%d pitch=%lld)", (int)r, n, (long long)x_row_pitch);
  }
  // side-pure tiles skip the K steps that are zero for all their rows (score_tile_class); BX_SCORE_NO_SKIP for A/B runs
  static const bool no_skip = getenv("BX_SCORE_NO_SKIP") != nullptr;
  for (int m0 = 0; m0 < n_models; m0 += 256) {
    const int m = std::min(256, n_models - m0), n_pad = score_group_pad(m);
    ScoreParams p = {};
    p.skip_zero_steps = no_skip ? 0 : 1;
    p.x = x_dev; p.ld_x = x_row_pitch; p.dist = dist_dev; p.strand = strand_dev; p.shifts = shifts_dev;
    p.offset = offset_dev + m0;
    p.out = out_dev + m0;
    p.ld_out = n_models;
    p.n = n; p.m = m; p.n_pad = n_pad;
    p.tiles = (n + 255) / 256;
    p.idesc = make_idesc_f16(0, 256, n_pad);
    const size_t b_stage = (size_t)2 * (n_pad / 2) * 128;
    size_t smem;
    if (x_tma) {   // x stages + A ring (2) + K stage, the rest of the 227 KB for the weight ring
      const size_t fixed = (size_t)kScXStages * kScXStage + (size_t)3 * 2 * kScATile + 256 + 1024;
      p.sb_stages = (int)std::min<size_t>(4, (232448 - fixed) / b_stage);
      if (sb_cap > 0) p.sb_stages = std::min(p.sb_stages, sb_cap);
      smem = fixed + p.sb_stages * b_stage;
    } else {
      p.sb_stages = 3;
      smem = (size_t)(2 + 2) * 2 * kScATile + (size_t)p.sb_stages * b_stage + 256 + 1024;
    }
    CUtensorMap tmW;
    if (int r = make_tmap(&tmW, BX_DTYPE_F16, wp, (uint64_t)n_pad, (uint64_t)2 * kScK, (uint64_t)2 * kScK, (uint32_t)(n_pad / 2)))
      return r;
    static std::atomic<size_t> attr_smem[2][kMaxDevices];
    if (attr_smem[x_tma][dev].load(std::memory_order_acquire) < smem) {
      if (x_tma) CK(cudaFuncSetAttribute(score_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      else CK(cudaFuncSetAttribute(score_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr_smem[x_tma][dev].store(smem, std::memory_order_release);
    }
    int clusters = std::max(1, std::min(p.tiles, sms / 2));
    if (cluster_cap > 0) clusters = std::min(clusters, cluster_cap);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(clusters * 2);
    cfg.blockDim = dim3(x_tma ? kScTmaThreads : kScThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (x_tma) CK(cudaLaunchKernelEx(&cfg, score_tma_kernel, tmW, tmX, p));
    else CK(cudaLaunchKernelEx(&cfg, score_tc_kernel, tmW, p));
    g_launches++;
    wp += (size_t)n_pad * 2 * kScK;
  }
  return 0;
}

int bx_segment_capacity(bx_ctx* c, int seg_len) {
  if (!c || seg_len < 2000) return fail(BX_E_ARG, "invalid arguments");
  return max_segments(c, make_geom(seg_len));
}

}  // extern "C"

// conv1..conv6 over ns segments of L positions described by d: conv6 output of pooling phase A -> out6a and
// (two_phases) of phase B -> out6b, [ns, g.p6, 640] each.  Uses the context's activation scratch act[0..5].
static int run_segment_convs(bx_ctx* c, const bx_window_desc* d, const uint8_t* alleles_dev, int ns, int L,
                             const Geom& g, void* out6a, void* out6b, bool two_phases, cudaStream_t st) {
  const bool f16 = c->dtype == BX_DTYPE_F16;
  EncParams e = {};
  e.descs = d; e.alleles = alleles_dev; e.bases2 = c->bases2; e.nmask = c->nmask; e.n_bases = c->n_bases;
  e.L = L; e.n_seg = ns;
  const bool gen = conv1_gen_enabled(c);
  if (!gen) {
    StageTimer t(c, 0, st);
    dim3 grid((g.p1 * 8 + 255) / 256, ns);
    if (f16)
      encode_im2col_seg_kernel<__half><<<grid, 256, 0, st>>>(d, alleles_dev, c->bases2, c->nmask, c->n_bases,
                                                             (__half*)c->act[0], L, g.p1);
    else
      encode_im2col_seg_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(d, alleles_dev, c->bases2, c->nmask, c->n_bases,
                                                                    (__nv_bfloat16*)c->act[0], L, g.p1);
    g_launches++;
    CK(cudaGetLastError());
  }
  for (int i = 0; i < 3; ++i) {  // conv1..conv3
    StageTimer t(c, 1 + i, st);
    if (int r = run_conv_geom(c, i, ns, g, c->act[i], c->act[i + 1], false, st, gen ? &e : nullptr)) return r;
  }
  {
    // conv4 once, pooled over row pairs; both phases of the second MaxPool(1,4) are finished from that
    // (phase A: groups from row 0 -> act[4]; phase B: groups from row 2 -> act4b)
    StageTimer t(c, 4, st);
    if (int r = run_conv_geom(c, 3, ns, g, c->act[3], c->act4h, false, st, nullptr, true)) return r;
    const long long total = (long long)ns * g.p4 * (480 / 8);
    const unsigned blocks = (unsigned)((total + 255) / 256);
    if (f16 && c->split_in[4])
      pool_phases_kernel<__half, true><<<blocks, 256, 0, st>>>((const __half*)c->act4h, (__half*)c->act[4],
                                                               (__half*)c->act4b, g.ph, g.p4, g.l4b, 480 / 8, total);
    else if (f16)
      pool_phases_kernel<__half, false><<<blocks, 256, 0, st>>>((const __half*)c->act4h, (__half*)c->act[4],
                                                                (__half*)c->act4b, g.ph, g.p4, g.l4b, 480 / 8, total);
    else
      pool_phases_kernel<__nv_bfloat16, false><<<blocks, 256, 0, st>>>(
          (const __nv_bfloat16*)c->act4h, (__nv_bfloat16*)c->act[4], (__nv_bfloat16*)c->act4b, g.ph, g.p4, g.l4b,
          480 / 8, total);
    g_launches++;
    CK(cudaGetLastError());
  }
  for (int phase = 0; phase < 2; ++phase) {
    if (phase == 1 && !two_phases) break;
    { StageTimer t(c, 5, st);
      if (int r = run_conv_geom(c, 4, ns, g, phase ? c->act4b : c->act[4], c->act[5], false, st)) return r; }
    { StageTimer t(c, 6, st);
      if (int r = run_conv_geom(c, 5, ns, g, c->act[5], phase ? out6b : out6a, false, st)) return r; }
  }
  return 0;
}

// BX_DELTA_DENSE=0: alt windows go through FC1 in full (comparisons).
static bool delta_dense_enabled() {
  static const bool on = !(getenv("BX_DELTA_DENSE") && atoi(getenv("BX_DELTA_DENSE")) == 0);
  return on;
}

// Dense layers of one delta-mode chunk (kernels_misc.cuh, "delta mode, dense layers"): FC1 over all slices for the
// ref windows, over the slices that hold patched rows for the alt windows, then FC2 and the scatter to the final order.
static int delta_dense_phase(bx_ctx* c, const SegGather& sg, const Geom& gp, int ng, int n_shifts, int patch_row0,
                             int shift_major, float* out, cudaStream_t st, const ChromSink* sink = nullptr) {
  (void)gp;
  const bool f16 = c->dtype == BX_DTYPE_F16;
  DeltaDense dd = {};
  dd.ng = ng; dd.S = n_shifts; dd.shift_major = shift_major ? 1 : 0;
  dd.p6 = sg.p6; dd.p6p = sg.p6p; dd.row0 = patch_row0; dd.valid_a = sg.valid_a; dd.valid_b = sg.valid_b;
  dd.kb_per_split = kFc1KbPerSplit;
  dd.rw = sg.rw;
  for (int i = 0; i < n_shifts; ++i) { dd.k_of_shift[i] = sg.k_of_shift[i]; dd.si_of_k[i] = -1; }
  for (int i = 0; i < n_shifts; ++i) dd.si_of_k[sg.k_of_shift[i]] = i;
  for (int k = 0; k < n_shifts; ++k) {
    if (dd.si_of_k[k] < 0) return fail(BX_E_ARG, "k_of_shift is not a permutation (offset %d missing)", k);
  }
  const size_t n_half = (size_t)ng * n_shifts * 2;   // ref windows of the chunk = alt windows of the chunk
  const size_t ref_bytes = (size_t)kFc1Splits * n_half * 2048 * sizeof(float);
  if (ref_bytes > c->acc_ref_bytes) {
    CK(cudaStreamSynchronize(st));
    cudaFree(c->acc_ref);
    c->acc_ref = nullptr; c->acc_ref_bytes = 0;
    CK(cudaMalloc(&c->acc_ref, ref_bytes));
    c->acc_ref_bytes = ref_bytes;
  }
  if (!c->probs_dense) CK(cudaMalloc(&c->probs_dense, (size_t)c->cap * BX_FEATURES * sizeof(float)));
  const size_t ref_stride = n_half * 2048;
  void* saved = c->act[6];
  auto gather = [&](int alt, long long d0, int m) {
    if (f16)
      gather_fc1_delta_kernel<__half><<<m, 256, 0, st>>>((const __half*)saved, (const __half*)c->act6b,
                                                         (const __half*)c->patch6a, (const __half*)c->patch6b, dd, alt,
                                                         d0, (__half*)c->fc_in);
    else
      gather_fc1_delta_kernel<__nv_bfloat16><<<m, 256, 0, st>>>(
          (const __nv_bfloat16*)saved, (const __nv_bfloat16*)c->act6b, (const __nv_bfloat16*)c->patch6a,
          (const __nv_bfloat16*)c->patch6b, dd, alt, d0, (__nv_bfloat16*)c->fc_in);
    g_launches++;
  };
  if (sink && sink->average) {   // per-strand ref probabilities for the alt launches' alt - ref
    const size_t need = (size_t)n_shifts * 2 * ng * BX_FEATURES * sizeof(float);
    if (need > c->ref_scratch_bytes) {
      CK(cudaStreamSynchronize(st));
      cudaFree(c->ref_scratch);
      c->ref_scratch = nullptr; c->ref_scratch_bytes = 0;
      CK(cudaMalloc(&c->ref_scratch, need));
      c->ref_scratch_bytes = need;
    }
  }
  auto finish = [&](int alt, long long d0, int m) -> int {   // FC2 + scatter of a piece whose hidden layer is ready
    StageTimer t(c, 8, st);
    if (sink) {   // FC2's epilogue writes ref / alt / diff in their final layout: no dense copy, no scatter
      const ChromEpi ce = make_chrom(c, *sink, alt ? 2 : 1, ng, n_shifts, d0);
      return run_fc2(c, m, nullptr, st, &ce);
    }
    if (int r = run_fc2(c, m, c->probs_dense, st)) return r;
    scatter_delta_rows_kernel<<<m, 256, 0, st>>>(c->probs_dense, dd, alt, d0, out);
    g_launches++;
    CK(cudaGetLastError());
    return 0;
  };
  // ref windows: all slices, partials kept in acc_ref for the alt windows of the same chunk
  const size_t ref_piece = (size_t)(c->cap & ~3);   // pieces start at multiples of 4 rows (strand pairs / quads stay in a warp)
  for (size_t d0 = 0; d0 < n_half; d0 += ref_piece) {
    const int m = (int)std::min<size_t>(ref_piece, n_half - d0);
    {
      StageTimer t(c, 7, st);
      gather(0, (long long)d0, m);
      CK(cudaGetLastError());
      c->act[6] = c->fc_in;
      const int r = run_fc1(c, m, st, c->acc_ref + d0 * 2048, ref_stride);
      c->act[6] = saved;
      if (r) return r;
    }
    if (int r = finish(0, (long long)d0, m)) return r;
  }
  // alt windows: per M tile only the slices its rows' classes need (at most 32 tiles per launch: GemmParams)
  const int tile_rows = 128 * c->cg;
  const size_t piece = std::min<size_t>(ref_piece, (size_t)32 * tile_rows);
  for (size_t d0 = 0; d0 < n_half; d0 += piece) {
    const int m = (int)std::min<size_t>(piece, n_half - d0);
    GemmParams ranges = {};
    ranges.n_tile_ranges = (m + tile_rows - 1) / tile_rows;
    if (ranges.n_tile_ranges > 32) return fail(BX_E_STATE, "delta dense phase: %d M tiles per piece", ranges.n_tile_ranges);
    for (int t = 0; t < ranges.n_tile_ranges; ++t) {
      const size_t first = d0 + (size_t)t * tile_rows, last = std::min(d0 + m, first + tile_rows) - 1;
      int lo = kFc1Splits, hi = -1;
      for (int cls = (int)(first / (2 * (size_t)ng)); cls <= (int)(last / (2 * (size_t)ng)); ++cls) {
        const DeltaClass cl = delta_class(dd, cls);
        if (cl.lo >= cl.hi) continue;  // the patch does not reach this window: every slice comes from the ref window
        lo = std::min(lo, cl.ks_lo); hi = std::max(hi, cl.ks_hi);
      }
      ranges.tile_lo[t] = (unsigned char)(hi < lo ? 0 : lo);
      ranges.tile_n[t] = (unsigned char)(hi < lo ? 0 : hi - lo + 1);
    }
    {
      StageTimer t(c, 7, st);
      gather(1, (long long)d0, m);
      CK(cudaGetLastError());
      c->act[6] = c->fc_in;
      const size_t stride = (size_t)m * 2048;
      const int r = launch_fc1_gemm(c, m, c->acc1, stride, &ranges, st);
      c->act[6] = saved;
      if (r) return r;
      const size_t total = (size_t)m * 2048;
      const unsigned blocks = (unsigned)((total / 4 + 255) / 256);
      if (f16)
        fc1_finalize_delta_kernel<__half><<<blocks, 256, 0, st>>>(c->acc1, stride, c->acc_ref, ref_stride, dd,
                                                                  (long long)d0, kFc1Splits, c->bfc1, (__half*)c->hid,
                                                                  total, c->split_in[7] ? 1 : 0);
      else
        fc1_finalize_delta_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(c->acc1, stride, c->acc_ref, ref_stride, dd,
                                                                         (long long)d0, kFc1Splits, c->bfc1,
                                                                         (__nv_bfloat16*)c->hid, total, 0);
      g_launches++;
      CK(cudaGetLastError());
    }
    if (int r = finish(1, (long long)d0, m)) return r;
  }
  return 0;
}

// Segment mode proper.  patch_descs_dev == nullptr: descs_dev holds segs_out segments per group.  Delta mode
// (patch_descs_dev != nullptr, segs_out == 4): descs_dev holds the group's two REF segments (fwd, rc) and
// patch_descs_dev two short ALT patches (fwd, rc) of patch_len positions that start 16*patch_row0 positions into
// the alt segments; alt windows read the conv6 rows the substitution can reach from the patch and all others from
// the ref segment (identical there).
static int forward_segments_impl(bx_ctx* c, const bx_window_desc* descs_dev, const bx_window_desc* patch_descs_dev,
                                 const uint8_t* alleles_dev, int n_groups, int segs_out, int seg_len, int patch_len,
                                 int patch_row0, int n_shifts, const int32_t* k_of_shift, int shift_major,
                                 float* probs_dev, cudaStream_t st, const ChromSink* sink = nullptr) {
  const bool delta = patch_descs_dev != nullptr;
  if (c->cap < 4) return fail(BX_E_ARG, "context capacity must be >= 4 windows");
  const int segs_in = delta ? 2 : segs_out;
  const Geom g = make_geom(seg_len);
  const int seg_cap = max_segments(c, g);
  const int grp_cap = seg_cap / segs_in;
  if (grp_cap < 1)
    return fail(BX_E_ARG, "context capacity (%d windows) is too small for %d segments of %d positions", c->cap,
                segs_in, seg_len);
  if (!c->act6b) CK(cudaMalloc(&c->act6b, (size_t)c->cap * act_rows(6) * act_width(c, 6) * 2));
  if (!c->fc_in) CK(cudaMalloc(&c->fc_in, (size_t)c->cap * 106 * act_width(c, 6) * 2));
  // ph <= 2*p4 + 1 and n_seg <= cap, n_seg*p4 <= cap*120 (max_segments): n_seg*ph <= cap*241 rows
  if (!c->act4h) CK(cudaMalloc(&c->act4h, (size_t)c->cap * 241 * act_width(c, 4) * 2));
  if (!c->act4b) CK(cudaMalloc(&c->act4b, (size_t)c->cap * act_rows(4) * act_width(c, 4) * 2));
  SegGather sg = {};
  sg.segs_per_group = segs_out;
  sg.n_shifts = n_shifts;
  sg.shift_major = shift_major ? 1 : 0;
  sg.p6 = g.p6;
  sg.rw = (int)act_width(c, 6);
  Geom gp = {};
  if (delta) {
    gp = make_geom(patch_len);
    sg.delta = 1;
    sg.p6p = gp.p6;
    sg.row0 = patch_row0;
    sg.valid_a = gp.p6;             // patch conv6 rows computed from complete receptive fields, phase A / phase B
    sg.valid_b = gp.l4b - 14;
    if (sg.valid_b < 1 || gp.p6 < 1) return fail(BX_E_ARG, "patch_len %d is too short", patch_len);
    const size_t need = (size_t)std::min(grp_cap, n_groups) * 2 * gp.p6 * act_width(c, 6) * 2;
    if (need > c->patch6_bytes) {
      CK(cudaStreamSynchronize(st));
      cudaFree(c->patch6a); cudaFree(c->patch6b);
      c->patch6a = c->patch6b = nullptr;
      c->patch6_bytes = 0;
      CK(cudaMalloc(&c->patch6a, need));
      CK(cudaMalloc(&c->patch6b, need));
      c->patch6_bytes = need;
    }
  }
  for (int i = 0; i < n_shifts; ++i) {
    if (k_of_shift[i] < 0 || k_of_shift[i] >= n_shifts) return fail(BX_E_ARG, "k_of_shift[%d] out of range", i);
    sg.k_of_shift[i] = k_of_shift[i];
  }
  const bool f16 = c->dtype == BX_DTYPE_F16;
  // Groups are processed in chunks that fit the activation scratch; with shift-major output order a chunk's
  // windows are not contiguous in probs_dev, so chunks are only allowed there when one chunk covers everything.
  if (shift_major && n_groups > grp_cap)
    return fail(BX_E_ARG, "shift-major output needs all %d groups in one chunk (capacity %d): split the call", n_groups,
                grp_cap);
  for (int g0 = 0; g0 < n_groups; g0 += grp_cap) {
    const int ng = std::min(grp_cap, n_groups - g0);
    const bx_window_desc* d = descs_dev + (size_t)g0 * segs_in;
    if (int r = run_segment_convs(c, d, alleles_dev, ng * segs_in, seg_len, g, c->act[6], c->act6b, n_shifts > 1, st))
      return r;
    if (delta)
      if (int r = run_segment_convs(c, patch_descs_dev + (size_t)g0 * 2, alleles_dev, ng * 2, patch_len, gp,
                                    c->patch6a, c->patch6b, n_shifts > 1, st))
        return r;
    if (delta && delta_dense_enabled()) {
      if (int r = delta_dense_phase(c, sg, gp, ng, n_shifts, patch_row0, shift_major,
                                    sink ? nullptr : probs_dev + (size_t)g0 * segs_out * n_shifts * BX_FEATURES, st,
                                    sink))
        return r;
      continue;
    }
    if (sink && (delta || segs_out != 4 || !shift_major))
      return fail(BX_E_STATE, "the fused ref/alt/diff epilogue needs shift-major quads (or the delta dense phase)");
    // dense layers per window, in final output order, `cap` windows at a time
    sg.n_groups = ng;
    const long long n_win = (long long)ng * segs_out * n_shifts;
    float* out = sink ? nullptr : probs_dev + (size_t)g0 * segs_out * n_shifts * BX_FEATURES;  // group-major chunks are contiguous
    void* saved = c->act[6];
    const long long step = sink ? (long long)(c->cap & ~3) : (long long)c->cap;   // quads stay inside one launch
    for (long long w0 = 0; w0 < n_win; w0 += step) {
      const int m = (int)std::min<long long>(step, n_win - w0);
      {
        StageTimer t(c, 7, st);
        if (f16)
          gather_fc1_kernel<__half><<<m, 256, 0, st>>>((const __half*)saved, (const __half*)c->act6b,
                                                       (const __half*)c->patch6a, (const __half*)c->patch6b, d, sg, w0,
                                                       (__half*)c->fc_in);
        else
          gather_fc1_kernel<__nv_bfloat16><<<m, 256, 0, st>>>(
              (const __nv_bfloat16*)saved, (const __nv_bfloat16*)c->act6b, (const __nv_bfloat16*)c->patch6a,
              (const __nv_bfloat16*)c->patch6b, d, sg, w0, (__nv_bfloat16*)c->fc_in);
        g_launches++;
        CK(cudaGetLastError());
        c->act[6] = c->fc_in;  // run_fc1 reads its operand from act[6]
        const int r = run_fc1(c, m, st);
        c->act[6] = saved;
        if (r) return r;
      }
      StageTimer t(c, 8, st);
      if (sink) {
        const ChromEpi ce = make_chrom(c, *sink, 0, ng, n_shifts, w0);
        if (int r = run_fc2(c, m, nullptr, st, &ce)) return r;
      } else if (int r = run_fc2(c, m, out + (size_t)w0 * BX_FEATURES, st)) return r;
    }
  }
  return 0;
}

extern "C" {

int bx_forward_segments(bx_ctx* c, const bx_window_desc* descs_dev, const uint8_t* alleles_dev, int n_groups,
                        int segs_per_group, int seg_len, int n_shifts, const int32_t* k_of_shift, int shift_major,
                        float* probs_dev, void* stream) {
  if (!c || n_groups < 0 || !k_of_shift) return fail(BX_E_ARG, "invalid arguments");
  if (n_groups == 0) return 0;
  if (!descs_dev || !alleles_dev || !probs_dev) return fail(BX_E_ARG, "invalid arguments");
  if (segs_per_group < 1 || n_shifts < 1 || n_shifts > 200 || seg_len != 2000 + 200 * (n_shifts - 1))
    return fail(BX_E_ARG, "segment mode needs seg_len == 2000 + 200*(n_shifts-1) and 1 <= n_shifts <= 200");
  if (!c->bases2) return fail(BX_E_STATE, "genome not set (bx_ctx_set_genome)");
  if (!c->weights_ready) return fail(BX_E_STATE, "weights not loaded (bx_ctx_load_weights)");
  ON_DEVICE(c->device);
  return forward_segments_impl(c, descs_dev, nullptr, alleles_dev, n_groups, segs_per_group, seg_len, 0, 0, n_shifts,
                               k_of_shift, shift_major, probs_dev, (cudaStream_t)stream);
}

int bx_forward_segments_delta(bx_ctx* c, const bx_window_desc* ref_descs_dev, const bx_window_desc* patch_descs_dev,
                              const uint8_t* alleles_dev, int n_groups, int seg_len, int patch_len, int patch_row0,
                              int n_shifts, const int32_t* k_of_shift, int shift_major, float* probs_dev,
                              void* stream) {
  if (!c || n_groups < 0 || !k_of_shift) return fail(BX_E_ARG, "invalid arguments");
  if (n_groups == 0) return 0;
  if (!ref_descs_dev || !patch_descs_dev || !alleles_dev || !probs_dev) return fail(BX_E_ARG, "invalid arguments");
  if (n_shifts < 1 || n_shifts > 200 || seg_len != 2000 + 200 * (n_shifts - 1))
    return fail(BX_E_ARG, "segment mode needs seg_len == 2000 + 200*(n_shifts-1) and 1 <= n_shifts <= 200");
  if (patch_len < 336 || patch_len % 16 || patch_row0 < 0 || 16 * patch_row0 + patch_len > seg_len)
    return fail(BX_E_ARG, "patch must be a multiple of 16 positions (>= 336) inside the segment, starting at a multiple "
                          "of 16 (patch_len=%d patch_row0=%d seg_len=%d)", patch_len, patch_row0, seg_len);
  if (!c->bases2) return fail(BX_E_STATE, "genome not set (bx_ctx_set_genome)");
  if (!c->weights_ready) return fail(BX_E_STATE, "weights not loaded (bx_ctx_load_weights)");
  ON_DEVICE(c->device);
  return forward_segments_impl(c, ref_descs_dev, patch_descs_dev, alleles_dev, n_groups, 4, seg_len, patch_len,
                               patch_row0, n_shifts, k_of_shift, shift_major, probs_dev, (cudaStream_t)stream);
}

// ------------------------------------------------------------------ FASTA text in HBM -> packed genome
static int check_device() {
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(BX_E_CUDA, "no CUDA device (%s): this library has no CPU fallback", cudaGetErrorString(e));
  return 0;
}
static unsigned grid_for(long long n_items, int per_block) {
  return (unsigned)std::max<long long>(1, std::min<long long>((n_items + per_block - 1) / per_block, 148LL * 16));
}

int64_t bx_fasta_blocks(int64_t body_start, int64_t body_end) {
  if (body_start < 0 || body_end <= body_start) return 0;
  const long long a0 = body_start & ~15LL;
  return (body_end - a0 + kFaBlockBytes - 1) / kFaBlockBytes;
}

int bx_fasta_find_headers(const uint8_t* text_dev, int64_t n_text, int64_t* pos_dev, int64_t cap, int64_t* count_dev,
                          void* stream) {
  if (n_text < 0 || cap < 0 || !count_dev || (cap > 0 && !pos_dev)) return fail(BX_E_ARG, "invalid arguments");
  if (int r = check_device()) return r;
  cudaStream_t st = (cudaStream_t)stream;
  CK(cudaMemsetAsync(count_dev, 0, sizeof(int64_t), st));
  if (n_text == 0) return 0;
  if (!text_dev || (reinterpret_cast<uintptr_t>(text_dev) & 15)) return fail(BX_E_ARG, "text_dev must be 16-byte aligned");
  fasta_headers_kernel<<<grid_for((n_text + 15) / 16, 256), 256, 0, st>>>(
      text_dev, n_text, reinterpret_cast<long long*>(pos_dev), cap, reinterpret_cast<unsigned long long*>(count_dev));
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

int bx_fasta_count(const uint8_t* text_dev, int64_t body_start, int64_t body_end, int64_t* block_offsets_dev,
                   int64_t* n_bases_host, void* stream) {
  if (body_start < 0 || body_end < body_start || !block_offsets_dev || !n_bases_host)
    return fail(BX_E_ARG, "invalid arguments");
  if (int r = check_device()) return r;
  if (!text_dev || (reinterpret_cast<uintptr_t>(text_dev) & 15)) return fail(BX_E_ARG, "text_dev must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const long long n_blocks = bx_fasta_blocks(body_start, body_end);
  long long* off = reinterpret_cast<long long*>(block_offsets_dev);
  CK(cudaMemsetAsync(off + n_blocks, 0, sizeof(long long), st));  // the scan turns this extra entry into the total
  if (n_blocks > 0) {
    if (n_blocks + 1 > 0x7fffffffLL) return fail(BX_E_ARG, "contig body too large");
    fasta_count_kernel<<<grid_for(n_blocks, kFaThreads / 32), kFaThreads, 0, st>>>(text_dev, body_start & ~15LL, body_start, body_end,
                                                                      n_blocks, off);
    g_launches++;
    CK(cudaGetLastError());
    size_t tmp_bytes = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, off, off, (int)(n_blocks + 1), st));
    void* tmp = nullptr;
    CK(cudaMallocAsync(&tmp, tmp_bytes ? tmp_bytes : 1, st));
    cudaError_t e = cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, off, off, (int)(n_blocks + 1), st);
    g_launches++;
    cudaFreeAsync(tmp, st);
    CK(e);
  }
  long long total = 0;
  CK(cudaMemcpyAsync(&total, off + n_blocks, sizeof(long long), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  *n_bases_host = total;
  return 0;
}

int bx_fasta_pack(const uint8_t* text_dev, int64_t body_start, int64_t body_end, const int64_t* block_offsets_dev,
                  int64_t base_offset, uint32_t* bases2_dev, uint32_t* nmask_dev, int64_t* invalid_dev,
                  int64_t invalid_cap, int64_t* invalid_count_dev, void* stream) {
  if (body_start < 0 || body_end < body_start || base_offset < 0 || (base_offset & 31) || invalid_cap < 0)
    return fail(BX_E_ARG, "invalid arguments (base_offset must be a multiple of 32)");
  const long long n_blocks = bx_fasta_blocks(body_start, body_end);
  if (n_blocks == 0) return 0;
  if (!text_dev || !block_offsets_dev || !bases2_dev || !nmask_dev || !invalid_count_dev ||
      (invalid_cap > 0 && !invalid_dev))
    return fail(BX_E_ARG, "invalid arguments");
  if (reinterpret_cast<uintptr_t>(text_dev) & 15) return fail(BX_E_ARG, "text_dev must be 16-byte aligned");
  if (int r = check_device()) return r;
  cudaStream_t st = (cudaStream_t)stream;
  fasta_pack_kernel<<<grid_for(n_blocks, 1), kFaThreads, 0, st>>>(
      text_dev, body_start & ~15LL, body_start, body_end, n_blocks, reinterpret_cast<const long long*>(block_offsets_dev),
      base_offset, bases2_dev, nmask_dev, reinterpret_cast<long long*>(invalid_dev), invalid_cap,
      reinterpret_cast<unsigned long long*>(invalid_count_dev));
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}

}  // extern "C"

// ------------------------------------------------------------------ variant pipeline (chromatin.py:243-286 in the library)
// bx_chromatin_open takes whole-VCF column arrays once; the library classifies the variants (delta / full-segment /
// per-window path), builds every descriptor on the device (variants.cuh) and runs the chunk loop with its own
// streams: compute of chunk k+1 overlaps the device -> host copy of chunk k.  FC2's epilogue writes ref / alt /
// diff = alt - ref straight into the strand-stacked layout of chromatin.py:281-286 (EPI_CHROMATIN).
struct ChromPlan {
  int64_t n = 0;
  int maxshift = 0, S = 0, seg_len = 0, flags = 0, chunk = 0, n_chunks = 0;
  bool segments = false;           // segment mode usable (maxshift a multiple of 200, <= 1000)
  int patch_row0 = 0, patch_len = 0;
  int shifts[200] = {}, k_of_shift[200] = {};
  // device copies of the columns, the allele table and the per-chunk group lists
  long long* gpos = nullptr;
  int *ref_len = nullptr, *alt_len = nullptr, *ref_aoff = nullptr, *alt_aoff = nullptr, *idx = nullptr;
  uint8_t* alleles = nullptr;
  std::vector<int> n_delta, n_full, n_win;   // per chunk; idx holds [delta | full | window] local indices per chunk
  bx_window_desc* descs = nullptr;           // 4 * chunk descriptors of scratch
  float* stage[2][3] = {};                   // ref / alt / diff staging, [S][2][chunk][2002] each
  cudaEvent_t ev_done[2] = {}, ev_copied[2] = {};
  int set_chunk[2] = {-1, -1};
  int64_t out_pitch = BX_FEATURES;           // row pitch of bx_chromatin_submit_device's outputs
  VariantCols cols() const { return VariantCols{gpos, ref_len, alt_len, ref_aoff, alt_aoff}; }
};

namespace {

// windows.py: delta_patch_geometry for substitutions of up to 16 bases.
bool delta_patch_geometry(int maxshift, int* row0_out, int* len_out) {
  const int seg_len = 2000 + 2 * maxshift, v0 = maxshift + 999, ma = 16, rf = 310;
  const int smin = std::min(v0, std::min(seg_len - v0 - ma, seg_len - v0 - 1));
  const int emax = std::max(v0 + ma, seg_len - v0);
  if (smin - (rf + 7) < 0) return false;
  const int row0 = (smin - (rf + 7)) / 16;
  const int rows = (emax - 1) / 16 - row0 + 1;
  int L = 16 * ((16 * (rows - 1) + rf + 8 + 15) / 16);
  auto valid = [](int len) {
    const int p2 = (len - 14) / 4;
    return std::min((p2 - 14) / 4 - 14, (p2 - 16) / 4 - 14);
  };
  while (valid(L) < rows) L += 16;
  if (16 * row0 + L > seg_len) return false;
  *row0_out = row0; *len_out = L;
  return true;
}

template <typename T>
int upload(T** dst, const T* src, size_t n, cudaStream_t st) {
  CK(cudaMalloc(dst, std::max<size_t>(n, 1) * sizeof(T)));
  if (n) CK(cudaMemcpyAsync(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice, st));
  return 0;
}

void free_plan(ChromPlan* p) {
  if (!p) return;
  cudaFree(p->gpos); cudaFree(p->ref_len); cudaFree(p->alt_len); cudaFree(p->ref_aoff); cudaFree(p->alt_aoff);
  cudaFree(p->idx); cudaFree(p->alleles); cudaFree(p->descs);
  for (int i = 0; i < 2; ++i) {
    for (int j = 0; j < 3; ++j) cudaFree(p->stage[i][j]);
    if (p->ev_done[i]) cudaEventDestroy(p->ev_done[i]);
    if (p->ev_copied[i]) cudaEventDestroy(p->ev_copied[i]);
  }
  delete p;
}

// One chunk of the plan through the forward; results go to `sink` (rows relative to sink.row0).
int run_plan_chunk(bx_ctx* c, int k, ChromSink sink, cudaStream_t st) {
  ChromPlan* p = c->plan;
  const long long v0 = (long long)k * p->chunk;
  const int nd = p->n_delta[k], nf = p->n_full[k], nw = p->n_win[k];
  const int* idx = p->idx + v0;
  const VariantCols cols = p->cols();
  if (nd > 0) {
    bx_window_desc* rd = p->descs;
    bx_window_desc* pd = p->descs + 2 * (size_t)nd;
    build_delta_descs_kernel<<<(nd + 127) / 128, 128, 0, st>>>(cols, idx, v0, nd, p->maxshift, p->patch_row0,
                                                               p->patch_len, rd, pd);
    g_launches++;
    CK(cudaGetLastError());
    sink.dst = idx;
    if (sink.average) {
      // delta mode's alt rows add their halves into zeroed outputs (ChromEpi); the chunk's other groups overwrite
      // their own rows afterwards
      const int64_t v0m = std::min<int64_t>(p->chunk, p->n - v0);
      for (float* t : {sink.alt, sink.diff})
        for (int si = 0; si < p->S; ++si)
          CK(cudaMemsetAsync(t + (size_t)(si * sink.shift_stride + sink.row0) * sink.ld, 0,
                             (size_t)v0m * sink.ld * sizeof(float), st));
    }
    if (int r = forward_segments_impl(c, rd, pd, p->alleles, nd, 4, p->seg_len, p->patch_len, p->patch_row0, p->S,
                                      p->k_of_shift, 1, nullptr, st, &sink))
      return r;
  }
  if (nf > 0) {
    build_segment_descs_kernel<<<(nf + 127) / 128, 128, 0, st>>>(cols, idx + nd, v0, nf, p->maxshift, p->descs);
    g_launches++;
    CK(cudaGetLastError());
    sink.dst = idx + nd;
    if (int r = forward_segments_impl(c, p->descs, nullptr, p->alleles, nf, 4, p->seg_len, 0, 0, p->S, p->k_of_shift,
                                      1, nullptr, st, &sink))
      return r;
  }
  if (nw > 0) {   // long indels, exotic shift ranges: every window separately, with the reference's slice semantics
    const int* widx = idx + nd + nf;
    const int step = (c->cap & ~3) / 4;   // variants per forward (4 windows each)
    for (int si = 0; si < p->S; ++si) {
      for (int g0 = 0; g0 < nw; g0 += step) {
        const int m = std::min(step, nw - g0);
        build_window_descs_kernel<<<(m + 127) / 128, 128, 0, st>>>(cols, widx + g0, v0, m, p->shifts[si], p->descs);
        g_launches++;
        CK(cudaGetLastError());
        sink.dst = widx + g0;
        const ChromEpi ce = make_chrom(c, sink, 3, m, p->S, 0, si);
        if (int r = forward_chunk_windows(c, p->descs, p->alleles, 4 * m, nullptr, st, &ce)) return r;
      }
    }
  }
  return 0;
}

}  // namespace

extern "C" {

int bx_chromatin_close(bx_ctx* c) {
  if (!c) return fail(BX_E_ARG, "ctx is null");
  if (c->plan) {
    DeviceGuard guard(c->device);
    if (c->s_comp) cudaStreamSynchronize(c->s_comp);
    if (c->s_out) cudaStreamSynchronize(c->s_out);
    free_plan(c->plan);
    c->plan = nullptr;
  }
  return 0;
}

int bx_chromatin_open(bx_ctx* c, int64_t n, const int64_t* gpos, const int32_t* ref_len, const int32_t* alt_len,
                      const int32_t* ref_aoff, const int32_t* alt_aoff, const uint8_t* alleles, int64_t n_alleles,
                      int maxshift, int flags, int chunk_variants, int* n_chunks_out, int* chunk_variants_out) {
  if (!c || n < 0 || n_alleles < 0 || maxshift < 0) return fail(BX_E_ARG, "invalid arguments");
  if (n > 0 && (!gpos || !ref_len || !alt_len || !ref_aoff || !alt_aoff)) return fail(BX_E_ARG, "null column");
  if (n_alleles > 0 && !alleles) return fail(BX_E_ARG, "null allele table");
  if (!c->bases2) return fail(BX_E_STATE, "genome not set (bx_ctx_set_genome)");
  if (!c->weights_ready) return fail(BX_E_STATE, "weights not loaded (bx_ctx_load_weights)");
  if (c->cap < 4) return fail(BX_E_ARG, "context capacity must be >= 4 windows");
  if (maxshift % 200 != 0 && !(flags & BX_CHROM_WINDOWS))
    flags |= BX_CHROM_WINDOWS;   // chromatin.py:243's range() then has no +-maxshift entry; windows one by one
  const int S = 1 + 2 * (maxshift / 200);
  if (S > 200) return fail(BX_E_ARG, "maxshift %d gives %d shifts (limit 200)", maxshift, S);
  ON_DEVICE(c->device);
  bx_chromatin_close(c);
  if (int r = ensure_pipe(c, 0, 0, 0)) return r;
  ChromPlan* p = new ChromPlan();
  c->plan = p;
  auto bail = [&](int rc) { free_plan(p); c->plan = nullptr; return rc; };
  p->n = n; p->maxshift = maxshift; p->S = S; p->seg_len = 2000 + 2 * maxshift; p->flags = flags;
  {  // chromatin.py:243: [0, -200 .. -maxshift, 200 .. maxshift]; window offset index k = (shift + maxshift) / 200
    int i = 0;
    p->shifts[i++] = 0;
    for (int s = -200; s >= -maxshift; s -= 200) p->shifts[i++] = s;
    for (int s = 200; s <= maxshift; s += 200) p->shifts[i++] = s;
    int si_of_k[200];
    for (int j = 0; j < S; ++j) { p->k_of_shift[j] = (p->shifts[j] + maxshift) / 200; si_of_k[p->k_of_shift[j]] = j; }
    if (!c->si_of_k_dev) { if (cudaMalloc(&c->si_of_k_dev, 200 * sizeof(int)) != cudaSuccess) return bail(fail(BX_E_CUDA, "cudaMalloc failed")); }
    if (cudaMemcpy(c->si_of_k_dev, si_of_k, S * sizeof(int), cudaMemcpyHostToDevice) != cudaSuccess)
      return bail(fail(BX_E_CUDA, "cudaMemcpy of the shift table failed"));
  }
  p->segments = !(flags & BX_CHROM_WINDOWS) && maxshift <= 1000;
  bool delta = p->segments && !(flags & BX_CHROM_NO_DELTA) && S > 1 &&
               delta_patch_geometry(maxshift, &p->patch_row0, &p->patch_len);
  int chunk = p->segments ? max_segments(c, make_geom(p->seg_len)) / 4 : (c->cap & ~3) / 4;
  if (chunk < 1) return bail(fail(BX_E_ARG, "context capacity (%d windows) is too small for %d-position segments", c->cap, p->seg_len));
  if (chunk_variants > 0) chunk = std::min(chunk, chunk_variants);
  p->chunk = chunk;
  p->n_chunks = (int)((n + chunk - 1) / chunk);
  // classification (windows.py: segment_eligible, delta_eligible) and per-chunk group lists
  std::vector<int> idx((size_t)n);
  p->n_delta.assign(p->n_chunks, 0); p->n_full.assign(p->n_chunks, 0); p->n_win.assign(p->n_chunks, 0);
  for (int k = 0; k < p->n_chunks; ++k) {
    const int64_t v0 = (int64_t)k * chunk, m = std::min<int64_t>(chunk, n - v0);
    std::vector<int> full, win;
    int nd = 0;
    for (int64_t j = 0; j < m; ++j) {
      const int64_t rl = ref_len[v0 + j], al = alt_len[v0 + j];
      if (rl < 0 || al < 0 || ref_aoff[v0 + j] < 0 || alt_aoff[v0 + j] < 0 || ref_aoff[v0 + j] + rl > n_alleles ||
          alt_aoff[v0 + j] + al > n_alleles)
        return bail(fail(BX_E_ARG, "variant %lld: allele offsets outside the allele table", (long long)(v0 + j)));
      const bool seg_ok = p->segments && rl <= 1050 - maxshift && al - rl >= -100 && al <= 1000 && rl >= 1;
      if (seg_ok && delta && rl == al && rl <= 16) idx[v0 + nd++] = (int)j;
      else if (seg_ok) full.push_back((int)j);
      else win.push_back((int)j);
    }
    std::copy(full.begin(), full.end(), idx.begin() + v0 + nd);
    std::copy(win.begin(), win.end(), idx.begin() + v0 + nd + full.size());
    p->n_delta[k] = nd; p->n_full[k] = (int)full.size(); p->n_win[k] = (int)win.size();
  }
  cudaStream_t st = c->s_in;
  if (upload(&p->gpos, reinterpret_cast<const long long*>(gpos), (size_t)n, st) || upload(&p->ref_len, ref_len, (size_t)n, st) ||
      upload(&p->alt_len, alt_len, (size_t)n, st) || upload(&p->ref_aoff, ref_aoff, (size_t)n, st) ||
      upload(&p->alt_aoff, alt_aoff, (size_t)n, st) || upload(&p->idx, idx.data(), (size_t)n, st) ||
      upload(&p->alleles, alleles, (size_t)n_alleles, st))
    return bail(BX_E_CUDA);
  if (cudaMalloc(&p->descs, (size_t)4 * chunk * sizeof(bx_window_desc)) != cudaSuccess)
    return bail(fail(BX_E_CUDA, "cudaMalloc of the descriptor scratch failed"));
  for (int i = 0; i < 2; ++i) {
    if (cudaEventCreateWithFlags(&p->ev_done[i], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&p->ev_copied[i], cudaEventDisableTiming) != cudaSuccess)
      return bail(fail(BX_E_CUDA, "cudaEventCreate failed"));
  }
  if (cudaStreamSynchronize(st) != cudaSuccess) return bail(fail(BX_E_CUDA, "column upload failed"));
  if (n_chunks_out) *n_chunks_out = p->n_chunks;
  if (chunk_variants_out) *chunk_variants_out = chunk;
  return 0;
}

int bx_chromatin_chunk_info(bx_ctx* c, int chunk, int64_t* v0, int* m, int* n_delta, int* n_full, int* n_window) {
  if (!c || !c->plan) return fail(BX_E_STATE, "no open variant plan (bx_chromatin_open)");
  ChromPlan* p = c->plan;
  if (chunk < 0 || chunk >= p->n_chunks) return fail(BX_E_ARG, "chunk %d out of range (%d chunks)", chunk, p->n_chunks);
  const int64_t first = (int64_t)chunk * p->chunk;
  if (v0) *v0 = first;
  if (m) *m = (int)std::min<int64_t>(p->chunk, p->n - first);
  if (n_delta) *n_delta = p->n_delta[chunk];
  if (n_full) *n_full = p->n_full[chunk];
  if (n_window) *n_window = p->n_win[chunk];
  return 0;
}

int bx_chromatin_submit(bx_ctx* c, int chunk, float* ref_host, float* alt_host, float* diff_host) {
  if (!c || !c->plan) return fail(BX_E_STATE, "no open variant plan (bx_chromatin_open)");
  ChromPlan* p = c->plan;
  if (chunk < 0 || chunk >= p->n_chunks || !ref_host || !alt_host || !diff_host) return fail(BX_E_ARG, "invalid arguments");
  ON_DEVICE(c->device);
  const int set = chunk & 1;
  const int m = (int)std::min<int64_t>(p->chunk, p->n - (int64_t)chunk * p->chunk);
  const bool avg = p->flags & BX_CHROM_AVERAGE;
  const size_t rows = (size_t)p->S * (avg ? 1 : 2) * m;
  for (int j = 0; j < 3; ++j)
    if (!p->stage[set][j])
      CK(cudaMalloc(&p->stage[set][j], (size_t)p->S * 2 * p->chunk * BX_FEATURES * sizeof(float)));
  // the staging set's previous contents (chunk - 2) must have left for the host
  CK(cudaStreamWaitEvent(c->s_comp, p->ev_copied[set], 0));
  ChromSink sink = {};
  sink.ref = p->stage[set][0]; sink.alt = p->stage[set][1]; sink.diff = p->stage[set][2];
  sink.shift_stride = avg ? m : 2 * (long long)m;
  sink.strand_stride = m;
  sink.row0 = 0;
  sink.average = avg ? 1 : 0;
  {
    StageTimer t(c, 9, c->s_comp);   // the whole chunk (the per-layer stages nest inside when profiling is on)
    if (int r = run_plan_chunk(c, chunk, sink, c->s_comp)) return r;
  }
  CK(cudaEventRecord(p->ev_done[set], c->s_comp));
  CK(cudaStreamWaitEvent(c->s_out, p->ev_done[set], 0));
  float* dst[3] = {ref_host, alt_host, diff_host};
  for (int j = 0; j < 3; ++j)
    CK(cudaMemcpyAsync(dst[j], p->stage[set][j], rows * BX_FEATURES * sizeof(float), cudaMemcpyDeviceToHost, c->s_out));
  CK(cudaEventRecord(p->ev_copied[set], c->s_out));
  p->set_chunk[set] = chunk;
  return 0;
}

int bx_chromatin_wait(bx_ctx* c, int chunk) {
  if (!c || !c->plan) return fail(BX_E_STATE, "no open variant plan (bx_chromatin_open)");
  ChromPlan* p = c->plan;
  const int set = chunk & 1;
  if (chunk < 0 || p->set_chunk[set] != chunk)
    return fail(BX_E_STATE, "chunk %d is not in flight (wait for a chunk before submitting chunk + 2)", chunk);
  ON_DEVICE(c->device);
  CK(cudaEventSynchronize(p->ev_copied[set]));
  return 0;
}

int bx_chromatin_submit_device(bx_ctx* c, int chunk, float* ref_dev, float* alt_dev, float* diff_dev,
                               int64_t shift_stride_rows, int64_t strand_stride_rows, int64_t row0, void* stream) {
  if (!c || !c->plan) return fail(BX_E_STATE, "no open variant plan (bx_chromatin_open)");
  ChromPlan* p = c->plan;
  if (chunk < 0 || chunk >= p->n_chunks || !ref_dev || !alt_dev || !diff_dev) return fail(BX_E_ARG, "invalid arguments");
  ON_DEVICE(c->device);
  ChromSink sink = {};
  sink.ref = ref_dev; sink.alt = alt_dev; sink.diff = diff_dev;
  sink.shift_stride = shift_stride_rows;
  sink.strand_stride = strand_stride_rows;
  sink.row0 = row0;
  sink.average = (p->flags & BX_CHROM_AVERAGE) ? 1 : 0;
  sink.ld = p->out_pitch;
  return run_plan_chunk(c, chunk, sink, (cudaStream_t)stream);
}

int bx_chromatin_set_row_pitch(bx_ctx* c, int64_t row_pitch) {
  if (!c || !c->plan) return fail(BX_E_STATE, "no open variant plan (bx_chromatin_open)");
  if (row_pitch < BX_FEATURES || (row_pitch & 1))
    return fail(BX_E_ARG, "row_pitch must be an even number of floats >= %d", BX_FEATURES);
  c->plan->out_pitch = row_pitch;
  return 0;
}

int bx_variant_descs_host(int kind, int64_t n, const int64_t* gpos, const int32_t* ref_len, const int32_t* alt_len,
                          const int32_t* ref_aoff, const int32_t* alt_aoff, int maxshift, int shift,
                          int* patch_row0, int* patch_len, bx_window_desc* out_host) {
  if (n < 0 || kind < 0 || kind > 2) return fail(BX_E_ARG, "invalid arguments");
  int row0 = 0, plen = 0;
  if (kind == 1) {
    if (!delta_patch_geometry(maxshift, &row0, &plen)) return fail(BX_E_ARG, "no delta patch for maxshift %d", maxshift);
    if (patch_row0) *patch_row0 = row0;
    if (patch_len) *patch_len = plen;
  }
  if (n == 0) return 0;
  if (!gpos || !ref_len || !alt_len || !ref_aoff || !alt_aoff || !out_host) return fail(BX_E_ARG, "null argument");
  if (int r = check_device()) return r;
  if (n > (1 << 26)) return fail(BX_E_ARG, "too many variants for the test entry point");
  ChromPlan p;
  std::vector<int> idx((size_t)n);
  for (int64_t i = 0; i < n; ++i) idx[i] = (int)i;
  bx_window_desc* out = nullptr;
  cudaStream_t st = 0;
  int rc = 0;
  if (upload(&p.gpos, reinterpret_cast<const long long*>(gpos), (size_t)n, st) || upload(&p.ref_len, ref_len, (size_t)n, st) ||
      upload(&p.alt_len, alt_len, (size_t)n, st) || upload(&p.ref_aoff, ref_aoff, (size_t)n, st) ||
      upload(&p.alt_aoff, alt_aoff, (size_t)n, st) || upload(&p.idx, idx.data(), (size_t)n, st) ||
      cudaMalloc(&out, (size_t)4 * n * sizeof(bx_window_desc)) != cudaSuccess)
    rc = fail(BX_E_CUDA, "device allocation failed");
  if (!rc) {
    const unsigned blocks = (unsigned)((n + 127) / 128);
    if (kind == 0) build_segment_descs_kernel<<<blocks, 128, 0, st>>>(p.cols(), p.idx, 0, (int)n, maxshift, out);
    else if (kind == 1) build_delta_descs_kernel<<<blocks, 128, 0, st>>>(p.cols(), p.idx, 0, (int)n, maxshift, row0, plen, out, out + 2 * n);
    else build_window_descs_kernel<<<blocks, 128, 0, st>>>(p.cols(), p.idx, 0, (int)n, shift, out);
    g_launches++;
    if (cudaMemcpy(out_host, out, (size_t)4 * n * sizeof(bx_window_desc), cudaMemcpyDeviceToHost) != cudaSuccess)
      rc = fail(BX_E_CUDA, "descriptor kernel failed: %s", cudaGetErrorString(cudaGetLastError()));
  }
  cudaFree(p.gpos); cudaFree(p.ref_len); cudaFree(p.alt_len); cudaFree(p.ref_aoff); cudaFree(p.alt_aoff); cudaFree(p.idx);
  cudaFree(out);
  return rc;
}

}  // extern "C"

// ------------------------------------------------------------------ bx_fasta_count without any allocation
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  bx_fasta_count takes
// the temporary storage of its CUB scan from cudaMallocAsync and hands the contig length back through a stream
// synchronise; this variant takes that scratch from a file-scope device array and leaves the length on the device.
// Every GPU that loads the library holds the array (256 KB) until exit.
constexpr size_t kFaScratchBytes = 256 * 1024;
constexpr long long kFaNoAllocMaxBlocks = 1LL << 20;   // 4096-byte blocks: contig bodies of up to 4 GiB of text
static __device__ __align__(256) unsigned char g_fa_scan_scratch[kFaScratchBytes];

namespace {
// Device address of g_fa_scan_scratch per device.  In host code the array's name is a host-side stand-in, not a
// device address, so it is resolved once per device, with that device current, through cudaGetSymbolAddress.
std::atomic<void*> g_fa_scratch_addr[kMaxDevices];

int fa_scratch_ptr(void** out) {
  int dev = 0;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(BX_E_ARG, "device index %d out of range", dev);
  void* p = g_fa_scratch_addr[dev].load(std::memory_order_acquire);
  if (!p) {
    CK(cudaGetSymbolAddress(&p, g_fa_scan_scratch));   // the current device's instance of the array
    if (!p) return fail(BX_E_CUDA, "cudaGetSymbolAddress returned a null address");
    g_fa_scratch_addr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return 0;
}
}  // namespace

extern "C" int bx_fasta_count_noalloc(const uint8_t* text_dev, int64_t body_start, int64_t body_end,
                                      int64_t* block_offsets_dev, int64_t* n_bases_dev, void* stream) {
  if (body_start < 0 || body_end < body_start || !block_offsets_dev || !n_bases_dev)
    return fail(BX_E_ARG, "invalid arguments");
  if (int r = check_device()) return r;
  if (!text_dev || (reinterpret_cast<uintptr_t>(text_dev) & 15)) return fail(BX_E_ARG, "text_dev must be 16-byte aligned");
  const long long n_blocks = bx_fasta_blocks(body_start, body_end);
  if (n_blocks > kFaNoAllocMaxBlocks)
    return fail(BX_E_ARG, "contig body of %lld 4096-byte blocks exceeds the %lld blocks (4 GiB of text) the static "
                          "scratch of bx_fasta_count_noalloc is sized for; use bx_fasta_count",
                n_blocks, kFaNoAllocMaxBlocks);
  void* scratch = nullptr;
  if (int r = fa_scratch_ptr(&scratch)) return r;
  cudaStream_t st = (cudaStream_t)stream;
  long long* off = reinterpret_cast<long long*>(block_offsets_dev);
  // the extra entry must start at zero: the scan turns it into the total.  (The scratch itself needs no clearing:
  // CUB's scan initialises the tile states it keeps there at the start of every call.)
  CK(cudaMemsetAsync(off + n_blocks, 0, sizeof(long long), st));
  if (n_blocks > 0) {
    fasta_count_kernel<<<grid_for(n_blocks, kFaThreads / 32), kFaThreads, 0, st>>>(text_dev, body_start & ~15LL, body_start,
                                                                                   body_end, n_blocks, off);
    g_launches++;
    CK(cudaGetLastError());
    size_t tmp_bytes = 0;
    CK(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, off, off, (int)(n_blocks + 1), st));
    if (tmp_bytes > kFaScratchBytes)
      return fail(BX_E_STATE, "the scan of %lld blocks needs %zu bytes of scratch, the static array holds %zu", n_blocks,
                  tmp_bytes, kFaScratchBytes);
    CK(cub::DeviceScan::ExclusiveSum(scratch, tmp_bytes, off, off, (int)(n_blocks + 1), st));
    g_launches++;
  }
  CK(cudaMemcpyAsync(n_bases_dev, off + n_blocks, sizeof(long long), cudaMemcpyDeviceToDevice, st));
  return 0;
}

#ifdef BX_SC_TRACE
extern "C" int bx_debug_score_trace(long long* out_host) {
  return cudaMemcpyFromSymbol(out_host, bx::g_sc_trace, sizeof(bx::g_sc_trace)) == cudaSuccess ? 0 : -1;
}
#endif
