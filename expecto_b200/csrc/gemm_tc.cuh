// Persistent, warp-specialised tcgen05 implicit-GEMM kernel used for every dense contraction of the
// Beluga forward (reference: Beluga.py:23-45 — six Conv2d(1,8) layers and two Linear layers).
//
// Data layout (HBM): activations are channels-last fp16/bf16 matrices [rows, C] where rows is the
// flattened (window, position) axis with a fixed per-window pitch.  A Conv2d(Cin,Cout,(1,8)) is then
//     out[r, co] = sum_{t<8} sum_{ci} act[r + t, ci] * W_t[co, ci]
// i.e. eight GEMMs whose A operands are the SAME activation tile shifted by t rows.  The producer
// TMA-loads one [128+8 rows x 64 ch] slab per K-block (SWIZZLE_128B) and the MMA thread addresses
// tap t as a +t*128-byte view of that slab, so activations cross L2->SMEM once, not eight times.
// Weights are stored tap-major [TAPS*Cout, Cin] (K-major) and stream through their own ring.
//
// Roles (192 threads): warp 0 = TMA producer, warp 1 = MMA issuer + TMEM allocator,
// warps 2..5 = epilogue (TMEM -> registers -> bias/ReLU/pool/sigmoid -> global).
// CTA_GROUP == 2 runs a CTA pair (cluster of 2): M = 256 per pair, each CTA stages its own 128 A
// rows and half of the B rows, the leader CTA issues tcgen05.mma.cta_group::2.
// Accumulators live in a TMEM ring of 512/BLOCK_N buffers so the epilogue overlaps the next tile.
//
// A_MODE != 0 (conv1 only; reference: fetchSeqs + encodeSeqs feeding model.forward, chromatin.py:138-209,270):
// the A operand is never in HBM.  The producer warp reads the 2-bit packed genome through the window / segment
// descriptors (A_MODE 1) or an explicit one-hot tensor (A_MODE 2), reverse-complements, and writes the
// 128 x [8 taps x 4 channels] operand tile straight into the swizzled A stage in shared memory; the weights
// still arrive by TMA.  The tile is built once per 128-row block and reused for every N chunk.
#pragma once
#include <type_traits>
#include "encode.cuh"
#include "ptx.cuh"

namespace bx {

enum EpiKind : int {
  EPI_BIAS_RELU = 0,       // y = relu(acc + b) -> 16-bit, rows remapped pitch_in -> pitch_out
  EPI_POOL4_BIAS_RELU = 1, // y = max over 4 consecutive rows of relu(acc + b) -> 16-bit (MaxPool(1,4))
  EPI_PARTIAL_F32 = 2,     // split-K partial sums: plain fp32 stores into slice ks of [k_splits, rows, ld_out]
                           // (summed in fixed order by the finaliser: deterministic, no atomics)
  EPI_BIAS_SIGMOID = 3,    // y = sigmoid(acc + b) -> fp32
  EPI_POOL2_BIAS_RELU = 4, // y = max over 2 consecutive rows of relu(acc + b) -> 16-bit: half of a MaxPool(1,4), so
                           // that segment mode can finish both pooling phases from one conv4 pass (pool_phases_kernel)
  EPI_CHROMATIN = 5,       // FC2 of the variant path: p = sigmoid(acc + b) written straight into chromatin.py's
                           // ref / alt / diff = alt - ref row layout (chromatin.py:281-286), optionally averaged over
                           // the forward / reverse-complement rows as predict.py:186-190 reads them (ChromEpi)
};

// Where the rows of an FC2 launch go in the variant path (EPI_CHROMATIN).  Every GEMM row is one window
// (variant, shift, q) with q = 0 ref fwd, 1 ref rc, 2 alt fwd, 3 alt rc; its probabilities land in
//   X[(si * shift_stride + (q & 1) * strand_stride + row0 + dst(grp)) * ld + feature],  X = ref (q < 2) or alt,
// and alt rows also produce diff = alt - ref of the same strand.  Row order of the launch:
//   mode 0: (si, grp, q) shift-major quads [segment mode, four full segments per variant]
//   mode 1: delta mode's ref rows  d = (si * ng + grp) * 2 + t
//   mode 2: delta mode's alt rows  d = c * 2 * ng + grp * 2 + t, si = si_of_k[t ? S - 1 - c : c]
//   mode 3: (grp, q) quads of one shift si_fixed [per-window path]
// Quads sit in four adjacent lanes of one warp (row offsets are multiples of 4), so alt - ref is a shfl.xor 2 away;
// delta mode's alt launch reads the ref probabilities its ref launch stored (same stream, earlier kernel).
// average: outputs are [.., row, 2002] without the strand axis: ref = (fwd + rc) / 2, alt likewise,
// diff = ((alt_f - ref_f) + (alt_r - ref_r)) / 2 — the exact expressions of predict.py:186-190 on chromatin.py:281's
// per-strand differences; per-strand ref probabilities of mode 1 go to ref_scratch for the mode-2 launch.  The two
// strands of one shift are neighbouring lanes in modes 0, 1 and 3.  In mode 2 they are not (a forward window of
// offset class c and the reverse-complement window of the same class belong to different shifts), so each alt row
// adds its half, x / 2, to outputs the caller zeroed: x / 2 is exact and a two-term sum does not depend on the
// order of its terms, so fl(a/2 + b/2) = fl(a + b) / 2 whichever row arrives first — deterministic and bit-identical.
struct ChromEpi {
  float* ref; float* alt; float* diff;
  float* ref_scratch;            // average && delta mode: per-strand ref rows [S][2][ng] (dense, dst ignored)
  const int* dst;                // [ng] local output row of each group member, or nullptr = identity
  const int* si_of_k;            // [S] device table (mode 2)
  long long shift_stride, strand_stride, row0;   // in rows of ld floats
  long long ld;                  // row pitch of ref / alt / diff: 2002, or padded (bx_chromatin_set_row_pitch)
  long long first;               // index of the launch's first row in its mode's row order
  int mode, ng, S, si_fixed, average;
};

struct GemmParams {
  int m_rows;        // valid rows of A (flattened window x position)
  int k_valid;       // K extent actually multiplied (per tap)
  int n_total;       // output channels (rows of B per tap)
  int m_tiles;       // ceil(m_rows / (128*CTA_GROUP))
  int n_chunks;      // n_total / BLOCK_N
  int k_splits;      // split-K factor (EPI_PARTIAL_F32 only, else 1)
  int k_blocks;      // ceil(k_valid / 64)
  int kb_per_split;  // ceil(k_blocks / k_splits)
  int pitch_in;      // rows per window in the accumulator row space
  int pitch_out;     // rows per window in the output
  int l_store;       // output rows stored per window
  int ld_out;        // output leading dimension in elements
  int n_store;       // columns >= n_store are not stored
  long long split_stride;  // elements between split-K slices of the output (EPI_PARTIAL_F32)
  int debug_flags;   // diagnostics only (BX_DEBUG_FLAGS): 1 = no B loads, 2 = no A loads, 4 = no epilogue work,
                     // 8 = B ring filled once and then reused (real weight values, no further L2->SMEM weight traffic)
  const float* bias; // [n_total] fp32 (unused for EPI_PARTIAL_F32)
  void* out;
  // Optional per-M-tile split-K ranges (EPI_PARTIAL_F32): tile mt computes slices tile_lo[mt] .. tile_lo[mt] +
  // tile_n[mt] - 1 only (delta mode's alt windows need just the slices whose K range holds changed rows).
  int n_tile_ranges;  // 0 = every tile computes all k_splits slices
  unsigned char tile_lo[32], tile_n[32];
  // Split-operand mode (BX_DTYPE_F16X3): A = Ah + Al and W = Wh + Wl as fp16 pairs, product = Ah*Wh + Ah*Wl + Al*Wh
  // accumulated in fp32 (the Al*Wl term is below fp32 resolution).  parts = 3 runs every logical K-block three times:
  // part 0 (Ah, Wh), part 1 (Ah, Wl: the A stage of part 0 is reused), part 2 (Al, Wh).  parts = 1: plain operands.
  // Column (K element) of logical K-block kb in the A / B tensor maps:
  //   A: (kb / a_grp) * a_grp_stride + (kb % a_grp) * 64 + (part == 2 ? a_lo_off : 0)
  //   B: kb * 64 + (part == 1 ? b_lo_off : 0)
  int parts, a_grp, a_grp_stride, a_lo_off, b_lo_off;
  int out_lo_off;    // SPLIT_OUT epilogues: the low half of an output goes out_lo_off elements after the high half
  ChromEpi chrom;    // EPI_CHROMATIN only
};

template <int TAPS, int BLOCK_N, int CG>
struct GemmCfg {
  static constexpr int A_ROWS = 128 + (TAPS > 1 ? 8 : 0);
  static constexpr int A_BYTES = A_ROWS * 128;
  static constexpr int B_ROWS = BLOCK_N / CG;
  // taps per B stage: with one tap (4 MMAs) per barrier round trip the single issuing thread, not the tensor
  // pipe, paced the N=160 layers (~95 cycles per MMA issued vs the 80-cycle MMA); two taps halve that overhead
  static constexpr int TPS = (TAPS > 1) ? 2 : 1;
  static constexpr int B_TAP_BYTES = B_ROWS * 128;
  static constexpr int B_BYTES = TPS * B_TAP_BYTES;
  static constexpr int BIAS_BYTES = 8192;                // fp32 bias vector staged once per CTA
  static constexpr int SMEM_BUDGET = 227 * 1024 - 2048 - BIAS_BYTES;  // barriers + 1 KB alignment slack + bias
  // Stages of the A ring.  The dense layers (one tap, N = 256: 16 KB + 16 KB per stage and CTA) stream every operand
  // byte exactly once, so their rate is bytes in flight / L2 latency: six stages instead of four took FC1 from 0.96
  // to 0.83 ms per 4096 windows (tools/dense_sweep.py); conv1 (one tap, N = 160) gains nothing from more than four.
  static constexpr int SA = (TAPS > 1) ? 3 : (BLOCK_N == 256 ? 6 : 4);
  static constexpr int SB_MAX = (SMEM_BUDGET - SA * A_BYTES) / B_BYTES;
  static constexpr int SB = (TAPS > 1) ? (SB_MAX > 8 ? 8 : SB_MAX) : (SB_MAX > SA ? SA : SB_MAX);
  static constexpr int NBUF = 512 / BLOCK_N;
  static constexpr int SMEM_BYTES = SA * A_BYTES + SB * B_BYTES + 1024 /*barriers*/ + BIAS_BYTES + 1024 /*align*/;
  static_assert(B_TAP_BYTES % 1024 == 0, "B tap tile must keep 1024-byte alignment (SWIZZLE_128B)");
  static_assert(TAPS % TPS == 0, "taps per stage must divide the tap count");
  static_assert(A_BYTES % 1024 == 0, "A stage must keep 1024-byte alignment (SWIZZLE_128B)");
  static_assert(SB >= 2 && NBUF >= 1, "pipeline too shallow");
  static_assert(BLOCK_N % 16 == 0 && BLOCK_N <= 256, "invalid UMMA N");
};

template <typename T>
__device__ __forceinline__ uint32_t pack2(float a, float b);
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  __half2 h = __floats2half2_rn(fminf(a, 65504.f), fminf(b, 65504.f));
  return *reinterpret_cast<uint32_t*>(&h);
}
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
// Two fp32 values -> packed 16-bit high parts (round to nearest) and packed 16-bit residuals (value - high part).
template <typename T>
__device__ __forceinline__ void split2(float a, float b, uint32_t& hi, uint32_t& lo);
template <>
__device__ __forceinline__ void split2<__half>(float a, float b, uint32_t& hi, uint32_t& lo) {
  a = fminf(a, 65504.f); b = fminf(b, 65504.f);
  const __half2 h = __floats2half2_rn(a, b);
  const float2 f = __half22float2(h);
  const __half2 l = __floats2half2_rn(a - f.x, b - f.y);
  hi = *reinterpret_cast<const uint32_t*>(&h);
  lo = *reinterpret_cast<const uint32_t*>(&l);
}
template <>
__device__ __forceinline__ void split2<__nv_bfloat16>(float a, float b, uint32_t& hi, uint32_t& lo) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  const float2 f = __bfloat1622float2(h);
  const __nv_bfloat162 l = __floats2bfloat162_rn(a - f.x, b - f.y);
  hi = *reinterpret_cast<const uint32_t*>(&h);
  lo = *reinterpret_cast<const uint32_t*>(&l);
}
template <typename T>
__device__ __forceinline__ uint32_t max2(uint32_t a, uint32_t b);
template <>
__device__ __forceinline__ uint32_t max2<__half>(uint32_t a, uint32_t b) {
  __half2 r = __hmax2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}
template <>
__device__ __forceinline__ uint32_t max2<__nv_bfloat16>(uint32_t a, uint32_t b) {
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}

struct RingState {
  uint32_t idx = 0, phase = 0;
  template <int N>
  __device__ __forceinline__ void advance() {
    if (++idx == N) { idx = 0; phase ^= 1; }
  }
};

// PARTS: 1 = plain operands; 3 = split operands (compile-time so that the plain instances keep the exact loop
// structure they were tuned with: FC1's producer / issuer loops are sensitive to a few instructions per K-block).
template <typename T, int TAPS, int BLOCK_N, int CG, int EPI, int A_MODE = 0, int SPLIT_OUT = 0, int PARTS = 1>
__global__ void __launch_bounds__(192, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const GemmParams p, const EncParams enc) {
  static_assert(A_MODE == 0 || TAPS == 1, "in-kernel operand generation is the conv1 (im2col, one K block) form");
  static_assert(PARTS == 1 || (PARTS == 3 && A_MODE == 0), "split operands: three parts, TMA-loaded A");
  using Cfg = GemmCfg<TAPS, BLOCK_N, CG>;
  constexpr int SA = Cfg::SA, SB = Cfg::SB, NBUF = Cfg::NBUF;
  constexpr int FMT = sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value ? 1 : 0;
  constexpr uint32_t IDESC = make_idesc_f16(FMT, 128 * CG, BLOCK_N);

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;  // identical in both CTAs of a pair
  const uint32_t sA = base;
  const uint32_t sB = sA + SA * Cfg::A_BYTES;
  const uint32_t bars = sB + SB * Cfg::B_BYTES;
  auto fullA = [&](uint32_t i) { return bars + 8u * i; };
  auto emptyA = [&](uint32_t i) { return bars + 8u * (SA + i); };
  auto fullB = [&](uint32_t i) { return bars + 8u * (2 * SA + i); };
  auto emptyB = [&](uint32_t i) { return bars + 8u * (2 * SA + SB + i); };
  auto tfull = [&](uint32_t i) { return bars + 8u * (2 * SA + 2 * SB + i); };
  auto tempty = [&](uint32_t i) { return bars + 8u * (2 * SA + 2 * SB + NBUF + i); };
  const uint32_t tmem_slot = bars + 8u * (2 * SA + 2 * SB + 2 * NBUF);
  float* sbias = reinterpret_cast<float*>(smem_raw + (bars + 1024 - raw));  // [n_total] fp32, n_total <= 2048
  // A_MODE scratch: 144 per-position operand patterns (upper half of the bias area; n_total <= 1024 there)
  unsigned long long* s_pat = reinterpret_cast<unsigned long long*>(smem_raw + (bars + 1024 + 4096 - raw));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = (CG == 1) ? 0u : cluster_ctarank();
  const int num_clusters = gridDim.x / CG;
  const int cluster_id = blockIdx.x / CG;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&tmA);
    prefetch_tensormap(&tmB);
    // A_MODE: one arrival per CTA of the pair (each producer warp fills its own 128 rows) instead of TMA bytes
    for (int i = 0; i < SA; ++i) { mbar_init(fullA(i), A_MODE ? CG : 1); mbar_init(emptyA(i), 1); }
    for (int i = 0; i < SB; ++i) { mbar_init(fullB(i), 1); mbar_init(emptyB(i), 1); }
    for (int i = 0; i < NBUF; ++i) { mbar_init(tfull(i), 1); mbar_init(tempty(i), 4 * CG); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<CG>(tmem_slot, 512);
  if (p.bias != nullptr) {  // n_total is a multiple of 16 and the bias vectors are cudaMalloc'd: 16-byte loads
    const float4* b4 = reinterpret_cast<const float4*>(p.bias);
    float4* s4 = reinterpret_cast<float4*>(sbias);
    for (int i = threadIdx.x; i < (p.n_total >> 2); i += blockDim.x) s4[i] = __ldg(b4 + i);
  }
  tc_fence_before();
  if constexpr (CG == 1) __syncthreads(); else cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int tiles_mn = p.m_tiles * p.n_chunks;
  const int total_units = tiles_mn * p.k_splits;
  // Work unit `it` of this cluster.  Default: units strided over the clusters.  A_MODE: 128-row blocks strided
  // over the clusters, all N chunks of a block back to back so the generated A tile is reused.
  auto unit = [&](int it, int& ks, int& mt, int& nc) -> bool {
    if constexpr (A_MODE != 0) {
      const int g = it / p.n_chunks;
      ks = 0;
      mt = cluster_id + g * num_clusters;
      nc = it - g * p.n_chunks;
      return mt < p.m_tiles;
    } else {
      const int u = cluster_id + it * num_clusters;
      if (p.n_tile_ranges > 0) {  // units ordered (tile, slice, N chunk)
        int first = 0;
        for (int t = 0; t < p.n_tile_ranges; ++t) {
          const int cnt = (int)p.tile_n[t] * p.n_chunks;
          if (u < first + cnt) {
            const int l = u - first;
            mt = t;
            ks = (int)p.tile_lo[t] + l / p.n_chunks;
            nc = l % p.n_chunks;
            return true;
          }
          first += cnt;
        }
        return false;
      }
      if (u >= total_units) return false;
      ks = u / tiles_mn;
      const int rem = u - ks * tiles_mn;
      mt = rem / p.n_chunks;
      nc = rem - mt * p.n_chunks;
      return true;
    }
  };

  // Producer and MMA warps run their loops warp-uniformly (all 32 lanes wait on the barriers) and only
  // the elected lane issues the async instruction: descriptors, coordinates and barrier addresses then
  // live in uniform registers and one tcgen05.mma costs a handful of issue slots.  (A first version
  // ran these loops under `if (lane == 0)`; ncu showed ~40 SASS instructions per MMA from the
  // divergent-code UR waterfall and the single thread became the bottleneck at ~195 cycles/MMA.)
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    RingState ra, rb;
    int ks, mt, nc;
    int b_fills = 0;  // debug flag 8
    // A_MODE: raw genome / one-hot words of the NEXT 128-row block, loaded one block ahead (encode.cuh)
    EncRaw<A_MODE> enc_raw;
    auto load_block = [&](int mtile) {
      const int r0 = mtile * 128 * CG + (int)rank * 128;
      const int w = r0 / p.pitch_in, l = r0 - w * p.pitch_in;
      enc_tile_load<A_MODE>(enc, w, l, p.pitch_in - l, lane, enc_raw);
    };
    if constexpr (A_MODE != 0) {
      if (cluster_id < p.m_tiles) load_block(cluster_id);
    }
    for (int it = 0; unit(it, ks, mt, nc); ++it) {
      const int kb0 = ks * p.kb_per_split;
      const int kb1 = min(p.k_blocks, kb0 + p.kb_per_split);
      const int row0 = mt * 128 * CG + (int)rank * 128;
      const int ncol0 = nc * BLOCK_N + (int)rank * Cfg::B_ROWS;
      // A column of logical K-block kb, kept incrementally (no division in the loop): groups of a_grp K-blocks are
      // a_grp_stride elements apart
      int a_in_grp = 0, a_col0 = 0;
      if constexpr (PARTS > 1) {
        a_in_grp = kb0 % p.a_grp;
        a_col0 = (kb0 / p.a_grp) * p.a_grp_stride + a_in_grp * 64;
      }
      for (int kb = kb0; kb < kb1; ++kb) {
        if constexpr (A_MODE != 0) {
          if (nc == 0) {
            mbar_wait(emptyA(ra.idx), ra.phase ^ 1, 1);
            if (!(p.debug_flags & 2)) {
              // Slots of s_pat: see enc_tile_load (pitch_in >= 136, so a tile touches at most two windows).
              const int jB = p.pitch_in - row0 % p.pitch_in;
              int codes[5];
              enc_tile_codes<A_MODE>(enc_raw, codes);
              if (mt + num_clusters < p.m_tiles) load_block(mt + num_clusters);  // in flight during the stores below
#pragma unroll
              for (int k = 0; k < 5; ++k)  // 64-bit operand pattern of one position: k = tap*4 + channel
                if (lane + 32 * k < 144)
                  s_pat[lane + 32 * k] =
                      (codes[k] < 4) ? ((unsigned long long)one_bits<T>() << (16 * codes[k])) : 0ull;
              __syncwarp();
              uint8_t* stage = smem_raw + (sA - raw) + ra.idx * Cfg::A_BYTES;
#pragma unroll
              for (int jj = 0; jj < 4; ++jj) {
                const int j = jj * 32 + lane;
                const int o = (j < jB) ? j : j + 8;
#pragma unroll
                for (int q = 0; q < 4; ++q) {  // 16-byte piece q = taps 2q, 2q+1; SWIZZLE_128B: piece ^ (row & 7)
                  const unsigned long long lo = s_pat[o + 2 * q], hi = s_pat[o + 2 * q + 1];
                  const uint4 piece = make_uint4((uint32_t)lo, (uint32_t)(lo >> 32), (uint32_t)hi, (uint32_t)(hi >> 32));
                  *reinterpret_cast<uint4*>(stage + j * 128 + ((q ^ (j & 7)) << 4)) = piece;
                  // split weights (k_valid = 64): K elements 32..63 multiply the low halves of the weights by the
                  // same one-hot pattern
                  if (p.k_valid > 32) *reinterpret_cast<uint4*>(stage + j * 128 + (((q + 4) ^ (j & 7)) << 4)) = piece;
                }
              }
              fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
            }
            __syncwarp();
            if (elect_one()) {
              if constexpr (CG == 1) mbar_arrive_local(fullA(ra.idx));
              else mbar_arrive_remote(mapa(fullA(ra.idx), 0));
            }
            __syncwarp();
            ra.advance<SA>();
          }
        }
#pragma unroll
        for (int part = 0; part < PARTS; ++part) {
        if constexpr (A_MODE == 0) {
          if (part != 1) {  // part 1 (Ah x Wl) reuses the A stage of part 0
          mbar_wait(emptyA(ra.idx), ra.phase ^ 1, 1);
          const int a_col = (PARTS == 1) ? kb * 64 : a_col0 + (part == 2 ? p.a_lo_off : 0);
          if (p.debug_flags & 2) {
            if (elect_one() && rank == 0) mbar_arrive_local(fullA(ra.idx));
          } else if (elect_one()) {
            if constexpr (CG == 1) {
              mbar_arrive_expect_tx(fullA(ra.idx), Cfg::A_BYTES);
              tma_load_2d(sA + ra.idx * Cfg::A_BYTES, &tmA, fullA(ra.idx), a_col, row0);
            } else {
              if (rank == 0) mbar_arrive_expect_tx(fullA(ra.idx), 2 * Cfg::A_BYTES);
              tma_load_2d_pair(sA + ra.idx * Cfg::A_BYTES, &tmA, mapa(fullA(ra.idx), 0), a_col, row0);
            }
          }
          __syncwarp();
          ra.advance<SA>();
          }
        }
        const int b_col = kb * 64 + ((PARTS > 1 && part == 1) ? p.b_lo_off : 0);
#pragma unroll 1
        for (int t = 0; t < TAPS; t += Cfg::TPS) {
          mbar_wait(emptyB(rb.idx), rb.phase ^ 1, 2);
          if ((p.debug_flags & 1) || ((p.debug_flags & 8) && b_fills >= SB)) {
            if (elect_one() && rank == 0) mbar_arrive_local(fullB(rb.idx));
          } else if (elect_one()) {
            const uint32_t dst = sB + rb.idx * Cfg::B_BYTES;
            if constexpr (CG == 1) {
              mbar_arrive_expect_tx(fullB(rb.idx), Cfg::B_BYTES);
#pragma unroll
              for (int j = 0; j < Cfg::TPS; ++j)
                tma_load_2d(dst + j * Cfg::B_TAP_BYTES, &tmB, fullB(rb.idx), b_col, (t + j) * p.n_total + ncol0);
            } else {
              if (rank == 0) mbar_arrive_expect_tx(fullB(rb.idx), 2 * Cfg::B_BYTES);
              const uint32_t bar = mapa(fullB(rb.idx), 0);
#pragma unroll
              for (int j = 0; j < Cfg::TPS; ++j)
                tma_load_2d_pair(dst + j * Cfg::B_TAP_BYTES, &tmB, bar, b_col, (t + j) * p.n_total + ncol0);
            }
          }
          __syncwarp();
          rb.advance<SB>();
          ++b_fills;
        }
        }  // part
        if constexpr (PARTS > 1) {
          if (++a_in_grp == p.a_grp) { a_in_grp = 0; a_col0 += p.a_grp_stride - (p.a_grp - 1) * 64; }
          else a_col0 += 64;
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA, elected lane)
    if (rank == 0) {
      RingState ra, rb, rt;
      int ks, mt, nc;
      for (int it = 0; unit(it, ks, mt, nc); ++it) {
        const int kb0 = ks * p.kb_per_split;
        const int kb1 = min(p.k_blocks, kb0 + p.kb_per_split);
        mbar_wait(tempty(rt.idx), rt.phase ^ 1, 3);
        tc_fence_after();
        const uint32_t d = tmem_base + rt.idx * BLOCK_N;
        uint32_t accum = 0;
        for (int kb = kb0; kb < kb1; ++kb) {
          const int kcount = min(4, (p.k_valid - kb * 64 + 15) >> 4);
#pragma unroll
          for (int part = 0; part < PARTS; ++part) {
          if ((A_MODE == 0 && part != 1) || (A_MODE != 0 && nc == 0)) mbar_wait(fullA(ra.idx), ra.phase, 4);
          const uint32_t a_lo = sw128_desc_lo(sA + ra.idx * Cfg::A_BYTES);
#pragma unroll 1
          for (int t = 0; t < TAPS; t += Cfg::TPS) {
            mbar_wait(fullB(rb.idx), rb.phase, 5);
            tc_fence_after();
            // tap t = the A slab viewed t rows (t*128 B) further down; K advances 32 B per UMMA_K=16
            const uint32_t b0_lo = sw128_desc_lo(sB + rb.idx * Cfg::B_BYTES);
            if (elect_one()) {
#pragma unroll
              for (int j = 0; j < Cfg::TPS; ++j) {
                const uint32_t at_lo = a_lo + (t + j) * 8;
                const uint32_t b_lo = b0_lo + j * (Cfg::B_TAP_BYTES >> 4);
                if (kcount == 4) {
                  umma_f16<CG>(d, sw128_desc(at_lo), sw128_desc(b_lo), IDESC, accum);
                  umma_f16<CG>(d, sw128_desc(at_lo + 2), sw128_desc(b_lo + 2), IDESC, 1u);
                  umma_f16<CG>(d, sw128_desc(at_lo + 4), sw128_desc(b_lo + 4), IDESC, 1u);
                  umma_f16<CG>(d, sw128_desc(at_lo + 6), sw128_desc(b_lo + 6), IDESC, 1u);
                } else if (kcount == 2) {
                  umma_f16<CG>(d, sw128_desc(at_lo), sw128_desc(b_lo), IDESC, accum);
                  umma_f16<CG>(d, sw128_desc(at_lo + 2), sw128_desc(b_lo + 2), IDESC, 1u);
                } else {
                  for (int k = 0; k < kcount; ++k)
                    umma_f16<CG>(d, sw128_desc(at_lo + 2 * k), sw128_desc(b_lo + 2 * k), IDESC, k ? 1u : accum);
                }
                accum = 1;
              }
              umma_commit<CG>(emptyB(rb.idx));
            }
            __syncwarp();
            accum = 1;
            rb.advance<SB>();
          }
          // release the A stage after its last use: split mode keeps Ah for parts 0 and 1; A_MODE: the generated
          // tile serves every N chunk
          if ((A_MODE == 0 && (PARTS == 1 || part >= 1)) || (A_MODE != 0 && nc == p.n_chunks - 1)) {
            if (elect_one()) umma_commit<CG>(emptyA(ra.idx));
            __syncwarp();
            ra.advance<SA>();
          }
          }  // part
        }
        if (elect_one()) umma_commit<CG>(tfull(rt.idx));
        __syncwarp();
        rt.advance<NBUF>();
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps 2..5
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    RingState rt;
    int ks, mt, nc;
    for (int it = 0; unit(it, ks, mt, nc); ++it) {
      const int r = mt * 128 * CG + (int)rank * 128 + q * 32 + lane;
      const int wb = r / p.pitch_in;
      const int l = r - wb * p.pitch_in;
      mbar_wait(tfull(rt.idx), rt.phase, 6);
      tc_fence_after();
      const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + rt.idx * BLOCK_N;
      const int n0 = nc * BLOCK_N;

      // Per-thread output row and validity.
      int orow_idx;
      bool valid;
      if constexpr (EPI == EPI_BIAS_RELU) {
        valid = (r < p.m_rows) && (l < p.l_store);
        orow_idx = wb * p.pitch_out + l;
      } else if constexpr (EPI == EPI_POOL4_BIAS_RELU) {
        valid = (r < p.m_rows) && ((l >> 2) < p.l_store);
        orow_idx = wb * p.pitch_out + (l >> 2);
      } else if constexpr (EPI == EPI_POOL2_BIAS_RELU) {
        valid = (r < p.m_rows) && ((l >> 1) < p.l_store);
        orow_idx = wb * p.pitch_out + (l >> 1);
      } else {
        valid = r < p.m_rows;
        orow_idx = r;
      }
      const size_t obase = (size_t)orow_idx * p.ld_out + n0;
      const int j = lane & 3;
      // EPI_CHROMATIN: this thread's window -> (shift, group member, q) -> output rows
      int cq = 0;
      size_t crow = 0, cscr = 0;
      if constexpr (EPI == EPI_CHROMATIN) {
        const ChromEpi& ce = p.chrom;
        const long long d = ce.first + r;
        int si, grp;
        if (ce.mode == 0) { cq = (int)(d & 3); const long long x = d >> 2; si = (int)(x / ce.ng); grp = (int)(x % ce.ng); }
        else if (ce.mode == 1) { cq = (int)(d & 1); const long long x = d >> 1; si = (int)(x / ce.ng); grp = (int)(x % ce.ng); }
        else if (ce.mode == 2) {
          const int cls = (int)(d / (2 * ce.ng)), x = (int)(d % (2 * ce.ng));
          grp = x >> 1; cq = 2 + (x & 1);
          si = valid ? ce.si_of_k[(x & 1) ? ce.S - 1 - cls : cls] : 0;
        } else { cq = (int)(d & 3); grp = (int)(d >> 2); si = ce.si_fixed; }
        const int t = cq & 1;
        const long long v = ce.row0 + ((ce.dst != nullptr && valid) ? ce.dst[grp] : grp);
        crow = (size_t)(si * ce.shift_stride + (ce.average ? 0 : t * ce.strand_stride) + v) * (size_t)ce.ld;
        cscr = ((size_t)((size_t)si * 2 + t) * ce.ng + grp) * 2002;
      }

      // One 16-column chunk: registers -> bias/activation -> global.
      auto process = [&](const uint32_t (&v)[16], int c, const float4 (&bq)[4]) {
        if constexpr (SPLIT_OUT && (EPI == EPI_BIAS_RELU || EPI == EPI_POOL4_BIAS_RELU || EPI == EPI_POOL2_BIAS_RELU)) {
          // split-operand mode: the next layer reads this activation as an fp16 pair (hi = rounded value, lo = the
          // rounding residual, out_lo_off elements further along the row); pooling runs on the fp32 values
          float y[16];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 b4 = bq[i];
            y[4 * i] = fmaxf(__uint_as_float(v[4 * i]) + b4.x, 0.f);
            y[4 * i + 1] = fmaxf(__uint_as_float(v[4 * i + 1]) + b4.y, 0.f);
            y[4 * i + 2] = fmaxf(__uint_as_float(v[4 * i + 2]) + b4.z, 0.f);
            y[4 * i + 3] = fmaxf(__uint_as_float(v[4 * i + 3]) + b4.w, 0.f);
          }
          if constexpr (EPI != EPI_BIAS_RELU) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              y[i] = fmaxf(y[i], __shfl_xor_sync(0xffffffffu, y[i], 1));
              if constexpr (EPI == EPI_POOL4_BIAS_RELU) y[i] = fmaxf(y[i], __shfl_xor_sync(0xffffffffu, y[i], 2));
            }
          }
          uint32_t hi[8], lo[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) split2<T>(y[2 * i], y[2 * i + 1], hi[i], lo[i]);
          T* o = reinterpret_cast<T*>(p.out) + obase + c * 16;
          T* ol = o + p.out_lo_off;
          if constexpr (EPI == EPI_BIAS_RELU) {
            if (valid) { st_global_v8(o, hi); st_global_v8(ol, lo); }
          } else if constexpr (EPI == EPI_POOL2_BIAS_RELU) {
            const bool odd = lane & 1;
            if (valid) {
              *reinterpret_cast<uint4*>(o + (odd ? 8 : 0)) =
                  odd ? make_uint4(hi[4], hi[5], hi[6], hi[7]) : make_uint4(hi[0], hi[1], hi[2], hi[3]);
              *reinterpret_cast<uint4*>(ol + (odd ? 8 : 0)) =
                  odd ? make_uint4(lo[4], lo[5], lo[6], lo[7]) : make_uint4(lo[0], lo[1], lo[2], lo[3]);
            }
          } else {
            uint32_t h0 = hi[0], h1 = hi[1], l0 = lo[0], l1 = lo[1];
            if (j == 1) { h0 = hi[2]; h1 = hi[3]; l0 = lo[2]; l1 = lo[3]; }
            if (j == 2) { h0 = hi[4]; h1 = hi[5]; l0 = lo[4]; l1 = lo[5]; }
            if (j == 3) { h0 = hi[6]; h1 = hi[7]; l0 = lo[6]; l1 = lo[7]; }
            if (valid) {
              *reinterpret_cast<uint2*>(o + 4 * j) = make_uint2(h0, h1);
              *reinterpret_cast<uint2*>(ol + 4 * j) = make_uint2(l0, l1);
            }
          }
        } else if constexpr (EPI == EPI_BIAS_RELU || EPI == EPI_POOL4_BIAS_RELU || EPI == EPI_POOL2_BIAS_RELU) {
          uint32_t pk[8];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 b4 = bq[i];
            pk[2 * i] = pack2<T>(fmaxf(__uint_as_float(v[4 * i]) + b4.x, 0.f),
                                 fmaxf(__uint_as_float(v[4 * i + 1]) + b4.y, 0.f));
            pk[2 * i + 1] = pack2<T>(fmaxf(__uint_as_float(v[4 * i + 2]) + b4.z, 0.f),
                                     fmaxf(__uint_as_float(v[4 * i + 3]) + b4.w, 0.f));
          }
          T* o = reinterpret_cast<T*>(p.out) + obase + c * 16;
          if constexpr (EPI == EPI_BIAS_RELU) {
            // one 256-bit store per lane (STG.256): each warp store touches 32 rows, so halving the instruction
            // count halves the LSU transactions that bound the conv1 epilogue
            if (valid) st_global_v8(o, pk);
          } else if constexpr (EPI == EPI_POOL2_BIAS_RELU) {
            // max over the row pair held by lanes 2g, 2g+1; the even lane stores columns 0..7, the odd one 8..15
#pragma unroll
            for (int i = 0; i < 8; ++i) pk[i] = max2<T>(pk[i], __shfl_xor_sync(0xffffffffu, pk[i], 1));
            const bool odd = lane & 1;
            const uint4 w = odd ? make_uint4(pk[4], pk[5], pk[6], pk[7]) : make_uint4(pk[0], pk[1], pk[2], pk[3]);
            if (valid) *reinterpret_cast<uint4*>(o + (odd ? 8 : 0)) = w;
          } else {
            // MaxPool over the 4 rows held by lanes 4g..4g+3 (rounding is monotonic, so max after
            // rounding equals rounding after max); lane j of the group stores columns 4j..4j+3.
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              pk[i] = max2<T>(pk[i], __shfl_xor_sync(0xffffffffu, pk[i], 1));
              pk[i] = max2<T>(pk[i], __shfl_xor_sync(0xffffffffu, pk[i], 2));
            }
            uint32_t w0 = pk[0], w1 = pk[1];
            if (j == 1) { w0 = pk[2]; w1 = pk[3]; }
            if (j == 2) { w0 = pk[4]; w1 = pk[5]; }
            if (j == 3) { w0 = pk[6]; w1 = pk[7]; }
            if (valid) *reinterpret_cast<uint2*>(o + 4 * j) = make_uint2(w0, w1);
          }
        } else if constexpr (EPI == EPI_PARTIAL_F32) {
          float4* o = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + (size_t)ks * p.split_stride +
                                                obase + c * 16);
          if (valid) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              o[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                 __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
          }
        } else if constexpr (EPI == EPI_CHROMATIN) {
          const ChromEpi& ce = p.chrom;
          const int col0 = n0 + c * 16;
#pragma unroll
          for (int i = 0; i < 16; i += 2) {
            const bool in = col0 + i < p.n_store;   // warp-uniform (n_store is even)
            float2 pr = make_float2(0.f, 0.f);
            if (in) {
              const float z0 = __uint_as_float(v[i]) + sbias[col0 + i];
              const float z1 = __uint_as_float(v[i + 1]) + sbias[col0 + i + 1];
              pr = make_float2(1.f / (1.f + __expf(-z0)), 1.f / (1.f + __expf(-z1)));
            }
            const size_t o = crow + col0 + i;
            // the same strand's ref probability: two lanes down in a quad, or what the ref launch stored
            float2 rf;
            if (ce.mode == 2) {
              rf = make_float2(0.f, 0.f);
              if (in && valid) rf = *reinterpret_cast<const float2*>((ce.average ? ce.ref_scratch + cscr : ce.ref + crow) + col0 + i);
            } else {
              rf.x = __shfl_xor_sync(0xffffffffu, pr.x, 2);
              rf.y = __shfl_xor_sync(0xffffffffu, pr.y, 2);
            }
            const float2 df = make_float2(pr.x - rf.x, pr.y - rf.y);   // meaningful on alt rows
            if (!ce.average) {
              if (in && valid) {
                if (cq < 2) *reinterpret_cast<float2*>(ce.ref + o) = pr;
                else {
                  *reinterpret_cast<float2*>(ce.alt + o) = pr;
                  *reinterpret_cast<float2*>(ce.diff + o) = df;
                }
              }
            } else if (ce.mode == 2) {
              if (in && valid) {
                atomicAdd(ce.alt + o, pr.x / 2.0f);
                atomicAdd(ce.alt + o + 1, pr.y / 2.0f);
                atomicAdd(ce.diff + o, df.x / 2.0f);
                atomicAdd(ce.diff + o + 1, df.y / 2.0f);
              }
            } else {
              // forward / reverse-complement rows of one shift are neighbouring lanes
              const float2 po = make_float2(__shfl_xor_sync(0xffffffffu, pr.x, 1), __shfl_xor_sync(0xffffffffu, pr.y, 1));
              const float2 dn = make_float2(__shfl_xor_sync(0xffffffffu, df.x, 1), __shfl_xor_sync(0xffffffffu, df.y, 1));
              if (in && valid) {
                if (ce.mode == 1) *reinterpret_cast<float2*>(ce.ref_scratch + cscr + col0 + i) = pr;
                if ((cq & 1) == 0) {   // the forward lane writes the averages: (fwd + rc) / 2.0
                  const float2 avg = make_float2((pr.x + po.x) / 2.0f, (pr.y + po.y) / 2.0f);
                  if (cq < 2) *reinterpret_cast<float2*>(ce.ref + o) = avg;
                  else {
                    *reinterpret_cast<float2*>(ce.alt + o) = avg;
                    *reinterpret_cast<float2*>(ce.diff + o) = make_float2((df.x + dn.x) / 2.0f, (df.y + dn.y) / 2.0f);
                  }
                }
              }
            }
          }
        } else {  // EPI_BIAS_SIGMOID
          // rows are ld_out = 2002 floats apart (8-byte aligned) and n_store is even: float2 stores, half the
          // store instructions of the scalar form (each one touches 32 rows, so the LSU sees 32 sectors either way)
          float2* o = reinterpret_cast<float2*>(reinterpret_cast<float*>(p.out) + obase + c * 16);
          if (valid) {
#pragma unroll
            for (int i = 0; i < 16; i += 2) {
              if (n0 + c * 16 + i < p.n_store) {
                const float z0 = __uint_as_float(v[i]) + sbias[n0 + c * 16 + i];
                const float z1 = __uint_as_float(v[i + 1]) + sbias[n0 + c * 16 + i + 1];
                o[i >> 1] = make_float2(1.f / (1.f + __expf(-z0)), 1.f / (1.f + __expf(-z1)));
              }
            }
          }
        }
      };

      // Software-pipelined TMEM reads: the load of chunk c+1 / c+2 is in flight while chunk c is processed.
      constexpr int NC = BLOCK_N / 16;
      uint32_t va[16], vb[16];
      float4 qa[4], qb[4];
      // bias of a chunk is read from shared memory BEFORE the TMEM wait so the two latencies overlap (ncu:
      // short-scoreboard stalls on these loads were 52 % of the conv1 epilogue's cycles)
      auto load_bias = [&](float4 (&q)[4], int c) {
        if constexpr (EPI == EPI_BIAS_RELU || EPI == EPI_POOL4_BIAS_RELU || EPI == EPI_POOL2_BIAS_RELU) {
          const float4* bp = reinterpret_cast<const float4*>(sbias + n0 + c * 16);
#pragma unroll
          for (int i = 0; i < 4; ++i) q[i] = bp[i];
        }
      };
      if (!(p.debug_flags & 4)) {
      tmem_ld16(t0, va);
#pragma unroll 1
      for (int c = 0; c < NC; c += 2) {
        if (c + 1 < NC) tmem_ld16(t0 + (c + 1) * 16, vb);
        load_bias(qa, c);
        if (c + 1 < NC) load_bias(qb, c + 1);
        tmem_ld_wait();
        process(va, c, qa);
        if (c + 2 < NC) tmem_ld16(t0 + (c + 2) * 16, va);
        if (c + 1 < NC) process(vb, c + 1, qb);
      }
      }
      // Release the TMEM buffer to the MMA thread (leader CTA's barrier).
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(CG == 1 ? tempty(rt.idx) : mapa(tempty(rt.idx), 0));
      rt.advance<NBUF>();
    }
  }

  // ------------------------------------------------------------------ teardown
  __syncwarp();
  tc_fence_before();
  if constexpr (CG == 1) __syncthreads(); else cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<CG>(tmem_base, 512);
  }
}

}  // namespace bx
