// Per-variant decay transform fused with gblinear scoring as ONE tcgen05 GEMM (reference:
// compute_effects, predict.py:87-160 — float64 [n, 20020] feature matrices built from np.tile / np.repeat temporaries,
// then xgboost Booster.predict = base_score + bias + w . x per model).
//
//   out[n, m] = offset[m] + sum_{k<10} sum_{f<2002} feat[n, k, f] * W[m, k*2002 + f],
//   feat[n, k, f] = sum_{j<9} w[n, j, k] * x[j, n, f]          (w from the variant's distance and strand)
//
// The [256 variants x 64] feature tiles are generated in shared memory straight from the nine shift rows of x and
// never exist in HBM: x is read once (72 KB per variant and tensor — the kernel's HBM roofline), the weight matrix
// streams from L2 (18 MB for 224 models).  The GEMM's K axis is a permutation of (k, f) chosen so that what one
// thread computes from one 32-byte piece of each shift row is one swizzled row of one operand stage: per
// "super-block" of 32 features there are five K-blocks of 64 columns,
//   K-block c < 4 :  column k*8 + i          = feature 32*sb + 8*c + i,  basis k < 8
//   K-block 4     :  column (k-8)*32 + 8*c+i = feature 32*sb + 8*c + i,  basis k = 8, 9
// and the model weights are permuted the same way once (score_prep_weights_kernel).  Features and weights are fp16
// (hi, lo) pairs and every product is Ah*Wh + Ah*Wl + Al*Wh in fp32 (three MMAs per K step): REF and ALT scores keep
// fp32-level accuracy, which SED = ALT - REF (predict.py:264) needs because the two scores differ in their 4th digit.
// The tensor core's fp32 accumulation truncates, so a chain of n MMAs into one accumulator loses ~n * 2^-24 of its
// magnitude, always towards zero (measured: 7.9e-5 relative with all 3780 MMAs of a tile in one accumulator).  The two
// correction products therefore go to a second TMEM accumulator — 2^-11 the magnitude, so their truncation is
// negligible — and the epilogue adds the two in fp32: the main chain is 1260 MMAs long instead of 3780.
//
// CTA pair (cta_group::2): M = 256 variants (128 per CTA), N = models padded to a multiple of 16 (each CTA stages half
// of the weight rows).  Warps 0-7: feature generators (four lanes per variant row, each lane one feature pair of every
// 8-feature chunk and all ten bases: coalesced 32-byte row pieces in, 4-byte swizzled stores out; a first version with
// one thread per row was bound by L1 wavefronts — 32 sectors per load instruction — at 5.9 us per K-block) and, at
// the end of a tile, the epilogue; warp 8: TMA producer of the weight tiles; warp 9: MMA issuer (leader CTA) and TMEM
// allocator.
#pragma once
#include "gemm_tc.cuh"

namespace bx {

constexpr int kScSuper = 63;                 // ceil(2002 / 32) feature super-blocks
constexpr int kScKBlocks = kScSuper * 5;     // 315 K-blocks of 64 columns
constexpr int kScK = kScKBlocks * 64;        // 20160 permuted K columns (2016 features x 10 bases)
constexpr int kScShifts = 9;                 // chromatin.py:243 with the default --maxshift 800
constexpr int kScATile = 128 * 128;          // one [128 x 64] fp16 operand tile
constexpr int kScThreads = 320;
#ifndef BX_SC_XPF
#define BX_SC_XPF 0   // experiment: L2 prefetch of the x chunk this many chunks ahead of its TMA load
#endif

struct ScoreParams {
  const float* x;          // [9, n, ld_x]: row pitch ld_x >= 2002 floats (columns >= 2002 are never read)
  long long ld_x;
  const double* dist;      // [n] variant - TSS (predict.py:241)
  const int8_t* strand;    // [n] +1 / -1
  const int32_t* shifts;   // [9]
  const float* offset;     // [m] base_score + bias
  int skip_zero_steps;     // score_tma_kernel: skip the K steps that are zero for every row of a tile (score_tile_class)
  float* out;              // [n, ld_out]: this launch writes columns [0, m)
  int n, m, n_pad, tiles;  // m models in this launch, n_pad: padded to a multiple of 16 (<= 256); tiles = ceil(n / 256)
  int ld_out;
  uint32_t idesc;
  int sb_stages;           // weight-ring depth (2 or 3, by shared-memory budget)
};

// Permuted K column -> (basis k, feature f); f >= 2002 is padding.
__host__ __device__ inline void score_col_to_kf(int col, int& k, int& f) {
  const int kb = col >> 6, c = kb % 5, sb = kb / 5, w = col & 63;
  if (c < 4) { k = w >> 3; f = 32 * sb + 8 * c + (w & 7); }
  else { k = 8 + (w >> 5); f = 32 * sb + (w & 31); }
}

// Model weights fp32 [m, 20020] -> fp16 pairs [n_pad, 2 * kScK]: high parts in columns [0, kScK), residuals after.
__global__ void score_prep_weights_kernel(const float* __restrict__ w, int m, int n_pad, __half* __restrict__ out) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)n_pad * kScK) return;
  const int row = (int)(idx / kScK), col = (int)(idx % kScK);
  int k, f;
  score_col_to_kf(col, k, f);
  const float v = (row < m && f < 2002) ? w[(size_t)row * 20020 + (size_t)k * 2002 + f] : 0.f;
  const __half h = __float2half_rn(v);
  out[(size_t)row * 2 * kScK + col] = h;
  out[(size_t)row * 2 * kScK + kScK + col] = __float2half_rn(v - __half2float(h));
}

// This kernel loads x through registers and serves rows that are NOT on 16-byte boundaries (a caller's dense
// [9, n, 2002] array: 8008-byte rows); score_tma_kernel below is the one the drop-in's own path runs.  Memory behaviour
// (profiles/r02_ncu_score_tc.txt): a tile walks 256 variants x 9 shift rows in parallel, 32 bytes of each per chunk;
// with 8008-byte rows most pieces straddle two DRAM sectors (1.5 TB/s, long-scoreboard stalls: 3.9 ms per 75,776
// variants; 2.8 ms once the rows are 2016 floats apart).  L2::128B / L2::256B fill hints on the loads made it slower
// (4.6 / 5.2 ms); an L2 prefetch of each row's next 128 bytes one super-block ahead plus a one-chunk register double
// buffer is the best of what 168 registers allow.
__global__ void __launch_bounds__(kScThreads, 1)
score_tc_kernel(const __grid_constant__ CUtensorMap tmW, const ScoreParams p) {
  constexpr int SA = 2, SK = 2;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t b_half = (uint32_t)(p.n_pad / 2) * 128u;          // one weight tile (hi or lo) of this CTA
  const uint32_t sA = base;                                        // SA x (hi, lo)
  const uint32_t sK = sA + SA * 2 * kScATile;                      // SK x (hi, lo): the k = 8, 9 K-blocks
  const uint32_t sB = sK + SK * 2 * kScATile;                      // sb_stages x (hi, lo)
  const uint32_t bars = sB + (uint32_t)p.sb_stages * 2 * b_half;
  auto fullA = [&](uint32_t i) { return bars + 8u * i; };
  auto emptyA = [&](uint32_t i) { return bars + 8u * (2 + i); };
  auto fullK = [&](uint32_t i) { return bars + 8u * (4 + i); };
  auto emptyK = [&](uint32_t i) { return bars + 8u * (6 + i); };
  auto fullB = [&](uint32_t i) { return bars + 8u * (8 + i); };
  auto emptyB = [&](uint32_t i) { return bars + 8u * (12 + i); };
  const uint32_t tfull = bars + 8u * 16, tempty = bars + 8u * 17, tmem_slot = bars + 8u * 18;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int num_clusters = gridDim.x / 2, cluster_id = blockIdx.x / 2;

  if (warp == 8 && lane == 0) {
    prefetch_tensormap(&tmW);
    for (int i = 0; i < 2; ++i) {
      mbar_init(fullA(i), 16); mbar_init(emptyA(i), 1);     // 8 generator warps in each CTA of the pair
      mbar_init(fullK(i), 16); mbar_init(emptyK(i), 1);
    }
    for (int i = 0; i < 4; ++i) { mbar_init(fullB(i), 1); mbar_init(emptyB(i), 1); }
    mbar_init(tfull, 1);
    mbar_init(tempty, 16);
    fence_barrier_init();
  }
  if (warp == 9) tmem_alloc<2>(tmem_slot, 512);   // columns [0, n_pad): Ah*Wh; [256, 256 + n_pad): Ah*Wl + Al*Wh
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  if (warp < 8) {
    // ------------------------------------------------------------------ feature generators + epilogue
    // Four lanes share a variant row: lane (rw, sub) of warp w handles rows w*16 + {0, 8} + rw and, in every chunk of
    // eight features, the feature pair 2*sub, 2*sub + 1 — so a warp's loads cover eight rows x 32 contiguous bytes
    // (full sectors), and the four lanes' 4-byte stores assemble each 16-byte swizzled piece of the operand row.
    const int sub = lane & 3, rw = lane >> 2;
    RingState ra, rk;
    uint32_t tphase = 0;
    for (int t = cluster_id; t < p.tiles; t += num_clusters) {
      // y = exp(-0.01 * floor(|pos| / 200)) per shift (fp64 like numpy, predict.py:87-109) and which side of the TSS
      // each shift falls on, for this lane's two rows.  The five decay rates 0.01, 0.02, 0.05, 0.1, 0.2 are the powers
      // y, y^2, y^5, y^10, y^20, formed by five multiplications where they are used: keeping all 90 weights of the two
      // rows in registers left no room to hold the next chunk's loads, and the kernel then ran at the load latency.
      float y[2][kScShifts];
      uint32_t negm[2], posm[2];
      const float* xrow[2];
      bool valid[2];
#pragma unroll
      for (int ps = 0; ps < 2; ++ps) {
        const long long row = (long long)t * 256 + rank * 128 + warp * 16 + ps * 8 + rw;
        valid[ps] = row < p.n;
        xrow[ps] = p.x + (size_t)(valid[ps] ? row : 0) * (size_t)p.ld_x;
        const double d = valid[ps] ? p.dist[row] : 0.0;
        const double sgn = (valid[ps] && p.strand[row] > 0) ? 1.0 : -1.0;
        negm[ps] = posm[ps] = 0;
#pragma unroll
        for (int j = 0; j < kScShifts; ++j) {
          const double pos = d * sgn + (double)p.shifts[j] * sgn;
          const double b = floor(fabs(pos / 200.0));
          if (valid[ps] && pos <= 0.0) negm[ps] |= 1u << j;
          if (valid[ps] && pos >= 0.0) posm[ps] |= 1u << j;
          y[ps][j] = (float)exp(-0.01 * b);
        }
      }
      const size_t shift_stride = (size_t)p.n * (size_t)p.ld_x;
      auto load_chunk = [&](int f0, float2 (&xv)[2][kScShifts]) {
#pragma unroll
        for (int ps = 0; ps < 2; ++ps)
#pragma unroll
          for (int j = 0; j < kScShifts; ++j)
            xv[ps][j] = (valid[ps] && f0 + 2 * sub < 2002)
                            ? __ldg(reinterpret_cast<const float2*>(xrow[ps] + j * shift_stride + f0) + sub)
                            : make_float2(0.f, 0.f);
      };
      auto prefetch_super = [&](int sb) {
        if (32 * sb >= 2002) return;
#pragma unroll
        for (int ps = 0; ps < 2; ++ps)
#pragma unroll
          for (int j = 0; j < kScShifts; ++j)
            if ((j & 3) == sub && valid[ps])
              asm volatile("prefetch.global.L2 [%0];" ::"l"(xrow[ps] + j * shift_stride + 32 * sb));
      };
      prefetch_super(0);
      prefetch_super(1);
      float2 xn[2][kScShifts];
      load_chunk(0, xn);
      for (int sb = 0; sb < kScSuper; ++sb) {
        prefetch_super(sb + 2);
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          float2 xv[2][kScShifts];
#pragma unroll
          for (int ps = 0; ps < 2; ++ps)
#pragma unroll
            for (int j = 0; j < kScShifts; ++j) xv[ps][j] = xn[ps][j];
          const int fnext = 32 * sb + 8 * c + 8;        // the next chunk's loads fly during this chunk's arithmetic
          if (fnext < 32 * kScSuper) load_chunk(fnext, xn);
          mbar_wait(emptyA(ra.idx), ra.phase ^ 1, 11);
          if (c == 0) mbar_wait(emptyK(rk.idx), rk.phase ^ 1, 12);
#pragma unroll
          for (int ps = 0; ps < 2; ++ps) {
            float2 acc[10];
#pragma unroll
            for (int k = 0; k < 10; ++k) acc[k] = make_float2(0.f, 0.f);
#pragma unroll
            for (int j = 0; j < kScShifts; ++j) {
              const bool ng = (negm[ps] >> j) & 1u, ps_ = (posm[ps] >> j) & 1u;
              const float2 xm = ng ? xv[ps][j] : make_float2(0.f, 0.f);
              const float2 xp = ps_ ? xv[ps][j] : make_float2(0.f, 0.f);
              const float y1 = y[ps][j], y2 = y1 * y1, y4 = y2 * y2, y5 = y4 * y1, y10 = y5 * y5, y20 = y10 * y10;
              const float e[5] = {y1, y2, y5, y10, y20};
#pragma unroll
              for (int q = 0; q < 5; ++q) {
                acc[q].x = fmaf(e[q], xm.x, acc[q].x);
                acc[q].y = fmaf(e[q], xm.y, acc[q].y);
                acc[5 + q].x = fmaf(e[q], xp.x, acc[5 + q].x);
                acc[5 + q].y = fmaf(e[q], xp.y, acc[5 + q].y);
              }
            }
            const int r = warp * 16 + ps * 8 + rw;
            const uint32_t swz = (uint32_t)(r & 7);
            uint8_t* st = smem_raw + (sA - raw) + ra.idx * 2 * kScATile + r * 128 + sub * 4;
            uint8_t* sk = smem_raw + (sK - raw) + rk.idx * 2 * kScATile + r * 128 + sub * 4;
#pragma unroll
            for (int k = 0; k < 10; ++k) {
              uint32_t hi, lo;
              split2<__half>(acc[k].x, acc[k].y, hi, lo);
              // K-block 5*sb + c holds bases 0..7 (piece k); K-block 5*sb + 4 bases 8, 9 (piece (k - 8) * 4 + c)
              uint8_t* dst = k < 8 ? st + (((uint32_t)k ^ swz) << 4) : sk + (((uint32_t)((k - 8) * 4 + c) ^ swz) << 4);
              *reinterpret_cast<uint32_t*>(dst) = hi;
              *reinterpret_cast<uint32_t*>(dst + kScATile) = lo;
            }
          }
          fence_proxy_async_smem();   // generic-proxy stores -> visible to the tensor core's async-proxy reads
          __syncwarp();
          if (lane == 0) {
            mbar_arrive_remote(mapa(fullA(ra.idx), 0));
            if (c == 3) mbar_arrive_remote(mapa(fullK(rk.idx), 0));
          }
          __syncwarp();
          ra.advance<SA>();
        }
        rk.advance<SK>();
      }
      // ---- epilogue: TMEM -> + offset -> out[row, model]; warp handles its lane quarter and every other 16-column chunk
      mbar_wait(tfull, tphase, 13);
      tphase ^= 1;
      tc_fence_after();
      {
        const int q = warp & 3;
        const long long orow = (long long)t * 256 + rank * 128 + q * 32 + lane;
        const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16);
        for (int c16 = warp >> 2; c16 < p.n_pad / 16; c16 += 2) {
          uint32_t v[16], u[16];
          tmem_ld16(t0 + c16 * 16, v);
          tmem_ld16(t0 + 256 + c16 * 16, u);
          tmem_ld_wait();
          if (orow < p.n) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int col = c16 * 16 + i;
              if (col < p.m)
                p.out[(size_t)orow * p.ld_out + col] = (__uint_as_float(v[i]) + __uint_as_float(u[i])) + __ldg(p.offset + col);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(mapa(tempty, 0));
    }
  } else if (warp == 8) {
    // ------------------------------------------------------------------ weight tiles by TMA (both CTAs, own half of N)
    RingState rb;
    const int row0 = (int)rank * (p.n_pad / 2);
    for (int t = cluster_id; t < p.tiles; t += num_clusters) {
      for (int kb = 0; kb < kScKBlocks; ++kb) {
        mbar_wait(emptyB(rb.idx), rb.phase ^ 1, 14);
        if (elect_one()) {
          if (rank == 0) mbar_arrive_expect_tx(fullB(rb.idx), 4 * b_half);
          const uint32_t bar = mapa(fullB(rb.idx), 0);
          const uint32_t dst = sB + rb.idx * 2 * b_half;
          tma_load_2d_pair(dst, &tmW, bar, kb * 64, row0);
          tma_load_2d_pair(dst + b_half, &tmW, bar, kScK + kb * 64, row0);
        }
        __syncwarp();
        if (++rb.idx == (uint32_t)p.sb_stages) { rb.idx = 0; rb.phase ^= 1; }
      }
    }
  } else if (rank == 0) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA)
    RingState ra, rk, rb;
    uint32_t tphase = 0;
    for (int t = cluster_id; t < p.tiles; t += num_clusters) {
      mbar_wait(tempty, tphase ^ 1, 15);   // the previous tile's epilogue has read the accumulator
      tphase ^= 1;
      tc_fence_after();
      uint32_t accum = 0;
      for (int kb = 0; kb < kScKBlocks; ++kb) {
        const bool k89 = (kb % 5) == 4;
        uint32_t a_hi;
        if (k89) { mbar_wait(fullK(rk.idx), rk.phase, 16); a_hi = sK + rk.idx * 2 * kScATile; }
        else { mbar_wait(fullA(ra.idx), ra.phase, 17); a_hi = sA + ra.idx * 2 * kScATile; }
        mbar_wait(fullB(rb.idx), rb.phase, 18);
        tc_fence_after();
        const uint32_t b_hi = sB + rb.idx * 2 * b_half;
        if (elect_one()) {
          const uint32_t ah = sw128_desc_lo(a_hi), al = sw128_desc_lo(a_hi + kScATile);
          const uint32_t bh = sw128_desc_lo(b_hi), bl = sw128_desc_lo(b_hi + b_half);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            umma_f16<2>(tmem_base, sw128_desc(ah + 2 * ks), sw128_desc(bh + 2 * ks), p.idesc, accum);
            umma_f16<2>(tmem_base + 256, sw128_desc(ah + 2 * ks), sw128_desc(bl + 2 * ks), p.idesc, accum);
            umma_f16<2>(tmem_base + 256, sw128_desc(al + 2 * ks), sw128_desc(bh + 2 * ks), p.idesc, 1u);
            accum = 1;
          }
          umma_commit<2>(emptyB(rb.idx));
          umma_commit<2>(k89 ? emptyK(rk.idx) : emptyA(ra.idx));
        }
        __syncwarp();
        accum = 1;
        if (++rb.idx == (uint32_t)p.sb_stages) { rb.idx = 0; rb.phase ^= 1; }
        if (k89) rk.advance<SK>(); else ra.advance<SA>();
      }
      if (elect_one()) umma_commit<2>(tfull);
      __syncwarp();
    }
  }

  __syncwarp();
  tc_fence_before();
  cluster_sync_all();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc<2>(tmem_base, 512);
  }
}


// Side of the TSS of every row of a 256-variant tile (the masks of the generators' prologue, same expressions): 1 = every
// row has all nine shifts upstream (or is padding), 2 = all downstream, 0 = anything else.  For a one-sided row the five
// bases of the other side are exact zeros, so in a tile of class 1 or 2 whole K steps of the operand are zero for every
// row: score_tma_kernel neither stores, loads nor multiplies them (adding an MMA of zeros changes no bit).  Random row
// orders give class 0 throughout; callers that order their rows by side (transform.side_order: predict does, for free,
// wherever it gathers rows anyway) get classes 1 and 2 for all but a few tiles.  Every warp that needs the class
// (generators, weight producer, MMA issuer, in both CTAs) evaluates it itself from dist / strand — eight rows per lane,
// about a microsecond per 450 us tile — so the roles need no extra synchronisation and run ahead of each other as before.
__device__ __forceinline__ int score_tile_class(const ScoreParams& p, int t, int lane) {
  bool neg = false, posi = false;
  for (int i = 0; i < 8; ++i) {
    const long long row = (long long)t * 256 + i * 32 + lane;
    if (row >= p.n) continue;
    const double d = p.dist[row];
    const double sgn = p.strand[row] > 0 ? 1.0 : -1.0;
#pragma unroll
    for (int j = 0; j < kScShifts; ++j) {
      const double pos = d * sgn + (double)p.shifts[j] * sgn;
      if (pos <= 0.0) neg = true;
      if (pos >= 0.0) posi = true;
    }
  }
  const bool any_pos = __any_sync(0xffffffffu, posi), any_neg = __any_sync(0xffffffffu, neg);
  return !any_pos ? 1 : (!any_neg ? 2 : 0);
}

// ====================================================================================================================
// The same GEMM with x staged through shared memory by TMA (rows on 16-byte boundaries: row pitch a multiple of four
// floats — what the drop-in's own producers write, transform.padded_rows).  A producer lane copies each chunk — the box
// [8 features x 128 rows x 9 shifts] = 36 KB per CTA — with one cp.async.bulk.tensor two chunks ahead of its use, at no
// register or issue cost to the generators (tools/tma_box_bench.cu: one SM streams such boxes at 5.7 TB/s chip-wide with
// two in flight, so x delivery is not the limit); rows past n and features past 2002 arrive as zeros (TMA bounds).  The
// shared-memory budget (227 KB) is met by a one-stage ring for the k = 8, 9 K-blocks — chunk 0 keeps its four pieces of
// that block in registers for one chunk, so nobody waits for the stage — and a two-stage weight ring.
//   * Sixteen generator warps, one 8-row set each (four warps per scheduler instead of two, 94 registers).
//   * Packed arithmetic (FFMA2, scalar powers broadcast), y and y^5 kept, y^2, y^10, y^20 formed where used.
//   * Rows whose shifts straddle the TSS (|distance| <= 800, ~2 % of variants) need sums for both sides.  Pass A sums
//     the larger side of such a row — for every other row simply all nine shifts, unmasked when the set has no such
//     row — and pass B the smaller side over just the shifts the set has there.  Terms left out are exact zeros of the
//     two-sided formulation, so a row's bits never depend on the rows it shares a warp with.
// Operand layout, weight permutation, MMA order and therefore every bit of the result are those of the kernel above
// (tests/test_gpu_transform.py compares the two).  What bounds it now (kernel 5;
// profiles/r02_ncu_score_tma.txt, r02_score_trace.txt): shared-memory bandwidth.  Per chunk and CTA the tensor core reads
// 112 KB of operands (three MMAs per K step), TMA writes 35 KB of weights and 36 KB of x, the generators read those 36 KB
// and write 32 KB of operand tiles: 251 KB per 2,460 cycles = 102 of the SM's 128 B/clk — at the ~1.35 GHz the board's
// power cap leaves once the kernel runs for more than a burst.
constexpr int kScXStage = kScShifts * 128 * 32;   // one chunk of one CTA: [9 shifts][128 rows][8 floats]
#ifndef BX_SC_XSTAGES
#define BX_SC_XSTAGES 2
#endif
constexpr int kScXStages = BX_SC_XSTAGES;
#ifndef BX_SC_SETS
#define BX_SC_SETS 1
#endif
constexpr int kScSets = BX_SC_SETS;               // 8-row sets per generator warp: 1 -> sixteen generator warps, 2 -> eight
constexpr int kScGenWarps = 16 / kScSets;
constexpr int kScTmaThreads = (kScGenWarps + 3) * 32;   // + W producer, MMA issuer, x producer

__device__ __forceinline__ uint64_t pack2f(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ float2 unpack2f(uint64_t v) {
  float2 r;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
  return r;
}
// Packed fp32 pair FMA (FFMA2 on sm_100): each half is the scalar fma.rn.
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
// The value, opaque to the optimiser: pins what is computed from it (the powers of y) to the place it is used —
// hoisted out of the chunk loop they would cost 180 registers.
__device__ __forceinline__ float pinned(float v) {
  asm volatile("" : "+f"(v));
  return v;
}

#ifdef BX_SC_TRACE   // diagnostic build: cycle stamps of one CTA's pipeline events (tools/score_trace.py)
__device__ long long g_sc_trace[4][512][8];
#define SC_TRACE(role, idx, ev) do { if (blockIdx.x == 0 && (idx) < 512) g_sc_trace[role][idx][ev] = clock64(); } while (0)
#else
#define SC_TRACE(role, idx, ev) do { } while (0)
#endif

__global__ void __launch_bounds__(kScTmaThreads, 1)
score_tma_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX, const ScoreParams p) {
  constexpr int SA = 2;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t b_half = (uint32_t)(p.n_pad / 2) * 128u;          // one weight tile (hi or lo) of this CTA
  const uint32_t sX = base;                                        // kScXStages chunks of x
  const uint32_t sA = sX + kScXStages * kScXStage;                 // SA x (hi, lo)
  const uint32_t sK = sA + SA * 2 * kScATile;                      // (hi, lo): the k = 8, 9 K-block of a super-block
  const uint32_t sB = sK + 2 * kScATile;                           // sb_stages x (hi, lo)
  const uint32_t bars = sB + (uint32_t)p.sb_stages * 2 * b_half;
  auto fullA = [&](uint32_t i) { return bars + 8u * i; };
  auto emptyA = [&](uint32_t i) { return bars + 8u * (2 + i); };
  const uint32_t fullK = bars + 8u * 4, emptyK = bars + 8u * 5;
  auto fullX = [&](uint32_t i) { return bars + 8u * (6 + i); };
  auto emptyX = [&](uint32_t i) { return bars + 8u * (10 + i); };
  auto fullB = [&](uint32_t i) { return bars + 8u * (14 + i); };
  auto emptyB = [&](uint32_t i) { return bars + 8u * (18 + i); };
  const uint32_t tfull = bars + 8u * 22, tempty = bars + 8u * 23, tmem_slot = bars + 8u * 24;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int num_clusters = gridDim.x / 2, cluster_id = blockIdx.x / 2;

  if (warp == kScGenWarps && lane == 0) {
    prefetch_tensormap(&tmW);
    prefetch_tensormap(&tmX);
    for (int i = 0; i < 2; ++i) {
      mbar_init(fullA(i), 2 * kScGenWarps); mbar_init(emptyA(i), 1);     // the generator warps of both CTAs
    }
    for (int i = 0; i < kScXStages; ++i) { mbar_init(fullX(i), 1); mbar_init(emptyX(i), kScGenWarps); }   // per CTA: its own 128 rows of x
    mbar_init(fullK, 2 * kScGenWarps); mbar_init(emptyK, 1);
    for (int i = 0; i < 4; ++i) { mbar_init(fullB(i), 1); mbar_init(emptyB(i), 1); }
    mbar_init(tfull, 1);
    mbar_init(tempty, 2 * kScGenWarps);
    fence_barrier_init();
  }
  if (warp == kScGenWarps + 1) tmem_alloc<2>(tmem_slot, 512);   // columns [0, n_pad): Ah*Wh; [256, 256 + n_pad): Ah*Wl + Al*Wh
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  if (warp < kScGenWarps) {
    // ------------------------------------------------------------------ feature generators + epilogue
    // Lane (rw, sub) of warp w: rows w * 8 * kScSets + ps * 8 + rw of this CTA (ps < kScSets), feature pair 2*sub,
    // 2*sub + 1 of every chunk.
    const int sub = lane & 3, rw = lane >> 2;
    RingState ra, rx;
    uint32_t kphase = 0, tphase = 0;
    for (int t = cluster_id; t < p.tiles; t += num_clusters) {
      // y = exp(-0.01 * floor(|pos| / 200)) per shift (fp64 like numpy, predict.py:87-109); the five decay rates are
      // its powers y, y^2, y^5, y^10, y^20.  y and y^5 are kept, the others formed where they are used (same products
      // in the same order as the kernel above: y2 = y*y, y4 = y2*y2, y5 = y4*y, y10 = y5*y5, y20 = y10*y10).
      float y[kScSets][kScShifts], y5[kScSets][kScShifts];
      uint32_t negm[kScSets], posm[kScSets];
#pragma unroll
      for (int ps = 0; ps < kScSets; ++ps) {
        const long long row = (long long)t * 256 + rank * 128 + warp * (8 * kScSets) + ps * 8 + rw;
        const bool valid = row < p.n;
        const double d = valid ? p.dist[row] : 0.0;
        const double sgn = (valid && p.strand[row] > 0) ? 1.0 : -1.0;
        negm[ps] = posm[ps] = 0;
#pragma unroll
        for (int j = 0; j < kScShifts; ++j) {
          const double pos = d * sgn + (double)p.shifts[j] * sgn;
          const double b = floor(fabs(pos / 200.0));
          if (valid && pos <= 0.0) negm[ps] |= 1u << j;
          if (valid && pos >= 0.0) posm[ps] |= 1u << j;
          y[ps][j] = (float)exp(-0.01 * b);
          const float y2 = y[ps][j] * y[ps][j], y4 = y2 * y2;
          y5[ps][j] = y4 * y[ps][j];
        }
      }
      // A row whose shifts straddle the TSS (|distance| <= 800, ~2 % of variants) needs two sets of sums.  Pass A sums
      // the larger side of such a row — for every other row simply all nine shifts — and pass B the smaller side, over
      // just the shifts some row of this 8-row set has there (warp-uniform skip).  A set without such rows runs pass A
      // unmasked and nothing else.  Terms left out are exact zeros in the two-sided formulation, so the sums are the
      // same bits whichever way a row is processed and never depend on the rows it shares a warp with.
      uint32_t maskA[kScSets], maskB[kScSets], unionB[kScSets];
      bool up[kScSets], mixed[kScSets];     // up: pass A's sums belong to the upstream bases 0..4
#pragma unroll
      for (int ps = 0; ps < kScSets; ++ps) {
        const bool two = negm[ps] != 0 && posm[ps] != 0;
        const bool neg_major = __popc(negm[ps]) >= __popc(posm[ps]);
        up[ps] = two ? neg_major : negm[ps] != 0;
        maskA[ps] = two ? (neg_major ? negm[ps] : posm[ps]) : 0x1ffu;
        maskB[ps] = two ? (neg_major ? posm[ps] : negm[ps]) : 0u;
        unionB[ps] = __reduce_or_sync(0xffffffffu, maskB[ps]);
        mixed[ps] = unionB[ps] != 0;
      }
      bool any_mixed = false;
#pragma unroll
      for (int ps = 0; ps < kScSets; ++ps) any_mixed = any_mixed || mixed[ps];
      uint32_t late[kScSets][2][2] = {};
      // Tile class (score_tile_class): bases that are zero for every row of the tile are not stored, and a
      // class-1 tile has no k = 8, 9 K-block at all (the issuer and the weight producer skip it the same way).
      const int cls = p.skip_zero_steps ? score_tile_class(p, t, lane) : 0;
      const bool use_k = cls != 1;
      for (int sb = 0; sb < kScSuper; ++sb) {
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          // this chunk of x: shared memory -> registers, and the stage goes back to the producer at once
          uint64_t xv[kScSets][kScShifts];
          const int tr_i = (t == cluster_id && lane == 0 && (warp == 0 || warp == kScGenWarps - 1)) ? sb * 4 + c : 1 << 20;
          const int tr_r = warp == 0 ? 0 : 1;
          SC_TRACE(tr_r, tr_i, 0);
          mbar_wait(fullX(rx.idx), rx.phase, 21);
          SC_TRACE(tr_r, tr_i, 1);
          {
            const uint8_t* xs = smem_raw + (sX - raw) + rx.idx * kScXStage + (warp * (8 * kScSets) + rw) * 32 + sub * 8;
#pragma unroll
            for (int ps = 0; ps < kScSets; ++ps)
#pragma unroll
              for (int j = 0; j < kScShifts; ++j)
                xv[ps][j] = *reinterpret_cast<const uint64_t*>(xs + j * (128 * 32) + ps * (8 * 32));
          }
          __syncwarp();
          if (lane == 0) mbar_arrive_local(emptyX(rx.idx));
          rx.advance<kScXStages>();

          uint8_t* st[kScSets];
          uint8_t* sk[kScSets];
          uint32_t swz[kScSets];
#pragma unroll
          for (int ps = 0; ps < kScSets; ++ps) {
            const int r = warp * (8 * kScSets) + ps * 8 + rw;
            swz[ps] = (uint32_t)(r & 7);
            st[ps] = smem_raw + (sA - raw) + ra.idx * 2 * kScATile + r * 128 + sub * 4;
            sk[ps] = smem_raw + (sK - raw) + r * 128 + sub * 4;
          }
          // K-block 5*sb + c holds bases 0..7 (piece k); K-block 5*sb + 4 bases 8, 9 (piece (k - 8) * 4 + chunk)
          auto put = [&](int ps, int k, uint32_t hi, uint32_t lo) {
            uint8_t* dst = st[ps] + (((uint32_t)k ^ swz[ps]) << 4);
            *reinterpret_cast<uint32_t*>(dst) = hi;
            *reinterpret_cast<uint32_t*>(dst + kScATile) = lo;
          };
          auto put89 = [&](int ps, int k, int chunk, uint32_t hi, uint32_t lo) {
            uint8_t* dst = sk[ps] + (((uint32_t)((k - 8) * 4 + chunk) ^ swz[ps]) << 4);
            *reinterpret_cast<uint32_t*>(dst) = hi;
            *reinterpret_cast<uint32_t*>(dst + kScATile) = lo;
          };
          // The k = 8, 9 K-block has ONE stage: the MMA hands it back two K-blocks after the previous super-block's
          // last chunk.  Chunk 0 therefore keeps its four k = 8, 9 pieces in registers (late[]) and chunk 1 stores
          // them, by which time the stage is free and nobody waits.
          auto out89 = [&](int ps, int k, uint32_t hi, uint32_t lo) {
            if (c == 0) { late[ps][k - 8][0] = hi; late[ps][k - 8][1] = lo; }
            else put89(ps, k, c, hi, lo);
          };
          // chunk 1, before any of its own k = 8, 9 stores: wait for the stage, then chunk 0's pieces
          auto flush_late = [&]() {
            if (c != 1 || !use_k) return;
            mbar_wait(emptyK, kphase ^ 1, 12);
#pragma unroll
            for (int ps = 0; ps < kScSets; ++ps)
#pragma unroll
              for (int k = 8; k < 10; ++k) put89(ps, k, 0, late[ps][k - 8][0], late[ps][k - 8][1]);
          };
          // One shift's contribution to the five sums (decay rates) of row set ps; y2, y10, y20 are formed here from
          // the kept y and y^5.
          auto add_shift = [&](int ps, int j, uint64_t xsel, uint64_t (&acc)[5]) {
            const float y1 = pinned(y[ps][j]), y5f = pinned(y5[ps][j]);
            const float y2 = y1 * y1, y10 = y5f * y5f, y20 = y10 * y10;
            const float e[5] = {y1, y2, y5f, y10, y20};
#pragma unroll
            for (int q = 0; q < 5; ++q) acc[q] = ffma2(pack2f(e[q], e[q]), xsel, acc[q]);
          };

          auto split5 = [&](const uint64_t (&acc)[5], uint32_t (&hi)[5], uint32_t (&lo)[5]) {
#pragma unroll
            for (int q = 0; q < 5; ++q) {
              const float2 a = unpack2f(acc[q]);
              split2<__half>(a.x, a.y, hi[q], lo[q]);
            }
          };
          // Pass A of both row sets in one instruction stream: ten independent accumulator chains (with two warps per
          // scheduler the chains' latency is what the generators run at).
          uint32_t hiA[kScSets][5], loA[kScSets][5];
          {
            uint64_t acc[kScSets][5];
#pragma unroll
            for (int ps = 0; ps < kScSets; ++ps)
#pragma unroll
              for (int q = 0; q < 5; ++q) acc[ps][q] = 0ull;
            if (any_mixed) {
#pragma unroll
              for (int j = 0; j < kScShifts; ++j)
#pragma unroll
                for (int ps = 0; ps < kScSets; ++ps) add_shift(ps, j, ((maskA[ps] >> j) & 1u) ? xv[ps][j] : 0ull, acc[ps]);
            } else {
#pragma unroll
              for (int j = 0; j < kScShifts; ++j)
#pragma unroll
                for (int ps = 0; ps < kScSets; ++ps) add_shift(ps, j, xv[ps][j], acc[ps]);
            }
#pragma unroll
            for (int ps = 0; ps < kScSets; ++ps) split5(acc[ps], hiA[ps], loA[ps]);
          }
          // the waits come after the arithmetic: the MMA has that long to hand the stages back
          SC_TRACE(tr_r, tr_i, 2);
          mbar_wait(emptyA(ra.idx), ra.phase ^ 1, 11);
          SC_TRACE(tr_r, tr_i, 3);
          flush_late();
          SC_TRACE(tr_r, tr_i, 4);
#pragma unroll
          for (int ps = 0; ps < kScSets; ++ps) {
            uint32_t hiB[5] = {0, 0, 0, 0, 0}, loB[5] = {0, 0, 0, 0, 0};
            if (mixed[ps]) {   // pass B: the smaller side of this set's two-sided rows, over the shifts the set has there
              uint64_t acc[5] = {0ull, 0ull, 0ull, 0ull, 0ull};
#pragma unroll
              for (int j = 0; j < kScShifts; ++j)
                if ((unionB[ps] >> j) & 1u) add_shift(ps, j, ((maskB[ps] >> j) & 1u) ? xv[ps][j] : 0ull, acc);
              split5(acc, hiB, loB);
            }
            if (cls == 1) {          // every row upstream: bases 0..4 are the sums, base 5 completes the K step with zeros
#pragma unroll
              for (int k = 0; k < 5; ++k) put(ps, k, hiA[ps][k], loA[ps][k]);
              put(ps, 5, 0u, 0u);
            } else if (cls == 2) {   // every row downstream: base 4 zero, bases 5..9 the sums
              put(ps, 4, 0u, 0u);
#pragma unroll
              for (int k = 5; k < 8; ++k) put(ps, k, hiA[ps][k - 5], loA[ps][k - 5]);
#pragma unroll
              for (int k = 8; k < 10; ++k) out89(ps, k, hiA[ps][k - 5], loA[ps][k - 5]);
            } else {
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                const bool a_side = (k < 5) == up[ps];
                put(ps, k, a_side ? hiA[ps][k % 5] : hiB[k % 5], a_side ? loA[ps][k % 5] : loB[k % 5]);
              }
#pragma unroll
              for (int k = 8; k < 10; ++k)
                out89(ps, k, up[ps] ? hiB[k - 5] : hiA[ps][k - 5], up[ps] ? loB[k - 5] : loA[ps][k - 5]);
            }
          }
          SC_TRACE(tr_r, tr_i, 6);
          fence_proxy_async_smem();   // generic-proxy stores -> visible to the tensor core's async-proxy reads
          SC_TRACE(tr_r, tr_i, 7);
          __syncwarp();
          if (lane == 0) {
            mbar_arrive_remote(mapa(fullA(ra.idx), 0));
            if (c == 3 && use_k) mbar_arrive_remote(mapa(fullK, 0));
          }
          __syncwarp();
          SC_TRACE(tr_r, tr_i, 5);
          ra.advance<SA>();
        }
        if (use_k) kphase ^= 1;
      }
      // ---- epilogue: TMEM -> + offset -> out[row, model]; warp handles its lane quarter and every other 16-column chunk
      mbar_wait(tfull, tphase, 13);
      tphase ^= 1;
      tc_fence_after();
      {
        const int q = warp & 3;
        const long long orow = (long long)t * 256 + rank * 128 + q * 32 + lane;
        const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16);
        for (int c16 = warp >> 2; c16 < p.n_pad / 16; c16 += kScGenWarps / 4) {
          uint32_t v[16], u[16];
          tmem_ld16(t0 + c16 * 16, v);
          tmem_ld16(t0 + 256 + c16 * 16, u);
          tmem_ld_wait();
          if (orow < p.n) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int col = c16 * 16 + i;
              if (col < p.m)
                p.out[(size_t)orow * p.ld_out + col] = (__uint_as_float(v[i]) + __uint_as_float(u[i])) + __ldg(p.offset + col);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(mapa(tempty, 0));
    }
  } else if (warp == kScGenWarps) {
    // ------------------------------------------------------------------ weight tiles by TMA (both CTAs, own half of N)
    RingState rb;
    const int row0 = (int)rank * (p.n_pad / 2);
    for (int t = cluster_id; t < p.tiles; t += num_clusters) {
      const bool use_k = !(p.skip_zero_steps && score_tile_class(p, t, lane) == 1);
      for (int kb = 0; kb < kScKBlocks; ++kb) {
        if (!use_k && kb % 5 == 4) continue;
        mbar_wait(emptyB(rb.idx), rb.phase ^ 1, 14);
        if (elect_one()) {
          if (rank == 0) mbar_arrive_expect_tx(fullB(rb.idx), 4 * b_half);
          const uint32_t bar = mapa(fullB(rb.idx), 0);
          const uint32_t dst = sB + rb.idx * 2 * b_half;
          tma_load_2d_pair(dst, &tmW, bar, kb * 64, row0);
          tma_load_2d_pair(dst + b_half, &tmW, bar, kScK + kb * 64, row0);
        }
        __syncwarp();
        if (++rb.idx == (uint32_t)p.sb_stages) { rb.idx = 0; rb.phase ^= 1; }
      }
    }
  } else if (warp == kScGenWarps + 2) {
    // ------------------------------------------------------------------ x chunks by TMA (each CTA its own 128 rows)
    RingState rx;
    for (int t = cluster_id; t < p.tiles; t += num_clusters) {
      const int row0 = t * 256 + (int)rank * 128;
      for (int ch = 0; ch < kScSuper * 4; ++ch) {
        SC_TRACE(3, (t == cluster_id && lane == 0) ? ch : 1 << 20, 0);
        mbar_wait(emptyX(rx.idx), rx.phase ^ 1, 22);
        SC_TRACE(3, (t == cluster_id && lane == 0) ? ch : 1 << 20, 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(fullX(rx.idx), kScXStage);
          tma_load_3d(sX + rx.idx * kScXStage, &tmX, fullX(rx.idx), ch * 8, row0, 0);
#if BX_SC_XPF > 0
          if (ch + BX_SC_XPF < kScSuper * 4) tma_prefetch_3d(&tmX, (ch + BX_SC_XPF) * 8, row0, 0);
#endif
        }
        __syncwarp();
        rx.advance<kScXStages>();
      }
    }
  } else if (warp == kScGenWarps + 1 && rank == 0) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA)
    RingState ra, rb;
    uint32_t kphase = 0, tphase = 0;
    for (int t = cluster_id; t < p.tiles; t += num_clusters) {
      mbar_wait(tempty, tphase ^ 1, 15);   // the previous tile's epilogue has read the accumulator
      tphase ^= 1;
      tc_fence_after();
      uint32_t accum = 0;
      const int cls = p.skip_zero_steps ? score_tile_class(p, t, lane) : 0;
      for (int kb = 0; kb < kScKBlocks; ++kb) {
        const bool k89 = (kb % 5) == 4;
        if (k89 && cls == 1) continue;
        // K steps whose bases are zero for every row of the tile: bases 6..9 in class 1, bases 0..3 in class 2
        const int ks_lo = (cls == 2 && !k89) ? 2 : 0, ks_hi = (cls == 1) ? 3 : 4;
        const int tr_k = (t == cluster_id && lane == 0) ? kb : 1 << 20;
        SC_TRACE(2, tr_k, 0);
        uint32_t a_hi;
        if (k89) { mbar_wait(fullK, kphase, 16); a_hi = sK; }
        else { mbar_wait(fullA(ra.idx), ra.phase, 17); a_hi = sA + ra.idx * 2 * kScATile; }
        SC_TRACE(2, tr_k, 1);
        mbar_wait(fullB(rb.idx), rb.phase, 18);
        SC_TRACE(2, tr_k, 2);
        tc_fence_after();
        const uint32_t b_hi = sB + rb.idx * 2 * b_half;
        if (elect_one()) {
          const uint32_t ah = sw128_desc_lo(a_hi), al = sw128_desc_lo(a_hi + kScATile);
          const uint32_t bh = sw128_desc_lo(b_hi), bl = sw128_desc_lo(b_hi + b_half);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            if (ks < ks_lo || ks >= ks_hi) continue;
            umma_f16<2>(tmem_base, sw128_desc(ah + 2 * ks), sw128_desc(bh + 2 * ks), p.idesc, accum);
            umma_f16<2>(tmem_base + 256, sw128_desc(ah + 2 * ks), sw128_desc(bl + 2 * ks), p.idesc, accum);
            umma_f16<2>(tmem_base + 256, sw128_desc(al + 2 * ks), sw128_desc(bh + 2 * ks), p.idesc, 1u);
            accum = 1;
          }
          umma_commit<2>(emptyB(rb.idx));
          umma_commit<2>(k89 ? emptyK : emptyA(ra.idx));
        }
        __syncwarp();
        SC_TRACE(2, tr_k, 3);
        accum = 1;
        if (++rb.idx == (uint32_t)p.sb_stages) { rb.idx = 0; rb.phase ^= 1; }
        if (k89) kphase ^= 1; else ra.advance<SA>();
      }
      if (elect_one()) umma_commit<2>(tfull);
      __syncwarp();
    }
  }

  __syncwarp();
  tc_fence_before();
  cluster_sync_all();
  if (warp == kScGenWarps + 1) {
    tc_fence_after();
    tmem_dealloc<2>(tmem_base, 512);
  }
}

}  // namespace bx
