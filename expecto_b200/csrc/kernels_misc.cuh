// Non-GEMM kernels of the hot path: window extraction / one-hot / reverse complement from the 2-bit
// packed genome (reference: fetchSeqs chromatin.py:175-209, encodeSeqs chromatin.py:138-172),
// weight re-layout, small element-wise finalisers and the decay-weighted shift reductions
// (reference: compute_expecto_features.py:123-124, predict.py:87-136).  All are HBM-bound byte/float
// streaming kernels: coalesced along the innermost axis, no tensor cores.
#pragma once
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include "../../include/beluga_b200.h"
#include "encode.cuh"

namespace bx {

constexpr int kWin = 2000;
constexpr int kP1 = 1996;   // row pitch of the conv1 operand / conv1 output per window (1993 valid)
constexpr int kFeat = 2002;

// Codes of one window in strand orientation into shared memory (2000 + 8 pad entries of 4).
__device__ __forceinline__ void load_window_codes(uint8_t* codes, const bx_window_desc& d,
                                                  const uint32_t* __restrict__ bases2,
                                                  const uint32_t* __restrict__ nmask,
                                                  const uint8_t* __restrict__ alleles, int64_t n_bases) {
  const bool rc = d.flags & 1;
  for (int i = threadIdx.x; i < kWin + 8; i += blockDim.x) {
    int c = 4;
    if (i < kWin) {
      c = window_code(d, rc ? (kWin - 1 - i) : i, bases2, nmask, alleles, n_bases);
      if (rc && c < 4) c = 3 - c;  // A<->T, G<->C: encodeSeqs' [:, ::-1, ::-1]
    }
    codes[i] = (uint8_t)c;
  }
}

// encodeSeqs materialised: out uint8 [n,4,2000].  One block per window.
__global__ void encode_onehot_kernel(const bx_window_desc* __restrict__ descs, const uint8_t* __restrict__ alleles,
                                     const uint32_t* __restrict__ bases2, const uint32_t* __restrict__ nmask,
                                     int64_t n_bases, uint8_t* __restrict__ out) {
  __shared__ uint8_t codes[kWin + 8];
  const bx_window_desc d = descs[blockIdx.x];
  load_window_codes(codes, d, bases2, nmask, alleles, n_bases);
  __syncthreads();
  uint8_t* o = out + (size_t)blockIdx.x * 4 * kWin;
  for (int idx = threadIdx.x; idx < 4 * kWin; idx += blockDim.x) {
    const int ch = idx / kWin, i = idx - ch * kWin;
    o[idx] = codes[i] == ch ? 1 : 0;
  }
}

// Write the conv1 operand rows of one window from its codes: row l = [onehot(l), ..., onehot(l+7), 0 x32]
// (k = tap*4 + channel), 128 bytes per row.  Eight lanes share a row (16 bytes each), so one warp store covers
// four whole rows = 512 contiguous bytes.
template <typename T>
__device__ __forceinline__ void write_im2col_rows(const uint8_t* codes, uint4* __restrict__ dst_rows) {
  const uint32_t one = one_bits<T>();
  const int q = threadIdx.x & 7;  // 16-byte piece of the row: pieces 0..3 hold taps 2q, 2q+1; 4..7 are zero
  for (int l = threadIdx.x >> 3; l < kP1; l += blockDim.x >> 3) {
    uint4 v = make_uint4(0, 0, 0, 0);
    if (q < 4) {
      const int c0 = (l + 2 * q < kWin + 8) ? codes[l + 2 * q] : 4;
      const int c1 = (l + 2 * q + 1 < kWin + 8) ? codes[l + 2 * q + 1] : 4;
      // channels 0,1 -> first word of the tap ; channels 2,3 -> second word
      v.x = (c0 == 0) ? one : (c0 == 1) ? (one << 16) : 0u;
      v.y = (c0 == 2) ? one : (c0 == 3) ? (one << 16) : 0u;
      v.z = (c1 == 0) ? one : (c1 == 1) ? (one << 16) : 0u;
      v.w = (c1 == 2) ? one : (c1 == 3) ? (one << 16) : 0u;
    }
    dst_rows[(size_t)l * 8 + q] = v;
  }
}

// Kernel 1 fused form: packed genome + descriptors -> conv1 operand [n*1996, 64] (never a one-hot in HBM).
template <typename T>
__global__ void encode_im2col_kernel(const bx_window_desc* __restrict__ descs, const uint8_t* __restrict__ alleles,
                                     const uint32_t* __restrict__ bases2, const uint32_t* __restrict__ nmask,
                                     int64_t n_bases, T* __restrict__ a0) {
  __shared__ uint8_t codes[kWin + 8];
  const bx_window_desc d = descs[blockIdx.x];
  load_window_codes(codes, d, bases2, nmask, alleles, n_bases);
  __syncthreads();
  write_im2col_rows<T>(codes, reinterpret_cast<uint4*>(a0) + (size_t)blockIdx.x * kP1 * 8);
}

// Same from an explicit one-hot uint8 [n,4,2000] (all-zero column -> code 4).
template <typename T>
__global__ void onehot_im2col_kernel(const uint8_t* __restrict__ onehot, T* __restrict__ a0) {
  __shared__ uint8_t codes[kWin + 8];
  const uint8_t* x = onehot + (size_t)blockIdx.x * 4 * kWin;
  for (int i = threadIdx.x; i < kWin + 8; i += blockDim.x) {
    int c = 4;
    if (i < kWin) {
      if (x[i]) c = 0;
      else if (x[kWin + i]) c = 1;
      else if (x[2 * kWin + i]) c = 2;
      else if (x[3 * kWin + i]) c = 3;
    }
    codes[i] = (uint8_t)c;
  }
  __syncthreads();
  write_im2col_rows<T>(codes, reinterpret_cast<uint4*>(a0) + (size_t)blockIdx.x * kP1 * 8);
}

// ------------------------------------------------------------------ segment mode (shared activations)
// The 9 (chromatin.py:243) or 200 (compute_expecto_features.py:88) shifted windows of one variant / gene overlap
// almost entirely: they are 2000-bp substrings, 200 bp apart, of ONE strand-oriented sequence (the "segment":
// 3600 or 41800 positions; for an alt allele the spliced sequence).  A valid convolution's output at a sequence
// position does not depend on which window asked for it, and both MaxPool(1,4) stages keep their alignment for
// offsets that are multiples of 16 (even multiples of the 200-bp step) — odd multiples fall on the second pool's
// phase 2.  So the conv stack runs once per segment (twice from conv4's pooling on: phase A / phase B) and only
// the dense layers run per window, on 106-row slices of the segment's conv6 output.  Every window's result is
// bit-identical to the per-window path (same dot products, same order).
//
// conv1 operand rows of segments: a0 [n_seg * pitch, 64]; grid (ceil(pitch*8/256), n_seg); one thread per
// 16-byte piece of a row (two taps), codes fetched straight from the packed genome (L2-resident re-reads).
template <typename T>
__global__ void __launch_bounds__(256)
encode_im2col_seg_kernel(const bx_window_desc* __restrict__ descs, const uint8_t* __restrict__ alleles,
                         const uint32_t* __restrict__ bases2, const uint32_t* __restrict__ nmask, int64_t n_bases,
                         T* __restrict__ a0, int L, int pitch) {
  const int idx = blockIdx.x * 256 + threadIdx.x;
  const int l = idx >> 3, q = idx & 7;
  if (l >= pitch) return;
  const bx_window_desc d = descs[blockIdx.y];
  const uint32_t one = one_bits<T>();
  uint4 v = make_uint4(0, 0, 0, 0);
  if (q < 4) {
    const int c0 = segment_code(d, l + 2 * q, L, bases2, nmask, alleles, n_bases);
    const int c1 = segment_code(d, l + 2 * q + 1, L, bases2, nmask, alleles, n_bases);
    v.x = (c0 == 0) ? one : (c0 == 1) ? (one << 16) : 0u;
    v.y = (c0 == 2) ? one : (c0 == 3) ? (one << 16) : 0u;
    v.z = (c1 == 0) ? one : (c1 == 1) ? (one << 16) : 0u;
    v.w = (c1 == 2) ? one : (c1 == 3) ? (one << 16) : 0u;
  }
  reinterpret_cast<uint4*>(a0)[((size_t)blockIdx.y * pitch + l) * 8 + q] = v;
}

// Both phases of the second MaxPool(1,4) from conv4's half-pooled output h[n_seg, ph, C] (h[j] = max of conv4 rows
// 2j, 2j+1): phase A pools rows 4g..4g+3 = max(h[2g], h[2g+1]); phase B (windows whose offset is an odd multiple
// of 200 bp) pools rows 4g+2..4g+5 = max(h[2g+1], h[2g+2]).  max is exact, so both equal the fused MaxPool4
// epilogue bit for bit.  One thread per 8 channels of one pooled row; rows >= l4b of phase B are zero-filled.
//
// SPLIT (split-operand mode): rows hold [hi(C) | lo(C)] fp16 pairs (value = hi + lo, |lo| <= ulp(hi)/2); the larger
// value is the pair that is larger lexicographically, and because the hi/lo split is monotonic this equals the split
// of the fp32 maximum the fused epilogue forms — still bit-identical to the per-window path.
template <typename T>
__device__ __forceinline__ void pair_max8(const uint4& ah, const uint4& al, const uint4& bh, const uint4& bl, uint4& oh,
                                          uint4& ol) {
  const T* xh = reinterpret_cast<const T*>(&ah); const T* xl = reinterpret_cast<const T*>(&al);
  const T* yh = reinterpret_cast<const T*>(&bh); const T* yl = reinterpret_cast<const T*>(&bl);
  T* rh = reinterpret_cast<T*>(&oh); T* rl = reinterpret_cast<T*>(&ol);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float a1 = (float)xh[i], b1 = (float)yh[i];
    const bool first = a1 > b1 || (a1 == b1 && (float)xl[i] >= (float)yl[i]);
    rh[i] = first ? xh[i] : yh[i];
    rl[i] = first ? xl[i] : yl[i];
  }
}

template <typename T, bool SPLIT>
__global__ void __launch_bounds__(256)
pool_phases_kernel(const T* __restrict__ h, T* __restrict__ a, T* __restrict__ b, int ph, int p4, int l4b, int c8,
                   long long total) {
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= total) return;
  const int c = (int)(idx % c8);
  const long long row = idx / c8;
  const int g = (int)(row % p4);
  const long long seg = row / p4;
  if constexpr (SPLIT) {
    // c8 = C/8 uint4 of high parts per row, followed by as many of low parts (row width 2*c8 uint4)
    const int rw = 2 * c8;
    const uint4* hp = reinterpret_cast<const uint4*>(h) + (seg * ph + 2 * g) * rw + c;
    const uint4 h0 = hp[0], l0 = hp[c8], h1 = hp[rw], l1 = hp[rw + c8];
    uint4 oh, ol;
    pair_max8<T>(h0, l0, h1, l1, oh, ol);
    uint4* ap = reinterpret_cast<uint4*>(a) + row * rw + c;
    ap[0] = oh; ap[c8] = ol;
    oh = ol = make_uint4(0, 0, 0, 0);
    if (g < l4b) pair_max8<T>(h1, l1, hp[2 * (size_t)rw], hp[2 * (size_t)rw + c8], oh, ol);
    uint4* bp = reinterpret_cast<uint4*>(b) + row * rw + c;
    bp[0] = oh; bp[c8] = ol;
    return;
  }
  const uint4* hp = reinterpret_cast<const uint4*>(h) + (seg * ph + 2 * g) * c8 + c;
  auto mx = [](uint4 x, uint4 y) {
    uint4 r;
    r.x = max2<T>(x.x, y.x); r.y = max2<T>(x.y, y.y); r.z = max2<T>(x.z, y.z); r.w = max2<T>(x.w, y.w);
    return r;
  };
  const uint4 h0 = hp[0], h1 = hp[c8];
  reinterpret_cast<uint4*>(a)[idx] = mx(h0, h1);
  uint4 vb = make_uint4(0, 0, 0, 0);
  if (g < l4b) vb = mx(h1, hp[2 * (size_t)c8]);
  reinterpret_cast<uint4*>(b)[idx] = vb;
}

// Geometry of the window -> segment mapping handed to the gather kernel.
struct SegGather {
  int n_groups;       // variants / genes in this chunk
  int segs_per_group; // windows' segments per group: 4 (ref fwd, ref rc, alt fwd, alt rc) or 2 (fwd, rc)
  int n_shifts;       // windows per segment
  int shift_major;    // 1: window index = (shift, group, seg) [chromatin]; 0: (group, shift, seg) [TSS]
  int p6;             // rows per segment of the conv6 output
  // delta mode: only the two ref segments of a group were run in full; alt windows take conv6 rows
  // [row0, row0 + valid) from the group's alt patch and every other row from the ref segment of the same strand
  int delta, p6p, row0, valid_a, valid_b;
  int rw;             // elements per conv6 row: 640, or 1280 in split-operand mode ([hi | lo])
  int k_of_shift[200]; // offset index (window start = 200*k positions into a non-reversed segment) per shift index
};

// FC1 operand rows: dense [n_windows, 106*640] gathered from the segment-level conv6 outputs (phase A / B).
// One block per window; 135,680 contiguous bytes each.  Window w0 + blockIdx.x of the final output order.
template <typename T>
__global__ void __launch_bounds__(256)
gather_fc1_kernel(const T* __restrict__ act6a, const T* __restrict__ act6b, const T* __restrict__ patch6a,
                  const T* __restrict__ patch6b, const bx_window_desc* __restrict__ descs, const SegGather g,
                  long long w0, T* __restrict__ dst) {
  const long long w = w0 + blockIdx.x;
  const int q = (int)(w % g.segs_per_group);
  const long long r = w / g.segs_per_group;
  int grp, si;
  if (g.shift_major) { si = (int)(r / g.n_groups); grp = (int)(r % g.n_groups); }
  else { grp = (int)(r / g.n_shifts); si = (int)(r % g.n_shifts); }
  // segment that was run in full for this window (delta mode: the ref segment of the same strand)
  const long long seg = g.delta ? (long long)grp * 2 + (q & 1) : (long long)grp * g.segs_per_group + q;
  int k = g.k_of_shift[si];
  if (descs[seg].flags & 2) k = g.n_shifts - 1 - k;  // reversed offset order (rc segments, '-' strand genes)
  const int j = k >> 1;
  const int off = ((k & 1) ? 12 : 0) + 25 * j;       // first conv6 row of the window within its phase
  const T* src = ((k & 1) ? act6b : act6a) + ((size_t)seg * g.p6 + off) * g.rw;
  const uint4* s4 = reinterpret_cast<const uint4*>(src);
  uint4* d4 = reinterpret_cast<uint4*>(dst + (size_t)blockIdx.x * 106 * g.rw);
  const int r16 = g.rw / 8;                           // 16-byte pieces per conv6 row (80, or 160 in split mode)
  const int N16 = 106 * r16;
  if (g.delta && q >= 2) {
    const int valid = (k & 1) ? g.valid_b : g.valid_a;
    const int lo = g.row0 - off;                      // window rows [lo, lo + valid) come from the patch
    const uint4* p4 = reinterpret_cast<const uint4*>(((k & 1) ? patch6b : patch6a) + (size_t)seg * g.p6p * g.rw);
    for (int i = threadIdx.x; i < N16; i += 256) {
      const int l = i / r16;
      const int pr = l - lo;
      d4[i] = (pr >= 0 && pr < valid) ? __ldg(p4 + (size_t)pr * r16 + (i - l * r16)) : __ldg(s4 + i);
    }
  } else {
    for (int i = threadIdx.x; i < N16; i += 256) d4[i] = __ldg(s4 + i);
  }
}

// ------------------------------------------------------------------ weight re-layout (run once)
template <typename T> __device__ __forceinline__ T to_t(float v);
template <> __device__ __forceinline__ __half to_t<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 to_t<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// Split-operand mode stores every weight as an fp16 pair: the rounded value and, `lo_off` columns further along the
// row, the rounding residual (value - rounded value, itself rounded to 16 bits).
template <typename T>
__device__ __forceinline__ void store_weight(T* __restrict__ row, size_t col, float v, int lo_off) {
  const T h = to_t<T>(v);
  row[col] = h;
  if (lo_off > 0) row[col + lo_off] = to_t<T>(v - (float)h);
}
// conv1: src [320,4,1,8] -> dst [320,64], k = t*4 + c; k >= 32 zero, or (split) the low halves.
template <typename T>
__global__ void prep_conv1_kernel(const float* __restrict__ src, T* __restrict__ dst, int cout, int split) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= cout * 32) return;
  const int co = idx >> 5, k = idx & 31;
  const int t = k >> 2, c = k & 3;
  const float v = src[(co * 4 + c) * 8 + t];
  T* row = dst + (size_t)co * 64;
  row[k + 32] = to_t<T>(0.f);
  store_weight<T>(row, k, v, split ? 32 : 0);
}
// conv2..6: src [Cout,Cin,1,8] -> dst [8*Cout, ld] (tap-major, K-major rows; ld = Cin, or 2*Cin in split mode).
template <typename T>
__global__ void prep_conv_kernel(const float* __restrict__ src, T* __restrict__ dst, int cout, int cin, int split) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)8 * cout * cin) return;
  const int ci = idx % cin;
  const int co = (idx / cin) % cout;
  const int t = idx / ((size_t)cin * cout);
  const size_t ld = split ? 2 * (size_t)cin : (size_t)cin;
  store_weight<T>(dst + ((size_t)t * cout + co) * ld, ci, src[((size_t)co * cin + ci) * 8 + t], split ? cin : 0);
}
// fc1: src [2003, 67840] with column c*106+l -> dst [2048, ld] with column l*640+c, rows >= 2003 zero.
template <typename T>
__global__ void prep_fc1_kernel(const float* __restrict__ src, T* __restrict__ dst, int split) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)2048 * 67840) return;
  const int col = idx % 67840;
  const int n = idx / 67840;
  const int l = col / 640, c = col - l * 640;
  const size_t ld = split ? 2 * (size_t)67840 : (size_t)67840;
  store_weight<T>(dst + (size_t)n * ld, col, n < 2003 ? src[(size_t)n * 67840 + c * 106 + l] : 0.f, split ? 67840 : 0);
}
// fc2: src [2002, 2003] -> dst [2048, ld] zero padded (ld = 2048, or 4096 in split mode).
template <typename T>
__global__ void prep_fc2_kernel(const float* __restrict__ src, T* __restrict__ dst, int split) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= 2048 * 2048) return;
  const int k = idx & 2047, n = idx >> 11;
  const size_t ld = split ? 4096 : 2048;
  store_weight<T>(dst + (size_t)n * ld, k, (n < 2002 && k < 2003) ? src[(size_t)n * 2003 + k] : 0.f, split ? 2048 : 0);
}

// relu + round of four consecutive hidden units starting at flat index idx of [n, 2048]; split-operand mode writes
// rows of 4096 = [hi(2048) | lo(2048)].
template <typename T>
__device__ __forceinline__ void store_hidden(T* __restrict__ h, size_t idx, int split, float x0, float x1, float x2,
                                             float x3) {
  const float v[4] = {fminf(fmaxf(x0, 0.f), 65504.f), fminf(fmaxf(x1, 0.f), 65504.f), fminf(fmaxf(x2, 0.f), 65504.f),
                      fminf(fmaxf(x3, 0.f), 65504.f)};
  if (!split) {
#pragma unroll
    for (int i = 0; i < 4; ++i) h[idx + i] = to_t<T>(v[i]);
    return;
  }
  T* row = h + (idx >> 11) * 4096 + (idx & 2047);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const T hi = to_t<T>(v[i]);
    row[i] = hi;
    row[2048 + i] = to_t<T>(v[i] - (float)hi);
  }
}

// fc1 finaliser: h = relu(sum_ks part[ks] + b1) -> 16-bit [n, 2048]; slices are summed in a fixed order so
// the result does not depend on scheduling (the split-K GEMM writes plain stores, no atomics).
template <typename T>
__global__ void fc1_finalize_kernel(const float* __restrict__ part, int k_splits, size_t split_stride,
                                    const float* __restrict__ bias, T* __restrict__ h, size_t total, int split) {
  const size_t idx = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (idx >= total) return;
  float4 a = *reinterpret_cast<const float4*>(part + idx);
  for (int k = 1; k < k_splits; ++k) {
    const float4 b = *reinterpret_cast<const float4*>(part + (size_t)k * split_stride + idx);
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
  }
  const float4 b4 = *reinterpret_cast<const float4*>(bias + (idx & 2047));
  store_hidden<T>(h, idx, split, a.x + b4.x, a.y + b4.y, a.z + b4.z, a.w + b4.w);
}

// ------------------------------------------------------------------ delta mode, dense layers
// In delta mode an alt window's FC1 operand equals the ref window's (same shift, same strand) outside the conv6
// rows the patch supplies.  FC1's split-K slices are fixed K ranges (8 slices of 133 K-blocks = 13.3 rows), and the
// finaliser adds the slices in index order, so every slice whose K range holds no patched row has, for the alt
// window, bit for bit the partial sum the ref window already produced.  The dense phase therefore runs FC1 over
// all slices for the ref windows only (partials kept for the whole chunk) and, for the alt windows, only over the
// 1-3 slices their patched rows fall into; the finaliser takes the remaining slices from the ref window's partials.
// Dense row orders: ref row rr = (si*ng + grp)*2 + t; alt row ar = c*2*ng + grp*2 + t with t = strand (0 fwd, 1 rc)
// and class c = the window's offset index k (fwd) or S-1-k (rc), so that all rows of a class share one offset.
struct DeltaDense {
  int ng, S, shift_major;
  int p6, p6p, row0, valid_a, valid_b;   // as SegGather
  int kb_per_split;                       // K-blocks (64 elements, a tenth of a 640-channel row) per FC1 slice
  int rw;                                 // elements per conv6 row: 640, or 1280 in split-operand mode ([hi | lo])
  int k_of_shift[200], si_of_k[200];
};

struct DeltaClass { int off, lo, hi, ks_lo, ks_hi; };  // conv6 row offset, patched window rows [lo, hi), slice range

__host__ __device__ inline DeltaClass delta_class(const DeltaDense& g, int c) {
  DeltaClass r;
  r.off = ((c & 1) ? 12 : 0) + 25 * (c >> 1);
  const int valid = (c & 1) ? g.valid_b : g.valid_a;
  r.lo = g.row0 - r.off < 0 ? 0 : g.row0 - r.off;
  r.hi = g.row0 - r.off + valid > 106 ? 106 : g.row0 - r.off + valid;
  if (r.lo >= r.hi) {  // the patch does not reach this window: no slice of its own
    r.lo = r.hi = 0; r.ks_lo = 1; r.ks_hi = 0;
    return r;
  }
  r.ks_lo = r.lo * 10 / g.kb_per_split;
  r.ks_hi = (r.hi * 10 - 1) / g.kb_per_split;
  return r;
}

// FC1 operand rows of a piece of the ref (alt == 0) or alt (alt == 1) dense order starting at dense row d0.
// Alt rows copy only the conv6 rows their FC1 slices read: patched rows from the patch, the others from the ref
// segment.  One block per dense row.
template <typename T>
__global__ void __launch_bounds__(256)
gather_fc1_delta_kernel(const T* __restrict__ act6a, const T* __restrict__ act6b, const T* __restrict__ patch6a,
                        const T* __restrict__ patch6b, const DeltaDense g, int alt, long long d0, T* __restrict__ dst) {
  const long long d = d0 + blockIdx.x;
  int c, grp, t;
  if (alt) {
    c = (int)(d / (2 * g.ng));
    const int x = (int)(d % (2 * g.ng));
    grp = x >> 1; t = x & 1;
  } else {
    t = (int)(d & 1);
    const long long x = d >> 1;
    const int si = (int)(x / g.ng);
    grp = (int)(x % g.ng);
    const int k = g.k_of_shift[si];
    c = t ? g.S - 1 - k : k;
  }
  const DeltaClass cl = delta_class(g, c);
  const long long seg = (long long)grp * 2 + t;
  const bool pb = c & 1;
  const uint4* s4 = reinterpret_cast<const uint4*>((pb ? act6b : act6a) + ((size_t)seg * g.p6 + cl.off) * g.rw);
  uint4* d4 = reinterpret_cast<uint4*>(dst + (size_t)blockIdx.x * 106 * g.rw);
  const int r16 = g.rw / 8;                                       // 16-byte pieces per conv6 row
  if (!alt) {
    const int N16 = 106 * r16;
    for (int i = threadIdx.x; i < N16; i += 256) d4[i] = __ldg(s4 + i);
    return;
  }
  const uint4* p4 = reinterpret_cast<const uint4*>((pb ? patch6b : patch6a) + (size_t)seg * g.p6p * g.rw);
  if (cl.ks_lo > cl.ks_hi) return;                               // nothing of this window is recomputed
  const int plo = g.row0 - cl.off;                               // window row of patch row 0
  const int r_lo = cl.ks_lo * g.kb_per_split / 10;               // conv6 rows the computed slices read
  int r_hi = ((cl.ks_hi + 1) * g.kb_per_split + 9) / 10;
  if (r_hi > 106) r_hi = 106;
  for (int i = r_lo * r16 + threadIdx.x; i < r_hi * r16; i += 256) {
    const int l = i / r16;
    d4[i] = (l >= cl.lo && l < cl.hi) ? __ldg(p4 + (size_t)(l - plo) * r16 + (i - l * r16)) : __ldg(s4 + i);
  }
}

// fc1 finaliser for a piece of alt rows: slices inside the row's class range come from this piece's partials,
// the others from the ref window's; same summation order as fc1_finalize_kernel.
template <typename T>
__global__ void fc1_finalize_delta_kernel(const float* __restrict__ part, size_t split_stride,
                                          const float* __restrict__ ref_part, size_t ref_stride, const DeltaDense g,
                                          long long a0, int k_splits, const float* __restrict__ bias,
                                          T* __restrict__ h, size_t total, int split) {
  const size_t idx = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (idx >= total) return;
  const long long ar = a0 + (long long)(idx >> 11);
  const int c = (int)(ar / (2 * g.ng));
  const int x = (int)(ar % (2 * g.ng));
  const int grp = x >> 1, t = x & 1;
  const int k = t ? g.S - 1 - c : c;
  const size_t ref_idx = ((size_t)((size_t)g.si_of_k[k] * g.ng + grp) * 2 + t) * 2048 + (idx & 2047);
  const DeltaClass cl = delta_class(g, c);
  auto slice = [&](int s) {
    return (s >= cl.ks_lo && s <= cl.ks_hi) ? *reinterpret_cast<const float4*>(part + (size_t)s * split_stride + idx)
                                            : *reinterpret_cast<const float4*>(ref_part + (size_t)s * ref_stride + ref_idx);
  };
  float4 a = slice(0);
  for (int s = 1; s < k_splits; ++s) {
    const float4 b = slice(s);
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
  }
  const float4 b4 = *reinterpret_cast<const float4*>(bias + (idx & 2047));
  store_hidden<T>(h, idx, split, a.x + b4.x, a.y + b4.y, a.z + b4.z, a.w + b4.w);
}

// Probabilities of a piece of dense rows -> their windows' rows of the final output (shift-major or group-major).
__global__ void __launch_bounds__(256)
scatter_delta_rows_kernel(const float* __restrict__ src, const DeltaDense g, int alt, long long d0,
                          float* __restrict__ out) {
  const long long d = d0 + blockIdx.x;
  int si, grp, t;
  if (alt) {
    const int c = (int)(d / (2 * g.ng));
    const int x = (int)(d % (2 * g.ng));
    grp = x >> 1; t = x & 1;
    si = g.si_of_k[t ? g.S - 1 - c : c];
  } else {
    t = (int)(d & 1);
    const long long x = d >> 1;
    si = (int)(x / g.ng);
    grp = (int)(x % g.ng);
  }
  const int q = alt ? 2 + t : t;
  const size_t w = g.shift_major ? ((size_t)si * g.ng + grp) * 4 + q : ((size_t)grp * g.S + si) * 4 + q;
  const float* s = src + (size_t)blockIdx.x * kFeat;
  float* o = out + w * kFeat;
  for (int f = threadIdx.x; f < kFeat; f += 256) o[f] = s[f];
}

// ------------------------------------------------------------------ chromatin.py:281-286 outputs
// probs [4N, 2002] in (variant, {ref fwd, ref rc, alt fwd, alt rc}) order.
__global__ void chromatin_outputs_kernel(const float* __restrict__ probs, int n_var, float* __restrict__ ref,
                                         float* __restrict__ alt, float* __restrict__ diff) {
  const int v = blockIdx.x;
  const float* p = probs + (size_t)v * 4 * kFeat;
  for (int f = threadIdx.x; f < kFeat; f += blockDim.x) {
    const float rf = p[f], rr = p[kFeat + f], af = p[2 * kFeat + f], ar = p[3 * kFeat + f];
    const size_t of = (size_t)v * kFeat + f, orc = (size_t)(n_var + v) * kFeat + f;
    ref[of] = rf; ref[orc] = rr;
    alt[of] = af; alt[orc] = ar;
    diff[of] = af - rf; diff[orc] = ar - rr;
  }
}
// predict.py:186-190 folded in: (fwd + rc) / 2 of diff/ref/alt, where diff halves are formed first as the
// reference does (alt - ref per strand, chromatin.py:281).
__global__ void chromatin_outputs_avg_kernel(const float* __restrict__ probs, int n_var, float* __restrict__ ref,
                                             float* __restrict__ alt, float* __restrict__ diff) {
  const int v = blockIdx.x;
  const float* p = probs + (size_t)v * 4 * kFeat;
  for (int f = threadIdx.x; f < kFeat; f += blockDim.x) {
    const float rf = p[f], rr = p[kFeat + f], af = p[2 * kFeat + f], ar = p[3 * kFeat + f];
    const size_t o = (size_t)v * kFeat + f;
    ref[o] = (rf + rr) / 2.0f;
    alt[o] = (af + ar) / 2.0f;
    diff[o] = ((af - rf) + (ar - rr)) / 2.0f;
  }
}

// predict.py:183-190: (rows 0..n-1 + rows n..2n-1) / 2.0 of a strand-stacked [2n, 2002] matrix -> [n, 2002].
__global__ void __launch_bounds__(256)
strand_average_kernel(const float* __restrict__ x, int n, float* __restrict__ out) {
  const size_t total2 = (size_t)n * (kFeat / 2);
  const float2* f = reinterpret_cast<const float2*>(x);
  const float2* r = f + total2;
  float2* o = reinterpret_cast<float2*>(out);
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total2; i += (size_t)gridDim.x * 256) {
    const float2 a = __ldg(f + i), b = __ldg(r + i);
    o[i] = make_float2((a.x + b.x) / 2.0f, (a.y + b.y) / 2.0f);
  }
}

// ------------------------------------------------------------------ kernel 4: decay-weighted shift reductions
// HBM-bound streaming reductions, coalesced along the 2002-feature axis.  Rows are 8008 bytes, so every row is
// 8-byte aligned and each thread moves float2 (two features); loads of several shifts are issued back to back
// so that enough bytes are in flight per SM to cover HBM latency.
//
// Fixed weights (compute_expecto_features.py:123-124).  Block = (gene, 256-feature tile) x 4 shift slices; each
// thread keeps 2 x 10 partial sums, slices are combined through shared memory.
// probs [G, S, 2, 2002], w [10, S], out [G, 10, 2002].
constexpr int kTfF = 128, kTfSlices = 4;   // 128 float2 lanes = 256 features per block
constexpr int kFeat2 = kFeat / 2;          // 1001 float2 per row
__global__ void __launch_bounds__(kTfF * kTfSlices)
decay_fixed_kernel(const float* __restrict__ probs, const float* __restrict__ w, int n_shifts,
                   float* __restrict__ out) {
  extern __shared__ float sm[];                 // w[10*S] then partials [slices-1][20][128]
  float* sw = sm;
  float* part = sm + 10 * n_shifts;
  const int g = blockIdx.y;
  const int f2 = blockIdx.x * kTfF + threadIdx.x;  // float2 index within the row
  const int slice = threadIdx.y;
  for (int i = threadIdx.y * kTfF + threadIdx.x; i < 10 * n_shifts; i += kTfF * kTfSlices) sw[i] = w[i];
  __syncthreads();
  float2 acc[10];
#pragma unroll
  for (int k = 0; k < 10; ++k) acc[k] = make_float2(0.f, 0.f);
  if (f2 < kFeat2) {
    const float2* p = reinterpret_cast<const float2*>(probs + (size_t)g * n_shifts * 2 * kFeat) + f2;
    int s = slice;
    // two shifts (four rows) per iteration: 4 independent 8-byte loads in flight per thread
    for (; s + kTfSlices < n_shifts; s += 2 * kTfSlices) {
      const float2 a0 = __ldg(p + (size_t)(2 * s) * kFeat2), a1 = __ldg(p + (size_t)(2 * s + 1) * kFeat2);
      const int s2 = s + kTfSlices;
      const float2 b0 = __ldg(p + (size_t)(2 * s2) * kFeat2), b1 = __ldg(p + (size_t)(2 * s2 + 1) * kFeat2);
      const float xa0 = 0.5f * (a0.x + a1.x), xa1 = 0.5f * (a0.y + a1.y);
      const float xb0 = 0.5f * (b0.x + b1.x), xb1 = 0.5f * (b0.y + b1.y);
#pragma unroll
      for (int k = 0; k < 10; ++k) {
        const float wa = sw[k * n_shifts + s], wb = sw[k * n_shifts + s2];
        acc[k].x = fmaf(wa, xa0, acc[k].x);
        acc[k].y = fmaf(wa, xa1, acc[k].y);
        acc[k].x = fmaf(wb, xb0, acc[k].x);
        acc[k].y = fmaf(wb, xb1, acc[k].y);
      }
    }
    for (; s < n_shifts; s += kTfSlices) {
      const float2 a0 = __ldg(p + (size_t)(2 * s) * kFeat2), a1 = __ldg(p + (size_t)(2 * s + 1) * kFeat2);
      const float xa0 = 0.5f * (a0.x + a1.x), xa1 = 0.5f * (a0.y + a1.y);
#pragma unroll
      for (int k = 0; k < 10; ++k) {
        const float wa = sw[k * n_shifts + s];
        acc[k].x = fmaf(wa, xa0, acc[k].x);
        acc[k].y = fmaf(wa, xa1, acc[k].y);
      }
    }
  }
  float2* part2 = reinterpret_cast<float2*>(part);
  if (slice > 0) {
#pragma unroll
    for (int k = 0; k < 10; ++k) part2[((slice - 1) * 10 + k) * kTfF + threadIdx.x] = acc[k];
  }
  __syncthreads();
  if (slice == 0 && f2 < kFeat2) {
    float2* o = reinterpret_cast<float2*>(out + (size_t)g * 10 * kFeat) + f2;
#pragma unroll
    for (int k = 0; k < 10; ++k) {
      float2 v = acc[k];
      for (int sl = 0; sl < kTfSlices - 1; ++sl) {
        const float2 q = part2[(sl * 10 + k) * kTfF + threadIdx.x];
        v.x += q.x;
        v.y += q.y;
      }
      o[(size_t)k * kFeat2] = v;
    }
  }
}

// Weights of one variant for every (shift, basis): predict.py:87-109, evaluated in fp64 like numpy does.
__device__ __forceinline__ void variant_weights_to_smem(float* sw, double dist, int strand,
                                                        const int32_t* __restrict__ shifts, int n_shifts) {
  if (threadIdx.x < n_shifts * 10) {
    const int j = threadIdx.x / 10, k = threadIdx.x - j * 10;
    const double sgn = strand > 0 ? 1.0 : -1.0;
    const double pos = dist * sgn + (double)shifts[j] * sgn;
    const double rate[5] = {0.01, 0.02, 0.05, 0.1, 0.2};
    const double e = exp(-rate[k % 5] * floor(fabs(pos / 200.0)));
    const bool on = (k < 5) ? (pos <= 0.0) : (pos >= 0.0);
    sw[j * 10 + k] = on ? (float)e : 0.f;
  }
}

// Per-variant weights, then the same reduction over S shifts.  x [S, n, 2002], out [n, 10, 2002].
// One block per variant (the fp64 weight prologue is paid once per 152 KB of traffic); each thread streams
// float2 columns tid, tid+256, ... with all S loads of a column issued before the FMAs.
constexpr int kTvMaxShifts = 25;
template <int S_STATIC>
__global__ void __launch_bounds__(256)
decay_variant_kernel(const float* __restrict__ x, const double* __restrict__ dist, const int8_t* __restrict__ strand,
                     const int32_t* __restrict__ shifts, int n, int n_shifts_rt, float* __restrict__ out) {
  __shared__ float sw[kTvMaxShifts * 10];  // [S][10]
  const int n_shifts = S_STATIC > 0 ? S_STATIC : n_shifts_rt;
  const int v = blockIdx.x;
  variant_weights_to_smem(sw, dist[v], strand[v], shifts, n_shifts);
  __syncthreads();
  const float2* xp = reinterpret_cast<const float2*>(x + (size_t)v * kFeat);
  float2* op = reinterpret_cast<float2*>(out + (size_t)v * 10 * kFeat);
  const size_t shift_stride = (size_t)n * kFeat2;
  for (int f2 = threadIdx.x; f2 < kFeat2; f2 += 256) {
    float2 acc[10];
#pragma unroll
    for (int k = 0; k < 10; ++k) acc[k] = make_float2(0.f, 0.f);
    if constexpr (S_STATIC > 0) {
      float2 xv[S_STATIC];
#pragma unroll
      for (int j = 0; j < S_STATIC; ++j) xv[j] = __ldg(xp + (size_t)j * shift_stride + f2);
#pragma unroll
      for (int j = 0; j < S_STATIC; ++j) {
#pragma unroll
        for (int k = 0; k < 10; ++k) {
          acc[k].x = fmaf(sw[j * 10 + k], xv[j].x, acc[k].x);
          acc[k].y = fmaf(sw[j * 10 + k], xv[j].y, acc[k].y);
        }
      }
    } else {
      for (int j = 0; j < n_shifts; ++j) {
        const float2 xv = __ldg(xp + (size_t)j * shift_stride + f2);
#pragma unroll
        for (int k = 0; k < 10; ++k) {
          acc[k].x = fmaf(sw[j * 10 + k], xv.x, acc[k].x);
          acc[k].y = fmaf(sw[j * 10 + k], xv.y, acc[k].y);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 10; ++k) op[(size_t)k * kFeat2 + f2] = acc[k];
  }
}

// Per-variant transform fused with gblinear scoring (predict.py:123-160): the 20020 features of a variant are
// formed in registers and immediately contracted with the model weights, so the float64[n,20020] feature
// matrices of the reference (480 KB per variant for diff/ref/alt) never exist.
// x [S, n, 2002]; w [m, 10*2002] (column k*2002+f; columns dropped by the keep mask are zero);
// offset [m] = base_score + bias; out [n, m].  Block = (variant, group of kScoreModels models).
constexpr int kScoreModels = 8;
__global__ void __launch_bounds__(256)
decay_score_variant_kernel(const float* __restrict__ x, const double* __restrict__ dist,
                           const int8_t* __restrict__ strand, const int32_t* __restrict__ shifts,
                           const float* __restrict__ w, const float* __restrict__ offset, int n, int n_shifts, int m,
                           float* __restrict__ out) {
  extern __shared__ float sw[];  // [S][10] then [8 warps][kScoreModels]
  float* red = sw + n_shifts * 10;
  const int v = blockIdx.x;
  const int m0 = blockIdx.y * kScoreModels;
  variant_weights_to_smem(sw, dist[v], strand[v], shifts, n_shifts);
  __syncthreads();
  float acc[kScoreModels];
#pragma unroll
  for (int i = 0; i < kScoreModels; ++i) acc[i] = 0.f;
  for (int f = threadIdx.x; f < kFeat; f += 256) {
    float feat[10];
#pragma unroll
    for (int k = 0; k < 10; ++k) feat[k] = 0.f;
    for (int j = 0; j < n_shifts; ++j) {
      const float xv = __ldg(x + ((size_t)j * n + v) * kFeat + f);
#pragma unroll
      for (int k = 0; k < 10; ++k) feat[k] = fmaf(sw[j * 10 + k], xv, feat[k]);
    }
#pragma unroll
    for (int i = 0; i < kScoreModels; ++i) {
      if (m0 + i < m) {
        const float* wi = w + (size_t)(m0 + i) * 10 * kFeat + f;
#pragma unroll
        for (int k = 0; k < 10; ++k) acc[i] = fmaf(feat[k], __ldg(wi + k * kFeat), acc[i]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < kScoreModels; ++i) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0)
#pragma unroll
    for (int i = 0; i < kScoreModels; ++i) red[warp * kScoreModels + i] = acc[i];
  __syncthreads();
  if (threadIdx.x < kScoreModels && m0 + threadIdx.x < m) {
    float s = 0.f;
    for (int wp = 0; wp < 8; ++wp) s += red[wp * kScoreModels + threadIdx.x];
    out[(size_t)v * m + m0 + threadIdx.x] = s + offset[m0 + threadIdx.x];
  }
}

}  // namespace bx
