// Kernel 1's core: the channel code of one position of a window / segment read straight from the 2-bit packed
// genome (reference: fetchSeqs chromatin.py:175-209 + encodeSeqs chromatin.py:138-172, as index arithmetic).
// Shared by the stand-alone encode kernels (kernels_misc.cuh) and by the GEMM producer that builds conv1's
// operand tile in shared memory (gemm_tc.cuh, A_MODE != 0).
#pragma once
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include "../../include/beluga_b200.h"

namespace bx {

// 16-bit 1.0 in the operand format.
template <typename T> __device__ __forceinline__ uint32_t one_bits();
template <> __device__ __forceinline__ uint32_t one_bits<__half>() { return 0x3C00u; }
template <> __device__ __forceinline__ uint32_t one_bits<__nv_bfloat16>() { return 0x3F80u; }

// Channel code (A=0,G=1,C=2,T=3; 4 = all-zero column) of forward-strand position i of a window.
__device__ __forceinline__ int window_code(const bx_window_desc& d, int i, const uint32_t* __restrict__ bases2,
                                           const uint32_t* __restrict__ nmask, const uint8_t* __restrict__ alleles,
                                           int64_t n_bases) {
  int64_t g;
  if (i < d.len1) {
    g = d.g1 + i;
  } else if (i < d.len1 + d.alen) {
    return alleles[d.aoff + (i - d.len1)];
  } else if (i < d.len1 + d.alen + d.len2) {
    g = d.g2 + (i - d.len1 - d.alen);
  } else {
    return 4;
  }
  if (g < 0 || g >= n_bases) return 4;
  if ((__ldg(nmask + (g >> 5)) >> (g & 31)) & 1u) return 4;
  return (__ldg(bases2 + (g >> 4)) >> ((g & 15) * 2)) & 3u;
}

// Strand-oriented code of position i of a segment of length L described like a window (bx_window_desc).
__device__ __forceinline__ int segment_code(const bx_window_desc& d, int i, int L, const uint32_t* __restrict__ bases2,
                                            const uint32_t* __restrict__ nmask, const uint8_t* __restrict__ alleles,
                                            int64_t n_bases) {
  if (i < 0 || i >= L) return 4;
  const bool rc = d.flags & 1;
  int c = window_code(d, rc ? (L - 1 - i) : i, bases2, nmask, alleles, n_bases);
  if (rc && c < 4) c = 3 - c;
  return c;
}

// Where the conv1 operand comes from when the GEMM producer builds it in shared memory instead of TMA-loading
// an im2col matrix: window / segment descriptors over the packed genome, or an explicit one-hot tensor.
struct EncParams {
  const bx_window_desc* descs;   // [n_seg] (A_MODE 1)
  const uint8_t* alleles;
  const uint32_t* bases2;
  const uint32_t* nmask;
  long long n_bases;
  const uint8_t* onehot;         // uint8 [n_seg, 4, L] (A_MODE 2)
  int L;                         // positions per window / segment
  int n_seg;
};

// Codes of the 144 operand slots of one 128-row conv1 tile, five slots per lane (slot = lane + 32 k), for the
// GEMM producer warp.  Tile rows j = 0..127 are (window wA, position lA + j) up to the window's last row and
// (wA + 1, j - jB) after it; slots [0, jB + 7) hold wA's positions lA.., slots [jB + 8, 143) the next window's
// positions 0...  Same arithmetic as segment_code, split in two so that the global loads of a tile are all
// issued back to back (enc_tile_load: addresses clamped, no load depends on another slot's) and consumed one
// tile later (enc_tile_codes), with the previous tile's shared-memory work and the weight TMA issue in between.
template <int A_MODE>
struct EncRaw {
  uint32_t nm[5], bs[5];   // A_MODE 1: N-mask word, 2-bit bases word; A_MODE 2: the four one-hot bytes packed
  int al[5], sh[5];        // allele code (4 = none), bit index of the base within its mask word
  uint32_t useg, rcv, ok;  // per-slot bit masks
};

template <int A_MODE>
__device__ __forceinline__ void enc_tile_load(const EncParams& e, int wA, int lA, int jB, int lane, EncRaw<A_MODE>& r) {
  const bool hasA = wA < e.n_seg, hasB = wA + 1 < e.n_seg;
  r.useg = r.rcv = r.ok = 0;
  if constexpr (A_MODE == 1) {
    const bx_window_desc dA = e.descs[hasA ? wA : 0];       // warp-uniform loads
    const bx_window_desc dB = e.descs[hasB ? wA + 1 : 0];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      const int i = lane + 32 * k;
      const bool first = i < jB + 7;
      const int pos = first ? lA + i : i - 8 - jB;
      const bool ok = (first ? hasA : hasB) && pos >= 0 && pos < e.L && i < 144;
      const int len1 = first ? dA.len1 : dB.len1, alen = first ? dA.alen : dB.alen;
      const int len2 = first ? dA.len2 : dB.len2;
      const int64_t g1 = first ? dA.g1 : dB.g1, g2 = first ? dA.g2 : dB.g2, aoff = first ? dA.aoff : dB.aoff;
      const bool rc = (first ? dA.flags : dB.flags) & 1;
      const int i2 = rc ? (e.L - 1 - pos) : pos;
      const bool in1 = i2 < len1;
      const bool inA = !in1 && i2 < len1 + alen;
      const bool in2 = !in1 && !inA && i2 < len1 + alen + len2;
      const int64_t g = in1 ? g1 + i2 : g2 + (i2 - len1 - alen);
      const bool useg = ok && (in1 || in2) && g >= 0 && g < e.n_bases;
      r.useg |= (uint32_t)useg << k;
      r.rcv |= (uint32_t)rc << k;
      const int64_t gg = useg ? g : 0;
      r.nm[k] = __ldg(e.nmask + (gg >> 5));
      r.bs[k] = __ldg(e.bases2 + (gg >> 4));
      r.sh[k] = (int)(gg & 31);
      r.al[k] = (ok && inA) ? (int)__ldg(e.alleles + aoff + (i2 - len1)) : 4;
    }
  } else {
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      const int i = lane + 32 * k;
      const bool first = i < jB + 7;
      const int pos = first ? lA + i : i - 8 - jB;
      const bool ok = (first ? hasA : hasB) && pos >= 0 && pos < e.L && i < 144;
      r.ok |= (uint32_t)ok << k;
      const uint8_t* x = e.onehot + (ok ? (size_t)(first ? wA : wA + 1) * 4 * e.L + pos : 0);
      r.nm[k] = __ldg(x);
      r.bs[k] = __ldg(x + (size_t)e.L);
      r.al[k] = __ldg(x + (size_t)2 * e.L);
      r.sh[k] = __ldg(x + (size_t)3 * e.L);
    }
  }
}

template <int A_MODE>
__device__ __forceinline__ void enc_tile_codes(const EncRaw<A_MODE>& r, int (&codes)[5]) {
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    if constexpr (A_MODE == 1) {
      int c = r.al[k];
      if ((r.useg >> k) & 1u) c = ((r.nm[k] >> r.sh[k]) & 1u) ? 4 : (int)((r.bs[k] >> ((r.sh[k] & 15) * 2)) & 3u);
      if (((r.rcv >> k) & 1u) && c < 4) c = 3 - c;  // A<->T, G<->C: encodeSeqs' [:, ::-1, ::-1]
      codes[k] = c;
    } else {
      codes[k] = !((r.ok >> k) & 1u) ? 4 : r.nm[k] ? 0 : r.bs[k] ? 1 : r.al[k] ? 2 : r.sh[k] ? 3 : 4;
    }
  }
}

}  // namespace bx
