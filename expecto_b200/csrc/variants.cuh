// Device-side descriptor builders of the variant path: the index arithmetic of fetchSeqs (chromatin.py:175-209:
// 2100-bp fetch around pos + shift, alleles spliced in at mutpos = 1049 - shift) and encodeSeqs' centre crop
// (chromatin.py:164), evaluated per variant on the GPU from whole-VCF column arrays, so that the host hands over the
// columns once and never loops over variants.  The host restatement of the same arithmetic (expecto_b200/windows.py:
// variant_descs, _allele_segment, sub_descs) is the oracle of these kernels: tests compare them descriptor for
// descriptor.
#pragma once
#include <cstdint>
#include "../../include/beluga_b200.h"

namespace bx {

// Whole-VCF columns resident in HBM (bx_chromatin_open).
struct VariantCols {
  const long long* gpos;   // global packed-genome index of the variant's first base: contig offset + pos - 1
  const int* ref_len;      // len(ref): the span that is spliced out (chromatin.py:209)
  const int* alt_len;
  const int* ref_aoff;     // offsets of the ref / alt allele codes in the allele table
  const int* alt_aoff;
};

__device__ __forceinline__ long long floor_div2(long long x) { return x >= 0 ? x / 2 : -((-x + 1) / 2); }

// One strand-oriented segment of 2000 + 2*maxshift positions holding every shifted window of one allele
// (windows.py: _allele_segment): left flank, allele, right flank.
__device__ __forceinline__ bx_window_desc allele_segment(long long v, int ref_len, int la, int aoff, int maxshift,
                                                         bool rc) {
  const int seg_len = 2000 + 2 * maxshift;
  const int a = (int)floor_div2(100 + (la - ref_len));   // crop start, chromatin.py:164
  const int len1 = maxshift + 1049 - a;
  bx_window_desc d;
  d.g1 = v - len1;
  d.len1 = len1;
  d.alen = la;
  d.aoff = aoff;
  d.g2 = v + ref_len;
  d.len2 = seg_len - len1 - la;
  d.flags = rc ? 3 : 0;   // rc segments: reversed strand AND reversed offset order
  d.reserved = 0;
  return d;
}

// Strand-oriented positions [start, start + length) of a segment (windows.py: sub_descs).
__device__ __forceinline__ bx_window_desc sub_segment(const bx_window_desc& d, int start, int length, int seg_len) {
  const bool rc = d.flags & 1;
  const long long f0 = rc ? (long long)seg_len - start - length : start;   // first forward-strand position kept
  const long long a0 = d.len1, a1 = (long long)d.len1 + d.alen;
  bx_window_desc o = d;
  long long len1 = a0 - f0;
  len1 = len1 < 0 ? 0 : (len1 > length ? length : len1);
  long long alen = (a1 < f0 + length ? a1 : f0 + length) - (a0 > f0 ? a0 : f0);
  if (alen < 0) alen = 0;
  o.g1 = d.g1 + f0;
  o.len1 = (int)len1;
  o.alen = (int)alen;
  o.aoff = d.aoff + (int)(f0 - a0 > 0 ? f0 - a0 : 0);
  o.g2 = d.g2 + (f0 - a1 > 0 ? f0 - a1 : 0);
  o.len2 = (int)(length - len1 - alen);
  return o;
}

// Four full segments per group member: {ref fwd, ref rc, alt fwd, alt rc} (windows.py: chromatin_segment_descs).
__global__ void build_segment_descs_kernel(VariantCols c, const int* __restrict__ idx, long long v0, int n,
                                           int maxshift, bx_window_desc* __restrict__ out) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n) return;
  const long long i = v0 + idx[g];
  const long long v = c.gpos[i];
  const int rl = c.ref_len[i], al = c.alt_len[i];
  out[4 * g + 0] = allele_segment(v, rl, rl, c.ref_aoff[i], maxshift, false);
  out[4 * g + 1] = allele_segment(v, rl, rl, c.ref_aoff[i], maxshift, true);
  out[4 * g + 2] = allele_segment(v, rl, al, c.alt_aoff[i], maxshift, false);
  out[4 * g + 3] = allele_segment(v, rl, al, c.alt_aoff[i], maxshift, true);
}

// Delta mode (substitutions): two full ref segments and two alt patches of patch_len positions starting
// 16 * patch_row0 positions into the alt segments (windows.py: chromatin_delta_descs).
__global__ void build_delta_descs_kernel(VariantCols c, const int* __restrict__ idx, long long v0, int n, int maxshift,
                                         int patch_row0, int patch_len, bx_window_desc* __restrict__ ref_out,
                                         bx_window_desc* __restrict__ patch_out) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n) return;
  const long long i = v0 + idx[g];
  const long long v = c.gpos[i];
  const int rl = c.ref_len[i], al = c.alt_len[i];
  const int seg_len = 2000 + 2 * maxshift;
  ref_out[2 * g + 0] = allele_segment(v, rl, rl, c.ref_aoff[i], maxshift, false);
  ref_out[2 * g + 1] = allele_segment(v, rl, rl, c.ref_aoff[i], maxshift, true);
  patch_out[2 * g + 0] = sub_segment(allele_segment(v, rl, al, c.alt_aoff[i], maxshift, false), 16 * patch_row0,
                                     patch_len, seg_len);
  patch_out[2 * g + 1] = sub_segment(allele_segment(v, rl, al, c.alt_aoff[i], maxshift, true), 16 * patch_row0,
                                     patch_len, seg_len);
}

// Python slice semantics of seq[:x] / seq[x:] on a string of n characters: negative indices count from the end.
__device__ __forceinline__ long long py_norm(long long x, long long n) {
  return x < 0 ? (n + x > 0 ? n + x : 0) : (x < n ? x : n);
}

// One window of fetchSeqs + encodeSeqs for one allele at one shift (windows.py: variant_descs), with the
// reference's literal slice semantics beyond |shift| = 1049 and for deletions longer than the 100-bp margin.
__device__ __forceinline__ bx_window_desc variant_window(long long v, int ref_len, int la, int aoff, int shift,
                                                         bool rc) {
  const long long W = 2100, I = 2000;
  const long long mutpos = 1049 - shift;               // chromatin.py:203
  const long long s0 = v + shift - 1049;                // global index of seq[0]
  const long long len_p = py_norm(mutpos, W);           // seq[:mutpos]
  const long long suffix0 = py_norm(mutpos + ref_len, W);
  const long long len_s = W - suffix0;                  // seq[mutpos + len(ref):]
  const long long total = len_p + la + len_s;
  // line[floor((len - 2000) / 2) : floor(len - (len - 2000) / 2)], chromatin.py:164
  long long a = floor_div2(total - I), b = floor_div2(total + I);
  if (a < 0) a = (a + total > 0) ? a + total : 0;
  if (a > total) a = total;
  b = b < 0 ? 0 : (b > total ? total : b);
  if (b < a) b = a;
  auto mn = [](long long x, long long y) { return x < y ? x : y; };
  auto clip = [](long long x, long long lo, long long hi) { return x < lo ? lo : (x > hi ? hi : x); };
  bx_window_desc d;
  const long long p_lo = mn(a, len_p), p_hi = mn(b, len_p);
  d.len1 = (int)(p_hi - p_lo > 0 ? p_hi - p_lo : 0);
  d.g1 = s0 + p_lo;
  const long long al_lo = clip(a - len_p, 0, la), al_hi = clip(b - len_p, 0, la);
  d.alen = (int)(al_hi - al_lo > 0 ? al_hi - al_lo : 0);
  d.aoff = aoff + (int)al_lo;
  const long long s_lo = clip(a - len_p - la, 0, len_s), s_hi = clip(b - len_p - la, 0, len_s);
  d.len2 = (int)(s_hi - s_lo > 0 ? s_hi - s_lo : 0);
  d.g2 = s0 + suffix0 + s_lo;
  d.flags = rc ? 1 : 0;
  d.reserved = 0;
  return d;
}

// The four windows {ref fwd, ref rc, alt fwd, alt rc} of each group member at one shift
// (windows.py: chromatin_window_descs).
__global__ void build_window_descs_kernel(VariantCols c, const int* __restrict__ idx, long long v0, int n, int shift,
                                          bx_window_desc* __restrict__ out) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n) return;
  const long long i = v0 + (idx ? idx[g] : g);
  const long long v = c.gpos[i];
  const int rl = c.ref_len[i], al = c.alt_len[i];
  out[4 * g + 0] = variant_window(v, rl, rl, c.ref_aoff[i], shift, false);
  out[4 * g + 1] = variant_window(v, rl, rl, c.ref_aoff[i], shift, true);
  out[4 * g + 2] = variant_window(v, rl, al, c.alt_aoff[i], shift, false);
  out[4 * g + 3] = variant_window(v, rl, al, c.alt_aoff[i], shift, true);
}

}  // namespace bx
