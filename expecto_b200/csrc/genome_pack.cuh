// FASTA text resident in HBM -> 2-bit packed genome + zero-column mask (the layout of bx_ctx_set_genome).
// Replaces pyfasta's flat index build + string slicing (chromatin.py:44,205; compute_expecto_features.py:31,108)
// and encodeSeqs' per-character `mydict` lookup (chromatin.py:155-160) for the genome side of kernel 1.
//
// Byte / integer work, HBM-bound.  A contig body is the bytes between its header line and the next header; line
// ends (LF, CR) may sit anywhere (ragged line widths, CRLF files), so the position of a base in the contig is a
// prefix count of the non-line-end bytes before it:
//   pass 1  fasta_count_kernel   bases per 4096-byte block            (text read once, 8 B written per block)
//           cub::DeviceScan      exclusive sum over the blocks; the last entry is the contig length
//   pass 2  fasta_pack_kernel    block-local scan, the block's bases compacted in shared memory at their position
//                                modulo 32, then one thread per 16 bases: one bases2 word + half a mask word.
//                                Groups a block covers completely are plain stores; the (at most two) groups it
//                                shares with its neighbours are merged with atomicOr / atomicAnd into the planes the
//                                caller pre-filled (bases2 = 0, nmask = ones).  Characters outside mydict (the
//                                reference raises KeyError when a window touches them) pack as zero columns and
//                                are appended to a bounded list.
// Algorithmic bytes per base: 1 + 1 (text read by both passes) + 0.375 (planes written) = 2.375.
// Every load of the text is a 16-byte vector load on a 16-byte boundary of the file offset.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace bx {

constexpr int kFaThreads = 256;
constexpr int kFaBlockBytes = kFaThreads * 16;
constexpr uint8_t kFaEol = 254, kFaInvalid = 255;

// mydict (chromatin.py:155-160): A/a -> row 0, G/g -> 1, C/c -> 2, T/t -> 3; N, n, H, - -> all-zero column (4).
__device__ __forceinline__ uint8_t fasta_code_of(uint8_t c) {
  switch (c) {
    case 'A': case 'a': return 0;
    case 'G': case 'g': return 1;
    case 'C': case 'c': return 2;
    case 'T': case 't': return 3;
    case 'N': case 'n': case 'H': case '-': return 4;
    case '\n': case '\r': return kFaEol;
    default: return kFaInvalid;
  }
}

// The 16 text bytes of this thread in block `blk` (absolute offsets a0 + blk*4096 + tid*16 ...), with bytes outside
// the body [b0, b1) replaced by LF so they count as line ends.
__device__ __forceinline__ void fasta_load16(const uint8_t* __restrict__ text, long long a0, long long b0, long long b1,
                                             long long blk, uint8_t (&c)[16]) {
  const long long at = a0 + blk * kFaBlockBytes + (long long)threadIdx.x * 16;
  uint4 v = make_uint4(0x0a0a0a0au, 0x0a0a0a0au, 0x0a0a0a0au, 0x0a0a0a0au);
  if (at + 16 > b0 && at < b1) v = __ldg(reinterpret_cast<const uint4*>(text + at));
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  if (at >= b0 && at + 16 <= b1) {  // the common case: no per-byte bounds
#pragma unroll
    for (int i = 0; i < 16; ++i) c[i] = (uint8_t)(w[i >> 2] >> (8 * (i & 3)));
    return;
  }
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const long long p = at + i;
    c[i] = (p >= b0 && p < b1) ? (uint8_t)(w[i >> 2] >> (8 * (i & 3))) : (uint8_t)'\n';
  }
}

// Exclusive prefix of `v` over the block's 256 threads and the block total.
__device__ __forceinline__ int block_exclusive_scan(int v, int* s_warp, int& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  int before = 0, sum = 0;
#pragma unroll
  for (int w = 0; w < kFaThreads / 32; ++w) {
    const int x = s_warp[w];
    if (w < warp) before += x;
    sum += x;
  }
  total = sum;
  __syncthreads();
  return before + incl - v;
}

// Line-end bytes (LF, CR) among the 16 bytes of a vector, by byte-wise SIMD compares.
__device__ __forceinline__ int eol_count16(const uint4& v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  int n = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    n += __popc(__vcmpeq4(w[i], 0x0a0a0a0au) | __vcmpeq4(w[i], 0x0d0d0d0du));
  return n >> 3;  // 0xff per matching byte
}

// One WARP per 4096-byte block: eight independent 16-byte loads per lane in flight, a shuffle reduction, no block
// barrier (the kernel is a pure streaming read).
__global__ void __launch_bounds__(kFaThreads)
fasta_count_kernel(const uint8_t* __restrict__ text, long long a0, long long b0, long long b1, long long n_blocks,
                   long long* __restrict__ counts) {
  const int lane = threadIdx.x & 31;
  const long long warp0 = (long long)blockIdx.x * (kFaThreads / 32) + (threadIdx.x >> 5);
  const long long n_warps = (long long)gridDim.x * (kFaThreads / 32);
  for (long long blk = warp0; blk < n_blocks; blk += n_warps) {
    const long long base = a0 + blk * kFaBlockBytes;
    int n = 0;
    if (base >= b0 && base + kFaBlockBytes <= b1) {
      uint4 v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = __ldg(reinterpret_cast<const uint4*>(text + base) + j * 32 + lane);
#pragma unroll
      for (int j = 0; j < 8; ++j) n += 16 - eol_count16(v[j]);
    } else {  // first / last block of the body: byte-wise bounds
#pragma unroll 1
      for (int j = 0; j < 8; ++j) {
        const long long at = base + (long long)(j * 32 + lane) * 16;
        if (at + 16 > b0 && at < b1) {
          const uint4 q = __ldg(reinterpret_cast<const uint4*>(text + at));
          const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const uint32_t ch = (w[i >> 2] >> (8 * (i & 3))) & 0xffu;
            n += (at + i >= b0 && at + i < b1 && ch != '\n' && ch != '\r');
          }
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if (lane == 0) counts[blk] = n;
  }
}

constexpr uint8_t kFaUncovered = 253;  // a slot of a boundary group that belongs to a neighbouring block

__global__ void __launch_bounds__(kFaThreads)
fasta_pack_kernel(const uint8_t* __restrict__ text, long long a0, long long b0, long long b1, long long n_blocks,
                  const long long* __restrict__ offsets, long long base_off, uint32_t* __restrict__ bases2,
                  uint32_t* __restrict__ nmask, long long* __restrict__ invalid, long long invalid_cap,
                  unsigned long long* __restrict__ invalid_count) {
  __shared__ uint8_t s_tab[256];
  __shared__ int s_warp[kFaThreads / 32];
  __shared__ __align__(16) uint8_t s_codes[kFaBlockBytes + 64];
  s_tab[threadIdx.x] = fasta_code_of((uint8_t)threadIdx.x);
  __syncthreads();
  for (long long blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
    uint8_t c[16];
    fasta_load16(text, a0, b0, b1, blk, c);
    uint8_t k[16];
    int n = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      k[i] = s_tab[c[i]];
      n += k[i] != kFaEol;
    }
    int total;
    const int local = block_exclusive_scan(n, s_warp, total);
    if (total == 0) continue;  // uniform over the block
    const long long G = base_off + offsets[blk];  // global index of the block's first base
    const int lead = (int)(G & 31);
    const int n_groups = (lead + total + 31) >> 5;
    // compacted codes at their position modulo 32, so that group t is s_codes[32t .. 32t+31]
    if (threadIdx.x < lead) s_codes[threadIdx.x] = kFaUncovered;
    if (threadIdx.x < 32 && lead + total + (int)threadIdx.x < n_groups * 32)
      s_codes[lead + total + threadIdx.x] = kFaUncovered;
    int at = lead + local;
#pragma unroll
    for (int i = 0; i < 16; ++i)
      if (k[i] != kFaEol) s_codes[at++] = k[i];
    __syncthreads();
    // One thread per HALF group (16 bases -> one bases2 word and 16 mask bits); the two halves of a group sit in
    // neighbouring lanes and combine their mask bits with one shuffle.
    // (lead + total can reach 129 groups = 258 halves: the loop runs twice for such a block, uniformly.)
    for (int t0 = 0; t0 < 2 * n_groups; t0 += kFaThreads) {
    const int t = t0 + (int)threadIdx.x;
    const int half = t & 1;
    const long long g = (G >> 5) + (t >> 1);
    uint32_t two = 0u, base_bits = 0u, covered = 0u;
    const bool active = (t >> 1) < n_groups;
    if (active) {
      const uint4 q = *reinterpret_cast<const uint4*>(s_codes + 16 * t);
      const uint32_t w[4] = {q.x, q.y, q.z, q.w};
      if (((w[0] | w[1] | w[2] | w[3]) & 0xfcfcfcfcu) == 0u) {  // sixteen plain bases: pack four codes per step
#pragma unroll
        for (int j = 0; j < 4; ++j) two |= ((w[j] | (w[j] >> 6) | (w[j] >> 12) | (w[j] >> 18)) & 0xffu) << (8 * j);
        base_bits = covered = 0xffffu;
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const uint32_t code = (w[i >> 2] >> (8 * (i & 3))) & 0xffu;
          if (code == kFaUncovered) continue;
          covered |= 1u << i;
          if (code < 4) {
            base_bits |= 1u << i;
            two |= code << (2 * i);
          } else if (code == kFaInvalid) {
            const unsigned long long slot = atomicAdd(invalid_count, 1ull);
            if ((long long)slot < invalid_cap) invalid[slot] = g * 32 + half * 16 + i;
          }
        }
      }
      if (covered == 0xffffu) bases2[2 * g + half] = two;
      else if (two) atomicOr(bases2 + 2 * g + half, two);  // shared with a neighbouring block / the contig's tail
    }
    const uint32_t my_bits = base_bits << (16 * half), my_cov = covered << (16 * half);
    const uint32_t other_bits = __shfl_xor_sync(0xffffffffu, my_bits, 1);
    const uint32_t other_cov = __shfl_xor_sync(0xffffffffu, my_cov, 1);
    if (active) {
      if ((my_cov | other_cov) == 0xffffffffu) {
        if (half == 0) nmask[g] = ~(my_bits | other_bits);
      } else if (my_bits) {
        atomicAnd(nmask + g, ~my_bits);
      }
    }
    }
    __syncthreads();
  }
}

// '>' at a line start (file offset 0 or after LF): the contig headers.  Unordered; the host sorts.
__global__ void __launch_bounds__(256)
fasta_headers_kernel(const uint8_t* __restrict__ text, long long n_text, long long* __restrict__ pos, long long cap,
                     unsigned long long* __restrict__ count) {
  const long long n_vec = (n_text + 15) >> 4;
  for (long long v = (long long)blockIdx.x * 256 + threadIdx.x; v < n_vec; v += (long long)gridDim.x * 256) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(text) + v);
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
    uint32_t hits = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i)
      if (((w[i >> 2] >> (8 * (i & 3))) & 0xffu) == (uint32_t)'>') hits |= 1u << i;
    while (hits) {
      const int i = __ffs(hits) - 1;
      hits &= hits - 1;
      const long long p = v * 16 + i;
      if (p < n_text && (p == 0 || text[p - 1] == '\n')) {
        const unsigned long long slot = atomicAdd(count, 1ull);
        if ((long long)slot < cap) pos[slot] = p;
      }
    }
  }
}

}  // namespace bx
