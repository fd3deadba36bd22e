/* beluga_b200 — C ABI of the B200-native Beluga chromatin-effect hot path.
 *
 * The reference (richardshuai/ExPecto) has no FFI: its boundary is Python scripts + files.  This
 * header is what a binding for that path would call; each entry point cites the reference code it
 * replaces.  All functions return 0 on success or a negative code (see BX_E_*); bx_last_error()
 * gives the message.  No exceptions cross the boundary, the caller owns every buffer, device
 * pointers are plain `void*` (e.g. torch.Tensor.data_ptr()), streams are cudaStream_t as `void*`.
 * Entry points that take a context run on that context's device and restore the caller's current device before
 * returning; the context-free ones (bx_chromatin_outputs*, bx_strand_average, bx_decay_*, bx_fasta_*) launch on the
 * caller's current device, which must own the buffers.
 * There is NO CPU fallback: every compute entry point fails with BX_E_CUDA when no sm_100 device
 * is present.
 */
#ifndef BELUGA_B200_H
#define BELUGA_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define BX_OK 0
#define BX_E_ARG (-1)     /* invalid argument */
#define BX_E_CUDA (-2)    /* CUDA runtime/driver error, or no sm_100 device */
#define BX_E_STATE (-3)   /* weights / genome not set */
#define BX_E_BASE (-4)    /* a window touched a base the reference's encodeSeqs would raise KeyError on */

#define BX_DTYPE_F16 0
#define BX_DTYPE_BF16 1
/* fp16 operands split into (hi, lo) pairs: every product is Ah*Wh + Ah*Wl + Al*Wh with fp32 accumulation (three
 * tcgen05.mma per K step), which carries ~21 significant bits of both operands through all eight layers.  Three
 * times the tensor work of BX_DTYPE_F16; for callers that need the reference's fp32 results to ~1e-5 on the
 * probabilities rather than the 2e-3 bar. */
#define BX_DTYPE_F16X3 2

#define BX_WINDOW 2000     /* Beluga input length, chromatin.py:35-36 */
#define BX_FEATURES 2002   /* Beluga outputs, Beluga.py:45 */
#define BX_NBASIS 10       /* decay bases, predict.py:88-109 */
#define BX_NSTAGES 10      /* timed stages: 0 encode, 1..6 conv1..6, 7 fc1, 8 fc2, 9 ref/alt/diff outputs */

/* One 2000-bp model input, described as  genome[g1 : g1+len1] + allele[aoff : aoff+alen] +
 * genome[g2 : g2+len2]  (positions past the total are all-zero columns).  This is the result of
 * fetchSeqs' splice (chromatin.py:175-209) followed by encodeSeqs' centre crop (chromatin.py:164),
 * or a plain TSS window (compute_expecto_features.py:107-111, alen = len2 = 0).
 * g1/g2 index the packed genome set with bx_ctx_set_genome.  flags bit 0: emit the reverse
 * complement (encodeSeqs' [:, ::-1, ::-1], chromatin.py:170). */
typedef struct bx_window_desc {
  int64_t g1;
  int64_t g2;
  int32_t len1;
  int32_t alen;
  int32_t len2;
  int32_t aoff;
  int32_t flags;
  int32_t reserved;
} bx_window_desc;

typedef struct bx_ctx bx_ctx;

int bx_version(void);
const char* bx_last_error(void);
/* Number of kernels this library has launched since load (bench.py's gpu_launches). */
int64_t bx_launch_count(void);

/* One context per GPU; a context is single-threaded.  capacity_windows bounds the windows per
 * forward call chunk (activation scratch is sized once here).  cta_group: 1 or 2 (CTA pairs). */
int bx_ctx_create(int device, int dtype, int cta_group, int capacity_windows, bx_ctx** out);
int bx_ctx_destroy(bx_ctx* ctx);

/* Weights in the reference's state_dict order and layouts (Beluga.py:23-45), HOST fp32:
 * w[0..5]  conv weights [Cout,Cin,1,8], b[0..5] conv biases,
 * w[6] Linear(67840,2003).weight [2003,67840] (flatten index c*106+l), b[6] its bias,
 * w[7] Linear(2003,2002).weight [2002,2003], b[7] its bias.
 * Replaces model.load_state_dict(...) + .cuda(), chromatin.py:102-106. */
int bx_ctx_load_weights(bx_ctx* ctx, const float* const w[8], const float* const b[8]);

/* Packed genome resident in HBM: bases2 holds 16 bases per uint32 (2 bits each, code A=0 G=1 C=2 T=3,
 * base i in bits 2*(i%16)); nmask holds 1 bit per base (1 = all-zero column: N, n, H, -), 32 per uint32.
 * Replaces pyfasta.Fasta('./resources/hg19.fa'), chromatin.py:44. */
int bx_ctx_set_genome(bx_ctx* ctx, const uint32_t* bases2_dev, const uint32_t* nmask_dev, int64_t n_bases);

/* encodeSeqs (chromatin.py:138-172 / expecto_utils.py:5-39), materialised: out_dev is uint8 [n,4,2000],
 * channel order A,G,C,T.  alleles_dev: codes 0..3 = A,G,C,T, 4 = zero column. */
int bx_encode_windows(bx_ctx* ctx, const bx_window_desc* descs_dev, const uint8_t* alleles_dev, int n,
                      uint8_t* out_dev, void* stream);

/* Beluga.forward (Beluga.py:50-51) over n windows, encode fused in front: probs_dev is fp32 [n,2002]. */
int bx_forward_windows(bx_ctx* ctx, const bx_window_desc* descs_dev, const uint8_t* alleles_dev, int n,
                       float* probs_dev, void* stream);
/* Same, from an explicit one-hot uint8 [n,4,2000] (the tensor chromatin.py:267 feeds the model). */
int bx_forward_onehot(bx_ctx* ctx, const uint8_t* onehot_dev, int n, float* probs_dev, void* stream);

/* Host-buffer form of bx_forward_onehot — the call chromatin.py:266-270 makes per batch
 * (torch.from_numpy(...).cuda() -> model.forward -> .cpu()): onehot_host uint8 [n,4,2000] and probs_host fp32
 * [n,2002] are HOST memory (pinned memory makes the copies asynchronous).  Chunks of `capacity_windows` are
 * pipelined over three streams (H2D of chunk i+1 and D2H of chunk i-1 overlap the compute of chunk i); the
 * call returns when probs_host is complete. */
int bx_forward_onehot_host(bx_ctx* ctx, const uint8_t* onehot_host, int n, float* probs_host);

/* One iteration of chromatin.py:243-286 for n_var variants from HOST descriptors (4 per variant, order
 * {ref fwd, ref rc, alt fwd, alt rc}) to HOST ref/alt/diff arrays: fp32 [2*n_var,2002] each (forward rows
 * then reverse-complement rows, diff = alt - ref), or, with average != 0, the strand-averaged fp32
 * [n_var,2002] predict.py:186-190 derives from them.  Same three-stream pipeline as above. */
int bx_chromatin_shift_host(bx_ctx* ctx, const bx_window_desc* descs_host, const uint8_t* alleles_host,
                            int64_t n_alleles, int n_var, int average, float* ref_host, float* alt_host,
                            float* diff_host);

/* Per-stage device timing with CUDA events on the launching stream (bench.py's roofline numbers).
 * bx_ctx_profile_read fills ms_out[BX_NSTAGES] (summed milliseconds) and count_out[BX_NSTAGES] (launch groups). */
int bx_ctx_profile(bx_ctx* ctx, int enable);
int bx_ctx_profile_read(bx_ctx* ctx, double* ms_out, int64_t* count_out, int reset);

/* Copy an internal activation (test hook): layer 0 = im2col one-hot, 1..6 = conv outputs, 7 = fc1 hidden.
 * Writes rows*cols 16-bit elements for n windows to dst_dev; returns the element count through *count. */
int bx_debug_activation(bx_ctx* ctx, int layer, int n, void* dst_dev, int64_t* count, void* stream);

/* Re-run one stage on the scratch left by the last forward (profiling hook): 0..5 = conv1..conv6,
 * 6 = fc1 (+ReLU), 7 = fc2 (+sigmoid, writes probs_dev fp32 [n,2002]). */
int bx_debug_run_layer(bx_ctx* ctx, int layer, int n, float* probs_dev, void* stream);

/* chromatin.py:281-286 layout: probs_dev is fp32 [4*n_var,2002] in window order
 * (variant, {ref fwd, ref rc, alt fwd, alt rc}); writes ref/alt/diff fp32 [2*n_var,2002]
 * (rows 0..n_var-1 forward, n_var..2n_var-1 reverse complement), diff = alt - ref. */
int bx_chromatin_outputs(const float* probs_dev, int n_var, float* ref_dev, float* alt_dev, float* diff_dev,
                         void* stream);
/* predict.py:186-190 fused on device: strand-averaged diff/ref/alt fp32 [n_var,2002] from the same probs. */
int bx_chromatin_outputs_avg(const float* probs_dev, int n_var, float* ref_dev, float* alt_dev, float* diff_dev,
                             void* stream);

/* predict.py:183-190: x_dev fp32 [2n,2002] (forward rows then reverse-complement rows, as chromatin.py writes
 * them) -> out_dev fp32 [n,2002] = (x[0:n] + x[n:2n]) / 2.0. */
int bx_strand_average(const float* x_dev, int n, float* out_dev, void* stream);

/* compute_expecto_features.py:123-124: probs_dev fp32 [n_genes, S, 2, 2002] (shift-major, fwd then rc),
 * weights_dev fp32 [10,S]; out_dev fp32 [n_genes, 10*2002] = sum_s w[k,s] * 0.5*(fwd+rc)[s,f]. */
int bx_decay_transform_fixed(const float* probs_dev, const float* weights_dev, int n_genes, int n_shifts,
                             float* out_dev, void* stream);
/* predict.py:87-136 (compute_effects): x_dev fp32 [S, n, 2002]; dist_dev fp64 [n] (variant - TSS);
 * strand_dev int8 [n] (+1/-1); shifts_dev int32 [S]; out_dev fp32 [n, 10*2002]. */
int bx_decay_transform_variant(const float* x_dev, const double* dist_dev, const int8_t* strand_dev,
                               const int32_t* shifts_dev, int n, int n_shifts, float* out_dev, void* stream);

/* Segment mode — the whole shift loop of one variant / gene with shared activations.  The 9 windows of
 * chromatin.py:243 (or the 200 of compute_expecto_features.py:88) are 2000-bp substrings, 200 bp apart, of one
 * strand-oriented sequence of seg_len = 2000 + 200*(n_shifts-1) positions; the conv stack runs once per segment
 * and only the dense layers run per window.  Results are bit-identical to bx_forward_windows on the same windows.
 * descs_dev: n_groups*segs_per_group segment descriptors (same struct, len1+alen+len2 = seg_len; flags bit 0 =
 * reverse complement, bit 1 = window offsets run backwards: window of shift index i starts at 200*(n_shifts-1-k)
 * instead of 200*k, k = k_of_shift[i]).  probs_dev fp32 [n_windows, 2002] ordered (shift, group, seg) when
 * shift_major != 0 (chromatin.py's per-shift files) or (group, shift, seg) (compute_expecto_features.py). */
/* How many seg_len-position segments fit one chunk of the context's activation scratch (>= 0), or a negative code. */
int bx_segment_capacity(bx_ctx* ctx, int seg_len);
int bx_forward_segments(bx_ctx* ctx, const bx_window_desc* descs_dev, const uint8_t* alleles_dev, int n_groups,
                        int segs_per_group, int seg_len, int n_shifts, const int32_t* k_of_shift, int shift_major,
                        float* probs_dev, void* stream);

/* Segment mode for substitutions (len(ref) == len(alt)): the alt sequence differs from the ref sequence in a few
 * positions at the segment centre, so only the conv6 rows whose 310-position receptive field reaches them can
 * differ.  ref_descs_dev: 2 segments per group (ref fwd, ref rc), run in full; patch_descs_dev: 2 per group (alt
 * fwd, alt rc), patch_len positions (multiple of 16) that start 16*patch_row0 positions into the alt segments and
 * must cover every affected row's receptive field in both pooling phases.  Alt windows take those rows from the
 * patch and all others from the ref segment.  Output as bx_forward_segments with segs_per_group = 4 (ref fwd, ref
 * rc, alt fwd, alt rc), bit-identical to it (reference: the ref/alt loop bodies of chromatin.py:262-279). */
int bx_forward_segments_delta(bx_ctx* ctx, const bx_window_desc* ref_descs_dev, const bx_window_desc* patch_descs_dev,
                              const uint8_t* alleles_dev, int n_groups, int seg_len, int patch_len, int patch_row0,
                              int n_shifts, const int32_t* k_of_shift, int shift_major, float* probs_dev,
                              void* stream);

/* predict.py:123-160 fused: per-variant decay features contracted with n_models gblinear weight vectors
 * (xgboost Booster.predict = base_score + bias + w.x) without materialising the [n,20020] feature matrix.
 * w_dev fp32 [n_models, 10*2002] (columns removed by cluster_utils.get_keep_mask set to 0), offset_dev fp32
 * [n_models] = base_score + bias, out_dev fp32 [n, n_models]. */
int bx_decay_score_variant(const float* x_dev, const double* dist_dev, const int8_t* strand_dev,
                            const int32_t* shifts_dev, const float* w_dev, const float* offset_dev, int n,
                            int n_shifts, int n_models, float* out_dev, void* stream);

/* ---- The variant loop of chromatin.py:243-286 inside the library -------------------------------------------------
 * The reference walks the VCF row by row in Python (chromatin.py:248-254: fetchSeqs per variant and shift, then
 * encodeSeqs, two batched forwards, diff = alt - ref, one HDF5 file per shift).  Here the caller hands over whole-VCF
 * column arrays once (HOST memory, one entry per variant after chromatin.py:240-241's filter):
 *   gpos      global index, in the packed genome of bx_ctx_set_genome, of the variant's first base
 *             (contig offset + POS - 1);
 *   ref_len / alt_len    len(REF), len(ALT);  ref_aoff / alt_aoff   offsets of their codes in `alleles`
 *             (0..3 = A,G,C,T, 4 = zero column: the values of chromatin.py:155-160's mydict).
 * The library classifies every variant (substitutions of up to 16 bases -> delta mode; other variants whose shifted
 * windows are substrings of one spliced sequence -> four full segments; the rest -> window by window with the
 * reference's literal slice semantics), builds all descriptors on the device, and runs chunks of *chunk_variants_out
 * consecutive variants.  FC2's epilogue writes each chunk's ref / alt / diff = alt - ref in chromatin.py:281-286's
 * strand-stacked order: [S][2][m][2002] floats per tensor, S shifts in chromatin.py:243's order
 * (0, -200 .. -maxshift, 200 .. maxshift), forward rows then reverse-complement rows, m variants of the chunk.
 * Results are bit-identical to bx_forward_windows + bx_chromatin_outputs on the same variants.
 * flags: BX_CHROM_NO_DELTA  substitutions run four full segments (same bits; for comparisons);
 *        BX_CHROM_AVERAGE   outputs are [S][m][2002], averaged over the two strands as predict.py:186-190 reads them:
 *                           ref = (fwd + rc) / 2, alt likewise, diff = ((alt_f - ref_f) + (alt_r - ref_r)) / 2;
 *        BX_CHROM_WINDOWS   every variant window by window (the reference's schedule).
 * One plan per context; bx_chromatin_open replaces an open plan. */
#define BX_CHROM_NO_DELTA 1
#define BX_CHROM_AVERAGE 2
#define BX_CHROM_WINDOWS 4
int bx_chromatin_open(bx_ctx* ctx, int64_t n, const int64_t* gpos, const int32_t* ref_len, const int32_t* alt_len,
                      const int32_t* ref_aoff, const int32_t* alt_aoff, const uint8_t* alleles, int64_t n_alleles,
                      int maxshift, int flags, int chunk_variants, int* n_chunks_out, int* chunk_variants_out);
/* First variant, size and path counts of a chunk (any output pointer may be null). */
int bx_chromatin_chunk_info(bx_ctx* ctx, int chunk, int64_t* v0, int* m, int* n_delta, int* n_full, int* n_window);
/* Enqueue chunk `chunk`: compute on the library's stream, then copy the three tensors into the caller's HOST buffers
 * (pinned for overlap) in the layout above.  Returns at once; chunks must be submitted in any order but at most two
 * may be in flight: call bx_chromatin_wait(chunk) before submitting chunk + 2.  While chunk k's results travel to
 * the host, chunk k + 1 computes. */
int bx_chromatin_submit(bx_ctx* ctx, int chunk, float* ref_host, float* alt_host, float* diff_host);
int bx_chromatin_wait(bx_ctx* ctx, int chunk);
/* Same computation with the results left in HBM, enqueued on `stream`: row (si, strand t, variant v of the chunk) of
 * each tensor goes to row  si * shift_stride_rows + t * strand_stride_rows + row0 + v  of the caller's device array
 * (rows of 2002 floats; t = 0 with BX_CHROM_AVERAGE) — e.g. the operands of an NCCL gather, or the [S, n, 2002]
 * tensors the decay transform reads. */
int bx_chromatin_submit_device(bx_ctx* ctx, int chunk, float* ref_dev, float* alt_dev, float* diff_dev,
                               int64_t shift_stride_rows, int64_t strand_stride_rows, int64_t row0, void* stream);
/* Row pitch, in floats, of the device arrays bx_chromatin_submit_device writes (default 2002 = dense rows; even,
 * >= 2002; the padding columns are left untouched, except that averaged plans zero whole rows first).  The drop-in's
 * predict path keeps its averaged effects at 2016 so that rows start on 128-byte boundaries, which is what
 * bx_decay_score_variant_tc reads fastest.  Applies to the open plan; the host-buffer path is always dense. */
int bx_chromatin_set_row_pitch(bx_ctx* ctx, int64_t row_pitch);
int bx_chromatin_close(bx_ctx* ctx);
/* Test hook: the device descriptor builders on HOST columns.  kind 0: four full segments per variant (ref fwd, ref
 * rc, alt fwd, alt rc); kind 1: delta mode, out_host[0 .. 2n) the ref segments and out_host[2n .. 4n) the alt patches,
 * *patch_row0 / *patch_len the patch geometry; kind 2: the four windows of each variant at `shift`.  out_host holds
 * 4 n descriptors.  The host restatement in expecto_b200/windows.py is their oracle. */
int bx_variant_descs_host(int kind, int64_t n, const int64_t* gpos, const int32_t* ref_len, const int32_t* alt_len,
                          const int32_t* ref_aoff, const int32_t* alt_aoff, int maxshift, int shift, int* patch_row0,
                          int* patch_len, bx_window_desc* out_host);

/* The same computation as bx_decay_score_variant as ONE tcgen05 GEMM ("[N', 20020] x [20020, 218]"
 * with the features never leaving shared memory; reference predict.py:120-160).  The model weights are converted once
 * to the kernel's operand layout — fp16 (hi, lo) pairs, K axis permuted to the order the feature tiles are generated
 * in — with bx_score_prepare_weights into a caller-owned device buffer of bx_score_weights_bytes(n_models) bytes;
 * bx_decay_score_variant_tc then takes that buffer.  n_shifts must be 9 (--maxshift 800); every product is
 * Ah*Wh + Ah*Wl + Al*Wh with fp32 accumulation, so scores agree with the float64 reference to ~5e-5 of their scale.
 * x_dev is [9, n, x_row_pitch] floats, x_row_pitch >= 2002 (even): rows of exactly 2002 floats work, but they are
 * 8008 bytes apart, so three in four of the kernel's 32-byte reads straddle two DRAM sectors; a pitch of 2016 floats
 * (rows on 128-byte boundaries) is what the drop-in's own producers use: rows on 16-byte boundaries are staged by TMA
 * (1.8 ms per 75,776 variants x 218 models against 3.9 ms for dense rows).  The order of the variants does not change any
 * score, but it changes the time: in a 256-variant tile whose rows all lie upstream (or all downstream) of the TSS at
 * every shift, the operand columns of the other side are zero for every row and are skipped, so callers that group
 * their rows by side (expecto_b200/transform.py: side_order) get 1.6 ms. */
int64_t bx_score_weights_bytes(int n_models);
int bx_score_prepare_weights(const float* w_dev, int n_models, void* wp_dev, void* stream);
int bx_decay_score_variant_tc(const float* x_dev, int64_t x_row_pitch, const double* dist_dev, const int8_t* strand_dev,
                              const int32_t* shifts_dev, const void* wp_dev, const float* offset_dev, int n,
                              int n_shifts, int n_models, float* out_dev, void* stream);

/* ---- FASTA text resident in HBM -> packed genome (the arrays of bx_ctx_set_genome) --------------------------------
 * Replaces the genome side of the reference's input handling: pyfasta.Fasta('./resources/hg19.fa') builds a flat
 * index and serves string slices (chromatin.py:44,205; compute_expecto_features.py:31,108), and encodeSeqs maps each
 * character through `mydict` (chromatin.py:155-160).  Here the whole file is copied to the device once and packed
 * there.  text_dev: the file's bytes, 16-byte aligned, readable up to the next multiple of 16 past n_text.
 *
 * 1. bx_fasta_find_headers: file offsets of every '>' that starts a line -> pos_dev[0..min(count,cap)) (unordered),
 *    *count_dev = how many there are (device int64; the caller reads it back and sorts).
 * 2. per contig body [body_start, body_end) = the bytes after its header line up to the next header (line ends may
 *    sit anywhere: ragged widths, CRLF): bx_fasta_count fills block_offsets_dev[0 .. bx_fasta_blocks()] with the
 *    exclusive prefix of the bases per 4096-byte block and returns the contig length in *n_bases_host (synchronises
 *    the stream).
 * 3. bx_fasta_pack writes the contig at global base index base_offset (a multiple of 32).  The caller lays the
 *    contigs out and PRE-FILLS bases2 with 0 and nmask with ones (the padding between contigs reads as zero columns;
 *    32-base groups shared by two 4096-byte blocks are merged into the planes with atomic OR / AND).  Characters
 *    outside mydict (the reference raises KeyError when a window touches one) are packed as zero columns and their
 *    global indices appended to invalid_dev[0..invalid_cap) (unordered); *invalid_count_dev (device int64, zeroed by
 *    the caller) counts all of them. */
int64_t bx_fasta_blocks(int64_t body_start, int64_t body_end);
int bx_fasta_find_headers(const uint8_t* text_dev, int64_t n_text, int64_t* pos_dev, int64_t cap, int64_t* count_dev,
                          void* stream);
int bx_fasta_count(const uint8_t* text_dev, int64_t body_start, int64_t body_end, int64_t* block_offsets_dev,
                   int64_t* n_bases_host, void* stream);
int bx_fasta_pack(const uint8_t* text_dev, int64_t body_start, int64_t body_end, const int64_t* block_offsets_dev,
                  int64_t base_offset, uint32_t* bases2_dev, uint32_t* nmask_dev, int64_t* invalid_dev,
                  int64_t invalid_cap, int64_t* invalid_count_dev, void* stream);

/* bx_fasta_count for callers that cannot allocate.  This entry point has no counterpart in the reference; it was
 * added (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  Same result as bx_fasta_count in block_offsets_dev[0 .. bx_fasta_blocks()], but it calls no
 * cudaMalloc / cudaFree / allocating routine and does not synchronise: the scratch of its scan lives in a file-scope
 * device array of 256 KB that every GPU holds from the moment it loads the library until exit, and the contig length
 * is left in *n_bases_dev (a DEVICE int64 of the caller's) instead of being read back to the host.
 * Limit: contig bodies of at most 2^20 4096-byte blocks (4 GiB of FASTA text); larger ones are refused with BX_E_ARG
 * — use bx_fasta_count.  Every call on a device shares the one scratch array: calls that overlap on the GPU (two
 * streams, two threads) race; calls on one stream are safe.  The array keeps the last call's contents between
 * calls; nothing in it has to be cleared (the scan initialises what it uses), and the one value that must start at
 * zero, block_offsets_dev[bx_fasta_blocks()], is cleared by the call itself. */
int bx_fasta_count_noalloc(const uint8_t* text_dev, int64_t body_start, int64_t body_end, int64_t* block_offsets_dev,
                           int64_t* n_bases_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif
