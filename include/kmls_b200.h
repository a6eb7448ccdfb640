/* kmls_b200 — C ABI of the B200-native rule-generation hot path.
 *
 * Drop-in boundary for the hot path of diogoneiss/kubernetes-machine-learning-server:
 * machine-learning/main.py:262-313 (calculate_and_save_fp_growth_fast), i.e. the work that
 * the reference delegates to mlxtend 0.23.3's TransactionEncoder and fpgrowth
 * (main.py:267-272) plus its own dict fold (main.py:282-296). The reference has no FFI of
 * its own (it is two Python scripts); the binding a maintainer adds is the ctypes stub in
 * INTEGRATION.md, which is what kubernetes-machine-learning-server_b200/_lib.py implements.
 *
 * Conventions
 *  - plain pointers and sizes only; `d_` pointers are device memory of the current CUDA
 *    device, everything else is host memory; buffers are owned by the caller, inputs are
 *    never written;
 *  - no entry point allocates or frees device memory: scratch is passed in;
 *  - work is enqueued on `stream` (a cudaStream_t passed as void*), nothing synchronises;
 *  - return 0 = ok, > 0 = cudaError_t, < 0 = argument/capacity error (text from
 *    kmls_last_error()); an output that would not fit is an error, never a truncation;
 *  - "row" = a kept track of the vertical bitmap (bit p of row r = playlist p contains
 *    track r); rows are `row_words` 32-bit words long, see kmls_bitmap_row_words().
 */
#ifndef KMLS_B200_H_
#define KMLS_B200_H_

#include <stddef.h>
#include <stdint.h>

#define KMLS_B200_VERSION 100

#ifdef __cplusplus
extern "C" {
#endif

typedef void* kmls_stream_t; /* cudaStream_t */

const char* kmls_last_error(void);
int kmls_version(void);

/* ---- level 1: TransactionEncoder + column sums (main.py:267-269; mlxtend
 *      TransactionEncoder.fit/transform, fpcommon.setup_fptree's np.sum(axis=0)) ---------- */

/* d_occ[t] = occurrences of track id t in d_items (an upper bound of its support; exact
 * when no playlist repeats a track). d_occ is zeroed here. d_items 16-byte aligned. */
int kmls_item_histogram(const int32_t* d_items, int64_t nnz, int32_t n_items, int32_t* d_occ,
                        kmls_stream_t stream);

/* Words per bitmap row for n_playlists columns (padded to 256 playlists = 8 words). */
int64_t kmls_bitmap_row_words(int64_t n_playlists);

/* TransactionEncoder.transform, transposed and bit-packed: playlists in CSR
 * (d_offsets[n_playlists+1], d_items[nnz]); track t goes to row d_row_of_item[t] (or is
 * dropped when that is < 0 or >= n_rows). Writes every word of d_bitmap[n_rows][row_words].
 * A track repeated inside a playlist sets its bit once. */
int kmls_encode_bitmap(const int64_t* d_offsets, const int32_t* d_items, int64_t n_playlists,
                       const int32_t* d_row_of_item, int32_t n_items, int32_t n_rows,
                       uint32_t* d_bitmap, int64_t row_words, kmls_stream_t stream);

/* setup_fptree's item counts: d_counts[r] = popcount of row r. */
int kmls_row_popcount(const uint32_t* d_bitmap, int32_t n_rows, int64_t row_words,
                      int32_t* d_counts, kmls_stream_t stream);

/* d_dst row r = d_src row d_src_row[r] (puts rows into their final order). */
int kmls_gather_rows(const uint32_t* d_src, const int32_t* d_src_row, int32_t n_rows_out,
                     int64_t row_words, uint32_t* d_dst, kmls_stream_t stream);

/* ---- level 2: pair supports (the k = 2 part of fpgrowth.fpg_step /
 *      FPTree.conditional_tree; feeds the fold of main.py:282-296) ------------------------ */

/* d_pair[i * n_rows + j] = |row i AND row j| for all i, j (symmetric, diagonal = support). */
int kmls_pair_counts(const uint32_t* d_bitmap, int32_t n_rows, int64_t row_words,
                     int32_t* d_pair, kmls_stream_t stream);

/* The same counts from the playlists themselves (sparse form): every playlist adds 1 to the
 * counter of each pair of distinct kept tracks it holds (a repeated track counts once) and to
 * the diagonal. Cost = sum over playlists of m(m+1)/2 updates, m = kept tracks of the playlist,
 * instead of n_rows^2 * n_playlists / 2 multiply-adds: the cheaper route when the one-hot
 * matrix is sparse. d_row_of_item[track id] = row or -1 (track not kept). Writes the upper
 * triangle and the diagonal of d_pair (zeroed here). *d_overflow (zeroed here) counts the
 * playlists with more than 1024 distinct kept tracks, which are SKIPPED: if it is not 0 the
 * result is incomplete and the caller must use one of the dense kernels. n_rows <= 16384. */
int kmls_pair_counts_csr(const int64_t* d_offsets, const int32_t* d_items, int64_t n_playlists,
                         const int32_t* d_row_of_item, int32_t n_rows, int32_t* d_pair,
                         int32_t* d_overflow, kmls_stream_t stream);

/* The same contraction on the tensor cores (tcgen05.mma kind::i8, int32 accumulators in
 * TMEM, bit slices fetched by TMA and expanded to int8 inside the SM). Only entries with
 * j >= i (upper triangle, diagonal included) are guaranteed to be written. d_scratch:
 * kmls_pair_gemm_scratch_bytes() bytes, 16-byte aligned (holds a K-block-major copy of the
 * bitmap). */
size_t kmls_pair_gemm_scratch_bytes(int32_t n_rows, int64_t row_words);
int kmls_pair_counts_i8(const uint32_t* d_bitmap, int32_t n_rows, int64_t row_words,
                        int32_t* d_pair, void* d_scratch, size_t scratch_bytes,
                        kmls_stream_t stream);

/* Hybrid of the two for a Zipf-like catalogue (what kmls_mine runs by default on a sparse
 * input with enough rows): rows are in ascending-support order, so the LAST n_hot rows are
 * the most frequent tracks and their n_hot x n_hot block takes a large share of all counter
 * updates. kmls_pair_counts_csr_cold counts everything EXCEPT the pairs and diagonal entries
 * inside that block (which it leaves zero); kmls_pair_counts_i8_block then fills the block on
 * the tensor cores: d_bitmap = first hot row, row_words = words of the playlist range counted
 * (a multiple of 4), row_stride = words between rows, d_pair = the block's top-left entry,
 * ld = the whole matrix's row length. The block must be zero on entry (counts are stored, or
 * added with atomics when the playlist dimension is split). Same kernels as
 * k_pair_counts_csr / k_pair_gemm_i8_ts (reference: k = 2 of fpgrowth.fpg_step). */
int kmls_pair_counts_csr_cold(const int64_t* d_offsets, const int32_t* d_items, int64_t n_playlists,
                              const int32_t* d_row_of_item, int32_t n_rows, int32_t n_hot,
                              int32_t* d_pair, int32_t* d_overflow, kmls_stream_t stream);
int kmls_pair_counts_i8_block(const uint32_t* d_bitmap, int32_t n_rows, int64_t row_words,
                              int64_t row_stride, int32_t* d_pair, int64_t ld, void* d_scratch,
                              size_t scratch_bytes, kmls_stream_t stream);

/* Frequent-pair table: all (a, b), a < b, with d_pair[a][b] >= thr, sorted by (a, b).
 * d_row_off[n_rows+1] receives the exclusive scan of per-row counts (d_row_off[n_rows] =
 * number of pairs). With capacity == 0 only d_row_off is produced (sizing pass); otherwise
 * d_items2[2*i], d_items2[2*i+1], d_counts[i] are written for rows that fit in capacity. */
int kmls_pair_select(const int32_t* d_pair, int32_t n_rows, int32_t thr, int64_t capacity,
                     int64_t* d_row_off, int32_t* d_items2, int32_t* d_counts,
                     kmls_stream_t stream);

/* The fold of main.py:282-296 as a CSR, from the frequent-pair table (d_items2[n_pairs][2]
 * bitmap rows, d_counts2[n_pairs]): the entries of row r (track d_row_items[r]) are
 * d_indptr[r] .. d_indptr[r+1]-1; entry e says "d_other[e] (a TRACK ID) occurs with this track
 * in a frequent pair of support d_support[e] = count / N". Both directions of every pair are
 * present (2 * n_pairs entries); inside a row entries are ordered by support descending, ties
 * by track id ascending, so that the dict the host builds from it (and rest_api's stable
 * top-K over it, rest_api/app/main.py:250) is deterministic. d_indptr[n_rows + 1],
 * d_other / d_support[2 * n_pairs]; d_scratch: kmls_pair_map_scratch_bytes() bytes, 256-aligned. */
size_t kmls_pair_map_scratch_bytes(int64_t n_pairs, int32_t n_rows);
int kmls_pair_map(const int32_t* d_items2, const int32_t* d_counts2, int64_t n_pairs,
                  const int32_t* d_row_items, int32_t n_rows, int64_t n_playlists_total,
                  int64_t* d_indptr, int32_t* d_other, double* d_support, void* d_scratch,
                  size_t scratch_bytes, kmls_stream_t stream);

/* fpcommon.generate_itemsets: support = count / N (IEEE double division). */
int kmls_supports_f64(const int32_t* d_counts, int64_t n, int64_t n_playlists_total,
                      double* d_out, kmls_stream_t stream);

/* No counterpart in the reference (plumbing). Copies n_words (<= 4096) 32-bit words from
 * device memory to pinned host memory that the device can address (cudaHostAlloc /
 * cudaHostRegister memory under unified addressing), with a kernel rather than the copy
 * engine: level sizes read back by the host between levels then never wait behind a bulk
 * device-to-host transfer. The host reads h_mapped_dst after synchronising the stream. */
int kmls_words_to_host(const void* d_src, int32_t n_words, void* h_mapped_dst, kmls_stream_t stream);

/* ---- levels k >= 3 (fpgrowth.fpg_step recursion, restated level-wise) ------------------- */

size_t kmls_scan_scratch_bytes(int64_t n);
/* d_out[n+1] = exclusive prefix sums of d_in[n] in 64 bits; d_out[n] = total. */
int kmls_exclusive_scan_i32(const int32_t* d_in, int64_t n, int64_t* d_out, void* d_scratch,
                            size_t scratch_bytes, kmls_stream_t stream);

/* No counterpart in the reference. Same result as kmls_exclusive_scan_i32 for callers that
 * cannot supply scratch because they cannot allocate device memory (they only hold buffers
 * handed to them: a plugin inside a pre-budgeted CUDA context, a loop that must avoid the
 * device-wide stall of cudaMalloc/cudaFree, an FFI caller without a device allocator). The
 * scratch is a 1 MiB array inside the library image (held on every GPU that loads the
 * library until exit); n <= KMLS_NOALLOC_SCAN_MAX_N or the call fails with an error. The
 * first call on a device resolves the array's address: issue it before any stream capture.
 * Calls share the array, so calls that overlap on the GPU race - keep them on one stream. */
#define KMLS_NOALLOC_SCAN_MAX_N 4194304
int kmls_exclusive_scan_i32_noalloc(const int32_t* d_in, int64_t n, int64_t* d_out,
                                    kmls_stream_t stream);

/* Fingerprint set of a level table d_items[n][k] (used for Apriori pruning). */
int64_t kmls_set_capacity(int64_t n);
int kmls_itemset_set_build(const int32_t* d_items, int64_t n, int32_t k, uint64_t* d_slots,
                           int64_t capacity, kmls_stream_t stream);

/* Candidate generation for level k+1 from level table d_items[n][k] (sorted): join of
 * itemsets sharing their first k-1 members, Apriori-pruned (k == 2: against d_pair with
 * threshold thr; k >= 3: against the fingerprint set of this level).
 * kmls_join_count: d_kept_cnt[n], d_kept_mask[n] and the scan d_kept_off[n+1]
 *                  (d_kept_off[n] = number of candidates).
 * kmls_join_write: candidates (d_cand_p[c], d_cand_q[c]) = indices of the two parents,
 *                  grouped by p in ascending order; d_grp_p / *d_n_groups = the parents that
 *                  have at least one candidate (skipped when d_grp_p is NULL).
 * A NULL d_pair (k == 2) or NULL d_slots (k >= 3) switches the pruning off. d_par (may be
 * NULL) = each itemset's parent index in level k-1: with it, "same class" is one integer
 * compare instead of k-1. */
int kmls_join_count(const int32_t* d_items, int64_t n, int32_t k, const uint64_t* d_slots,
                    int64_t capacity, const int32_t* d_pair, int32_t n_rows, int32_t thr,
                    const int32_t* d_par, int32_t* d_kept_cnt, uint32_t* d_kept_mask, int64_t* d_kept_off,
                    void* d_scratch, size_t scratch_bytes, kmls_stream_t stream);
int kmls_join_write(const int32_t* d_items, int64_t n, int32_t k, const uint64_t* d_slots,
                    int64_t capacity, const int32_t* d_pair, int32_t n_rows, int32_t thr,
                    const int32_t* d_par, const int32_t* d_kept_cnt, const uint32_t* d_kept_mask,
                    const int64_t* d_kept_off, int32_t* d_cand_p, int32_t* d_cand_q,
                    int32_t* d_grp_p, int64_t* d_n_groups, void* d_scratch, size_t scratch_bytes,
                    kmls_stream_t stream);

/* Support counts of the candidates: d_cand_cnt[c] = |AND of the k+1 member rows|. */
int kmls_count_candidates(const uint32_t* d_bitmap, int64_t row_words, const int32_t* d_items,
                          int32_t k, const int32_t* d_grp_p, int64_t n_groups,
                          const int64_t* d_kept_off, const int32_t* d_cand_q, int64_t n_cand,
                          int32_t* d_cand_cnt, kmls_stream_t stream);

/* Filter by the integer threshold Tk = ceil(min_support * N) (fpgrowth.py) and emit the
 * level k+1 table in order. kmls_emit_count: d_out_off[n_cand+1] (last = new level size). */
int kmls_emit_count(const int32_t* d_cand_cnt, int64_t n_cand, int32_t thr, int64_t* d_out_off,
                    void* d_scratch, size_t scratch_bytes, kmls_stream_t stream);
int kmls_emit_level(const int32_t* d_items, int32_t k, const int32_t* d_cand_p,
                    const int32_t* d_cand_q, const int32_t* d_cand_cnt, int64_t n_cand,
                    int32_t thr, const int64_t* d_out_off, int32_t* d_out_items,
                    int32_t* d_out_counts, kmls_stream_t stream);

/* ---- conditional (projected) bitmaps: the bitmap analogue of FPTree.conditional_tree ---
 * Every frequent itemset X = (P, x) of a level may own a bit row over the playlists that
 * contain its class prefix P (columns compacted): row i starts at d_arena + d_rowptr[i] and
 * is d_roww[i] words long (0 = no row; multiples of 4). Level 1 = the full bitmap
 * (d_rowptr[r] = r * row_words, d_roww[r] = row_words). */

/* d_cand_cnt[c] = |row(d_cand_p[c]) AND row(d_cand_q[c])| (this rank's playlists). */
int kmls_count_pairs(const uint32_t* d_arena, const int64_t* d_rowptr, const int32_t* d_roww,
                     const int32_t* d_cand_p, const int32_t* d_cand_q, int64_t n_cand,
                     int32_t* d_cand_cnt, kmls_stream_t stream);
/* kmls_emit_level that also records each new itemset's parents and its local count
 * (d_cand_cnt = global counts used for the threshold, d_cand_cnt_local = this rank's). */
int kmls_emit_level_px(const int32_t* d_items, int32_t k, const int32_t* d_cand_p,
                       const int32_t* d_cand_q, const int32_t* d_cand_cnt,
                       const int32_t* d_cand_cnt_local, int64_t n_cand, int32_t thr,
                       const int64_t* d_out_off, int32_t* d_out_items, int32_t* d_out_counts,
                       int32_t* d_out_counts_local, int32_t* d_out_par_p, int32_t* d_out_par_q,
                       kmls_stream_t stream);
/* Row widths of a new level: 0 if no other itemset shares o's first parent p; else the
 * parent's width d_parent_roww[p] if that is <= keep_words words (the row then stays in the parent's
 * column space); else ceil(d_parent_cnt_local[p] / 32) padded to 4 words. d_parent_roww may be
 * NULL (always compact). keep_words: the limit above (-1 = default, 256; pass the same value to
 * kmls_project_rows). */
int kmls_row_alloc(const int32_t* d_par_p, int64_t n, const int32_t* d_parent_cnt_local,
                   const int32_t* d_parent_roww, int32_t keep_words, int32_t* d_roww,
                   kmls_stream_t stream);
/* New rows: row(o) = row_old(d_par_q[o]) restricted to the columns set in
 * row_old(d_par_p[o]), compacted (parallel bit extract) - or, when the parents' rows are
 * <= 128 words, row_old(p) AND row_old(q) uncompacted (widths as set by kmls_row_alloc). */
int kmls_project_rows(const uint32_t* d_arena_old, const int64_t* d_rowptr_old,
                      const int32_t* d_roww_old, const int32_t* d_par_p, const int32_t* d_par_q,
                      int64_t n, uint32_t* d_arena_new, const int64_t* d_rowptr_new,
                      const int32_t* d_roww_new, int32_t keep_words, kmls_stream_t stream);

/* First projection (level 1 -> 2), where the masks are sparse single-track rows:
 * kmls_row_tidlists turns every bitmap row r into the ascending list of its set columns at
 * d_tids[d_tid_off[r] .. d_tid_off[r+1]) (d_tid_off = exclusive scan of the local row
 * counts); kmls_project_rows_gather then builds row(a, b) by sampling bitmap row b at the
 * columns of a (total_words = d_rowptr_new[n], the size of the new arena). Same output as
 * kmls_project_rows on the full bitmap. Rows 16-byte aligned (row_words a multiple of 4). */
int kmls_row_tidlists(const uint32_t* d_bitmap, int32_t n_rows, int64_t row_words,
                      const int64_t* d_tid_off, int32_t* d_tids, kmls_stream_t stream);
int kmls_project_rows_gather(const uint32_t* d_bitmap, int64_t row_words, const int32_t* d_par_p,
                             const int32_t* d_par_q, int64_t n, const int32_t* d_tids,
                             const int64_t* d_tid_off, uint32_t* d_arena_new,
                             const int64_t* d_rowptr_new, int64_t total_words,
                             kmls_stream_t stream);

/* ---- rules (mlxtend association_rules semantics; the dormant path main.py:224-260 is the
 *      only reference site of a min_confidence knob) -------------------------------------- */

/* All level tables concatenated ("arena"): level k's members start at
 * sum_{j<k} level_sizes[j]*j in d_items and its counts at sum_{j<k} level_sizes[j] in
 * d_counts (= its first global id); level_sizes is a HOST array indexed 1..max_k (index 0
 * unused). d_par[g - level_sizes[1]] = index, inside level k-1, of the parent (itemset
 * without its last member) of the itemset with global id g (sizes >= 2 only).
 *
 * Exact subset index: one 16-byte node per itemset (all levels, indexed by global id):
 * {global id of its first child, number of children, last member, count}. The children of an
 * itemset (it plus one later member) are contiguous in the next level because the tables are
 * sorted. d_nodes: (total number of itemsets) * 16 bytes, 16-byte aligned. */
int kmls_trie_index_build(const int32_t* d_items, const int32_t* d_counts, const int32_t* d_par,
                          const int64_t* level_sizes, int32_t max_k, void* d_nodes,
                          kmls_stream_t stream);
/* Slots per k-itemset in pass 1's output = 2^(k-1): slot 0 plus one per (antecedent,
 * consequent) split (each split yields up to two rules); -1 if k is not supported (k > 15).
 * Pass 1's output for n itemsets of level k, S = that number: two int32 planes, d_part[0 ..
 * n*S) = count(A), d_part[n*S .. 2*n*S) = count(C); slot s >= 1 of an itemset is the split whose
 * antecedent mask over the itemset's first k-1 members is the bit reversal (over k-1 bits) of
 * s, C = the rest of the itemset (it always holds the last member); slot 0 = the itemset's own
 * count in both planes. So d_part needs n * S * 8 bytes. */
int64_t kmls_rules_parts_per_itemset(int32_t k);
/* Pass 1 for the itemsets lvl_begin .. lvl_begin+lvl_count-1 of level k (ranks may split a
 * level): d_part (layout above, n = lvl_count) receives the subset counts of every split,
 * d_rule_cnt[lvl_count] the number of kept rules per itemset; *d_missing counts failed subset
 * lookups (0 for a downward-closed table; caller zeroes it). */
int kmls_rules_count(const int32_t* d_items, const int32_t* d_counts, const int64_t* level_sizes,
                     int32_t max_k, int32_t k, int64_t lvl_begin, int64_t lvl_count,
                     const void* d_nodes, int64_t n_playlists_total, double min_confidence,
                     void* d_part, int32_t* d_rule_cnt, int32_t* d_missing, kmls_stream_t stream);
/* Pass 1 without a search per split, for a caller that runs pass 1 over WHOLE levels in
 * ascending order on one stream (one GPU holding the whole table). Two calls per level k:
 *  kmls_rules_subsets: d_sub[level_sizes[k] * (k-1)]: entry (x, i) = index inside level k-1 of
 *    itemset x without its i-th member (i < k-1, i.e. the subsets that keep the last member);
 *    d_sub_prev = the same array of level k-1 (unused for k = 2). One search per entry.
 *  kmls_rules_count_gather: fills d_part / d_rule_cnt of the whole level k like
 *    kmls_rules_count, reading every subset count from d_part_prev = the d_part of the whole
 *    level k-1 (unused for k = 2): count(A) from the parent's row, count(C) from the row of
 *    the subset that drops A's first member.
 * Same results as kmls_rules_count; *d_missing as there. */
int kmls_rules_subsets(const int32_t* d_items, const int32_t* d_par, const int64_t* level_sizes,
                       int32_t max_k, int32_t k, const void* d_nodes, const int32_t* d_sub_prev,
                       int32_t* d_sub, int32_t* d_missing, kmls_stream_t stream);
int kmls_rules_count_gather(const int32_t* d_items, const int32_t* d_counts, const int32_t* d_par,
                            const int64_t* level_sizes, int32_t max_k, int32_t k,
                            const int32_t* d_sub, const void* d_part_prev,
                            int64_t n_playlists_total, double min_confidence, void* d_part,
                            int32_t* d_rule_cnt, int32_t* d_missing, kmls_stream_t stream);
/* Pass 2: d_rule_off[lvl_count + 1] = first output slot of each itemset (exclusive scan of the
 * kept counts over everything this rank emits; the entry after the last itemset is read too).
 * Rule r: d_out_rule[2r] = itemset (global id), d_out_rule[2r+1] = antecedent mask (bit t =
 * t-th member; the consequent is the rest), d_out_confidence[r], d_out_lift[r].
 * *d_mismatch (NOT zeroed here) is incremented for every itemset for which this pass keeps a
 * different number of rules than pass 1 sized: must stay 0, anything else is an error. */
int kmls_rules_write(const int32_t* d_counts, const int64_t* level_sizes, int32_t max_k,
                     int32_t k, int64_t lvl_begin, int64_t lvl_count, int64_t n_playlists_total,
                     double min_confidence, const void* d_part, const int64_t* d_rule_off,
                     int32_t* d_out_rule, double* d_out_confidence, double* d_out_lift,
                     int32_t* d_mismatch, kmls_stream_t stream);
/* kmls_rules_write with the two planes of pass 1 passed separately: the rows lvl_begin ..
 * lvl_begin + lvl_count - 1 of planes that pass 1 wrote for a LARGER range (d_plane_a /
 * d_plane_c point at the first slot of row lvl_begin in each plane). This is how a rank of a
 * small multi-GPU job writes its slice of the rules after a gather pass 1 over whole levels. */
int kmls_rules_write_planes(const int32_t* d_counts, const int64_t* level_sizes, int32_t max_k,
                            int32_t k, int64_t lvl_begin, int64_t lvl_count,
                            int64_t n_playlists_total, double min_confidence,
                            const int32_t* d_plane_a, const int32_t* d_plane_c,
                            const int64_t* d_rule_off, int32_t* d_out_rule, double* d_out_confidence,
                            double* d_out_lift, int32_t* d_mismatch, kmls_stream_t stream);

/* No counterpart in the reference (parity plumbing). Adds to *d_out (NOT zeroed here) the
 * order-independent checksum of one level table d_items[n][k] (bitmap rows) / d_counts[n]:
 * sum over itemsets of mix(sum_i mix(track id_i + 1) ^ mix(count * 0x100000001B3 + k)) mod
 * 2^64, track ids through d_row_items. Same arithmetic as the CPU oracle's, so a full-size
 * result can be compared with a committed golden value without leaving the device. */
int kmls_itemset_checksum(const int32_t* d_items, const int32_t* d_counts, int64_t n, int32_t k,
                          const int32_t* d_row_items, uint64_t* d_out, kmls_stream_t stream);

/* The support column of the rule table, on demand: d_out[r] = count of rule r's itemset / N
 * (d_counts = the arena's counts). Not written by pass 2: it would be a quarter of its bytes
 * and is one gather away. */
int kmls_rule_supports(const int32_t* d_rule, int64_t n_rules, const int32_t* d_counts,
                       int64_t n_playlists_total, double* d_out, kmls_stream_t stream);

/* No counterpart in the reference (verification aid). Order-independent 64-bit checksum of a
 * rule table over (antecedent track ids, consequent track ids, bits of confidence, lift and
 * support) - the arithmetic is documented at k_rule_checksum in rules.cu and restated by the
 * CPU oracle, so two implementations' full tables can be compared without moving them.
 * d_row_items maps bitmap rows to track ids; *d_out is zeroed here. */
int kmls_rule_checksum(const int32_t* d_items, const int32_t* d_counts, const int64_t* level_sizes,
                       int32_t max_k, const int32_t* d_rule, const double* d_confidence,
                       const double* d_lift, int64_t n_rules, const int32_t* d_row_items,
                       int64_t n_playlists_total, uint64_t* d_out, kmls_stream_t stream);

/* No counterpart in the reference (verification aid). Pass 2 computes support, confidence and
 * lift without a division instruction sequence per quotient (reciprocal + FMA corrections,
 * rules.cu div_by_rcp). This draws n_samples pseudo-random count triples cX <= cA, cC <= N
 * from `seed`, evaluates the quotients of a split both ways (that path and __ddiv_rn) and
 * adds the number of quotients whose bits differ to *d_mismatches (zeroed here); it must stay
 * 0. */
int kmls_rule_division_selftest(int64_t n_playlists_total, int64_t n_samples, uint64_t seed,
                                uint64_t* d_mismatches, kmls_stream_t stream);

/* ---- the whole path in one call (what stands behind
 *      TransactionEncoder().fit().transform() + fpgrowth(df, min_support, use_colnames=True),
 *      machine-learning/main.py:267-272, and behind association_rules for the rule table) ---
 *
 * kmls_mine runs every level on the device and returns when the tables are complete; the
 * level loop, its sizing decisions and its buffers live inside the library. Device memory
 * comes from ONE caller-supplied workspace (no cudaMalloc here): tables that make up the
 * result are placed at its low end and stay valid until the caller reuses the workspace,
 * per-level buffers come and go above them. When the workspace is too small the call fails
 * with a capacity error whose text says how much was missing; nothing is ever truncated -
 * call again with a larger workspace. The host reads a few sizes per level from a pinned
 * mailbox (a small cudaHostAlloc per host thread, made on the thread's first call; host
 * memory, but make that first call outside stream capture). */
#define KMLS_MAX_LEVELS 32

typedef struct kmls_mine_options {
  int32_t max_len;        /* longest itemset to mine; 0 = no limit (fpgrowth's max_len=None) */
  int32_t pair_impl;      /* level 2: 0 = choose by cost, 1 = per playlist (sparse), 2 = CUDA-core
                             AND+popc, 3 = tcgen05 int8 GEMM */
  int32_t keep_words;     /* conditional rows at most this wide keep their column space (AND
                             instead of bit extract); 0 = default (256), at most 1024 */
  int32_t n_parts;        /* class partition for multi-GPU jobs: this call mines only the      */
  int32_t part;           /* subtrees of the frequent pairs in blocks part, part + n_parts, ... */
  int32_t part_block;     /* of part_block consecutive pairs; n_parts <= 1 = everything        */
  int32_t squeeze_pct;    /* a row wider than keep_words is squeezed (bit extract) only if that
                             leaves at most this percentage of its width; 0 = default (50)     */
  int32_t hot_rows;       /* level 2, per-playlist route: the block of the hot_rows most frequent
                             rows is counted on the tensor cores instead (kmls_pair_counts_csr_cold
                             + kmls_pair_counts_i8_block). 0 = choose by cost (1024 rows or none),
                             > 0 = this many (rounded up to 128), < 0 = never                   */
  int64_t max_itemsets;   /* guard: fail (never truncate) beyond this many itemsets; 0 = 2e8   */
  void* h_tables;         /* optional PINNED host buffer: every level table is copied there on */
  size_t h_tables_bytes;  /* a side stream as soon as it is final, while mining continues      */
  /* Optional, for the ranks of a class-partitioned job (n_parts > 1, every rank holding all
   * playlists): sum-allreduce d_buf[n] (int32, device memory inside the workspace) over the
   * ranks, ordered on `stream`; return 0. With it the level-2 pair counts are taken from
   * playlists part*N/n_parts .. (part+1)*N/n_parts only and summed over the ranks, instead of
   * every rank counting all playlists. The host side of the job supplies it (NCCL through
   * torch.distributed in this repository); the library itself links no communication library. */
  int (*allreduce_i32)(void* ctx, int32_t* d_buf, int64_t n, kmls_stream_t stream);
  void* allreduce_ctx;
} kmls_mine_options;

typedef struct kmls_mine_result {
  int64_t n_playlists_total;
  int32_t t1, tk, thr_k;  /* thresholds used: level 1, levels >= 2, and the one applied */
  int32_t n_rows;         /* frequent tracks = rows of the bitmap, in ascending-support order */
  int32_t max_k;          /* longest frequent itemset */
  int32_t pair_impl_used; /* 1, 2, 3 as in the options; 4 = 1 with the hot block on the tensor cores */
  int64_t n_itemsets;
  int64_t kernel_launches;
  size_t workspace_peak_bytes;
  /* index k = itemset size (0 unused). Tables of level k, as BYTE OFFSETS into the workspace
   * (-1 = none): items int32[n][k] (bitmap rows, ascending inside an itemset, table sorted),
   * counts int32[n], parent int32[n] (index in level k-1 of the itemset minus its last member,
   * k >= 2; for k == 2 the first member's row). */
  int64_t level_size[KMLS_MAX_LEVELS + 2];
  int64_t level_candidates[KMLS_MAX_LEVELS + 2];
  int64_t level_row_words[KMLS_MAX_LEVELS + 2];  /* words of conditional rows built for the level */
  int64_t items_off[KMLS_MAX_LEVELS + 2];
  int64_t counts_off[KMLS_MAX_LEVELS + 2];
  int64_t parent_off[KMLS_MAX_LEVELS + 2];
  int64_t row_items_off;    /* int32[n_rows]: row -> track id */
  int64_t occurrences_off;  /* int32[n_items]: rows of the input per track (with repeats) */
  int64_t pair_off;         /* int32[n_rows][n_rows] pair counts (upper triangle valid), or -1 */
  /* with options.h_tables: byte offsets into that host buffer, same indexing */
  int64_t h_items_off[KMLS_MAX_LEVELS + 2];
  int64_t h_counts_off[KMLS_MAX_LEVELS + 2];
  int64_t h_row_items_off;
  int64_t h_bytes;
} kmls_mine_result;

/* The two integer thresholds mlxtend's float tests amount to, computed with
 * the same IEEE double operations as Python: *t1 = min{c : c / (double)n >= min_support}
 * (fpcommon.setup_fptree), *tk = ceil(min_support * n) (fpgrowth.fpgrowth). Returns an
 * argument error for min_support <= 0 (mlxtend raises ValueError there). */
int kmls_thresholds(double min_support, int64_t n_playlists_total, int64_t* t1, int64_t* tk);

/* Playlists in CSR (this process's: d_offsets[n_playlists + 1], d_items[nnz], track ids
 * 0 .. n_items-1) -> all frequent itemsets. n_playlists_total is N of the whole job (it may
 * exceed n_playlists only when the caller merges counts itself; single process: equal).
 * d_workspace: 256-byte aligned device memory. All work is enqueued on `stream`; the call
 * synchronises that stream several times per level and returns with the tables complete. */
int kmls_mine(const int64_t* d_offsets, const int32_t* d_items, int64_t n_playlists, int32_t n_items,
              int64_t n_playlists_total, int32_t t1, int32_t tk, const kmls_mine_options* options,
              void* d_workspace, size_t workspace_bytes, kmls_mine_result* result,
              kmls_stream_t stream);

typedef struct kmls_rules_result {
  int64_t n_rules;          /* rules this call wrote (its part of the table) */
  int64_t n_itemsets;       /* itemsets of all levels (= length of the arena's counts) */
  int64_t n_splits;         /* pass-1 slots of this part: sum of 2^(k-1) over its itemsets */
  /* byte offsets into the workspace (-1 = none): rule int32[n_rules][2] = (itemset global id =
   * position in the level tables concatenated in level order, antecedent bit mask over its
   * members), confidence double[n_rules], lift double[n_rules]; and the concatenated level
   * tables the ids refer to: arena_items (members, level after level), arena_counts int32 */
  int64_t rule_off, confidence_off, lift_off, arena_items_off, arena_counts_off;
  size_t workspace_needed_bytes;  /* set as soon as it is known, also when the call fails for lack of room */
  size_t workspace_peak_bytes;
  int64_t kernel_launches;
  int32_t pass1_used;       /* 1 = gather route, 2 = search route */
  int32_t reserved;
} kmls_rules_result;

/* association_rules(metric="confidence", min_threshold=min_confidence) over the tables of a
 * kmls_mine result (mined->*_off are byte offsets from d_tables_base, normally the mining
 * workspace; a caller that assembled the tables itself may pass absolute addresses and a NULL
 * base): confidence = (cX/N)/(cA/N), lift = confidence/(cC/N) in mlxtend's operation order,
 * kept iff confidence >= min_confidence; support is the itemset's own count / N
 * (kmls_rule_supports). part / n_parts: write only the part-th of n_parts contiguous slices of
 * the table, cut by work (how the ranks of a multi-GPU job divide it; every rank needs the
 * whole itemset table). pass1: 0 = choose, 1 = gather (whole table only), 2 = search.
 * d_workspace must not overlap the tables. Capacity errors name the bytes needed
 * (result->workspace_needed_bytes); max_rules <= 0 = 4e9. Itemsets of more than 15 members
 * are refused. */
int kmls_rules(const kmls_mine_result* mined, const void* d_tables_base, double min_confidence,
               int32_t part, int32_t n_parts, int32_t pass1, int64_t max_rules, void* d_workspace,
               size_t workspace_bytes, kmls_rules_result* result, kmls_stream_t stream);

/* ---- multi-GPU: merging class-partitioned results (no counterpart in the reference, which
 *      has no parallelism). Each rank mined the subtrees of its own frequent pairs (kmls_mine
 *      with n_parts > 1); the job's table of level k >= 3 is the ranks' itemsets merged in the
 *      order of their root pair, which is the sorted order a single GPU produces. ----------- */

/* d_root[o] = the frequent pair (index into level 2) under which local itemset o of level k
 * lives: its parent for k == 3 (d_root_prev NULL), else d_root_prev[parent] (roots of the local
 * level k-1). d_per_root[r] (zeroed by the caller, one int per frequent pair) += itemsets. */
int kmls_subtree_roots(const int32_t* d_par, int64_t n, const int32_t* d_root_prev, int32_t* d_root,
                       int32_t* d_per_root, kmls_stream_t stream);
/* Writes the n local itemsets of level k at their job-wide positions g = d_global_start[root] +
 * (local index - d_local_start[root]) (exclusive scans over the pairs of the job-wide and of
 * this rank's per-root counts) into job-wide tables that the caller zero-initialised and
 * afterwards completes with a sum-allreduce; the parent column is translated to job-wide
 * positions through d_gprev (= d_g_out of level k-1; NULL for k == 3, whose parents are pair
 * indices already). d_g_out[o] = g. */
int kmls_place_level(const int32_t* d_items, const int32_t* d_counts, const int32_t* d_par, int64_t n,
                     int32_t k, const int32_t* d_root, const int64_t* d_global_start,
                     const int64_t* d_local_start, const int32_t* d_gprev, int32_t* d_items_out,
                     int32_t* d_counts_out, int32_t* d_par_out, int32_t* d_g_out, kmls_stream_t stream);

/* The same merge with 12 bytes per itemset on the wire instead of 4k + 8: only (last member,
 * count, parent) are placed and summed over the ranks; kmls_rebuild_items then recreates the
 * member lists of level k on every rank from the merged parent links, level after level:
 * d_items[g] = d_items_prev[d_par[g]] (k-1 members) followed by d_last[g]. */
int kmls_place_level_compact(const int32_t* d_items, const int32_t* d_counts, const int32_t* d_par,
                             int64_t n, int32_t k, const int32_t* d_root, const int64_t* d_global_start,
                             const int64_t* d_local_start, const int32_t* d_gprev, int32_t* d_last_out,
                             int32_t* d_counts_out, int32_t* d_par_out, int32_t* d_g_out,
                             kmls_stream_t stream);
int kmls_rebuild_items(const int32_t* d_items_prev, int32_t k, const int32_t* d_par, const int32_t* d_last,
                       int64_t n, int32_t* d_items, kmls_stream_t stream);

/* ---- measurement plumbing (no counterpart in the reference) --------------------------------
 * kmls_profile_begin switches per-kernel timing on for the calling host thread: from then on
 * kmls_mine and kmls_rules bracket every kernel they launch with two CUDA events on the
 * launching stream. kmls_profile_report switches it off again, waits for the events and writes
 * one text line per launch, "kernel,level,milliseconds,algorithmic_work", into buf (at most
 * capacity bytes, NUL-terminated); returns the bytes the full report needs. The work figure is
 * the algorithmic bytes of the launch as DESIGN.md section 3 defines them (operations for the
 * tensor-core GEMM). Where that figure depends on the data (rows actually read and written by
 * the counting kernels, counter updates, rules kept per level) finding it costs a few small
 * kernels and a host synchronisation per level: kmls_profile_begin_events records the events
 * only and leaves those figures at their static part, for the timed steps of a benchmark
 * (bench.py takes the work figures from one extra, untimed step with kmls_profile_begin). */
int kmls_profile_begin(void);
int kmls_profile_begin_events(void);
int64_t kmls_profile_report(char* buf, size_t capacity);

/* Measurement plumbing: n_updates 32-bit RED.ADD updates scattered uniformly over
 * d_buf[n_words] - the micro-benchmark that gives k_pair_counts_csr's roofline its peak
 * (counter updates per second; tools/redg_peak.py times it). */
int kmls_redg_microbench(int32_t* d_buf, int64_t n_words, int64_t n_updates, uint64_t seed,
                         kmls_stream_t stream);
/* The same with the addresses of one warp instruction confined to one stretch of `span`
 * consecutive counters at a random multiple of span (the pattern of k_pair_counts_csr, whose
 * warp keeps the matrix row while the column varies; span = 32: one 128-byte line). */
int kmls_redg_microbench_span(int32_t* d_buf, int64_t n_words, int64_t n_updates, int64_t span,
                              uint64_t seed, kmls_stream_t stream);
/* Measurement plumbing: a streaming kernel with `reads` (0..3) 8-byte loads and `writes`
 * (0..3) 8-byte stores per element over separate arrays of n elements - the achievable HBM
 * rate for the read:write mix of k_rules_write (1 : 3), which a copy (1 : 1) does not tell
 * (tools/stream_mix_peak.py times it). */
int kmls_stream_mix_microbench(const void* d_a, const void* d_b, const void* d_c, int64_t n,
                               int32_t reads, int32_t writes, void* d_x, void* d_y, void* d_z,
                               kmls_stream_t stream);


#ifdef __cplusplus
}
#endif
#endif /* KMLS_B200_H_ */
