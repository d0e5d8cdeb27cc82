/*
 * salsa_b200.h -- C ABI of libsalsa_b200.so: SALSA's Hi-C link-scoring hot
 * path as hand-written CUDA for B200 (sm_100a).
 *
 * The reference (marbl/SALSA) has NO in-process plugin / FFI interface: its
 * hot path sits behind process + file boundaries driven by run_pipeline.py
 * (SURVEY.md section 8(b)).  Each entry point below therefore cites the
 * reference PROGRAM (file:line of the loop) whose work it replaces; the
 * drop-in step programs in salsa_b200/bin/ (make_links.py, RE_sites.py,
 * fast_scaled_scores.py, layout_unitigs.py, refactor_breaks.py through
 * ctypes; break_contigs, break_contigs_start as C++ mains) keep the
 * reference's argv and file formats and call these functions.
 * INTEGRATION.md shows the binding a maintainer of the reference would add.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / C++ types.
 *   - "d_" arguments are DEVICE pointers owned by the caller (e.g.
 *     torch.Tensor.data_ptr()); "h_" arguments are HOST pointers.  The
 *     library never frees caller memory.
 *   - every function returns 0 on success, < 0 on error; the message is in
 *     salsa_last_error() (thread local).
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream).
 *     Work is enqueued asynchronously unless the function is documented to
 *     synchronise.
 *   - there is NO CPU fallback: without a CUDA device every compute entry
 *     point fails with SALSA_ERR_CUDA.
 */
#ifndef SALSA_B200_H
#define SALSA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SALSA_OK 0
#define SALSA_ERR_CUDA (-1)
#define SALSA_ERR_ARG (-2)
#define SALSA_ERR_NOMEM (-3)
#define SALSA_ERR_UNSUPPORTED (-4)

/* Data errors found by the BED kernels, mirroring the Python exception the
 * reference would die with (run_pipeline.py:166-170 turns any of them into
 * "step failed"). Reported in salsa_ingest_stats.err_code. */
#define SALSA_DATA_OK 0
#define SALSA_DATA_INDEX_ERROR 1   /* < 1 or < 4 fields   make_links.py:71,76-77  */
#define SALSA_DATA_VALUE_ERROR 2   /* bad start/end text  make_links.py:81-82     */
#define SALSA_DATA_KEY_ERROR 3     /* unknown contig / no RE counts  make_links.py:101-113 */
#define SALSA_DATA_ZERO_DIV 4      /* zero-length contig  make_links.py:106-107   */
#define SALSA_DATA_LINE_TOO_LONG 5 /* a line's first four fields, or a whole line
                                      that precedes a tile boundary, exceed
                                      SALSA_MAX_LINE bytes (library limit)       */

#define SALSA_DATA_AVOID_CHAIN 9   /* no longer produced: the avoid-mode walk-back has no limit any more */

#define SALSA_MAX_LINE 1024        /* bytes; see SALSA_DATA_LINE_TOO_LONG */

typedef struct salsa_ctx salsa_ctx;

const char *salsa_last_error(void);
int salsa_version(void);

/* One context per (process, device).  Holds the device-side contig table. */
int salsa_ctx_create(int device, salsa_ctx **out);
void salsa_ctx_destroy(salsa_ctx *ctx);
/* Tuning / test knob of the BED link scan (no counterpart in the reference; results never depend on it).  The scan
 * marks "special" bytes either exactly (<= 0x20 or '/') or with a cheaper superset test (<= 0x2F); with the superset,
 * lines whose first four fields hold one of the 14 characters 0x21..0x2E take the slow full parser.  mode 0 (default):
 * a sample of the text decides per call; 1: always exact; 2: always the superset. */
int salsa_ctx_scan_mode(salsa_ctx *ctx, int mode);
/* Device timers (the reference itself has no timers: run_pipeline.py only logs commands).
 * After salsa_timers_enable(ctx, 1) every BED link scan launched through this context is bracketed by a pair of CUDA
 * events on its stream.  salsa_timers waits for the events recorded so far and reports accumulated milliseconds and
 * launch counts: slot 0 = BED link scans (salsa_bed_ingest_*); slots 1..3 are reserved and read 0.  n = number of
 * slots to write; reset != 0 clears the accumulators afterwards.  Synchronises on the recorded events only. */
int salsa_timers_enable(salsa_ctx *ctx, int on);
int salsa_timers(salsa_ctx *ctx, double *h_out_ms, uint64_t *h_out_launches, int32_t n, int reset);

/*
 * Contig table: replaces the three host dicts of make_links.py:37-48
 * (contig_lengths, RF_counts_left, RF_counts_right) and the string compare of
 * make_links.py:90.
 *   h_names_blob / h_name_off[n+1] : contig names, STRICTLY INCREASING in byte
 *       order (memcmp, shorter prefix first) so that id order == Python-2
 *       string order; the id of a contig is its index.
 *   h_len_floor[i] : floor(float(length)); the B/E rule of make_links.py:108-139
 *       is exactly  B <=> start+end <= floor(len)  (DESIGN.md).
 *   h_flags[i]     : SALSA_CTG_* bits.
 * Names that appear only in links_avoid_iteration_N are listed too, without
 * SALSA_CTG_HAS_LEN.  Synchronises.  Device memory held by the context: the
 * open-addressing name table (32 bytes per slot; 64 slots per name up to 2^20
 * slots, at least 2 per name beyond that -- sparse on purpose, a second probe
 * costs the scanner a whole L2 round trip), 16 bytes per contig and the names.
 */
#define SALSA_CTG_HAS_LEN 1u
#define SALSA_CTG_HAS_RE 2u
int salsa_names_build(salsa_ctx *ctx, const char *h_names_blob, const int64_t *h_name_off, int32_t n,
                      const int64_t *h_len_floor, const uint32_t *h_flags);
int32_t salsa_names_count(const salsa_ctx *ctx);

/* links_avoid_iteration_N (make_links.py:29-35): id pairs, both orders implied.
 * n == 0 clears the set.  Synchronises. */
int salsa_avoid_set(salsa_ctx *ctx, const int32_t *h_id_pairs, int32_t npairs);

/*
 * Ingest statistics, 4 x uint64 in DEVICE memory, accumulated across calls.
 * Reset with salsa_ingest_stats_reset before the first chunk.
 *   [0] err   : min over errors of (byte offset of the offending line << 8 | code);
 *               UINT64_MAX when clean
 *   [1] lines : lines seen as "current" (make_links.py:60 minus the first line)
 *   [2] pairs : adjacent-line pairs counted into the table
 *   [3] reserved
 */
typedef struct {
    uint64_t err_offset;
    int32_t err_code;
    uint64_t lines;
    uint64_t pairs;
} salsa_ingest_stats;
int salsa_ingest_stats_reset(uint64_t *d_stats, void *stream);
/* copies to the host and decodes; synchronises the stream */
int salsa_ingest_stats_read(const uint64_t *d_stats, salsa_ingest_stats *h_out, void *stream);

/*
 * BED ingest -> dense contig-end link histogram.
 * Replaces make_links.py:60-146 (the per-line loop: split, adjacent-line
 * pairing, read-name test, end classification, contig_links[key] += 1).
 *
 *   d_bed[0..nbytes) : raw BED text whose FIRST byte starts a line.  Every
 *       adjacent-line pair inside the buffer is evaluated exactly as the
 *       reference evaluates it; the first line only ever plays "previous"
 *       (make_links.py:66-68).  To ingest a file in chunks (or shards on
 *       several GPUs) start each chunk at the first byte of the previous
 *       chunk's LAST line: chunks overlap by exactly one line.  With a
 *       non-empty avoid set (salsa_avoid_set) the scan reproduces the
 *       reference's stale prev_line rule (make_links.py:72-74: an avoided
 *       pair `continue`s WITHOUT advancing prev_line), and a chunk must
 *       additionally start on a line whose contig is in no avoid pair.
 *   d_table : uint64 [n*n*4], n = salsa_names_count(); cell
 *       ((first*n + second)*4 + ends) with first < second, ends bit1 = first
 *       is 'E', bit0 = second is 'E'.  Counts are ADDED (zero it first).
 *   base_offset : byte offset of d_bed[0] in the whole file (only used to
 *       report error positions).
 * Asynchronous on `stream`.  No scratch memory, no allocation.
 */
int salsa_bed_ingest_dense(salsa_ctx *ctx, const void *d_bed, size_t nbytes, uint64_t base_offset,
                           uint64_t *d_table, uint64_t *d_stats, void *stream);

/*
 * Same scan, sparse link table for high-cardinality assemblies (dense table
 * would not fit): open-addressing hash table in caller memory.
 *   d_keys   : uint64 [capacity] (capacity a power of two), initialised to
 *              UINT64_MAX by salsa_sparse_reset; key = (first*n + second)*4 + ends
 *   d_counts : uint64 [capacity], zeroed by salsa_sparse_reset
 * If the table fills up the err code is SALSA_DATA_TABLE_FULL.
 */
#define SALSA_DATA_TABLE_FULL 6
int salsa_sparse_reset(uint64_t *d_keys, uint64_t *d_counts, size_t capacity, void *stream);
int salsa_bed_ingest_sparse(salsa_ctx *ctx, const void *d_bed, size_t nbytes, uint64_t base_offset,
                            uint64_t *d_keys, uint64_t *d_counts, size_t capacity,
                            uint64_t *d_stats, void *stream);

/*
 * Convenience for host-resident BEDs (what make_links.py does with a file):
 * streams h_bed through two device staging buffers of `chunk_bytes` each
 * (caller-owned, d_stage[2*chunk_bytes]) with copies overlapped with the
 * kernel, splitting at line starts (with an avoid set: at lines whose contig is
 * in no avoid pair).  h_bed should be pinned for full PCIe speed.  The work is
 * ordered after what the caller enqueued on `stream` (zeroing the table) and
 * `stream` waits for it; the call returns when h_bed has been consumed.
 */
int salsa_bed_ingest_dense_host(salsa_ctx *ctx, const void *h_bed, size_t nbytes, void *d_stage,
                                size_t chunk_bytes, uint64_t *d_table, uint64_t *d_stats, void *stream);

/*
 * Table -> ordered edge list (replaces the dict walk of make_links.py:161-169;
 * the reference's own order is Python-2 dict order, which no consumer uses).
 * Writes the (key, count) of every non-empty cell in ascending cell order.
 *   dense table : d_keys == NULL, the cell index is the key -> output sorted
 *   sparse table: d_keys = the hash table's keys -> output in slot order; sort
 *                 it with salsa_sort_pairs_u64
 * d_scratch: salsa_links_compact_scratch_bytes(ncells) bytes.  With
 * d_out_keys == NULL only counts.  *h_n receives the number of links.
 * Synchronises the stream.
 */
size_t salsa_links_compact_scratch_bytes(size_t ncells);
int salsa_links_compact(const uint64_t *d_keys, const uint64_t *d_counts, size_t ncells,
                        uint64_t *d_out_keys, uint64_t *d_out_counts, size_t out_capacity,
                        void *d_scratch, uint64_t *h_n, void *stream);
/* Radix sort of (key, value) pairs by key (CUB; off the hot path).  Call with
 * d_tmp == NULL to get the scratch size in *tmp_bytes. */
int salsa_sort_pairs_u64(uint64_t *d_keys_in, uint64_t *d_vals_in, uint64_t *d_keys_out,
                         uint64_t *d_vals_out, size_t n, void *d_tmp, size_t *tmp_bytes, void *stream);

/*
 * Multi-GPU merge of per-GPU link tables (no counterpart in the
 * single-process reference; the link key space is what make_links.py:140-144
 * counts into).  Read pairs shard by BED byte range, so every GPU holds
 * partial counts for arbitrary keys:
 *   dense tables : one all-reduce (sum) of the uint64 table;
 *   sparse tables: salsa_links_compact, then salsa_links_partition buckets the
 *       (key, count) pairs by owner = mix64(key) % nparts (h_part_counts =
 *       send counts of the all-to-all; d_scratch = 64 uint64), the caller
 *       exchanges them (NCCL all-to-all), and salsa_sparse_add adds the
 *       received pairs into the owner's table (*d_full_flag becomes 1 when the
 *       table has no free slot).
 * Integer sums: the merged result does not depend on the number of shards.
 */
int salsa_links_partition(const uint64_t *d_keys, const uint64_t *d_counts, size_t n, int32_t nparts,
                          uint64_t *d_out_keys, uint64_t *d_out_counts, uint64_t *d_scratch,
                          uint64_t *h_part_counts, void *stream);
int salsa_sparse_add(const uint64_t *d_in_keys, const uint64_t *d_in_counts, size_t n, uint64_t *d_keys,
                     uint64_t *d_counts, size_t capacity, uint32_t *d_full_flag, void *stream);
/*
 * The same merge without compaction, sort or host round trip (the link key space of make_links.py:140-144; integer
 * sums, any shard order).  salsa_sparse_partition walks the sender's hash table ONCE and appends every link to the
 * segment of its owner = mix64(key) % nparts.  A segment is [u64 count][u64 pad][u64 keys[slot_cap]]
 * [u32 counts[slot_cap]] -- 12 bytes per link and salsa_sparse_segment_bytes(slot_cap) bytes in all; d_segments holds
 * nparts of them back to back, so the exchange is ONE all-to-all with equal splits.  slot_cap (a multiple of 4) bounds
 * the links per owner: *d_overflow (caller-zeroed) becomes 1 when a segment was too small or a count did not fit 32
 * bits -- retry with a larger slot_cap.  salsa_sparse_add_segments adds the nparts segments a rank received into its
 * table (*d_full_flag as in salsa_sparse_add).  Both are asynchronous on `stream`.
 */
size_t salsa_sparse_segment_bytes(size_t slot_cap);
int salsa_sparse_partition(const uint64_t *d_keys, const uint64_t *d_counts, size_t capacity, int32_t nparts,
                           size_t slot_cap, void *d_segments, uint32_t *d_overflow, void *stream);
int salsa_sparse_add_segments(const void *d_segments, int32_t nparts, size_t slot_cap, uint64_t *d_keys,
                              uint64_t *d_counts, size_t capacity, uint32_t *d_full_flag, void *stream);

/*
 * Python-2 str(float) on the device.  The reference passes scores between
 * programs as text written with str(float) = '%.12g' (make_links.py:169,
 * fast_scaled_scores.py:86; 12 digits evidenced by fast_scaled_scores.py:39-40),
 * so the next program sees float(str(x)).  For each x:
 *   digits : the 12-digit integer D of '%.12g' (sign carried; 0 for x == 0)
 *   exp10  : decimal exponent k of the leading digit (x ~ D * 10^(k-11))
 *   w12    : the double nearest to that decimal, i.e. float(str(x))
 * Exact (round-half-even on the binary value: 128-bit integers for 1e-8 <= |x| < 1e12, 320-bit ones beyond).
 * Values that have no digits: x = +inf / -inf / nan give digits = INT64_MIN + 1 / + 2 / + 3 and print as Python
 * prints them ('inf', '-inf', 'nan'); a finite |x| outside [1e-30, 1e30] gives digits = INT64_MIN, which the
 * salsa_format_* functions refuse with SALSA_ERR_UNSUPPORTED (the pipeline's own scores cannot get there).
 */
int salsa_g12(const double *d_x, size_t n, int64_t *d_digits, int32_t *d_exp10, double *d_w12, void *stream);

/*
 * Link scores.  Replaces make_links.py:104-137 (norm = RE*0.5 + RE*0.5 with
 * left counts for a 'B' end and right counts for an 'E' end) and :165-168
 * (score = 0 if norm == 0 else count*1.0/norm), plus the decimal round trip
 * into fast_scaled_scores.py:48,71.
 *   d_keys/d_counts : compacted table (salsa_links_compact), key = (first*n+second)*4+ends
 *   d_re_left/right : int64 [ncontig] by contig id
 * Outputs per link: score (raw double), is_int0 (the reference prints the int
 * literal 0), digits/exp10/w12 as in salsa_g12, and the two node ids
 * u = first*2 + (first end is E), v = second*2 + (second end is E).
 */
int salsa_links_score(const uint64_t *d_keys, const uint64_t *d_counts, size_t n, int32_t ncontig,
                      const int64_t *d_re_left, const int64_t *d_re_right, double *d_score,
                      uint8_t *d_is_int0, int64_t *d_digits, int32_t *d_exp10, double *d_w12,
                      int32_t *d_u, int32_t *d_v, void *stream);

/*
 * Best-buddy ratios.  Replaces fast_scaled_scores.py:35-89: node2weight (max
 * weight per contig end, :47-57), bestAlt = max of the two ends' best weights,
 * replaced -- when this edge IS the best at both ends -- by the best weight
 * among the OTHER edges at either end (:13-33,79-82), 0 -> int 1 (:83-84),
 * ratio = w / bestAlt (:86).  Weights must be >= 0.
 *   d_u, d_v : node ids in [0, nnodes); the graph must be simple
 *   d_w      : float(text of column 3)  (w12 of salsa_links_score)
 *   d_scratch: salsa_scaled_scratch_bytes(nnodes) bytes of device memory
 * Outputs per edge: best_alt, is_int1, ratio, and the str(float) digits of
 * ratio and of best_alt.
 */
size_t salsa_scaled_scratch_bytes(int32_t nnodes);
int salsa_scaled_scores(const int32_t *d_u, const int32_t *d_v, const double *d_w, size_t n, int32_t nnodes,
                        void *d_scratch, double *d_best_alt, uint8_t *d_is_int1, double *d_ratio,
                        int64_t *d_ratio_digits, int32_t *d_ratio_exp10, int64_t *d_best_digits,
                        int32_t *d_best_exp10, void *stream);

/*
 * salsa_scaled_scores for callers that cannot allocate.  (No counterpart in
 * the reference: this entry point exists for the hosts of this library, not
 * for parity with a reference function.)
 *
 * Who needs it: code that is handed only its input and output buffers and has
 * no way to come up with a device scratch buffer at call time -- e.g. a plugin
 * called inside a host program's stream-ordered section where cudaMalloc (a
 * device-wide synchronising call) is forbidden, a pipeline stage whose memory
 * pool is frozen while it runs, or a minimal cgo/JNI-style binding that has no
 * allocator on the device side.  Same inputs, outputs and results as
 * salsa_scaled_scores; the per-node working arrays live in a fixed array that
 * is part of the loaded library image.
 *
 *   - Limit: nnodes <= salsa_scaled_scores_static_max_nodes() (65536 contig
 *     ends = 32768 contigs); more is rejected with SALSA_ERR_ARG and a message
 *     that names the limit.  The array is about 1.5 MB and stays resident on
 *     every GPU that loads the library until the process exits.
 *   - Calls cudaMalloc / cudaFree / any allocating routine: never.
 *   - The first call on each device looks up the array's device address (a
 *     one-time, per-device set-up done with that device current).  Make that
 *     first call before starting any stream capture on that device.
 *   - All calls on a device share the one array.  Its contents after a call
 *     are leftovers; each call clears what it needs before use.  Two calls
 *     whose kernels can overlap on the GPU (different streams, or different
 *     host threads without ordering) race on it: serialise them.
 */
int32_t salsa_scaled_scores_static_max_nodes(void);
int salsa_scaled_scores_static(const int32_t *d_u, const int32_t *d_v, const double *d_w, size_t n, int32_t nnodes,
                               double *d_best_alt, uint8_t *d_is_int1, double *d_ratio,
                               int64_t *d_ratio_digits, int32_t *d_ratio_exp10, int64_t *d_best_digits,
                               int32_t *d_best_exp10, void *stream);

/*
 * Restriction-site counting.  Replaces RE_sites.py:6-28 (parse_fasta) and
 * RE_sites.py:45-88 (per contig, per expanded enzyme pattern, the starts of
 * non-overlapping re.finditer matches, split at len/2).  The FASTA text is in
 * device memory; record names are read by the host from the header lines.
 *
 *  1. salsa_fasta_headers: byte offsets of every '>' that starts a line
 *     (RE_sites.py:11), unordered; *d_count may exceed `capacity` (then retry
 *     with a larger buffer).
 *  2. The host sorts them, reads the header lines and forms the sequence
 *     spans [beg, end) = text between a header line's '\n' and the next header.
 *  3. salsa_fasta_count: per 4 KiB text tile (salsa_fasta_tiles(nbytes) tiles)
 *     the number of KEPT bytes: bytes of a span that are not '\n' and not
 *     trailing whitespace of their line (ln.rstrip(), RE_sites.py:22).
 *     salsa_exclusive_scan_u64 turns d_tile_cnt[tiles + 1] into offsets.
 *  4. salsa_fasta_span_lengths: kept-byte prefix at each span's start
 *     (d_pbeg) and the span's sequence length (d_plen = len(seq)).
 *  5. salsa_fasta_pack: sequences -> bit planes.  Record r starts at 64-base
 *     block d_blk_off[r] (blocks = ceil(len/64), caller-computed prefix,
 *     nspan + 1 entries); block b holds three uint64: hi, lo, valid with
 *     A=00 C=01 G=10 T=11 and valid=0 for every other character (sequence N,
 *     lower case and stray bytes never match, RE_sites.py:66).  d_planes must
 *     be zeroed by the caller (3 * nblk uint64).
 *  6. salsa_re_count: all patterns in one pass.  `patterns` = npat expanded
 *     patterns ([ACGT]+, each <= 63 chars, NUL-terminated, concatenated; the
 *     N expansion of RE_sites.py:52-56 is done by the host; a pattern listed
 *     twice counts twice).  d_left/d_right: uint64 [nrec], ADDED to.
 *     d_occ_scratch: nblk uint64 per self-overlapping pattern (e.g. AAAA), may
 *     be NULL when no pattern overlaps itself.
 * d_err (int32, caller-zeroed) becomes non-zero when a run of trailing
 * whitespace longer than 1024 bytes is met (library limit).
 */
int salsa_fasta_headers(const void *d_fasta, size_t nbytes, uint64_t *d_pos, uint32_t *d_count, uint32_t capacity,
                        void *stream);
size_t salsa_fasta_tiles(size_t nbytes);
int salsa_fasta_count(const void *d_fasta, size_t nbytes, const uint64_t *d_span_beg, const uint64_t *d_span_end,
                      int32_t nspan, uint64_t *d_tile_cnt, int32_t *d_err, void *stream);
int salsa_exclusive_scan_u64(uint64_t *d_vals, size_t n, void *stream);
int salsa_fasta_span_lengths(const void *d_fasta, size_t nbytes, const uint64_t *d_span_beg,
                             const uint64_t *d_span_end, int32_t nspan, const uint64_t *d_tile_excl,
                             uint64_t *d_pbeg, uint64_t *d_plen, int32_t *d_err, void *stream);
int salsa_fasta_pack(const void *d_fasta, size_t nbytes, const uint64_t *d_span_beg, const uint64_t *d_span_end,
                     int32_t nspan, const uint64_t *d_tile_excl, const uint64_t *d_pbeg, const uint64_t *d_blk_off,
                     uint64_t *d_planes, int32_t *d_err, void *stream);
int salsa_re_count(const uint64_t *d_planes, const uint64_t *d_blk_off, const uint64_t *d_seqlen, int32_t nrec,
                   uint64_t nblk, const char *patterns, int32_t npat, uint64_t *d_left, uint64_t *d_right,
                   uint64_t *d_occ_scratch, void *stream);

/*
 * BED rewrite of the cleaning step (-m yes).  Replaces the streaming loop of correct.py:62-92: a line on a contig that
 * was not cut is copied as its first four fields; a line on a cut contig is dropped when it straddles the cut, keeps
 * its coordinates as `contig_1` left of it, and becomes `contig_2` with both coordinates shifted otherwise.
 *   d_cut[id]       : cut position of the contig (salsa_names_build ids), -1 = not cut
 *   d_line_start[i] : byte offset of line i in d_bed (every line, in file order)
 * salsa_clean_bed_plan writes what each line will emit (bytes; 0 = dropped) and the first error as
 * (offset << 8 | code) in *d_err (UINT64_MAX when clean; 1 IndexError, 2 ValueError, 3 KeyError -- where the reference
 * would die).  The caller turns the lengths into offsets -- and zeroes the lengths of the lines the reference loses:
 * its last, partial buffer of 1,000,000 lines is never written (correct.py:87-92) -- and salsa_clean_bed_emit writes.
 */
int salsa_clean_bed_plan(salsa_ctx *ctx, const void *d_bed, size_t nbytes, const int64_t *d_cut, const int64_t *d_line_start,
                         size_t nlines, uint32_t *d_out_len, uint64_t *d_err, void *stream);
int salsa_clean_bed_emit(salsa_ctx *ctx, const void *d_bed, size_t nbytes, const int64_t *d_cut, const int64_t *d_line_start,
                         size_t nlines, const uint32_t *d_out_len, const int64_t *d_out_off, void *d_out, void *stream);

/*
 * RE counts of the scaffolds of iterations >= 2.  Replaces layout_unitigs.py:736-814 and refactor_breaks.py:63-131:
 * the contigs of a scaffold lie end to end; contigs wholly left / right of the midpoint (integer length/2) give all
 * their sites to that side, the contig(s) touching the midpoint are split in proportion
 * (float(left+right)/len * part, summed in path order in double precision, int() at the end); a single-contig
 * scaffold keeps (left, right) of its contig, swapped when it is reversed.
 *   d_path_off[nscaf+1], d_path_contig : the contigs of each scaffold in path order (CSR)
 *   d_first_is_b[s] : the scaffold's first node is the contig's 'B' end (only read for single-contig scaffolds)
 *   d_contig_len, d_re_left, d_re_right : int64 per contig id (re_counts_iteration_1)
 * Results are bit-identical to the Python loop while the sums stay below 2^53.
 */
int salsa_scaffold_re(const int64_t *d_path_off, const int32_t *d_path_contig, const uint8_t *d_first_is_b,
                      const int64_t *d_contig_len, const int64_t *d_re_left, const int64_t *d_re_right, int32_t nscaf,
                      int64_t *d_out_left, int64_t *d_out_right, void *stream);

/*
 * Hi-C physical coverage for misassembly detection.
 * Replaces break_contigs_start.cpp:307-460 (run once with -m yes,
 * run_pipeline.py:120-121) and break_contigs.cpp:199-356 (every iteration,
 * run_pipeline.py:222-230,327-335).
 *
 * Step 1, BED scan -> coverage difference arrays (one pass over the BED text,
 * same chunking rule as salsa_bed_ingest_dense: chunks overlap by one line):
 *   contig table: salsa_names_build with h_len_floor = the length read as
 *       `long` (break_contigs_start.cpp:286-294); d_len holds the same values
 *       by contig id on the device.
 *   d_covoff[id] : element offset of that contig's int32 difference array
 *       (len + 10 elements, break_contigs_start.cpp:335) inside d_diff, or
 *       UINT64_MAX for contigs shorter than 1,000,000 (never analysed, :325).
 *   d_diff       : zeroed by the caller; diff[span_start] += 1,
 *       diff[span_end + 1] -= 1 per accepted read pair (:375-376 / :242-243).
 *   d_touched[id]: set to 1 when any line lands on an analysed contig (the
 *       reference creates its coverage vector then, :333-337).
 *   first_chunk  : non-zero when d_bed starts at the FIRST line of the file:
 *       the reference never parses that line, so the pair (line 0, line 1) is
 *       not evaluated (:315-319).
 *   joins variant: d_join_off[n+1] / d_join_pos = CSR of the KEPT join
 *       positions per scaffold (pos - 2,000,000 >= 0 and pos + 2,000,000 <= len,
 *       break_contigs.cpp:84-87), ascending; a pair counts only when
 *       prev_end <= start and [prev_start, end + 1] lies inside some join's
 *       +-2 Mb window (:235-238).
 * Data conditions reported through d_stats (salsa_ingest_stats):
 *   SALSA_DATA_MALFORMED: a line on which `iss >> contig >> start >> end >>
 *       read` fails; the reference's loop stops there (:321-322 / :212-213), so
 *       the caller must zero d_diff / d_touched and rescan the first
 *       err_offset bytes only.
 *   SALSA_DATA_RANGE: an index the reference's vector::at() would reject
 *       (it aborts).
 *
 * Step 2, salsa_cov_analyse: for each segment an inclusive prefix sum of the
 * difference values (coverage), its exact integer total, and for the 11
 * thresholds cutoff = (total/len)/div, div = 5..15, the [start, end] of
 * Kadane's maximum-sum window over delta = (cov < cutoff ? +5 : -5) restricted
 * to [lo, hi), with the reference's tie rules (break_contigs_start.cpp:412-449,
 * break_contigs.cpp:298-335).  A segment is a whole contig (src_off ==
 * dst_off: in place; lo = sz/100, hi = 9*sz/10, sz = len + 10) or the window
 * [pos - 2 Mb, pos + 2 Mb) of a join (seed_idx = index of diff[start - 1], the
 * RAW difference value the reference seeds its local prefix with,
 * break_contigs.cpp:283-285; dst in a separate buffer; lo = sz/3, hi = 2*sz/3).
 * h_start/h_end: [nseg * 11], relative to the segment.  Synchronises.
 */
#define SALSA_DATA_MALFORMED 7
#define SALSA_DATA_RANGE 8
typedef struct {
    uint64_t src_off;   /* first element in d_diff */
    uint64_t dst_off;   /* first element in d_cov */
    uint64_t seed_idx;  /* element of d_diff added before the first element; UINT64_MAX = none */
    int64_t len;        /* elements */
    int64_t lo, hi;     /* Kadane range [lo, hi) */
} salsa_cov_seg;
int salsa_cov_scan_start(salsa_ctx *ctx, const void *d_bed, size_t nbytes, uint64_t base_offset, int32_t first_chunk,
                         const uint64_t *d_covoff, const int64_t *d_len, int32_t *d_diff, uint8_t *d_touched,
                         uint64_t *d_stats, void *stream);
int salsa_cov_scan_joins(salsa_ctx *ctx, const void *d_bed, size_t nbytes, uint64_t base_offset, int32_t first_chunk,
                         const uint64_t *d_covoff, const int64_t *d_len, const uint64_t *d_join_off,
                         const int64_t *d_join_pos, int32_t *d_diff, uint8_t *d_touched, uint64_t *d_stats,
                         void *stream);
size_t salsa_cov_scratch_bytes(const salsa_cov_seg *h_segs, int32_t nseg);
int salsa_cov_analyse(const int32_t *d_diff, int32_t *d_cov, const salsa_cov_seg *h_segs, int32_t nseg,
                      void *d_scratch, size_t scratch_bytes, int64_t *h_sum, int64_t *h_start, int64_t *h_end,
                      void *stream);

/*
 * The two coverage programs as one call: main() of break_contigs_start.cpp:271-463
 * (joins = 0; argv = -a BED -l LEN -s N) and of break_contigs.cpp:162-358
 * (joins = 1; argv = -a BED -b BREAKPOINTS -l LEN -i N -s N), argv WITHOUT the
 * program name, cmdline.h's short / --long=value / --long value forms.  The
 * report the reference prints to stdout goes to out_path (NULL = stdout).
 * d_bed != NULL: the BED (all nbytes of the -a file) is already in device
 * memory and the file is not read.  Returns the exit status of the program
 * (134 where the reference aborts on vector::at()).
 */
int salsa_break_contigs_run(int32_t joins, int32_t argc, const char *const *argv, const void *d_bed, size_t nbytes,
                            const char *out_path);

/*
 * Re-projection of the iteration-1 BED into scaffold coordinates: the BED
 * pass of update_bed, layout_unitigs.py:816-886 (refactor = 0) and
 * refactor_breaks.py:133-203 (refactor = 1).  The whole BED is resident at
 * d_bed; contigs are those of salsa_names_build; per contig id:
 * d_scaffold_of (scaffold index, -1 = in no scaffold), d_offset (start of
 * the contig inside its scaffold), d_length, d_reversed (path orientation
 * E..B).  Scaffold names: concatenated bytes + d_scaf_name_off[nscaf + 1].
 *  1. salsa_reproject_candidates: one scan of the BED; appends (unordered)
 *     the file offset of every line that closes a kept pair (refactor = 0)
 *     or opens a same-contig run (refactor = 1) to d_cand; *d_count gets the
 *     number found, which may exceed `capacity` -- then grow and repeat.
 *     Data errors (missing fields or '/', bad integers) land in d_stats as
 *     for the ingest; the reference raises IndexError / ValueError there.
 *  2. salsa_sort_u64 (CUB) puts them in file order = output order; max_key
 *     (0 = unknown) bounds the keys, so that only the bits a key can have
 *     are sorted (offsets into a 10 GB BED: 34 bits, 5 radix passes of 8).
 *  3. salsa_reproject_emit works in blocks of 256 consecutive candidates
 *     (nblk = salsa_reproject_blocks(ncand)).  With d_out == NULL it writes
 *     each candidate's output byte count to d_len[ncand] and each block's to
 *     d_blk[nblk] (and 0 to d_blk[nblk]); after salsa_exclusive_scan_u64 over
 *     nblk + 1 entries (d_blk[nblk] = total) the same call with d_out
 *     writes "scaffold\tstart\tend\tread\n" twice per kept
 *     pair (:864-871).
 */
int salsa_reproject_candidates(salsa_ctx *ctx, const void *d_bed, size_t nbytes, int32_t refactor,
                               const int32_t *d_scaffold_of, uint64_t *d_cand, size_t capacity,
                               uint64_t *d_count, uint64_t *d_stats, void *stream);
int salsa_sort_u64(uint64_t *d_in, uint64_t *d_out, size_t n, uint64_t max_key, void *d_tmp, size_t *tmp_bytes, void *stream);
size_t salsa_reproject_blocks(size_t ncand);
int salsa_reproject_emit(salsa_ctx *ctx, const void *d_bed, size_t nbytes, int32_t refactor,
                         const uint64_t *d_cand, size_t ncand, const int32_t *d_scaffold_of,
                         const int64_t *d_offset, const int64_t *d_length, const uint8_t *d_reversed,
                         const char *d_scaf_names, const uint32_t *d_scaf_name_off, uint32_t *d_len,
                         uint64_t *d_blk, void *d_out, size_t out_capacity, uint64_t *d_stats, void *stream);

/*
 * Host-side text assembly (no float conversion happens here: text is built
 * from the integer digits the device produced).
 * salsa_format_g12   : str(float) strings, concatenated; h_off[n+1] offsets
 * salsa_format_links : contig_links_iteration_N lines (make_links.py:169);
 *                      returns bytes written or -1
 * salsa_format_scaled: contig_links_scaled_iteration_N lines
 *                      (fast_scaled_scores.py:86); returns bytes written or -1
 */
int salsa_format_g12(const int64_t *h_digits, const int32_t *h_exp10, size_t n, char *h_out, size_t out_cap,
                     int64_t *h_off);
int64_t salsa_format_links(const char *names_blob, const int64_t *name_off, const int64_t *first,
                           const int64_t *second, const uint8_t *ends, const int64_t *count,
                           const int64_t *digits, const int32_t *exp10, const uint8_t *is_int0, size_t n,
                           char *out, size_t out_cap);
int64_t salsa_format_scaled(const char *node_blob, const int64_t *node_off, const int32_t *u, const int32_t *v,
                            const int64_t *w_digits, const int32_t *w_exp10, const int64_t *b_digits,
                            const int32_t *b_exp10, const uint8_t *b_is_int1, const int64_t *r_digits,
                            const int32_t *r_exp10, const int64_t *links, size_t n, char *out, size_t out_cap);

#ifdef __cplusplus
}
#endif
#endif /* SALSA_B200_H */
