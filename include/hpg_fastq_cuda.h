/*
 * hpg_fastq_cuda.h -- C ABI of the B200-native per-read QC hot path of hpg-fastq.
 *
 * Plain C linkage, plain pointers and sizes, no C++ / torch types.  Every entry
 * point returns 0 on success or a negative hpgq_status code; nothing inside the
 * library exits, aborts or falls back to the CPU.
 *
 * What each entry point replaces in the reference (paths under the reference tree):
 *
 *   hpgq_batch_submit / hpgq_batch_wait     the three workflow *stage functions*
 *       fastq_stats_worker   src/stats_fastq.c:202-251   (fastq_filter + fastq_reads_stats)
 *       fastq_filter_worker  src/filter_fastq.c:134-155  (fastq_filter)
 *       fastq_edit_worker    src/edit_fastq.c:142-178    (fastq_edit, then fastq_filter)
 *   device-side counter accumulation         fastq_stats_consumer src/stats_fastq.c:257-417
 *   hpgq_counters_download                   the stats_counters_t handed to stats_report()
 *                                            src/stats_fastq.h:35-73, src/stats_report.h:18
 *   hpgq_params_t.filter / .edit             fastq_filter_options_new(12 ints) src/stats_fastq.c:209-214,
 *                                            fastq_edit_options_new(9 ints)    src/edit_fastq.c:148-151
 *   hpgq_batch_t                             fastq_batch_t as used by old/chaos_game.c:174-193
 *                                            (seq, quality, data_indices[num_reads+1])
 *   k-mer tables                             kmer_t counters, src/stats_fastq.c:385-410
 *   chaos-game tables                        chaos_game_fill_tables, old/chaos_game.c:165-267
 */
#ifndef HPG_FASTQ_CUDA_H
#define HPG_FASTQ_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ status */
typedef enum hpgq_status {
  HPGQ_OK = 0,
  HPGQ_ERR_INVALID = -1,   /* bad argument / limits exceeded            */
  HPGQ_ERR_CUDA = -2,      /* a CUDA runtime call failed (see strerror) */
  HPGQ_ERR_NOMEM = -3,
  HPGQ_ERR_STATE = -4,     /* wait without submit, slot busy, ...       */
  HPGQ_ERR_NO_DEVICE = -5, /* no usable sm_100 device: there is NO CPU fallback */
  HPGQ_ERR_NCCL = -6
} hpgq_status;

const char *hpgq_strerror(int status);
/* last CUDA error text recorded by this thread's most recent failing call */
const char *hpgq_last_error(void);

/* ------------------------------------------------------------------ options */
/* "unset" values of the reference CLI: src/commons_fastq.h:21-23 */
#define HPGQ_NO_VALUE (-1)
#define HPGQ_MIN_VALUE 0
#define HPGQ_MAX_VALUE 100000

/* The 12-int tuple of fastq_filter_options_new(), in call-site order
 * (src/stats_fastq.c:209-214, src/filter_fastq.c:140-145, src/edit_fastq.c:159-164). */
typedef struct hpgq_filter_opts {
  int32_t min_read_length, max_read_length;
  int32_t min_read_quality, max_read_quality; /* Phred scale, inclusive */
  int32_t max_out_of_quality;
  int32_t left_length, min_left_quality, max_left_quality;
  int32_t right_length, min_right_quality, max_right_quality;
  int32_t max_N;
} hpgq_filter_opts_t;

/* The 9-int tuple of fastq_edit_options_new() (src/edit_fastq.c:148-151);
 * the last three are always 0 in the reference program. */
typedef struct hpgq_edit_opts {
  int32_t left_length, min_left_quality, max_left_quality;
  int32_t right_length, min_right_quality, max_right_quality;
  int32_t reserved0, reserved1, reserved2;
} hpgq_edit_opts_t;

/* work selection bits */
#define HPGQ_WORK_STATS 1u  /* per-read + per-position statistics merged into the counters */
#define HPGQ_WORK_FILTER 2u /* pass/fail mask; with STATS, statistics cover passed reads only */
#define HPGQ_WORK_EDIT 4u   /* 5'/3' window trim (then FILTER on the trimmed read if set)   */
#define HPGQ_WORK_KMERS 8u  /* 5-mer totals + per-position, and k2-mer totals if k2 > 0      */
#define HPGQ_WORK_CHAOS 16u /* chaos-game frequency/quality matrices, word size cg_k         */
#define HPGQ_WORK_COMPACT 32u /* order-preserving partition of (trimmed) reads into passed / failed */

typedef struct hpgq_params {
  uint32_t work;           /* OR of HPGQ_WORK_*                                     */
  int32_t phred;           /* 33 or 64; ASCII offset applied to Phred thresholds     */
  int32_t k2;              /* second k-mer size (BASELINE: 11), 0 = off, <= 13       */
  int32_t cg_k;            /* chaos-game word size, 1..12                            */
  int32_t cg_base_quality; /* chaos_game_data_t.base_quality (old/chaos_game.h:92)   */
  int32_t kernel_variant;  /* 0 = auto (tiled fast path), 1 = force simple warp-per-read path,
                            * 2 = auto + one-pass edit / filter / compaction kernel (HPGQ_WORK_COMPACT) */
  hpgq_filter_opts_t filter;
  hpgq_edit_opts_t edit;
} hpgq_params_t;

/* ------------------------------------------------------------------ batches */
/* SoA batch == fastq_batch_t: read i is bytes [idx[i], idx[i+1]) of seq and of qual. */
typedef struct hpgq_batch {
  uint32_t n_reads;
  uint32_t reserved; /* device batches: length hint, see hpgq_batch_submit_device; host batches: ignored */
  const uint8_t *seq;
  const uint8_t *qual;
  const uint32_t *idx; /* n_reads + 1 offsets, idx[0] == 0 for host batches */
} hpgq_batch_t;

typedef struct hpgq_limits {
  uint32_t max_reads_per_batch; /* device/pinned buffers are sized for this          */
  uint64_t max_bytes_per_batch; /* total bases per batch, < 2^32 - 64                 */
  uint32_t max_read_len;        /* per-position tables have this many entries (Lcap)  */
  uint32_t n_slots;             /* 1..4 double-buffered stream slots (default 2)      */
  uint32_t host_staging;        /* 1: allocate pinned staging for hpgq_batch_submit   */
  uint32_t k2, cg_k;            /* table sizes to reserve: 4^k2 and 4^cg_k words      */
  uint32_t kmers;               /* 1: reserve the 5-mer (and per-position) tables     */
} hpgq_limits_t;

/* per-read trim record produced by EDIT: the kept range is [start, start+len) */
typedef struct hpgq_trim {
  uint32_t start;
  uint32_t len;
} hpgq_trim_t;

/* Result views; host pointers are valid until the slot is submitted again. */
typedef struct hpgq_batch_result {
  uint32_t n_reads;
  uint32_t n_passed, n_failed, n_edited;
  const uint8_t *mask;     /* n_reads bytes: 1 = passed, 0 = failed (all 1 without FILTER) */
  const hpgq_trim_t *trim; /* n_reads records (EDIT), else NULL                            */
  /* COMPACT: passed reads then failed reads, each in input order */
  const uint8_t *out_seq, *out_qual; /* passed bytes at [0, passed_bytes), failed after     */
  const uint32_t *out_idx;           /* (n_passed + n_failed + 1) offsets into out_seq/qual */
  const uint32_t *out_src;           /* source read index of every output read              */
  uint64_t passed_bytes, failed_bytes;
} hpgq_batch_result_t;

/* ------------------------------------------------------------------ counters */
/* Dense u64 mirror of stats_counters_t (src/stats_fastq.h:35-73) as ONE flat block so
 * that multi-GPU merge is a single sum-allreduce.  Section offsets (in u64 words)
 * come from hpgq_layout(); sections that are not reserved have size 0. */
enum hpgq_section {
  HPGQ_SEC_SCALARS = 0, /* HPGQ_S_* below, 32 words                                  */
  HPGQ_SEC_LEN_HIST,    /* Lcap+1: reads of each length      (kh_length_histogram)   */
  HPGQ_SEC_SUMQ_BY_LEN, /* Lcap+1: sum of raw quality bytes of reads of that length   */
  HPGQ_SEC_QUAL_HIST,   /* 256: round(mean raw quality)      (kh_quality_histogram)  */
  HPGQ_SEC_GC_HIST,     /* 104: 100*(G+C)/len                (kh_gc_histogram)       */
  HPGQ_SEC_POS_ACCQ,    /* Lcap: sum of raw quality at pos   (kh_acc_quality_per_nt) */
  HPGQ_SEC_POS_A,       /* Lcap each                          (kh_num_As_per_nt ...) */
  HPGQ_SEC_POS_C,
  HPGQ_SEC_POS_G,
  HPGQ_SEC_POS_T,
  HPGQ_SEC_POS_N,
  HPGQ_SEC_K5,     /* 1024 5-mer totals (kmer_t.counter); row sums of K5_POS, derived at download */
  HPGQ_SEC_K5_POS, /* 1024 x Lcap, [id*Lcap + start_pos]  (kmer_t.counter_by_pos)     */
  HPGQ_SEC_K2,     /* 4^k2 totals                                                     */
  HPGQ_SEC_CG_SEQ, /* 2^cg_k x 2^cg_k, [co_x*dim + co_y]  (table_seq)                 */
  HPGQ_SEC_CG_Q,   /* same shape                           (table_q)                  */
  HPGQ_SEC_COUNT
};

enum hpgq_scalar {
  HPGQ_S_NUM_READS = 0, /* reads merged into the statistics (passed only under FILTER) */
  HPGQ_S_NUM_PASSED,
  HPGQ_S_NUM_FAILED,
  HPGQ_S_NUM_EDITED,
  HPGQ_S_ACC_LENGTH,
  HPGQ_S_NUM_A, /* base totals, derived at download from the per-position tables */
  HPGQ_S_NUM_C,
  HPGQ_S_NUM_G,
  HPGQ_S_NUM_T,
  HPGQ_S_NUM_N,
  HPGQ_S_MIN_LENGTH, /* derived at download from LEN_HIST (100000 if no reads: src/stats_fastq.c:24) */
  HPGQ_S_MAX_LENGTH,
  HPGQ_S_CG_WORDS,      /* fq_word_count                                             */
  HPGQ_S_CG_LONG_RUNS,  /* informational: axis runs >= 21 bases / single-axis reads seen (re-emulated exactly) */
  HPGQ_S_SLOW_READS,    /* reads that took the exact per-byte path (non-ACGTN bytes)  */
  HPGQ_S_BATCHES,
  HPGQ_S_BAD_BATCH,     /* non-zero: a batch broke the length statement it was submitted with (outputs invalid) */
  HPGQ_S_CG_FIXED_READS, /* informational: reads re-run in the reference's double arithmetic (old/chaos_game.c:200-251) */
  HPGQ_S_CG_MOVED_WORDS, /* informational: chaos-game words whose cell differs between the double and the integer form */
  HPGQ_S_COUNT = 32
};

typedef struct hpgq_layout {
  uint32_t max_read_len; /* Lcap */
  uint32_t kmers, k2, cg_k;
  uint64_t offset[HPGQ_SEC_COUNT + 1]; /* offset[HPGQ_SEC_COUNT] == total words */
} hpgq_layout_t;

/* pure host arithmetic, usable without a GPU */
int hpgq_layout(uint32_t max_read_len, uint32_t kmers, uint32_t k2, uint32_t cg_k, hpgq_layout_t *out);

/* ------------------------------------------------------------------ context */
typedef struct hpgq_ctx hpgq_ctx_t;

int hpgq_device_count(int *n);
int hpgq_ctx_create(int device, const hpgq_limits_t *limits, hpgq_ctx_t **out);
int hpgq_ctx_destroy(hpgq_ctx_t *ctx);
int hpgq_set_params(hpgq_ctx_t *ctx, const hpgq_params_t *params);
int hpgq_get_layout(hpgq_ctx_t *ctx, hpgq_layout_t *out);

/* Host batch: packs seq/qual/idx into the slot's pinned staging buffers, ships them with
 * cudaMemcpyAsync on the slot's stream and launches the kernels behind the copy. */
int hpgq_batch_submit(hpgq_ctx_t *ctx, const hpgq_batch_t *host_batch, int slot);
/* Device-resident batch: pointers are device pointers, 16-byte aligned, each array readable
 * 32 bytes past its logical end; idx need not start at 0.  No copies are made.  The batch's byte span
 * idx[n] - idx[0] must not exceed hpgq_limits_t.max_bytes_per_batch.
 * dev_batch->reserved: 0 = the library measures the shortest / longest read on the device first (one stream
 * synchronisation); L > 0 = the caller states that no read is longer than L (no synchronisation);
 * L | HPGQ_BATCH_UNIFORM = the caller states that EVERY read has exactly L bases, which selects the
 * uniform-length kernels.  A statement that turns out false is detected on the device: the batch's outputs are
 * then invalid and hpgq_batch_wait / hpgq_batch_sync / hpgq_counters_download return HPGQ_ERR_INVALID. */
#define HPGQ_BATCH_UNIFORM 0x80000000u
int hpgq_batch_submit_device(hpgq_ctx_t *ctx, const hpgq_batch_t *dev_batch, int slot);
/* FASTQ text batch: the device-side replacement of the record scan of fastq_fread_se()
 * (reference call sites src/stats_fastq.c:183, src/filter_fastq.c:114, src/edit_fastq.c:121).  `text` is host memory
 * (pinned: copied asynchronously; pageable: staged through the slot's pinned buffer) holding n_bytes of FASTQ that
 * start at a header line and end after a quality line (less than 4 GiB; a chunk of at most 2 * max_bytes_per_batch
 * bytes can never exceed the ctx's base limit).  The text is copied
 * to the device, cut into records there (strict four-line records, LF or CRLF), packed into the slot's seq / qual /
 * idx arrays and processed exactly as hpgq_batch_submit would process that batch.  The call returns after the
 * parse (one stream synchronisation) and fills *info.  info->status != 0 means the chunk was NOT submitted (the
 * slot stays free): HPGQ_TEXT_LINES (line count not a multiple of four, e.g. blank lines), HPGQ_TEXT_RECORD
 * (record info->bad_record, counted from 0 in this chunk, has no '@' / '+' or sequence and quality differ in
 * length), HPGQ_TEXT_TOO_MANY (more reads or bases than the ctx limits); the caller hands such a chunk to its
 * host-side reader, which owns the diagnostics. */
enum { HPGQ_TEXT_LINES = 1, HPGQ_TEXT_RECORD = 2, HPGQ_TEXT_TOO_MANY = 4 };
typedef struct hpgq_text_info {
  uint32_t n_reads, max_len;
  uint64_t n_bases;
  uint32_t n_lines;
  uint32_t status;     /* 0 or HPGQ_TEXT_* bits */
  uint32_t bad_record; /* first malformed record (HPGQ_TEXT_RECORD), else 0xFFFFFFFF */
  uint32_t reserved;
} hpgq_text_info_t;
int hpgq_batch_submit_text(hpgq_ctx_t *ctx, const uint8_t *text, uint64_t n_bytes, int slot, hpgq_text_info_t *info);
/* The same in two halves, so that a reader can stage the NEXT chunk while the device still parses this one.
 * hpgq_batch_submit_text_fd reads n_bytes at file_offset of an open file with pread() (a pool of host threads,
 * straight into a small ring of pinned pieces the ctx owns: no page faults of a mapped file, no extra copy, no
 * chunk-sized pinned buffer), queues the host -> device copy of every piece and the parse, and returns without waiting.  hpgq_batch_text_launch then waits for that parse only, fills *info as
 * hpgq_batch_submit_text does and launches the QC kernels when info->status == 0.  Replaces the fastq_fread_se()
 * call of the producers (src/stats_fastq.c:183, src/filter_fastq.c:114, src/edit_fastq.c:121). */
int hpgq_batch_submit_text_fd(hpgq_ctx_t *ctx, int fd, uint64_t file_offset, uint64_t n_bytes, int slot);
int hpgq_batch_text_launch(hpgq_ctx_t *ctx, int slot, hpgq_text_info_t *info);
/* drops a chunk submitted with hpgq_batch_submit_text_fd without launching anything for it (the slot is free again) */
int hpgq_batch_text_cancel(hpgq_ctx_t *ctx, int slot);
/* Header line of every read of the slot's text batch: byte offset inside `text` and length without the line
 * terminator (n_reads entries each).  Valid between hpgq_batch_submit_text and the next submit on the slot. */
int hpgq_batch_text_headers(hpgq_ctx_t *ctx, int slot, uint32_t *hdr_off, uint32_t *hdr_len);
/* hpgq_batch_wait for a text batch submitted with HPGQ_WORK_COMPACT: instead of the compacted seq / qual / idx / src
 * arrays the device writes the output FASTQ records themselves (header of the source read, sequence, "+", quality,
 * each '\n'-terminated: the layout of fastq_fwrite(), reference call sites src/filter_fastq.c:165-169,
 * src/edit_fastq.c:190-198) and only that text comes back: *text (pinned, valid until the slot is reused) holds the
 * passed reads' records in input order in its first *passed_bytes bytes, then *failed_bytes bytes of failed reads'
 * records.  res gets the counts (n_reads, n_passed, n_failed, n_edited) and no array views. */
int hpgq_batch_wait_text(hpgq_ctx_t *ctx, int slot, hpgq_batch_result_t *res, const uint8_t **text,
                         uint64_t *passed_bytes, uint64_t *failed_bytes);
/* The same with the writing included (the whole fastq_fwrite() stage): the passed reads' records are appended to
 * fd_passed and the failed reads' records to fd_failed, each at the descriptor's file position, while the rest of
 * the text is still on its way back through the ctx's pinned output ring -- no chunk-sized host buffer.  A regular
 * file opened read-write is extended and filled through a mapping by the pool threads; any other descriptor gets
 * write().  (The mapping is used only while the file system reports room for the bytes plus a margin: a store into a
 * mapped page that cannot be backed would raise SIGBUS where write() returns ENOSPC.)  A descriptor may be -1 when the
 * batch has no records for it (edit without filter: no failed reads). */
int hpgq_batch_wait_text_fd(hpgq_ctx_t *ctx, int slot, hpgq_batch_result_t *res, int fd_passed, int fd_failed,
                            uint64_t *passed_bytes, uint64_t *failed_bytes);
/* Optional: allocates the text path's buffers (pinned rings, per-slot device chunk + parser scratch, and with
 * with_output != 0 the output text buffers) for chunks of up to chunk_bytes now instead of on the first chunk. */
int hpgq_ctx_reserve_text(hpgq_ctx_t *ctx, uint64_t chunk_bytes, int with_output);
/* Blocks until the slot's work is finished and fills the result views. */
int hpgq_batch_wait(hpgq_ctx_t *ctx, int slot, hpgq_batch_result_t *res);
/* Same but copies nothing back (device-resident measurement); counters keep accumulating. */
int hpgq_batch_sync(hpgq_ctx_t *ctx, int slot);
/* device views of the slot's outputs (mask / trim / compacted arrays), valid until resubmit */
int hpgq_batch_device_views(hpgq_ctx_t *ctx, int slot, hpgq_batch_result_t *res);

/* Compaction without any allocation.  This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  It is the order-preserving partition of HPGQ_WORK_COMPACT (passed
 * reads, then failed reads, each in input order, trimmed to d_trim when given) on caller-owned device arrays: the same
 * kernels, with their scratch (the exclusive scan of the per-read records, 16 bytes per read, and its block sums) in a
 * static device array inside the library instead of buffers from cudaMalloc.  It calls no cudaMalloc / cudaFree and
 * no allocating library routine, needs no ctx, and works on the device that is current on the calling thread (all
 * pointers must belong to it).
 *   inputs   d_seq, d_qual, d_idx[n_reads + 1]: the batch as in hpgq_batch_t (device pointers);
 *            d_mask[n_reads]: 1 = passed, 0 = failed, NULL = every read passed;
 *            d_trim[n_reads]: kept range of every read, NULL = whole reads;
 *            max_len: 0, or a length no read exceeds (selects the staged copy kernel for reads up to 22 000 bases;
 *            a value SMALLER than the longest read is a caller error with undefined output -- 0 is always safe);
 *   outputs  d_out_seq, d_out_qual: (idx[n] - idx[0]) + 64 bytes each; d_out_idx[n_reads + 1]; d_out_src[n_reads];
 *            d_totals[4] = passed reads, failed reads, passed bytes, failed bytes -- as hpgq_batch_result_t's views;
 *   stream   a cudaStream_t (NULL = the legacy default stream); the call only enqueues, the caller synchronises.
 * Limit: n_reads <= HPGQ_NOALLOC_MAX_READS; more is rejected with HPGQ_ERR_INVALID and a message.  The scratch array
 * (4 MB + 4 KB) is held by every GPU that loads the library until the process exits.  ALL calls on a device share it:
 * calls whose kernels may overlap on the GPU (different streams without an ordering between them, or two host threads)
 * race and produce garbage -- order them on one stream or with events.  Nothing in it has to be zero at entry (every
 * word a call reads it has written before), so the contents a previous call left behind do no harm. */
#define HPGQ_NOALLOC_MAX_READS 262144u
int hpgq_compact_noalloc(const uint8_t *d_seq, const uint8_t *d_qual, const uint32_t *d_idx, const uint8_t *d_mask,
                         const hpgq_trim_t *d_trim, uint32_t n_reads, uint32_t max_len, uint8_t *d_out_seq,
                         uint8_t *d_out_qual, uint32_t *d_out_idx, uint32_t *d_out_src, uint32_t *d_totals, void *stream);

/* Order-sensitive 64-bit checksum of the compacted output (HPGQ_WORK_COMPACT) a slot holds, computed on the
 * device; the slot must be idle (after hpgq_batch_wait / hpgq_batch_sync).  hpgq_checksum_host is the same sum over
 * host arrays: sum of hpgq_checksum_mix(position, value) (include/hpgq_synth.h) over out_seq[0..T), out_qual[0..T) (positions T + i),
 * out_idx[0..n] (2T + j) and out_src[0..n) (2T + n + 1 + j), T = passed + failed bytes.  Measurement / test aid
 * (SURVEY.md 8d "64-bit order-sensitive checksums"); replaces nothing in the reference. */
int hpgq_batch_checksum(hpgq_ctx_t *ctx, int slot, uint64_t out[4]); /* checksum, passed bytes, failed bytes, passed reads */
uint64_t hpgq_checksum_host(const uint8_t *out_seq, const uint8_t *out_qual, const uint32_t *out_idx,
                            const uint32_t *out_src, uint32_t n_reads, uint64_t total_bytes);

int hpgq_counters_reset(hpgq_ctx_t *ctx);
/* Copies the flat counter block (layout words) to host memory and fills the derived scalars. */
int hpgq_counters_download(hpgq_ctx_t *ctx, uint64_t *dst, size_t n_words);
/* Device pointer of the flat block, for an external sum-allreduce (NCCL / torch.distributed). */
int hpgq_counters_device_ptr(hpgq_ctx_t *ctx, uint64_t **dptr, size_t *n_words);
/* Multi-GPU merge: ONE ncclAllReduce(ncclSum, ncclUint64) over the flat blocks of n contexts that
 * live on n different devices of this process; afterwards every context holds the sum.  Replaces nothing in the
 * reference (its consumer is a single thread, src/stats_fastq.c:257-417); NCCL is bound with dlopen at the
 * first call.  n == 1 is a no-op. */
int hpgq_counters_reduce(hpgq_ctx_t **ctxs, int n);
/* Optional: creates the communicators hpgq_counters_reduce needs for contexts on these devices (in this order) now;
 * they are cached per device list for the life of the process.  May run in a host thread of its own while contexts
 * are created and batches submitted: creating communicators takes 0.5-1 s, the allreduce itself microseconds. */
int hpgq_comm_prepare(const int *devices, int n);
/* Fill the derived scalars (totals, min/max length) of a host copy, e.g. after an allreduce. */
int hpgq_counters_finalize_host(const hpgq_layout_t *layout, uint64_t *block);

/* CUDA stream (cudaStream_t) of a slot, so callers can time with events on it. */
int hpgq_slot_stream(hpgq_ctx_t *ctx, int slot, void **stream);
/* number of kernels this ctx has launched since creation (bench "gpu_launches") */
int hpgq_launch_count(hpgq_ctx_t *ctx, uint64_t *n);
/* milliseconds spent in the dominant QC kernel of the slot's last batch (CUDA events) */
int hpgq_last_kernel_ms(hpgq_ctx_t *ctx, int slot, float *ms);

/* ------------------------------------------------------------------ synthetic reads */
/* Counter-based generator (include/hpgq_synth.h); the device kernel and the host function
 * emit identical bytes for the same (cfg, first_read). */
typedef struct hpgq_synth_cfg {
  uint64_t seed;
  uint32_t min_len, max_len; /* fixed length when equal */
  uint32_t short_pct;        /* percent of reads drawn from [short_min, max_len) when min_len==max_len */
  uint32_t short_min;
  uint32_t cap_axis_runs;    /* 1: cap A/T and G/T runs (chaos-game configs)  */
  uint32_t qual_shift;       /* subtracted from every Phred value (mate 2)    */
  uint32_t lowercase_ppm;    /* parts-per-million of bases emitted lower-case (exactness tests) */
  uint32_t reserved;
} hpgq_synth_cfg_t;

/* Fills device arrays for reads [first_read, first_read + n_reads): d_idx gets n_reads+1 offsets
 * starting at 0; d_seq/d_qual must hold the bases plus 32 bytes.  total_bytes returns idx[n]. */
int hpgq_synth_device(hpgq_ctx_t *ctx, const hpgq_synth_cfg_t *cfg, uint64_t first_read, uint32_t n_reads,
                      uint8_t *d_seq, uint8_t *d_qual, uint32_t *d_idx, uint64_t capacity_bytes,
                      uint64_t *total_bytes);
/* Host twin (no GPU needed). */
int hpgq_synth_host(const hpgq_synth_cfg_t *cfg, uint64_t first_read, uint32_t n_reads, uint8_t *seq,
                    uint8_t *qual, uint32_t *idx, uint64_t capacity_bytes, uint64_t *total_bytes);

/* raw device memory helpers so a C host (or ctypes) can hold device-resident batches */
int hpgq_dev_alloc(hpgq_ctx_t *ctx, size_t bytes, void **dptr);
int hpgq_dev_free(hpgq_ctx_t *ctx, void *dptr);
int hpgq_dev_download(hpgq_ctx_t *ctx, void *dst_host, const void *src_dev, size_t bytes);
int hpgq_dev_upload(hpgq_ctx_t *ctx, void *dst_dev, const void *src_host, size_t bytes);
int hpgq_host_alloc_pinned(size_t bytes, void **ptr);
int hpgq_host_free_pinned(void *ptr);

const char *hpgq_version(void);

#ifdef __cplusplus
}
#endif

#endif /* HPG_FASTQ_CUDA_H */
