// hpgq_internal.h -- kernel argument block and launcher prototypes shared by the .cu files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/hpg_fastq_cuda.h"

// Everything a QC kernel needs, passed by value (fits the 4 KB kernel-parameter space easily).
struct hq_kargs {
  // batch (device pointers, see hpgq_batch_submit_device for alignment/padding rules)
  const uint8_t *seq;
  const uint8_t *qual;
  const uint32_t *idx;
  uint32_t n_reads;
  uint32_t work; // HPGQ_WORK_*
  // filter thresholds, already mapped to raw-ASCII means (Phred + offset) and clamped
  int32_t min_len, max_len;
  int32_t q_lo, q_hi; // inclusive bounds on the mean raw quality byte
  int32_t max_ooq, max_n;
  uint32_t c_lo, c_hi; // hq_ooq_const_lo/hi of the same bounds
  int32_t f_ll, f_lq_lo, f_lq_hi; // filter left window (length, raw-ASCII mean bounds)
  int32_t f_rl, f_rq_lo, f_rq_hi;
  int32_t e_ll, e_lq_lo, e_lq_hi; // edit (trim) left window
  int32_t e_rl, e_rq_lo, e_rq_hi;
  int32_t phred;
  int32_t k2, cg_k, cg_base_quality;
  uint32_t Lcap;
  // flat counter block and section offsets (u64 words)
  unsigned long long *ctr;
  uint32_t o_len_hist, o_sumq, o_qual_hist, o_gc_hist, o_accq, o_A, o_C, o_G, o_T, o_N;
  uint64_t o_k5, o_k5pos, o_k2, o_cg_seq, o_cg_q;
  unsigned long long *cg_pack; // packed chaos-game shadow table (count << 38 | quality sum), NULL = not used
  unsigned long long *kc_pack; // packed k2-mer + chaos-game table (4^k2 cells, same packing), NULL = not used
  // per-read outputs
  uint8_t *mask;     // may be NULL
  hpgq_trim_t *trim; // may be NULL
  uint8_t *cgflag;   // chaos game: reads to re-emulate in double arithmetic (n_reads bytes), may be NULL
  uint32_t *flist;   // one-pass statistics of long reads: [0] = number of failed reads, then their indices (n_reads + 1 words)
  uint32_t *rsum;    // one-pass statistics of long reads: quality sum, G+C | N << 16 per read (all zero between batches), may be NULL
  // tiled-kernel geometry
  uint32_t RT;          // reads per tile (multiple of 4)
  uint32_t G;           // lanes per read group
  uint32_t stage_bytes; // capacity of one seq (and one qual) stage buffer
  uint32_t n_tiles;
  uint32_t n_stages;    // depth of the shared-memory ring
  uint32_t uni_len;     // uniform-length kernels: the length every read of the batch has
  uint32_t kc_bytes;    // k-mer / chaos kernel: 1 = byte loop for every read (HPGQ_KC_BYTES, A/B and tests)
};

// launchers (return cudaGetLastError()); `launches` is incremented per kernel launched
cudaError_t hq_launch_qc_simple(const hq_kargs &a, cudaStream_t st, uint64_t *launches);
// returns cudaErrorInvalidValue if the batch does not fit the tiled kernel (max read length > 640)
// uniform = true: the caller states that every read has exactly max_len_in_batch bases (the kernel verifies it and
// raises HPGQ_S_BAD_BATCH otherwise); the aligned-word variant is used when the geometry allows it
cudaError_t hq_launch_qc_tiled(hq_kargs a, uint32_t max_len_in_batch, bool uniform, int sm_count, cudaStream_t st,
                               uint64_t *launches);
bool hq_tiled_supports(uint32_t max_len_in_batch, uint32_t work);
// vectorised any-length path (hpgq_group.cu): decisions + per-read histograms, then per-position tables
cudaError_t hq_launch_qc_group(const hq_kargs &a, uint32_t max_len_in_batch, int sm_count, cudaStream_t st,
                               uint64_t *launches);
cudaError_t hq_launch_pos(const hq_kargs &a, uint32_t max_len_in_batch, int sm_count, cudaStream_t st,
                          uint64_t *launches);
// statistics (+ simple filter) of long reads in one pass over the batch (hpgq_group.cu); false = decision kernel + hq_launch_pos
bool hq_pos_fused_supports(const hq_kargs &a, uint32_t max_len_in_batch);
cudaError_t hq_launch_pos_fused(const hq_kargs &a, uint32_t max_len_in_batch, int sm_count, cudaStream_t st,
                                uint64_t *launches);
cudaError_t hq_launch_kmers_cg(const hq_kargs &a, uint32_t max_len_in_batch, cudaStream_t st, uint64_t *launches);
// d_out[0] = longest read, d_out[1] = shortest read, d_out[2] = 1 if some offset decreases
cudaError_t hq_launch_maxlen(const uint32_t *idx, uint32_t n, uint32_t *d_out, cudaStream_t st, uint64_t *launches);

struct hq_compact_bufs {
  uint4 *scan;       // n+1 entries: x = passed bytes, y = failed bytes, z = passed reads, w = failed reads (exclusive scan)
  uint4 *block_sums; // scratch, >= ceil((n+1)/1024) + 1 entries
  uint8_t *out_seq, *out_qual;
  uint32_t *out_idx, *out_src; // n+1 / n
  uint32_t *totals;            // [0]=n_passed [1]=n_failed [2]=passed_bytes [3]=failed_bytes (device)
  // one-pass (fused) path: per-tile look-back records and the side buffers of the failed stream
  uint4 *fz_agg, *fz_incl;
  uint32_t *fz_status;         // fz_tiles_cap + 1 words (the last one is the tile ticket)
  uint32_t fz_tiles_cap;
  uint8_t *fail_seq, *fail_qual;
  uint32_t *fail_idx, *fail_src;
};
// error reporting of hpgq_api.cu (thread-local hpgq_last_error text) for entry points that live in other files
int hq_fail_msg(int code, const char *what);
int hq_fail_cuda(cudaError_t e, const char *what);

// decisions (trim, filter) and compaction in one pass over the batch; false = this batch needs the two-pass path
bool hq_fz_supports(uint32_t work, uint32_t max_len, uint32_t n_reads, const hq_compact_bufs &b);
// decisions only (edit / filter without statistics), word frame; cudaErrorInvalidValue = use hq_launch_qc_group
cudaError_t hq_launch_dec_fast(const hq_kargs &a, uint32_t max_len, int sm_count, cudaStream_t st, uint64_t *launches);
cudaError_t hq_launch_fz(const hq_kargs &a, const hq_compact_bufs &b, uint32_t max_len, int sm_count, cudaStream_t st,
                         uint64_t *launches);
cudaError_t hq_launch_compact(const hq_kargs &a, const hq_compact_bufs &b, uint32_t mean_len, uint32_t max_len,
                              cudaStream_t st, uint64_t *launches);

cudaError_t hq_launch_checksum(const hq_compact_bufs &b, uint32_t n, unsigned long long *d_out, cudaStream_t st,
                               uint64_t *launches);

cudaError_t hq_launch_synth(const hpgq_synth_cfg_t &cfg, uint64_t first_read, uint32_t n, uint8_t *d_seq,
                            uint8_t *d_qual, const uint32_t *d_idx, cudaStream_t st, uint64_t *launches);

// ------------------------------------------------------------------ device FASTQ text parser (hpgq_parse.cu)
enum { HQ_PINFO_LINES = 0, HQ_PINFO_READS, HQ_PINFO_MAXLEN, HQ_PINFO_BAD, HQ_PINFO_BASES, HQ_PINFO_STATUS, HQ_PINFO_MINLEN, HQ_PINFO_COUNT = 8 };
// bits of info[HQ_PINFO_STATUS] = hpgq_text_info_t.status (include/hpg_fastq_cuda.h HPGQ_TEXT_*)
enum { HQ_PSTAT_LINES = 1, HQ_PSTAT_RECORD = 2, HQ_PSTAT_TOO_MANY = 4 };
struct hq_parse_bufs {
  const uint8_t *text;  // chunk on the device, 16-byte aligned, readable to n_bytes + 16
  uint32_t n_bytes;
  uint32_t *tile_cnt;   // ceil(n_bytes / 16384) + 1
  uint32_t *cta_sum;    // 8 * sm_count + 1
  uint32_t *nl;         // cap_lines positions of the line breaks
  uint32_t cap_lines;
  uint32_t *info;       // HQ_PINFO_COUNT words
  uint32_t *seq_off, *qual_off, *hdr_off, *hdr_len;  // cap_reads each
  uint32_t cap_reads;
  uint64_t cap_bytes;
  uint8_t *seq, *qual;  // outputs: the SoA batch
  uint32_t *idx;
};
cudaError_t hq_launch_parse(const hq_parse_bufs &b, int sm_count, cudaStream_t st, uint64_t *launches);
// FASTQ text of a text batch's compacted output (passed records, then failed); rec: n + 1 words of scratch,
// oinfo: [0] = bytes of the passed records, [1] = bytes of all records
cudaError_t hq_launch_compose(const hq_parse_bufs &b, uint32_t n, const hq_compact_bufs &cb, uint32_t *rec, uint32_t *oinfo,
                              uint8_t *otext, int sm_count, cudaStream_t st, uint64_t *launches);
