/*
 * hpgq_host.h -- C host side of the B200-native hpg-fastq: command-line options, FASTQ text reader
 * (straight into pinned SoA batches), ordered FASTQ writer and the stats_report text writer.
 *
 * Mirrors, above the C ABI of include/hpg_fastq_cuda.h, the reference's host pieces (paths under the
 * reference tree):
 *   options      src/stats_options.c, src/filter_options.c, src/edit_options.c, src/commons_fastq.c
 *   reader       fastq_fread_se()  called at src/stats_fastq.c:183, src/filter_fastq.c:114, src/edit_fastq.c:121
 *   writer       fastq_fwrite()    called at src/filter_fastq.c:165-169, src/edit_fastq.c:190-198
 *   report       src/stats_report.c:32-563
 * Nothing here computes per-read statistics: that is the CUDA library's job and there is no CPU fallback.
 */
#ifndef HPGQ_HOST_H
#define HPGQ_HOST_H

#include <stddef.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/hpg_fastq_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ options */
enum hq_command { HQ_CMD_STATS = 0, HQ_CMD_FILTER = 1, HQ_CMD_EDIT = 2 };

#define HQ_DEFAULT_NUM_THREADS 2       /* src/stats_options.c:21 */
#define HQ_DEFAULT_BATCH_SIZE 262144   /* reference: 10000 (src/stats_options.c:22); results do not depend on it */
#define HQ_PHRED33_NAME "phred33"
#define HQ_PHRED64_NAME "phred64"

/* union of stats_options_t / filter_options_t / edit_options_t (src/stats_options.h:30-64 and siblings) */
typedef struct hq_options {
  int command;
  int kmers_on, filter_on, edit_on;
  int help, num_threads, batch_size;
  int quality_encoding_value;
  int min_read_length, max_read_length;
  int min_read_quality, max_read_quality;
  int left_length, min_left_quality, max_left_quality;
  int right_length, min_right_quality, max_right_quality;
  int max_N, max_out_of_quality;
  char *quality_encoding_name;
  char *read_length_range, *read_quality_range, *left_quality_range, *right_quality_range;
  char *in_filename, *out_dirname;
  char *in_filename1, *in_filename2; /* paired end (--fq1 / --fq2): the two mate files, run one after the other */
  int paired;
  char *exec_name, *command_name;
  /* extensions of this implementation (not in the reference CLI) */
  int gpus;            /* --gpus N: batches dealt round-robin over N devices (default 1)            */
  int max_read_len;    /* --max-read-length L: skip the pre-scan that sizes the per-position tables */
  int k2;              /* --kmers2 K: second k-mer size (BASELINE.json: 11), stats only             */
  int cg_k;            /* --chaos-game K: chaos-game word size (old/chaos_game.c), stats only       */
  int host_parse;      /* --host-parse: records are cut by hq_reader_next, not by the device parser */
  char *cg_signature;  /* --chaos-game-signature FILE: reference genomic signature for the difference image */
} hq_options_t;

/* Parses argv (argv[0] = command name, as in src/hpg-fastq.c:55-57).  Returns NULL after printing the
 * usage text when the command line is invalid or --help was given (the reference exits there). */
hq_options_t *hq_options_parse(const char *exec_name, const char *command_name, int argc, char **argv);
/* validate: 1 ok, 0 = invalid (message + usage already printed) */
int hq_options_validate(hq_options_t *o);
/* display: prints PARAMETERS CONFIGURATION, sets filter_on / edit_on; 0 = "nothing to do" */
int hq_options_display(hq_options_t *o, FILE *out);
/* NO_VALUE -> MIN_VALUE / MAX_VALUE (src/stats_fastq.c:431-444, filter_fastq.c:195-206, edit_fastq.c:230-241) */
void hq_options_fill_unset(hq_options_t *o);
void hq_options_free(hq_options_t *o);
void hq_usage(const char *exec_name, const char *command_name, int command, FILE *out);
/* src/commons_fastq.c:31-103 */
int hq_parse_range(int *min, int *max, const char *range, const char *msg);
/* options -> hpgq_params_t for the command (tuple orders of src/stats_fastq.c:209-214, src/edit_fastq.c:148-164) */
void hq_options_to_params(const hq_options_t *o, hpgq_params_t *p);

/* ------------------------------------------------------------------ FASTQ reader */
typedef struct hq_reader hq_reader_t;

/* SoA batch in caller-provided (pinned) memory + where each read's header line lives in the input */
typedef struct hq_host_batch {
  uint32_t n_reads;
  uint32_t max_len;
  uint8_t *seq, *qual; /* cap_bytes + 64 */
  uint32_t *idx;       /* cap_reads + 1  */
  uint64_t *hdr_off;   /* cap_reads: offset of the header line in the input file */
  uint32_t *hdr_len;   /* cap_reads: its length without the line terminator      */
  uint32_t cap_reads;
  uint64_t cap_bytes;
} hq_host_batch_t;

int hq_reader_open(const char *path, hq_reader_t **out);
void hq_reader_close(hq_reader_t *r);
const uint8_t *hq_reader_data(const hq_reader_t *r); /* the mapped input (headers are written from it) */
uint64_t hq_reader_size(const hq_reader_t *r);
/* longest line of the file and number of lines, one memchr pass (sizes Lcap before the first batch) */
int hq_reader_scan(hq_reader_t *r, uint32_t *max_line, uint64_t *n_lines);
/* same over the lines that start in the first `limit` bytes (filter / edit keep no per-position tables: a sample
 * is enough to size their batches) */
int hq_reader_scan_prefix(hq_reader_t *r, uint64_t limit, uint32_t *max_line, uint64_t *n_lines);
/* Fills b with up to b->cap_reads records / b->cap_bytes bases.  Returns 1 if reads were produced, 0 at EOF,
 * < 0 on a malformed record (message in hq_reader_error). */
int hq_reader_next(hq_reader_t *r, hq_host_batch_t *b);
const char *hq_reader_error(const hq_reader_t *r);
/* Device-parse path (hpgq_batch_submit_text): the next run of whole records, at most max_bytes long, as a span of
 * the mapped file -- nothing is scanned or copied except a short backward search for a record boundary (a line
 * that starts with '@' whose second successor starts with '+').  Returns 1 and sets *off / *len, 0 at end of
 * file, -1 when no boundary was found inside the budget (the caller falls back to hq_reader_next).  The reader
 * does not move until hq_reader_advance. */
int hq_reader_peek_text(const hq_reader_t *r, uint64_t max_bytes, uint64_t *off, uint64_t *len);
void hq_reader_advance(hq_reader_t *r, uint64_t bytes, uint64_t records);
int hq_reader_fd(const hq_reader_t *r);                 /* the open input, for hpgq_batch_submit_text_fd */
void hq_reader_rewind(hq_reader_t *r, uint64_t pos);    /* back to a chunk the device parser handed back */

/* ------------------------------------------------------------------ FASTQ writer */
typedef struct hq_writer hq_writer_t;
int hq_writer_open(const char *path, hq_writer_t **out);
int hq_writer_close(hq_writer_t *w); /* 0 ok, < 0 if any write failed */
/* Writes output reads [from, to) of a (compacted) SoA result: record k is the header of source read
 * src[k] (src == NULL: k itself), then bytes [idx[k], idx[k+1]) of seq and qual.  Record layout:
 * header, sequence, "+", quality, each '\n'-terminated. */
int hq_writer_put(hq_writer_t *w, const uint8_t *file_data, const hq_host_batch_t *b, const uint8_t *seq,
                  const uint8_t *qual, const uint32_t *idx, const uint32_t *src, uint32_t from, uint32_t to);

/* finished FASTQ text (records composed on the device, hpgq_batch_wait_text) */
int hq_writer_put_raw(hq_writer_t *w, const uint8_t *text, uint64_t n);
/* flushes the buffered records and returns the descriptor (positioned at the end of what was written), < 0 after a
 * failed write: for hpgq_batch_wait_text_fd, which appends a device-composed batch itself */
int hq_writer_fd(hq_writer_t *w);

/* ------------------------------------------------------------------ report */
typedef struct hq_kmer {
  int id;
  char string[8];
  uint64_t counter;
  const uint64_t *counter_by_pos;
  int counter_by_pos_size;
} hq_kmer_t;

/* k-mer id -> text, A=0 C=1 G=2 T=3, first base most significant */
void hq_kmer_string(int id, int k, char *out);
/* descending by count, ties by ascending id */
void hq_kmers_sorted(const hpgq_layout_t *lay, const uint64_t *block, hq_kmer_t out[1024]);

/* stats_report() (src/stats_report.c:32-56): writes <outdir>/<basename(in)>.summary.txt and the data /
 * gnuplot files from a finalized counter block.  Returns 0 or -1 (errno set). */
int hq_stats_report(const hpgq_layout_t *lay, const uint64_t *block, const hq_options_t *opts);

/* ------------------------------------------------------------------ chaos-game post-processing (chaos_post.c) */
typedef struct hq_cg_diff_stats {
  int highest, lowest; /* chaos_game_data_t.highest_table_dif_value / lowest_table_dif_value */
  double mean, sd;     /* mean_table_dif_value / standard_deviation_table_dif_value          */
  double fq_norm, gs_norm;
} hq_cg_diff_stats_t;

/* flat row-major dim x dim tables (dim = 2^k), u64 cells as in the counter block; pix = dim * dim grey values */
void hq_cg_pixels_seq(const uint64_t *table_seq, uint32_t dim, uint64_t fq_word_count, uint8_t *pix);
void hq_cg_pixels_quality(const uint64_t *table_seq, const uint64_t *table_q, uint32_t dim, int k, uint8_t *pix);
void hq_cg_pixels_diff(const int32_t *dif, uint32_t dim, uint8_t *pix);
/* genomic-signature file: 196-byte header_gs_t, then dim rows of dim 32-bit cells (old/chaos_game.c:298-320) */
int hq_cg_load_signature(const char *path, uint32_t dim, uint32_t *table_gs, uint32_t *ref_word_count);
/* normalised difference matrix + its range, mean and standard deviation (old/chaos_game.c:322-405); -1 if a
 * word count is 0 (the reference aborts there) */
int hq_cg_diff(const uint64_t *table_seq, const uint32_t *table_gs, uint32_t dim, uint64_t fq_word_count,
               uint32_t ref_word_count, int32_t *dif, hq_cg_diff_stats_t *st);
/* binary PGM; k < 7 is zoomed to 128 x 128 (old/chaos_game.c:514-593) */
int hq_cg_write_pgm(const char *path, const uint8_t *pix, uint32_t dim, int k);
/* <out_dir>/<basename(fq_path)>_k=<k>_FG.pgm, _QQ.pgm and, with a signature file, _FG_dif.pgm
 * (old/chaos_game.c:407-465).  0 ok, -1 I/O error, -2 no words counted, -3 bad signature file */
int hq_cg_write_images(const hpgq_layout_t *lay, const uint64_t *block, const char *fq_path, const char *out_dir,
                       const char *signature_path, hq_cg_diff_stats_t *st_out);

#ifdef __cplusplus
}
#endif
#endif
