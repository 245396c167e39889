/*
 * report.c -- the stats_report text files, written from the flat u64 counter block.
 *
 * Follows src/stats_report.c line by line for WHAT is printed (file names, formats, float32 expressions, loop
 * bounds, the "%i" of a size_t, rows that are skipped); the khash tables it reads are the dense sections of
 * hpgq_counters_download (SURVEY.md 8a row 11 and the quirk list below the table):
 *   kh_length_histogram  -> HPGQ_SEC_LEN_HIST        kh_quality_histogram -> HPGQ_SEC_QUAL_HIST
 *   kh_gc_histogram      -> HPGQ_SEC_GC_HIST         kh_acc_quality_per_nt -> HPGQ_SEC_POS_ACCQ
 *   kh_count_quality_per_nt[j] = #reads longer than j = sum of LEN_HIST[L], L > j
 *   kh_num_{A,C,G,T,N}s_per_nt -> HPGQ_SEC_POS_{A,C,G,T,N}
 *   acc_quality (float sum of per-read means) = sum over L of SUMQ_BY_LEN[L] / L, exact integers, one divide per length
 * Deviations, all where the reference is undefined: GC row 100 and mean-quality keys >= 100 overflow its stack
 * arrays (src/stats_report.c:217-226,406-418) -- here they are simply not printed / printed from a 256 array;
 * the .gnuplot text is the intended text (the reference passes it as a printf format, :606-647).
 */
#include "hpgq_host.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define SEC(lay, blk, s) ((blk) + (lay)->offset[s])

void hq_kmer_string(int id, int k, char *out) {
  static const char nt[4] = {'A', 'C', 'G', 'T'};
  for (int i = 0; i < k; i++) out[i] = nt[(id >> (2 * (k - 1 - i))) & 3];
  out[k] = 0;
}

static int kmer_cmp(const void *a, const void *b) {
  const hq_kmer_t *x = (const hq_kmer_t *)a, *y = (const hq_kmer_t *)b;
  if (x->counter != y->counter) return x->counter > y->counter ? -1 : 1;
  return x->id - y->id;
}

void hq_kmers_sorted(const hpgq_layout_t *lay, const uint64_t *block, hq_kmer_t out[1024]) {
  /* src/stats_fastq.c:475-481: ids and strings are filled, then qsort by count (ties: ascending id) */
  const uint64_t *k5 = SEC(lay, block, HPGQ_SEC_K5), *kp = SEC(lay, block, HPGQ_SEC_K5_POS);
  const int has = lay->offset[HPGQ_SEC_K5 + 1] > lay->offset[HPGQ_SEC_K5];
  for (int i = 0; i < 1024; i++) {
    out[i].id = i;
    hq_kmer_string(i, 5, out[i].string);
    out[i].counter = has ? k5[i] : 0;
    out[i].counter_by_pos = has ? kp + (uint64_t)i * lay->max_read_len : NULL;
    out[i].counter_by_pos_size = 0;
  }
  qsort(out, 1024, sizeof out[0], kmer_cmp);
}

typedef struct graph {
  char title[128], xlabel[128], ylabel[128], type[32];
  int x_autoscale, x_start, x_end, y_autoscale, x_column, num_y_columns;
  int y_columns[5];
  char y_titles[5][32];
} graph_t;

static void graph_init(graph_t *g, const char *title, const char *xl, const char *yl, const char *type, int x_end) {
  memset(g, 0, sizeof *g);
  snprintf(g->title, sizeof g->title, "%s", title);
  snprintf(g->xlabel, sizeof g->xlabel, "%s", xl);
  snprintf(g->ylabel, sizeof g->ylabel, "%s", yl);
  snprintf(g->type, sizeof g->type, "%s", type);
  g->x_autoscale = 0;
  g->x_start = 0;
  g->x_end = x_end;
  g->y_autoscale = 1;
  g->x_column = 1;
  g->num_y_columns = 1;
  g->y_columns[0] = 2;
}

/* _generate_gnuplot_image(), src/stats_report.c:592-654, minus the system("gnuplot") call (no gnuplot here) */
static int write_gnuplot(const graph_t *g, const char *data_filename, const char *prefix) {
  char path[4096];
  snprintf(path, sizeof path, "%s.gnuplot", prefix);
  FILE *f = fopen(path, "w");
  if (!f) return -1;
  fprintf(f, "set output '%s.png'\n", prefix);
  fprintf(f, "set terminal png nocrop enhanced font arial 10 size 640,360\n");
  fprintf(f, "set ylabel '%s'\n", g->ylabel);
  fprintf(f, "set xlabel '%s'\n", g->xlabel);
  fprintf(f, "set ytics border in scale 1,0.5 mirror norotate  offset character 0, 0, 0\n");
  fprintf(f, "set title '%s'\n", g->title);
  if (g->x_autoscale == 1) fprintf(f, "set autoscale x\n");
  else fprintf(f, "set xrange [ %i : %i ] noreverse nowriteback\n", g->x_start, g->x_end);
  if (g->y_autoscale == 1) fprintf(f, "set autoscale y\n");
  else fprintf(f, "set yrange [ %i : %i ] noreverse nowriteback\n", g->x_start, g->x_end);
  fprintf(f, "set lmargin '%i'\n", 10);
  fprintf(f, "set rmargin '%i'\n", 4);
  fprintf(f, "set tmargin '%i'\n", 3);
  fprintf(f, "set bmargin '%i'\n", 4);
  fprintf(f, "plot ");
  for (int i = 0; i < g->num_y_columns; i++)
    fprintf(f, "%s '%s' using %i:%i title '%s' with %s", i == 0 ? "" : ", ", data_filename, g->x_column, g->y_columns[i],
            g->y_titles[i], g->type);
  fprintf(f, "\n");
  return fclose(f);
}

/* the reference prints size_t values with "%i" (src/stats_report.c:179,230,423): the low 32 bits as an int */
static int as_i(uint64_t v) { return (int)(uint32_t)v; }

/* _normalize_quality, src/stats_report.c:26 */
static double normalize_quality(float quality, int phred) { return round(quality - phred); }

typedef struct ctx {
  const hpgq_layout_t *lay;
  const uint64_t *blk;
  const hq_options_t *o;
  const char *name; /* basename of the input */
  int max_length, phred;
  uint64_t *count_per_nt; /* kh_count_quality_per_nt */
} ctx_t;

static FILE *open_out(const ctx_t *c, const char *suffix, char *path, size_t n) {
  snprintf(path, n, "%s/%s.%s", c->o->out_dirname, c->name, suffix);
  return fopen(path, "w");
}

static int report_summary(const ctx_t *c, const hq_kmer_t *kmers) {
  const uint64_t *S = c->blk;
  char path[4096];
  FILE *f = open_out(c, "summary.txt", path, sizeof path);
  if (!f) return -1;
  const hq_options_t *opts = c->o;
  const uint64_t num_reads = S[HPGQ_S_NUM_READS], num_passed = S[HPGQ_S_NUM_PASSED], num_failed = S[HPGQ_S_NUM_FAILED];
  const uint64_t nA = S[HPGQ_S_NUM_A], nC = S[HPGQ_S_NUM_C], nG = S[HPGQ_S_NUM_G], nT = S[HPGQ_S_NUM_T], nN = S[HPGQ_S_NUM_N];
  fprintf(f, "-----------------------------------\n");
  fprintf(f, "      FastQ quality report\n");
  fprintf(f, "-----------------------------------\n");
  const uint64_t num_nucleotides = nA + nC + nG + nT + nN;
  fprintf(f, "FastQ filename: %s\n", c->name);
  fprintf(f, "\n");
  if (opts->filter_on) {
    fprintf(f, "Filter options:\n");
    if (opts->read_length_range) fprintf(f, "\tRead length range   : %s\n", opts->read_length_range);
    if (opts->read_quality_range) fprintf(f, "\tRead quality range  : %s\n", opts->read_quality_range);
    if (opts->left_length != HPGQ_MIN_VALUE && opts->left_quality_range) {
      fprintf(f, "\tLeft length         : %i nucleotides\n", opts->left_length);
      fprintf(f, "\tLeft quality range  : %s\n", opts->left_quality_range);
    }
    if (opts->right_length != HPGQ_MIN_VALUE && opts->right_quality_range) {
      fprintf(f, "\tRight length        : %i nucleotides\n", opts->right_length);
      fprintf(f, "\tRight quality range : %s\n", opts->right_quality_range);
    }
    if (opts->max_N != HPGQ_MAX_VALUE) fprintf(f, "\tMax. number of Ns   : %i\n", opts->max_N);
    if (opts->max_out_of_quality != HPGQ_MAX_VALUE && opts->read_quality_range)
      fprintf(f, "\tMax. out of quality : %i nucletotides\n", opts->max_out_of_quality);
    fprintf(f, "\n");
    fprintf(f, "Number of reads in file  : %lu\n", (unsigned long)(num_passed + num_failed));
    /* "%0.2f %)" prints a literal '%' on glibc (SURVEY.md Appendix C probe 2) */
    fprintf(f, "Number of processed reads: %lu (%0.2f %%)\n", (unsigned long)num_reads, 100.0f * num_reads / (num_passed + num_failed));
  } else {
    fprintf(f, "Filter         : Disabled\n");
    fprintf(f, "Number of reads: %lu\n", (unsigned long)num_reads);
  }
  fprintf(f, "\n");
  fprintf(f, "Read length (min., mean, max.): (%i, %0.2f, %i)\n", (int)S[HPGQ_S_MIN_LENGTH], 1.0f * S[HPGQ_S_ACC_LENGTH] / num_reads,
          c->max_length);
  fprintf(f, "\n");
  /* acc_quality: the reference adds each read's float mean into a float (src/stats_fastq.c:297); here the exact
   * sum of S/L by length, rounded to float once */
  const uint64_t *lh = SEC(c->lay, c->blk, HPGQ_SEC_LEN_HIST), *sq = SEC(c->lay, c->blk, HPGQ_SEC_SUMQ_BY_LEN);
  double accq = 0.0;
  for (int L = 1; L <= c->max_length; L++)
    if (lh[L]) accq += (double)sq[L] / (double)L;
  const float acc_quality = (float)accq;
  int qual = (int)normalize_quality(1.0f * acc_quality / num_reads, c->phred);
  fprintf(f, "Mean quality = %i [%c]\n", qual, qual + c->phred);
  fprintf(f, "\n");
  fprintf(f, "Nucleotide content (A, C, G, T, N)\n");
  fprintf(f, "\tA: %0.2f %%\n", 100.0f * nA / num_nucleotides);
  fprintf(f, "\tT: %0.2f %%\n", 100.0f * nT / num_nucleotides);
  fprintf(f, "\tG: %0.2f %%\n", 100.0f * nG / num_nucleotides);
  fprintf(f, "\tC: %0.2f %%\n", 100.0f * nC / num_nucleotides);
  fprintf(f, "\tN: %0.2f %%\n", 100.0f * nN / num_nucleotides);
  fprintf(f, "GC content\n");
  fprintf(f, "\tCG: %0.2f %%\n", 100.0f * (nG + nC) / num_nucleotides);
  fprintf(f, "\n");
  fprintf(f, "Mean quality per nucleotide position\n");
  const uint64_t *acc = SEC(c->lay, c->blk, HPGQ_SEC_POS_ACCQ);
  for (int k = 0; k < c->max_length; k++) {
    qual = (int)normalize_quality(1.0f * acc[k] / c->count_per_nt[k], c->phred);
    fprintf(f, "\tpos. %i: %i [%c]\t", k + 1, qual, qual + c->phred);
    if ((k + 1) % 5 == 0) fprintf(f, "\n");
  }
  fprintf(f, "\n");
  if (opts->kmers_on) {
    fprintf(f, "K-mers (top 20)\n");
    fprintf(f, "\tSequence\tCount\n");
    for (int i = 0; i < 21; i++) fprintf(f, "\t%s\t\t%lu\n", kmers[i].string, (unsigned long)kmers[i].counter);
  }
  return fclose(f);
}

/* report_length(), src/stats_report.c:160-203 */
static int report_length(const ctx_t *c) {
  char data[4096], prefix[4096];
  FILE *f = open_out(c, "length.histogram.data", data, sizeof data);
  if (!f) return -1;
  const uint64_t *lh = SEC(c->lay, c->blk, HPGQ_SEC_LEN_HIST);
  for (int i = 1; i <= c->max_length; i++) fprintf(f, "%i\t%i\n", i, as_i(lh[i]));
  if (fclose(f)) return -1;
  snprintf(prefix, sizeof prefix, "%s/%s.length.histogram", c->o->out_dirname, c->name);
  graph_t g;
  graph_init(&g, "Read Length Histogram", "Read length", "Number of reads", "boxes", c->max_length + 1);
  return write_gnuplot(&g, data, prefix);
}

/* report_quality(), src/stats_report.c:396-485.  Its quality.per.nt.data is overwritten afterwards by
 * report_nt_content() (call order :49-50), so only the histogram and the gnuplot file survive from here. */
static int report_quality(const ctx_t *c) {
  char data[4096], prefix[4096], str[64];
  FILE *f = open_out(c, "read.quality.histogram.data", data, sizeof data);
  if (!f) return -1;
  const uint64_t *qh = SEC(c->lay, c->blk, HPGQ_SEC_QUAL_HIST);
  int min_qual = 1000, max_qual = 0;
  for (int q = 0; q < 256; q++)
    if (qh[q]) {
      if (min_qual > q) min_qual = q;
      if (max_qual < q) max_qual = q;
    }
  for (int i = min_qual; i <= max_qual; i++) fprintf(f, "%i\t%i\n", i - c->phred, as_i(qh[i]));
  if (fclose(f)) return -1;
  snprintf(prefix, sizeof prefix, "%s/%s.read.quality.histogram", c->o->out_dirname, c->name);
  snprintf(str, sizeof str, "Read Quality (Phred%i scale)", c->phred);
  graph_t g;
  graph_init(&g, "Avg. Read Quality Histogram", str, "Number of reads", "boxes", max_qual - min_qual + 5);
  return write_gnuplot(&g, data, prefix);
}

/* report_nt_content(), src/stats_report.c:207-392 */
static int report_nt_content(const ctx_t *c) {
  char data[4096], prefix[4096], str[100];
  graph_t g;
  const uint64_t *gc = SEC(c->lay, c->blk, HPGQ_SEC_GC_HIST);
  FILE *f = open_out(c, "GC.histogram.data", data, sizeof data);
  if (!f) return -1;
  for (int i = 1; i < 100; i++)
    if (gc[i]) fprintf(f, "%i\t%i\n", i, as_i(gc[i]));
  if (fclose(f)) return -1;
  snprintf(prefix, sizeof prefix, "%s/%s.GC.histogram", c->o->out_dirname, c->name);
  graph_init(&g, "GC Content Histogram", "GC content (%)", "Number of reads", "boxes", 100);
  if (write_gnuplot(&g, data, prefix)) return -1;

  const uint64_t *hA = SEC(c->lay, c->blk, HPGQ_SEC_POS_A), *hC = SEC(c->lay, c->blk, HPGQ_SEC_POS_C);
  const uint64_t *hG = SEC(c->lay, c->blk, HPGQ_SEC_POS_G), *hT = SEC(c->lay, c->blk, HPGQ_SEC_POS_T);
  const uint64_t *hN = SEC(c->lay, c->blk, HPGQ_SEC_POS_N), *acc = SEC(c->lay, c->blk, HPGQ_SEC_POS_ACCQ);

  f = open_out(c, "GC.per.nt.data", data, sizeof data);
  if (!f) return -1;
  for (int k = 0; k < c->max_length; k++) {
    const uint64_t num_total = hA[k] + hT[k] + hG[k] + hC[k] + hN[k];
    float v;
    if ((v = 100.0f * (hG[k] + hC[k]) / num_total) > 1.0f) fprintf(f, "%i\t%0.2f\n", k + 1, v);
  }
  if (fclose(f)) return -1;
  snprintf(prefix, sizeof prefix, "%s/%s.GC.per.nt", c->o->out_dirname, c->name);
  graph_init(&g, "GC Content per Nucleotide Position", "Nucleotide position", "GC content (%)", "lines", c->max_length + 1);
  if (write_gnuplot(&g, data, prefix)) return -1;

  f = open_out(c, "quality.per.nt.data", data, sizeof data);
  if (!f) return -1;
  for (int k = 0; k < c->max_length; k++) {
    const int qual = (int)normalize_quality(1.0f * acc[k] / c->count_per_nt[k], c->phred);
    fprintf(f, "%i\t%i\n", k, qual);
  }
  if (fclose(f)) return -1;
  snprintf(prefix, sizeof prefix, "%s/%s.quality.per.nt", c->o->out_dirname, c->name);
  snprintf(str, sizeof str, "Quality (Phred%i scale)", c->phred);
  graph_init(&g, "Quality per Nucleotide Position", "Nucleotide position", str, "lines", c->max_length + 1);
  if (write_gnuplot(&g, data, prefix)) return -1;

  f = open_out(c, "nucleotides.data", data, sizeof data);
  if (!f) return -1;
  for (int k = 0; k < c->max_length; k++) {
    const uint64_t num_total = hA[k] + hT[k] + hG[k] + hC[k] + hN[k];
    fprintf(f, "%i\t%0.2f\t%0.2f\t%0.2f\t%0.2f\t%0.2f\n", k + 1, 100.0f * hA[k] / num_total, 100.0f * hT[k] / num_total,
            100.0f * hG[k] / num_total, 100.0f * hC[k] / num_total, 100.0f * hN[k] / num_total);
  }
  if (fclose(f)) return -1;
  snprintf(prefix, sizeof prefix, "%s/%s.nucleotides", c->o->out_dirname, c->name);
  graph_init(&g, "Nucleotide Content per Position", "Nucleotide position", "Nucleotide content (%)", "lines", c->max_length + 1);
  g.num_y_columns = 5;
  static const char *tt[5] = {"A %", "T %", "G %", "C %", "N %"};
  for (int i = 0; i < 5; i++) {
    g.y_columns[i] = 2 + i;
    snprintf(g.y_titles[i], sizeof g.y_titles[i], "%s", tt[i]);
  }
  return write_gnuplot(&g, data, prefix);
}

/* report_kmers(), src/stats_report.c:489-563 */
static int report_kmers(const ctx_t *c, hq_kmer_t *kmers) {
  char data[4096], prefix[4096];
  FILE *f = open_out(c, "kmers.txt", data, sizeof data);
  if (!f) return -1;
  fprintf(f, "# Sequence\tCount\n");
  for (int i = 0; i < 1024; i++) fprintf(f, "%s\t%lu\n", kmers[i].string, (unsigned long)kmers[i].counter);
  if (fclose(f)) return -1;
  f = open_out(c, "kmers.per.nt.data", data, sizeof data);
  if (!f) return -1;
  /* counter_by_pos_size: one entry per read position (the reference sizes it by the read, src/stats_fastq.c:396-405) */
  int num_cols = 0;
  for (int i = 0; i < 5; i++) {
    kmers[i].counter_by_pos_size = kmers[i].counter ? c->max_length : 0;
    if (num_cols < kmers[i].counter_by_pos_size) num_cols = kmers[i].counter_by_pos_size;
  }
  for (int i = 0; i < num_cols; i++) {
    fprintf(f, "%i", i + 1);
    for (int j = 0; j < 5; j++)
      fprintf(f, "\t%lu", (unsigned long)(kmers[j].counter_by_pos_size <= i ? 0 : kmers[j].counter_by_pos[i]));
    fprintf(f, "\n");
  }
  if (fclose(f)) return -1;
  snprintf(prefix, sizeof prefix, "%s/%s.kmers.per.nt", c->o->out_dirname, c->name);
  graph_t g;
  graph_init(&g, "Relative Enrichment over Read Length", "Nucleotide position", "Number of K-mers", "lines", num_cols + 1);
  g.num_y_columns = 5;
  for (int i = 0; i < 5; i++) {
    g.y_columns[i] = 2 + i;
    snprintf(g.y_titles[i], sizeof g.y_titles[i], "%s", kmers[i].string);
  }
  return write_gnuplot(&g, data, prefix);
}

/* extensions: second k-mer size and chaos-game matrices (BASELINE.json configs[3]); not reference files */
static int report_extensions(const ctx_t *c) {
  char data[4096];
  const hpgq_layout_t *lay = c->lay;
  if (lay->k2 && lay->offset[HPGQ_SEC_K2 + 1] > lay->offset[HPGQ_SEC_K2]) {
    char suffix[64], s[16];
    snprintf(suffix, sizeof suffix, "kmers.k%u.txt", lay->k2);
    FILE *f = open_out(c, suffix, data, sizeof data);
    if (!f) return -1;
    const uint64_t *t = SEC(lay, c->blk, HPGQ_SEC_K2);
    const uint64_t n = 1ull << (2 * lay->k2);
    fprintf(f, "# Sequence\tCount (non-zero %u-mers, id order)\n", lay->k2);
    for (uint64_t i = 0; i < n; i++)
      if (t[i]) {
        hq_kmer_string((int)i, (int)lay->k2, s);
        fprintf(f, "%s\t%lu\n", s, (unsigned long)t[i]);
      }
    if (fclose(f)) return -1;
  }
  if (lay->cg_k && lay->offset[HPGQ_SEC_CG_SEQ + 1] > lay->offset[HPGQ_SEC_CG_SEQ]) {
    /* table_seq / table_q of chaos_game_fill_tables (old/chaos_game.c:253-259), one matrix row per line */
    const uint32_t dim = 1u << lay->cg_k;
    for (int which = 0; which < 2; which++) {
      char suffix[64];
      snprintf(suffix, sizeof suffix, which ? "chaos.k%u.quality.data" : "chaos.k%u.freq.data", lay->cg_k);
      FILE *f = open_out(c, suffix, data, sizeof data);
      if (!f) return -1;
      const uint64_t *t = SEC(lay, c->blk, which ? HPGQ_SEC_CG_Q : HPGQ_SEC_CG_SEQ);
      for (uint32_t x = 0; x < dim; x++) {
        for (uint32_t y = 0; y < dim; y++) fprintf(f, y ? "\t%lu" : "%lu", (unsigned long)t[(uint64_t)x * dim + y]);
        fprintf(f, "\n");
      }
      if (fclose(f)) return -1;
    }
    /* the reference's own chaos-game outputs: the PGM images (old/chaos_game.c:407-465) and, with a signature
     * file, the figures of the difference matrix (:322-405) */
    hq_cg_diff_stats_t st;
    const int rc = hq_cg_write_images(lay, c->blk, c->o->in_filename, c->o->out_dirname, c->o->cg_signature, &st);
    if (rc == -1 || rc == -3) return -1;
    if (rc == 0 && c->o->cg_signature) {
      char suffix[64];
      snprintf(suffix, sizeof suffix, "chaos.k%u.diff.txt", lay->cg_k);
      FILE *f = open_out(c, suffix, data, sizeof data);
      if (!f) return -1;
      fprintf(f, "Fastq file normalization ratio = %12.6f\n", st.fq_norm);       /* old/chaos_game.c:339 */
      fprintf(f, "Genomic signature normalization ratio = %12.6f\n", st.gs_norm); /* :351 */
      fprintf(f, "Interval of variation of diff matrix values = [%d, %d]\n", st.highest, st.lowest); /* :375 */
      fprintf(f, "Mean of diff matrix values = %f\nStandard deviation of diff matrix values = %f\n", st.mean, st.sd);
      if (fclose(f)) return -1;
    }
  }
  return 0;
}

int hq_stats_report(const hpgq_layout_t *lay, const uint64_t *block, const hq_options_t *opts) {
  /* stats_report(), src/stats_report.c:32-56: basename of the input, then the five writers in this order */
  ctx_t c;
  memset(&c, 0, sizeof c);
  c.lay = lay;
  c.blk = block;
  c.o = opts;
  const char *p = strrchr(opts->in_filename, '/');
  c.name = p ? p + 1 : opts->in_filename;
  c.max_length = (int)block[HPGQ_S_MAX_LENGTH];
  c.phred = opts->quality_encoding_value;
  if ((uint32_t)c.max_length > lay->max_read_len) return -1;
  c.count_per_nt = (uint64_t *)calloc((size_t)c.max_length + 2, sizeof(uint64_t));
  if (!c.count_per_nt) return -1;
  const uint64_t *lh = SEC(lay, block, HPGQ_SEC_LEN_HIST);
  uint64_t run = 0;
  for (int j = c.max_length - 1; j >= 0; j--) {
    run += lh[j + 1];
    c.count_per_nt[j] = run;
  }
  hq_kmer_t *kmers = (hq_kmer_t *)calloc(1024, sizeof(hq_kmer_t));
  int rc = kmers ? 0 : -1;
  if (!rc) hq_kmers_sorted(lay, block, kmers);
  if (!rc) rc = report_summary(&c, kmers);
  if (!rc) rc = report_length(&c);
  if (!rc) rc = report_quality(&c);
  if (!rc) rc = report_nt_content(&c);
  if (!rc && opts->kmers_on) rc = report_kmers(&c, kmers);
  if (!rc) rc = report_extensions(&c);
  free(kmers);
  free(c.count_per_nt);
  return rc ? -1 : 0;
}
