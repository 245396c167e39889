/*
 * hpgq_oracle.c -- CPU restatement of hpg-fastq's per-read QC hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this.  The product (the CUDA library
 * behind include/hpg_fastq_cuda.h) never links or calls it.
 *
 * PARITY STATUS
 *   pinned   : chaos game (oracle_chaos_fill) -- checked in tests against the reference's own
 *              old/chaos_game.c compiled into oracle/_ref (see oracle/Makefile, target ref);
 *              merge semantics, option defaults and tuple orders -- restated from in-tree
 *              code cited below.
 *   UNPINNED : fastq_reads_stats / fastq_filter / fastq_edit / k-mer ids live in the
 *              bioinfo-libs submodule @5840c6781924d30a37cf58279d27a4b898cfb91e which is an
 *              empty directory in the reference tree (.SUBMODULES.json:8-15); the reference
 *              has no tests or golden vectors.  Their semantics here are
 *              the decisions written down in SURVEY.md section 8c (1)-(7) and DESIGN.md;
 *              "parity unpinned" for those rows.
 *
 * Every function cites the reference lines it follows (paths relative to the reference tree).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/hpg_fastq_cuda.h"

/* ------------------------------------------------------------------ layout */
/* Dense mirror of stats_counters_t (src/stats_fastq.h:35-73): every khash keyed by position
 * or value becomes a dense u64 array (the report indexes them densely). */
int oracle_layout(uint32_t Lcap, uint32_t kmers, uint32_t k2, uint32_t cg_k, hpgq_layout_t *out) {
  if (!out || Lcap == 0 || k2 > 13 || cg_k > 12) return -1;
  uint64_t sz[HPGQ_SEC_COUNT];
  sz[HPGQ_SEC_SCALARS] = HPGQ_S_COUNT;
  sz[HPGQ_SEC_LEN_HIST] = (uint64_t)Lcap + 1;
  sz[HPGQ_SEC_SUMQ_BY_LEN] = (uint64_t)Lcap + 1;
  sz[HPGQ_SEC_QUAL_HIST] = 256;
  sz[HPGQ_SEC_GC_HIST] = 104;
  sz[HPGQ_SEC_POS_ACCQ] = Lcap;
  sz[HPGQ_SEC_POS_A] = Lcap;
  sz[HPGQ_SEC_POS_C] = Lcap;
  sz[HPGQ_SEC_POS_G] = Lcap;
  sz[HPGQ_SEC_POS_T] = Lcap;
  sz[HPGQ_SEC_POS_N] = Lcap;
  sz[HPGQ_SEC_K5] = kmers ? 1024 : 0;
  sz[HPGQ_SEC_K5_POS] = kmers ? 1024ull * Lcap : 0;
  sz[HPGQ_SEC_K2] = (kmers && k2) ? (1ull << (2 * k2)) : 0;
  sz[HPGQ_SEC_CG_SEQ] = cg_k ? (1ull << (2 * cg_k)) : 0;
  sz[HPGQ_SEC_CG_Q] = cg_k ? (1ull << (2 * cg_k)) : 0;
  out->max_read_len = Lcap;
  out->kmers = kmers;
  out->k2 = k2;
  out->cg_k = cg_k;
  uint64_t o = 0;
  for (int s = 0; s < HPGQ_SEC_COUNT; ++s) {
    out->offset[s] = o;
    o += (sz[s] + 3) & ~3ull; /* sections start on 32-byte boundaries */
  }
  out->offset[HPGQ_SEC_COUNT] = o;
  return 0;
}

/* ------------------------------------------------------------------ per-read pieces */
typedef struct {
  uint32_t len;
  uint64_t sumq; /* sum of RAW quality bytes: stats_report.c:26,114 subtract phred afterwards */
  uint32_t nA, nC, nG, nT, nN;
} read_stats_t;

/* fastq_read_stats_t fields consumed at src/stats_fastq.c:289-304 (definition in the missing
 * bioinfo-libs): exact upper-case byte matches only (src/stats_fastq.c:363-374). */
static void read_stats(const uint8_t *s, const uint8_t *q, uint32_t L, read_stats_t *r) {
  memset(r, 0, sizeof *r);
  r->len = L;
  for (uint32_t j = 0; j < L; ++j) {
    r->sumq += q[j];
    switch (s[j]) {
      case 'A': r->nA++; break;
      case 'C': r->nC++; break;
      case 'G': r->nG++; break;
      case 'T': r->nT++; break;
      case 'N': r->nN++; break;
      default: break;
    }
  }
}

/* inclusive mean test on the Phred scale with exact integers: lo*w <= S - off*w <= hi*w */
static int mean_in_range(uint64_t S, uint32_t w, int32_t lo, int32_t hi, int32_t off) {
  int64_t v = (int64_t)S - (int64_t)off * (int64_t)w;
  return v >= (int64_t)lo * (int64_t)w && v <= (int64_t)hi * (int64_t)w;
}

/* fastq_filter for one read: 12-int tuple in call-site order (src/stats_fastq.c:209-214);
 * semantics SURVEY.md 8c(3) [UNPINNED]; hints old/README:125-130. Returns 1 = passed. */
static int filter_read(const uint8_t *s, const uint8_t *q, uint32_t L, const hpgq_filter_opts_t *f, int32_t off) {
  if ((int64_t)L < f->min_read_length || (int64_t)L > f->max_read_length) return 0;
  uint64_t S = 0;
  uint32_t ooq = 0, nN = 0;
  for (uint32_t j = 0; j < L; ++j) {
    S += q[j];
    int32_t ph = (int32_t)q[j] - off;
    if (ph < f->min_read_quality || ph > f->max_read_quality) ooq++;
    if (s[j] == 'N') nN++;
  }
  if (!mean_in_range(S, L, f->min_read_quality, f->max_read_quality, off)) return 0;
  if ((int64_t)ooq > f->max_out_of_quality) return 0;
  if (f->left_length > 0) {
    uint32_t w = (uint32_t)f->left_length < L ? (uint32_t)f->left_length : L;
    uint64_t Sw = 0;
    for (uint32_t j = 0; j < w; ++j) Sw += q[j];
    if (w && !mean_in_range(Sw, w, f->min_left_quality, f->max_left_quality, off)) return 0;
  }
  if (f->right_length > 0) {
    uint32_t w = (uint32_t)f->right_length < L ? (uint32_t)f->right_length : L;
    uint64_t Sw = 0;
    for (uint32_t j = L - w; j < L; ++j) Sw += q[j];
    if (w && !mean_in_range(Sw, w, f->min_right_quality, f->max_right_quality, off)) return 0;
  }
  if ((int64_t)nN > f->max_N) return 0;
  return 1;
}

/* fastq_edit for one read: 9-int tuple (src/edit_fastq.c:148-151); SURVEY.md 8c(4) [UNPINNED];
 * hints old/README:48-49,81-82,97-98.  Returns 1 if the read was trimmed on either side. */
static int edit_read(const uint8_t *q, uint32_t L, const hpgq_edit_opts_t *e, int32_t off, hpgq_trim_t *t) {
  uint32_t start = 0, len = L;
  int edited = 0;
  if (e->left_length > 0) {
    uint32_t w = (uint32_t)e->left_length < len ? (uint32_t)e->left_length : len;
    uint64_t Sw = 0;
    for (uint32_t j = 0; j < w; ++j) Sw += q[start + j];
    if (w && !mean_in_range(Sw, w, e->min_left_quality, e->max_left_quality, off)) {
      start += w;
      len -= w;
      edited = 1;
    }
  }
  if (e->right_length > 0) {
    uint32_t w = (uint32_t)e->right_length < len ? (uint32_t)e->right_length : len;
    uint64_t Sw = 0;
    for (uint32_t j = 0; j < w; ++j) Sw += q[start + len - w + j];
    if (w && !mean_in_range(Sw, w, e->min_right_quality, e->max_right_quality, off)) {
      len -= w;
      edited = 1;
    }
  }
  t->start = start;
  t->len = len;
  return edited;
}

static inline int base_code(uint8_t c) { /* A=0 C=1 G=2 T=3 (old/chaos_game.c:56-66), else -1 */
  switch (c) {
    case 'A': return 0;
    case 'C': return 1;
    case 'G': return 2;
    case 'T': return 3;
    default: return -1;
  }
}

/* k-mer counting, SURVEY.md 8c(5) [UNPINNED]: first base most significant; a window holding any
 * other byte is skipped; by_pos is indexed by the window's 0-based start (src/stats_fastq.c:385-410
 * merges counter and counter_by_pos[]). */
static void kmers_read(const uint8_t *s, uint32_t L, int k, uint64_t *tot, uint64_t *by_pos, uint32_t Lcap,
                       int atomic) {
  if (k <= 0 || L < (uint32_t)k) return;
  uint32_t code = 0, valid = 0;
  const uint32_t mask = (k >= 16) ? 0xFFFFFFFFu : ((1u << (2 * k)) - 1u);
  for (uint32_t j = 0; j < L; ++j) {
    int c = base_code(s[j]);
    if (c < 0) {
      valid = 0;
      code = 0;
      continue;
    }
    code = ((code << 2) | (uint32_t)c) & mask;
    if (++valid >= (uint32_t)k) {
      if (atomic) {
#pragma omp atomic
        tot[code]++;
      } else {
        tot[code]++;
      }
      if (by_pos) by_pos[(uint64_t)code * Lcap + (j + 1 - (uint32_t)k)]++;
    }
  }
}

/* ------------------------------------------------------------------ chaos game */
/* Line-for-line restatement of chaos_game_fill_tables (old/chaos_game.c:165-267) with the
 * tables widened to u64 (SURVEY.md 7 "hard parts": unsigned int cells overflow at C4) and
 * acc_word_quality kept 32-bit unsigned exactly as in the reference (wrap-around included).
 * f_x,f_y start at dim_n/2 on every call (old/chaos_game.c:180-181, never written back) and
 * carry across the reads of the batch.  status == NULL means "all reads". */
int oracle_chaos_fill(uint32_t n_reads, const uint8_t *seq, const uint8_t *quality, const uint32_t *data_indices,
                      const uint8_t *status, int k, int base_quality, uint64_t *table_seq, uint64_t *table_q,
                      uint64_t *fq_word_count) {
  if (k < 1 || k > 12) return -1;
  const int dim_n = 1 << k;
  const int word_size = k;
  const int word_quality_substract = base_quality * word_size;
  int nt_word_count = 0;
  unsigned int acc_word_quality = 0;
  double f_x = (double)(dim_n * 0.5), f_y = f_x;
  for (uint32_t i = 0; i < n_reads; i++) {
    if (status && !status[i]) continue;
    uint32_t read_position = data_indices[i];
    const uint32_t read_length = data_indices[i + 1] - data_indices[i];
    uint32_t quality_position = read_position;
    for (uint32_t j = 0; j < read_length; j++) {
      const char quality_character = (char)quality[quality_position++];
      switch (seq[read_position++]) {
        case 65: /* A */
          f_x = f_x + ((dim_n - f_x) * 0.5);
          f_y = f_y * 0.5;
          nt_word_count++;
          acc_word_quality += quality_character;
          break;
        case 67: /* C */
          f_x = f_x * 0.5;
          f_y = f_y * 0.5;
          nt_word_count++;
          acc_word_quality += quality_character;
          break;
        case 71: /* G */
          f_x = f_x * 0.5;
          f_y = f_y + ((dim_n - f_y) * 0.5);
          nt_word_count++;
          acc_word_quality += quality_character;
          break;
        case 84: /* T */
          f_x = f_x + ((dim_n - f_x) * 0.5);
          f_y = f_y + ((dim_n - f_y) * 0.5);
          nt_word_count++;
          acc_word_quality += quality_character;
          break;
        case 78: /* N */
          nt_word_count = 0;
          acc_word_quality = 0;
          break;
        default: break;
      }
      if (nt_word_count == word_size) {
        int co_x = (int)f_x, co_y = (int)f_y;
        if (co_x == dim_n) {
          co_x = dim_n - 1;
          f_x = f_x - 0.00001;
        }
        if (co_y == dim_n) {
          co_y = dim_n - 1;
          f_y = f_y - 0.00001;
        }
        table_seq[(size_t)co_x * dim_n + co_y]++;
        (*fq_word_count)++;
        nt_word_count--;
        table_q[(size_t)co_x * dim_n + co_y] += (unsigned int)(acc_word_quality - (unsigned int)word_quality_substract);
        acc_word_quality = acc_word_quality - (unsigned int)(char)quality[quality_position - word_size];
      }
    }
    nt_word_count = 0;
    acc_word_quality = 0;
  }
  return 0;
}

/* Integer-index form of the same tables: co_x = sum xbit(n_i) 2^(i-1), latest
 * base = MSB; equal to the double form unless an axis-run of >= 53-k bases occurred.  Used by tests
 * to show where (and only where) the two differ; long_runs counts positions where an A/T or G/T run
 * (N and other bytes do not break it, runs span reads) reaches 41. */
int oracle_chaos_fill_int(uint32_t n_reads, const uint8_t *seq, const uint8_t *quality, const uint32_t *idx, int k,
                          int base_quality, uint64_t *table_seq, uint64_t *table_q, uint64_t *words,
                          uint64_t *long_runs) {
  if (k < 1 || k > 12) return -1;
  const uint32_t dim = 1u << k;
  uint32_t ix = dim >> 1, iy = dim >> 1; /* f = dim/2: bit k-1 set, older bits zero */
  uint32_t run_x = 0, run_y = 0;
  for (uint32_t i = 0; i < n_reads; ++i) {
    int cnt = 0;
    unsigned int acc = 0;
    const uint32_t o = idx[i], L = idx[i + 1] - idx[i];
    for (uint32_t j = 0; j < L; ++j) {
      const uint8_t c = seq[o + j];
      int xb = -1, yb = 0;
      if (c == 'A') { xb = 1; yb = 0; }
      else if (c == 'C') { xb = 0; yb = 0; }
      else if (c == 'G') { xb = 0; yb = 1; }
      else if (c == 'T') { xb = 1; yb = 1; }
      if (xb >= 0) {
        ix = (ix >> 1) | ((uint32_t)xb << (k - 1));
        iy = (iy >> 1) | ((uint32_t)yb << (k - 1));
        run_x = xb ? run_x + 1 : 0;
        run_y = yb ? run_y + 1 : 0;
        if (run_x == 41 || run_y == 41) (*long_runs)++;
        cnt++;
        acc += (unsigned int)(char)quality[o + j];
      } else if (c == 'N') {
        cnt = 0;
        acc = 0;
      }
      if (cnt == k) {
        table_seq[(size_t)ix * dim + iy]++;
        (*words)++;
        cnt--;
        table_q[(size_t)ix * dim + iy] += (unsigned int)(acc - (unsigned int)(base_quality * k));
        acc -= (unsigned int)(char)quality[o + j + 1 - (uint32_t)k];
      }
    }
  }
  return 0;
}

/* ------------------------------------------------------------------ one batch through the workers */
/* fastq_stats_worker (src/stats_fastq.c:202-251), fastq_filter_worker (src/filter_fastq.c:134-155),
 * fastq_edit_worker (src/edit_fastq.c:142-178) and the consumer merge (src/stats_fastq.c:257-417)
 * for one SoA batch.  mask / trim may be NULL.  Batch-parallel like the reference's worker pool:
 * thread-private dense counters merged at the end. */
int oracle_process_batch(const hpgq_params_t *p, const hpgq_layout_t *lay, uint32_t n, const uint8_t *seq,
                         const uint8_t *qual, const uint32_t *idx, uint64_t *counters, uint8_t *mask,
                         hpgq_trim_t *trim, int num_threads) {
  if (!p || !lay || !counters) return -1;
  const uint32_t Lcap = lay->max_read_len;
  const int do_stats = (p->work & HPGQ_WORK_STATS) != 0, do_filter = (p->work & HPGQ_WORK_FILTER) != 0;
  const int do_edit = (p->work & HPGQ_WORK_EDIT) != 0;
  const int do_kmers = (p->work & HPGQ_WORK_KMERS) != 0 && lay->kmers;
  const int do_cg = (p->work & HPGQ_WORK_CHAOS) != 0 && lay->cg_k;
  const uint64_t priv_words = lay->offset[HPGQ_SEC_K2]; /* everything before the big tables */
  int err = 0;
  if (num_threads <= 0) {
#ifdef _OPENMP
    num_threads = omp_get_max_threads();
#else
    num_threads = 1;
#endif
  }
  uint64_t *priv = (uint64_t *)calloc((size_t)num_threads * priv_words, sizeof(uint64_t));
  if (!priv) return -3;
  uint64_t *k2tab = counters + lay->offset[HPGQ_SEC_K2];

#pragma omp parallel num_threads(num_threads)
  {
#ifdef _OPENMP
    const int tid = omp_get_thread_num();
#else
    const int tid = 0;
#endif
    uint64_t *c = priv + (size_t)tid * priv_words;
    uint64_t *S = c + lay->offset[HPGQ_SEC_SCALARS];
#pragma omp for schedule(dynamic, 10000) /* reference batch size: src/stats_options.c:22 */
    for (int64_t ii = 0; ii < (int64_t)n; ++ii) {
      const uint32_t i = (uint32_t)ii;
      const uint32_t o = idx[i];
      uint32_t L = idx[i + 1] - idx[i];
      const uint8_t *s = seq + o, *q = qual + o;
      hpgq_trim_t t = {0, L};
      /* fastq_edit first: src/edit_fastq.c:154, the filter then sees the trimmed read (:159-167) */
      if (do_edit) {
        if (edit_read(q, L, &p->edit, p->phred, &t)) S[HPGQ_S_NUM_EDITED]++;
        s += t.start;
        q += t.start;
        L = t.len;
      }
      if (trim) trim[i] = t;
      int pass = 1;
      if (do_filter) {
        pass = filter_read(s, q, L, &p->filter, p->phred);
        if (pass) S[HPGQ_S_NUM_PASSED]++; /* src/stats_fastq.c:268-272 */
        else S[HPGQ_S_NUM_FAILED]++;
      }
      if (mask) mask[i] = (uint8_t)pass;
      /* statistics only for reads that passed: src/stats_fastq.c:226-232 */
      if (do_stats && pass) {
        if (L > Lcap) {
#pragma omp atomic write
          err = -1;
          continue;
        }
        read_stats_t r;
        read_stats(s, q, L, &r);
        S[HPGQ_S_NUM_READS]++;                    /* :286 */
        S[HPGQ_S_ACC_LENGTH] += L;                /* :290 */
        c[lay->offset[HPGQ_SEC_LEN_HIST] + L]++;  /* :307-314 */
        c[lay->offset[HPGQ_SEC_SUMQ_BY_LEN] + L] += r.sumq; /* exact stand-in for float acc_quality :297 */
        if (L > 0) {
          /* round((float)S/L), :317 ; integer form identical for L <= 131071 */
          c[lay->offset[HPGQ_SEC_QUAL_HIST] + (2 * r.sumq + L) / (2ull * L)]++;
          c[lay->offset[HPGQ_SEC_GC_HIST] + (100ull * (r.nG + r.nC)) / L]++; /* :327 */
        }
        uint64_t *accq = c + lay->offset[HPGQ_SEC_POS_ACCQ];
        uint64_t *pa = c + lay->offset[HPGQ_SEC_POS_A], *pc = c + lay->offset[HPGQ_SEC_POS_C];
        uint64_t *pg = c + lay->offset[HPGQ_SEC_POS_G], *pt = c + lay->offset[HPGQ_SEC_POS_T];
        uint64_t *pn = c + lay->offset[HPGQ_SEC_POS_N];
        for (uint32_t j = 0; j < L; ++j) { /* :338-382 */
          accq[j] += q[j];
          switch (s[j]) {
            case 'A': pa[j]++; break;
            case 'T': pt[j]++; break;
            case 'C': pc[j]++; break;
            case 'G': pg[j]++; break;
            case 'N': pn[j]++; break;
            default: break;
          }
        }
      }
      if (do_kmers && pass) { /* k-mers are part of fastq_reads_stats: passed reads only */
        if (L > Lcap) {
#pragma omp atomic write
          err = -1;
          continue;
        }
        kmers_read(s, L, 5, c + lay->offset[HPGQ_SEC_K5], c + lay->offset[HPGQ_SEC_K5_POS], Lcap, 0);
        if (lay->k2) kmers_read(s, L, (int)lay->k2, k2tab, NULL, 0, 1);
      }
    }
  }
  for (int t = 0; t < num_threads; ++t) {
    const uint64_t *c = priv + (size_t)t * priv_words;
    for (uint64_t w = 0; w < priv_words; ++w) counters[w] += c[w];
  }
  free(priv);
  if (do_cg) { /* legacy path: serial per batch, all reads (old/main_hpg_fastq_old.c:726) */
    uint64_t words = 0;
    int rc = oracle_chaos_fill(n, seq, qual, idx, NULL, (int)lay->cg_k, p->cg_base_quality,
                               counters + lay->offset[HPGQ_SEC_CG_SEQ], counters + lay->offset[HPGQ_SEC_CG_Q], &words);
    if (rc) return rc;
    counters[HPGQ_S_CG_WORDS] += words;
  }
  counters[HPGQ_S_BATCHES]++;
  return err;
}

/* derived scalars: base totals (src/stats_fastq.c:300-304 summed over reads == column sums of the
 * per-position tables) and min/max length (:291-294; min starts at 100000, :24) */
int oracle_finalize(const hpgq_layout_t *lay, uint64_t *c) {
  const uint32_t Lcap = lay->max_read_len;
  const int secs[5] = {HPGQ_SEC_POS_A, HPGQ_SEC_POS_C, HPGQ_SEC_POS_G, HPGQ_SEC_POS_T, HPGQ_SEC_POS_N};
  const int dst[5] = {HPGQ_S_NUM_A, HPGQ_S_NUM_C, HPGQ_S_NUM_G, HPGQ_S_NUM_T, HPGQ_S_NUM_N};
  for (int k = 0; k < 5; ++k) {
    uint64_t t = 0;
    for (uint32_t j = 0; j < Lcap; ++j) t += c[lay->offset[secs[k]] + j];
    c[dst[k]] = t;
  }
  uint64_t mn = 100000, mx = 0;
  for (uint32_t L = 0; L <= Lcap; ++L)
    if (c[lay->offset[HPGQ_SEC_LEN_HIST] + L]) {
      if (L < mn) mn = L;
      if (L > mx) mx = L;
    }
  c[HPGQ_S_MIN_LENGTH] = mn;
  c[HPGQ_S_MAX_LENGTH] = mx;
  return 0;
}

/* Order-preserving partition of the (trimmed) reads: passed reads first, failed after, each in
 * input order -- what fastq_fwrite(passed) / fastq_fwrite(failed) emit (src/filter_fastq.c:165-169,
 * src/edit_fastq.c:190-198) with --num-threads 1.  out_idx has n+1 entries. */
int oracle_compact(uint32_t n, const uint8_t *seq, const uint8_t *qual, const uint32_t *idx, const uint8_t *mask,
                   const hpgq_trim_t *trim, uint8_t *out_seq, uint8_t *out_qual, uint32_t *out_idx,
                   uint32_t *out_src, uint32_t *n_passed, uint64_t *passed_bytes, uint64_t *failed_bytes) {
  uint32_t k = 0;
  uint64_t o = 0;
  out_idx[0] = 0;
  for (int want = 1; want >= 0; --want) {
    for (uint32_t i = 0; i < n; ++i) {
      const int pass = mask ? (mask[i] != 0) : 1;
      if (pass != want) continue;
      const uint32_t st = trim ? trim[i].start : 0, ln = trim ? trim[i].len : idx[i + 1] - idx[i];
      memcpy(out_seq + o, seq + idx[i] + st, ln);
      memcpy(out_qual + o, qual + idx[i] + st, ln);
      o += ln;
      out_src[k] = i;
      out_idx[++k] = (uint32_t)o;
    }
    if (want == 1) {
      *n_passed = k;
      *passed_bytes = o;
    }
  }
  *failed_bytes = o - *passed_bytes;
  return 0;
}

/* "Reference-faithful merge" timing variant: the single consumer thread doing one
 * open-addressing hash insertion per counter per base, as src/stats_fastq.c:338-382 does with khash
 * (int keys, identity hash).  Produces the same per-position tables as the dense path. */
typedef struct {
  uint32_t cap, used;
  int32_t *keys;
  uint64_t *vals;
} ihash_t;
static void ih_init(ihash_t *h) {
  h->cap = 16;
  h->used = 0;
  h->keys = (int32_t *)malloc(sizeof(int32_t) * h->cap);
  h->vals = (uint64_t *)calloc(h->cap, sizeof(uint64_t));
  for (uint32_t i = 0; i < h->cap; ++i) h->keys[i] = -1;
}
static uint64_t *ih_put(ihash_t *h, int32_t key) {
  if ((h->used + 1) * 4 > h->cap * 3) {
    ihash_t n;
    n.cap = h->cap * 2;
    n.used = 0;
    n.keys = (int32_t *)malloc(sizeof(int32_t) * n.cap);
    n.vals = (uint64_t *)calloc(n.cap, sizeof(uint64_t));
    for (uint32_t i = 0; i < n.cap; ++i) n.keys[i] = -1;
    for (uint32_t i = 0; i < h->cap; ++i)
      if (h->keys[i] >= 0) {
        uint32_t s = (uint32_t)h->keys[i] & (n.cap - 1);
        while (n.keys[s] >= 0) s = (s + 1) & (n.cap - 1);
        n.keys[s] = h->keys[i];
        n.vals[s] = h->vals[i];
        n.used++;
      }
    free(h->keys);
    free(h->vals);
    *h = n;
  }
  uint32_t s = (uint32_t)key & (h->cap - 1);
  while (h->keys[s] >= 0 && h->keys[s] != key) s = (s + 1) & (h->cap - 1);
  if (h->keys[s] < 0) {
    h->keys[s] = key;
    h->used++;
  }
  return &h->vals[s];
}
int oracle_merge_hash_faithful(const hpgq_layout_t *lay, uint32_t n, const uint8_t *seq, const uint8_t *qual,
                               const uint32_t *idx, uint64_t *counters) {
  ihash_t cnt, acc, hA, hC, hG, hT, hN;
  ih_init(&cnt); ih_init(&acc); ih_init(&hA); ih_init(&hC); ih_init(&hG); ih_init(&hT); ih_init(&hN);
  for (uint32_t i = 0; i < n; ++i) {
    const uint32_t o = idx[i], L = idx[i + 1] - idx[i];
    for (uint32_t j = 0; j < L; ++j) {
      (*ih_put(&cnt, (int32_t)j))++;
      *ih_put(&acc, (int32_t)j) += qual[o + j];
      ihash_t *h = NULL;
      switch (seq[o + j]) {
        case 'A': h = &hA; break;
        case 'T': h = &hT; break;
        case 'C': h = &hC; break;
        case 'G': h = &hG; break;
        case 'N': h = &hN; break;
        default: break;
      }
      if (h) (*ih_put(h, (int32_t)j))++;
    }
  }
  ihash_t *hs[6] = {&acc, &hA, &hC, &hG, &hT, &hN};
  const int secs[6] = {HPGQ_SEC_POS_ACCQ, HPGQ_SEC_POS_A, HPGQ_SEC_POS_C, HPGQ_SEC_POS_G, HPGQ_SEC_POS_T, HPGQ_SEC_POS_N};
  for (int k = 0; k < 6; ++k)
    for (uint32_t s = 0; s < hs[k]->cap; ++s)
      if (hs[k]->keys[s] >= 0 && (uint32_t)hs[k]->keys[s] < lay->max_read_len)
        counters[lay->offset[secs[k]] + (uint32_t)hs[k]->keys[s]] += hs[k]->vals[s];
  ihash_t *all[7] = {&cnt, &acc, &hA, &hC, &hG, &hT, &hN};
  for (int k = 0; k < 7; ++k) {
    free(all[k]->keys);
    free(all[k]->vals);
  }
  return 0;
}

int oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
