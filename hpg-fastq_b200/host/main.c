/*
 * main.c -- `hpg-fastq stats|filter|edit`: the reference's command line (src/hpg-fastq.c:38-127) over the
 * B200 C ABI.  The workflow of each command keeps the reference's shape
 *   producer  fastq_*_producer  (src/stats_fastq.c:174-196)  -> hq_reader_peek_text: a span of whole records of the
 *                                                                mapped file (records are cut on the GPU), or
 *                                                                hq_reader_next: FASTQ text -> pinned SoA batch
 *                                                                (--host-parse, and input the device parser
 *                                                                hands back: blank lines, malformed records)
 *   worker    fastq_*_worker    (src/stats_fastq.c:202-251, src/filter_fastq.c:134-155, src/edit_fastq.c:142-178)
 *                                                             -> hpgq_batch_submit_text_fd + _text_launch /
 *                                                                hpgq_batch_submit (H2D + kernels, async)
 *   consumer  fastq_*_consumer  (src/stats_fastq.c:257-417, src/filter_fastq.c:161-176, src/edit_fastq.c:184-208)
 *                                                             -> hpgq_batch_wait_text_fd (records composed on
 *                                                                the GPU, appended to the files) / hpgq_batch_wait
 *                                                                + FASTQ writer; the statistics merge stays on
 *                                                                the device
 * with batch b on GPU b mod N, stream slot (b / N) mod 2, so that parsing batch b+1 overlaps the copies and
 * kernels of batch b, and outputs are written in ascending batch order (= input order).
 * There is no CPU fallback: without a B200 the program stops with the library's error.
 */
#include "hpgq_host.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define HQ_SLOTS 2
#define HQ_MAX_GPUS 16

static FILE *g_log;

static void logf_(const char *level, const char *msg) {
  if (g_log) {
    fprintf(g_log, "%s: %s\n", level, msg);
    fflush(g_log);
  }
}

static int die_hpgq(const char *what, int rc) {
  char buf[768];
  snprintf(buf, sizeof buf, "%s: %s [%d] -- %s", what, hpgq_strerror(rc), rc, hpgq_last_error());
  fprintf(stderr, "\nFATAL: %s\n", buf);
  logf_("FATAL", buf);
  return 1;
}

static void usage(const char *exec_name) {
  /* usage(), src/hpg-fastq.c:17-31 */
  printf("Program: %s (High-performance tools for handling FastQ files)\n", exec_name);
  printf("Version: 1.0.0\n");
  printf("\n");
  printf("Usage: %s <command> [options]\n", exec_name);
  printf("\n");
  printf("Command: stats\t\tstatistics summary\n");
  printf("         filter\t\tfilter a FastQ file by using advanced criteria\n");
  printf("         edit\t\tedit a FastQ file according the specified options\n");
  printf("\n");
  printf("For more information about a certain command, type %s <command> --help\n", exec_name);
}

typedef struct run {
  hq_options_t *o;
  hq_reader_t *reader;
  int n_gpus;
  hpgq_ctx_t *ctx[HQ_MAX_GPUS];
  hq_host_batch_t hb[HQ_MAX_GPUS][HQ_SLOTS];
  int busy[HQ_MAX_GPUS][HQ_SLOTS];
  int is_text[HQ_MAX_GPUS][HQ_SLOTS]; /* the slot's batch went through the device parser */
  hq_writer_t *w_pass, *w_fail;
  uint64_t num_passed, num_failed, num_edited, num_reads;
  int device_parse;      /* records are cut on the GPU until a chunk is handed back */
  uint64_t text_budget;  /* bytes of FASTQ text per device-parsed batch */
  uint64_t n_text_batches, n_host_batches;
  /* the text chunk whose parse is in flight: its QC kernels are launched one chunk later (resolve_pending), so that
   * the next chunk is staged and copied while the device parses this one */
  int pend, pend_g, pend_s;
  uint64_t pend_off, pend_len;
} run_t;

static double now_s(void);
static void trace_(const char *what, double since);

/* the pinned SoA arrays are only needed by the host reader: with the device parser they are allocated when (if)
 * the input is handed back (pinning hundreds of MB costs more than parsing them) */
static int alloc_host_soa(hq_host_batch_t *b) {
  int rc;
  if (b->seq) return 0;
  if ((rc = hpgq_host_alloc_pinned(b->cap_bytes + 64, (void **)&b->seq))) return rc;
  if ((rc = hpgq_host_alloc_pinned(b->cap_bytes + 64, (void **)&b->qual))) return rc;
  return hpgq_host_alloc_pinned(((size_t)b->cap_reads + 1) * 4 + 64, (void **)&b->idx);
}

static int alloc_host_batch(hq_host_batch_t *b, uint32_t cap_reads, uint64_t cap_bytes, int with_soa) {
  memset(b, 0, sizeof *b);
  b->cap_reads = cap_reads;
  b->cap_bytes = cap_bytes;
  b->hdr_off = (uint64_t *)malloc((size_t)cap_reads * sizeof(uint64_t));
  b->hdr_len = (uint32_t *)malloc((size_t)cap_reads * sizeof(uint32_t));
  if (!b->hdr_off || !b->hdr_len) return HPGQ_ERR_NOMEM;
  return with_soa ? alloc_host_soa(b) : 0;
}

static void free_host_batch(hq_host_batch_t *b) {
  if (b->seq) hpgq_host_free_pinned(b->seq);
  if (b->qual) hpgq_host_free_pinned(b->qual);
  if (b->idx) hpgq_host_free_pinned(b->idx);
  free(b->hdr_off);
  free(b->hdr_len);
  memset(b, 0, sizeof *b);
}

/* setup of one GPU: context, parameters, host batch records, text buffers */
typedef struct gpu_setup {
  run_t *r;
  int g;
  const hpgq_limits_t *lim;
  const hpgq_params_t *par;
  uint64_t cap_bytes;
  int rc;
  const char *what;
  int no_thread;
  pthread_t th;
  char msg[256];
} gpu_setup_t;

static void *gpu_setup_main(void *arg) {
  gpu_setup_t *a = (gpu_setup_t *)arg;
  run_t *r = a->r;
  const int g = a->g;
  a->what = "hpgq_ctx_create";
  a->rc = hpgq_ctx_create(g, a->lim, &r->ctx[g]);
  if (!a->rc) {
    a->what = "hpgq_set_params";
    a->rc = hpgq_set_params(r->ctx[g], a->par);
  }
  for (int s = 0; s < HQ_SLOTS && !a->rc; s++) {
    a->what = "pinned batch allocation";
    a->rc = alloc_host_batch(&r->hb[g][s], a->lim->max_reads_per_batch, a->cap_bytes, r->o->host_parse);
  }
  if (!a->rc && r->device_parse) {
    a->what = "hpgq_ctx_reserve_text";
    a->rc = hpgq_ctx_reserve_text(r->ctx[g], 2 * a->cap_bytes, r->o->command != HQ_CMD_STATS);
  }
  if (a->rc) snprintf(a->msg, sizeof a->msg, "%s", hpgq_last_error()); /* the message is per thread */
  return NULL;
}

/* the NCCL communicators of the end-of-run merge are created by a thread of their own while the batches run */
static void *comm_prepare_main(void *arg) {
  const int n = *(const int *)arg;
  int devs[HQ_MAX_GPUS];
  for (int g = 0; g < n; g++) devs[g] = g;
  return (void *)(intptr_t)hpgq_comm_prepare(devs, n);
}

/* consumer stage for one retired slot */
static int consume(run_t *r, int g, int s) {
  hpgq_batch_result_t res;
  const hq_options_t *o = r->o;
  int rc;
  if (r->is_text[g][s] && o->command != HQ_CMD_STATS) {
    /* device-parsed batch: the output records are composed on the GPU and appended to the files by the library,
     * piece by piece while the rest is still on its way back (hpgq_batch_wait_text_fd) */
    uint64_t pbytes = 0, fbytes = 0;
    const int fdp = r->w_pass ? hq_writer_fd(r->w_pass) : -1, fdf = r->w_fail ? hq_writer_fd(r->w_fail) : -1;
    if ((r->w_pass && fdp < 0) || (r->w_fail && fdf < 0)) {
      fprintf(stderr, "\nFATAL: write error on the output FASTQ files\n");
      return 1;
    }
    if ((rc = hpgq_batch_wait_text_fd(r->ctx[g], s, &res, fdp, fdf, &pbytes, &fbytes))) return die_hpgq("hpgq_batch_wait_text_fd", rc);
    r->busy[g][s] = 0;
    r->num_reads += res.n_reads;
    r->num_edited += res.n_edited;
    if (o->command == HQ_CMD_FILTER || o->filter_on) {
      r->num_passed += res.n_passed;
      r->num_failed += res.n_failed;
    }
    return 0;
  }
  /* host-parsed batch: the records are formatted here */
  rc = hpgq_batch_wait(r->ctx[g], s, &res);
  if (rc) return die_hpgq("hpgq_batch_wait", rc);
  r->busy[g][s] = 0;
  const hq_host_batch_t *b = &r->hb[g][s];
  const uint8_t *fd = hq_reader_data(r->reader);
  r->num_reads += res.n_reads;
  if (o->command == HQ_CMD_STATS) return 0; /* counters stay on the device */
  r->num_edited += res.n_edited;
  const int filtering = o->command == HQ_CMD_FILTER || o->filter_on;
  if (filtering) {
    /* passed reads [0, n_passed) then failed reads, each in input order (src/filter_fastq.c:165-169) */
    r->num_passed += res.n_passed;
    r->num_failed += res.n_failed;
    if (hq_writer_put(r->w_pass, fd, b, res.out_seq, res.out_qual, res.out_idx, res.out_src, 0, res.n_passed) ||
        hq_writer_put(r->w_fail, fd, b, res.out_seq, res.out_qual, res.out_idx, res.out_src, res.n_passed,
                      res.n_passed + res.n_failed)) {
      fprintf(stderr, "\nFATAL: write error on the output FASTQ files\n");
      return 1;
    }
  } else {
    /* edit without filter options: every (trimmed) read goes to edit.fq (src/edit_fastq.c:196-199) */
    if (hq_writer_put(r->w_pass, fd, b, res.out_seq, res.out_qual, res.out_idx, res.out_src, 0, res.n_reads)) {
      fprintf(stderr, "\nFATAL: write error on the output FASTQ file\n");
      return 1;
    }
  }
  return 0;
}

/* Launches the QC kernels of the pending text chunk.  0 = done; 1 = fatal; 2 = the device parser handed the chunk
 * back (the reader stands at its start again, and either the chunk budget was halved or the host reader takes over). */
static int resolve_pending(run_t *r) {
  if (!r->pend) return 0;
  r->pend = 0;
  const int g = r->pend_g, s = r->pend_s;
  hpgq_text_info_t info;
  int rc = hpgq_batch_text_launch(r->ctx[g], s, &info);
  if (rc) return die_hpgq("hpgq_batch_text_launch", rc);
  if (info.status == 0) {
    r->hb[g][s].n_reads = info.n_reads;
    r->hb[g][s].max_len = info.max_len;
    hq_reader_advance(r->reader, 0, info.n_reads);
    r->busy[g][s] = 1;
    r->is_text[g][s] = 1;
    r->n_text_batches++;
    return 0;
  }
  hq_reader_rewind(r->reader, r->pend_off);
  if (info.status == HPGQ_TEXT_TOO_MANY && r->text_budget > (1u << 16)) {
    r->text_budget /= 2; /* shorter reads than the sizing assumed: same chunk again, smaller */
    return 2;
  }
  /* blank lines, a malformed record: the host reader takes over (and owns the messages) */
  r->device_parse = 0;
  logf_("INFO", "device FASTQ parser handed the input back to the host reader");
  return 2;
}

static int run_pipeline(run_t *r) {
  const int N = r->n_gpus;
  uint64_t b = 0;
  for (;; b++) {
    const int g = (int)(b % (uint64_t)N), s = (int)((b / (uint64_t)N) % HQ_SLOTS);
    if (r->pend && r->pend_g == g && r->pend_s == s) { /* a single slot in all: nothing to overlap */
      const int st = resolve_pending(r);
      if (st == 1) return 1;
      if (st == 2) {
        b--;
        continue;
      }
    }
    double tt = now_s();
    if (r->busy[g][s] && consume(r, g, s)) return 1;
    trace_("  consume", tt);
    hq_host_batch_t *hb = &r->hb[g][s];
    if (r->device_parse) {
      uint64_t off = 0, len = 0;
      const int pk = hq_reader_peek_text(r->reader, r->text_budget, &off, &len);
      if (pk == 0) {
        const int st = resolve_pending(r);
        if (st == 1) return 1;
        if (st == 2) {
          b--;
          continue;
        }
        break;
      }
      if (pk > 0) {
        /* chunk b is read (pread threads, pinned staging), copied and parsed while chunk b - 1 is still being parsed;
         * only then are the kernels of chunk b - 1 launched */
        int rc;
        tt = now_s();
        rc = hpgq_batch_submit_text_fd(r->ctx[g], hq_reader_fd(r->reader), off, len, s);
        if (rc) return die_hpgq("hpgq_batch_submit_text_fd", rc);
        trace_("  submit_text_fd", tt);
        hq_reader_advance(r->reader, len, 0);
        tt = now_s();
        const int st = resolve_pending(r);
        trace_("  text_launch (previous)", tt);
        if (st == 1) return 1;
        if (st == 2) {
          /* chunk b - 1 was handed back: chunk b (behind it) is dropped unseen -- its kernels were never launched */
          if ((rc = hpgq_batch_text_cancel(r->ctx[g], s))) return die_hpgq("hpgq_batch_text_cancel", rc);
          b -= 2; /* the handed-back chunk takes its slot again */
          continue;
        }
        r->pend = 1;
        r->pend_g = g;
        r->pend_s = s;
        r->pend_off = off;
        r->pend_len = len;
        continue;
      }
      /* no record boundary found inside the budget: the host reader takes over (and owns the messages) */
      {
        const int st = resolve_pending(r);
        if (st == 1) return 1;
        if (st == 2) {
          b--;
          continue;
        }
      }
      r->device_parse = 0;
      logf_("INFO", "device FASTQ parser handed the input back to the host reader");
    }
    int arc = alloc_host_soa(hb);
    if (arc) return die_hpgq("pinned batch allocation", arc);
    int got = hq_reader_next(r->reader, hb);
    if (got < 0) {
      fprintf(stderr, "\nFATAL: %s: %s\n", r->o->in_filename, hq_reader_error(r->reader));
      logf_("FATAL", hq_reader_error(r->reader));
      return 1;
    }
    if (!got) break;
    hpgq_batch_t hbatch = {hb->n_reads, 0, hb->seq, hb->qual, hb->idx};
    int rc = hpgq_batch_submit(r->ctx[g], &hbatch, s);
    if (rc) return die_hpgq("hpgq_batch_submit", rc);
    r->busy[g][s] = 1;
    r->is_text[g][s] = 0;
    r->n_host_batches++;
  }
  /* drain in batch order: the oldest outstanding batch is b - N*HQ_SLOTS */
  const uint64_t first = b > (uint64_t)N * HQ_SLOTS ? b - (uint64_t)N * HQ_SLOTS : 0;
  for (uint64_t k = first; k < b; k++) {
    const int g = (int)(k % (uint64_t)N), s = (int)((k / (uint64_t)N) % HQ_SLOTS);
    if (r->busy[g][s] && consume(r, g, s)) return 1;
  }
  return 0;
}

static double now_s(void);
static void trace_(const char *what, double since) {
  static int on = -1;
  if (on < 0) on = getenv("HPGQ_TRACE") != NULL; /* developer knob: where the wall time of a run goes */
  if (on) {
    struct timespec ts;
    clock_gettime(CLOCK_REALTIME, &ts);
    fprintf(stderr, "[trace] %-28s %8.3f ms   (epoch %.3f)\n", what, (now_s() - since) * 1e3, (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec);
  }
}

static double now_s(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static int run_command(hq_options_t *o) {
  run_t *r = (run_t *)calloc(1, sizeof *r);
  if (!r) return 1;
  r->o = o;
  int status = 1, rc;
  char pass_name[4096] = "", fail_name[4096] = "";
  pthread_t comm_th;
  int comm_bg = 0;
  const double t0 = now_s();

  /* "update opts": NO_VALUE -> MIN/MAX (stats only when a filter option was given, src/stats_fastq.c:430-444) */
  if (o->command != HQ_CMD_STATS || o->filter_on) hq_options_fill_unset(o);

  if (hq_reader_open(o->in_filename, &r->reader)) {
    fprintf(stderr, "\nFATAL: cannot open %s\n", o->in_filename);
    goto out;
  }
  trace_("reader open", t0);
  /* Lcap (entries of the per-position tables) and the batch byte capacity come from the longest line */
  uint32_t max_line = (uint32_t)o->max_read_len;
  if (max_line == 0) {
    uint64_t n_lines = 0;
    /* stats sizes the per-position tables from the longest read of the whole file; filter / edit only size their
     * batches (cut by bytes anyway), so the first 64 MB are sample enough */
    const uint64_t scan_bytes = o->command == HQ_CMD_STATS ? hq_reader_size(r->reader) : (64ull << 20);
    if (hq_reader_scan_prefix(r->reader, scan_bytes, &max_line, &n_lines)) {
      fprintf(stderr, "\nFATAL: %s holds a line longer than 4 GiB\n", o->in_filename);
      goto out;
    }
  }
  trace_("longest-line scan (cum.)", t0);
  if (max_line == 0) max_line = 1;
  uint64_t cap_bytes = (uint64_t)o->batch_size * max_line;
  if (cap_bytes > (1ull << 30)) cap_bytes = 1ull << 30; /* batches are cut by bytes first for long reads */
  if (cap_bytes < max_line) cap_bytes = max_line;
  if (cap_bytes < 4096) cap_bytes = 4096;

  int ndev = 0;
  rc = hpgq_device_count(&ndev);
  if (rc || ndev < 1) {
    die_hpgq("hpgq_device_count", rc ? rc : HPGQ_ERR_NO_DEVICE);
    goto out;
  }
  trace_("device count (cum.)", t0);
  r->n_gpus = o->gpus;
  if (r->n_gpus > ndev) {
    fprintf(stderr, "\nFATAL: --gpus %d but only %d CUDA device(s) are visible\n", o->gpus, ndev);
    goto out;
  }
  comm_bg = o->command == HQ_CMD_STATS && r->n_gpus > 1 && pthread_create(&comm_th, NULL, comm_prepare_main, &r->n_gpus) == 0;
  hpgq_params_t par;
  hq_options_to_params(o, &par);
  hpgq_limits_t lim;
  memset(&lim, 0, sizeof lim);
  lim.max_reads_per_batch = (uint32_t)o->batch_size;
  lim.max_bytes_per_batch = cap_bytes;
  lim.max_read_len = (max_line + 63u) & ~63u;
  lim.n_slots = HQ_SLOTS;
  lim.host_staging = 1;
  lim.kmers = (uint32_t)(o->command == HQ_CMD_STATS && o->kmers_on);
  lim.k2 = lim.kmers ? (uint32_t)o->k2 : 0;
  lim.cg_k = o->command == HQ_CMD_STATS ? (uint32_t)o->cg_k : 0;
  /* one thread per GPU: creating a CUDA context takes 0.2-0.35 s and pinning the text rings 0.05 s each */
  r->device_parse = !o->host_parse;
  gpu_setup_t gs[HQ_MAX_GPUS];
  for (int g = 0; g < r->n_gpus; g++) {
    memset(&gs[g], 0, sizeof gs[g]);
    gs[g].r = r;
    gs[g].g = g;
    gs[g].lim = &lim;
    gs[g].par = &par;
    gs[g].cap_bytes = cap_bytes;
    if (g == 0 || pthread_create(&gs[g].th, NULL, gpu_setup_main, &gs[g])) {
      gs[g].no_thread = 1;
      if (g) gpu_setup_main(&gs[g]);
    }
  }
  gpu_setup_main(&gs[0]);
  for (int g = 1; g < r->n_gpus; g++)
    if (!gs[g].no_thread) pthread_join(gs[g].th, NULL);
  for (int g = 0; g < r->n_gpus; g++)
    if (gs[g].rc) {
      fprintf(stderr, "\nFATAL: %s: %s (%s)\n", gs[g].what, hpgq_strerror(gs[g].rc), gs[g].msg);
      logf_("FATAL", gs[g].what);
      goto out;
    }
  trace_("text buffers (cum.)", t0);
  r->text_budget = 2 * cap_bytes; /* bases are less than half of the text, so a chunk never exceeds cap_bytes bases */
  if (o->paired && o->command != HQ_CMD_STATS) {
    /* paired end: one .valid (and .invalid) file per mate file (old/README:78,115-118) */
    const char *base = strrchr(o->in_filename, '/');
    base = base ? base + 1 : o->in_filename;
    snprintf(pass_name, sizeof pass_name, "%s/%s.valid", o->out_dirname, base);
    if (o->command == HQ_CMD_FILTER || o->filter_on) snprintf(fail_name, sizeof fail_name, "%s/%s.invalid", o->out_dirname, base);
  } else if (o->command == HQ_CMD_FILTER) {
    snprintf(pass_name, sizeof pass_name, "%s/passed.fq", o->out_dirname); /* src/filter_fastq.c:185 */
    snprintf(fail_name, sizeof fail_name, "%s/failed.fq", o->out_dirname); /* src/filter_fastq.c:189 */
  } else if (o->command == HQ_CMD_EDIT) {
    snprintf(pass_name, sizeof pass_name, "%s/edit.fq", o->out_dirname);   /* src/edit_fastq.c:217 */
    if (o->filter_on) snprintf(fail_name, sizeof fail_name, "%s/failed.fq", o->out_dirname); /* :223 */
  }
  if (pass_name[0] && hq_writer_open(pass_name, &r->w_pass)) {
    fprintf(stderr, "\nFATAL: cannot create %s\n", pass_name);
    goto out;
  }
  if (fail_name[0] && hq_writer_open(fail_name, &r->w_fail)) {
    fprintf(stderr, "\nFATAL: cannot create %s\n", fail_name);
    goto out;
  }

  const double t1 = now_s();
  const int prc = run_pipeline(r);
  if (comm_bg) {
    pthread_join(comm_th, NULL); /* a failure shows up again (with its message) in hpgq_counters_reduce */
    comm_bg = 0;
  }
  if (prc) goto out;
  const double t2 = now_s();

  if (r->w_pass && hq_writer_close(r->w_pass)) {
    r->w_pass = NULL;
    fprintf(stderr, "\nFATAL: write error on %s\n", pass_name);
    goto out;
  }
  r->w_pass = NULL;
  if (r->w_fail && hq_writer_close(r->w_fail)) {
    r->w_fail = NULL;
    fprintf(stderr, "\nFATAL: write error on %s\n", fail_name);
    goto out;
  }
  r->w_fail = NULL;

  if (o->command == HQ_CMD_STATS) {
    /* end of run: one sum-allreduce of the counter blocks, one download, then stats_report() */
    if ((rc = hpgq_counters_reduce(r->ctx, r->n_gpus))) {
      die_hpgq("hpgq_counters_reduce", rc);
      goto out;
    }
    hpgq_layout_t lay;
    hpgq_get_layout(r->ctx[0], &lay);
    const size_t words = lay.offset[HPGQ_SEC_COUNT];
    uint64_t *blk = (uint64_t *)malloc(words * sizeof(uint64_t));
    if (!blk) goto out;
    if ((rc = hpgq_counters_download(r->ctx[0], blk, words))) {
      free(blk);
      die_hpgq("hpgq_counters_download", rc);
      goto out;
    }
    if (hq_stats_report(&lay, blk, o)) {
      free(blk);
      fprintf(stderr, "\nFATAL: cannot write the report files into %s\n", o->out_dirname);
      goto out;
    }
    /* RESULTS block, src/stats_fastq.c:486-498 */
    printf("\n\n");
    printf("RESULTS\n");
    printf("=================================================\n");
    printf("Report files and images were stored in '%s' directory\n", o->out_dirname);
    if (o->filter_on) {
      printf("\nFiltering: enabled\n");
      printf("\tSo, statistics were computed for %lu of %lu reads.\n", (unsigned long)blk[HPGQ_S_NUM_PASSED],
             (unsigned long)(blk[HPGQ_S_NUM_PASSED] + blk[HPGQ_S_NUM_FAILED]));
    } else {
      printf("\nFiltering: disabled\n");
      printf("\tSo, statistics were computed for the whole input file.\n");
    }
    printf("=================================================\n");
    free(blk);
  } else if (o->command == HQ_CMD_FILTER) {
    /* src/filter_fastq.c:226-231 */
    printf("\n\n");
    printf("RESULTS\n");
    printf("=================================================\n");
    printf("Num. passed reads: %lu (%s)\n", (unsigned long)r->num_passed, pass_name);
    printf("Num. failed reads: %lu (%s)\n", (unsigned long)r->num_failed, fail_name);
    printf("=================================================\n");
  } else {
    /* src/edit_fastq.c:262-272 */
    printf("\n\n");
    printf("RESULTS\n");
    printf("=================================================\n");
    printf("Num. edited reads : %lu\n", (unsigned long)r->num_edited);
    printf("Output file       : %s\n", pass_name);
    if (o->filter_on) {
      printf("\nFiltering : Enabled\n");
      printf("\tNum. passed reads : %lu (%s)\n", (unsigned long)r->num_passed, pass_name);
      printf("\tNum. failed reads : %lu (%s)\n", (unsigned long)r->num_failed, fail_name);
    }
    printf("=================================================\n");
  }
  {
    char msg[256];
    const double dt = now_s() - t0;
    snprintf(msg, sizeof msg, "%s: %lu reads on %d GPU(s) in %.3f s (%.3g reads/s, %lu device-parsed + %lu host-parsed batches, %s)",
             o->command_name, (unsigned long)r->num_reads, r->n_gpus, dt, dt > 0 ? (double)r->num_reads / dt : 0.0,
             (unsigned long)r->n_text_batches, (unsigned long)r->n_host_batches, hpgq_version());
    logf_("INFO", msg);
    snprintf(msg, sizeof msg, "phases: setup %.3f s, batches %.3f s, merge + report %.3f s", t1 - t0, t2 - t1, now_s() - t2);
    logf_("INFO", msg);
  }
  status = 0;
out:
  if (comm_bg) pthread_join(comm_th, NULL);
  trace_("run (cum.)", t0);
  if (r->w_pass) hq_writer_close(r->w_pass);
  if (r->w_fail) hq_writer_close(r->w_fail);
  trace_("writers closed (cum.)", t0);
  for (int g = 0; g < HQ_MAX_GPUS; g++) {
    if (r->ctx[g]) hpgq_ctx_destroy(r->ctx[g]);
    for (int s = 0; s < HQ_SLOTS; s++) free_host_batch(&r->hb[g][s]);
  }
  trace_("contexts destroyed (cum.)", t0);
  if (r->reader) hq_reader_close(r->reader);
  free(r);
  trace_("reader closed (cum.)", t0);
  return status;
}

int main(int argc, char *argv[]) {
  const char *exec_name = argv[0];
  if (argc == 1 || !strcmp(argv[1], "-h") || !strcmp(argv[1], "--help")) {
    usage(exec_name);
    return 255; /* the reference exits with -1 here, src/hpg-fastq.c:30 */
  }
  const char *command_name = argv[1];
  /* Action flags of the legacy generation (old/main_hpg_fastq_old.c:131-140, old/README:52,85,119): `hpg-fastq --qc`,
   * `--filter`, `--prep` select stats / filter / edit; "if not flag at all only quality control is performed". */
  char **largv = NULL;
  if (command_name[0] == '-' && command_name[1] == '-') {
    command_name = "stats";
    largv = (char **)calloc((size_t)argc + 1, sizeof *largv);
    if (!largv) return 255;
    int n = 1;
    for (int a = 1; a < argc; a++) {
      if (!strcmp(argv[a], "--qc") || !strcmp(argv[a], "--quality-control")) command_name = "stats";
      else if (!strcmp(argv[a], "--filter")) command_name = "filter";
      else if (!strcmp(argv[a], "--prep") || !strcmp(argv[a], "--preprocessing")) command_name = "edit";
      else largv[n++] = argv[a];
    }
    largv[0] = (char *)command_name;
  }
  int pargc = argc - 1;
  char **pargv = argv + 1;
  if (largv) {
    pargc = 1;
    while (largv[pargc]) pargc++;
    pargv = largv;
  }
  if (strcmp(command_name, "stats") && strcmp(command_name, "filter") && strcmp(command_name, "edit")) {
    usage(exec_name);
    return 255;
  }
  hq_options_t *o = hq_options_parse(exec_name, command_name, pargc, pargv);
  free(largv);
  if (!o) return 255;
  if (!hq_options_validate(o)) {
    hq_options_free(o);
    return 255;
  }
  if (!hq_options_display(o, stdout)) {
    hq_options_free(o);
    return 255;
  }
  g_log = fopen("hpg-fastq.log", "w"); /* init_log_custom(..., "hpg-fastq.log", "w"), src/hpg-fastq.c:66 */
  int rc = run_command(o);
  if (!rc && o->paired) {
    /* second mate: an independent run of the same command ("the outputs are calculated seperately for both files",
     * old/README:181).  A synchronised drop of both mates is described nowhere in the reference tree. */
    free(o->in_filename);
    o->in_filename = strdup(o->in_filename2);
    rc = o->in_filename ? run_command(o) : 1;
  }
  if (!rc) logf_("INFO", "Done !");
  if (g_log) fclose(g_log);
  hq_options_free(o);
  return rc ? 1 : 0;
}
