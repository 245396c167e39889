// hpgq_api.cu -- implementation of the C ABI in include/hpg_fastq_cuda.h.
// One ctx per GPU, used by one host thread; n_slots streams double-buffer H2D copies against
// kernels; the flat u64 counter block lives on the device for the whole run.
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <new>

#include "hpgq_dev.cuh"
#include "hpgq_internal.h"
#include "../../include/hpgq_synth.h"

#define HPGQ_MAX_SLOTS 4
#define HQ_RING_MAX 8

static thread_local char g_last_error[512] = "";

static int fail_cuda(cudaError_t e, const char *what) {
  snprintf(g_last_error, sizeof g_last_error, "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
  return HPGQ_ERR_CUDA;
}
static int fail_msg(int code, const char *what) {
  snprintf(g_last_error, sizeof g_last_error, "%s", what);
  return code;
}
int hq_fail_msg(int code, const char *what) { return fail_msg(code, what); }
int hq_fail_cuda(cudaError_t e, const char *what) { return fail_cuda(e, what); }
#define CK(call)                                     \
  do {                                               \
    cudaError_t e__ = (call);                        \
    if (e__ != cudaSuccess) return fail_cuda(e__, #call); \
  } while (0)

struct hq_ring {
  uint8_t *base = nullptr;
  uint64_t piece = 0;
  int n = 0;
  uint64_t seq = 0;  // pieces handed out so far
  cudaEvent_t ev[HQ_RING_MAX] = {};
  bool used[HQ_RING_MAX] = {};
};

struct hpgq_slot {
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool busy = false, timed = false;
  bool dev_batch = false;
  bool uniform = false;  // every read of the current batch has max_len bases (proven or stated by the caller)
  uint32_t n_reads = 0, max_len = 0;
  uint64_t n_bytes = 0;
  uint32_t work = 0;
  // device input buffers (host-batch path only)
  uint8_t *d_seq = nullptr, *d_qual = nullptr;
  uint32_t *d_idx = nullptr;
  // pinned staging for pageable callers
  uint8_t *h_seq = nullptr, *h_qual = nullptr;
  uint32_t *h_idx = nullptr;
  // current batch pointers (device)
  const uint8_t *c_seq = nullptr, *c_qual = nullptr;
  const uint32_t *c_idx = nullptr;
  // outputs
  uint8_t *d_mask = nullptr, *h_mask = nullptr;
  uint8_t *d_cgflag = nullptr;  // chaos game: reads to re-emulate in double arithmetic
  uint32_t *d_rsum = nullptr;   // per-read sums of the one-pass statistics of long reads (zero between batches)
  hpgq_trim_t *d_trim = nullptr, *h_trim = nullptr;
  uint32_t *d_maxlen = nullptr, *h_small = nullptr; // h_small: [0..2]=max/min length, order flag [4..7]=totals [8..9]=HPGQ_S_BAD_BATCH (pinned)
  hq_compact_bufs cb = {};
  uint64_t cb_reads = 0, cb_bytes = 0;
  uint8_t *h_out_seq = nullptr, *h_out_qual = nullptr;
  uint32_t *h_out_idx = nullptr, *h_out_src = nullptr, *h_lens = nullptr;
  uint64_t h_out_bytes = 0, h_out_reads = 0;
  // FASTQ text path (hpgq_batch_submit_text): device chunk, parser scratch (the pinned staging is the ctx's ring)
  uint8_t *d_text = nullptr;
  uint64_t d_text_cap = 0;
  hq_parse_bufs pb = {};
  uint32_t *h_pinfo = nullptr;  // pinned copy of pb.info
  bool text_batch = false;
  bool text_pending = false;         // a chunk is parsed / being parsed, hpgq_batch_text_launch not called yet
  cudaEvent_t ev_parse = nullptr;    // end of the parse chain + info block copy
  uint8_t *d_otext = nullptr, *h_otext = nullptr;  // FASTQ text of the compacted output (hpgq_batch_wait_text)
  uint64_t otext_cap = 0, h_otext_cap = 0;
};

struct hpgq_ctx {
  int device = 0, sm_count = 0;
  hpgq_limits_t lim = {};
  hpgq_params_t par = {};
  bool have_params = false;
  hpgq_layout_t lay = {};
  unsigned long long *d_ctr = nullptr;
  unsigned long long *d_cgpack = nullptr;  // packed chaos-game shadow table, always all-zero between batches
  unsigned long long *d_kcpack = nullptr;  // packed k2-mer + chaos-game table (both reserved, k2 >= cg_k), all-zero between batches
  hpgq_slot slot[HPGQ_MAX_SLOTS];
  uint64_t launches = 0;
  uint64_t batches = 0;
  // ev1 of the most recent batch on any slot: the next batch's kernels wait for it, so that kernels of
  // different slots never queue behind each other inside an ev0..ev1 bracket (each QC kernel fills every SM
  // anyway).  Copies of the next batch still overlap the running kernel.
  cudaEvent_t last_done = nullptr;
  // FASTQ text path: pinned staging ring shared by the slots.  Pinning costs 0.39 ms per MB on the GPU box
  // (profiles/probe_alloc_r2.log), so a chunk goes through a few pieces that are reused as soon as their copy is done
  // instead of through a pinned buffer of its own size.
  hq_ring ring_in, ring_out;
};

// ------------------------------------------------------------------ misc
extern "C" const char *hpgq_version(void) { return "hpg-fastq-b200 0.1 (sm_100a)"; }
extern "C" const char *hpgq_last_error(void) { return g_last_error; }
extern "C" const char *hpgq_strerror(int s) {
  switch (s) {
    case HPGQ_OK: return "ok";
    case HPGQ_ERR_INVALID: return "invalid argument or limit exceeded";
    case HPGQ_ERR_CUDA: return "CUDA runtime error";
    case HPGQ_ERR_NOMEM: return "out of memory";
    case HPGQ_ERR_STATE: return "bad slot state";
    case HPGQ_ERR_NO_DEVICE: return "no usable CUDA device (there is no CPU fallback)";
    case HPGQ_ERR_NCCL: return "NCCL error";
    default: return "unknown status";
  }
}

extern "C" int hpgq_layout(uint32_t Lcap, uint32_t kmers, uint32_t k2, uint32_t cg_k, hpgq_layout_t *out) {
  if (!out || Lcap == 0 || k2 > 13 || cg_k > 12) return fail_msg(HPGQ_ERR_INVALID, "hpgq_layout: bad sizes");
  uint64_t sz[HPGQ_SEC_COUNT];
  sz[HPGQ_SEC_SCALARS] = HPGQ_S_COUNT;
  sz[HPGQ_SEC_LEN_HIST] = (uint64_t)Lcap + 1;
  sz[HPGQ_SEC_SUMQ_BY_LEN] = (uint64_t)Lcap + 1;
  sz[HPGQ_SEC_QUAL_HIST] = 256;
  sz[HPGQ_SEC_GC_HIST] = 104;
  for (int s = HPGQ_SEC_POS_ACCQ; s <= HPGQ_SEC_POS_N; ++s) sz[s] = Lcap;
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
    o += (sz[s] + 3) & ~3ull;
  }
  out->offset[HPGQ_SEC_COUNT] = o;
  return HPGQ_OK;
}

extern "C" int hpgq_counters_finalize_host(const hpgq_layout_t *lay, uint64_t *c) {
  if (!lay || !c) return fail_msg(HPGQ_ERR_INVALID, "finalize: null");
  const uint32_t Lcap = lay->max_read_len;
  const int secs[5] = {HPGQ_SEC_POS_A, HPGQ_SEC_POS_C, HPGQ_SEC_POS_G, HPGQ_SEC_POS_T, HPGQ_SEC_POS_N};
  const int dst[5] = {HPGQ_S_NUM_A, HPGQ_S_NUM_C, HPGQ_S_NUM_G, HPGQ_S_NUM_T, HPGQ_S_NUM_N};
  for (int k = 0; k < 5; ++k) {
    uint64_t t = 0;
    const uint64_t *p = c + lay->offset[secs[k]];
    for (uint32_t j = 0; j < Lcap; ++j) t += p[j];
    c[dst[k]] = t;
  }
  // 5-mer totals = row sums of the 5-mer x start-position table (the device keeps only the table)
  if (lay->kmers && lay->offset[HPGQ_SEC_K5 + 1] > lay->offset[HPGQ_SEC_K5]) {
    uint64_t *k5 = c + lay->offset[HPGQ_SEC_K5];
    const uint64_t *kp = c + lay->offset[HPGQ_SEC_K5_POS];
    for (uint32_t id = 0; id < 1024; ++id) {
      uint64_t t = 0;
      for (uint32_t j = 0; j < Lcap; ++j) t += kp[(uint64_t)id * Lcap + j];
      k5[id] = t;
    }
  }
  uint64_t mn = 100000, mx = 0;
  const uint64_t *h = c + lay->offset[HPGQ_SEC_LEN_HIST];
  for (uint32_t L = 0; L <= Lcap; ++L)
    if (h[L]) {
      if (L < mn) mn = L;
      if (L > mx) mx = L;
    }
  c[HPGQ_S_MIN_LENGTH] = mn;
  c[HPGQ_S_MAX_LENGTH] = mx;
  return HPGQ_OK;
}

extern "C" int hpgq_device_count(int *n) {
  if (!n) return HPGQ_ERR_INVALID;
  cudaError_t e = cudaGetDeviceCount(n);
  if (e != cudaSuccess) {
    *n = 0;
    fail_cuda(e, "cudaGetDeviceCount");
    return HPGQ_ERR_NO_DEVICE;
  }
  return HPGQ_OK;
}

// ------------------------------------------------------------------ context
static void ring_free(hq_ring &r) {
  cudaFreeHost(r.base);
  for (int i = 0; i < HQ_RING_MAX; ++i)
    if (r.ev[i]) cudaEventDestroy(r.ev[i]);
  r = hq_ring();
}

static void slot_free(hpgq_slot &s) {
  cudaFree(s.d_seq); cudaFree(s.d_qual); cudaFree(s.d_idx);
  cudaFreeHost(s.h_seq); cudaFreeHost(s.h_qual); cudaFreeHost(s.h_idx);
  cudaFree(s.d_mask); cudaFreeHost(s.h_mask);
  cudaFree(s.d_cgflag);
  cudaFree(s.d_rsum);
  cudaFree(s.d_trim); cudaFreeHost(s.h_trim);
  cudaFree(s.d_maxlen); cudaFreeHost(s.h_small);
  cudaFree(s.cb.scan); cudaFree(s.cb.block_sums); cudaFree(s.cb.out_seq); cudaFree(s.cb.out_qual);
  cudaFree(s.cb.out_idx); cudaFree(s.cb.out_src); cudaFree(s.cb.totals);
  cudaFree(s.cb.fz_agg); cudaFree(s.cb.fz_incl); cudaFree(s.cb.fz_status);
  cudaFree(s.cb.fail_seq); cudaFree(s.cb.fail_qual); cudaFree(s.cb.fail_idx); cudaFree(s.cb.fail_src);
  cudaFreeHost(s.h_out_seq); cudaFreeHost(s.h_out_qual); cudaFreeHost(s.h_out_idx); cudaFreeHost(s.h_out_src);
  cudaFreeHost(s.h_lens);
  cudaFree(s.d_text); cudaFreeHost(s.h_pinfo);
  cudaFree(s.d_otext); cudaFreeHost(s.h_otext);
  cudaFree(s.pb.tile_cnt); cudaFree(s.pb.cta_sum); cudaFree(s.pb.nl); cudaFree(s.pb.info);
  cudaFree(s.pb.seq_off); cudaFree(s.pb.qual_off); cudaFree(s.pb.hdr_off); cudaFree(s.pb.hdr_len);
  if (s.ev_parse) cudaEventDestroy(s.ev_parse);
  if (s.ev0) cudaEventDestroy(s.ev0);
  if (s.ev1) cudaEventDestroy(s.ev1);
  if (s.stream) cudaStreamDestroy(s.stream);
  s = hpgq_slot();
}

extern "C" int hpgq_ctx_destroy(hpgq_ctx_t *ctx) {
  if (!ctx) return HPGQ_OK;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  for (int i = 0; i < HPGQ_MAX_SLOTS; ++i) slot_free(ctx->slot[i]);
  ring_free(ctx->ring_in);
  ring_free(ctx->ring_out);
  cudaFree(ctx->d_ctr);
  cudaFree(ctx->d_cgpack);
  cudaFree(ctx->d_kcpack);
  delete ctx;
  return HPGQ_OK;
}

extern "C" int hpgq_ctx_create(int device, const hpgq_limits_t *lim, hpgq_ctx_t **out) {
  if (!lim || !out) return fail_msg(HPGQ_ERR_INVALID, "ctx_create: null argument");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev)
    return fail_msg(HPGQ_ERR_NO_DEVICE, "ctx_create: no such CUDA device; this library has no CPU fallback");
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10)
    return fail_msg(HPGQ_ERR_NO_DEVICE, "ctx_create: kernels are built for sm_100a only (device is older)");
  if (lim->max_read_len == 0 || lim->max_bytes_per_batch == 0 || lim->max_bytes_per_batch >= 0xFFFFFFC0ull ||
      lim->max_reads_per_batch == 0 || lim->n_slots > HPGQ_MAX_SLOTS)
    return fail_msg(HPGQ_ERR_INVALID, "ctx_create: bad limits");
  CK(cudaSetDevice(device));
  hpgq_ctx *c = new (std::nothrow) hpgq_ctx();
  if (!c) return HPGQ_ERR_NOMEM;
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  c->lim = *lim;
  if (c->lim.n_slots == 0) c->lim.n_slots = 2;
  int rc = hpgq_layout(lim->max_read_len, lim->kmers, lim->kmers ? lim->k2 : 0, lim->cg_k, &c->lay);
  if (rc) {
    delete c;
    return rc;
  }
  const size_t words = c->lay.offset[HPGQ_SEC_COUNT];
  cudaError_t e = cudaMalloc(&c->d_ctr, words * 8);
  if (e == cudaSuccess) e = cudaMemset(c->d_ctr, 0, words * 8);
  if (e == cudaSuccess && lim->cg_k) {
    e = cudaMalloc(&c->d_cgpack, (8ull << (2 * lim->cg_k)));
    if (e == cudaSuccess) e = cudaMemset(c->d_cgpack, 0, (8ull << (2 * lim->cg_k)));
    if (e == cudaSuccess && lim->kmers && lim->k2 >= lim->cg_k) {
      e = cudaMalloc(&c->d_kcpack, (8ull << (2 * lim->k2)));
      if (e == cudaSuccess) e = cudaMemset(c->d_kcpack, 0, (8ull << (2 * lim->k2)));
    }
  }
  const size_t nb = (size_t)lim->max_bytes_per_batch + 64, nr = (size_t)lim->max_reads_per_batch;
  for (uint32_t i = 0; i < c->lim.n_slots && e == cudaSuccess; ++i) {
    hpgq_slot &s = c->slot[i];
    e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreate(&s.ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&s.ev1);
    if (e == cudaSuccess) e = cudaMalloc(&s.d_maxlen, 64);
    if (e == cudaSuccess) e = cudaMallocHost(&s.h_small, 64);
    if (e == cudaSuccess) e = cudaMalloc(&s.d_mask, nr + 64);
    if (e == cudaSuccess && lim->cg_k) e = cudaMalloc(&s.d_cgflag, nr + 64);
    if (lim->host_staging) {
      if (e == cudaSuccess) e = cudaMalloc(&s.d_seq, nb);
      if (e == cudaSuccess) e = cudaMalloc(&s.d_qual, nb);
      if (e == cudaSuccess) e = cudaMalloc(&s.d_idx, (nr + 1) * 4 + 64);
    }
  }
  if (e != cudaSuccess) {
    fail_cuda(e, "ctx_create allocation");
    hpgq_ctx_destroy(c);
    return e == cudaErrorMemoryAllocation ? HPGQ_ERR_NOMEM : HPGQ_ERR_CUDA;
  }
  *out = c;
  return HPGQ_OK;
}

extern "C" int hpgq_get_layout(hpgq_ctx_t *ctx, hpgq_layout_t *out) {
  if (!ctx || !out) return HPGQ_ERR_INVALID;
  *out = ctx->lay;
  return HPGQ_OK;
}

extern "C" int hpgq_set_params(hpgq_ctx_t *ctx, const hpgq_params_t *p) {
  if (!ctx || !p) return fail_msg(HPGQ_ERR_INVALID, "set_params: null");
  if (p->phred != 33 && p->phred != 64) return fail_msg(HPGQ_ERR_INVALID, "set_params: phred must be 33 or 64");
  if ((p->work & HPGQ_WORK_KMERS) && !ctx->lay.kmers)
    return fail_msg(HPGQ_ERR_INVALID, "set_params: k-mer tables were not reserved in hpgq_limits_t");
  if ((p->work & HPGQ_WORK_KMERS) && (uint32_t)p->k2 != ctx->lay.k2)
    return fail_msg(HPGQ_ERR_INVALID, "set_params: k2 differs from the reserved table size");
  if ((p->work & HPGQ_WORK_CHAOS) && (p->cg_k < 1 || (uint32_t)p->cg_k != ctx->lay.cg_k))
    return fail_msg(HPGQ_ERR_INVALID, "set_params: cg_k differs from the reserved table size");
  if ((p->work & HPGQ_WORK_EDIT) && (p->work & (HPGQ_WORK_STATS | HPGQ_WORK_KMERS)))
    return fail_msg(HPGQ_ERR_INVALID, "set_params: EDIT cannot be combined with STATS/KMERS (the reference's edit command computes no statistics)");
  ctx->par = *p;
  ctx->have_params = true;
  return HPGQ_OK;
}

extern "C" int hpgq_counters_reset(hpgq_ctx_t *ctx) {
  if (!ctx) return HPGQ_ERR_INVALID;
  CK(cudaSetDevice(ctx->device));
  CK(cudaDeviceSynchronize());
  CK(cudaMemset(ctx->d_ctr, 0, ctx->lay.offset[HPGQ_SEC_COUNT] * 8));
  ctx->batches = 0;
  return HPGQ_OK;
}

extern "C" int hpgq_counters_download(hpgq_ctx_t *ctx, uint64_t *dst, size_t n_words) {
  if (!ctx || !dst || n_words < ctx->lay.offset[HPGQ_SEC_COUNT])
    return fail_msg(HPGQ_ERR_INVALID, "counters_download: buffer too small");
  CK(cudaSetDevice(ctx->device));
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(dst, ctx->d_ctr, ctx->lay.offset[HPGQ_SEC_COUNT] * 8, cudaMemcpyDeviceToHost));
  dst[HPGQ_S_BATCHES] = ctx->batches;
  int rc = hpgq_counters_finalize_host(&ctx->lay, dst);
  if (rc == HPGQ_OK && dst[HPGQ_S_BAD_BATCH])
    return fail_msg(HPGQ_ERR_INVALID, "a batch broke the length statement it was submitted with (hpgq_batch_t.reserved): counters are invalid");
  return rc;
}

extern "C" int hpgq_counters_device_ptr(hpgq_ctx_t *ctx, uint64_t **dptr, size_t *n_words) {
  if (!ctx || !dptr || !n_words) return HPGQ_ERR_INVALID;
  *dptr = reinterpret_cast<uint64_t *>(ctx->d_ctr);
  *n_words = ctx->lay.offset[HPGQ_SEC_COUNT];
  return HPGQ_OK;
}

extern "C" int hpgq_slot_stream(hpgq_ctx_t *ctx, int slot, void **stream) {
  if (!ctx || !stream || slot < 0 || (uint32_t)slot >= ctx->lim.n_slots) return HPGQ_ERR_INVALID;
  *stream = (void *)ctx->slot[slot].stream;
  return HPGQ_OK;
}

extern "C" int hpgq_launch_count(hpgq_ctx_t *ctx, uint64_t *n) {
  if (!ctx || !n) return HPGQ_ERR_INVALID;
  *n = ctx->launches;
  return HPGQ_OK;
}

// ------------------------------------------------------------------ raw memory helpers
extern "C" int hpgq_dev_alloc(hpgq_ctx_t *ctx, size_t bytes, void **dptr) {
  if (!ctx || !dptr) return HPGQ_ERR_INVALID;
  CK(cudaSetDevice(ctx->device));
  cudaError_t e = cudaMalloc(dptr, bytes ? bytes : 1);
  if (e != cudaSuccess) {
    fail_cuda(e, "cudaMalloc");
    return HPGQ_ERR_NOMEM;
  }
  return HPGQ_OK;
}
extern "C" int hpgq_dev_free(hpgq_ctx_t *ctx, void *dptr) {
  if (!ctx) return HPGQ_ERR_INVALID;
  CK(cudaSetDevice(ctx->device));
  CK(cudaFree(dptr));
  return HPGQ_OK;
}
extern "C" int hpgq_dev_download(hpgq_ctx_t *ctx, void *dst, const void *src, size_t bytes) {
  if (!ctx) return HPGQ_ERR_INVALID;
  CK(cudaSetDevice(ctx->device));
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
  return HPGQ_OK;
}
extern "C" int hpgq_dev_upload(hpgq_ctx_t *ctx, void *dst, const void *src, size_t bytes) {
  if (!ctx) return HPGQ_ERR_INVALID;
  CK(cudaSetDevice(ctx->device));
  CK(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
  return HPGQ_OK;
}
extern "C" int hpgq_host_alloc_pinned(size_t bytes, void **ptr) {
  if (!ptr) return HPGQ_ERR_INVALID;
  cudaError_t e = cudaMallocHost(ptr, bytes ? bytes : 1);
  if (e != cudaSuccess) {
    fail_cuda(e, "cudaMallocHost");
    return HPGQ_ERR_NOMEM;
  }
  return HPGQ_OK;
}
extern "C" int hpgq_host_free_pinned(void *ptr) {
  CK(cudaFreeHost(ptr));
  return HPGQ_OK;
}

#include "hpgq_api_batch.inc"
#include "hpgq_api_reduce.inc"
