/*
 * fastq_io.c -- FASTQ text reader (straight into SoA batches = fastq_batch_t layout) and ordered writer.
 *
 * Replaces fastq_fread_se() (called at src/stats_fastq.c:183, src/filter_fastq.c:114, src/edit_fastq.c:121) and
 * fastq_fwrite() (src/filter_fastq.c:165-169, src/edit_fastq.c:190-198), whose definitions live in the absent
 * bioinfo-libs submodule.  The reference builds an array_list of heap fastq_read_t per batch (AoS, SURVEY.md 8a
 * row 2); here the four lines of each record are scanned with memchr in the mapped file and the sequence /
 * quality bytes are copied once, directly into the pinned seq/qual/idx arrays that cudaMemcpyAsync ships.
 * Header lines never leave the mapping: the writer emits them from there.
 */
#define _GNU_SOURCE
#include "hpgq_host.h"

#include <errno.h>
#include <fcntl.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

struct hq_reader {
  int fd;
  const uint8_t *data;
  uint64_t size, pos;
  uint64_t record; /* records consumed so far (for messages) */
  char err[256];
};

int hq_reader_open(const char *path, hq_reader_t **out) {
  *out = NULL;
  hq_reader_t *r = (hq_reader_t *)calloc(1, sizeof *r);
  if (!r) return -1;
  r->fd = open(path, O_RDONLY);
  if (r->fd < 0) {
    free(r);
    return -1;
  }
  struct stat st;
  if (fstat(r->fd, &st) != 0) {
    close(r->fd);
    free(r);
    return -1;
  }
  r->size = (uint64_t)st.st_size;
  if (r->size) {
    void *p = mmap(NULL, r->size, PROT_READ, MAP_PRIVATE, r->fd, 0);
    if (p == MAP_FAILED) {
      close(r->fd);
      free(r);
      return -1;
    }
    madvise(p, r->size, MADV_SEQUENTIAL);
    r->data = (const uint8_t *)p;
  }
  *out = r;
  return 0;
}

void hq_reader_close(hq_reader_t *r) {
  if (!r) return;
  if (r->data) munmap((void *)r->data, r->size);
  if (r->fd >= 0) close(r->fd);
  free(r);
}

const uint8_t *hq_reader_data(const hq_reader_t *r) { return r->data; }
uint64_t hq_reader_size(const hq_reader_t *r) { return r->size; }
const char *hq_reader_error(const hq_reader_t *r) { return r->err; }

int hq_reader_scan(hq_reader_t *r, uint32_t *max_line, uint64_t *n_lines) {
  return hq_reader_scan_prefix(r, r->size, max_line, n_lines);
}

/* One memchr pass over [lo, hi): number of line breaks, first and last break, longest gap between two consecutive
 * breaks inside the range (as a line length without '\r').  Ranges are scanned by a few threads and stitched. */
typedef struct scan_part {
  const uint8_t *data;
  uint64_t lo, hi;
  uint64_t n_breaks, first, last, max_inner;
  pthread_t th;
} scan_part_t;

static void *scan_range(void *arg) {
  scan_part_t *p = (scan_part_t *)arg;
  uint64_t pos = p->lo, prev = UINT64_MAX;
  p->n_breaks = 0;
  p->first = p->last = UINT64_MAX;
  p->max_inner = 0;
  while (pos < p->hi) {
    const uint8_t *nl = (const uint8_t *)memchr(p->data + pos, '\n', p->hi - pos);
    if (!nl) break;
    const uint64_t e = (uint64_t)(nl - p->data);
    if (prev == UINT64_MAX) {
      p->first = e;
    } else {
      uint64_t len = e - prev - 1;
      if (len && p->data[e - 1] == '\r') len--;
      if (len > p->max_inner) p->max_inner = len;
    }
    prev = e;
    p->n_breaks++;
    pos = e + 1;
  }
  p->last = prev;
  return NULL;
}

int hq_reader_scan_prefix(hq_reader_t *r, uint64_t limit, uint32_t *max_line, uint64_t *n_lines) {
  if (limit > r->size) limit = r->size;
  /* a line that starts before `limit` counts in full: extend the range to the end of that line */
  uint64_t hi = limit;
  if (hi < r->size && hi > 0) {
    const uint8_t *nl = (const uint8_t *)memchr(r->data + hi - 1, '\n', r->size - (hi - 1));
    hi = nl ? (uint64_t)(nl - r->data) + 1 : r->size;
  }
  enum { T = 16 };
  scan_part_t part[T];
  long ncpu = sysconf(_SC_NPROCESSORS_ONLN);
  if (getenv("HPGQ_SCAN_THREADS")) ncpu = atol(getenv("HPGQ_SCAN_THREADS")); /* developer knob */
  const int want = ncpu < 1 ? 1 : (ncpu > T ? T : (int)ncpu);
  const int nt = hi >= (64u << 20) ? want : (hi >= (1u << 20) ? (want < 4 ? want : 4) : 1);
  const uint64_t per = (hi + (uint64_t)nt - 1) / (uint64_t)nt;
  for (int t = 0; t < nt; t++) {
    part[t].data = r->data;
    part[t].lo = (uint64_t)t * per < hi ? (uint64_t)t * per : hi;
    part[t].hi = part[t].lo + per < hi ? part[t].lo + per : hi;
  }
  for (int t = 1; t < nt; t++)
    if (pthread_create(&part[t].th, NULL, scan_range, &part[t])) {
      scan_range(&part[t]);
      part[t].th = 0;
    }
  scan_range(&part[0]);
  for (int t = 1; t < nt; t++)
    if (part[t].th) pthread_join(part[t].th, NULL);
  uint64_t lines = 0, mx = 0, prev = UINT64_MAX; /* prev: last break seen so far (none: the first line starts at 0) */
  for (int t = 0; t < nt; t++) {
    if (!part[t].n_breaks) continue;
    uint64_t len = prev == UINT64_MAX ? part[t].first : part[t].first - prev - 1;
    if (len && r->data[part[t].first - 1] == '\r') len--;
    if (len > mx) mx = len;
    if (part[t].max_inner > mx) mx = part[t].max_inner;
    lines += part[t].n_breaks;
    prev = part[t].last;
  }
  const uint64_t tail_from = prev == UINT64_MAX ? 0 : prev + 1;
  if (tail_from < hi) { /* last line without a line break */
    uint64_t len = hi - tail_from;
    if (r->data[hi - 1] == '\r') len--;
    if (len > mx) mx = len;
    lines++;
  }
  if (mx > 0xFFFFFFF0ull) return -1;
  *max_line = (uint32_t)mx;
  *n_lines = lines;
  return 0;
}

/* one line [*b, *e) without terminator; advances *pos past the '\n'.  0 at end of data. */
static inline int next_line(const hq_reader_t *r, uint64_t *pos, uint64_t *b, uint64_t *e) {
  if (*pos >= r->size) return 0;
  const uint8_t *nl = (const uint8_t *)memchr(r->data + *pos, '\n', r->size - *pos);
  uint64_t end = nl ? (uint64_t)(nl - r->data) : r->size;
  *b = *pos;
  *e = (end > *pos && r->data[end - 1] == '\r') ? end - 1 : end;
  *pos = end + 1;
  return 1;
}

int hq_reader_next(hq_reader_t *r, hq_host_batch_t *bt) {
  uint32_t n = 0, mx = 0;
  uint64_t o = 0;
  bt->idx[0] = 0;
  while (n < bt->cap_reads) {
    uint64_t save = r->pos, hb, he, sb, se, pb, pe, qb, qe;
    /* blank lines between records (and a trailing one) are skipped */
    int got;
    while ((got = next_line(r, &r->pos, &hb, &he)) && he == hb) save = r->pos;
    if (!got) break;
    if (r->data[hb] != '@') {
      snprintf(r->err, sizeof r->err, "record %llu: header line does not start with '@' (file offset %llu)",
               (unsigned long long)(r->record + 1), (unsigned long long)hb);
      return -2;
    }
    if (!next_line(r, &r->pos, &sb, &se) || !next_line(r, &r->pos, &pb, &pe) || !next_line(r, &r->pos, &qb, &qe)) {
      snprintf(r->err, sizeof r->err, "record %llu: truncated record at end of file", (unsigned long long)(r->record + 1));
      return -2;
    }
    if (pe == pb || r->data[pb] != '+') {
      snprintf(r->err, sizeof r->err, "record %llu: third line does not start with '+'", (unsigned long long)(r->record + 1));
      return -2;
    }
    const uint64_t L = se - sb;
    if (qe - qb != L) {
      snprintf(r->err, sizeof r->err, "record %llu: sequence has %llu bases but quality has %llu", (unsigned long long)(r->record + 1),
               (unsigned long long)L, (unsigned long long)(qe - qb));
      return -2;
    }
    if (L > bt->cap_bytes) {
      snprintf(r->err, sizeof r->err, "record %llu: read of %llu bases exceeds the batch capacity", (unsigned long long)(r->record + 1),
               (unsigned long long)L);
      return -2;
    }
    if (o + L > bt->cap_bytes) { /* batch is full by bytes: leave the record for the next call */
      r->pos = save;
      break;
    }
    memcpy(bt->seq + o, r->data + sb, L);
    memcpy(bt->qual + o, r->data + qb, L);
    bt->hdr_off[n] = hb;
    bt->hdr_len[n] = (uint32_t)(he - hb);
    o += L;
    bt->idx[++n] = (uint32_t)o;
    if (L > mx) mx = (uint32_t)L;
    r->record++;
  }
  bt->n_reads = n;
  bt->max_len = mx;
  return n ? 1 : 0;
}

/* start of the line after the one that begins at p, or size if there is none */
static uint64_t line_after(const hq_reader_t *r, uint64_t p) {
  if (p >= r->size) return r->size;
  const uint8_t *nl = (const uint8_t *)memchr(r->data + p, '\n', r->size - p);
  return nl ? (uint64_t)(nl - r->data) + 1 : r->size;
}

int hq_reader_peek_text(const hq_reader_t *r, uint64_t max_bytes, uint64_t *off, uint64_t *len) {
  if (r->pos >= r->size) return 0;
  *off = r->pos;
  if (r->size - r->pos <= max_bytes) {
    *len = r->size - r->pos;
    return 1;
  }
  /* walk back over line starts from pos + max_bytes; a header is the only line that starts with '@' and is
   * followed, two lines on, by a line that starts with '+' (a quality line may start with '@', but then the line
   * two further is a sequence) */
  uint64_t p = r->pos + max_bytes;
  const uint64_t floor = r->pos + max_bytes / 2;
  while (p > floor) {
    const uint8_t *nl = (const uint8_t *)memrchr(r->data + floor, '\n', p - floor);
    if (!nl) break;
    const uint64_t ls = (uint64_t)(nl - r->data) + 1;
    if (ls < r->size && ls <= r->pos + max_bytes && r->data[ls] == '@') {
      const uint64_t l2 = line_after(r, line_after(r, ls));
      if (l2 < r->size && r->data[l2] == '+') {
        *len = ls - r->pos;
        return 1;
      }
    }
    p = (uint64_t)(nl - r->data);
  }
  return -1;
}

void hq_reader_advance(hq_reader_t *r, uint64_t bytes, uint64_t records) {
  r->pos += bytes;
  r->record += records;
}

int hq_reader_fd(const hq_reader_t *r) { return r->fd; }

void hq_reader_rewind(hq_reader_t *r, uint64_t pos) { r->pos = pos; }

/* ------------------------------------------------------------------ writer */
struct hq_writer {
  int fd;
  uint8_t *buf;
  size_t cap, len;
  int failed;
};

#define HQ_WBUF (8u << 20)

int hq_writer_open(const char *path, hq_writer_t **out) {
  *out = NULL;
  hq_writer_t *w = (hq_writer_t *)calloc(1, sizeof *w);
  if (!w) return -1;
  w->fd = open(path, O_RDWR | O_CREAT | O_TRUNC, 0644); /* read-write: hpgq_batch_wait_text_fd fills it through a mapping */
  w->buf = (uint8_t *)malloc(HQ_WBUF);
  if (w->fd < 0 || !w->buf) {
    if (w->fd >= 0) close(w->fd);
    free(w->buf);
    free(w);
    return -1;
  }
  w->cap = HQ_WBUF;
  *out = w;
  return 0;
}

static void wflush(hq_writer_t *w) {
  size_t off = 0;
  while (off < w->len) {
    ssize_t k = write(w->fd, w->buf + off, w->len - off);
    if (k < 0) {
      if (errno == EINTR) continue;
      w->failed = 1;
      break;
    }
    off += (size_t)k;
  }
  w->len = 0;
}

static inline void wput(hq_writer_t *w, const uint8_t *p, size_t n) {
  if (w->len + n + 8 > w->cap) {
    wflush(w);
    if (n + 8 > w->cap) { /* a single line longer than the buffer */
      size_t off = 0;
      while (off < n) {
        ssize_t k = write(w->fd, p + off, n - off);
        if (k < 0) {
          if (errno == EINTR) continue;
          w->failed = 1;
          return;
        }
        off += (size_t)k;
      }
      return;
    }
  }
  memcpy(w->buf + w->len, p, n);
  w->len += n;
}

int hq_writer_put(hq_writer_t *w, const uint8_t *file_data, const hq_host_batch_t *b, const uint8_t *seq,
                  const uint8_t *qual, const uint32_t *idx, const uint32_t *src, uint32_t from, uint32_t to) {
  for (uint32_t k = from; k < to; k++) {
    const uint32_t s = src ? src[k] : k;
    const uint32_t o = idx[k], L = idx[k + 1] - o;
    wput(w, file_data + b->hdr_off[s], b->hdr_len[s]);
    w->buf[w->len++] = '\n';
    wput(w, seq + o, L);
    w->buf[w->len++] = '\n';
    w->buf[w->len++] = '+';
    w->buf[w->len++] = '\n';
    wput(w, qual + o, L);
    w->buf[w->len++] = '\n';
  }
  return w->failed ? -1 : 0;
}

int hq_writer_fd(hq_writer_t *w) {
  wflush(w);
  return w->failed ? -1 : w->fd;
}

int hq_writer_put_raw(hq_writer_t *w, const uint8_t *text, uint64_t n) {
  /* large blocks go to write() directly (after the buffered bytes), without another copy */
  wput(w, text, (size_t)n);
  return w->failed ? -1 : 0;
}

int hq_writer_close(hq_writer_t *w) {
  if (!w) return 0;
  wflush(w);
  int rc = w->failed ? -1 : 0;
  if (close(w->fd) != 0) rc = -1;
  free(w->buf);
  free(w);
  return rc;
}
