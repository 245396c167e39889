// hpgq_parse.cu -- FASTQ text -> SoA batch on the device.
//
// Replaces the per-record scan of fastq_fread_se() (bioinfo-libs, called at reference src/stats_fastq.c:183,
// src/filter_fastq.c:114, src/edit_fastq.c:121) for well-formed input: strict four-line records, LF or CRLF line
// ends.  The host ships the raw text of a chunk that starts at a header line and ends after a quality line; the
// kernels find the line breaks, cut the records and pack sequence / quality bytes into the fastq_batch_t layout
// (seq, qual, idx[n+1]) that every QC kernel of this library consumes, plus the (offset, length) of each header
// line so that the writer can emit it from the host's copy of the text.  Anything else (blank lines, a record that
// is not four lines, '@' / '+' missing, sequence and quality of different length) is reported in hq_parse_info and
// left to the host reader, which owns the error messages.
//
//   1. hq_parse_count   line breaks per 16 KB tile                       (text read once, 16 B per load)
//   2. hq_parse_scan    exclusive scan of the tile counts (one CTA)
//   3. hq_parse_fill    position of every line break, in order            (text read a second time)
//   4. hq_parse_cut     per record: four line spans, validation, length; per-CTA sums of the lengths
//   5. hq_parse_scan    exclusive scan of the per-CTA sums
//   6. hq_parse_idx     lengths -> offsets idx[n+1]
//   7. hq_parse_gather  sequence and quality bytes -> packed arrays       (one warp per read)
// All grids are fixed (device-side loop bounds come from hq_parse_info), so the chain needs no host round trip;
// the host reads hq_parse_info once, after the last kernel.
// Bound: HBM, about 3 x text bytes read + 2 x bases written; in practice the chunk arrives over PCIe (55 GB/s), and
// the whole chain runs at a multiple of that.
#include "hpgq_dev.cuh"
#include "hpgq_internal.h"

namespace {

constexpr int PT = 256;                 // threads per CTA
constexpr uint32_t PER_THREAD = 64;     // text bytes per thread (4 x uint4)
constexpr uint32_t TILE = PT * PER_THREAD;

// bit 7 of every byte of w that equals '\n'
__device__ __forceinline__ uint32_t nl_flags(uint32_t w) { return __vcmpeq4(w, 0x0A0A0A0Au) & 0x80808080u; }

__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *warp_sums, uint32_t *total) {
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
    if (lane >= (uint32_t)o) inc += t;
  }
  __syncthreads();  // protects warp_sums against the previous call
  if (lane == 31) warp_sums[warp] = inc;
  __syncthreads();
  uint32_t base = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < PT / 32; ++w) {
    const uint32_t s = warp_sums[w];
    if ((uint32_t)w < warp) base += s;
    tot += s;
  }
  if (total) *total = tot;
  return base + inc - v;
}

// the 64 text bytes of this thread as 16 words; bytes at or past n read as 0 (text is readable to the next 16 B)
__device__ __forceinline__ void load_thread_words(const uint8_t *text, uint32_t n, uint32_t p0, uint32_t w[16]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t p = p0 + 16u * i;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (p < n) v = __ldg(reinterpret_cast<const uint4 *>(text + p));
    w[4 * i + 0] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
  }
  // mask the bytes past the end of the text in the last partial 16-byte granule
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const uint32_t p = p0 + 4u * i;
    if (p + 4u > n) w[i] = p >= n ? 0u : (w[i] & (0xFFFFFFFFu >> (8u * (p + 4u - n))));
  }
}

__global__ void __launch_bounds__(PT) hq_parse_count(const uint8_t *text, uint32_t n, uint32_t n_tiles, uint32_t *tile_cnt) {
  __shared__ uint32_t ws[PT / 32];
  for (uint32_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    uint32_t w[16];
    load_thread_words(text, n, t * TILE + threadIdx.x * PER_THREAD, w);
    uint32_t c = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) c += __popc(nl_flags(w[i]));
    uint32_t tot;
    (void)block_excl_scan(c, ws, &tot);
    if (threadIdx.x == 0) tile_cnt[t] = tot;
  }
}

// in-place exclusive scan of v[0..m) by one CTA; the total goes to *total_out.  When `text` is given (line-break
// pass) a chunk that does not end in '\n' gets a virtual line break at n, and info[INFO_LINES] is the line count.
__global__ void __launch_bounds__(1024) hq_parse_scan(uint32_t *v, uint32_t m, uint32_t *total_out, const uint8_t *text,
                                                     uint32_t n, uint32_t *nl, uint32_t cap_lines) {
  __shared__ uint32_t ws[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  for (uint32_t b = 0; b < m; b += 1024u) {
    const uint32_t i = b + threadIdx.x;
    const uint32_t x = i < m ? v[i] : 0u;
    uint32_t inc = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
      if (lane >= (uint32_t)o) inc += t;
    }
    if (lane == 31) ws[warp] = inc;
    __syncthreads();
    uint32_t base = carry;
    for (uint32_t w = 0; w < warp; ++w) base += ws[w];
    if (i < m) v[i] = base + inc - x;
    __syncthreads();
    if (threadIdx.x == 1023) carry = base + inc;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    uint32_t tot = carry;
    if (text && n && text[n - 1] != '\n') {
      if (tot < cap_lines) nl[tot] = n;
      tot++;
    }
    *total_out = tot;
  }
}

__global__ void __launch_bounds__(PT) hq_parse_fill(const uint8_t *text, uint32_t n, uint32_t n_tiles, const uint32_t *tile_off,
                                                  uint32_t *nl, uint32_t cap_lines) {
  __shared__ uint32_t ws[PT / 32];
  for (uint32_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    uint32_t w[16];
    const uint32_t p0 = t * TILE + threadIdx.x * PER_THREAD;
    load_thread_words(text, n, p0, w);
    uint32_t c = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) c += __popc(nl_flags(w[i]));
    uint32_t k = tile_off[t] + block_excl_scan(c, ws, nullptr);
    if (c) {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        uint32_t f = nl_flags(w[i]);
        while (f) {
          const uint32_t byte = (uint32_t)(__ffs((int)f) - 1) >> 3;
          f &= f - 1u;
          if (k < cap_lines) nl[k] = p0 + 4u * i + byte;
          ++k;
        }
      }
    }
  }
}

// Record r is lines 4r .. 4r+3; line j spans (nl[j-1], nl[j]) exclusive of the breaks.  A '\r' before the break is
// not part of the line.  Per-CTA sums of the read lengths go to cta_sum[blockIdx.x] (ranges of hq_parse_range).
__device__ __forceinline__ void hq_parse_range(uint32_t R, uint32_t *b, uint32_t *e) {
  const uint32_t per = (R + gridDim.x - 1u) / gridDim.x;
  const uint32_t lo = min(R, blockIdx.x * per);
  *b = lo;
  *e = min(R, lo + per);
}

__global__ void __launch_bounds__(PT) hq_parse_cut(const uint8_t *text, const uint32_t *nl, uint32_t *info, uint32_t cap_reads,
                                                 uint32_t cap_lines, uint32_t *idx, uint32_t *seq_off, uint32_t *qual_off,
                                                 uint32_t *hdr_off, uint32_t *hdr_len, uint32_t *cta_sum) {
  __shared__ uint32_t ws[PT / 32];
  const uint32_t lines = info[HQ_PINFO_LINES];
  uint32_t R = lines / 4u;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    info[HQ_PINFO_READS] = R;
    if (lines % 4u) atomicOr(&info[HQ_PINFO_STATUS], HQ_PSTAT_LINES);
    if (R > cap_reads || lines > cap_lines) atomicOr(&info[HQ_PINFO_STATUS], HQ_PSTAT_TOO_MANY);
  }
  if (R > cap_reads || lines > cap_lines) R = 0;  // nothing below may index past the arrays
  uint32_t rb, re;
  hq_parse_range(R, &rb, &re);
  uint32_t sum = 0, mx = 0, mn = 0xFFFFFFFFu, bad = 0xFFFFFFFFu;
  for (uint32_t r = rb + threadIdx.x; r < re; r += PT) {
    const uint32_t hb = r ? nl[4u * r - 1u] + 1u : 0u;
    uint32_t he = nl[4u * r], se = nl[4u * r + 1u], pe = nl[4u * r + 2u], qe = nl[4u * r + 3u];
    const uint32_t sb = he + 1u, pb = se + 1u, qb = pe + 1u;
    if (he > hb && text[he - 1u] == '\r') --he;
    if (se > sb && text[se - 1u] == '\r') --se;
    if (pe > pb && text[pe - 1u] == '\r') --pe;
    if (qe > qb && text[qe - 1u] == '\r') --qe;
    const uint32_t L = se - sb;
    const bool ok = he > hb && text[hb] == '@' && pe > pb && text[pb] == '+' && (qe - qb) == L;
    if (!ok) bad = min(bad, r);
    idx[r + 1u] = L;
    seq_off[r] = sb;
    qual_off[r] = qb;
    hdr_off[r] = hb;
    hdr_len[r] = he - hb;
    sum += L;
    mx = max(mx, L);
    mn = min(mn, L);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
    mn = min(mn, __shfl_xor_sync(0xFFFFFFFFu, mn, o));
    bad = min(bad, __shfl_xor_sync(0xFFFFFFFFu, bad, o));
  }
  if ((threadIdx.x & 31u) == 0) {
    if (mx) atomicMax(&info[HQ_PINFO_MAXLEN], mx);
    atomicMin(&info[HQ_PINFO_MINLEN], mn);
    if (bad != 0xFFFFFFFFu) {
      atomicMin(&info[HQ_PINFO_BAD], bad);
      atomicOr(&info[HQ_PINFO_STATUS], HQ_PSTAT_RECORD);
    }
  }
  uint32_t tot;
  (void)block_excl_scan(sum, ws, &tot);
  if (threadIdx.x == 0) cta_sum[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(PT) hq_parse_idx(uint32_t *info, uint32_t cap_reads, uint64_t cap_bytes, uint32_t *idx,
                                                 const uint32_t *cta_off) {
  __shared__ uint32_t ws[PT / 32];
  uint32_t R = info[HQ_PINFO_READS];
  if (R > cap_reads) R = 0;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    idx[0] = 0u;
    const uint32_t bases = info[HQ_PINFO_BASES];
    if (bases > cap_bytes) atomicOr(&info[HQ_PINFO_STATUS], HQ_PSTAT_TOO_MANY);
  }
  uint32_t rb, re;
  hq_parse_range(R, &rb, &re);
  uint32_t run = cta_off[blockIdx.x];
  for (uint32_t r0 = rb; r0 < re; r0 += PT) {
    const uint32_t r = r0 + threadIdx.x;
    const uint32_t L = r < re ? idx[r + 1u] : 0u;
    uint32_t tot;
    const uint32_t ex = block_excl_scan(L, ws, &tot);
    if (r < re) idx[r + 1u] = run + ex + L;
    run += tot;
  }
}

// One warp per read: lanes stride the bytes.  Source spans start at arbitrary byte phases of the text and the
// destination at arbitrary phases of the packed arrays, so this is a byte copy; it moves 2 x bases per chunk and is
// far from limiting (the chunk took 40 x longer to arrive over PCIe).
__global__ void __launch_bounds__(PT) hq_parse_gather(const uint8_t *text, const uint32_t *info, uint32_t cap_reads,
                                                    const uint32_t *idx, const uint32_t *seq_off, const uint32_t *qual_off,
                                                    uint8_t *seq, uint8_t *qual) {
  if (info[HQ_PINFO_STATUS]) return;
  uint32_t R = info[HQ_PINFO_READS];
  if (R > cap_reads) return;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t warps = gridDim.x * (PT / 32);
  for (uint32_t r = blockIdx.x * (PT / 32) + (threadIdx.x >> 5); r < R; r += warps) {
    const uint32_t o = idx[r], L = idx[r + 1u] - o;
    const uint8_t *s = text + seq_off[r], *q = text + qual_off[r];
    for (uint32_t j = lane; j < L; j += 32u) {
      seq[o + j] = s[j];
      qual[o + j] = q[j];
    }
  }
}

// ------------------------------------------------------------------ FASTQ text of the compacted output
// Output read k (passed reads first, then failed, each in input order: hq_compact_*) becomes the record
//   header of source read out_src[k] '\n' sequence '\n' '+' '\n' quality '\n'
// (the reference's fastq_fwrite layout).  rec[k+1] first holds the record length, then (in place)
// the end offset; oinfo[0] = bytes of the passed records, oinfo[1] = bytes of all records.
__global__ void __launch_bounds__(PT) hq_compose_len(uint32_t n, const uint32_t *out_idx, const uint32_t *out_src,
                                                   const uint32_t *hdr_len, uint32_t *rec, uint32_t *cta_sum) {
  __shared__ uint32_t ws[PT / 32];
  uint32_t rb, re;
  hq_parse_range(n, &rb, &re);
  uint32_t sum = 0;
  for (uint32_t k = rb + threadIdx.x; k < re; k += PT) {
    const uint32_t len = hdr_len[out_src[k]] + 2u * (out_idx[k + 1u] - out_idx[k]) + 5u;
    rec[k + 1u] = len;
    sum += len;
  }
  uint32_t tot;
  (void)block_excl_scan(sum, ws, &tot);
  if (threadIdx.x == 0) cta_sum[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(PT) hq_compose_off(uint32_t n, uint32_t *rec, const uint32_t *cta_off) {
  __shared__ uint32_t ws[PT / 32];
  if (blockIdx.x == 0 && threadIdx.x == 0) rec[0] = 0u;
  uint32_t rb, re;
  hq_parse_range(n, &rb, &re);
  uint32_t run = cta_off[blockIdx.x];
  for (uint32_t k0 = rb; k0 < re; k0 += PT) {
    const uint32_t k = k0 + threadIdx.x;
    const uint32_t len = k < re ? rec[k + 1u] : 0u;
    uint32_t tot;
    const uint32_t ex = block_excl_scan(len, ws, &tot);
    if (k < re) rec[k + 1u] = run + ex + len;
    run += tot;
  }
}

__global__ void __launch_bounds__(PT) hq_compose_gather(uint32_t n, const uint8_t *text, const uint32_t *hdr_off,
                                                      const uint32_t *hdr_len, const uint8_t *out_seq, const uint8_t *out_qual,
                                                      const uint32_t *out_idx, const uint32_t *out_src, const uint32_t *rec,
                                                      const uint32_t *totals, uint32_t *oinfo, uint8_t *otext) {
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    oinfo[0] = rec[min(totals[0], n)];
    oinfo[1] = rec[n];
  }
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t warps = gridDim.x * (PT / 32);
  for (uint32_t k = blockIdx.x * (PT / 32) + (threadIdx.x >> 5); k < n; k += warps) {
    const uint32_t src = out_src[k], o = out_idx[k], L = out_idx[k + 1u] - o, H = hdr_len[src];
    uint8_t *d = otext + rec[k];
    const uint8_t *h = text + hdr_off[src];
    for (uint32_t j = lane; j < H; j += 32u) d[j] = h[j];
    d += H;
    const uint8_t *sq = out_seq + o, *ql = out_qual + o;
    for (uint32_t j = lane; j < L; j += 32u) {
      d[1u + j] = sq[j];
      d[L + 4u + j] = ql[j];
    }
    if (lane == 0) {
      d[0] = '\n';
      d[L + 1u] = '\n';
      d[L + 2u] = '+';
      d[L + 3u] = '\n';
      d[2u * L + 4u] = '\n';
    }
  }
}

}  // namespace

cudaError_t hq_launch_compose(const hq_parse_bufs &b, uint32_t n, const hq_compact_bufs &cb, uint32_t *rec, uint32_t *oinfo,
                              uint8_t *otext, int sm_count, cudaStream_t st, uint64_t *launches) {
  if (!n) return cudaMemsetAsync(oinfo, 0, 8, st);
  const int grid = sm_count * 8;
  hq_compose_len<<<grid, PT, 0, st>>>(n, cb.out_idx, cb.out_src, b.hdr_len, rec, b.cta_sum);
  hq_parse_scan<<<1, 1024, 0, st>>>(b.cta_sum, (uint32_t)grid, oinfo + 1, nullptr, 0, nullptr, 0);
  hq_compose_off<<<grid, PT, 0, st>>>(n, rec, b.cta_sum);
  hq_compose_gather<<<grid, PT, 0, st>>>(n, b.text, b.hdr_off, b.hdr_len, cb.out_seq, cb.out_qual, cb.out_idx, cb.out_src, rec,
                                        cb.totals, oinfo, otext);
  *launches += 4;
  return cudaGetLastError();
}

cudaError_t hq_launch_parse(const hq_parse_bufs &b, int sm_count, cudaStream_t st, uint64_t *launches) {
  const uint32_t n = b.n_bytes;
  const uint32_t n_tiles = (n + TILE - 1u) / TILE;
  const int grid = sm_count * 8;
  cudaError_t e = cudaMemsetAsync(b.info, 0, HQ_PINFO_COUNT * 4, st);
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(b.info + HQ_PINFO_BAD, 0xFF, 4, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(b.info + HQ_PINFO_MINLEN, 0xFF, 4, st);
  if (e != cudaSuccess) return e;
  if (n_tiles) {
    hq_parse_count<<<min((uint32_t)grid, n_tiles), PT, 0, st>>>(b.text, n, n_tiles, b.tile_cnt);
    hq_parse_scan<<<1, 1024, 0, st>>>(b.tile_cnt, n_tiles, b.info + HQ_PINFO_LINES, b.text, n, b.nl, b.cap_lines);
    hq_parse_fill<<<min((uint32_t)grid, n_tiles), PT, 0, st>>>(b.text, n, n_tiles, b.tile_cnt, b.nl, b.cap_lines);
    *launches += 3;
  }
  hq_parse_cut<<<grid, PT, 0, st>>>(b.text, b.nl, b.info, b.cap_reads, b.cap_lines, b.idx, b.seq_off, b.qual_off, b.hdr_off,
                                   b.hdr_len, b.cta_sum);
  hq_parse_scan<<<1, 1024, 0, st>>>(b.cta_sum, (uint32_t)grid, b.info + HQ_PINFO_BASES, nullptr, 0, nullptr, 0);
  hq_parse_idx<<<grid, PT, 0, st>>>(b.info, b.cap_reads, b.cap_bytes, b.idx, b.cta_sum);
  hq_parse_gather<<<grid, PT, 0, st>>>(b.text, b.info, b.cap_reads, b.idx, b.seq_off, b.qual_off, b.seq, b.qual);
  *launches += 4;
  return cudaGetLastError();
}
