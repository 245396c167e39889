// hpgq_compact.cu -- order-preserving partition of the (trimmed) reads of a batch into
// "passed" then "failed", each in input order: what fastq_fwrite(passed_reads) and
// fastq_fwrite(failed_reads) emit in the reference consumers (src/filter_fastq.c:161-174,
// src/edit_fastq.c:184-206) when run with one worker thread.
//
// Three steps: per-read (bytes, count) records -> exclusive scan (block scan + block-sum scan +
// add) -> warp-per-read copy of seq and qual bytes to their final offsets.
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include "hpgq_dev.cuh"
#include "hpgq_internal.h"
#include "hpgq_regs.cuh"
#include "../../include/hpgq_synth.h"

namespace {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 4; // 1024 records per block

__device__ __forceinline__ uint4 add4(uint4 a, uint4 b) { return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }

__device__ __forceinline__ uint4 shfl_up4(uint4 v, int d) {
  return make_uint4(__shfl_up_sync(0xFFFFFFFFu, v.x, d), __shfl_up_sync(0xFFFFFFFFu, v.y, d),
                    __shfl_up_sync(0xFFFFFFFFu, v.z, d), __shfl_up_sync(0xFFFFFFFFu, v.w, d));
}

// record of read i (i == n gives the zero sentinel so that scan[n] holds the totals)
__device__ __forceinline__ uint4 make_record(const hq_kargs &a, uint32_t i) {
  if (i >= a.n_reads) return make_uint4(0, 0, 0, 0);
  const uint32_t ln = a.trim ? a.trim[i].len : (a.idx[i + 1] - a.idx[i]);
  const bool pass = a.mask ? (a.mask[i] != 0) : true;
  return pass ? make_uint4(ln, 0, 1, 0) : make_uint4(0, ln, 0, 1);
}

// inclusive block scan of one uint4 per thread; returns the thread's inclusive value, total in *tot
__device__ __forceinline__ uint4 block_scan_incl(uint4 v, uint4 *tot) {
  __shared__ uint4 wsum[SCAN_THREADS / 32];
  const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint4 u = shfl_up4(v, d);
    if ((int)lane >= d) v = add4(v, u);
  }
  if (lane == 31) wsum[w] = v;
  __syncthreads();
  if (w == 0) {
    uint4 s = lane < SCAN_THREADS / 32 ? wsum[lane] : make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int d = 1; d < SCAN_THREADS / 32; d <<= 1) {
      const uint4 u = shfl_up4(s, d);
      if ((int)lane >= d) s = add4(s, u);
    }
    if (lane < SCAN_THREADS / 32) wsum[lane] = s;
  }
  __syncthreads();
  if (w > 0) v = add4(v, wsum[w - 1]);
  *tot = wsum[SCAN_THREADS / 32 - 1];
  __syncthreads();
  return v;
}

// pass 1: per-block totals
__global__ void __launch_bounds__(SCAN_THREADS) hq_compact_sums(hq_kargs a, uint4 *block_sums) {
  const uint32_t base = blockIdx.x * (SCAN_THREADS * SCAN_ITEMS) + threadIdx.x * SCAN_ITEMS;
  uint4 v = make_uint4(0, 0, 0, 0);
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) v = add4(v, make_record(a, base + k));
  uint4 tot;
  block_scan_incl(v, &tot);
  if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

// pass 2: exclusive scan of the block totals by ONE block (nb <= a few thousand)
__global__ void __launch_bounds__(SCAN_THREADS) hq_compact_scan_sums(uint4 *block_sums, uint32_t nb) {
  uint4 carry = make_uint4(0, 0, 0, 0);
  for (uint32_t b0 = 0; b0 < nb; b0 += SCAN_THREADS) {
    const uint32_t i = b0 + threadIdx.x;
    const uint4 v = i < nb ? block_sums[i] : make_uint4(0, 0, 0, 0);
    uint4 tot;
    const uint4 inc = block_scan_incl(v, &tot);
    if (i < nb) block_sums[i] = make_uint4(carry.x + inc.x - v.x, carry.y + inc.y - v.y, carry.z + inc.z - v.z,
                                           carry.w + inc.w - v.w);
    carry = add4(carry, tot);
  }
}

// pass 3: exclusive scan written to scan[0..n]
__global__ void __launch_bounds__(SCAN_THREADS) hq_compact_scan(hq_kargs a, const uint4 *block_sums, uint4 *scan,
                                                                uint32_t *totals) {
  const uint32_t base = blockIdx.x * (SCAN_THREADS * SCAN_ITEMS) + threadIdx.x * SCAN_ITEMS;
  uint4 r[SCAN_ITEMS];
  uint4 v = make_uint4(0, 0, 0, 0);
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    r[k] = make_record(a, base + k);
    v = add4(v, r[k]);
  }
  uint4 tot;
  const uint4 inc = block_scan_incl(v, &tot);
  const uint4 off = block_sums[blockIdx.x];
  uint4 run = make_uint4(off.x + inc.x - v.x, off.y + inc.y - v.y, off.z + inc.z - v.z, off.w + inc.w - v.w);
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    const uint32_t i = base + k;
    if (i <= a.n_reads) {
      scan[i] = run;
      if (i == a.n_reads) {
        totals[0] = run.z;
        totals[1] = run.w;
        totals[2] = run.x;
        totals[3] = run.y;
      }
    }
    run = add4(run, r[k]);
  }
}

// pass 4: copy.  G lanes per read; a lane moves DESTINATION-aligned 16-byte chunks.  The source of such a chunk
// starts at an arbitrary byte, always with the same phase inside a read, so every lane loads ONE aligned 16-byte
// source chunk per array, takes the following one from its neighbour lane by shuffle, and funnel-shifts the pair
// into the 16 bytes it stores.  The partial chunks at the two ends of a read are moved byte by byte.
__device__ __forceinline__ uint4 shift_pair(const uint4 &A, const uint4 &B, uint32_t sh) {
  // bytes [sh, sh + 16) of the 32 bytes A:B
  const uint32_t bs = (sh & 3u) * 8u;
  uint32_t w0, w1, w2, w3, w4;
  switch (sh >> 2) {
    case 0: w0 = A.x; w1 = A.y; w2 = A.z; w3 = A.w; w4 = B.x; break;
    case 1: w0 = A.y; w1 = A.z; w2 = A.w; w3 = B.x; w4 = B.y; break;
    case 2: w0 = A.z; w1 = A.w; w2 = B.x; w3 = B.y; w4 = B.z; break;
    default: w0 = A.w; w1 = B.x; w2 = B.y; w3 = B.z; w4 = B.w; break;
  }
  return make_uint4(__funnelshift_r(w0, w1, bs), __funnelshift_r(w1, w2, bs), __funnelshift_r(w2, w3, bs),
                    __funnelshift_r(w3, w4, bs));
}

__device__ __forceinline__ uint4 shfl_down4(const uint4 &v, unsigned gmask) {
  return make_uint4(__shfl_down_sync(gmask, v.x, 1), __shfl_down_sync(gmask, v.y, 1), __shfl_down_sync(gmask, v.z, 1),
                    __shfl_down_sync(gmask, v.w, 1));
}

template <int G>
__global__ void __launch_bounds__(256) hq_compact_copy(hq_kargs a, hq_compact_bufs b) {
  const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t lig = gtid % G, grp = gtid / G, ngrp = (gridDim.x * blockDim.x) / G;
  const unsigned gmask = (G == 32 ? 0xFFFFFFFFu : ((1u << G) - 1u)) << ((threadIdx.x & 31u) / G * G);
  const uint32_t n = a.n_reads;
  const uint4 tot = b.scan[n];
  for (uint32_t i = grp; i < n; i += ngrp) {  // group-uniform loop: shuffles below name the group's lanes only
    const uint4 sc = b.scan[i];
    const bool pass = a.mask ? (a.mask[i] != 0) : true;
    const uint32_t st = a.trim ? a.trim[i].start : 0u;
    const uint32_t ln = a.trim ? a.trim[i].len : (a.idx[i + 1] - a.idx[i]);
    const uint32_t dst = pass ? sc.x : tot.x + sc.y;
    const uint32_t slot = pass ? sc.z : tot.z + sc.w;
    if (lig == 0) {
      b.out_src[slot] = i;
      b.out_idx[slot] = dst;
    }
    if (ln == 0) continue;
    const uint32_t so = a.idx[i] + st;
    const uint32_t db = dst & ~15u, de = dst + ln;
    const uint32_t nch = (de - db + 15u) >> 4;
    const uint32_t sh = (so - dst) & 15u;  // byte phase of the source inside its aligned chunk, same for all chunks
    for (uint32_t k0 = 0; k0 < nch; k0 += G) {
      const uint32_t k = k0 + lig;
      const uint32_t dc = db + 16u * k;
      // valid bytes [vlo, vhi) of this destination chunk
      const uint32_t vlo = (k < nch && dc < dst) ? dst - dc : 0u;
      const uint32_t vhi = k >= nch ? 0u : (de - dc >= 16u ? 16u : de - dc);
      // source byte that lands on byte 0 of the chunk; for the head chunk it may lie before the read (and, for
      // the first read of the batch, before the array: then the A half is not loaded, none of its bytes is used)
      const uint32_t sa = so + (dc - dst);
      const bool under = dc < dst && so < dst - dc;
      uint4 As = make_uint4(0, 0, 0, 0), Aq = As;
      if (k <= nch && !under) {  // chunk nch only feeds the B half of chunk nch-1
        As = __ldg(reinterpret_cast<const uint4 *>(a.seq + (sa & ~15u)));
        Aq = __ldg(reinterpret_cast<const uint4 *>(a.qual + (sa & ~15u)));
      }
      uint4 Bs = shfl_down4(As, gmask), Bq = shfl_down4(Aq, gmask);
      if (lig == G - 1 && k < nch && sh) {  // the neighbour belongs to the next round
        Bs = __ldg(reinterpret_cast<const uint4 *>(a.seq + (sa & ~15u) + 16u));
        Bq = __ldg(reinterpret_cast<const uint4 *>(a.qual + (sa & ~15u) + 16u));
      }
      if (vhi > vlo) {
        const uint4 vs = shift_pair(As, Bs, sh), vq = shift_pair(Aq, Bq, sh);
        if (vhi - vlo == 16u) {
          *reinterpret_cast<uint4 *>(b.out_seq + dc) = vs;
          *reinterpret_cast<uint4 *>(b.out_qual + dc) = vq;
        } else {  // partial chunk at an end of the read: whole words where possible, bytes for the rest
          const uint32_t ws[4] = {vs.x, vs.y, vs.z, vs.w}, wq[4] = {vq.x, vq.y, vq.z, vq.w};
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            const uint32_t w0 = 4u * w;
            if (vlo <= w0 && vhi >= w0 + 4u) {
              *reinterpret_cast<uint32_t *>(b.out_seq + dc + w0) = ws[w];
              *reinterpret_cast<uint32_t *>(b.out_qual + dc + w0) = wq[w];
            } else if (vlo < w0 + 4u && vhi > w0) {
#pragma unroll
              for (int t = 0; t < 4; ++t)
                if (w0 + t >= vlo && w0 + t < vhi) {
                  b.out_seq[dc + w0 + t] = (uint8_t)(ws[w] >> (8 * t));
                  b.out_qual[dc + w0 + t] = (uint8_t)(wq[w] >> (8 * t));
                }
            }
          }
        }
      }
    }
  }
  if (gtid == 0) b.out_idx[n] = tot.x + tot.y;
}


// pass 4, tiled form (reads up to CT_SPAN bytes): DESTINATION-chunk ownership.  A CTA takes a tile of consecutive
// input reads.  Their source bytes are one contiguous span of the batch, copied into shared memory with aligned
// 16-byte loads; their output is two contiguous spans (the passed reads of the tile, and its failed reads behind
// all passed reads of the batch).  Every thread then builds whole aligned 16-byte chunks of the OUTPUT: it finds the
// read that holds the chunk's first byte (binary search over the tile's scan values in shared memory), pulls the
// bytes from the staged source at that read's phase, and continues into the following read(s) when the chunk
// straddles a boundary.  Global stores are full aligned 16-byte chunks except at the two ends of each span.
constexpr uint32_t CT_RPT = 128;      // reads per tile at most
constexpr uint32_t CT_SPAN = 22000;   // staged source bytes per array at most (keeps the CTA under 48 KB)
constexpr uint32_t CT_PAD = 16;       // staged data starts 16 bytes in: a chunk may begin before its read

struct ct_hdr {
  uint32_t key[2][CT_RPT + 1];  // output offset of read r in the passed / failed stream (tile-local exclusive scan)
  uint32_t soff[CT_RPT + 1];    // staged offset of the first (trimmed) byte of read r
};

// bytes [sh, sh + 16) of the 32 bytes A:B, sh in [0, 16), branch-free
__device__ __forceinline__ uint4 shift_sel(const uint4 &A, const uint4 &B, uint32_t sh) {
  const bool o1 = (sh & 4u) != 0u, o2 = (sh & 8u) != 0u;
  const uint32_t t0 = o1 ? A.y : A.x, t1 = o1 ? A.z : A.y, t2 = o1 ? A.w : A.z, t3 = o1 ? B.x : A.w;
  const uint32_t t4 = o1 ? B.y : B.x, t5 = o1 ? B.z : B.y, t6 = o1 ? B.w : B.z;
  const uint32_t u0 = o2 ? t2 : t0, u1 = o2 ? t3 : t1, u2 = o2 ? t4 : t2, u3 = o2 ? t5 : t3, u4 = o2 ? t6 : t4;
  const uint32_t bs = (sh & 3u) * 8u;
  return make_uint4(__funnelshift_r(u0, u1, bs), __funnelshift_r(u1, u2, bs), __funnelshift_r(u2, u3, bs),
                    __funnelshift_r(u3, u4, bs));
}

// 0xFF in every byte j of the 16 with lo <= j < hi, as four words
__device__ __forceinline__ uint4 range_mask16(uint32_t lo, uint32_t hi) {
  const uint32_t bits = ((1u << hi) - 1u) & ~((1u << lo) - 1u);  // hi <= 16
  auto ex = [](uint32_t nib) { return (((nib & 0xFu) * 0x00204081u) & 0x01010101u) * 0xFFu; };
  return make_uint4(ex(bits), ex(bits >> 4), ex(bits >> 8), ex(bits >> 12));
}

__host__ __device__ constexpr uint32_t ct_rmap_bytes(uint32_t cap) { return ((cap >> 4) + 8u + 15u) & ~15u; }

__device__ __forceinline__ void fz_cp16(void *dst_smem, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_smem)), "l"(src)
               : "memory");
}

// Builds the aligned 16-byte chunks of a tile's two output spans (stream 0 = passed, 1 = failed) from the staged
// source bytes and stores them to oseq[stream] / oqual[stream]; h.key[stream][r] is the offset of read r inside that
// stream's buffer, rmap the chunk -> read map (failed span's entries mo_f behind the passed span's).
__device__ __forceinline__ void ct_emit(const ct_hdr &h, const uint8_t *rmap, uint32_t mo_f, const uint8_t *sseq,
                                        const uint8_t *squal, uint32_t nt, uint8_t *const oseq[2], uint8_t *const oqual[2]) {
  const uint32_t tid = threadIdx.x;
  // ONE loop over the chunks of both spans (passed span first): the failed span of a tile is a few chunks, and a loop of
  // its own cost every warp a round with one or two active lanes (19 of 32 lanes were active on average)
  const uint32_t D00 = h.key[0][0], D01 = h.key[0][nt], D10 = h.key[1][0], D11 = h.key[1][nt];
  const uint32_t c0 = D00 >> 4, n0 = D01 > D00 ? ((D01 + 15u) >> 4) - c0 : 0u;
  const uint32_t c1 = D10 >> 4, n1 = D11 > D10 ? ((D11 + 15u) >> 4) - c1 : 0u;
  {
    for (uint32_t v = tid; v < n0 + n1; v += blockDim.x) {
      const bool s1 = v >= n0;
      const uint32_t *key = &h.key[0][0] + (s1 ? CT_RPT + 1u : 0u);  // offsets, not pointer selects: stays an LDS
      const uint32_t D0 = s1 ? D10 : D00, D1 = s1 ? D11 : D01;
      const uint32_t c = s1 ? c1 + (v - n0) : c0 + v, c_hi = (D1 + 15u) >> 4;
      const uint8_t *mp = rmap + (int32_t)(s1 ? mo_f - c1 : 0u - c0);
      // pointers selected into registers (as two-entry arrays indexed by a loop variable they lived in local memory:
      // a dependent LDL in front of every store, 18 % of the stall samples of the gather kernel)
      uint8_t *const os = s1 ? oseq[1] : oseq[0], *const oq = s1 ? oqual[1] : oqual[0];
      const uint32_t dc = c << 4, d0 = max(dc, D0), d1 = min(dc + 16u, D1);
      const uint32_t r = mp[c];
      const uint32_t kend = key[r + 1];
      // first source: read r; second source: the next non-empty read of this stream when the chunk straddles
      const uint32_t e1 = min(d1, kend);
      const uint32_t x1 = h.soff[r] + (dc - key[r]);  // staged byte that lands on byte 0 of the chunk (mod 2^32)
      uint32_t x2 = x1, e2 = d1, r2 = r;
      if (e1 < d1) {
        r2 = c + 1u < c_hi ? (uint32_t)mp[c + 1u] : r + 1u;  // usually the owner of the next chunk
        uint32_t k2 = key[r2 + 1];
        if (__builtin_expect(key[r2] != kend || k2 <= kend, 0)) {  // a read shorter than the rest of this chunk
          r2 = r + 1u;
          k2 = key[r2 + 1];
          while (k2 <= e1) k2 = key[++r2 + 1];
        }
        x2 = h.soff[r2] + (dc - key[r2]);
        e2 = min(d1, k2);
      }
      // Reads that follow each other in the input AND in this stream untrimmed keep their distance: then the chunk's
      // second source is the continuation of the first (x2 == x1) and one shifted 32-byte window gives all 16 bytes.
      uint4 vs, vq;
      {
        const uint32_t xa = x1 & ~15u;
        const uint4 As = *reinterpret_cast<const uint4 *>(sseq + xa), Bs = *reinterpret_cast<const uint4 *>(sseq + xa + 16u);
        const uint4 Aq = *reinterpret_cast<const uint4 *>(squal + xa), Bq = *reinterpret_cast<const uint4 *>(squal + xa + 16u);
        vs = shift_sel(As, Bs, x1 & 15u);
        vq = shift_sel(Aq, Bq, x1 & 15u);
      }
      if (__builtin_expect(x2 != x1, 0)) {
        const uint32_t xb = x2 & ~15u;
        const uint4 Cs = *reinterpret_cast<const uint4 *>(sseq + xb), Ds = *reinterpret_cast<const uint4 *>(sseq + xb + 16u);
        const uint4 Cq = *reinterpret_cast<const uint4 *>(squal + xb), Dq = *reinterpret_cast<const uint4 *>(squal + xb + 16u);
        const uint4 ts = shift_sel(Cs, Ds, x2 & 15u), tq = shift_sel(Cq, Dq, x2 & 15u);
        const uint4 m = range_mask16(0u, e1 - dc);  // bytes of the first source
        vs = make_uint4((vs.x & m.x) | (ts.x & ~m.x), (vs.y & m.y) | (ts.y & ~m.y), (vs.z & m.z) | (ts.z & ~m.z),
                        (vs.w & m.w) | (ts.w & ~m.w));
        vq = make_uint4((vq.x & m.x) | (tq.x & ~m.x), (vq.y & m.y) | (tq.y & ~m.y), (vq.z & m.z) | (tq.z & ~m.z),
                        (vq.w & m.w) | (tq.w & ~m.w));
      }
      if (__builtin_expect(e2 < d1, 0)) {  // three or more reads in one chunk (reads shorter than 16 bytes)
        uint32_t d = e2, rr = r2 + 1u;
        while (d < d1) {
          uint32_t k3 = key[rr + 1];
          while (k3 <= d) k3 = key[++rr + 1];
          const uint32_t e = min(d1, k3);
          const uint32_t x = h.soff[rr] + (dc - key[rr]);
          const uint32_t xa = x & ~15u;
          const uint4 As = *reinterpret_cast<const uint4 *>(sseq + xa), Bs = *reinterpret_cast<const uint4 *>(sseq + xa + 16u);
          const uint4 Aq = *reinterpret_cast<const uint4 *>(squal + xa), Bq = *reinterpret_cast<const uint4 *>(squal + xa + 16u);
          const uint4 ps = shift_sel(As, Bs, x & 15u), pq = shift_sel(Aq, Bq, x & 15u);
          const uint4 m = range_mask16(d - dc, e - dc);
          vs.x = (vs.x & ~m.x) | (ps.x & m.x); vs.y = (vs.y & ~m.y) | (ps.y & m.y);
          vs.z = (vs.z & ~m.z) | (ps.z & m.z); vs.w = (vs.w & ~m.w) | (ps.w & m.w);
          vq.x = (vq.x & ~m.x) | (pq.x & m.x); vq.y = (vq.y & ~m.y) | (pq.y & m.y);
          vq.z = (vq.z & ~m.z) | (pq.z & m.z); vq.w = (vq.w & ~m.w) | (pq.w & m.w);
          d = e;
          ++rr;
        }
      }
      if (d1 - d0 == 16u) {
        *reinterpret_cast<uint4 *>(os + dc) = vs;
        *reinterpret_cast<uint4 *>(oq + dc) = vq;
      } else {  // first / last chunk of the span: its other bytes belong to a neighbouring tile
        const uint32_t ws[4] = {vs.x, vs.y, vs.z, vs.w}, wq[4] = {vq.x, vq.y, vq.z, vq.w};
        const uint32_t vlo = d0 - dc, vhi = d1 - dc;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
          const uint32_t w0 = 4u * w;
          if (vlo <= w0 && vhi >= w0 + 4u) {
            *reinterpret_cast<uint32_t *>(os + dc + w0) = ws[w];
            *reinterpret_cast<uint32_t *>(oq + dc + w0) = wq[w];
          } else if (vlo < w0 + 4u && vhi > w0) {
#pragma unroll
            for (int t = 0; t < 4; ++t)
              if (w0 + t >= vlo && w0 + t < vhi) {
                os[dc + w0 + t] = (uint8_t)(ws[w] >> (8 * t));
                oq[dc + w0 + t] = (uint8_t)(wq[w] >> (8 * t));
              }
          }
        }
      }
    }
  }
}

__global__ void __launch_bounds__(256) hq_compact_copy_tile(hq_kargs a, hq_compact_bufs b, uint32_t rpt, uint32_t cap) {
  extern __shared__ __align__(16) uint8_t ct_raw[];
  ct_hdr &h = *reinterpret_cast<ct_hdr *>(ct_raw);
  constexpr uint32_t HDR = (sizeof(ct_hdr) + 15u) & ~15u;
  uint8_t *rmap = ct_raw + HDR;  // per output chunk of the tile (passed span, then failed span): the read holding its first byte
  uint8_t *sseq = rmap + ct_rmap_bytes(cap), *squal = sseq + cap;
  const uint32_t tid = threadIdx.x, n = a.n_reads;
  const uint4 tot = b.scan[n];
  for (uint32_t i0 = blockIdx.x * rpt; i0 < n; i0 += gridDim.x * rpt) {
    const uint32_t i1 = min(n, i0 + rpt), nt = i1 - i0;
    __syncthreads();  // the previous tile's staged bytes are no longer read
    const uint32_t S0a = a.idx[i0] & ~15u, S1 = a.idx[i1];
    const uint4 sc0 = b.scan[i0], sc1 = b.scan[i1];
    const uint32_t mo_f = ((sc1.x + 15u) >> 4) - (sc0.x >> 4) + 1u;  // the failed span's chunks follow the passed span's
    for (uint32_t r = tid; r <= nt; r += 256u) {
      const uint32_t i = i0 + r;
      const uint4 sc = b.scan[i];
      h.key[0][r] = sc.x;
      h.key[1][r] = tot.x + sc.y;
      if (r < nt) {
        const bool pass = a.mask ? (a.mask[i] != 0) : true;
        const uint32_t st = a.trim ? a.trim[i].start : 0u;
        const uint32_t slot = pass ? sc.z : tot.z + sc.w;
        b.out_src[slot] = i;
        b.out_idx[slot] = pass ? sc.x : tot.x + sc.y;
        h.soff[r] = a.idx[i] + st - S0a + CT_PAD;
        // chunk -> read map: this read owns the chunks whose first in-span byte it holds
        const uint4 scn = b.scan[i + 1u];
        const uint32_t k0 = pass ? sc.x : tot.x + sc.y, k1 = pass ? scn.x : tot.x + scn.y;
        const uint32_t D0s = pass ? sc0.x : tot.x + sc0.y;
        if (k1 > k0) {
          const uint32_t cb = D0s >> 4;
          const uint32_t cf = k0 == D0s ? cb : (k0 + 15u) >> 4, cl = (k1 - 1u) >> 4;
          uint8_t *mp = rmap + (pass ? 0u : mo_f) - cb;
          for (uint32_t c = cf; c <= cl; ++c) mp[c] = (uint8_t)r;
        }
      }
    }
    const uint32_t nb16 = (S1 - S0a + 15u) >> 4;
    for (uint32_t c = tid; c < nb16; c += 256u) {
      reinterpret_cast<uint4 *>(sseq + CT_PAD)[c] = __ldg(reinterpret_cast<const uint4 *>(a.seq + S0a) + c);
      reinterpret_cast<uint4 *>(squal + CT_PAD)[c] = __ldg(reinterpret_cast<const uint4 *>(a.qual + S0a) + c);
    }
    __syncthreads();
    uint8_t *const oseq[2] = {b.out_seq, b.out_seq}, *const oqual[2] = {b.out_qual, b.out_qual};
    ct_emit(h, rmap, mo_f, sseq, squal, nt, oseq, oqual);
  }
  if (blockIdx.x == 0 && tid == 0) b.out_idx[n] = tot.x + tot.y;
}

// The same with two tile buffers per CTA and NO synchronous global load on a tile's path.  The single-buffer kernel
// pays, per tile, three dependent round trips to global memory (tile offsets -> scan / mask / trim records -> staged
// span), ~2 us each under load, before it can build a chunk: 12.7 stall cycles per issued instruction on global loads,
// issue slots 32 % busy (profiles/compact_r2_summary.md).  Here everything tile t + 1 needs -- its scan records, offsets,
// mask bytes, trim records and its source span -- is sent to shared memory with cp.async while tile t is being built,
// and the two offsets that size the copies of tile t + 2 are loaded into registers one iteration ahead.
struct ct_raw_t {
  uint4 scan[CT_RPT + 2];
  uint32_t idx[CT_RPT + 4];
  hpgq_trim_t trim[CT_RPT];
  uint32_t mask[CT_RPT / 4 + 2];  // the aligned words that hold the tile's mask bytes
};

__device__ __forceinline__ void ct_cp(void *dst_smem, const void *src, int bytes) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst_smem);
  if (bytes == 4) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(src) : "memory");
  else asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(src) : "memory");
}

__global__ void __launch_bounds__(256) hq_compact_copy_tile2(hq_kargs a, hq_compact_bufs b, uint32_t rpt, uint32_t cap) {
  extern __shared__ __align__(16) uint8_t ct_raw[];
  constexpr uint32_t HDR = (sizeof(ct_hdr) + 15u) & ~15u, RAW = (sizeof(ct_raw_t) + 15u) & ~15u;
  const uint32_t bufsz = RAW + ct_rmap_bytes(cap) + 2u * cap;  // a multiple of 16; the header is shared by both buffers
  ct_hdr &h = *reinterpret_cast<ct_hdr *>(ct_raw);
  const uint32_t tid = threadIdx.x, n = a.n_reads;
  const uint4 tot = b.scan[n];
  // copies of tile [i0, i1) into buffer k; s0 / s1 = idx[i0] / idx[i1]
  auto send = [&](uint32_t i0, uint32_t k, uint32_t s0, uint32_t s1) {
    uint8_t *base = ct_raw + HDR + k * bufsz;
    ct_raw_t &w = *reinterpret_cast<ct_raw_t *>(base);
    uint8_t *sseq = base + RAW + ct_rmap_bytes(cap), *squal = sseq + cap;
    const uint32_t nt = min(n, i0 + rpt) - i0;
    if (tid <= nt) {
      fz_cp16(&w.scan[tid], &b.scan[i0 + tid]);
      ct_cp(&w.idx[tid], &a.idx[i0 + tid], 4);
      if (tid < nt && a.trim) ct_cp(&w.trim[tid], &a.trim[i0 + tid], 8);
    }
    if (a.mask && tid >= 128u) {
      const uint8_t *m0 = a.mask + i0;
      const uint8_t *ma = m0 - ((uintptr_t)m0 & 3u);
      const uint32_t words = ((uint32_t)(m0 - ma) + nt + 3u) >> 2, t = tid - 128u;
      if (t < words) ct_cp(&w.mask[t], ma + 4u * t, 4);
    }
    const uint32_t S0a = s0 & ~15u, nb16 = (s1 - S0a + 15u) >> 4;
    for (uint32_t c = tid; c < nb16; c += 256u) {
      fz_cp16(sseq + CT_PAD + 16u * c, a.seq + S0a + 16u * c);
      fz_cp16(squal + CT_PAD + 16u * c, a.qual + S0a + 16u * c);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  const uint32_t step = gridDim.x * rpt;
  uint32_t i0 = blockIdx.x * rpt, k = 0;
  uint32_t n0 = 0, n1 = 0;  // idx[i0 + step], idx[min(n, i0 + step + rpt)]: loaded one iteration before they are used
  if (i0 < n) {
    send(i0, 0, a.idx[i0], a.idx[min(n, i0 + rpt)]);
    if (i0 + step < n) {
      n0 = a.idx[i0 + step];
      n1 = a.idx[min(n, i0 + step + rpt)];
    }
  }
  for (; i0 < n; i0 += step, k ^= 1u) {
    const bool more = i0 + step < n;
    if (more) {
      send(i0 + step, k ^ 1u, n0, n1);  // its buffer was released by the barrier that closed tile t - 1
      if (i0 + 2u * step < n) {
        n0 = a.idx[i0 + 2u * step];
        n1 = a.idx[min(n, i0 + 2u * step + rpt)];
      }
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();  // tile t's records and bytes have landed
    uint8_t *base = ct_raw + HDR + k * bufsz;
    const ct_raw_t &w = *reinterpret_cast<const ct_raw_t *>(base);
    uint8_t *rmap = base + RAW;
    const uint8_t *sseq = rmap + ct_rmap_bytes(cap), *squal = sseq + cap;
    const uint32_t nt = min(n, i0 + rpt) - i0;
    const uint4 sc0 = w.scan[0], sc1 = w.scan[nt];
    const uint32_t mo_f = ((sc1.x + 15u) >> 4) - (sc0.x >> 4) + 1u;  // the failed span's chunks follow the passed span's
    if (tid <= nt) {
      const uint32_t r = tid, i = i0 + r;
      const uint4 sc = w.scan[r];
      h.key[0][r] = sc.x;
      h.key[1][r] = tot.x + sc.y;
      if (r < nt) {
        const uint32_t mb = (uint32_t)((uintptr_t)(a.mask + i0) & 3u) + r;
        const bool pass = a.mask ? ((w.mask[mb >> 2] >> (8u * (mb & 3u))) & 0xFFu) != 0u : true;
        const uint32_t st = a.trim ? w.trim[r].start : 0u;
        const uint32_t slot = pass ? sc.z : tot.z + sc.w;
        b.out_src[slot] = i;
        b.out_idx[slot] = pass ? sc.x : tot.x + sc.y;
        h.soff[r] = w.idx[r] + st - (w.idx[0] & ~15u) + CT_PAD;
        const uint4 scn = w.scan[r + 1u];
        const uint32_t k0 = pass ? sc.x : tot.x + sc.y, k1 = pass ? scn.x : tot.x + scn.y;
        const uint32_t D0s = pass ? sc0.x : tot.x + sc0.y;
        if (k1 > k0) {
          const uint32_t cb = D0s >> 4;
          const uint32_t cf = k0 == D0s ? cb : (k0 + 15u) >> 4, cl = (k1 - 1u) >> 4;
          uint8_t *mp = rmap + (pass ? 0u : mo_f) - cb;
          for (uint32_t c = cf; c <= cl; ++c) mp[c] = (uint8_t)r;
        }
      }
    }
    __syncthreads();  // header and chunk map of tile t
    uint8_t *const oseq[2] = {b.out_seq, b.out_seq}, *const oqual[2] = {b.out_qual, b.out_qual};
    ct_emit(h, rmap, mo_f, sseq, squal, nt, oseq, oqual);
    __syncthreads();  // tile t's buffer and the header are free again
  }
  if (blockIdx.x == 0 && tid == 0) b.out_idx[n] = tot.x + tot.y;
}

// ---------------------------------------------------------------- edit + filter + compaction in ONE pass over HBM
// fastq_edit_worker / fastq_filter_worker (src/edit_fastq.c:142-178, src/filter_fastq.c:134-155) and the consumers'
// fastq_fwrite of the passed / failed lists (src/edit_fastq.c:184-206, src/filter_fastq.c:161-174) for one batch.
// A CTA draws tiles of <= 128 consecutive reads in input order (ticket counter).  Per tile: (A) the tile's source span
// is staged in shared memory -- the only time the batch is read from HBM; (B) lane groups take the trim and filter
// decisions from the staged bytes (the register-resident code of hpgq_regs.cuh); (C) one warp scans the tile's
// (passed bytes, failed bytes, passed reads, failed reads); (D) a decoupled look-back over the preceding tiles'
// published aggregates turns that into the tile's place in the output; (E, F) the chunks of the passed span go
// straight to their final place, those of the failed span to a side buffer, because the failed reads follow ALL
// passed reads of the batch and that total is only known at the end: hq_fz_finish moves the side buffer (a few
// per cent of the bytes) behind the passed reads.  DRAM traffic: 2 L + 4 in, 2 L' + 17 out per read.
// Look-back records.  A tile publishes its own totals and, once it knows them, the totals of everything up to and
// including itself, each as four 64-bit words {value, 1}: a word is written once and carries its own validity, so
// neither side needs a memory fence (a __threadfence here waits for the CTA's output stores: 40 % of the kernel).
struct fz_rec {
  unsigned long long w[4];
};
__device__ __forceinline__ void fz_publish(fz_rec *r, uint4 v) {
  volatile unsigned long long *p = r->w;
  p[0] = (1ull << 32) | v.x;
  p[1] = (1ull << 32) | v.y;
  p[2] = (1ull << 32) | v.z;
  p[3] = (1ull << 32) | v.w;
}
__device__ __forceinline__ bool fz_read(const fz_rec *r, uint4 *v) {
  const volatile unsigned long long *p = r->w;
  const unsigned long long a0 = p[0], a1 = p[1], a2 = p[2], a3 = p[3];
  *v = make_uint4((uint32_t)a0, (uint32_t)a1, (uint32_t)a2, (uint32_t)a3);
  return ((a0 & a1 & a2 & a3) >> 32) != 0ull;
}
struct fz_bufs {
  fz_rec *agg, *incl;       // agg: a tile's own totals; incl: [0] the batch's totals, then the level-2 and level-3 records
  uint32_t *status;         // status[0] = ticket counter
  uint8_t *fail_seq, *fail_qual;
  uint32_t *fail_idx, *fail_src;
  uint32_t n_tiles, rpt, cap;
};

struct fz_hdr {
  ct_hdr ct;
  uint32_t sidx[CT_RPT + 1];   // offsets of the tile's reads
  uint32_t start[CT_RPT], len[CT_RPT];
  uint8_t pass[CT_RPT];
  uint4 prefix;                // totals of all preceding tiles
  uint32_t tile;
};

// Trim + filter decision of one read by the G lanes of its group, the read held in registers (32 bytes per lane):
// takes every option (filter windows, out-of-quality count).  o = batch offset of the read, L = its length; the
// staged copy of batch byte x lies at squal / sseq [x - S0a + CT_PAD].
struct fz_dec {
  uint32_t start, len;
  bool pass, edited;
};
template <int G, bool OOQ>
__device__ __forceinline__ fz_dec fz_decide_regs(const hq_kargs &a, const uint8_t *sseq, const uint8_t *squal, uint32_t S0a,
                                             uint32_t o, uint32_t L, uint32_t lig, bool do_edit, bool do_filter,
                                             bool zero_is_ooq) {
  bool edited_any = false;
  rg_read d;
  d.c = (o & ~15u) + RG_LB * lig;
#pragma unroll
  for (int k = 0; k < RG_CPL; ++k) {
    uint4 q4 = make_uint4(0, 0, 0, 0), s4 = q4;
    if (d.c + 16u * k < o + L) {
      const uint32_t x = d.c + 16u * k - S0a + CT_PAD;
      q4 = *reinterpret_cast<const uint4 *>(squal + x);
      s4 = *reinterpret_cast<const uint4 *>(sseq + x);
    }
    d.q[4 * k + 0] = q4.x; d.q[4 * k + 1] = q4.y; d.q[4 * k + 2] = q4.z; d.q[4 * k + 3] = q4.w;
    d.s[4 * k + 0] = s4.x; d.s[4 * k + 1] = s4.y; d.s[4 * k + 2] = s4.z; d.s[4 * k + 3] = s4.w;
  }
  uint32_t start = 0, len = L;
  if (do_edit) {  // fastq_edit: a window whose mean quality is out of range is cut off
    bool edited = false;
    if (a.e_ll > 0) {
      const uint32_t w = min((uint32_t)a.e_ll, len);
      const uint32_t sw = group_sum<G>(rg_window_lane(d, o, o + w));
      if (w && !mean_ok32(sw, w, a.e_lq_lo, a.e_lq_hi)) {
        start += w;
        len -= w;
        edited = true;
      }
    }
    if (a.e_rl > 0) {
      const uint32_t w = min((uint32_t)a.e_rl, len);
      const uint32_t e = o + start + len;
      const uint32_t sw = group_sum<G>(rg_window_lane(d, e - w, e));
      if (w && !mean_ok32(sw, w, a.e_rq_lo, a.e_rq_hi)) {
        len -= w;
        edited = true;
      }
    }
    edited_any = edited;
  }
  rg_clip(d, o + start, o + start + len);
  const uint32_t b0 = o + start, b1 = b0 + len;
  bool pass = true;
  if (do_filter) {
    uint32_t S = 0, fN = 0, fO = 0, hib = 0;
#pragma unroll
    for (int j = 0; j < RG_W; ++j) {
      const uint32_t qw = d.q[j], sw = d.s[j];
      S = __dp4a(qw, 0x01010101u, S);
      fN += rg_zero_flags(sw ^ 0x4E4E4E4Eu) >> 7;
      if (OOQ) {
        fO += (~((qw + a.c_lo) & ~(qw + a.c_hi)) & 0x80808080u) >> 7;
        hib |= qw;
      }
    }
    uint32_t nN = __dp4a(fN, 0x01010101u, 0u), ooq = __dp4a(fO, 0x01010101u, 0u);
    if (OOQ && zero_is_ooq) {  // remove the lane's zeroed bytes from the out-of-range count
      const uint32_t lo = max(b0, d.c), hi = min(b1, d.c + RG_LB);
      ooq -= RG_LB - (hi > lo ? hi - lo : 0u);
    }
    if (OOQ && (hib & 0x80808080u)) {  // a quality byte >= 128: exact per-byte count for this lane
      ooq = 0;
      for (int j = 0; j < RG_W; ++j)
        for (int i = 0; i < 4; ++i) {
          const uint32_t pos = d.c + 4u * j + i;
          const int32_t qq = (int32_t)((d.q[j] >> (8 * i)) & 0xFFu);
          if (pos >= b0 && pos < b1) ooq += (qq < a.q_lo) | (qq > a.q_hi);
        }
    }
    uint32_t p1 = group_sum<G>(S | (nN << 20));
    if (OOQ) ooq = group_sum<G>(ooq);
    S = p1 & 0xFFFFFu;
    nN = p1 >> 20;
    pass = (int32_t)len >= a.min_len && (int32_t)len <= a.max_len && mean_ok32(S, len, a.q_lo, a.q_hi) &&
           (!OOQ || (int32_t)ooq <= a.max_ooq) && (int32_t)nN <= a.max_n;
    if (a.f_ll > 0) {
      const uint32_t wl = min((uint32_t)a.f_ll, len);
      const uint32_t SL = group_sum<G>(rg_window_lane(d, b0, b0 + wl));
      if (wl) pass = pass && mean_ok32(SL, wl, a.f_lq_lo, a.f_lq_hi);
    }
    if (a.f_rl > 0) {
      const uint32_t wr = min((uint32_t)a.f_rl, len);
      const uint32_t SR = group_sum<G>(rg_window_lane(d, b1 - wr, b1));
      if (wr) pass = pass && mean_ok32(SR, wr, a.f_rq_lo, a.f_rq_hi);
    }
  }
  return fz_dec{start, len, pass, edited_any};
}

// The same decision for the common case -- no filter windows, no out-of-quality limit, trim windows of <= 255 bases
// that do not overlap (L >= left + right window) -- in the word frame of the tiled statistics kernel: lane lig owns
// the words {lig + k G} of the read, byte masks of the read end and of the two trim windows are per-LENGTH constants
// held in registers, and every sum is a dp4a.  Returns false in `ok` for a read it does not take (the caller then
// runs fz_decide_regs for the warp).
template <int K, int G>
struct fz_len_consts {
  uint32_t L, wl, wr;
  uint32_t m[K], m01[K], ml[K], mr[K];  // read bytes (0xFF form, 0x01 form), left / right trim window (0x01 form)
};
template <int K, int G>
__device__ __forceinline__ void fz_len_update(const hq_kargs &a, fz_len_consts<K, G> &c, uint32_t L, uint32_t lig) {
  c.L = L;
  c.wl = a.e_ll > 0 ? min((uint32_t)a.e_ll, L) : 0u;
  c.wr = a.e_rl > 0 ? min((uint32_t)a.e_rl, L) : 0u;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const uint32_t p = 4u * ((uint32_t)k * G + lig);
    // three clamp + funnel-shift masks per word (hq_tail_mask_shf) instead of five compare / select ones: this block
    // runs twice per short read of an otherwise uniform batch, for the whole warp, and was 12 % of the decision
    // kernel's instructions on C3 (2 % short reads; capture dec_r2x)
    c.m[k] = hq_tail_mask_shf(p, L);
    c.m01[k] = c.m[k] & 0x01010101u;
    c.ml[k] = hq_tail_mask_shf(p, c.wl) & 0x01010101u;            // positions [0, wl)
    c.mr[k] = c.m01[k] & ~hq_tail_mask_shf(p, L - c.wr);          // positions [L - wr, L)
  }
}
template <int K, int G>
__device__ __forceinline__ fz_dec fz_decide_fast(const hq_kargs &a, const fz_len_consts<K, G> &c, const uint32_t *ps,
                                                 const uint32_t *pq, uint32_t bsel, bool do_edit, bool do_filter) {
  uint32_t S = 0, nN = 0, SL = 0, SR = 0, NL = 0, NR = 0;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const uint32_t sw = hq_prmt_raw(ps[k * G], ps[k * G + 1], bsel);
    const uint32_t q = hq_prmt_raw(pq[k * G], pq[k * G + 1], bsel) & c.m[k];
    const uint32_t z = (rg_zero_flags(sw ^ 0x4E4E4E4Eu) >> 7) & c.m01[k];  // 0x01 per 'N' of the read
    S = __dp4a(q, 0x01010101u, S);
    nN = __dp4a(z, 0x01010101u, nN);
    SL = __dp4a(q, c.ml[k], SL);
    SR = __dp4a(q, c.mr[k], SR);
    NL = __dp4a(z, c.ml[k], NL);
    NR = __dp4a(z, c.mr[k], NR);
  }
  // reads of at most 4 K G <= 640 bases, windows of at most 255: S < 2^18, counts < 2^10, window sums < 2^16
  const uint32_t p1 = group_sum<G>(S | (nN << 20));
  S = p1 & 0xFFFFFu;
  nN = p1 >> 20;
  uint32_t start = 0, len = c.L;
  bool edited = false;
  if (do_edit) {
    const uint32_t p2 = group_sum<G>(SL | (NL << 16)), p3 = group_sum<G>(SR | (NR << 16));
    if (c.wl && !mean_ok32(p2 & 0xFFFFu, c.wl, a.e_lq_lo, a.e_lq_hi)) {
      start = c.wl;
      len -= c.wl;
      S -= p2 & 0xFFFFu;
      nN -= p2 >> 16;
      edited = true;
    }
    if (c.wr && !mean_ok32(p3 & 0xFFFFu, c.wr, a.e_rq_lo, a.e_rq_hi)) {
      len -= c.wr;
      S -= p3 & 0xFFFFu;
      nN -= p3 >> 16;
      edited = true;
    }
  }
  bool pass = true;
  if (do_filter)
    pass = (int32_t)len >= a.min_len && (int32_t)len <= a.max_len && mean_ok32(S, len, a.q_lo, a.q_hi) && (int32_t)nN <= a.max_n;
  return fz_dec{start, len, pass, edited};
}

// FAST: fz_decide_fast takes the reads (K words per lane), fz_decide_regs only those it declines.
template <int K, int G, bool OOQ, bool FAST>
__global__ void __launch_bounds__(256, 4) hq_fz_kernel(hq_kargs a, hq_compact_bufs b, fz_bufs z) {
  extern __shared__ __align__(16) uint8_t fz_raw[];
  fz_hdr &h = *reinterpret_cast<fz_hdr *>(fz_raw);
  constexpr uint32_t HDR = (sizeof(fz_hdr) + 15u) & ~15u;
  uint8_t *rmap = fz_raw + HDR;
  uint8_t *sseq = rmap + ct_rmap_bytes(z.cap), *squal = sseq + z.cap;
  constexpr uint32_t NG = 32u / G;
  const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5, lig = lane % G, giw = lane / G;
  const uint32_t n = a.n_reads, rpt = z.rpt;
  const bool do_filter = a.work & HPGQ_WORK_FILTER, do_edit = a.work & HPGQ_WORK_EDIT;
  const bool zero_is_ooq = a.q_lo > 0 || a.q_hi < 0;
  uint32_t n_pass = 0, n_fail = 0, n_edit = 0;
  fz_len_consts<K, G> lc;
  lc.L = 0xFFFFFFFFu;
  lc.wl = lc.wr = 0u;
  for (;;) {
    __syncthreads();  // the previous tile's shared memory is no longer read
    if (tid == 0) h.tile = atomicAdd(&z.status[0], 1u);
    __syncthreads();
    const uint32_t tile = h.tile;
    if (tile >= z.n_tiles) break;
    const uint32_t i0 = tile * rpt, i1 = min(n, i0 + rpt), nt = i1 - i0;
    // ---- A: offsets and source span into shared memory
    for (uint32_t r = tid; r <= nt; r += 256u) h.sidx[r] = a.idx[i0 + r];
    const uint32_t S0a = a.idx[i0] & ~15u, S1 = a.idx[i1];
    const uint32_t nb16 = (S1 - S0a + 15u) >> 4;
    for (uint32_t c = tid; c < nb16; c += 256u) {  // 16-byte cp.async: every chunk of the tile in flight at once
      fz_cp16(sseq + CT_PAD + 16u * c, a.seq + S0a + 16u * c);
      fz_cp16(squal + CT_PAD + 16u * c, a.qual + S0a + 16u * c);
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();
    // ---- B: trim + filter decisions, G lanes per read, from the staged bytes
    for (uint32_t rb = warp * NG; rb < nt; rb += 8u * NG) {  // warp-uniform trip count (shuffles below)
      const uint32_t r = rb + giw;
      const bool live = r < nt;
      const uint32_t o = live ? h.sidx[r] : S0a, L = live ? h.sidx[r + 1] - o : 0u;
      fz_dec dec;
      bool slow = !FAST;
      if (FAST) {
        if (L != lc.L) fz_len_update<K, G>(a, lc, L, lig);
        slow = lc.wl + lc.wr > L;  // overlapping trim windows: the right one moves with the left trim
        const uint32_t so = o - S0a + CT_PAD;
        const uint32_t *ps = reinterpret_cast<const uint32_t *>(sseq) + (so >> 2) + lig;
        const uint32_t *pq = reinterpret_cast<const uint32_t *>(squal) + (so >> 2) + lig;
        dec = fz_decide_fast<K, G>(a, lc, ps, pq, 0x3210u + 0x1111u * (so & 3u), do_edit, do_filter);
      }
      if (__any_sync(0xFFFFFFFFu, slow)) {
        const fz_dec d2 = fz_decide_regs<G, OOQ>(a, sseq, squal, S0a, o, L, lig, do_edit, do_filter, zero_is_ooq);
        if (slow) dec = d2;
      }
      const uint32_t start = dec.start, len = dec.len;
      const bool pass = dec.pass;
      if (live && lig == 0) {
        n_edit += dec.edited;
        if (do_filter) {
          n_pass += pass;
          n_fail += !pass;
        }
      }
      if (live && lig == 0) {
        h.start[r] = start;
        h.len[r] = len;
        h.pass[r] = pass ? 1 : 0;
        if (a.trim) a.trim[i0 + r] = hpgq_trim_t{start, len};
        if (a.mask) a.mask[i0 + r] = pass ? 1 : 0;
      }
    }
    __syncthreads();
    // ---- C + D: tile scan (warp 0, four reads per lane) and decoupled look-back
    if (warp == 0) {
      uint4 v[4], sum = make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t r = 4u * lane + k;
        v[k] = make_uint4(0, 0, 0, 0);
        if (r < nt) v[k] = h.pass[r] ? make_uint4(h.len[r], 0, 1, 0) : make_uint4(0, h.len[r], 0, 1);
        sum = add4(sum, v[k]);
      }
      uint4 inc = sum;
#pragma unroll
      for (int dd = 1; dd < 32; dd <<= 1) {
        const uint4 u = shfl_up4(inc, dd);
        if ((int)lane >= dd) inc = add4(inc, u);
      }
      const uint4 agg = make_uint4(__shfl_sync(0xFFFFFFFFu, inc.x, 31), __shfl_sync(0xFFFFFFFFu, inc.y, 31),
                                   __shfl_sync(0xFFFFFFFFu, inc.z, 31), __shfl_sync(0xFFFFFFFFu, inc.w, 31));
      // ---- D: the tile's place in the output from a three-level tree of TOTALS (no chain of inclusive prefixes).
      // Level 1 = a tile's own totals; the last tile of every group of 32 also sums its group into a level-2 record,
      // the last tile of every 32 groups those into a level-3 record.  A tile's exclusive prefix is then the sum of
      // the level-3 records in front of its 1024-group, the level-2 records in front of its group and the level-1
      // records in front of it inside its group: three sets of independent loads, all requested at once, ONE round
      // trip to L2 -- and every record it needs was published by a tile that started before it and never waits for
      // anything.  (The classic look-back walks back window by window to the nearest tile that already knows its
      // inclusive prefix: with 592 CTAs in flight that cost ~ 17 ns per tile chip-wide, profiles/fused_r2_summary.md.)
      fz_rec *L1 = z.agg, *L2 = z.incl + 1, *L3 = z.incl + 1 + ((z.n_tiles + 31u) >> 5);
      if (lane == 0) fz_publish(&L1[tile], agg);
      auto sum_recs = [&](const fz_rec *base, uint32_t count) -> uint4 {  // sum of base[0 .. count), count <= 32
        uint4 val = make_uint4(0, 0, 0, 0);
        bool ok;
        do {
          ok = true;
          if (lane < count) ok = fz_read(&base[lane], &val);
        } while (!__all_sync(0xFFFFFFFFu, ok));
        if (lane >= count) val = make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int dd = 16; dd > 0; dd >>= 1) {
          val.x += __shfl_xor_sync(0xFFFFFFFFu, val.x, dd);
          val.y += __shfl_xor_sync(0xFFFFFFFFu, val.y, dd);
          val.z += __shfl_xor_sync(0xFFFFFFFFu, val.z, dd);
          val.w += __shfl_xor_sync(0xFFFFFFFFu, val.w, dd);
        }
        return val;
      };
      const uint32_t t1 = tile & 31u, g2 = tile >> 5, t2 = g2 & 31u, g3 = tile >> 10;
      const uint4 in_group = sum_recs(&L1[tile - t1], t1);            // tiles of my group in front of me
      if (t1 == 31u) {                                                // duty: my group's level-2 record
        const uint4 gsum = add4(in_group, agg);
        if (lane == 0) fz_publish(&L2[g2], gsum);
        if (t2 == 31u) {                                              // and the level-3 record of my 32 groups
          const uint4 ssum = add4(sum_recs(&L2[g2 - 31u], 31u), gsum);
          if (lane == 0) fz_publish(&L3[g3], ssum);
        }
      }
      uint4 pre = add4(in_group, sum_recs(&L2[g2 - t2], t2));
      for (uint32_t k = 0; k < g3; k += 32u) pre = add4(pre, sum_recs(&L3[k], min(32u, g3 - k)));
      if (tile == z.n_tiles - 1u && lane == 0) fz_publish(&z.incl[0], add4(pre, agg));  // the batch's totals
      // ---- E: keys of the tile's reads in their streams, source offsets, output offsets / source indices
      uint4 run = make_uint4(pre.x + inc.x - sum.x, pre.y + inc.y - sum.y, pre.z + inc.z - sum.z, pre.w + inc.w - sum.w);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t r = 4u * lane + k;
        if (r <= nt) {
          h.ct.key[0][r] = run.x;
          h.ct.key[1][r] = run.y;
        }
        if (r < nt) {
          h.ct.soff[r] = h.sidx[r] + h.start[r] - S0a + CT_PAD;
          if (h.pass[r]) {
            b.out_src[run.z] = i0 + r;
            b.out_idx[run.z] = run.x;
          } else {
            z.fail_src[run.w] = i0 + r;
            z.fail_idx[run.w] = run.y;
          }
        }
        run = add4(run, v[k]);
      }
      if (lane == 31 && nt == 4u * 32u) {  // a full tile: the end of both spans lies behind the last lane's reads
        h.ct.key[0][nt] = run.x;
        h.ct.key[1][nt] = run.y;
      }
      if (lane == 0) h.prefix = pre;
    }
    __syncthreads();
    // chunk -> read map of both spans
    const uint32_t D0p = h.ct.key[0][0], D1p = h.ct.key[0][nt], D0f = h.ct.key[1][0];
    const uint32_t mo_f = ((D1p + 15u) >> 4) - (D0p >> 4) + 1u;
    for (uint32_t r = tid; r < nt; r += 256u) {
      const uint32_t st = h.pass[r] ? 0u : 1u;
      const uint32_t k0 = h.ct.key[st][r], k1 = h.ct.key[st][r + 1u], D0s = st ? D0f : D0p;
      if (k1 > k0) {
        const uint32_t cb = D0s >> 4;
        const uint32_t cf = k0 == D0s ? cb : (k0 + 15u) >> 4, cl = (k1 - 1u) >> 4;
        uint8_t *mp = rmap + (st ? mo_f : 0u) - cb;
        for (uint32_t c = cf; c <= cl; ++c) mp[c] = (uint8_t)r;
      }
    }
    __syncthreads();
    // ---- F: the two output spans
    uint8_t *const oseq[2] = {b.out_seq, z.fail_seq}, *const oqual[2] = {b.out_qual, z.fail_qual};
    ct_emit(h.ct, rmap, mo_f, sseq, squal, nt, oseq, oqual);
  }
  unsigned long long *C = a.ctr;
  n_pass = group_sum<32>(n_pass);
  n_fail = group_sum<32>(n_fail);
  n_edit = group_sum<32>(n_edit);
  if (lane == 0) {
    if (n_pass) atomicAdd(&C[HPGQ_S_NUM_PASSED], (unsigned long long)n_pass);
    if (n_fail) atomicAdd(&C[HPGQ_S_NUM_FAILED], (unsigned long long)n_fail);
    if (n_edit) atomicAdd(&C[HPGQ_S_NUM_EDITED], (unsigned long long)n_edit);
  }
}

// ---------------------------------------------------------------- decisions only, word frame, straight from global
// The decision half of the two-pass path (edit / filter without statistics) for the common case fz_decide_fast takes:
// G lanes per read, K = 5 words per lane read with 4-byte loads at the read's own byte phase (two loads + one PRMT per
// word; a group's loads cover consecutive 32-byte sectors).  ~ 40 instructions per read against ~ 100 for the
// register-resident kernel of hpgq_group.cu, which remains for filter windows / out-of-quality limits / longer reads.
template <int G>
__global__ void __launch_bounds__(256, 4) hq_dec_fast_kernel(hq_kargs a) {
  constexpr int K = 5;
  constexpr uint32_t NG = 32u / G;
  const uint32_t lane = threadIdx.x & 31u, lig = lane % G, giw = lane / G;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  const bool do_filter = a.work & HPGQ_WORK_FILTER, do_edit = a.work & HPGQ_WORK_EDIT;
  uint32_t n_pass = 0, n_fail = 0, n_edit = 0;
  fz_len_consts<K, G> lc;
  lc.L = 0xFFFFFFFFu;
  lc.wl = lc.wr = 0u;
  const uint32_t end_ok = a.idx[a.n_reads] + 32u;
  const uint32_t *dummy = reinterpret_cast<const uint32_t *>(a.ctr);  // >= 32 u64 scalars + histograms: always readable
  // the offsets of a round are loaded one round ahead, and the lines of its frames are asked for (prefetch into L1) as
  // soon as the offsets are there: the round's own loads then hit L1 instead of waiting for HBM behind the offset load
  // (45 % of the stall samples sat on that chain: profiles/compact_r2_summary.md)
  uint32_t po = 0, pe = 0;
  {
    const uint32_t r = warp * NG + giw;
    if (r < a.n_reads) {
      po = a.idx[r];
      pe = a.idx[r + 1];
    }
  }
  for (uint32_t r0 = warp * NG; r0 < a.n_reads; r0 += nwarps * NG) {  // warp-uniform trip count (shuffles below)
    const uint32_t r = r0 + giw;
    const bool live = r < a.n_reads;
    const uint32_t o = live ? po : a.idx[0], L = live ? pe - po : 0u;
    {
      const uint32_t rn = r + nwarps * NG;
      if (rn < a.n_reads) {
        po = a.idx[rn];
        pe = a.idx[rn + 1];
        // 32 bytes per lane of the group cover the frame (4 K G bytes + the phase) of the next read
        const uint32_t off = (po & ~31u) + 32u * lig;
        if (off < pe + 32u && off + 32u <= end_ok) {
          asm volatile("prefetch.global.L1 [%0];" ::"l"(a.seq + off));
          asm volatile("prefetch.global.L1 [%0];" ::"l"(a.qual + off));
        }
      }
    }
    if (L != lc.L) fz_len_update<K, G>(a, lc, L, lig);
    // the frame of the last reads of the batch reaches past the arrays' readable end (idx[n] + 32): those reads take
    // the register code, whose loads stop at the read's end, and the frame loads of their lanes go to the counter block
    const bool tail = o + 4u * K * G + 8u > end_ok;
    const bool slow = lc.wl + lc.wr > L || tail;
    const uint32_t *ps = tail ? dummy : reinterpret_cast<const uint32_t *>(a.seq) + (o >> 2) + lig;
    const uint32_t *pq = tail ? dummy : reinterpret_cast<const uint32_t *>(a.qual) + (o >> 2) + lig;
    fz_dec dec = fz_decide_fast<K, G>(a, lc, ps, pq, 0x3210u + 0x1111u * (o & 3u), do_edit, do_filter);
    if (__any_sync(0xFFFFFFFFu, slow)) {
      const fz_dec d2 = fz_decide_regs<G, false>(a, a.seq, a.qual, CT_PAD, o, L, lig, do_edit, do_filter, false);
      if (slow) dec = d2;
    }
    if (live && lig == 0) {
      n_edit += dec.edited;
      if (do_filter) {
        n_pass += dec.pass;
        n_fail += !dec.pass;
      }
      if (a.trim) a.trim[r] = hpgq_trim_t{dec.start, dec.len};
      if (a.mask) a.mask[r] = dec.pass ? 1 : 0;
    }
  }
  unsigned long long *C = a.ctr;
  n_pass = group_sum<32>(n_pass);
  n_fail = group_sum<32>(n_fail);
  n_edit = group_sum<32>(n_edit);
  if (lane == 0) {
    if (n_pass) atomicAdd(&C[HPGQ_S_NUM_PASSED], (unsigned long long)n_pass);
    if (n_fail) atomicAdd(&C[HPGQ_S_NUM_FAILED], (unsigned long long)n_fail);
    if (n_edit) atomicAdd(&C[HPGQ_S_NUM_EDITED], (unsigned long long)n_edit);
  }
}

// The failed reads' bytes, offsets and source indices move behind the passed reads; totals for the host.
__global__ void __launch_bounds__(256) hq_fz_finish(hq_kargs a, hq_compact_bufs b, fz_bufs z) {
  uint4 tot = make_uint4(0, 0, 0, 0);
  if (z.n_tiles) (void)fz_read(&z.incl[0], &tot);
  const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x, gstride = gridDim.x * blockDim.x;
  const uint32_t D0 = tot.x, D1 = tot.x + tot.y;
  for (uint32_t c = (D0 >> 4) + gtid; c < ((D1 + 15u) >> 4); c += gstride) {
    const uint32_t dc = c << 4, d0 = max(dc, D0), d1 = min(dc + 16u, D1);
    const uint32_t x = dc - D0;  // side-buffer byte that lands on byte 0 of the chunk (mod 2^32: may be "negative")
    const uint32_t xa = x & ~15u;
    uint4 As = make_uint4(0, 0, 0, 0), Aq = As, Bs = As, Bq = As;
    if (dc >= D0) {
      As = *reinterpret_cast<const uint4 *>(z.fail_seq + xa);
      Aq = *reinterpret_cast<const uint4 *>(z.fail_qual + xa);
    }
    if (xa + 16u < tot.y + 16u || dc < D0) {  // the B half exists (the side buffers are padded by 64 bytes)
      Bs = *reinterpret_cast<const uint4 *>(z.fail_seq + (dc >= D0 ? xa + 16u : 0u));
      Bq = *reinterpret_cast<const uint4 *>(z.fail_qual + (dc >= D0 ? xa + 16u : 0u));
    }
    const uint4 vs = shift_sel(As, Bs, x & 15u), vq = shift_sel(Aq, Bq, x & 15u);
    if (d1 - d0 == 16u) {
      *reinterpret_cast<uint4 *>(b.out_seq + dc) = vs;
      *reinterpret_cast<uint4 *>(b.out_qual + dc) = vq;
    } else {
      const uint32_t ws[4] = {vs.x, vs.y, vs.z, vs.w}, wq[4] = {vq.x, vq.y, vq.z, vq.w};
      for (uint32_t t = d0 - dc; t < d1 - dc; ++t) {
        b.out_seq[dc + t] = (uint8_t)(ws[t >> 2] >> (8u * (t & 3u)));
        b.out_qual[dc + t] = (uint8_t)(wq[t >> 2] >> (8u * (t & 3u)));
      }
    }
  }
  for (uint32_t j = gtid; j < tot.w; j += gstride) {
    b.out_idx[tot.z + j] = tot.x + z.fail_idx[j];
    b.out_src[tot.z + j] = z.fail_src[j];
  }
  if (gtid == 0) {
    b.out_idx[a.n_reads] = D1;
    b.totals[0] = tot.z;
    b.totals[1] = tot.w;
    b.totals[2] = tot.x;
    b.totals[3] = tot.y;
  }
}

// Order-sensitive 64-bit checksum of a slot's compacted output (bench.py compares it across GPU counts, tests with
// the host twin hpgq_checksum_host): sum over every output byte / offset / source index of mix(position, value).
__device__ __forceinline__ unsigned long long ck_mix(unsigned long long pos, unsigned long long v) {
  return hpgq_checksum_mix(pos, v);
}
__global__ void hq_checksum_kernel(hq_compact_bufs b, uint32_t n, unsigned long long *out) {
  const uint32_t T = b.totals[2] + b.totals[3];
  unsigned long long h = 0;
  const uint32_t stride = gridDim.x * blockDim.x, t0 = blockIdx.x * blockDim.x + threadIdx.x;
  for (uint32_t i = t0; i < T; i += stride) {
    h += ck_mix(i, b.out_seq[i]);
    h += ck_mix((unsigned long long)T + i, b.out_qual[i]);
  }
  for (uint32_t j = t0; j <= n; j += stride) h += ck_mix(2ull * T + j, b.out_idx[j]);
  for (uint32_t j = t0; j < n; j += stride) h += ck_mix(2ull * T + n + 1ull + j, b.out_src[j]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) h += __shfl_xor_sync(0xFFFFFFFFu, h, o);
  if ((threadIdx.x & 31u) == 0 && h) atomicAdd(out, h);
}

} // namespace

cudaError_t hq_launch_checksum(const hq_compact_bufs &b, uint32_t n, unsigned long long *d_out, cudaStream_t st,
                               uint64_t *launches) {
  cudaError_t e = cudaMemsetAsync(d_out, 0, 8, st);
  if (e != cudaSuccess) return e;
  hq_checksum_kernel<<<148 * 8, 256, 0, st>>>(b, n, d_out);
  ++*launches;
  return cudaGetLastError();
}

namespace {
uint32_t fz_rpt(uint32_t max_len) {
  uint32_t rpt = CT_SPAN / (max_len ? max_len : 1u);
  uint32_t cap = CT_RPT;
  const char *v = getenv("HPGQ_FZ_RPT");  // developer knob: reads per tile (16..128)
  if (v && atoi(v) >= 16 && atoi(v) <= (int)CT_RPT) cap = (uint32_t)atoi(v);
  return rpt > cap ? cap : rpt;
}
}  // namespace

bool hq_fz_supports(uint32_t work, uint32_t max_len, uint32_t n_reads, const hq_compact_bufs &b) {
  if (!(work & HPGQ_WORK_COMPACT) || (work & (HPGQ_WORK_STATS | HPGQ_WORK_KMERS | HPGQ_WORK_CHAOS))) return false;
  if (!(work & (HPGQ_WORK_EDIT | HPGQ_WORK_FILTER))) return false;
  if (max_len == 0 || max_len + 15u > 32u * RG_LB || !b.fz_status) return false;  // reads held in registers by <= 32 lanes
  const uint32_t rpt = fz_rpt(max_len);
  return rpt >= 16u && (n_reads + rpt - 1u) / rpt <= b.fz_tiles_cap;
}

cudaError_t hq_launch_fz(const hq_kargs &a, const hq_compact_bufs &b, uint32_t max_len, int sm_count, cudaStream_t st,
                         uint64_t *launches) {
  fz_bufs z;
  z.agg = reinterpret_cast<fz_rec *>(b.fz_agg);
  z.incl = reinterpret_cast<fz_rec *>(b.fz_incl);
  z.status = b.fz_status;
  z.fail_seq = b.fail_seq;
  z.fail_qual = b.fail_qual;
  z.fail_idx = b.fail_idx;
  z.fail_src = b.fail_src;
  z.rpt = fz_rpt(max_len);
  z.cap = ((z.rpt * max_len + 15u) & ~15u) + 64u;
  z.n_tiles = (a.n_reads + z.rpt - 1u) / z.rpt;
  cudaError_t e = cudaMemsetAsync(z.status, 0, 4, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(z.agg, 0, (size_t)z.n_tiles * sizeof(fz_rec), st);
  if (e == cudaSuccess) e = cudaMemsetAsync(z.incl, 0, (size_t)(2u + ((z.n_tiles + 31u) >> 5) + ((z.n_tiles + 1023u) >> 10)) * sizeof(fz_rec), st);
  if (e != cudaSuccess) return e;
  const uint32_t span = max_len + 15u;
  int G = span <= 4u * RG_LB ? 4 : span <= 8u * RG_LB ? 8 : span <= 16u * RG_LB ? 16 : 32;
  const bool ooq = (a.work & HPGQ_WORK_FILTER) != 0 && a.max_ooq < 32 * RG_LB;
  // the word-frame decision (fz_decide_fast) takes batches without filter windows or an out-of-quality limit, trim
  // windows of <= 255 bases and reads of <= 4 K G - 3 bases (K = 5); HPGQ_FZ_SLOW keeps the register code (tests)
  const bool windows = (a.work & HPGQ_WORK_FILTER) != 0 && (a.f_ll > 0 || a.f_rl > 0);
  bool fast = !ooq && !windows && a.e_ll <= 255 && a.e_rl <= 255 && max_len <= 20u * 32u - 3u && !getenv("HPGQ_FZ_SLOW");
  if (fast) G = max_len <= 77u ? 4 : max_len <= 157u ? 8 : max_len <= 317u ? 16 : 32;
  // fz_decide_fast loads whole frames (4 K G + 4 bytes from a read's first word) whatever the read's length: the last
  // reads of a tile that fills its staged span reach up to a frame past the quality array, the last thing in the CTA's
  // shared memory (found by tools/stress_parity.py: 60 reads x 363 bp per tile, illegal address).  One frame of slack.
  const uint32_t smem = (((uint32_t)sizeof(fz_hdr) + 15u) & ~15u) + ct_rmap_bytes(z.cap) + 2u * z.cap +
                        (fast ? 4u * 5u * (uint32_t)G + 16u : 0u);
  if (a.n_reads) {
#define HQ_FZ(KK, GG, OO, FF)                                                                                     \
  do {                                                                                                           \
    e = cudaFuncSetAttribute(hq_fz_kernel<KK, GG, OO, FF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    int per_sm = 1;                                                                                              \
    if (e == cudaSuccess)                                                                                        \
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, hq_fz_kernel<KK, GG, OO, FF>, 256, smem);       \
    if (e != cudaSuccess) return e;                                                                              \
    uint32_t grid = (uint32_t)sm_count * (uint32_t)(per_sm > 0 ? per_sm : 1);                                    \
    if (grid > z.n_tiles) grid = z.n_tiles;                                                                      \
    hq_fz_kernel<KK, GG, OO, FF><<<grid, 256, smem, st>>>(a, b, z);                                              \
  } while (0)
#define HQ_FZ_G(GG)                              \
  do {                                           \
    if (fast) HQ_FZ(5, GG, false, true);         \
    else if (ooq) HQ_FZ(1, GG, true, false);     \
    else HQ_FZ(1, GG, false, false);             \
  } while (0)
    switch (G) {
      case 4: HQ_FZ_G(4); break;
      case 8: HQ_FZ_G(8); break;
      case 16: HQ_FZ_G(16); break;
      default: HQ_FZ_G(32); break;
    }
#undef HQ_FZ_G
#undef HQ_FZ
    ++*launches;
  }
  hq_fz_finish<<<sm_count * 4, 256, 0, st>>>(a, b, z);
  ++*launches;
  return cudaGetLastError();
}

// returns cudaErrorInvalidValue when the batch / options need the kernels of hpgq_group.cu
cudaError_t hq_launch_dec_fast(const hq_kargs &a, uint32_t max_len, int sm_count, cudaStream_t st, uint64_t *launches) {
  if (a.n_reads == 0) return cudaSuccess;
  if ((a.work & (HPGQ_WORK_STATS | HPGQ_WORK_KMERS | HPGQ_WORK_CHAOS)) || !(a.work & (HPGQ_WORK_EDIT | HPGQ_WORK_FILTER)))
    return cudaErrorInvalidValue;
  const bool ooq = (a.work & HPGQ_WORK_FILTER) != 0 && a.max_ooq < 32 * RG_LB;
  const bool windows = (a.work & HPGQ_WORK_FILTER) != 0 && (a.f_ll > 0 || a.f_rl > 0);
  if (ooq || windows || a.e_ll > 255 || a.e_rl > 255 || max_len == 0 || max_len > 20u * 32u - 3u || getenv("HPGQ_FZ_SLOW"))
    return cudaErrorInvalidValue;
  const int G = max_len <= 77u ? 4 : max_len <= 157u ? 8 : max_len <= 317u ? 16 : 32;
  const uint32_t rpb = 256u / (uint32_t)G;
  uint32_t blocks = (a.n_reads + rpb - 1u) / rpb;
  const uint32_t cap = (uint32_t)sm_count * 8u * 4u;
  if (blocks > cap) blocks = cap;
  switch (G) {
    case 4: hq_dec_fast_kernel<4><<<blocks, 256, 0, st>>>(a); break;
    case 8: hq_dec_fast_kernel<8><<<blocks, 256, 0, st>>>(a); break;
    case 16: hq_dec_fast_kernel<16><<<blocks, 256, 0, st>>>(a); break;
    default: hq_dec_fast_kernel<32><<<blocks, 256, 0, st>>>(a); break;
  }
  ++*launches;
  return cudaGetLastError();
}

cudaError_t hq_launch_compact(const hq_kargs &a, const hq_compact_bufs &b, uint32_t mean_len, uint32_t max_len,
                              cudaStream_t st, uint64_t *launches) {
  const uint32_t n1 = a.n_reads + 1u;
  const uint32_t nb = (n1 + SCAN_THREADS * SCAN_ITEMS - 1u) / (SCAN_THREADS * SCAN_ITEMS);
  hq_compact_sums<<<nb, SCAN_THREADS, 0, st>>>(a, b.block_sums);
  hq_compact_scan_sums<<<1, SCAN_THREADS, 0, st>>>(b.block_sums, nb);
  hq_compact_scan<<<nb, SCAN_THREADS, 0, st>>>(a, b.block_sums, b.scan, b.totals);
  *launches += 3;
  if (a.n_reads) {
    // lanes per read from the mean read length of the batch (bytes / reads are known to the caller only as
    // capacities, so use the idx span): short reads take 8 or 16 lanes, long reads a whole warp
    if (max_len && max_len <= CT_SPAN) {
      uint32_t rpt = CT_SPAN / max_len;
      if (rpt > CT_RPT) rpt = CT_RPT;
      const uint32_t cap = ((rpt * max_len + 15u) & ~15u) + 64u;  // + pad in front, + the B half behind the span
      const uint32_t smem = (((uint32_t)sizeof(ct_hdr) + 15u) & ~15u) + ct_rmap_bytes(cap) + 2u * cap;
      uint32_t tiles = (a.n_reads + rpt - 1u) / rpt;
      if (tiles > 148u * 8u) tiles = 148u * 8u;
      // two buffers of half-size tiles when at least two such tiles per CTA exist (HPGQ_CT_SINGLE: developer A/B)
      uint32_t rpt2 = rpt >= 32u ? rpt / 2u : 0u;
      static const uint32_t knob_rpt = getenv("HPGQ_CT_RPT2") ? (uint32_t)atoi(getenv("HPGQ_CT_RPT2")) : 0u;    // developer knobs
      static const uint32_t knob_ctas = getenv("HPGQ_CT_CTAS") ? (uint32_t)atoi(getenv("HPGQ_CT_CTAS")) : 0u;
      if (rpt2 && knob_rpt >= 8u && knob_rpt <= CT_RPT && knob_rpt * max_len <= CT_SPAN) rpt2 = knob_rpt;
      if (rpt2 && !getenv("HPGQ_CT_SINGLE")) {
        const uint32_t cap2 = ((rpt2 * max_len + 15u) & ~15u) + 64u;
        const uint32_t smem2 = (((uint32_t)sizeof(ct_hdr) + 15u) & ~15u) +
                               2u * ((((uint32_t)sizeof(ct_raw_t) + 15u) & ~15u) + ct_rmap_bytes(cap2) + 2u * cap2);
        uint32_t tiles2 = (a.n_reads + rpt2 - 1u) / rpt2;
        if (smem2 <= 100u * 1024u) {
          if (smem2 > 48u * 1024u) {
            cudaError_t e = cudaFuncSetAttribute(hq_compact_copy_tile2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
            if (e != cudaSuccess) return e;
          }
          // a grid-stride loop over tiles: exactly as many CTAs as are resident at once (a fifth CTA per SM that has to
          // wait for a slot runs as a second wave: 1.08 against 0.99 ms per C3 batch, profiles/sweep_c3_r2f.log)
          int occ = 0;
          if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, hq_compact_copy_tile2, 256, smem2) != cudaSuccess || occ < 1) occ = 4;
          const uint32_t ctas = knob_ctas ? knob_ctas : (uint32_t)occ;
          if (tiles2 > 148u * ctas) tiles2 = 148u * ctas;
          hq_compact_copy_tile2<<<tiles2, 256, smem2, st>>>(a, b, rpt2, cap2);
          ++*launches;
          return cudaGetLastError();
        }
      }
      hq_compact_copy_tile<<<tiles, 256, smem, st>>>(a, b, rpt, cap);
      ++*launches;
      return cudaGetLastError();
    }
    const int G = mean_len <= 112u ? 8 : mean_len <= 400u ? 16 : 32;
    const uint32_t rpb = 256u / (uint32_t)G;
    uint32_t blocks = (a.n_reads + rpb - 1u) / rpb;
    if (blocks > 148u * 32u) blocks = 148u * 32u;
    if (G == 8) hq_compact_copy<8><<<blocks, 256, 0, st>>>(a, b);
    else if (G == 16) hq_compact_copy<16><<<blocks, 256, 0, st>>>(a, b);
    else hq_compact_copy<32><<<blocks, 256, 0, st>>>(a, b);
    ++*launches;
  }
  return cudaGetLastError();
}

// ------------------------------------------------------------------ compaction without any allocation
// hpgq_compact_noalloc (include/hpg_fastq_cuda.h): the scan + copy kernels above with their scratch (the exclusive
// scan of the per-read records and the block sums) in a file-scope device array instead of cudaMalloc'ed buffers,
// for callers that may not allocate (CUDA graph capture, serving loops).  Every GPU that loads the library holds
// the array (4 MB + 4 KB) until the process exits.
namespace {
constexpr uint32_t NOALLOC_SCAN = HPGQ_NOALLOC_MAX_READS + 2u;             // uint4 entries: scan[0..n]
constexpr uint32_t NOALLOC_SUMS = (HPGQ_NOALLOC_MAX_READS + 1u) / 1024u + 2u;  // uint4 entries: block sums
static __device__ __align__(256) uint4 g_noalloc_scratch[NOALLOC_SCAN + NOALLOC_SUMS];

// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is asked for
// once per device, with that device current, and cached.
constexpr int NOALLOC_MAX_DEVICES = 64;
uint4 *g_noalloc_ptr[NOALLOC_MAX_DEVICES];
bool g_noalloc_have[NOALLOC_MAX_DEVICES];
pthread_mutex_t g_noalloc_mu = PTHREAD_MUTEX_INITIALIZER;

cudaError_t noalloc_scratch(int device, uint4 **out) {
  cudaError_t e = cudaSuccess;
  pthread_mutex_lock(&g_noalloc_mu);
  if (!g_noalloc_have[device]) {
    uint4 *p = nullptr;
    e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
    if (e == cudaSuccess) {
      g_noalloc_ptr[device] = p;
      g_noalloc_have[device] = true;
    }
  }
  *out = g_noalloc_ptr[device];
  pthread_mutex_unlock(&g_noalloc_mu);
  return e;
}
}  // namespace

extern "C" int hpgq_compact_noalloc(const uint8_t *d_seq, const uint8_t *d_qual, const uint32_t *d_idx,
                                    const uint8_t *d_mask, const hpgq_trim_t *d_trim, uint32_t n_reads, uint32_t max_len,
                                    uint8_t *d_out_seq, uint8_t *d_out_qual, uint32_t *d_out_idx, uint32_t *d_out_src,
                                    uint32_t *d_totals, void *stream) {
  if (!d_seq || !d_qual || !d_idx || !d_out_seq || !d_out_qual || !d_out_idx || !d_out_src || !d_totals)
    return hq_fail_msg(HPGQ_ERR_INVALID, "compact_noalloc: null argument");
  if (n_reads > HPGQ_NOALLOC_MAX_READS)
    return hq_fail_msg(HPGQ_ERR_INVALID,
                       "compact_noalloc: n_reads exceeds HPGQ_NOALLOC_MAX_READS (262144), the size of the library's static "
                       "scratch; split the batch or use hpgq_batch_submit* with HPGQ_WORK_COMPACT");
  int device = -1;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess) return hq_fail_cuda(e, "compact_noalloc: cudaGetDevice");
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) return hq_fail_msg(HPGQ_ERR_INVALID, "compact_noalloc: device ordinal out of range");
  uint4 *scratch = nullptr;
  if ((e = noalloc_scratch(device, &scratch)) != cudaSuccess) return hq_fail_cuda(e, "compact_noalloc: cudaGetSymbolAddress");
  hq_kargs a;
  memset(&a, 0, sizeof a);
  a.seq = d_seq;
  a.qual = d_qual;
  a.idx = d_idx;
  a.n_reads = n_reads;
  a.mask = const_cast<uint8_t *>(d_mask);
  a.trim = const_cast<hpgq_trim_t *>(d_trim);
  hq_compact_bufs b;
  memset(&b, 0, sizeof b);
  // Nothing in the scratch has to start at zero: the three scan kernels write block_sums[0, nb) and scan[0, n] before
  // the copy kernel reads them, so what the previous call left behind is never read.
  b.scan = scratch;
  b.block_sums = scratch + NOALLOC_SCAN;
  b.out_seq = d_out_seq;
  b.out_qual = d_out_qual;
  b.out_idx = d_out_idx;
  b.out_src = d_out_src;
  b.totals = d_totals;
  uint64_t launches = 0;
  // lanes per read of the generic copy kernel: from the longest read when the caller states it, else a whole warp
  const uint32_t mean_len = max_len ? max_len : 401u;
  e = hq_launch_compact(a, b, mean_len, max_len, static_cast<cudaStream_t>(stream), &launches);
  if (e != cudaSuccess) return hq_fail_cuda(e, "compact_noalloc: kernel launch");
  if (n_reads == 0) {  // the copy kernel, which closes out_idx with the total, does not run for an empty batch
    e = cudaMemsetAsync(d_out_idx, 0, 4, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return hq_fail_cuda(e, "compact_noalloc: cudaMemsetAsync");
  }
  return HPGQ_OK;
}
