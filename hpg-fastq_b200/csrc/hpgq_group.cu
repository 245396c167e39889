// hpgq_group.cu -- vectorised QC kernels for everything the tiled kernel does not take: reads of ANY length
// (long-read config C5) and the EDIT path (config C3).
//
//   hq_qc_group_kernel<G>   G lanes own one read; 16-byte coalesced loads of the quality / sequence bytes, SIMD-in-word
//                           sums, shuffle reduction inside the group.  Computes the trim decision (fastq_edit,
//                           reference call src/edit_fastq.c:148-154), the pass/fail decision (fastq_filter,
//                           src/stats_fastq.c:209-224) and the per-read part of the consumer merge (length / mean
//                           quality / GC histograms, src/stats_fastq.c:286-334).
//   hq_pos_kernel           per-position part of the merge (src/stats_fastq.c:338-382) for long reads: a thread OWNS
//                           eight read positions and walks over the reads of its slab, so the A/C/G/T/N and quality
//                           sums are register accumulators -- no atomics in the loop; the grid is
//                           (position tiles) x (read slabs).
// Both read HBM with full 128-byte-and-wider warp requests; the second kernel re-reads the batch (2 passes).
#include <stdlib.h>
#include "hpgq_dev.cuh"
#include "hpgq_internal.h"
#include "hpgq_regs.cuh"

namespace {

struct hq_sums4 {
  uint32_t S, nN, nGC, ooq;
};

__device__ __forceinline__ bool mean_ok_g(uint32_t S, uint32_t w, int32_t lo, int32_t hi) {
  const long long s = (long long)S, ww = (long long)w;
  return s >= (long long)lo * ww && s <= (long long)hi * ww;
}

template <bool SEQ>
__device__ __forceinline__ void acc_word(const hq_kargs &a, uint32_t qw, uint32_t sw, uint32_t m, hq_sums4 &r) {
  const uint32_t qm = qw & m;
  r.S = __dp4a(qm, 0x01010101u, r.S);
  if (SEQ) {
    if (qm & 0x80808080u) {  // bytes >= 128 are not ASCII qualities: exact per-byte comparison
      for (int i = 0; i < 4; ++i)
        if ((m >> (8 * i)) & 1u) {
          const int32_t qq = (int32_t)((qw >> (8 * i)) & 0xFFu);
          r.ooq += (qq < a.q_lo) | (qq > a.q_hi);
        }
    } else {
      r.ooq += hq_ooq_count(qm, m, a.c_lo, a.c_hi);  // masked word: a neighbouring read's byte >= 128 must not carry in
    }
    r.nN += __popc(__vcmpeq4(sw, 0x4E4E4E4Eu) & m & 0x01010101u);
    r.nGC += __popc((__vcmpeq4(sw, 0x47474747u) | __vcmpeq4(sw, 0x43434343u)) & m & 0x01010101u);
  }
}

// sums over the bytes [lo, hi) (offsets into the batch arrays) by the G lanes of a group, NOT yet reduced
template <int G, bool SEQ>
__device__ __forceinline__ hq_sums4 range_sums(const hq_kargs &a, uint32_t lo, uint32_t hi, uint32_t lig) {
  hq_sums4 r = {0u, 0u, 0u, 0u};
  if (hi <= lo) return r;
  const uint32_t base = lo & ~15u;
  const uint32_t nch = (hi - base + 15u) >> 4;
#pragma unroll 2
  for (uint32_t k = lig; k < nch; k += G) {
    const uint32_t c = base + 16u * k;
    const uint4 q = *reinterpret_cast<const uint4 *>(a.qual + c);
    uint4 s = make_uint4(0, 0, 0, 0);
    if (SEQ) s = *reinterpret_cast<const uint4 *>(a.seq + c);
    if (c >= lo && hi - c >= 16u) {
      acc_word<SEQ>(a, q.x, s.x, 0xFFFFFFFFu, r);
      acc_word<SEQ>(a, q.y, s.y, 0xFFFFFFFFu, r);
      acc_word<SEQ>(a, q.z, s.z, 0xFFFFFFFFu, r);
      acc_word<SEQ>(a, q.w, s.w, 0xFFFFFFFFu, r);
    } else {
      acc_word<SEQ>(a, q.x, s.x, hq_tail_mask(c, hi) & ~hq_tail_mask(c, lo), r);
      acc_word<SEQ>(a, q.y, s.y, hq_tail_mask(c + 4u, hi) & ~hq_tail_mask(c + 4u, lo), r);
      acc_word<SEQ>(a, q.z, s.z, hq_tail_mask(c + 8u, hi) & ~hq_tail_mask(c + 8u, lo), r);
      acc_word<SEQ>(a, q.w, s.w, hq_tail_mask(c + 12u, hi) & ~hq_tail_mask(c + 12u, lo), r);
    }
  }
  return r;
}

template <int G>
__global__ void __launch_bounds__(256) hq_qc_group_kernel(hq_kargs a) {
  constexpr uint32_t NG = 32u / G;  // groups per warp
  const uint32_t lane = threadIdx.x & 31u, lig = lane % G, giw = lane / G;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
  const bool do_stats = a.work & HPGQ_WORK_STATS, do_filter = a.work & HPGQ_WORK_FILTER;
  const bool do_edit = a.work & HPGQ_WORK_EDIT;
  unsigned long long *C = a.ctr;
  uint32_t n_pass = 0, n_fail = 0, n_edit = 0, n_stat = 0;
  unsigned long long acc_len = 0;

  // warp-uniform trip count: every lane reaches every shuffle
  for (uint32_t r0 = warp * NG; r0 < a.n_reads; r0 += nwarps * NG) {
    const uint32_t r = r0 + giw;
    const bool live = r < a.n_reads;
    const uint32_t o = live ? a.idx[r] : 0u;
    const uint32_t L0 = live ? a.idx[r + 1] - o : 0u;
    uint32_t start = 0, len = L0;
    if (do_edit) {
      bool edited = false;
      if (a.e_ll > 0) {
        const uint32_t w = min((uint32_t)a.e_ll, len);
        const uint32_t sw = group_sum<G>(range_sums<G, false>(a, o, o + w, lig).S);
        if (w && !mean_ok_g(sw, w, a.e_lq_lo, a.e_lq_hi)) {
          start += w;
          len -= w;
          edited = true;
        }
      }
      if (a.e_rl > 0) {
        const uint32_t w = min((uint32_t)a.e_rl, len);
        const uint32_t e = o + start + len;
        const uint32_t sw = group_sum<G>(range_sums<G, false>(a, e - w, e, lig).S);
        if (w && !mean_ok_g(sw, w, a.e_rq_lo, a.e_rq_hi)) {
          len -= w;
          edited = true;
        }
      }
      n_edit += (edited && lig == 0);
    }
    if (a.trim && live && lig == 0) a.trim[r] = hpgq_trim_t{start, len};
    const uint32_t b0 = o + start, b1 = b0 + len;
    hq_sums4 t = range_sums<G, true>(a, b0, b1, lig);
    t.S = group_sum<G>(t.S);
    t.nN = group_sum<G>(t.nN);
    t.nGC = group_sum<G>(t.nGC);
    bool pass = true;
    if (do_filter) {
      t.ooq = group_sum<G>(t.ooq);
      const uint32_t wl = a.f_ll > 0 ? min((uint32_t)a.f_ll, len) : 0u;
      const uint32_t wr = a.f_rl > 0 ? min((uint32_t)a.f_rl, len) : 0u;
      pass = (long long)len >= a.min_len && (long long)len <= a.max_len && mean_ok_g(t.S, len, a.q_lo, a.q_hi) &&
             (long long)t.ooq <= a.max_ooq && (long long)t.nN <= a.max_n;
      if (a.f_ll > 0) {
        const uint32_t SL = group_sum<G>(range_sums<G, false>(a, b0, b0 + wl, lig).S);
        if (wl) pass = pass && mean_ok_g(SL, wl, a.f_lq_lo, a.f_lq_hi);
      }
      if (a.f_rl > 0) {
        const uint32_t SR = group_sum<G>(range_sums<G, false>(a, b1 - wr, b1, lig).S);
        if (wr) pass = pass && mean_ok_g(SR, wr, a.f_rq_lo, a.f_rq_hi);
      }
      if (live && lig == 0) {
        n_pass += pass;
        n_fail += !pass;
      }
    }
    if (live && lig == 0) {
      if (a.mask) a.mask[r] = pass ? 1 : 0;
      if (do_stats && pass && len <= a.Lcap) {
        n_stat++;
        acc_len += len;
        atomicAdd(&C[a.o_len_hist + len], 1ull);
        atomicAdd(&C[a.o_sumq + len], (unsigned long long)t.S);
        if (len) {
          atomicAdd(&C[a.o_qual_hist + (uint32_t)((2ull * t.S + len) / (2ull * len))], 1ull);
          atomicAdd(&C[a.o_gc_hist + (uint32_t)((100ull * t.nGC) / len)], 1ull);
        }
      }
    }
  }
  // one set of scalar atomics per warp
  n_pass = group_sum<32>(n_pass);
  n_fail = group_sum<32>(n_fail);
  n_edit = group_sum<32>(n_edit);
  n_stat = group_sum<32>(n_stat);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc_len += __shfl_xor_sync(0xFFFFFFFFu, acc_len, o);
  if (lane == 0) {
    if (n_pass) atomicAdd(&C[HPGQ_S_NUM_PASSED], (unsigned long long)n_pass);
    if (n_fail) atomicAdd(&C[HPGQ_S_NUM_FAILED], (unsigned long long)n_fail);
    if (n_edit) atomicAdd(&C[HPGQ_S_NUM_EDITED], (unsigned long long)n_edit);
    if (n_stat) atomicAdd(&C[HPGQ_S_NUM_READS], (unsigned long long)n_stat);
    if (acc_len) atomicAdd(&C[HPGQ_S_ACC_LENGTH], acc_len);
  }
}

// ---------------------------------------------------------------- register-resident variant (reads <= 32*G - 15 bytes)
// The whole read sits in registers: every lane owns 32 consecutive bytes (two 16-byte loads per array).  Bytes
// outside the read -- and, after the trim decision, outside the kept range -- are ZEROED once, so the main sums
// run without masks; the few lanes that overlap a quality window build byte masks with SIMD compares.
// GCS: the GC count is needed (statistics); OOQ: the out-of-quality count can fail a read (max_ooq below the
// longest read this kernel takes)
template <int G, bool GCS, bool OOQ>
__global__ void __launch_bounds__(256) hq_qc_regs_kernel(hq_kargs a) {
  constexpr uint32_t NG = 32u / G;
  constexpr int U = 2;
  const uint32_t lane = threadIdx.x & 31u, lig = lane % G, giw = lane / G;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
  const bool do_stats = a.work & HPGQ_WORK_STATS, do_filter = a.work & HPGQ_WORK_FILTER;
  const bool do_edit = a.work & HPGQ_WORK_EDIT;
  const bool zero_is_ooq = a.q_lo > 0 || a.q_hi < 0;  // a zeroed byte counts as out of range: corrected per lane
  unsigned long long *C = a.ctr;
  uint32_t n_pass = 0, n_fail = 0, n_edit = 0, n_stat = 0;
  unsigned long long acc_len = 0;

  for (uint32_t r0 = warp * NG * U; r0 < a.n_reads; r0 += nwarps * NG * U) {
    uint32_t rr[U], oo[U], LL[U];
    rg_read d[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      rr[u] = r0 + u * NG + giw;
      const bool live = rr[u] < a.n_reads;
      oo[u] = live ? a.idx[rr[u]] : 0u;
      LL[u] = live ? a.idx[rr[u] + 1] - oo[u] : 0u;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      d[u].c = (oo[u] & ~15u) + RG_LB * lig;
#pragma unroll
      for (int k = 0; k < RG_CPL; ++k) {
        uint4 q4 = make_uint4(0, 0, 0, 0), s4 = q4;
        if (d[u].c + 16u * k < oo[u] + LL[u]) {
          q4 = *reinterpret_cast<const uint4 *>(a.qual + d[u].c + 16u * k);
          s4 = *reinterpret_cast<const uint4 *>(a.seq + d[u].c + 16u * k);
        }
        d[u].q[4 * k + 0] = q4.x; d[u].q[4 * k + 1] = q4.y; d[u].q[4 * k + 2] = q4.z; d[u].q[4 * k + 3] = q4.w;
        d[u].s[4 * k + 0] = s4.x; d[u].s[4 * k + 1] = s4.y; d[u].s[4 * k + 2] = s4.z; d[u].s[4 * k + 3] = s4.w;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t r = rr[u], o = oo[u];
      const bool live = r < a.n_reads;
      uint32_t start = 0, len = LL[u];
      // the windows below mask their own byte range, so ONE clip (to the kept range) is enough
      if (do_edit) {
        bool edited = false;
        if (a.e_ll > 0) {
          const uint32_t w = min((uint32_t)a.e_ll, len);
          const uint32_t sw = group_sum<G>(rg_window_lane(d[u], o, o + w));
          if (w && !mean_ok32(sw, w, a.e_lq_lo, a.e_lq_hi)) {
            start += w;
            len -= w;
            edited = true;
          }
        }
        if (a.e_rl > 0) {
          const uint32_t w = min((uint32_t)a.e_rl, len);
          const uint32_t e = o + start + len;
          const uint32_t sw = group_sum<G>(rg_window_lane(d[u], e - w, e));
          if (w && !mean_ok32(sw, w, a.e_rq_lo, a.e_rq_hi)) {
            len -= w;
            edited = true;
          }
        }
        n_edit += (edited && lig == 0);
      }
      rg_clip(d[u], o + start, o + start + len);
      if (a.trim && live && lig == 0) a.trim[r] = hpgq_trim_t{start, len};
      const uint32_t b0 = o + start, b1 = b0 + len;
      // main sums over the registers (bytes outside [b0, b1) are zero)
      uint32_t S = 0, fN = 0, fGC = 0, fO = 0, hib = 0;
#pragma unroll
      for (int j = 0; j < RG_W; ++j) {
        const uint32_t qw = d[u].q[j], sw = d[u].s[j];
        S = __dp4a(qw, 0x01010101u, S);
        fN += rg_zero_flags(sw ^ 0x4E4E4E4Eu) >> 7;
        if (GCS) fGC += rg_zero_flags((sw & 0xFBFBFBFBu) ^ 0x43434343u) >> 7;  // 'C' = 0x43, 'G' = 0x47
        if (OOQ) {
          fO += (~((qw + a.c_lo) & ~(qw + a.c_hi)) & 0x80808080u) >> 7;
          hib |= qw;
        }
      }
      uint32_t nN = __dp4a(fN, 0x01010101u, 0u), nGC = __dp4a(fGC, 0x01010101u, 0u), ooq = __dp4a(fO, 0x01010101u, 0u);
      if (OOQ && zero_is_ooq) {  // remove the lane's zeroed bytes from the out-of-range count
        const uint32_t lo = max(b0, d[u].c), hi = min(b1, d[u].c + RG_LB);
        ooq -= RG_LB - (hi > lo ? hi - lo : 0u);
      }
      if (hib & 0x80808080u) {  // a quality byte >= 128: exact per-byte count for this lane
        ooq = 0;
        for (int j = 0; j < RG_W; ++j)
          for (int i = 0; i < 4; ++i) {
            const uint32_t pos = d[u].c + 4u * j + i;
            const int32_t qq = (int32_t)((d[u].q[j] >> (8 * i)) & 0xFFu);
            if (pos >= b0 && pos < b1) ooq += (qq < a.q_lo) | (qq > a.q_hi);
          }
      }
      // len < 32*G <= 1024: S < 2^18, counts < 2^10 -> two packed shuffle chains
      uint32_t p1 = S | (nN << 20), p2 = nGC | (ooq << 16);
      p1 = group_sum<G>(p1);
      if (GCS || OOQ) p2 = group_sum<G>(p2);
      S = p1 & 0xFFFFFu;
      nN = p1 >> 20;
      nGC = p2 & 0xFFFFu;
      ooq = p2 >> 16;
      bool pass = true;
      if (do_filter) {
        pass = (int32_t)len >= a.min_len && (int32_t)len <= a.max_len && mean_ok32(S, len, a.q_lo, a.q_hi) &&
               (int32_t)ooq <= a.max_ooq && (int32_t)nN <= a.max_n;
        if (a.f_ll > 0) {
          const uint32_t wl = min((uint32_t)a.f_ll, len);
          const uint32_t SL = group_sum<G>(rg_window_lane(d[u], b0, b0 + wl));
          if (wl) pass = pass && mean_ok32(SL, wl, a.f_lq_lo, a.f_lq_hi);
        }
        if (a.f_rl > 0) {
          const uint32_t wr = min((uint32_t)a.f_rl, len);
          const uint32_t SR = group_sum<G>(rg_window_lane(d[u], b1 - wr, b1));
          if (wr) pass = pass && mean_ok32(SR, wr, a.f_rq_lo, a.f_rq_hi);
        }
        if (live && lig == 0) {
          n_pass += pass;
          n_fail += !pass;
        }
      }
      if (live && lig == 0) {
        if (a.mask) a.mask[r] = pass ? 1 : 0;
        if (do_stats && pass && len <= a.Lcap) {
          n_stat++;
          acc_len += len;
          atomicAdd(&C[a.o_len_hist + len], 1ull);
          atomicAdd(&C[a.o_sumq + len], (unsigned long long)S);
          if (len) {
            atomicAdd(&C[a.o_qual_hist + (2u * S + len) / (2u * len)], 1ull);
            atomicAdd(&C[a.o_gc_hist + (100u * nGC) / len], 1ull);
          }
        }
      }
    }
  }
  n_pass = group_sum<32>(n_pass);
  n_fail = group_sum<32>(n_fail);
  n_edit = group_sum<32>(n_edit);
  n_stat = group_sum<32>(n_stat);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc_len += __shfl_xor_sync(0xFFFFFFFFu, acc_len, o);
  if (lane == 0) {
    if (n_pass) atomicAdd(&C[HPGQ_S_NUM_PASSED], (unsigned long long)n_pass);
    if (n_fail) atomicAdd(&C[HPGQ_S_NUM_FAILED], (unsigned long long)n_fail);
    if (n_edit) atomicAdd(&C[HPGQ_S_NUM_EDITED], (unsigned long long)n_edit);
    if (n_stat) atomicAdd(&C[HPGQ_S_NUM_READS], (unsigned long long)n_stat);
    if (acc_len) atomicAdd(&C[HPGQ_S_ACC_LENGTH], acc_len);
  }
}

// ---------------------------------------------------------------- per-position tables, column-owner form
#ifndef HQ_POS_THREADS
#define HQ_POS_THREADS 256
#endif
constexpr int POS_THREADS = HQ_POS_THREADS;
constexpr int POS_PER = 8;                       // positions owned by a thread (two 32-bit words)
constexpr int POS_TILE = POS_THREADS * POS_PER;  // positions per CTA
constexpr int POS_BATCH = 8;                     // reads in flight per thread (memory-level parallelism)

// eight bytes at an arbitrary byte offset: two aligned 8-byte loads and a 64-bit funnel shift
__device__ __forceinline__ void load8(const uint8_t *base, uint32_t off, uint32_t &v0, uint32_t &v1) {
  const unsigned long long *w = reinterpret_cast<const unsigned long long *>(base + (off & ~7u));
  const unsigned long long lo = __ldg(w), hi = __ldg(w + 1);
  const uint32_t sh = (off & 7u) * 8u;
  const unsigned long long v = sh ? ((lo >> sh) | (hi << (64u - sh))) : lo;
  v0 = (uint32_t)v;
  v1 = (uint32_t)(v >> 32);
}

// Shared-memory ring filled with 16-byte cp.async copies issued by ALL threads of the CTA: warp 0 looks up the next
// POS_RB reads that cover the tile and publishes their (offset, bytes, length) in the stage's header one iteration
// ahead; after the iteration's barrier every thread copies its share of the 16-byte aligned supersets of the tile's
// bytes and lets the copies arrive on the stage's mbarrier (cp.async.mbarrier.arrive.noinc); every warp then takes
// its eight positions of each read from shared memory.  NS - 1 stages (of POS_RB reads x 2 x 2 KB) are in flight per
// CTA, so HBM latency is hidden without spending registers on loads.  (One 1-D bulk copy per read and array was
// measured first: ~0.7 M copies of 2 KB per launch sustain only 1.9 TB/s, the TMA engine is op-rate bound there.)
// Four CTAs per SM (32 warps): the kernel is bound by latency, not by issue or HBM.  That takes <= 64 registers, so a
// thread keeps no wide per-position sums: its packed 8-bit / 16-bit lanes go to the global tables every ~240 reads
// (one reduction per non-zero lane), and a 50 KB ring of 3 stages x 4 reads.  C5 batch on B200: 0.745 ms with 2 CTAs,
// 8 x 3 ring and 48 wide registers; 0.80 with the direct flush alone; 0.73 / 0.72 with 3 CTAs (4 x 3 / 8 x 2 rings);
// **0.68** with 4 CTAs and 4 x 3 (6 x 2: 0.69, 3 x 4: 0.72); 5 CTAs at 48 registers (spills): 0.71; 512 threads x 2
// CTAs: 0.77 (profiles/pos_r2_summary.md).
#ifndef HQ_POS_CTAS
#define HQ_POS_CTAS 4
#endif
#ifndef HQ_POS_RB
#define HQ_POS_RB 4
#endif
#ifndef HQ_POS_NS
#define HQ_POS_NS 3
#endif
constexpr int POS_RB = HQ_POS_RB;              // reads per stage (developer knobs: tools/build_variants.sh)
constexpr int POS_NS = HQ_POS_NS;              // ring depth
constexpr int POS_SLOT = POS_TILE + 32;        // bytes per read and array in a stage (16-byte aligned superset)
constexpr int POS_CH = POS_SLOT / 16;          // 16-byte chunks per slot

struct pos_entry {
  uint32_t ga;     // batch offset of the 16-byte aligned superset of the tile's bytes of one read
  uint32_t gb;     // its size in bytes (multiple of 16, <= POS_SLOT) | offset of position p0 inside the superset << 16
  uint32_t L;      // read length
  uint32_t r;      // the read
};
constexpr int POS_FIFO = 128;  // descriptors; live range at any time < 4 * (NS + 2) + 32

struct pos_smem {
  uint64_t full[POS_NS];
  uint32_t total;  // number of descriptors of the slab once the scan is over, 0xFFFFFFFF before
  uint32_t pad[7];
  pos_entry fifo[POS_FIFO];
  alignas(16) uint8_t buf[POS_NS][POS_RB][2][POS_SLOT];
};

__device__ __forceinline__ void hq_cp_async16(void *dst_smem, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(hq_smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void hq_cp_async_arrive_noinc(uint64_t *bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(hq_smem_u32(bar)) : "memory");
}

// eight bytes at any byte offset of a shared-memory slot
__device__ __forceinline__ void lds8(const uint8_t *slot, uint32_t x, uint32_t &v0, uint32_t &v1) {
  const unsigned long long *w = reinterpret_cast<const unsigned long long *>(slot + (x & ~7u));
  const unsigned long long lo = w[0], hi = w[1];
  const uint32_t sh = (x & 7u) * 8u;
  const unsigned long long v = sh ? ((lo >> sh) | (hi << (64u - sh))) : lo;
  v0 = (uint32_t)v;
  v1 = (uint32_t)(v >> 32);
}

// MODE 0: the reads the mask lets through (every read without a mask).
// MODE 1: EVERY read, and the tile's share of each read's quality sum / G+C count / N count is added to a.rsum
//         (two words per read) -- the decisions are made from those sums afterwards (hq_pos_finish_kernel), so that
//         long reads are read once; MODE 2 then takes the reads that FAILED out of the tables again.
// MODE 2: the reads with mask == 0, subtracted.
template <int MODE>
__global__ void __launch_bounds__(POS_THREADS, HQ_POS_CTAS) hq_pos_kernel(hq_kargs a, uint32_t reads_per_slab) {
  extern __shared__ __align__(128) uint8_t pos_raw[];
  pos_smem &sm = *reinterpret_cast<pos_smem *>(pos_raw);
  const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const uint32_t p0 = blockIdx.x * POS_TILE;
  const uint32_t p = p0 + tid * POS_PER;
  const bool use_mask = MODE == 0 && (a.work & HPGQ_WORK_FILTER) != 0 && a.mask != nullptr;
  if (tid == 0) {
    for (int s = 0; s < POS_NS; ++s) hq_mbar_init(&sm.full[s], POS_THREADS);
    hq_mbar_init_fence();
  }

  if (tid == 0) sm.total = 0xFFFFFFFFu;
  // ---- warp 0: scan of the slab's reads, 32 per step; the lanes whose read covers the tile append its descriptor
  // to the FIFO at (appended + rank among the hits).  The idx / mask loads of the next step are already in flight.
  // MODE 2 walks the list of failed reads hq_pos_finish_kernel has written (a.flist[0] entries from a.flist[1] on)
  // instead of the batch: with few failures a CTA has next to nothing to scan
  const uint32_t n_scan = MODE == 2 ? a.flist[0] : a.n_reads;
  if (MODE == 2) {  // at least 64 reads per slab (the ring needs a few stages to hide its latency): the other CTAs leave
    const uint32_t slabs = min(gridDim.y, max(1u, n_scan / 64u));
    if (blockIdx.y >= slabs) return;
    reads_per_slab = (n_scan + slabs - 1u) / slabs;
  }
  uint32_t scan = blockIdx.y * reads_per_slab;
  const uint32_t scan_end = min(n_scan, scan + reads_per_slab);
  uint32_t appended = 0, po = 0, pe = 0, pm = 0, pr = 0;
  bool exhausted = scan >= scan_end;
  auto prefetch = [&]() {
    const uint32_t i = scan + lane;
    const bool valid = i < scan_end;
    pr = MODE == 2 ? (valid ? a.flist[1u + i] : 0u) : i;
    po = valid ? a.idx[pr] : 0u;
    pe = valid ? a.idx[pr + 1] : 0u;
    pm = valid ? (use_mask ? (uint32_t)(a.mask[pr] != 0) : 1u) : 0u;
  };
  if (warp == 0 && !exhausted) prefetch();
  auto produce = [&](uint32_t need) {
    while (appended < need && !exhausted) {
      const uint32_t o = po, L = pe - po, rd = pr;
      const bool ok = pm != 0u && L > p0 && L <= a.Lcap;
      scan += 32u;
      if (scan < scan_end) prefetch();
      else exhausted = true;
      const uint32_t bal = __ballot_sync(0xFFFFFFFFu, ok);
      if (ok) {
        const uint32_t first = o + p0, ga = first & ~15u;
        const uint32_t last = o + min(L, p0 + (uint32_t)POS_TILE);
        pos_entry e;
        e.ga = ga;
        e.gb = (((last - ga) + 15u) & ~15u) | ((first - ga) << 16);
        e.L = L;
        e.r = rd;
        sm.fifo[(appended + __popc(bal & ((1u << lane) - 1u))) % POS_FIFO] = e;
      }
      appended += __popc(bal);
    }
    if (exhausted && lane == 0) sm.total = appended;
  };
  // all threads: copies of the stage of iteration j, arriving on its barrier.  Warp w copies read w (w + 8, ...) of the
  // stage: one descriptor per warp (a broadcast load), a lane's chunks are lane, lane + 32, ... of both arrays, so a
  // copy instruction of a warp moves 512 contiguous bytes and the per-chunk index arithmetic is one add
  auto issue = [&](uint32_t j) {
    const uint32_t s = j % POS_NS, tot = sm.total;
    for (uint32_t k = warp; k < (uint32_t)POS_RB; k += POS_THREADS / 32) {
      const uint32_t ei = POS_RB * j + k;
      if (ei < tot) {
        const pos_entry e = sm.fifo[ei % POS_FIFO];
        const uint32_t gb = e.gb & 0xFFFFu;
        const uint8_t *gs = a.seq + e.ga, *gq = a.qual + e.ga;
        uint8_t *ds = sm.buf[s][k][0], *dq = sm.buf[s][k][1];
#pragma unroll
        for (uint32_t off = 16u * lane; off < 16u * (uint32_t)POS_CH; off += 512u)
          if (off < gb) {
            hq_cp_async16(ds + off, gs + off);
            hq_cp_async16(dq + off, gq + off);
          }
      }
    }
    hq_cp_async_arrive_noinc(&sm.full[s]);
  };

  uint32_t pk[5][2] = {{0u, 0u}, {0u, 0u}, {0u, 0u}, {0u, 0u}, {0u, 0u}};  // 8-bit lanes: A C G T N
  uint32_t q16[2][2] = {{0u, 0u}, {0u, 0u}};                               // 16-bit lanes: (byte0, byte2), (byte1, byte3)
  uint32_t nb[3][2] = {{0u, 0u}, {0u, 0u}, {0u, 0u}};  // 4-bit lanes (<= 15 reads): C | G << 4, A | T << 4, N
  uint32_t since = 0, nsince = 0;
  uint32_t tv[MODE == 1 ? POS_RB : 1];  // MODE 1: the thread's share of the sums of the stage's reads
  auto fold_nibbles = [&]() {
#pragma unroll
    for (int w = 0; w < 2; ++w) {
      pk[1][w] += nb[0][w] & 0x0F0F0F0Fu;
      pk[2][w] += (nb[0][w] >> 4) & 0x0F0F0F0Fu;
      pk[0][w] += nb[1][w] & 0x0F0F0F0Fu;
      pk[3][w] += (nb[1][w] >> 4) & 0x0F0F0F0Fu;
      pk[4][w] += nb[2][w];
      nb[0][w] = nb[1][w] = nb[2][w] = 0u;
    }
  };

  // 8-bit / 16-bit lanes -> the global tables (one reduction per non-zero lane), instead of 48 wide registers
  auto flush_wide = [&]() {
    unsigned long long *Cw = a.ctr;
    auto put = [&](uint32_t secw, uint32_t pos, uint32_t v) {
      if (v && pos < a.Lcap) atomicAdd(&Cw[secw + pos], MODE == 2 ? 0ull - (unsigned long long)v : (unsigned long long)v);
    };
    const uint32_t secs[5] = {a.o_A, a.o_C, a.o_G, a.o_T, a.o_N};
#pragma unroll
    for (int w = 0; w < 2; ++w) {
#pragma unroll
      for (int t = 0; t < 5; ++t) {
#pragma unroll
        for (int i = 0; i < 4; ++i) put(secs[t], p + 4 * w + i, (pk[t][w] >> (8 * i)) & 0xFFu);
        pk[t][w] = 0u;
      }
      put(a.o_accq, p + 4 * w + 0, q16[0][w] & 0xFFFFu);
      put(a.o_accq, p + 4 * w + 2, q16[0][w] >> 16);
      put(a.o_accq, p + 4 * w + 1, q16[1][w] & 0xFFFFu);
      put(a.o_accq, p + 4 * w + 3, q16[1][w] >> 16);
      q16[0][w] = q16[1][w] = 0u;
    }
  };
  if (warp == 0) produce(POS_RB * POS_NS);
  __syncthreads();  // barriers initialised, first descriptors visible
  for (uint32_t j = 0; j + 1 < POS_NS; ++j) issue(j);

  for (uint32_t it = 0;; ++it) {
    const uint32_t s = it % POS_NS;
    // descriptors of iteration it + NS (one iteration ahead of their copies), then the barrier that (a) publishes the
    // descriptors written in the previous iteration and (b) guarantees everybody is done with the stage to be refilled
    if (warp == 0) produce(POS_RB * (it + POS_NS + 1u));
    __syncthreads();
    issue(it + POS_NS - 1u);
    hq_mbar_wait(&sm.full[s], (it / POS_NS) & 1u);
    const uint32_t tot = sm.total;
    if (tot <= POS_RB * it) break;
#pragma unroll
    for (int k = 0; k < POS_RB; ++k) {
      const uint32_t ei = POS_RB * it + k;
      const pos_entry e = sm.fifo[ei % POS_FIFO];
      const uint32_t L = ei < tot ? e.L : 0u;
      if (MODE == 1) tv[k] = 0u;
      if (p < L && !(a.work & 0x8000u)) {
        // 16 bytes around the thread's eight: the byte phase (delta & 7) is the same for every thread of the CTA, so
        // the funnel is a uniform choice of three words and one PRMT per output word
        const uint32_t delta = e.gb >> 16;
        const uint32_t x = delta + tid * POS_PER, sh = delta & 7u;
        const uint2 *ws = reinterpret_cast<const uint2 *>(sm.buf[s][k][0] + (x & ~7u));
        const uint2 *wq = reinterpret_cast<const uint2 *>(sm.buf[s][k][1] + (x & ~7u));
        const uint2 s0 = ws[0], s1 = ws[1], q0 = wq[0], q1 = wq[1];
        const uint32_t fsel = 0x3210u + 0x1111u * (sh & 3u);
        uint32_t sv[2], qv[2];
        if (sh & 4u) {
          sv[0] = hq_prmt_raw(s0.y, s1.x, fsel);
          sv[1] = hq_prmt_raw(s1.x, s1.y, fsel);
          qv[0] = hq_prmt_raw(q0.y, q1.x, fsel);
          qv[1] = hq_prmt_raw(q1.x, q1.y, fsel);
        } else {
          sv[0] = hq_prmt_raw(s0.x, s0.y, fsel);
          sv[1] = hq_prmt_raw(s0.y, s1.x, fsel);
          qv[0] = hq_prmt_raw(q0.x, q0.y, fsel);
          qv[1] = hq_prmt_raw(q0.y, q1.x, fsel);
        }
        if (p + POS_PER > L) {  // the thread that holds the read's end: bytes past it become 0x00, no symbol
#pragma unroll
          for (int w = 0; w < 2; ++w) {
            const uint32_t m = hq_tail_mask(p + 4u * w, L);
            sv[w] &= m;
            qv[w] &= m;
          }
        }
#pragma unroll
        for (int w = 0; w < 2; ++w) {
          const uint32_t sw = sv[w], qm = qv[w];
          // four bases per PRMT: the low three bits of a byte select from 8-entry LUTs of 4-bit one-hots (C | G << 4,
          // A | T << 4, N) and of the expected byte; a word with a byte the LUT does not reproduce (lower case,
          // IUPAC, ...) is re-done byte by byte
          const uint32_t sel = hq_selector(hq_prmt(sw, 0u, 0x3120u));
          uint32_t cgv = hq_prmt_raw(0x01000000u, 0x10000000u, sel);
          uint32_t atv = hq_prmt_raw(0x00000100u, 0x00000010u, sel);
          uint32_t nv = hq_prmt_raw(0x00000000u, 0x00010000u, sel);
          if (__builtin_expect(hq_prmt_raw(0x43FF4100u, 0x474EFF54u, sel) != sw, 0)) {
            cgv = atv = nv = 0u;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint32_t c = (sw >> (8 * i)) & 0xFFu;
              cgv |= (c == 'C' ? 0x01u : c == 'G' ? 0x10u : 0u) << (8 * i);
              atv |= (c == 'A' ? 0x01u : c == 'T' ? 0x10u : 0u) << (8 * i);
              nv |= (c == 'N' ? 0x01u : 0u) << (8 * i);
            }
          }
          nb[0][w] += cgv;
          nb[1][w] += atv;
          nb[2][w] += nv;
          q16[0][w] += __byte_perm(qm, 0u, 0x4240u);
          q16[1][w] += __byte_perm(qm, 0u, 0x4341u);
          // the thread's share of the read's sums: quality (8 bytes: < 2^11) | G+C count << 11 | N count << 20
          if (MODE == 1) tv[k] = __dp4a(qm, 0x01010101u, tv[k]) + ((uint32_t)__popc(cgv) << 11) + ((uint32_t)__popc(nv) << 20);
        }
      }
    }
    if (MODE == 1) {
      // after the stage's reads, so that the reductions overlap: a warp covers 256 positions, its counts are <= 256
      // (9 bits each), its quality sum < 2^16.  Lane k keeps the sums of read k and adds them to the read's words.
      uint32_t mq = 0u, mc = 0u;
#pragma unroll
      for (int k = 0; k < POS_RB; ++k) {
        const uint32_t sq = __reduce_add_sync(0xFFFFFFFFu, tv[k] & 0x7FFu);
        const uint32_t sc = __reduce_add_sync(0xFFFFFFFFu, tv[k] >> 11);
        if (lane == (uint32_t)k) {
          mq = sq;
          mc = sc;
        }
      }
      const uint32_t ei = POS_RB * it + lane;
      if (lane < (uint32_t)POS_RB && ei < tot) {
        const pos_entry e = sm.fifo[ei % POS_FIFO];
        if (p0 + (warp << 8) < e.L) {
          atomicAdd(&a.rsum[2u * e.r], mq);
          atomicAdd(&a.rsum[2u * e.r + 1u], (mc & 0x1FFu) | ((mc >> 9) << 16));
        }
      }
    }
    since += POS_RB;
    nsince += POS_RB;
    if (nsince + POS_RB > 15u) {  // 4-bit lanes -> 8-bit lanes
      fold_nibbles();
      nsince = 0;
    }
    if (nsince == 0u && since + 15u > 255u) {  // before an 8-bit (or 16-bit) lane can overflow
      flush_wide();
      since = 0;
    }
  }
  fold_nibbles();
  flush_wide();
}

// Decisions and per-read statistics from the sums hq_pos_kernel<1> has collected (one thread per read); the sums are
// cleared for the next batch.  Same expressions as hq_qc_group_kernel; the options it cannot serve (out-of-quality
// count, window means, trimming) never get here (hq_pos_fused_supports).
__global__ void __launch_bounds__(256) hq_pos_finish_kernel(hq_kargs a, uint32_t stated_max_len) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31u;
  const bool do_filter = a.work & HPGQ_WORK_FILTER;
  unsigned long long *C = a.ctr;
  uint32_t n_pass = 0, n_fail = 0, n_stat = 0, qbin = 0, gbin = 0;
  bool failed = false, hist = false;
  unsigned long long acc_len = 0;
  if (r < a.n_reads) {
    const uint32_t len = a.idx[r + 1] - a.idx[r];
    // a read longer than the caller said: its tail was not covered, the batch's results are void (HPGQ_ERR_INVALID at wait)
    if (len > stated_max_len) atomicExch(&C[HPGQ_S_BAD_BATCH], 1ull);
    const uint32_t S = a.rsum[2u * r], pk = a.rsum[2u * r + 1u];
    a.rsum[2u * r] = 0u;
    a.rsum[2u * r + 1u] = 0u;
    const uint32_t nGC = pk & 0xFFFFu, nN = pk >> 16;
    bool pass = true;
    if (do_filter) {
      pass = (long long)len >= a.min_len && (long long)len <= a.max_len && mean_ok_g(S, len, a.q_lo, a.q_hi) &&
             (long long)nN <= a.max_n;
      n_pass = pass;
      n_fail = !pass;
    }
    if (a.mask) a.mask[r] = pass ? 1 : 0;
    failed = !pass;
    if (pass && len <= a.Lcap) {
      n_stat = 1;
      acc_len = len;
      atomicAdd(&C[a.o_len_hist + len], 1ull);
      atomicAdd(&C[a.o_sumq + len], (unsigned long long)S);
      hist = len != 0u;
      if (hist) {
        qbin = (uint32_t)((2ull * S + len) / (2ull * len));
        gbin = (uint32_t)((100ull * nGC) / len);
      }
    }
  }
  // the mean quality and the G+C percentage of a batch fall into a handful of bins: one atomic per bin and warp
  const unsigned hm = __ballot_sync(0xFFFFFFFFu, hist);
  if (hist) {
    const unsigned mq_ = __match_any_sync(hm, qbin), mg_ = __match_any_sync(hm, gbin);
    if (lane == (uint32_t)__ffs(mq_) - 1u) atomicAdd(&C[a.o_qual_hist + qbin], (unsigned long long)__popc(mq_));
    if (lane == (uint32_t)__ffs(mg_) - 1u) atomicAdd(&C[a.o_gc_hist + gbin], (unsigned long long)__popc(mg_));
  }
  // the failed reads, for hq_pos_kernel<2> (any order: it subtracts sums)
  const unsigned fm = __ballot_sync(0xFFFFFFFFu, failed);
  if (fm && a.flist) {
    uint32_t base = 0;
    if (lane == (uint32_t)__ffs(fm) - 1u) base = atomicAdd(&a.flist[0], (uint32_t)__popc(fm));
    base = __shfl_sync(0xFFFFFFFFu, base, __ffs(fm) - 1);
    if (failed) a.flist[1u + base + __popc(fm & ((1u << lane) - 1u))] = r;
  }
  n_pass = group_sum<32>(n_pass);
  n_fail = group_sum<32>(n_fail);
  n_stat = group_sum<32>(n_stat);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc_len += __shfl_xor_sync(0xFFFFFFFFu, acc_len, o);
  if (lane == 0) {
    if (n_pass) atomicAdd(&C[HPGQ_S_NUM_PASSED], (unsigned long long)n_pass);
    if (n_fail) atomicAdd(&C[HPGQ_S_NUM_FAILED], (unsigned long long)n_fail);
    if (n_stat) atomicAdd(&C[HPGQ_S_NUM_READS], (unsigned long long)n_stat);
    if (acc_len) atomicAdd(&C[HPGQ_S_ACC_LENGTH], acc_len);
  }
}

}  // namespace

cudaError_t hq_launch_qc_group(const hq_kargs &a, uint32_t max_len_in_batch, int sm_count, cudaStream_t st,
                               uint64_t *launches) {
  if (a.n_reads == 0) return cudaSuccess;
  // lanes per read: one 16-byte chunk per lane covers the longest read (at any byte phase) when it is short
  // enough -> register-resident kernel; longer reads loop over chunks with a whole warp per read
  const uint32_t span = max_len_in_batch + 15u;  // bytes from the 16-byte boundary below the read start to its end
  const int G = span <= 4u * RG_LB ? 4 : span <= 8u * RG_LB ? 8 : span <= 16u * RG_LB ? 16 : 32;
  const bool regs = span <= 32u * RG_LB;
  const uint32_t reads_per_block = (256u / (uint32_t)G) * (regs ? 2u : 1u);
  uint32_t blocks = (a.n_reads + reads_per_block - 1u) / reads_per_block;
  const uint32_t cap = (uint32_t)sm_count * 8u * 4u;
  if (blocks > cap) blocks = cap;
  if (regs) {
    const bool gcs = (a.work & HPGQ_WORK_STATS) != 0;
    const bool ooq = (a.work & HPGQ_WORK_FILTER) != 0 && a.max_ooq < 32 * RG_LB;  // else no read here can exceed it
#define HQ_REGS(GG)                                                                          \
  do {                                                                                       \
    if (gcs && ooq) hq_qc_regs_kernel<GG, true, true><<<blocks, 256, 0, st>>>(a);            \
    else if (gcs) hq_qc_regs_kernel<GG, true, false><<<blocks, 256, 0, st>>>(a);             \
    else if (ooq) hq_qc_regs_kernel<GG, false, true><<<blocks, 256, 0, st>>>(a);             \
    else hq_qc_regs_kernel<GG, false, false><<<blocks, 256, 0, st>>>(a);                     \
  } while (0)
    switch (G) {
      case 4: HQ_REGS(4); break;
      case 8: HQ_REGS(8); break;
      case 16: HQ_REGS(16); break;
      default: HQ_REGS(32); break;
    }
#undef HQ_REGS
  } else {
    hq_qc_group_kernel<32><<<blocks, 256, 0, st>>>(a);
  }
  ++*launches;
  return cudaGetLastError();
}

namespace {
template <int MODE>
cudaError_t launch_pos_mode(const hq_kargs &a, uint32_t max_len_in_batch, int sm_count, cudaStream_t st, uint64_t *launches) {
  const uint32_t cover = max_len_in_batch < a.Lcap ? max_len_in_batch : a.Lcap;
  const uint32_t tiles = (cover + POS_TILE - 1u) / POS_TILE;
  // HQ_POS_CTAS CTAs are resident per SM; `waves` grid waves (HPGQ_POS_WAVES, developer knob): the tiles near the
  // longest reads see few reads and finish early, and every CTA ends with one reduction per non-zero lane
  const char *wv = getenv("HPGQ_POS_WAVES");
  const uint32_t waves = wv && atoi(wv) > 0 ? (uint32_t)atoi(wv) : 6u;  // 2: 0.76, 3: 0.70, 4: 0.69, 6: 0.67, 8: 0.69 ms per C5 batch
  uint32_t slabs = ((uint32_t)sm_count * waves * (uint32_t)HQ_POS_CTAS) / tiles;
  // a thread empties its packed lanes every ~240 reads: any slab size is safe
  if (slabs > a.n_reads) slabs = a.n_reads;
  if (slabs < 1u) slabs = 1u;
  if (slabs > 65535u) slabs = 65535u;
  const uint32_t reads_per_slab = (a.n_reads + slabs - 1u) / slabs;
  hq_kargs a2 = a;
  if (getenv("HPGQ_DEBUG_LOADONLY")) a2.work |= 0x8000u;  // developer probe: ring traffic without the arithmetic
  // per device, so set on every launch (one process may drive several GPUs)
  cudaError_t e = cudaFuncSetAttribute(hq_pos_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(pos_smem));
  if (e != cudaSuccess) return e;
  hq_pos_kernel<MODE><<<dim3(tiles, slabs), POS_THREADS, sizeof(pos_smem), st>>>(a2, reads_per_slab);
  ++*launches;
  return cudaGetLastError();
}
}  // namespace

cudaError_t hq_launch_pos(const hq_kargs &a, uint32_t max_len_in_batch, int sm_count, cudaStream_t st,
                          uint64_t *launches) {
  if (a.n_reads == 0 || max_len_in_batch == 0) return cudaSuccess;
  return launch_pos_mode<0>(a, max_len_in_batch, sm_count, st, launches);
}

// Statistics (+ filter) of long reads with ONE pass over the batch: per-position tables of every read and the per-read
// sums (MODE 1), decisions and per-read histograms from the sums, then the failed reads are taken out of the tables
// again (MODE 2 reads only those).  Exact: the tables are integer sums.  Needs the per-read counts to fit 16 bits, every
// read inside the tables, and a filter made of length, mean quality and N count only.
bool hq_pos_fused_supports(const hq_kargs &a, uint32_t max_len_in_batch) {
  const bool off = getenv("HPGQ_NO_POSFUSE") != nullptr;  // developer A/B: decision kernel + masked position pass
  if (off || !a.rsum || !a.flist || !(a.work & HPGQ_WORK_STATS) || (a.work & HPGQ_WORK_EDIT)) return false;
  if (max_len_in_batch == 0 || max_len_in_batch > 65535u || max_len_in_batch > a.Lcap) return false;
  if (a.work & HPGQ_WORK_FILTER) {
    if (a.f_ll > 0 || a.f_rl > 0) return false;
    if ((long long)a.max_ooq < (long long)max_len_in_batch) return false;  // some read could fail on its out-of-quality count
    if (!a.mask) return false;
  }
  return true;
}

cudaError_t hq_launch_pos_fused(const hq_kargs &a, uint32_t max_len_in_batch, int sm_count, cudaStream_t st,
                                uint64_t *launches) {
  if (a.n_reads == 0) return cudaSuccess;
  cudaError_t e = launch_pos_mode<1>(a, max_len_in_batch, sm_count, st, launches);
  if (e != cudaSuccess) return e;
  if (a.work & HPGQ_WORK_FILTER) {
    if ((e = cudaMemsetAsync(a.flist, 0, 4, st)) != cudaSuccess) return e;
  }
  hq_pos_finish_kernel<<<(a.n_reads + 255u) / 256u, 256, 0, st>>>(a, max_len_in_batch);
  ++*launches;
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  if (a.work & HPGQ_WORK_FILTER) e = launch_pos_mode<2>(a, max_len_in_batch, sm_count, st, launches);
  return e;
}
