// hpgq_simple.cu -- warp-per-read QC kernel for reads of ANY length (long-read path and the
// cross-check twin of the tiled kernel), synthetic generator kernel and small utilities.
//
// Replaces, for one SoA batch: fastq_edit + fastq_filter + fastq_reads_stats and the consumer merge
// (reference src/edit_fastq.c:142-178, src/filter_fastq.c:134-155, src/stats_fastq.c:202-251,257-417).
#include "hpgq_dev.cuh"
#include "hpgq_internal.h"
#include "../../include/hpgq_synth.h"

namespace {

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
  return v;
}

__device__ __forceinline__ bool mean_ok(uint32_t S, uint32_t w, int32_t lo, int32_t hi) {
  const long long s = (long long)S, ww = (long long)w;
  return s >= (long long)lo * ww && s <= (long long)hi * ww;
}

__global__ void __launch_bounds__(256) hq_qc_simple_kernel(hq_kargs a) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
  const bool do_stats = a.work & HPGQ_WORK_STATS, do_filter = a.work & HPGQ_WORK_FILTER;
  const bool do_edit = a.work & HPGQ_WORK_EDIT;
  unsigned long long *C = a.ctr;
  uint32_t n_pass = 0, n_fail = 0, n_edit = 0, n_stat = 0;
  unsigned long long acc_len = 0;

  for (uint32_t r = warp; r < a.n_reads; r += nwarps) {
    const uint32_t o = a.idx[r];
    const uint32_t L0 = a.idx[r + 1] - o;
    uint32_t start = 0, len = L0;
    const uint8_t *q0 = a.qual + o;
    if (do_edit) {
      bool edited = false;
      if (a.e_ll > 0) {
        const uint32_t w = min((uint32_t)a.e_ll, len);
        uint32_t sw = 0;
        for (uint32_t j = lane; j < w; j += 32) sw += q0[start + j];
        sw = warp_sum(sw);
        if (w && !mean_ok(sw, w, a.e_lq_lo, a.e_lq_hi)) {
          start += w;
          len -= w;
          edited = true;
        }
      }
      if (a.e_rl > 0) {
        const uint32_t w = min((uint32_t)a.e_rl, len);
        uint32_t sw = 0;
        for (uint32_t j = lane; j < w; j += 32) sw += q0[start + len - w + j];
        sw = warp_sum(sw);
        if (w && !mean_ok(sw, w, a.e_rq_lo, a.e_rq_hi)) {
          len -= w;
          edited = true;
        }
      }
      n_edit += edited;
    }
    if (a.trim && lane == 0) a.trim[r] = hpgq_trim_t{start, len};
    const uint8_t *s = a.seq + o + start, *q = q0 + start;

    uint32_t S = 0, nN = 0, nGC = 0, ooq = 0, SL = 0, SR = 0;
    const uint32_t wl = a.f_ll > 0 ? min((uint32_t)a.f_ll, len) : 0u;
    const uint32_t wr = a.f_rl > 0 ? min((uint32_t)a.f_rl, len) : 0u;
    for (uint32_t j = lane; j < len; j += 32) {
      const uint32_t c = s[j], qq = q[j];
      S += qq;
      nN += (c == 'N');
      nGC += (c == 'G') | (c == 'C');
      ooq += ((int32_t)qq < a.q_lo) | ((int32_t)qq > a.q_hi);
      if (j < wl) SL += qq;
      if (j >= len - wr) SR += qq;
    }
    S = warp_sum(S);
    nN = warp_sum(nN);
    nGC = warp_sum(nGC);
    bool pass = true;
    if (do_filter) {
      ooq = warp_sum(ooq);
      SL = warp_sum(SL);
      SR = warp_sum(SR);
      pass = (long long)len >= a.min_len && (long long)len <= a.max_len && mean_ok(S, len, a.q_lo, a.q_hi) &&
             (long long)ooq <= a.max_ooq && (long long)nN <= a.max_n;
      if (wl) pass = pass && mean_ok(SL, wl, a.f_lq_lo, a.f_lq_hi);
      if (wr) pass = pass && mean_ok(SR, wr, a.f_rq_lo, a.f_rq_hi);
      n_pass += pass;
      n_fail += !pass;
    }
    if (a.mask && lane == 0) a.mask[r] = pass ? 1 : 0;
    if (do_stats && pass && len <= a.Lcap) {
      n_stat++;
      acc_len += len;
      if (lane == 0) {
        atomicAdd(&C[a.o_len_hist + len], 1ull);
        atomicAdd(&C[a.o_sumq + len], (unsigned long long)S);
        if (len) {
          atomicAdd(&C[a.o_qual_hist + (2u * S + len) / (2u * len)], 1ull);
          atomicAdd(&C[a.o_gc_hist + (100u * nGC) / len], 1ull);
        }
      }
      for (uint32_t j = lane; j < len; j += 32) {
        const uint32_t c = s[j];
        atomicAdd(&C[a.o_accq + j], (unsigned long long)q[j]);
        uint32_t sec = 0;
        if (c == 'A') sec = a.o_A;
        else if (c == 'C') sec = a.o_C;
        else if (c == 'G') sec = a.o_G;
        else if (c == 'T') sec = a.o_T;
        else if (c == 'N') sec = a.o_N;
        if (sec) atomicAdd(&C[sec + j], 1ull);
      }
    }
  }
  if (lane == 0) {
    if (n_pass) atomicAdd(&C[HPGQ_S_NUM_PASSED], (unsigned long long)n_pass);
    if (n_fail) atomicAdd(&C[HPGQ_S_NUM_FAILED], (unsigned long long)n_fail);
    if (n_edit) atomicAdd(&C[HPGQ_S_NUM_EDITED], (unsigned long long)n_edit);
    if (n_stat) atomicAdd(&C[HPGQ_S_NUM_READS], (unsigned long long)n_stat);
    if (acc_len) atomicAdd(&C[HPGQ_S_ACC_LENGTH], acc_len);
  }
}

// ---------------------------------------------------------------- k-mers + chaos game (first version)
// warp per read, each lane owns a contiguous chunk and warms its rolling state up over a look-back.
// k-mer ids: A=0 C=1 G=2 T=3, first base most significant, windows with any other byte skipped
// (SURVEY.md 8c(5)); chaos game = integer-index form of old/chaos_game.c:165-267.
constexpr uint32_t HQ_LOOKBACK = 20;
// packed (count, quality sum) cells: count << 37 | sum.  A sub-launch covers < 2^27 words of < 2^10 each, so neither
// field can overflow (2 x 27 + 10 = 64 bits); with 12-bit words the bound was 2^26 and a 2 M x 250 bp batch took eight
// sub-launches, each followed by an unpack pass over the 32 MB table, instead of four
constexpr int HQ_PK_SHIFT = 37;
constexpr uint32_t HQ_PK_QMAX = 1024u;
constexpr unsigned long long HQ_PK_WORDS = 1ull << 27;

// One chaos-game word: table_seq[cell]++ and table_q[cell] += qval (old/chaos_game.c:253-259).  With a packed
// shadow table both land in ONE 64-bit reduction, count in the top 27 bits and the quality sum in the low 37
// (a sub-launch covers < 2^27 words and qval < 2^10, so neither field can overflow); hq_cg_unpack_kernel adds the
// fields to the real tables afterwards.  Values outside [0, 1024) (qualities below the base quality wrap, as the
// reference's unsigned arithmetic does) go to the real tables directly.
__device__ __forceinline__ void cg_add(const hq_kargs &a, uint64_t cell, uint32_t qval) {
  unsigned long long *C = a.ctr;
  if (a.cg_pack && qval < HQ_PK_QMAX) {
    atomicAdd(&a.cg_pack[cell], (1ull << HQ_PK_SHIFT) | (unsigned long long)qval);
  } else {
    atomicAdd(&C[a.o_cg_seq + cell], 1ull);
    atomicAdd(&C[a.o_cg_q + cell], (unsigned long long)qval);
  }
}

__global__ void __launch_bounds__(256) hq_cg_unpack_kernel(hq_kargs a, uint32_t cells) {
  unsigned long long *C = a.ctr;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < cells; i += gridDim.x * blockDim.x) {
    const unsigned long long v = a.cg_pack[i];
    if (v) {
      C[a.o_cg_seq + i] += v >> HQ_PK_SHIFT;
      C[a.o_cg_q + i] += v & ((1ull << HQ_PK_SHIFT) - 1ull);
      a.cg_pack[i] = 0ull;
    }
  }
}

// Packed k2-mer / chaos-game table -> K2, CG_SEQ, CG_Q.  Cell c (a k2-mer, first base most significant) holds
// count << 37 | quality sum of the words that ended with it; the chaos-game cell of those words is
// co_x = sum xbit(base_i) 2^i, co_y = sum ybit(base_i) 2^i over the last cg_k bases, oldest = bit 0
// (old/chaos_game.c:200-228: A, T move x up; G, T move y up).
__global__ void __launch_bounds__(256) hq_kc_unpack_kernel(hq_kargs a, uint32_t cells) {
  // a thread owns one chaos-game cell = the 4^(k2 - cg_k) k2-mers that end with the same cg_k bases: no atomics
  unsigned long long *C = a.ctr;
  const int ck = a.cg_k;
  const uint32_t low = 1u << (2 * ck), tops = cells >> (2 * ck);
  for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < low; c += gridDim.x * blockDim.x) {
    unsigned long long cnt = 0, qs = 0;
    for (uint32_t t = 0; t < tops; ++t) {
      const uint32_t code = t * low + c;
      const unsigned long long v = a.kc_pack[code];
      if (!v) continue;
      a.kc_pack[code] = 0ull;
      C[a.o_k2 + code] += v >> HQ_PK_SHIFT;
      cnt += v >> HQ_PK_SHIFT;
      qs += v & ((1ull << HQ_PK_SHIFT) - 1ull);
    }
    if (!cnt) continue;
    uint32_t ix = 0, iy = 0;
    for (int i = 0; i < ck; ++i) {
      const uint32_t b = (c >> (2 * (ck - 1 - i))) & 3u;
      ix |= (uint32_t)(b == 0u || b == 3u) << i;
      iy |= (b >> 1) << i;
    }
    const uint64_t cell = ((uint64_t)ix << ck) + iy;
    C[a.o_cg_seq + cell] += cnt;
    C[a.o_cg_q + cell] += qs;
  }
}

__device__ __forceinline__ int base2(uint32_t c) {
  return c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : c == 'T' ? 3 : -1;
}

// MODE 0: every base checked against [jb, ce) and cb (the words at the chunk's ends and the one that holds cb);
// MODE 1: whole word inside the look-back (state only); MODE 2: whole word inside [cb, ce).
struct kc_state {
  uint32_t code, valid, X, Y, words, longruns;
  int32_t P;
};

template <int MODE>
__device__ __forceinline__ void kc_fast_word(const hq_kargs &a, kc_state &st, uint32_t r, uint32_t w, uint32_t o, uint32_t jb,
                                             uint32_t cb, uint32_t ce, int k2, int ck, uint32_t k2mask, uint32_t qoff,
                                             int32_t (*ring)[256]) {
  const uint32_t tid = threadIdx.x;
  const uint32_t s4 = __ldg(reinterpret_cast<const uint32_t *>(a.seq) + w);
  const uint32_t q4 = __ldg(reinterpret_cast<const uint32_t *>(a.qual) + w);
  const uint32_t sel = hq_selector(hq_prmt(s4, 0u, 0x3120u));
  const uint32_t b2w = hq_prmt_raw(0x01000000u, 0x02000003u, sel);       // A 0, C 1, G 2, T 3
  const uint32_t x = s4 ^ hq_prmt_raw(0x43FF41FFu, 0x47FFFF54u, sel);    // zero byte <=> one of A C G T
  const uint32_t zf = ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u;
  unsigned long long *C = a.ctr;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t j = 4u * w + (uint32_t)i - o;  // position in the read (wraps for the bytes in front of it)
    if (MODE == 0 && j - jb >= ce - jb) continue; // outside [jb, ce)
    // signed byte i of the quality word (the reference adds `char`s)
    const int32_t qb = (int32_t)hq_prmt_raw(q4, 0u, (uint32_t)(i | ((8 | i) << 4) | ((8 | i) << 8) | ((8 | i) << 12)));
    st.P += qb;
    int32_t pold = 0;
    if (MODE != 1) {
      pold = ring[(j - (uint32_t)ck) & 15u][tid];
      if (j < jb + (uint32_t)ck) pold = 0;
    }
    ring[j & 15u][tid] = st.P;
    const bool ok = ((zf >> (8 * i + 7)) & 1u) != 0u;
    const uint32_t b = (b2w >> (8 * i)) & 3u;
    // state by selection, not by branch
    st.code = ok ? ((st.code << 2) | b) : 0u;
    st.valid = ok ? st.valid + 1u : 0u;
    st.X = ok ? ((st.X << 1) | (~(b ^ (b >> 1)) & 1u)) : st.X;  // A or T
    st.Y = ok ? ((st.Y << 1) | (b >> 1)) : st.Y;                // G or T
    if (MODE == 1) continue;
    const bool emit = ok && (MODE == 2 || j >= cb);
    const uint32_t qval = (uint32_t)(st.P - pold) - qoff;
    const bool kev = st.valid >= (uint32_t)k2, cev = st.valid >= (uint32_t)ck;
    if (emit && kev && cev && qval < HQ_PK_QMAX) {
      atomicAdd(&a.kc_pack[st.code & k2mask], (1ull << HQ_PK_SHIFT) | (unsigned long long)qval);
      st.words++;
    } else if (emit && (kev || cev)) {
      if (kev) atomicAdd(&C[a.o_k2 + (st.code & k2mask)], 1ull);
      if (cev) {  // the cell from the last cg_k bases: newest base in the top bit
        uint32_t ix = 0, iy = 0;
        for (int t = 0; t < ck; ++t) {
          const uint32_t bb = (st.code >> (2 * t)) & 3u;
          ix |= (~(bb ^ (bb >> 1)) & 1u) << (ck - 1 - t);
          iy |= (bb >> 1) << (ck - 1 - t);
        }
        cg_add(a, ((uint64_t)ix << ck) + iy, qval);
        st.words++;
      }
    }
    // an axis run that reaches 21 bases
    if (emit && ((st.X & 0x3FFFFFu) == 0x1FFFFFu || (st.Y & 0x3FFFFFu) == 0x1FFFFFu)) {
      st.longruns++;
      if (a.cgflag) a.cgflag[r] = 1;
    }
  }
}

// Word-wise form of the rolling loop below for the common case: k-mers and chaos game together (one packed reduction
// per base), a read of A, C, G, T, N only.  Four bases per 32-bit load and PRMT lookup (2-bit codes, validity flags);
// the quality sum of the last cg_k bases is a difference of running sums, the older one taken from a 16-entry ring of
// the lane in shared memory; the two axis-run counters are shift registers of the x and y bits; the rolling state is
// updated by selection.  Same look-back, same events as the byte loop (which spends ~145 warp instructions per base
// step and held the kernel at 60 % of the L2 reduction rate, profiles/launches_c4_r2.md).
__device__ __forceinline__ void kc_fast_chunk(const hq_kargs &a, uint32_t r, uint32_t o, uint32_t cb, uint32_t ce, int k2,
                                              int ck, uint32_t k2mask, int32_t (*ring)[256], uint32_t &words_out,
                                              uint32_t &longruns_out) {
  const uint32_t jb = cb > HQ_LOOKBACK ? cb - HQ_LOOKBACK : 0u;
  const uint32_t qoff = (uint32_t)(a.cg_base_quality * ck);
  kc_state st = {0u, 0u, 0u, 0u, 0u, 0u, 0};
  // words: [w0, wa) generic (the partial first word), [wa, wc) look-back only, [wc, wd) generic (the word holding cb),
  // [wd, we) emit, [we, w1] generic (the partial last word)
  const uint32_t w0 = (o + jb) >> 2, w1 = (o + ce - 1u) >> 2;
  const uint32_t wa = min(w1 + 1u, (o + jb + 3u) >> 2);               // first word that lies wholly at or behind jb
  const uint32_t wc = max(wa, min(w1 + 1u, (o + cb) >> 2));           // word that holds cb
  const uint32_t wd = min(w1 + 1u, max(wc, (o + cb + 3u) >> 2));      // first word wholly at or behind cb
  const uint32_t we = max(wd, (o + ce) >> 2);                         // first word not wholly in front of ce
  uint32_t w = w0;
  for (; w < wa; ++w) kc_fast_word<0>(a, st, r, w, o, jb, cb, ce, k2, ck, k2mask, qoff, ring);
  for (; w < wc; ++w) kc_fast_word<1>(a, st, r, w, o, jb, cb, ce, k2, ck, k2mask, qoff, ring);
  for (; w < wd; ++w) kc_fast_word<0>(a, st, r, w, o, jb, cb, ce, k2, ck, k2mask, qoff, ring);
  for (; w < we; ++w) kc_fast_word<2>(a, st, r, w, o, jb, cb, ce, k2, ck, k2mask, qoff, ring);
  for (; w <= w1; ++w) kc_fast_word<0>(a, st, r, w, o, jb, cb, ce, k2, ck, k2mask, qoff, ring);
  words_out += st.words;
  longruns_out += st.longruns;
}

// G lanes share one read (G = 4 for short reads: the chunk of a lane is then much longer than the look-back that
// warms its rolling state up; G = 32 for long reads).
template <int G>
__global__ void __launch_bounds__(256) hq_kmers_cg_kernel(hq_kargs a) {
  constexpr uint32_t NG = 32u / G;
  const uint32_t lane = threadIdx.x & 31u, lig = lane % G, giw = lane / G;
  const uint32_t gmask = (G == 32 ? 0xFFFFFFFFu : ((1u << G) - 1u)) << (giw * G);
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
  const bool do_k = (a.work & HPGQ_WORK_KMERS) != 0, do_cg = (a.work & HPGQ_WORK_CHAOS) != 0 && a.cg_k > 0;
  const bool use_mask = (a.work & HPGQ_WORK_FILTER) != 0 && a.mask != nullptr;
  const int k2 = a.k2, ck = a.cg_k;
  const uint32_t k2mask = k2 > 0 ? ((k2 >= 16) ? 0xFFFFFFFFu : ((1u << (2 * k2)) - 1u)) : 0u;
  unsigned long long *C = a.ctr;
  unsigned long long words = 0, longruns = 0;
  const bool merged = a.kc_pack != nullptr && do_k && do_cg && k2 >= ck;
  __shared__ int32_t kc_ring[16][256];  // kc_fast_chunk: running quality sums of the lane's last 16 bases

  for (uint32_t r0 = warp * NG; r0 < a.n_reads; r0 += nwarps * NG) {  // warp-uniform trip count (ballots below)
    const uint32_t r = r0 + giw;
    const bool live = r < a.n_reads;
    const uint32_t o = live ? a.idx[r] : 0u, L = live ? a.idx[r + 1] - o : 0u;
    const uint8_t *s = a.seq + o, *q = a.qual + o;
    const bool kmers_here = do_k && live && (!use_mask || a.mask[r]) && L <= a.Lcap;
    // does the read hold bytes other than A,C,G,T,N?  (changes the chaos-game window arithmetic)
    bool other = false, single_x = true, single_y = true;
    if (L) {
      // word-wise: the lanes of the group take the aligned words of the read in turn; per word one PRMT lookup gives
      // "is A/C/G/T" (expected-byte compare) and the two flags that break an axis (C or G: not x; A or C: not y)
      const uint32_t *sw = reinterpret_cast<const uint32_t *>(a.seq);
      const uint32_t wl = (o + L - 1u) >> 2;
      uint32_t f_other = 0u, f_axis = 0u;
      for (uint32_t w = (o >> 2) + lig; w <= wl; w += G) {
        const uint32_t s4 = __ldg(sw + w);
        // bytes of the word inside [o, o + L)
        const uint32_t m = hq_tail_mask(4u * w, o + L) & ~hq_tail_mask(4u * w, o);
        const uint32_t sel = hq_selector(hq_prmt(s4, 0u, 0x3120u));
        const uint32_t x = s4 ^ hq_prmt_raw(0x43FF41FFu, 0x47FFFF54u, sel);   // zero byte <=> one of A C G T
        const uint32_t acgt = ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u;
        const uint32_t xn = s4 ^ 0x4E4E4E4Eu;
        const uint32_t isn = ~(((xn & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | xn) & 0x80808080u;
        f_other |= ~(acgt | isn) & 0x80808080u & m;
        // A (index 1): 0x02, C (3): 0x03, T (4): 0x00, G (7): 0x01
        f_axis |= hq_prmt_raw(0x03000200u, 0x01000000u, sel) & (acgt >> 7) * 3u & m;
      }
      other = f_other != 0u;
      single_x = (f_axis & 0x01010101u) == 0u;
      single_y = (f_axis & 0x02020202u) == 0u;
    }
    other = (__ballot_sync(0xFFFFFFFFu, other) & gmask) != 0u;
    single_x = (__ballot_sync(0xFFFFFFFFu, single_x) & gmask) == gmask;
    single_y = (__ballot_sync(0xFFFFFFFFu, single_y) & gmask) == gmask;
    // reads that may take the reference's double arithmetic off the integer form (hq_cg_fix_kernel re-emulates them):
    // an axis run of >= 21 bases inside the read, or a read that lies on one axis throughout (runs span reads)
    if (do_cg && lig == 0 && L != 0u && (single_x || single_y)) {
      longruns++;
      if (a.cgflag) a.cgflag[r] = 1;
    }

    const uint32_t chunk = (L + (uint32_t)G - 1u) / (uint32_t)G;
    const uint32_t cb = lig * chunk, ce = min(L, cb + chunk);
    // HPGQ_KC_BYTES (hq_kargs.kc_bytes): developer A/B and test switch, keeps the byte loop for every read
    const bool fast = merged && kmers_here && !other && k2 > 0 && ck <= 15 && !a.kc_bytes;
    if (cb < ce && fast) {
      uint32_t w32 = 0, l32 = 0;
      kc_fast_chunk(a, r, o, cb, ce, k2, ck, k2mask, kc_ring, w32, l32);
      words += w32;
      longruns += l32;
    } else if (cb < ce) {
      const uint32_t jb = cb > HQ_LOOKBACK ? cb - HQ_LOOKBACK : 0u;
      uint32_t code = 0, valid = 0; // k-mer rolling state
      uint32_t ix = 0, iy = 0, cnt = 0, accq = 0, run_x = 0, run_y = 0;
      for (uint32_t j = jb; j < ce; ++j) {
        const uint32_t c = s[j];
        const int b = base2(c);
        if (b >= 0) {
          code = (code << 2) | (uint32_t)b;
          valid++;
          const uint32_t xb = (b == 0 || b == 3), yb = (b == 2 || b == 3);
          if (do_cg && !other) {
            ix = (ix >> 1) | (xb << (ck - 1));
            iy = (iy >> 1) | (yb << (ck - 1));
            accq += (uint32_t)(int32_t)(signed char)q[j];
            if (cnt >= (uint32_t)ck) accq -= (uint32_t)(int32_t)(signed char)q[j - ck];
            cnt++;
          }
          run_x = xb ? run_x + 1 : 0;
          run_y = yb ? run_y + 1 : 0;
        } else {
          valid = 0;
          code = 0;
          if (c == 'N') {
            cnt = 0;
            accq = 0;
          }
        }
        if (j < cb) continue;
        if (b >= 0) {
          // One 64-bit RED per base when k-mers and chaos game run together: the chaos-game cell of a word is a bit
          // permutation of the low 2 cg_k bits of the k2-mer that ends at the same base, so the k2-mer count and the
          // word's quality sum travel as (1 << 37 | qsum) into ONE packed table indexed by the k2-mer;
          // hq_kc_unpack_kernel splits it into K2, CG_SEQ and CG_Q.  The two tables were 2 REDs per base, and the
          // kernel runs at the L2 atomic rate (~189 G RED/s).
          const uint32_t qval = accq - (uint32_t)(a.cg_base_quality * ck);
          const bool kev = kmers_here && k2 > 0 && valid >= (uint32_t)k2;
          const bool cev = do_cg && !other && cnt >= (uint32_t)ck;
          if (merged && kev && cev && qval < HQ_PK_QMAX) {
            atomicAdd(&a.kc_pack[code & k2mask], (1ull << HQ_PK_SHIFT) | (unsigned long long)qval);
            words++;
          } else {
            // 5-mer x start position: hq_k5pos_kernel (shared-memory slices); the 5-mer totals are the row sums of
            // that table, derived at download (hpgq_counters_finalize_host)
            if (kev) atomicAdd(&C[a.o_k2 + (code & k2mask)], 1ull);
            if (cev) {
              cg_add(a, ((uint64_t)ix << ck) + iy, qval);
              words++;
            }
          }
          if (do_cg && !other) {
            if (run_x == 21u || run_y == 21u) {
              longruns++;
              if (a.cgflag) a.cgflag[r] = 1;
            }
          }
        }
      }
    }
    // reads with other bytes: literal sequential emulation of the reference's positional window by lane 0
    if (do_cg && other && lig == 0) {
      uint32_t ix = 0, iy = 0, accq = 0, run_x = 0, run_y = 0;
      int cnt = 0;
      for (uint32_t j = 0; j < L; ++j) {
        const uint32_t c = s[j];
        const int b = base2(c);
        if (b >= 0) {
          const uint32_t xb = (b == 0 || b == 3), yb = (b == 2 || b == 3);
          ix = (ix >> 1) | (xb << (ck - 1));
          iy = (iy >> 1) | (yb << (ck - 1));
          run_x = xb ? run_x + 1 : 0;
          run_y = yb ? run_y + 1 : 0;
          if (run_x == 21u || run_y == 21u) {
            longruns++;
            if (a.cgflag) a.cgflag[r] = 1;
          }
          cnt++;
          accq += (uint32_t)(int32_t)(signed char)q[j];
        } else if (c == 'N') {
          cnt = 0;
          accq = 0;
        }
        if (cnt == ck) {
          cg_add(a, ((uint64_t)ix << ck) + iy, accq - (uint32_t)(a.cg_base_quality * ck));
          words++;
          cnt--;
          accq -= (uint32_t)(int32_t)(signed char)q[j + 1 - ck];
        }
      }
    }
  }
  words = words + __shfl_xor_sync(0xFFFFFFFFu, words, 16);
  words = words + __shfl_xor_sync(0xFFFFFFFFu, words, 8);
  words = words + __shfl_xor_sync(0xFFFFFFFFu, words, 4);
  words = words + __shfl_xor_sync(0xFFFFFFFFu, words, 2);
  words = words + __shfl_xor_sync(0xFFFFFFFFu, words, 1);
  longruns = longruns + __shfl_xor_sync(0xFFFFFFFFu, longruns, 16);
  longruns = longruns + __shfl_xor_sync(0xFFFFFFFFu, longruns, 8);
  longruns = longruns + __shfl_xor_sync(0xFFFFFFFFu, longruns, 4);
  longruns = longruns + __shfl_xor_sync(0xFFFFFFFFu, longruns, 2);
  longruns = longruns + __shfl_xor_sync(0xFFFFFFFFu, longruns, 1);
  if (lane == 0) {
    if (words) atomicAdd(&C[HPGQ_S_CG_WORDS], words);
    if (longruns) atomicAdd(&C[HPGQ_S_CG_LONG_RUNS], longruns);
  }
}

// ---------------------------------------------------------------- chaos game: exact double arithmetic where it matters
// The reference carries (f_x, f_y) as doubles across all reads of a batch (old/chaos_game.c:180-181,200-251).  Its
// cell index equals the integer form used above except after an axis run of >= 53 - k bases, where the rounding of
// f + (dim - f) * 0.5 carries through the run (SURVEY.md Appendix C probe 1), and where f reaches dim exactly while
// no word completes (the EPSILON clamp of :243-251 is only applied at word completions).  One thread per FLAGGED read
// re-runs the reference loop literally: it warms the double state up over the >= 128 bases in front of the read from
// the two extreme states 0 and dim -- the trajectories contract by half per base and must have met bit for bit at
// the read's first base, else the warm-up is lengthened, down to the batch start where the state is dim / 2 exactly --
// then walks the read (and the unflagged reads behind it while a run's tail is still within k + 1 bases) with the
// double and the integer index side by side and moves every word whose cells differ.
__device__ __forceinline__ double cg_up(double f, double dim) { return __dadd_rn(f, __dmul_rn(__dsub_rn(dim, f), 0.5)); }

__global__ void __launch_bounds__(128) hq_cg_fix_kernel(hq_kargs a, const uint8_t *flag) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= a.n_reads || !flag[r]) return;
  const int k = a.cg_k, dim = 1 << k;
  const double dimd = (double)dim, EPS = 0.00001;  // old/chaos_game.h:41
  const uint32_t qsub = (uint32_t)(a.cg_base_quality * k);
  unsigned long long *C = a.ctr;
  uint32_t patched = 0;
  for (uint32_t W = 128u;; W = W >= (1u << 24) ? 0xFFFFFFFFu : W * 8u) {
    uint32_t rs = r, have = 0;
    while (rs > 0u && have < W) {
      --rs;
      have += a.idx[rs + 1] - a.idx[rs];
    }
    const bool exact = rs == 0u;  // batch start: f = dim / 2 (old/chaos_game.c:107-108 via :180-181)
    double fx[2], fy[2];
    fx[0] = fy[0] = exact ? dimd * 0.5 : 0.0;
    fx[1] = fy[1] = exact ? dimd * 0.5 : dimd;
    uint32_t ix = 0, iy = 0, acc = 0, run_x = 0, run_y = 0, since = 0xFFFFu;
    int cnt = 0;
    bool patching = false, redo = false;
    for (uint32_t i = rs; i < a.n_reads; ++i) {
      if (i == r) {
        if (fx[0] != fx[1] || fy[0] != fy[1]) {
          redo = true;  // the two warm-up trajectories have not met yet: start further upstream
          break;
        }
        patching = true;
      } else if (i > r && (flag[i] || since > (uint32_t)k + 1u)) {
        break;  // the next flagged read has its own thread; or no run's tail reaches into this read
      }
      const uint32_t o = a.idx[i], L = a.idx[i + 1] - o;
      for (uint32_t j = 0; j < L; ++j) {
        const uint32_t c = a.seq[o + j];
        const int b = base2(c);
        if (b >= 0) {
          const uint32_t xb = (b == 0 || b == 3), yb = (b == 2 || b == 3);
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            fx[t] = xb ? cg_up(fx[t], dimd) : __dmul_rn(fx[t], 0.5);
            fy[t] = yb ? cg_up(fy[t], dimd) : __dmul_rn(fy[t], 0.5);
          }
          ix = (ix >> 1) | (xb << (k - 1));
          iy = (iy >> 1) | (yb << (k - 1));
          run_x = xb ? run_x + 1u : 0u;
          run_y = yb ? run_y + 1u : 0u;
          since = (run_x >= 21u || run_y >= 21u) ? 0u : min(since + 1u, 0xFFFFu);
          cnt++;
          acc += (uint32_t)(int32_t)(signed char)a.qual[o + j];
        } else if (c == 'N') {
          cnt = 0;
          acc = 0;
        }
        if (cnt == k) {
          int cx[2], cy[2];
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            cx[t] = __double2int_rz(fx[t]);
            cy[t] = __double2int_rz(fy[t]);
            if (cx[t] == dim) {
              cx[t] = dim - 1;
              fx[t] = __dsub_rn(fx[t], EPS);
            }
            if (cy[t] == dim) {
              cy[t] = dim - 1;
              fy[t] = __dsub_rn(fy[t], EPS);
            }
          }
          if (patching) {
            const uint64_t cd = ((uint64_t)cx[0] << k) + (uint32_t)cy[0], ci = ((uint64_t)ix << k) + iy;
            if (cd != ci) {
              const unsigned long long qv = (unsigned long long)(uint32_t)(acc - qsub);
              atomicAdd(&C[a.o_cg_seq + ci], ~0ull);
              atomicAdd(&C[a.o_cg_seq + cd], 1ull);
              atomicAdd(&C[a.o_cg_q + ci], 0ull - qv);
              atomicAdd(&C[a.o_cg_q + cd], qv);
              patched++;
            }
          }
          cnt--;
          acc -= (uint32_t)(int32_t)(signed char)a.qual[o + j + 1u - (uint32_t)k];
        }
      }
      cnt = 0;
      acc = 0;
    }
    if (!redo) break;
  }
  atomicAdd(&C[HPGQ_S_CG_FIXED_READS], 1ull);
  if (patched) atomicAdd(&C[HPGQ_S_CG_MOVED_WORDS], (unsigned long long)patched);
}

// ---------------------------------------------------------------- 5-mer x start-position table
// kmer_t.counter_by_pos (reference src/stats_fastq.c:396-405).  One global atomic per base is what bounds the naive
// kernel, so the table is privatised in shared memory by POSITION SLICE: a CTA owns K5_W start positions for a slab
// of <= 65535 reads and keeps all 1024 x K5_W cells as 16-bit counters (a read adds at most 1 to a cell, so a slab
// cannot overflow them); a warp stages the slice bytes of 32 reads with coalesced loads, then every lane rolls the
// 2-bit code over its own read and bumps the cells with shared-memory atomics.  The table is flushed once per CTA.
constexpr int K5_W = 48;                    // start positions per slice
constexpr int K5_SW = (K5_W + 4 + 3 + 3) / 4 + 1;  // staged words per read: slice + 4 bases, any byte phase (23)
constexpr int K5_STRIDE = K5_SW | 1;        // odd row stride: conflict-free lane-per-row access
constexpr int K5_THREADS = 1024;
constexpr int K5_ROW = K5_W / 2 + 1;        // words per 5-mer row, odd: rows of different 5-mers spread over all banks

struct k5_smem {
  uint32_t tab[1024 * K5_ROW];                          // u16 pairs
  uint32_t stage[2][K5_THREADS / 32][32 * K5_STRIDE];  // double-buffered: the next group is in flight (cp.async)
};

__global__ void __launch_bounds__(K5_THREADS, 1) hq_k5pos_kernel(hq_kargs a, uint32_t reads_per_slab) {
  extern __shared__ __align__(16) uint8_t k5_raw[];
  k5_smem &sm = *reinterpret_cast<k5_smem *>(k5_raw);
  const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const uint32_t p0 = blockIdx.x * K5_W;
  const bool use_mask = (a.work & HPGQ_WORK_FILTER) != 0 && a.mask != nullptr;
  for (uint32_t i = tid; i < 1024u * K5_ROW; i += K5_THREADS) sm.tab[i] = 0u;
  __syncthreads();
  const uint32_t r_begin = blockIdx.y * reads_per_slab, r_end = min(a.n_reads, r_begin + reads_per_slab);
  // my read of group g0: offset and length (0 = nothing to count in this slice)
  auto describe = [&](uint32_t g0, uint32_t &o, uint32_t &L) {
    const uint32_t r = g0 + lane;
    o = 0u;
    L = 0u;
    if (r < r_end) {
      o = a.idx[r];
      L = a.idx[r + 1] - o;
      if (L > a.Lcap || L < p0 + 5u || (use_mask && !a.mask[r])) L = 0u;
    }
  };
  // stage: two reads of the group per instruction -- lanes 0..15 copy consecutive aligned words of the slice of read i,
  // lanes 16..31 those of read i + 1 (K5_SW <= 16 words per read), asynchronously
  static_assert(K5_SW <= 16, "a half warp stages one read");
  auto stage = [&](uint32_t *buf, uint32_t o, uint32_t L) {
    const uint32_t half = lane >> 4, l16 = lane & 15u;
#pragma unroll 8
    for (int i = 0; i < 32; i += 2) {
      const uint32_t src = (uint32_t)i + half;
      const uint32_t oi = __shfl_sync(0xFFFFFFFFu, o, src), Li = __shfl_sync(0xFFFFFFFFu, L, src);
      const uint32_t first = oi + p0, wa = (first & ~3u) + 4u * l16;
      const uint32_t last = oi + min(Li, p0 + (uint32_t)K5_W + 4u);  // one past the last byte needed
      if (Li != 0u && l16 < (uint32_t)K5_SW && wa < last)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(hq_smem_u32(buf + src * K5_STRIDE + l16)), "l"(a.seq + wa) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  constexpr uint32_t GSTEP = (K5_THREADS / 32) * 32u;
  uint32_t g0 = r_begin + warp * 32u, cur = 0u, o = 0u, L = 0u;
  if (g0 < r_end) {
    describe(g0, o, L);
    stage(sm.stage[0][warp], o, L);
  }
  for (; g0 < r_end; g0 += GSTEP) {
    uint32_t on = 0u, Ln = 0u;
    describe(g0 + GSTEP, on, Ln);             // past r_end: all lanes get L = 0, an empty copy group
    stage(sm.stage[cur ^ 1u][warp], on, Ln);
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    __syncwarp();
    const uint32_t *st = sm.stage[cur][warp];
    if (L) {
      // Straight-line over the 13 words (52 bases) of the slice: four bases per PRMT (2-bit code LUT + expected-byte
      // LUT whose mismatch marks N / foreign bytes), a 64-bit validity mask whose five-fold AND gives the windows
      // to count, and the 10-bit code of each window cut out of two consecutive 8-bit word codes.  Position, cell
      // half and increment of every update are compile-time constants.
      const uint32_t ph8 = ((o + p0) & 3u) * 8u;               // byte phase of position p0 in the first staged word
      const uint32_t nb = min(L - p0, (uint32_t)K5_W + 4u);    // bases of this read in the slice
      const uint32_t *row = st + lane * K5_STRIDE;
      constexpr int NW = (K5_W + 4) / 4;                       // 13
      uint32_t c4[NW];
      uint32_t ok_lo = 0u, ok_hi = 0u;
      uint32_t wprev = row[0];
#pragma unroll
      for (int j = 0; j < NW; ++j) {
        const uint32_t wnext = row[j + 1];
        const uint32_t w = __funnelshift_r(wprev, wnext, ph8);
        wprev = wnext;
        const uint32_t sel = hq_selector(hq_prmt(w, 0u, 0x3120u));
        const uint32_t b2w = hq_prmt_raw(0x01000000u, 0x02000003u, sel);          // A 0, C 1, G 2, T 3
        const uint32_t x = w ^ hq_prmt_raw(0x43FF41FFu, 0x47FFFF54u, sel);        // zero byte <=> one of A C G T
        const uint32_t zf = ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u;
        const uint32_t nib = ((zf >> 7) * 0x10204080u) >> 28;                      // the four flags as one nibble
        if (j < 8) ok_lo |= nib << (4 * (j & 7));
        else ok_hi |= nib << (4 * (j & 7));
        c4[j] = __dp4a(b2w, 0x01041040u, 0u);                                      // b0 b1 b2 b3 -> 8-bit code
      }
      // bases beyond the read's share of the slice are not valid
      if (nb < 32u) {
        ok_lo &= (1u << nb) - 1u;
        ok_hi = 0u;
      } else if (nb < 64u) {
        ok_hi &= (1u << (nb - 32u)) - 1u;
      }
      unsigned long long m5 = ((unsigned long long)ok_hi << 32) | ok_lo;
      m5 = m5 & (m5 >> 1) & (m5 >> 2) & (m5 >> 3) & (m5 >> 4);                     // bit t: window starting at t
      const uint32_t m_lo = (uint32_t)m5, m_hi = (uint32_t)(m5 >> 32);
#pragma unroll
      for (int t = 0; t < K5_W; ++t) {
        const int u = t + 4, j = u >> 2, i = u & 3;                                // window ends at base u
        const uint32_t H = (c4[j - 1] << 8) | c4[j];
        const uint32_t code = (H >> (6 - 2 * i)) & 1023u;
        if (((t < 32 ? m_lo : m_hi) >> (t & 31)) & 1u)
          atomicAdd(&sm.tab[code * K5_ROW + (t >> 1)], (t & 1) ? 65536u : 1u);
      }
    }
    __syncwarp();
    o = on;
    L = Ln;
    cur ^= 1u;
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  unsigned long long *C = a.ctr;
  for (uint32_t i = tid; i < 1024u * K5_ROW; i += K5_THREADS) {
    const uint32_t v = sm.tab[i];
    if (v) {
      const uint32_t id = i / K5_ROW, wpos = 2u * (i - id * K5_ROW);
      if (v & 0xFFFFu) atomicAdd(&C[a.o_k5pos + (uint64_t)id * a.Lcap + p0 + wpos], (unsigned long long)(v & 0xFFFFu));
      if (v >> 16) atomicAdd(&C[a.o_k5pos + (uint64_t)id * a.Lcap + p0 + wpos + 1u], (unsigned long long)(v >> 16));
    }
  }
}

// ---------------------------------------------------------------- max read length of a batch
// out[0] = longest read, out[1] = shortest read, out[2] = 1 if some offset is smaller than its predecessor
__global__ void hq_maxlen_kernel(const uint32_t *idx, uint32_t n, uint32_t *out) {
  uint32_t m = 0, lo = 0xFFFFFFFFu, dec = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t a = idx[i], b = idx[i + 1];
    if (b < a) dec = 1u;
    m = max(m, b - a);
    lo = min(lo, b - a);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
    lo = min(lo, __shfl_xor_sync(0xFFFFFFFFu, lo, o));
    dec |= __shfl_xor_sync(0xFFFFFFFFu, dec, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (m) atomicMax(out, m);
    atomicMin(out + 1, lo);
    if (dec) atomicExch(out + 2, 1u);
  }
}

// ---------------------------------------------------------------- synthetic reads (thread per read)
__global__ void hq_synth_kernel(hpgq_synth_cfg_t cfg, uint64_t first_read, uint32_t n, uint8_t *seq, uint8_t *qual,
                                const uint32_t *idx) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint32_t o = idx[i], L = idx[i + 1] - o;
    hpgq_synth_read_fill(&cfg, first_read + i, L, seq + o, qual + o);
  }
}

} // namespace

cudaError_t hq_launch_qc_simple(const hq_kargs &a, cudaStream_t st, uint64_t *launches) {
  if (a.n_reads == 0) return cudaSuccess;
  const uint32_t warps_needed = a.n_reads;
  uint32_t blocks = (warps_needed + 7u) / 8u;
  if (blocks > 148u * 16u) blocks = 148u * 16u;
  hq_qc_simple_kernel<<<blocks, 256, 0, st>>>(a);
  ++*launches;
  return cudaGetLastError();
}

cudaError_t hq_launch_kmers_cg(const hq_kargs &a0, uint32_t max_len_in_batch, cudaStream_t st, uint64_t *launches) {
  if (a0.n_reads == 0) return cudaSuccess;
  const bool cg = (a0.work & HPGQ_WORK_CHAOS) != 0 && a0.cg_k > 0;
  // with the packed chaos-game table a sub-launch must stay below 2^27 words (see cg_add)
  uint32_t chunk = a0.n_reads;
  if (cg && a0.cg_pack) {
    const uint64_t per = (HQ_PK_WORDS - 1ull) / (max_len_in_batch ? max_len_in_batch : 1u);
    if (per == 0) return cudaErrorInvalidValue;
    if (per < chunk) chunk = (uint32_t)per;
  }
  if ((a0.work & HPGQ_WORK_KMERS) && max_len_in_batch >= 5u) {
    const uint32_t cover = (max_len_in_batch < a0.Lcap ? max_len_in_batch : a0.Lcap) - 4u;  // window start positions
    const uint32_t slices = (cover + K5_W - 1u) / K5_W;
    const uint32_t min_slabs = (a0.n_reads + 65534u) / 65535u;  // 16-bit cells: at most 65535 reads per slab
    // one CTA per SM is resident: make slices x slabs fill whole waves of 148 CTAs
    uint32_t waves = (min_slabs * slices + 147u) / 148u, slabs = (148u * waves) / slices;
    if (slabs < min_slabs) slabs = (148u * ++waves) / slices;
    if (slabs < min_slabs) slabs = min_slabs;
    if (slabs > a0.n_reads) slabs = a0.n_reads;
    if (slabs > 65535u) slabs = 65535u;
    const uint32_t reads_per_slab = (a0.n_reads + slabs - 1u) / slabs;
    cudaError_t e = cudaFuncSetAttribute(hq_k5pos_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(k5_smem));
    if (e != cudaSuccess) return e;
    hq_k5pos_kernel<<<dim3(slices, slabs), K5_THREADS, sizeof(k5_smem), st>>>(a0, reads_per_slab);
    ++*launches;
  }
  if (!(a0.work & HPGQ_WORK_CHAOS) && !((a0.work & HPGQ_WORK_KMERS) && a0.k2 > 0)) return cudaGetLastError();
  if (cg && a0.cgflag) {
    cudaError_t e = cudaMemsetAsync(a0.cgflag, 0, a0.n_reads, st);
    if (e != cudaSuccess) return e;
  }
  for (uint32_t first = 0; first < a0.n_reads; first += chunk) {
    hq_kargs a = a0;
    a.n_reads = a0.n_reads - first < chunk ? a0.n_reads - first : chunk;
    a.idx = a0.idx + first;
    if (a.mask) a.mask = a0.mask + first;
    if (a.cgflag) a.cgflag = a0.cgflag + first;
    const bool narrow = max_len_in_batch <= 1024u;
    uint32_t blocks = (a.n_reads + (narrow ? 63u : 7u)) / (narrow ? 64u : 8u);
    if (blocks > 148u * 16u) blocks = 148u * 16u;
    if (narrow) hq_kmers_cg_kernel<4><<<blocks, 256, 0, st>>>(a);
    else hq_kmers_cg_kernel<32><<<blocks, 256, 0, st>>>(a);
    ++*launches;
    if (cg && a.cg_pack) {
      const uint32_t cells = 1u << (2 * a.cg_k);
      hq_cg_unpack_kernel<<<(cells + 1023u) / 1024u < 592u ? (cells + 1023u) / 1024u : 592u, 256, 0, st>>>(a, cells);
      ++*launches;
    }
    if (cg && a.kc_pack && (a.work & HPGQ_WORK_KMERS) && a.k2 >= a.cg_k) {
      const uint32_t cells = 1u << (2 * a.k2);
      const uint32_t own = 1u << (2 * a.cg_k);
      hq_kc_unpack_kernel<<<(own + 255u) / 256u < 2368u ? (own + 255u) / 256u : 2368u, 256, 0, st>>>(a, cells);
      ++*launches;
    }
  }
  if (cg && a0.cgflag) {
    hq_cg_fix_kernel<<<(a0.n_reads + 127u) / 128u, 128, 0, st>>>(a0, a0.cgflag);
    ++*launches;
  }
  return cudaGetLastError();
}

cudaError_t hq_launch_maxlen(const uint32_t *idx, uint32_t n, uint32_t *d_out, cudaStream_t st, uint64_t *launches) {
  cudaError_t e = cudaMemsetAsync(d_out, 0, 3 * sizeof(uint32_t), st);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_out + 1, 0xFF, sizeof(uint32_t), st);
  if (e != cudaSuccess || n == 0) return e;
  uint32_t blocks = (n + 255u) / 256u;
  if (blocks > 148u * 8u) blocks = 148u * 8u;
  hq_maxlen_kernel<<<blocks, 256, 0, st>>>(idx, n, d_out);
  ++*launches;
  return cudaGetLastError();
}

cudaError_t hq_launch_synth(const hpgq_synth_cfg_t &cfg, uint64_t first_read, uint32_t n, uint8_t *d_seq,
                            uint8_t *d_qual, const uint32_t *d_idx, cudaStream_t st, uint64_t *launches) {
  if (n == 0) return cudaSuccess;
  uint32_t blocks = (n + 127u) / 128u;
  if (blocks > 148u * 32u) blocks = 148u * 32u;
  hq_synth_kernel<<<blocks, 128, 0, st>>>(cfg, first_read, n, d_seq, d_qual, d_idx);
  ++*launches;
  return cudaGetLastError();
}
