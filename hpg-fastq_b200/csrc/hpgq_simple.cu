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

__device__ __forceinline__ int base2(uint32_t c) {
  return c == 'A' ? 0 : c == 'C' ? 1 : c == 'G' ? 2 : c == 'T' ? 3 : -1;
}

__global__ void __launch_bounds__(256) hq_kmers_cg_kernel(hq_kargs a) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
  const bool do_k = (a.work & HPGQ_WORK_KMERS) != 0, do_cg = (a.work & HPGQ_WORK_CHAOS) != 0 && a.cg_k > 0;
  const bool use_mask = (a.work & HPGQ_WORK_FILTER) != 0 && a.mask != nullptr;
  const int k2 = a.k2, ck = a.cg_k;
  const uint32_t k2mask = k2 > 0 ? ((k2 >= 16) ? 0xFFFFFFFFu : ((1u << (2 * k2)) - 1u)) : 0u;
  unsigned long long *C = a.ctr;
  unsigned long long words = 0, longruns = 0;

  for (uint32_t r = warp; r < a.n_reads; r += nwarps) {
    const uint32_t o = a.idx[r], L = a.idx[r + 1] - o;
    if (L == 0) continue;
    const uint8_t *s = a.seq + o, *q = a.qual + o;
    const bool kmers_here = do_k && (!use_mask || a.mask[r]) && L <= a.Lcap;
    // does the read hold bytes other than A,C,G,T,N?  (changes the chaos-game window arithmetic)
    bool other = false, single_x = true, single_y = true;
    for (uint32_t j = lane; j < L; j += 32) {
      const uint32_t c = s[j];
      const int b = base2(c);
      if (b < 0 && c != 'N') other = true;
      if (b >= 0) {
        if (!(b == 0 || b == 3)) single_x = false;
        if (!(b == 2 || b == 3)) single_y = false;
      }
    }
    other = __any_sync(0xFFFFFFFFu, other);
    single_x = __all_sync(0xFFFFFFFFu, single_x);
    single_y = __all_sync(0xFFFFFFFFu, single_y);
    if (do_cg && lane == 0 && (single_x || single_y)) longruns++;

    const uint32_t chunk = (L + 31u) / 32u;
    const uint32_t cb = lane * chunk, ce = min(L, cb + chunk);
    if (cb < ce) {
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
          if (kmers_here) {
            if (valid >= 5u) {
              const uint32_t id = code & 1023u;
              atomicAdd(&C[a.o_k5 + id], 1ull);
              atomicAdd(&C[a.o_k5pos + (uint64_t)id * a.Lcap + (j - 4u)], 1ull);
            }
            if (k2 > 0 && valid >= (uint32_t)k2) atomicAdd(&C[a.o_k2 + (code & k2mask)], 1ull);
          }
          if (do_cg && !other) {
            if (cnt >= (uint32_t)ck) {
              const uint64_t cell = ((uint64_t)ix << ck) + iy;
              atomicAdd(&C[a.o_cg_seq + cell], 1ull);
              atomicAdd(&C[a.o_cg_q + cell], (unsigned long long)(uint32_t)(accq - (uint32_t)(a.cg_base_quality * ck)));
              words++;
            }
            if (run_x == 21u || run_y == 21u) longruns++;
          }
        }
      }
    }
    // reads with other bytes: literal sequential emulation of the reference's positional window by lane 0
    if (do_cg && other && lane == 0) {
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
          if (run_x == 21u || run_y == 21u) longruns++;
          cnt++;
          accq += (uint32_t)(int32_t)(signed char)q[j];
        } else if (c == 'N') {
          cnt = 0;
          accq = 0;
        }
        if (cnt == ck) {
          const uint64_t cell = ((uint64_t)ix << ck) + iy;
          atomicAdd(&C[a.o_cg_seq + cell], 1ull);
          atomicAdd(&C[a.o_cg_q + cell], (unsigned long long)(uint32_t)(accq - (uint32_t)(a.cg_base_quality * ck)));
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

// ---------------------------------------------------------------- max read length of a batch
__global__ void hq_maxlen_kernel(const uint32_t *idx, uint32_t n, uint32_t *out) {
  uint32_t m = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    m = max(m, idx[i + 1] - idx[i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
  if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
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

cudaError_t hq_launch_kmers_cg(const hq_kargs &a, cudaStream_t st, uint64_t *launches) {
  if (a.n_reads == 0) return cudaSuccess;
  uint32_t blocks = (a.n_reads + 7u) / 8u;
  if (blocks > 148u * 16u) blocks = 148u * 16u;
  hq_kmers_cg_kernel<<<blocks, 256, 0, st>>>(a);
  ++*launches;
  return cudaGetLastError();
}

cudaError_t hq_launch_maxlen(const uint32_t *idx, uint32_t n, uint32_t *d_out, cudaStream_t st, uint64_t *launches) {
  cudaError_t e = cudaMemsetAsync(d_out, 0, sizeof(uint32_t), st);
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
