// hpgq_tiled.cu -- the B200 hot kernel: fused per-read statistics + filter decision + per-position
// base/quality tables for short reads (max read length <= 640), one pass over HBM.
//
// Replaces fastq_filter + fastq_reads_stats (reference src/stats_fastq.c:202-251) and the serial
// consumer merge (src/stats_fastq.c:257-417) for one SoA batch.
//
// Shape (see DESIGN.md "stats+filter kernel"):
//   * persistent grid, one CTA per SM; a producer warp streams tiles of RT consecutive reads
//     (their seq bytes, qual bytes and offsets) into a 3-stage shared-memory ring with 1-D bulk
//     async copies (TMA engine, cp.async.bulk + mbarrier complete_tx) -- HBM is read exactly once;
//   * 16 consumer warps; a group of G lanes owns one read at a time and every lane owns K fixed
//     32-bit words (4 positions each) of the read frame, so the per-position counters live in
//     registers as packed byte / 16-bit lanes: no atomics in the loop;
//   * bases are classified four at a time with PRMT used as an 8-entry byte LUT (hpgq_dev.cuh);
//   * per-read sums are combined inside the group with REDUX (__reduce_add_sync on the group mask),
//     the pass/fail decision is taken before the (predicated) accumulation, so statistics cover
//     passed reads only, as in the reference;
//   * histogram keys (two integer divisions per read) are computed lane-per-read from a small
//     per-warp record buffer; tables are block-private in shared memory and reach the global u64
//     counter block with one atomic per non-zero cell per CTA.
#include "hpgq_dev.cuh"
#include "hpgq_internal.h"

namespace {

constexpr int HQ_NS = 3;                       // pipeline stages
constexpr int HQ_NCW = 16;                     // consumer warps
constexpr int HQ_THREADS = (HQ_NCW + 1) * 32;  // + 1 producer warp
constexpr int HQ_TAB = 640;                    // positions held by the block-private tables
constexpr int HQ_REC = 36;                     // per-warp record slots
constexpr int HQ_KMAX = 5;

// fixed shared-memory carve-up (bytes)
constexpr uint32_t SM_BAR = 0;                                    // full[NS], empty[NS]
constexpr uint32_t SM_TAB = 64;                                   // u32 [6][HQ_TAB]
constexpr uint32_t SM_LEN = SM_TAB + 4u * 6u * HQ_TAB;            // u32 [HQ_TAB + 4]
constexpr uint32_t SM_QH = SM_LEN + 4u * (HQ_TAB + 4);            // u32 [256]
constexpr uint32_t SM_GC = SM_QH + 4u * 256u;                     // u32 [104]
constexpr uint32_t SM_SUMQ = SM_GC + 4u * 104u;                   // u64 [HQ_TAB + 4]
constexpr uint32_t SM_RECS = SM_SUMQ + 8u * (HQ_TAB + 4);         // u32 [NCW][3][HQ_REC]
constexpr uint32_t SM_STAGES = (SM_RECS + 4u * HQ_NCW * 3u * HQ_REC + 127u) & ~127u;
static_assert(SM_SUMQ % 8 == 0, "sumq must be 8-byte aligned");

__host__ __device__ inline uint32_t hq_idx_cap_bytes(uint32_t RT) { return (((RT + 1u) * 4u + 15u) & ~15u) + 16u; }

template <int K, bool FULL>
__global__ void __launch_bounds__(HQ_THREADS, 1) hq_qc_tiled_kernel(const hq_kargs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t *full = reinterpret_cast<uint64_t *>(smem + SM_BAR);
  uint64_t *empty = full + HQ_NS;
  uint32_t *tab = reinterpret_cast<uint32_t *>(smem + SM_TAB);
  uint32_t *len_hist = reinterpret_cast<uint32_t *>(smem + SM_LEN);
  uint32_t *qual_hist = reinterpret_cast<uint32_t *>(smem + SM_QH);
  uint32_t *gc_hist = reinterpret_cast<uint32_t *>(smem + SM_GC);
  unsigned long long *sumq = reinterpret_cast<unsigned long long *>(smem + SM_SUMQ);
  uint32_t *recs = reinterpret_cast<uint32_t *>(smem + SM_RECS);
  const uint32_t idx_cap = hq_idx_cap_bytes(a.RT);
  const uint32_t stage_stride = idx_cap + 2u * a.stage_bytes;
  uint8_t *stages = smem + SM_STAGES;

  const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  // zero the block-private tables (everything between SM_TAB and SM_STAGES)
  for (uint32_t i = tid; i < (SM_STAGES - SM_TAB) / 4u; i += HQ_THREADS) reinterpret_cast<uint32_t *>(smem + SM_TAB)[i] = 0u;
  if (tid == 0) {
    for (int s = 0; s < HQ_NS; ++s) {
      hq_mbar_init(&full[s], 1u);
      hq_mbar_init(&empty[s], (uint32_t)HQ_NCW);
    }
    hq_mbar_init_fence();
  }
  __syncthreads();

  const uint32_t n = a.n_reads, RT = a.RT, n_tiles = a.n_tiles;

  if (warp == (uint32_t)HQ_NCW) {
    // ------------------------------------------------------------ producer (one elected thread)
    if (lane == 0) {
      uint32_t it = 0;
      for (uint32_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
        const uint32_t s = it % HQ_NS;
        if (it >= (uint32_t)HQ_NS) hq_mbar_wait(&empty[s], ((it / HQ_NS) - 1u) & 1u);
        const uint32_t r0 = t * RT, r1 = min(n, r0 + RT);
        const uint32_t i0 = a.idx[r0], i1 = a.idx[r1];
        const uint32_t b0 = i0 & ~15u;
        uint32_t bytes = ((i1 + 15u) & ~15u) - b0;
        if (bytes == 0u) bytes = 16u;
        const uint32_t idx_bytes = ((r1 - r0 + 1u) * 4u + 15u) & ~15u;
        uint8_t *st = stages + (size_t)s * stage_stride;
        hq_mbar_expect_tx(&full[s], 2u * bytes + idx_bytes);
        hq_bulk_g2s(st, a.idx + r0, idx_bytes, &full[s]);
        hq_bulk_g2s(st + idx_cap, a.seq + b0, bytes, &full[s]);
        hq_bulk_g2s(st + idx_cap + a.stage_bytes, a.qual + b0, bytes, &full[s]);
      }
    }
    return;
  }

  // -------------------------------------------------------------- consumers
  const bool do_stats = (a.work & HPGQ_WORK_STATS) != 0, do_filter = (a.work & HPGQ_WORK_FILTER) != 0;
  const uint32_t G = a.G, NG = 32u / G;
  const uint32_t g = lane / G, lig = lane - g * G;
  const bool lane_on = g < NG;
  const uint32_t gmask = lane_on ? ((G >= 32u ? 0xFFFFFFFFu : ((1u << G) - 1u)) << (g * G)) : 0u;
  const uint32_t NGB = HQ_NCW * NG;
  uint32_t *rec0 = recs + warp * 3u * HQ_REC, *rec1 = rec0 + HQ_REC, *rec2 = rec1 + HQ_REC;

  uint32_t accA[K], accC[K], accG[K], accT[K], accN[K], accQl[K], accQh[K], m[K];
  uint32_t mwl[FULL ? K : 1], mwr[FULL ? K : 1];
#pragma unroll
  for (int k = 0; k < K; ++k) accA[k] = accC[k] = accG[k] = accT[k] = accN[k] = accQl[k] = accQh[k] = m[k] = 0u;
  uint32_t curL = 0xFFFFFFFFu, wl = 0u, wr = 0u, nacc = 0u;
  uint32_t n_pass = 0, n_fail = 0, n_stat = 0, n_slow = 0, wcount = 0;
  unsigned long long acc_len = 0;

  auto flush_regs = [&]() {
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const uint32_t p = 4u * ((uint32_t)k * G + lig);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t pos = p + HQ_PERM(i);
        uint32_t v;
        v = (accA[k] >> (8 * i)) & 0xFFu; if (v) atomicAdd(&tab[1 * HQ_TAB + pos], v);
        v = (accC[k] >> (8 * i)) & 0xFFu; if (v) atomicAdd(&tab[2 * HQ_TAB + pos], v);
        v = (accG[k] >> (8 * i)) & 0xFFu; if (v) atomicAdd(&tab[3 * HQ_TAB + pos], v);
        v = (accT[k] >> (8 * i)) & 0xFFu; if (v) atomicAdd(&tab[4 * HQ_TAB + pos], v);
        v = (accN[k] >> (8 * i)) & 0xFFu; if (v) atomicAdd(&tab[5 * HQ_TAB + pos], v);
      }
      uint32_t v;
      v = accQl[k] & 0xFFFFu; if (v) atomicAdd(&tab[p + 0u], v);
      v = accQl[k] >> 16;     if (v) atomicAdd(&tab[p + 2u], v);
      v = accQh[k] & 0xFFFFu; if (v) atomicAdd(&tab[p + 1u], v);
      v = accQh[k] >> 16;     if (v) atomicAdd(&tab[p + 3u], v);
      accA[k] = accC[k] = accG[k] = accT[k] = accN[k] = accQl[k] = accQh[k] = 0u;
    }
    nacc = 0u;
  };

  auto epilogue = [&](uint32_t wc) {
    __syncwarp();
    if (lane < wc) {
      const uint32_t r = rec0[lane], y = rec1[lane], z = rec2[lane];
      const uint32_t L = z & 0xFFFFu, pass = (z >> 16) & 1u;
      const uint32_t S = y & 0x1FFFFu, GC = y >> 17;
      if (a.mask) a.mask[r] = (uint8_t)pass;
      if (do_filter) {
        n_pass += pass;
        n_fail += pass ^ 1u;
      }
      if (do_stats && pass) {
        n_stat++;
        acc_len += L;
        atomicAdd(&len_hist[L], 1u);
        atomicAdd(&sumq[L], (unsigned long long)S);
        if (L) {
          atomicAdd(&qual_hist[(2u * S + L) / (2u * L)], 1u);
          atomicAdd(&gc_hist[(100u * GC) / L], 1u);
        }
      }
    }
    __syncwarp();
  };

  uint32_t it = 0;
  for (uint32_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
    const uint32_t s = it % HQ_NS;
    hq_mbar_wait(&full[s], (it / HQ_NS) & 1u);
    const uint8_t *st = stages + (size_t)s * stage_stride;
    const uint32_t *idx_s = reinterpret_cast<const uint32_t *>(st);
    const uint32_t *seq_w = reinterpret_cast<const uint32_t *>(st + idx_cap);
    const uint32_t *qual_w = reinterpret_cast<const uint32_t *>(st + idx_cap + a.stage_bytes);
    const uint32_t r0 = t * RT, cnt = min(RT, n - r0);
    const uint32_t b0 = idx_s[0] & ~15u;

    for (uint32_t base = warp * NG; base < cnt; base += NGB) {
      const uint32_t rt = base + g;
      const uint32_t nact = min(NG, cnt - base);
      if (lane_on && rt < cnt) {
        const uint32_t i0 = idx_s[rt], L = idx_s[rt + 1] - i0, o = i0 - b0;
        const uint32_t bsel = 0x3210u + 0x1111u * (o & 3u);
        const uint32_t wb = (o >> 2) + lig;
        if (L != curL) {
          curL = L;
#pragma unroll
          for (int k = 0; k < K; ++k) m[k] = hq_tail_mask(4u * ((uint32_t)k * G + lig), L);
          if (FULL) {
            wl = a.f_ll > 0 ? min((uint32_t)a.f_ll, L) : 0u;
            wr = a.f_rl > 0 ? min((uint32_t)a.f_rl, L) : 0u;
#pragma unroll
            for (int k = 0; k < K; ++k) {
              const uint32_t p = 4u * ((uint32_t)k * G + lig);
              mwl[FULL ? k : 0] = hq_range_ones(p, 0u, wl);
              mwr[FULL ? k : 0] = hq_range_ones(p, L - wr, L);
            }
          }
        }
        hq_dec_t d[K];
        uint32_t qw[K];
        uint32_t S = 0, sumN = 0, sumGC = 0, bad = 0, oob = 0, SL = 0, SR = 0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const uint32_t wi = wb + (uint32_t)k * G;
          const uint32_t s0 = seq_w[wi], s1 = seq_w[wi + 1];
          const uint32_t q0 = qual_w[wi], q1 = qual_w[wi + 1];
          const uint32_t sw = hq_prmt(s0, s1, bsel) & m[k];
          qw[k] = hq_prmt(q0, q1, bsel) & m[k];
          d[k] = hq_decode(sw);
          bad |= d[k].bad | (qw[k] & 0x80808080u);
          S = hq_dp4a(qw[k], 0x01010101u, S);
          sumN += d[k].n;
          sumGC += d[k].c + d[k].g;
          if (FULL) {
            oob += hq_ooq_count(qw[k], m[k], a.c_lo, a.c_hi);
            SL = hq_dp4a(qw[k], mwl[FULL ? k : 0], SL);
            SR = hq_dp4a(qw[k], mwr[FULL ? k : 0], SR);
          }
        }
        uint32_t nN = hq_dp4a(sumN, 0x01010101u, 0u), nGC = hq_dp4a(sumGC, 0x01010101u, 0u);
        const uint32_t R1 = __reduce_add_sync(gmask, S | (bad ? (1u << 24) : 0u));
        S = R1 & 0xFFFFFFu;
        if (R1 >> 24) {
          // exact per-byte path: the read holds lower-case / IUPAC / non-ASCII bytes
          uint32_t oobx = 0;
          sumN = 0;
          sumGC = 0;
#pragma unroll
          for (int k = 0; k < K; ++k) {
            const uint32_t wi = wb + (uint32_t)k * G;
            const uint32_t sw = hq_prmt(seq_w[wi], seq_w[wi + 1], bsel) & m[k];
            d[k] = hq_decode_exact(sw);
            sumN += d[k].n;
            sumGC += d[k].c + d[k].g;
            if (FULL) {
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const int32_t qq = (int32_t)((qw[k] >> (8 * i)) & 0xFFu);
                if ((m[k] >> (8 * i)) & 1u) oobx += (qq < a.q_lo) | (qq > a.q_hi);
              }
            }
          }
          nN = hq_dp4a(sumN, 0x01010101u, 0u);
          nGC = hq_dp4a(sumGC, 0x01010101u, 0u);
          oob = oobx;
          if (lig == 0) n_slow++;
        }
        const uint32_t R2 = __reduce_add_sync(gmask, nN | (nGC << 10) | (oob << 20));
        nN = R2 & 1023u;
        nGC = (R2 >> 10) & 1023u;
        oob = R2 >> 20;
        bool pass = true;
        if (do_filter) {
          const int32_t Li = (int32_t)L, Si = (int32_t)S;
          pass = Li >= a.min_len && Li <= a.max_len && Si >= a.q_lo * Li && Si <= a.q_hi * Li && (int32_t)nN <= a.max_n;
          if (FULL) {
            pass = pass && (int32_t)oob <= a.max_ooq;
            if (wl) {
              const int32_t sl = (int32_t)__reduce_add_sync(gmask, SL), w = (int32_t)wl;
              pass = pass && sl >= a.f_lq_lo * w && sl <= a.f_lq_hi * w;
            }
            if (wr) {
              const int32_t sr = (int32_t)__reduce_add_sync(gmask, SR), w = (int32_t)wr;
              pass = pass && sr >= a.f_rq_lo * w && sr <= a.f_rq_hi * w;
            }
          }
        }
        if (do_stats && pass) {
#pragma unroll
          for (int k = 0; k < K; ++k) {
            accA[k] += d[k].a;
            accC[k] += d[k].c;
            accG[k] += d[k].g;
            accT[k] += d[k].t;
            accN[k] += d[k].n;
            accQl[k] += hq_q16_lo(qw[k]);
            accQh[k] += hq_q16_hi(qw[k]);
          }
          if (++nacc == 255u) flush_regs();
        }
        if (lig == 0) {
          rec0[wcount + g] = r0 + rt;
          rec1[wcount + g] = S | (nGC << 17);
          rec2[wcount + g] = L | ((pass ? 1u : 0u) << 16);
        }
      }
      wcount += nact;
      if (wcount + NG > 32u) {
        epilogue(wcount);
        wcount = 0;
      }
    }
    __syncwarp();
    if (lane == 0) hq_mbar_arrive(&empty[s]);
  }
  epilogue(wcount);
  if (lane_on) flush_regs();

  // per-lane scalar counters -> one atomic per warp
  unsigned long long *C = a.ctr;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n_pass += __shfl_xor_sync(0xFFFFFFFFu, n_pass, o);
    n_fail += __shfl_xor_sync(0xFFFFFFFFu, n_fail, o);
    n_stat += __shfl_xor_sync(0xFFFFFFFFu, n_stat, o);
    n_slow += __shfl_xor_sync(0xFFFFFFFFu, n_slow, o);
    acc_len += __shfl_xor_sync(0xFFFFFFFFu, acc_len, o);
  }
  if (lane == 0) {
    if (n_pass) atomicAdd(&C[HPGQ_S_NUM_PASSED], (unsigned long long)n_pass);
    if (n_fail) atomicAdd(&C[HPGQ_S_NUM_FAILED], (unsigned long long)n_fail);
    if (n_stat) atomicAdd(&C[HPGQ_S_NUM_READS], (unsigned long long)n_stat);
    if (n_slow) atomicAdd(&C[HPGQ_S_SLOW_READS], (unsigned long long)n_slow);
    if (acc_len) atomicAdd(&C[HPGQ_S_ACC_LENGTH], acc_len);
  }
  hq_named_bar_sync(1, HQ_NCW * 32);
  if (do_stats) {
    const uint32_t Lcap = a.Lcap;
    const uint32_t sec_off[6] = {a.o_accq, a.o_A, a.o_C, a.o_G, a.o_T, a.o_N};
    for (uint32_t i = tid; i < 6u * HQ_TAB; i += HQ_NCW * 32) {
      const uint32_t v = tab[i];
      const uint32_t sec = i / HQ_TAB, pos = i - sec * HQ_TAB;
      if (v && pos < Lcap) atomicAdd(&C[sec_off[sec] + pos], (unsigned long long)v);
    }
    for (uint32_t i = tid; i <= (uint32_t)HQ_TAB; i += HQ_NCW * 32) {
      if (i <= Lcap) {
        if (len_hist[i]) atomicAdd(&C[a.o_len_hist + i], (unsigned long long)len_hist[i]);
        if (sumq[i]) atomicAdd(&C[a.o_sumq + i], sumq[i]);
      }
    }
    for (uint32_t i = tid; i < 256u; i += HQ_NCW * 32)
      if (qual_hist[i]) atomicAdd(&C[a.o_qual_hist + i], (unsigned long long)qual_hist[i]);
    for (uint32_t i = tid; i < 104u; i += HQ_NCW * 32)
      if (gc_hist[i]) atomicAdd(&C[a.o_gc_hist + i], (unsigned long long)gc_hist[i]);
  }
}

struct hq_geom {
  int K;
  uint32_t G, RT, stage_bytes, smem_bytes;
};

// pick (K, G): lanes-per-read group and words per lane with the best lane utilisation
bool hq_pick_geom(uint32_t max_len, uint32_t n_reads, int sm_count, size_t smem_limit, hq_geom *out) {
  if (max_len == 0) max_len = 1;
  const uint32_t words = (max_len + 3u) / 4u;
  double best = -1.0;
  int bK = 0;
  uint32_t bG = 0;
  for (int K = 1; K <= HQ_KMAX; ++K) {
    const uint32_t G = (words + K - 1) / K;
    if (G > 32u) continue;
    const uint32_t NG = 32u / G;
    const double eff = ((double)words / (double)(K * G)) * ((double)(NG * G) / 32.0) + 0.002 * K; // ties: larger K
    if (eff > best) {
      best = eff;
      bK = K;
      bG = G;
    }
  }
  if (!bK) return false;
  const uint32_t NGB = HQ_NCW * (32u / bG);
  const uint32_t pad = 4u * bK * bG + 64u;
  // reads per tile: a multiple of lcm(NGB, 4) that fits the per-stage budget
  uint32_t unit = NGB;
  while (unit % 4u) unit += NGB;
  const size_t budget = (smem_limit - SM_STAGES) / HQ_NS;
  uint32_t RT = 0;
  for (uint32_t cand = unit; cand <= 65535u - unit; cand += unit) {
    const size_t need = hq_idx_cap_bytes(cand) + 2ull * ((((size_t)cand * max_len + 48u + pad) + 127u) & ~127ull);
    if (need > budget) break;
    RT = cand;
    if ((size_t)cand * max_len >= 48u * 1024u) break; // big enough to amortise the barriers
  }
  if (!RT) return false;
  // small batches: keep at least ~2 tiles per SM so the grid is balanced
  const uint32_t want_tiles = (uint32_t)sm_count * 2u;
  while (RT > unit && (n_reads + RT - 1u) / RT < want_tiles) RT -= unit;
  out->K = bK;
  out->G = bG;
  out->RT = RT;
  out->stage_bytes = (uint32_t)((((size_t)RT * max_len + 48u + pad) + 127u) & ~127ull);
  out->smem_bytes = SM_STAGES + HQ_NS * (hq_idx_cap_bytes(RT) + 2u * out->stage_bytes);
  return true;
}

template <int K, bool FULL>
cudaError_t launch_inst(const hq_kargs &a, int grid, uint32_t smem, cudaStream_t st) {
  // the attribute is per device; set it on every launch (one cheap driver call)
  cudaError_t e = cudaFuncSetAttribute(hq_qc_tiled_kernel<K, FULL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       227 * 1024);
  if (e != cudaSuccess) return e;
  hq_qc_tiled_kernel<K, FULL><<<grid, HQ_THREADS, smem, st>>>(a);
  return cudaGetLastError();
}

} // namespace

bool hq_tiled_supports(uint32_t max_len_in_batch, uint32_t work) {
  if (work & HPGQ_WORK_EDIT) return false;
  return max_len_in_batch <= (uint32_t)(4 * HQ_KMAX * 32);
}

cudaError_t hq_launch_qc_tiled(hq_kargs a, uint32_t max_len_in_batch, int sm_count, cudaStream_t st,
                               uint64_t *launches) {
  if (a.n_reads == 0) return cudaSuccess;
  hq_geom gm;
  if (!hq_tiled_supports(max_len_in_batch, a.work) ||
      !hq_pick_geom(max_len_in_batch, a.n_reads, sm_count, 227 * 1024, &gm))
    return cudaErrorInvalidValue;
  a.RT = gm.RT;
  a.G = gm.G;
  a.stage_bytes = gm.stage_bytes;
  a.n_tiles = (a.n_reads + gm.RT - 1u) / gm.RT;
  // every CTA keeps u32 block-private tables: bound the reads a CTA can see (255 * reads < 2^32)
  int grid = sm_count;
  if ((uint32_t)grid > a.n_tiles) grid = (int)a.n_tiles;
  if ((uint64_t)(a.n_tiles / grid + 1) * gm.RT >= 16000000ull) return cudaErrorInvalidValue;
  const bool full = (a.f_ll > 0) || (a.f_rl > 0) || (a.max_ooq < (int32_t)max_len_in_batch);
  cudaError_t e;
#define HQ_CASE(KK)                                                                             \
  case KK:                                                                                      \
    e = full ? launch_inst<KK, true>(a, grid, gm.smem_bytes, st) : launch_inst<KK, false>(a, grid, gm.smem_bytes, st); \
    break;
  switch (gm.K) {
    HQ_CASE(1)
    HQ_CASE(2)
    HQ_CASE(3)
    HQ_CASE(4)
    HQ_CASE(5)
    default: e = cudaErrorInvalidValue;
  }
#undef HQ_CASE
  if (e == cudaSuccess) ++*launches;
  return e;
}
