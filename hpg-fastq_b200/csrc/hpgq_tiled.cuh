// hpgq_tiled.cuh -- the B200 hot kernel (template body): fused per-read statistics + filter decision +
// per-position base/quality tables for short reads (max read length <= 640), one pass over HBM.
//
// Replaces fastq_filter + fastq_reads_stats (reference src/stats_fastq.c:202-251) and the serial
// consumer merge (src/stats_fastq.c:257-417) for one SoA batch.
//
// Shape (DESIGN.md "stats+filter kernel"):
//   * persistent grid, one CTA per SM; one producer thread streams tiles of RT consecutive reads (seq
//     bytes, qual bytes, offsets) into a 2-stage shared-memory ring with 1-D bulk async copies (TMA
//     engine, cp.async.bulk + mbarrier complete_tx): HBM is read exactly once, nothing is re-read;
//   * 15 consumer warps; a group of G lanes owns one read per iteration, lane `lig` owns the K words
//     {lig + k*G} (4 positions each) of the read frame.  K and G are template parameters, so all
//     shared-memory offsets are immediates and every accumulator lives in a register;
//   * bases are classified four at a time with PRMT as an 8-entry byte LUT giving packed 4-bit
//     one-hots (hq_decode2); per-position counts climb a 3-level ladder: 4-bit lanes (15 reads) ->
//     8-bit lanes (255 reads) -> block-private u32 tables in shared memory -> one u64 global atomic
//     per non-zero cell per CTA.  N is never counted: N[j] = reads covering j - A - C - G - T - other;
//   * per-read sums are combined inside the group with butterfly shuffles; the pass/fail decision is taken before
//     the predicated accumulation, so statistics cover passed reads only, as in the reference;
//   * histogram keys are computed lane-per-read from a per-warp record buffer (exact division by
//     float reciprocal + correction), with the length / sum-of-quality cells cached in registers.
#pragma once
#include "hpgq_dev.cuh"
#include "hpgq_internal.h"

namespace hqt {

constexpr int NSMAX = 8;                 // pipeline stages are a launch parameter (hq_kargs.n_stages <= NSMAX)
// Ring variants that were measured and lost on B200 (2^23 x 150 bp, K = 5, G = 8; profiles/sweep_r1l.log,
// sweep_r1m.log, sweep_r1n.log) and are no longer in the source: a producer-less ring in which the last warp to
// finish a tile refills the stage (3440 against 3750 GB/s), chunks drawn from a per-stage ticket counter (3280),
// the three warps that share a scheduler with the producer taking a tile's odd chunks (-5 %), read offsets
// prefetched one chunk ahead (-0.4 %), the 8-bit level of the ladder in shared-memory columns with 19 consumer
// warps (-1.6 %: the columns take 51 KB from the ring).
#ifndef HQ_NCW
#define HQ_NCW 15                        // developer knob (tools/build_variants.sh): consumer warps per CTA
#endif
constexpr int NCW = HQ_NCW;              // consumer warps (+ producer = 4 warps per SM sub-partition: 128 registers each)
constexpr int THREADS = (NCW + 1) * 32;
constexpr int KMAX = 5;

// Shared-memory carve-up (bytes) in front of the ring, sized by the frame of the geometry: POS = 4 K G positions
// (160 for 150 bp reads, at most 640), so that short-read geometries leave more room for ring stages.
struct sm_layout {
  uint32_t tab, len, qh, gc, sumq, cnt, recs, stages;
};
__host__ __device__ constexpr sm_layout sm_carve(uint32_t POS) {
  sm_layout l{};
  l.tab = 256;                               // [0, 256): full[NSMAX], empty[NSMAX] (u64)
  l.len = l.tab + 4u * 6u * POS;             // u32 [6][POS]: accq, A, C, G, T, other
  l.qh = l.len + 4u * (POS + 4u);            // u32 [POS + 4]: reads of each length
  l.gc = l.qh + 4u * 256u;                   // u32 [256]
  l.sumq = (l.gc + 4u * 104u + 7u) & ~7u;    // u32 [104]; then u64 [POS + 4]: sum of quality by length
  l.cnt = l.sumq + 8u * (POS + 4u);          // u32 [POS + 4]: reads covering position j
  l.recs = (l.cnt + 4u * (POS + 4u) + 15u) & ~15u;  // uint4 [NCW][32]
  l.stages = (l.recs + 16u * NCW * 32u + 127u) & ~127u;
  return l;
}
constexpr uint32_t SM_BAR = 0;

__host__ __device__ inline uint32_t idx_cap_bytes(uint32_t RT) { return (((RT + 1u) * 4u + 15u) & ~15u) + 16u; }

// FULL: window means and the out-of-quality count are part of the filter.  UNI: every read of the batch has
// hq_kargs.uni_len bases (verified on the device): the aligned-word loop without funnel shifts.
template <int K, int G, bool FULL, bool UNI>
__global__ void __launch_bounds__(THREADS, 1) qc_tiled_kernel(const hq_kargs a);

template <int K, int G, bool FULL, bool UNI>
cudaError_t launch_inst(const hq_kargs &a, int grid, uint32_t smem, cudaStream_t st);

}  // namespace hqt
