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
constexpr int TAB = 640;                 // positions held by the block-private tables
constexpr int KMAX = 5;

// fixed shared-memory carve-up (bytes)
constexpr uint32_t SM_BAR = 0;                               // full[NSMAX], empty[NSMAX] (u64)
constexpr uint32_t SM_TAB = 256;                             // u32 [6][TAB]: accq, A, C, G, T, other
constexpr uint32_t SM_LEN = SM_TAB + 4u * 6u * TAB;          // u32 [TAB + 4]
constexpr uint32_t SM_QH = SM_LEN + 4u * (TAB + 4);          // u32 [256]
constexpr uint32_t SM_GC = SM_QH + 4u * 256u;                // u32 [104]
constexpr uint32_t SM_SUMQ = SM_GC + 4u * 104u;              // u64 [TAB + 4]
constexpr uint32_t SM_CNT = SM_SUMQ + 8u * (TAB + 4);        // u32 [TAB + 4]: reads covering position j
constexpr uint32_t SM_RECS = SM_CNT + 4u * (TAB + 4);        // uint4 [NCW][32]
constexpr uint32_t SM_STAGES = (SM_RECS + 16u * NCW * 32u + 127u) & ~127u;
static_assert(SM_SUMQ % 8 == 0, "sumq must be 8-byte aligned");
static_assert(SM_RECS % 16 == 0, "records must be 16-byte aligned");

__host__ __device__ inline uint32_t idx_cap_bytes(uint32_t RT) { return (((RT + 1u) * 4u + 15u) & ~15u) + 16u; }

// FULL: window means and the out-of-quality count are part of the filter.  UNI: every read of the batch has
// hq_kargs.uni_len bases (verified on the device): the aligned-word loop without funnel shifts.
template <int K, int G, bool FULL, bool UNI>
__global__ void __launch_bounds__(THREADS, 1) qc_tiled_kernel(const hq_kargs a);

template <int K, int G, bool FULL, bool UNI>
cudaError_t launch_inst(const hq_kargs &a, int grid, uint32_t smem, cudaStream_t st);

}  // namespace hqt
