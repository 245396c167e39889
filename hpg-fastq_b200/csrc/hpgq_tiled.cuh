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
//   * per-read sums are combined inside the group with REDUX; the pass/fail decision is taken before
//     the predicated accumulation, so statistics cover passed reads only, as in the reference;
//   * histogram keys are computed lane-per-read from a per-warp record buffer (exact division by
//     float reciprocal + correction), with the length / sum-of-quality cells cached in registers.
#pragma once
#include "hpgq_dev.cuh"
#include "hpgq_internal.h"

namespace hqt {

constexpr int NSMAX = 8;                 // pipeline stages are a launch parameter (hq_kargs.n_stages <= NSMAX)
// Developer A/B switches of the ring (both off by default).
// HQ_NOPROD = 1: no producer warp.  All 16 warps consume (4 per SM sub-partition, so the four schedulers carry the
// same load); the warp that is LAST to finish a tile refills its ring stage, and the two boundary offsets that
// refill needs travel with the stage's previous load, so nobody spins or waits on a global load.
// HQ_NOPROD = 0: 15 consumer warps + 1 producer warp polling the `empty` barriers (the earlier design, kept for A/B).
// Measured on B200 (2^23 x 150 bp, K = 5, G = 8): producer warp 3750 GB/s, HQ_NOPROD 3440 GB/s (the warps wait
// twice as long for a full stage), so the producer warp stays the default.
#ifndef HQ_NOPROD
#define HQ_NOPROD 0
#endif
// HQ_TICKET = 1: the warps of a CTA draw the tile's chunks of NG reads from a per-stage ticket counter instead of
// owning a fixed share, so they finish a tile together and nobody runs ahead into a stage that is still loading.
// Measured slower in both designs (3750 -> 3280 GB/s with the producer warp): the draw is on every iteration's path.
#ifndef HQ_TICKET
#define HQ_TICKET 0
#endif
// HQ_PRANK = 1: the three consumer warps that share a scheduler with the producer own the first chunks of every
// round, i.e. the extra chunks of a tile whose chunk count is not a multiple of the consumer warps (HPGQ_RT=252: 63
// chunks = 4 per warp + one more for those three).  HQ_PRODHINT = ns: the producer parks on the `empty` barrier with a
// try_wait suspend-time hint instead of polling with nanosleep.
#ifndef HQ_PRANK
#define HQ_PRANK 0
#endif
#ifndef HQ_PRODHINT
#define HQ_PRODHINT 0
#endif
// HQ_IDXPF = 1: a chunk's two read offsets are loaded from the stage one chunk ahead (two registers).  Measured on
// B200: 3 874 GB/s against 3 891 without (profiles/sweep_r1m.log), and the K = 5 FULL kernels spill; off.
#ifndef HQ_IDXPF
#define HQ_IDXPF 0
#endif
// HQ_ACC8S = 1: the 8-bit level of the per-position ladder (4 K words per thread, touched once every 15 chunks) lives in
// shared memory instead of registers -- the registers it frees are what a fifth warp per scheduler (HQ_NCW = 19) needs:
// 96 registers, no spill in the FULL = 0 kernels.  Measured on B200 (profiles/sweep_r1n.log): 15 warps 3 890 GB/s (as
// without), 19 warps 3 827, 18 warps 3 684 -- the columns take 51 KB from the ring, a tile is then 228 reads = 3 chunks
// per warp and lasts about as long as its own load, so the fifth warp waits for data instead of hiding latency.  Off.
#ifndef HQ_ACC8S
#define HQ_ACC8S 0
#endif
#ifndef HQ_NCW
#define HQ_NCW (HQ_NOPROD ? 16 : 15)
#endif
constexpr int NCW = HQ_NCW;                 // consumer warps (4 warps per SM sub-partition: 128 registers each)
constexpr int THREADS = (NCW + (HQ_NOPROD ? 0 : 1)) * 32;
constexpr uint32_t NB_BYTES = HQ_NOPROD ? 32u : 0u;  // two 16-byte offset chunks at the end of a stage's idx area
constexpr int TAB = 640;                 // positions held by the block-private tables
constexpr int KMAX = 5;

// fixed shared-memory carve-up (bytes)
constexpr uint32_t SM_BAR = 0;                               // full[NSMAX], empty[NSMAX] (u64), ticket[NSMAX] (u32)
constexpr uint32_t SM_TICKET = 128;
constexpr uint32_t SM_TAB = 256;                             // u32 [6][TAB]: accq, A, C, G, T, other
constexpr uint32_t SM_LEN = SM_TAB + 4u * 6u * TAB;          // u32 [TAB + 4]
constexpr uint32_t SM_QH = SM_LEN + 4u * (TAB + 4);          // u32 [256]
constexpr uint32_t SM_GC = SM_QH + 4u * 256u;                // u32 [104]
constexpr uint32_t SM_SUMQ = SM_GC + 4u * 104u;              // u64 [TAB + 4]
constexpr uint32_t SM_CNT = SM_SUMQ + 8u * (TAB + 4);        // u32 [TAB + 4]: reads covering position j
constexpr uint32_t SM_RECS = SM_CNT + 4u * (TAB + 4);        // uint4 [NCW][32]
constexpr uint32_t SM_ACC8 = SM_RECS + 16u * NCW * 32u;     // u32 [4 * KMAX][THREADS] when HQ_ACC8S
constexpr uint32_t SM_STAGES = (SM_ACC8 + (HQ_ACC8S ? 16u * KMAX * THREADS : 0u) + 127u) & ~127u;
static_assert(SM_SUMQ % 8 == 0, "sumq must be 8-byte aligned");
static_assert(SM_RECS % 16 == 0, "records must be 16-byte aligned");

__host__ __device__ inline uint32_t idx_cap_bytes(uint32_t RT) { return (((RT + 1u) * 4u + 15u) & ~15u) + 16u + NB_BYTES; }

template <int K, int G, bool FULL>
__global__ void __launch_bounds__(THREADS, 1) qc_tiled_kernel(const hq_kargs a);

template <int K, int G, bool FULL>
cudaError_t launch_inst(const hq_kargs &a, int grid, uint32_t smem, cudaStream_t st);

}  // namespace hqt
