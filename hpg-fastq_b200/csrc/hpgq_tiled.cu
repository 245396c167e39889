// hpgq_tiled.cu -- geometry selection and dispatch of the tiled stats+filter kernel (hpgq_tiled.cuh).
// The kernel is instantiated for a fixed list of (K words per lane, G lanes per read) pairs in
// hpgq_tiled_g*.cu; the launcher picks the pair with the best lane utilisation for the batch's longest read.
#include <stdlib.h>
#include "hpgq_tiled.cuh"

namespace {

using namespace hqt;

struct geom_entry {
  int K, G;
};
// capacity 4*K*G bases; keep in sync with the HQT_INSTANTIATE lists in hpgq_tiled_g*.cu
const geom_entry kGeoms[] = {{2, 4}, {3, 4}, {4, 4}, {5, 4}, {7, 4}, {5, 5}, {3, 8}, {4, 8}, {5, 8},
                             {3, 16}, {4, 16}, {5, 16}, {3, 32}, {4, 32}, {5, 32}};

struct hq_geom {
  int K, G;
  uint32_t RT, NS, stage_bytes, smem_bytes;
};

// tuning knobs (defaults measured on B200): HPGQ_TILE_BYTES = bases per tile that stop
// the growth of RT, HPGQ_STAGES = ring depth
uint32_t env_u32(const char *name, uint32_t dflt, uint32_t lo, uint32_t hi) {
  const char *v = getenv(name);
  if (!v || !*v) return dflt;
  const long x = strtol(v, nullptr, 10);
  return x < (long)lo ? lo : (x > (long)hi ? hi : (uint32_t)x);
}

// uniform = true: geometry for the aligned-word kernel.  A read may start at byte phase 0..3 of its first word, so
// the frame is 3 bytes longer; the group count must keep the phase constant ((32 / G) * len % 4 == 0) and only a
// lane's first and last word may be partial (len >= 4 (K - 1) G).
bool pick_geom(uint32_t max_len, bool uniform, uint32_t n_reads, int sm_count, size_t smem_limit, hq_geom *out) {
  // two stages of the largest tiles that fit (300 reads of 150 bp) beat three stages of 180 reads by 2.3 %: the
  // per-tile entry / exit costs a warp about a fifth of an iteration (profiles/tiled_r1i_summary.md)
  const uint32_t NS = env_u32("HPGQ_STAGES", 2, 2, NSMAX);
  const uint32_t tile_bytes = env_u32("HPGQ_TILE_BYTES", 46 * 1000, 1024, 64 * 1024);
  if (max_len == 0) max_len = 1;
  const uint32_t words = (max_len + (uniform ? 6u : 3u)) / 4u;
  double best = -1.0;
  int bK = 0, bG = 0;
  for (const geom_entry &ge : kGeoms) {
    if ((uint32_t)(ge.K * ge.G) < words) continue;
    if (ge.K > 5 && !uniform) continue;  // 7 words per lane: only the aligned-word loop keeps them in registers
    if (4 * ge.K * ge.G <= 256 && max_len > 255u) continue;  // packed group sum holds 8-bit base counts
    const uint32_t NG = 32u / (uint32_t)ge.G;
    if (uniform && (((NG * max_len) & 3u) != 0u || max_len < 4u * (uint32_t)((ge.K - 1) * ge.G))) continue;
    // useful word slots per warp iteration / 32*K slots issued, ties go to the larger K (less per-read overhead)
    const double eff = ((double)words * NG) / (32.0 * ge.K) + 0.002 * ge.K;
    if (eff > best) {
      best = eff;
      bK = ge.K;
      bG = ge.G;
    }
  }
  if (!bK) return false;
  const uint32_t NGB = NCW * (32u / (uint32_t)bG);
  // bytes a lane may touch past the last read of a tile: one frame; the uniform loop also addresses the reads of
  // lane groups that have none left in the tile (32 / G + 1 reads further)
  const uint32_t pad = 4u * bK * bG + 64u + (uniform ? (32u / (uint32_t)bG + 1u) * max_len : 0u);
  // reads per tile: a multiple of lcm(NGB, 4) that fits the per-stage budget
  uint32_t unit = NGB;
  while (unit % 4u) unit += NGB;
  const uint32_t sm_front = sm_carve(4u * bK * bG).stages;  // tables of this geometry, then the ring
  const size_t budget = (smem_limit - sm_front) / NS;
  uint32_t RT = 0;
  for (uint32_t cand = unit; cand <= 65535u - unit; cand += unit) {
    const size_t need = idx_cap_bytes(cand) + 2ull * ((((size_t)cand * max_len + 48u + pad) + 127u) & ~127ull);
    if (need > budget) break;
    RT = cand;
    if ((size_t)cand * max_len >= tile_bytes) break;  // big enough to amortise the barriers
  }
  if (!RT) return false;
  // developer knob: reads per tile forced (a multiple of 4 that fits the stage budget), e.g. 252 with HQ_PRANK
  const uint32_t force = env_u32("HPGQ_RT", 0, 0, 65535u);
  if (force && force % 4u == 0u &&
      idx_cap_bytes(force) + 2ull * ((((size_t)force * max_len + 48u + pad) + 127u) & ~127ull) <= budget)
    RT = force;
  // small batches: keep at least ~2 tiles per SM so the grid is balanced
  const uint32_t want_tiles = (uint32_t)sm_count * 2u;
  while (RT > unit && (n_reads + RT - 1u) / RT < want_tiles) RT -= unit;
  // medium batches (a CTA sees fewer than 64 tiles): the CTA with the most tiles sets the time, so take the tile size
  // (down to half the largest) whose last grid wave is fullest; a tile costs about 16 reads' worth of entry / exit.
  // 1 M x 100 bp: 480 reads -> 14.07 waves = 15 x 480 reads on the longest CTA, 360 reads -> 19 x 360 (-5 %)
  if (!force && (n_reads + RT - 1u) / RT < (uint32_t)sm_count * 64u) {
    uint64_t best_cost = ~0ull;
    uint32_t best_rt = RT;
    for (uint32_t cand = RT; cand >= unit && 2u * cand >= RT; cand -= unit) {
      const uint32_t tiles = (n_reads + cand - 1u) / cand;
      if (tiles < want_tiles && cand != RT) continue;
      const uint64_t waves = (tiles + (uint32_t)sm_count - 1u) / (uint32_t)sm_count;
      const uint64_t cost = waves * (cand + 16u);
      if (cost < best_cost) {
        best_cost = cost;
        best_rt = cand;
      }
    }
    RT = best_rt;
  }
  out->K = bK;
  out->G = bG;
  out->RT = RT;
  out->NS = NS;
  out->stage_bytes = (uint32_t)((((size_t)RT * max_len + 48u + pad) + 127u) & ~127ull);
  out->smem_bytes = sm_front + NS * (idx_cap_bytes(RT) + 2u * out->stage_bytes);
  return true;
}

template <int K, int G>
cudaError_t launch_kg(const hq_kargs &a, bool full, bool uni, int grid, uint32_t smem, cudaStream_t st) {
  if (uni) return full ? launch_inst<K, G, true, true>(a, grid, smem, st) : launch_inst<K, G, false, true>(a, grid, smem, st);
  return full ? launch_inst<K, G, true, false>(a, grid, smem, st) : launch_inst<K, G, false, false>(a, grid, smem, st);
}

}  // namespace

bool hq_tiled_supports(uint32_t max_len_in_batch, uint32_t work) {
  if (work & HPGQ_WORK_EDIT) return false;
  return max_len_in_batch <= (uint32_t)(4 * KMAX * 32);
}

cudaError_t hq_launch_qc_tiled(hq_kargs a, uint32_t max_len_in_batch, bool uniform, int sm_count, cudaStream_t st,
                               uint64_t *launches) {
  if (a.n_reads == 0) return cudaSuccess;
  hq_geom gm;
  if (!hq_tiled_supports(max_len_in_batch, a.work)) return cudaErrorInvalidValue;
  // HPGQ_NO_UNI: developer / test knob that keeps uniform batches on the general kernel
  bool uni = uniform && max_len_in_batch > 0 && !getenv("HPGQ_NO_UNI");
  if (uni && !pick_geom(max_len_in_batch, true, a.n_reads, sm_count, 227 * 1024, &gm)) uni = false;
  if (!uni && !pick_geom(max_len_in_batch, false, a.n_reads, sm_count, 227 * 1024, &gm)) return cudaErrorInvalidValue;
  a.RT = gm.RT;
  if (getenv("HPGQ_DEBUG_LOADONLY")) a.work |= 0x8000u;  // developer probe: ring traffic without the arithmetic
  a.n_stages = gm.NS;
  a.uni_len = uni ? max_len_in_batch : 0u;
  a.G = (uint32_t)gm.G;
  a.stage_bytes = gm.stage_bytes;
  a.n_tiles = (a.n_reads + gm.RT - 1u) / gm.RT;
  // every CTA keeps u32 block-private tables: bound the reads a CTA can see (255 * reads < 2^32)
  // HPGQ_GRID_CAP: test knob -- few CTAs, so that a small batch drives every warp through the in-loop flushes
  int grid = (int)env_u32("HPGQ_GRID_CAP", (uint32_t)sm_count, 1, (uint32_t)sm_count);
  if ((uint32_t)grid > a.n_tiles) grid = (int)a.n_tiles;
  if ((uint64_t)(a.n_tiles / grid + 1) * gm.RT >= 16000000ull) return cudaErrorInvalidValue;
  const bool full = (a.f_ll > 0) || (a.f_rl > 0) || (a.max_ooq < (int32_t)max_len_in_batch);
  cudaError_t e = cudaErrorInvalidValue;
#define HQ_CASE(KK, GG) \
  if (gm.K == KK && gm.G == GG) e = launch_kg<KK, GG>(a, full, uni, grid, gm.smem_bytes, st);
  HQ_CASE(2, 4) HQ_CASE(3, 4) HQ_CASE(4, 4) HQ_CASE(5, 4) HQ_CASE(7, 4) HQ_CASE(5, 5)
  HQ_CASE(3, 8) HQ_CASE(4, 8) HQ_CASE(5, 8)
  HQ_CASE(3, 16) HQ_CASE(4, 16) HQ_CASE(5, 16)
  HQ_CASE(3, 32) HQ_CASE(4, 32) HQ_CASE(5, 32)
#undef HQ_CASE
  if (e == cudaSuccess) ++*launches;
  return e;
}
