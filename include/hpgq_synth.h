/*
 * hpgq_synth.h -- counter-based synthetic FASTQ read generator.
 *
 * Plain C, integer arithmetic only, so that the host oracle, the host packer and the
 * device generator kernel (compiled with nvcc, HPGQ_HD = __host__ __device__) emit
 * byte-identical reads for the same (cfg, read index).  Every read is a pure function
 * of (seed, read_index): any shard of any config can be regenerated anywhere.
 *
 * This is measurement infrastructure; it restates nothing from the reference (the
 * reference ships no inputs: SURVEY.md section 4).
 */
#ifndef HPGQ_SYNTH_H
#define HPGQ_SYNTH_H

#include <stdint.h>
#include "hpg_fastq_cuda.h"

#ifndef HPGQ_HD
#ifdef __CUDACC__
#define HPGQ_HD __host__ __device__ __forceinline__
#else
#define HPGQ_HD static inline
#endif
#endif

#define HPGQ_SYNTH_DEFAULT_SEED 0x48504751ULL
#define HPGQ_SYNTH_RUN_CAP 19u /* per-read cap; runs spanning two reads stay <= 38 < 53-12 */

HPGQ_HD uint64_t hpgq_mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}

HPGQ_HD uint64_t hpgq_synth_read_hash(const hpgq_synth_cfg_t *cfg, uint64_t read_index) {
  return hpgq_mix64(cfg->seed ^ (read_index * 0x9E3779B97F4A7C15ULL));
}

HPGQ_HD uint32_t hpgq_synth_read_len(const hpgq_synth_cfg_t *cfg, uint64_t read_index) {
  uint64_t h0 = hpgq_synth_read_hash(cfg, read_index);
  if (cfg->min_len == cfg->max_len) {
    if (cfg->short_pct > 0 && cfg->short_min < cfg->max_len && ((h0 >> 40) % 100u) < cfg->short_pct)
      return cfg->short_min + (uint32_t)((h0 >> 20) % (uint64_t)(cfg->max_len - cfg->short_min));
    return cfg->max_len;
  }
  return cfg->min_len + (uint32_t)((h0 >> 20) % (uint64_t)(cfg->max_len - cfg->min_len + 1u));
}

/* Writes the L = hpgq_synth_read_len() bases and qualities of one read. */
HPGQ_HD void hpgq_synth_read_fill(const hpgq_synth_cfg_t *cfg, uint64_t read_index, uint32_t L, uint8_t *seq,
                                  uint8_t *qual) {
  const uint64_t h0 = hpgq_synth_read_hash(cfg, read_index);
  const uint32_t bad = ((h0 >> 8) % 100u) < 3u;
  const uint32_t has_burst = ((h0 >> 48) % 100u) < 1u;
  const uint32_t burst_len = 5u + (uint32_t)((h0 >> 56) % 16u);
  const uint32_t burst_start = L ? (uint32_t)((h0 >> 4) % (uint64_t)L) : 0u;
  const uint64_t LL = (uint64_t)L * (uint64_t)L;
  uint32_t run_x = 0, run_y = 0;
  uint64_t r = 0, r2 = 0;
  for (uint32_t j = 0; j < L; ++j) {
    if ((j & 7u) == 0u) {
      r = hpgq_mix64(h0 + (uint64_t)((j >> 3) + 1u) * 0xD1B54A32D192ED03ULL);
      r2 = hpgq_mix64(r ^ 0xA0761D6478BD642FULL);
    }
    const uint32_t v = (uint32_t)(r >> (8u * (j & 7u))) & 0xFFu;
    const uint32_t w = (uint32_t)(r2 >> (8u * (j & 7u))) & 0xFFu;
    uint32_t code = v & 3u; /* 0 A, 1 C, 2 G, 3 T */
    uint32_t is_n = (w == 0u) || (has_burst && j >= burst_start && j < burst_start + burst_len);
    if (cfg->cap_axis_runs && !is_n) {
      /* x-upper: A,T ; y-upper: G,T (old/chaos_game.c:200-228); N does not break a run */
      uint32_t xb = (code == 0u) || (code == 3u), yb = (code == 2u) || (code == 3u);
      if ((xb && run_x >= HPGQ_SYNTH_RUN_CAP) || (yb && run_y >= HPGQ_SYNTH_RUN_CAP)) {
        code = 1u; /* C: lower on both axes */
        xb = 0u;
        yb = 0u;
      }
      run_x = xb ? run_x + 1u : 0u;
      run_y = yb ? run_y + 1u : 0u;
    }
    uint8_t b = is_n ? (uint8_t)'N' : (uint8_t)("ACGT"[code]);
    if (cfg->lowercase_ppm) {
      uint64_t t = hpgq_mix64(r2 + j);
      if ((t % 1000000u) < cfg->lowercase_ppm) b = (uint8_t)(b | 0x20u);
    }
    seq[j] = b;
    /* e ~ geometric(1/2): trailing ones of the upper six bits, capped at 6 */
    uint32_t u = (v >> 2), e = 0; /* six bits: e <= 6 */
    while (u & 1u) {
      ++e;
      u >>= 1;
    }
    int32_t q = 38 - (int32_t)((14ull * (uint64_t)j * (uint64_t)j) / (LL ? LL : 1ull)) - (int32_t)e -
                (bad ? 15 : 0) - (int32_t)cfg->qual_shift;
    if (q < 2) q = 2;
    if (q > 41) q = 41;
    qual[j] = (uint8_t)(q + 33);
  }
}

/* splitmix64 finaliser of (position, value); shared by the device kernel and hpgq_checksum_host */
#if defined(__CUDACC__)
__host__ __device__
#endif
static inline unsigned long long hpgq_checksum_mix(unsigned long long pos, unsigned long long v) {
  unsigned long long z = pos * 0x9E3779B97F4A7C15ull + (v + 1ull) * 0xD1B54A32D192ED03ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

#endif /* HPGQ_SYNTH_H */
