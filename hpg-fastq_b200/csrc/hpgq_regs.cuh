// hpgq_regs.cuh -- a short read held in the registers of a lane group: the SIMD-in-word pieces shared by the
// register-resident decision kernel (hpgq_group.cu) and the fused edit + filter + compaction kernel
// (hpgq_compact.cu).  Every lane owns 32 consecutive bytes of the read's 16-byte-aligned frame.
#pragma once
#include "hpgq_dev.cuh"
#include "hpgq_internal.h"

namespace {

template <int G>
__device__ __forceinline__ uint32_t group_sum(uint32_t v) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
  return v;
}

constexpr int RG_CPL = 2;            // 16-byte chunks per lane
constexpr int RG_W = 4 * RG_CPL;     // words per lane
constexpr int RG_LB = 16 * RG_CPL;   // bytes per lane

// 0xFF in the bytes of word j (byte positions 4j .. 4j+3 of the lane) that lie in [nlo, nhi), 0 <= nlo, nhi <= 32.
// tlo / thi are nlo / nhi broadcast to the four bytes.
__device__ __forceinline__ uint32_t rg_mask(int j, uint32_t tlo, uint32_t thi) {
  const uint32_t P = (0x03020100u + 0x04040404u * (uint32_t)j) | 0x80808080u;
  const uint32_t f = (P - tlo) & ~(P - thi) & 0x80808080u;  // bit 7: position >= nlo and not >= nhi
  return hq_prmt_raw(f, 0u, 0xBA98u);                        // PTX prmt, selector msb: replicate each byte's sign bit
}

// exact 0x80 flag per zero byte of x
__device__ __forceinline__ uint32_t rg_zero_flags(uint32_t x) {
  return ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u;
}

struct rg_read {
  uint32_t q[RG_W], s[RG_W];
  uint32_t c;  // batch offset of the lane's first byte
};

// zero every byte of the lane outside [lo, hi)
__device__ __forceinline__ void rg_clip(rg_read &d, uint32_t lo, uint32_t hi) {
  if (d.c >= lo && hi >= d.c + RG_LB) return;  // lane entirely inside
  const uint32_t nlo = lo > d.c ? min(lo - d.c, (uint32_t)RG_LB) : 0u;
  const uint32_t nhi = hi > d.c ? min(hi - d.c, (uint32_t)RG_LB) : 0u;
  const uint32_t tlo = nlo * 0x01010101u, thi = nhi * 0x01010101u;
#pragma unroll
  for (int j = 0; j < RG_W; ++j) {
    const uint32_t m = rg_mask(j, tlo, thi);
    d.q[j] &= m;
    d.s[j] &= m;
  }
}

// sum of the quality bytes in [lo, hi), this lane's share
__device__ __forceinline__ uint32_t rg_window_lane(const rg_read &d, uint32_t lo, uint32_t hi) {
  if (hi <= lo || d.c >= hi || d.c + RG_LB <= lo) return 0u;
  const uint32_t nlo = lo > d.c ? lo - d.c : 0u;
  const uint32_t nhi = min(hi - d.c, (uint32_t)RG_LB);
  const uint32_t tlo = nlo * 0x01010101u, thi = nhi * 0x01010101u;
  uint32_t S = 0;
#pragma unroll
  for (int j = 0; j < RG_W; ++j) S = __dp4a(d.q[j] & rg_mask(j, tlo, thi), 0x01010101u, S);
  return S;
}

__device__ __forceinline__ bool mean_ok32(uint32_t S, uint32_t w, int32_t lo, int32_t hi) {
  return (int32_t)S >= lo * (int32_t)w && (int32_t)S <= hi * (int32_t)w;
}

}  // namespace
