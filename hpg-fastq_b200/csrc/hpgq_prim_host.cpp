// hpgq_prim_host.cpp -- host build of the word-level primitives in hpgq_dev.cuh, so that the CPU
// test-suite can check the PRMT-LUT decoding, masks and SIMD-in-register quality tricks bit for bit
// (tests/test_primitives.py).  Not part of the product library.
#include <stdint.h>
#include "hpgq_dev.cuh"

extern "C" {
// out[0..4] = a,c,g,t,n counts words ; out[5] = bad
void hqp_decode(uint32_t sw, uint32_t *out) {
  hq_dec_t d = hq_decode(sw);
  out[0] = d.a; out[1] = d.c; out[2] = d.g; out[3] = d.t; out[4] = d.n; out[5] = d.bad;
}
void hqp_decode_exact(uint32_t sw, uint32_t *out) {
  hq_dec_t d = hq_decode_exact(sw);
  out[0] = d.a; out[1] = d.c; out[2] = d.g; out[3] = d.t; out[4] = d.n; out[5] = d.bad;
}
uint32_t hqp_perm(int i) { return HQ_PERM(i); }
uint32_t hqp_tail_mask(uint32_t p, uint32_t L) { return hq_tail_mask(p, L); }
uint32_t hqp_range_ones(uint32_t p, uint32_t lo, uint32_t hi) { return hq_range_ones(p, lo, hi); }
uint32_t hqp_q16_lo(uint32_t q) { return hq_q16_lo(q); }
uint32_t hqp_q16_hi(uint32_t q) { return hq_q16_hi(q); }
uint32_t hqp_dp4a(uint32_t a, uint32_t b, uint32_t c) { return hq_dp4a(a, b, c); }
uint32_t hqp_prmt(uint32_t a, uint32_t b, uint32_t s) { return hq_prmt(a, b, s); }
uint32_t hqp_ooq(uint32_t qw, uint32_t m, int32_t lo, int32_t hi) {
  return hq_ooq_count(qw, m, hq_ooq_const_lo(lo), hq_ooq_const_hi(hi));
}
}
