// hpgq_dev.cuh -- word-level SIMD-in-register primitives and sm_100a PTX wrappers shared by the
// QC kernels.  The primitives are __host__ __device__ so tests/ can exercise the exact same bit
// tricks on the CPU (tests/test_primitives.py drives csrc/hpgq_prim_host.cpp) before GPU time is spent.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define HQ_HDI __host__ __device__ __forceinline__
#define HQ_DI __device__ __forceinline__
#else
#define HQ_HDI static inline
#endif

// ------------------------------------------------------------------ generic helpers
HQ_HDI uint32_t hq_prmt(uint32_t a, uint32_t b, uint32_t s) {
#ifdef __CUDA_ARCH__
  return __byte_perm(a, b, s);
#else
  const uint64_t v = ((uint64_t)b << 32) | a;
  uint32_t r = 0;
  for (int i = 0; i < 4; ++i) {
    const uint32_t sel = (s >> (4 * i)) & 0xFu;
    uint32_t byte = (uint32_t)(v >> (8 * (sel & 7u))) & 0xFFu;
    if (sel & 8u) byte = (byte & 0x80u) ? 0xFFu : 0u;
    r |= byte << (8 * i);
  }
  return r;
#endif
}

// Raw PTX prmt (default mode): no (s & 0x7777) as __byte_perm's contract forces the compiler to emit, and a
// selector nibble with its msb set replicates the sign bit of the selected byte (bits above 15 are ignored)
HQ_HDI uint32_t hq_prmt_raw(uint32_t a, uint32_t b, uint32_t s) {
#ifdef __CUDA_ARCH__
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(s));
  return r;
#else
  return hq_prmt(a, b, s);
#endif
}

HQ_HDI uint32_t hq_dp4a(uint32_t a, uint32_t b, uint32_t c) {
#ifdef __CUDA_ARCH__
  return __dp4a(a, b, c);
#else
  for (int i = 0; i < 4; ++i) c += ((a >> (8 * i)) & 0xFFu) * ((b >> (8 * i)) & 0xFFu);
  return c;
#endif
}

HQ_HDI uint32_t hq_popc(uint32_t x) {
#ifdef __CUDA_ARCH__
  return (uint32_t)__popc(x);
#else
  return (uint32_t)__builtin_popcount(x);
#endif
}

// byte mask of the positions p, p+1, p+2, p+3 that are < L
HQ_HDI uint32_t hq_tail_mask(uint32_t p, uint32_t L) {
  const int32_t v = (int32_t)L - (int32_t)p;
  return v >= 4 ? 0xFFFFFFFFu : (v <= 0 ? 0u : ((1u << (8 * v)) - 1u));
}
#ifdef __CUDACC__
// The same mask in five instructions without a branch or select: t = bytes of the word at or past L, clamped to
// [0, 4]; all-ones shifted right by 8 t with the shift count clamped at 32 (SHF.R.CLAMP).  For the general tiled loop,
// which recomputes K of these whenever the read length changes -- every iteration on ragged reads, 15 % of its
// instructions with the form above (profiles/tiled_gen_r2w_summary.md).  Not used by the aligned-word loop: any change
// to that kernel's set-up code reshuffles ptxas' schedule of its hot loop (509 -> 523 us per 2^23 reads, same box).
__device__ __forceinline__ uint32_t hq_tail_mask_shf(uint32_t p, uint32_t L) {
  const int32_t t = min(max((int32_t)(p + 4u - L), 0), 4);
  return __funnelshift_rc(0xFFFFFFFFu, 0u, 8u * (uint32_t)t);
}
#endif
// 0x01 per byte for positions in [lo, hi)
HQ_HDI uint32_t hq_range_ones(uint32_t p, uint32_t lo, uint32_t hi) {
  return (hq_tail_mask(p, hi) & ~hq_tail_mask(p, lo)) & 0x01010101u;
}

// ------------------------------------------------------------------ base decoding
// One 32-bit word = 4 bases.  index = byte & 7 : A(0x41)->1  C(0x43)->3  T(0x54)->4  N(0x4E)->6
// G(0x47)->7 ; 0,2,5 are never a base.  PRMT is used as four parallel 8-entry byte LUTs whose
// selector nibbles are the four indices.  The selector t + (t >> 12) packs them in the order
// (byte0, byte2, byte1, byte3): output byte i of every LUT word belongs to input byte HQ_PERM(i).
#define HQ_PERM(i) ((i) == 1 ? 2 : ((i) == 2 ? 1 : (i)))

struct hq_dec_t {
  uint32_t a, c, g, t, n; // 0x01 in (permuted) byte lanes holding that base
  uint32_t bad;           // non-zero iff some byte is not one of A,C,G,T,N,0x00
};

HQ_HDI uint32_t hq_selector(uint32_t sw) {
  const uint32_t t = sw & 0x07070707u;
#ifdef __CUDA_ARCH__
  return __umulhi(t, 1u << 20) + t; // IMAD.HI: keeps the shift off the ALU pipe
#else
  return t + (t >> 12);
#endif
}

// sw must already have the bytes beyond the read end zeroed.
HQ_HDI hq_dec_t hq_decode(uint32_t sw) {
  hq_dec_t d;
  const uint32_t s = hq_selector(sw);
  d.a = hq_prmt(0x00000100u, 0u, s);
  d.c = hq_prmt(0x01000000u, 0u, s);
  d.t = hq_prmt(0u, 0x00000001u, s);
  d.n = hq_prmt(0u, 0x00010000u, s);
  d.g = hq_prmt(0u, 0x01000000u, s);
  // expected full byte for each index (0xFF = impossible), compared with the word in LUT order
  const uint32_t e = hq_prmt(0x43FF4100u, 0x474EFF54u, s);
  const uint32_t wp = hq_prmt(sw, 0u, 0x3120u);
  d.bad = wp ^ e;
  return d;
}

// exact per-byte decode (slow path for words holding lower-case / IUPAC / arbitrary bytes),
// same output convention as hq_decode
HQ_HDI hq_dec_t hq_decode_exact(uint32_t sw) {
  hq_dec_t d;
  d.a = d.c = d.g = d.t = d.n = d.bad = 0u;
  for (int i = 0; i < 4; ++i) {
    const uint32_t c = (sw >> (8 * HQ_PERM(i))) & 0xFFu;
    const uint32_t one = 1u << (8 * i);
    if (c == 'A') d.a |= one;
    else if (c == 'C') d.c |= one;
    else if (c == 'G') d.g |= one;
    else if (c == 'T') d.t |= one;
    else if (c == 'N') d.n |= one;
  }
  return d;
}

// ------------------------------------------------------------------ nibble-LUT decoding (tiled kernel v2)
// Natural byte order (output byte i <-> input byte i): the word is byte-swapped (1 <-> 2) BEFORE the
// selector is formed, so t + (t >> 12) yields the nibbles in input order.  Two LUTs give packed
// 4-bit one-hots: cg = C->0x01, G->0x10 ; at = A->0x01, T->0x10.  N (and only N) is a valid byte
// with no count: per-read N = L - #ACGT, so bytes past the read end are FILLED with 'N' (hq_fill_n)
// instead of zeroed, and 0x00 inside a read is flagged bad like any other foreign byte.
struct hq_dec2_t {
  uint32_t cg, at;
  uint32_t bad; // non-zero iff some byte is not one of A,C,G,T,N
};

// (sw & m) | (~m & "NNNN") : one LOP3
HQ_HDI uint32_t hq_fill_n(uint32_t sw, uint32_t m) { return (sw & m) | (~m & 0x4E4E4E4Eu); }

HQ_HDI hq_dec2_t hq_decode2(uint32_t swm) {
  hq_dec2_t d;
  const uint32_t wp = hq_prmt(swm, 0u, 0x3120u);
  const uint32_t s = hq_selector(wp);
  d.cg = hq_prmt_raw(0x01000000u, 0x10000000u, s);
  d.at = hq_prmt_raw(0x00000100u, 0x00000010u, s);
  const uint32_t e = hq_prmt_raw(0x43FF41FFu, 0x474EFF54u, s);
  d.bad = swm ^ e;
  return d;
}

// hq_decode2 with the low halves of the three LUTs held in registers by the caller (values pinned with an opaque
// asm so that the compiler keeps them live instead of re-materialising a MOV per use)
HQ_HDI hq_dec2_t hq_decode2_r(uint32_t swm, uint32_t lut_cg, uint32_t lut_at, uint32_t lut_e) {
  hq_dec2_t d;
  const uint32_t wp = hq_prmt(swm, 0u, 0x3120u);
  const uint32_t s = hq_selector(wp);
  d.cg = hq_prmt_raw(lut_cg, 0x10000000u, s);
  d.at = hq_prmt_raw(lut_at, 0x00000010u, s);
  const uint32_t e = hq_prmt_raw(lut_e, 0x474EFF54u, s);
  d.bad = swm ^ e;
  return d;
}

// same as hq_decode2 with the selector's shift-add issued as one IMAD.HI: p20 must be 2^20, passed through a
// kernel argument so that the compiler cannot turn it back into LEA.HI (ALU pipe)
HQ_HDI hq_dec2_t hq_decode2_fma(uint32_t swm, uint32_t p20) {
  hq_dec2_t d;
  const uint32_t wp = hq_prmt_raw(swm, 0u, 0x3120u);
  const uint32_t t = wp & 0x07070707u;
#ifdef __CUDA_ARCH__
  const uint32_t s = __umulhi(t, p20) + t;
#else
  const uint32_t s = (uint32_t)(((uint64_t)t * p20) >> 32) + t;
#endif
  d.cg = hq_prmt_raw(0x01000000u, 0x10000000u, s);
  d.at = hq_prmt_raw(0x00000100u, 0x00000010u, s);
  const uint32_t e = hq_prmt_raw(0x43FF41FFu, 0x474EFF54u, s);
  d.bad = swm ^ e;
  return d;
}

// exact per-byte twin for words holding foreign bytes; only bytes selected by m are looked at.
// other = 0x01 per in-read byte that is none of A,C,G,T,N
// 0x01 in every byte of x that equals the byte of pat (pat = one byte value replicated)
HQ_HDI uint32_t hq_eq_bytes(uint32_t x, uint32_t pat) {
  const uint32_t d = x ^ pat;
  return (~(((d & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | d) & 0x80808080u) >> 7;
}

// Exact classification of the bytes selected by m (0xFF per byte inside the read): C/G and A/T flags as in
// hq_decode2, *other = 0x01 per byte that is none of A, C, G, T, N.  Branch-free: every lane of the warp runs
// the same instructions whatever its bytes are.
HQ_HDI hq_dec2_t hq_decode2_exact(uint32_t sw, uint32_t m, uint32_t *other) {
  hq_dec2_t d;
  const uint32_t in = m & 0x01010101u;
  const uint32_t zc = hq_eq_bytes(sw, 0x43434343u), zg = hq_eq_bytes(sw, 0x47474747u);
  const uint32_t za = hq_eq_bytes(sw, 0x41414141u), zt = hq_eq_bytes(sw, 0x54545454u);
  const uint32_t zn = hq_eq_bytes(sw, 0x4E4E4E4Eu);
  d.cg = (zc | (zg << 4)) & (in * 0x11u);
  d.at = (za | (zt << 4)) & (in * 0x11u);
  d.bad = 0u;
  *other = in & ~(zc | zg | za | zt | zn);
  return d;
}

// sum over the four bytes of the low+high nibble of each byte (fields <= 7 each)
HQ_HDI uint32_t hq_nibble_total(uint32_t x) { return hq_dp4a((x + (x >> 4)) & 0x0F0F0F0Fu, 0x01010101u, 0u); }

// exact x / d for x < 2^22, d >= 1, from a float reciprocal of d (any rounding); ~8 instructions
HQ_HDI uint32_t hq_udiv_rcp(uint32_t x, uint32_t d, float rcp) {
  uint32_t q = (uint32_t)((float)x * rcp);
  const int32_t r = (int32_t)(x - q * d);
  if (r < 0) --q;
  else if (r >= (int32_t)d) ++q;
  return q;
}

// quality word -> two words of 16-bit lanes: lo = (byte0, byte2), hi = (byte1, byte3)
HQ_HDI uint32_t hq_q16_lo(uint32_t qw) { return hq_prmt(qw, 0u, 0x4240u); }
HQ_HDI uint32_t hq_q16_hi(uint32_t qw) { return hq_prmt(qw, 0u, 0x4341u); }

// number of bytes of qw (all < 128), among those selected by mask m, that are outside [lo, hi].
// c_lo = (128 - lo) * 0x01010101 with lo in [0,128]; c_hi = (127 - hi) * 0x01010101 with hi in [-1,127].
HQ_HDI uint32_t hq_ooq_count(uint32_t qw, uint32_t m, uint32_t c_lo, uint32_t c_hi) {
  const uint32_t ge = qw + c_lo; // bit7 of a byte set <=> q >= lo
  const uint32_t gt = qw + c_hi; // bit7 set <=> q >= hi + 1
  const uint32_t in = ge & ~gt;
  return hq_popc(m & 0x80808080u & ~in);
}
HQ_HDI uint32_t hq_ooq_const_lo(int32_t lo_ascii) {
  int32_t lo = lo_ascii < 0 ? 0 : (lo_ascii > 128 ? 128 : lo_ascii);
  return (uint32_t)(128 - lo) * 0x01010101u;
}
HQ_HDI uint32_t hq_ooq_const_hi(int32_t hi_ascii) {
  int32_t hi = hi_ascii < -1 ? -1 : (hi_ascii > 127 ? 127 : hi_ascii);
  return (uint32_t)(127 - hi) * 0x01010101u;
}

#ifdef __CUDACC__
// ------------------------------------------------------------------ sm_100a async-copy plumbing
HQ_DI uint32_t hq_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

HQ_DI void hq_mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(hq_smem_u32(bar)), "r"(count) : "memory");
}
HQ_DI void hq_mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
HQ_DI void hq_mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(hq_smem_u32(bar)), "r"(bytes) : "memory");
}
HQ_DI void hq_mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(hq_smem_u32(bar)) : "memory");
}
HQ_DI void hq_mbar_wait(uint64_t *bar, uint32_t parity) {
  const uint32_t a = hq_smem_u32(bar);
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(a), "r"(parity)
        : "memory");
  } while (!ok);
}
// same, for a lone thread that is expected to wait long (the producer): sleeps between polls so that it
// does not steal issue slots from the warps that share its scheduler
HQ_DI void hq_mbar_wait_backoff(uint64_t *bar, uint32_t parity) {
  const uint32_t a = hq_smem_u32(bar);
  uint32_t ok;
  for (;;) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(a), "r"(parity)
        : "memory");
    if (ok) break;
    __nanosleep(100);
  }
}
// try_wait with a suspend-time hint (ns): the thread is parked by the hardware until the phase completes or the
// hint runs out, instead of polling
HQ_DI void hq_mbar_wait_hint(uint64_t *bar, uint32_t parity, uint32_t hint_ns) {
  const uint32_t a = hq_smem_u32(bar);
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(a), "r"(parity), "r"(hint_ns)
        : "memory");
  } while (!ok);
}
// 1-D bulk async copy global -> shared (TMA engine, SASS UBLKCP); 16-byte aligned, size % 16 == 0
HQ_DI void hq_bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   hq_smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(hq_smem_u32(bar))
               : "memory");
}
HQ_DI void hq_named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
#endif
