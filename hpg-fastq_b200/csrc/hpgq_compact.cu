// hpgq_compact.cu -- order-preserving partition of the (trimmed) reads of a batch into
// "passed" then "failed", each in input order: what fastq_fwrite(passed_reads) and
// fastq_fwrite(failed_reads) emit in the reference consumers (src/filter_fastq.c:161-174,
// src/edit_fastq.c:184-206) when run with one worker thread.
//
// Three steps: per-read (bytes, count) records -> exclusive scan (block scan + block-sum scan +
// add) -> warp-per-read copy of seq and qual bytes to their final offsets.
#include "hpgq_dev.cuh"
#include "hpgq_internal.h"

namespace {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 4; // 1024 records per block

__device__ __forceinline__ uint4 add4(uint4 a, uint4 b) { return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }

__device__ __forceinline__ uint4 shfl_up4(uint4 v, int d) {
  return make_uint4(__shfl_up_sync(0xFFFFFFFFu, v.x, d), __shfl_up_sync(0xFFFFFFFFu, v.y, d),
                    __shfl_up_sync(0xFFFFFFFFu, v.z, d), __shfl_up_sync(0xFFFFFFFFu, v.w, d));
}

// record of read i (i == n gives the zero sentinel so that scan[n] holds the totals)
__device__ __forceinline__ uint4 make_record(const hq_kargs &a, uint32_t i) {
  if (i >= a.n_reads) return make_uint4(0, 0, 0, 0);
  const uint32_t ln = a.trim ? a.trim[i].len : (a.idx[i + 1] - a.idx[i]);
  const bool pass = a.mask ? (a.mask[i] != 0) : true;
  return pass ? make_uint4(ln, 0, 1, 0) : make_uint4(0, ln, 0, 1);
}

// inclusive block scan of one uint4 per thread; returns the thread's inclusive value, total in *tot
__device__ __forceinline__ uint4 block_scan_incl(uint4 v, uint4 *tot) {
  __shared__ uint4 wsum[SCAN_THREADS / 32];
  const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint4 u = shfl_up4(v, d);
    if ((int)lane >= d) v = add4(v, u);
  }
  if (lane == 31) wsum[w] = v;
  __syncthreads();
  if (w == 0) {
    uint4 s = lane < SCAN_THREADS / 32 ? wsum[lane] : make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int d = 1; d < SCAN_THREADS / 32; d <<= 1) {
      const uint4 u = shfl_up4(s, d);
      if ((int)lane >= d) s = add4(s, u);
    }
    if (lane < SCAN_THREADS / 32) wsum[lane] = s;
  }
  __syncthreads();
  if (w > 0) v = add4(v, wsum[w - 1]);
  *tot = wsum[SCAN_THREADS / 32 - 1];
  __syncthreads();
  return v;
}

// pass 1: per-block totals
__global__ void __launch_bounds__(SCAN_THREADS) hq_compact_sums(hq_kargs a, uint4 *block_sums) {
  const uint32_t base = blockIdx.x * (SCAN_THREADS * SCAN_ITEMS) + threadIdx.x * SCAN_ITEMS;
  uint4 v = make_uint4(0, 0, 0, 0);
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) v = add4(v, make_record(a, base + k));
  uint4 tot;
  block_scan_incl(v, &tot);
  if (threadIdx.x == 0) block_sums[blockIdx.x] = tot;
}

// pass 2: exclusive scan of the block totals by ONE block (nb <= a few thousand)
__global__ void __launch_bounds__(SCAN_THREADS) hq_compact_scan_sums(uint4 *block_sums, uint32_t nb) {
  uint4 carry = make_uint4(0, 0, 0, 0);
  for (uint32_t b0 = 0; b0 < nb; b0 += SCAN_THREADS) {
    const uint32_t i = b0 + threadIdx.x;
    const uint4 v = i < nb ? block_sums[i] : make_uint4(0, 0, 0, 0);
    uint4 tot;
    const uint4 inc = block_scan_incl(v, &tot);
    if (i < nb) block_sums[i] = make_uint4(carry.x + inc.x - v.x, carry.y + inc.y - v.y, carry.z + inc.z - v.z,
                                           carry.w + inc.w - v.w);
    carry = add4(carry, tot);
  }
}

// pass 3: exclusive scan written to scan[0..n]
__global__ void __launch_bounds__(SCAN_THREADS) hq_compact_scan(hq_kargs a, const uint4 *block_sums, uint4 *scan,
                                                                uint32_t *totals) {
  const uint32_t base = blockIdx.x * (SCAN_THREADS * SCAN_ITEMS) + threadIdx.x * SCAN_ITEMS;
  uint4 r[SCAN_ITEMS];
  uint4 v = make_uint4(0, 0, 0, 0);
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    r[k] = make_record(a, base + k);
    v = add4(v, r[k]);
  }
  uint4 tot;
  const uint4 inc = block_scan_incl(v, &tot);
  const uint4 off = block_sums[blockIdx.x];
  uint4 run = make_uint4(off.x + inc.x - v.x, off.y + inc.y - v.y, off.z + inc.z - v.z, off.w + inc.w - v.w);
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    const uint32_t i = base + k;
    if (i <= a.n_reads) {
      scan[i] = run;
      if (i == a.n_reads) {
        totals[0] = run.z;
        totals[1] = run.w;
        totals[2] = run.x;
        totals[3] = run.y;
      }
    }
    run = add4(run, r[k]);
  }
}

// pass 4: copy. One warp per read; output slot and byte offset come from the scan.
__global__ void __launch_bounds__(256) hq_compact_copy(hq_kargs a, hq_compact_bufs b) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
  const uint32_t n = a.n_reads;
  const uint4 tot = b.scan[n];
  for (uint32_t i = warp; i < n; i += nwarps) {
    const uint4 sc = b.scan[i];
    const bool pass = a.mask ? (a.mask[i] != 0) : true;
    const uint32_t st = a.trim ? a.trim[i].start : 0u;
    const uint32_t ln = a.trim ? a.trim[i].len : (a.idx[i + 1] - a.idx[i]);
    const uint32_t dst = pass ? sc.x : tot.x + sc.y;
    const uint32_t slot = pass ? sc.z : tot.z + sc.w;
    if (lane == 0) {
      b.out_src[slot] = i;
      b.out_idx[slot] = dst;
    }
    const uint8_t *s = a.seq + a.idx[i] + st, *q = a.qual + a.idx[i] + st;
    uint8_t *ds = b.out_seq + dst, *dq = b.out_qual + dst;
    for (uint32_t j = lane; j < ln; j += 32) {
      ds[j] = s[j];
      dq[j] = q[j];
    }
  }
  if (warp == 0 && lane == 0) b.out_idx[n] = tot.x + tot.y;
}

} // namespace

cudaError_t hq_launch_compact(const hq_kargs &a, const hq_compact_bufs &b, cudaStream_t st, uint64_t *launches) {
  const uint32_t n1 = a.n_reads + 1u;
  const uint32_t nb = (n1 + SCAN_THREADS * SCAN_ITEMS - 1u) / (SCAN_THREADS * SCAN_ITEMS);
  hq_compact_sums<<<nb, SCAN_THREADS, 0, st>>>(a, b.block_sums);
  hq_compact_scan_sums<<<1, SCAN_THREADS, 0, st>>>(b.block_sums, nb);
  hq_compact_scan<<<nb, SCAN_THREADS, 0, st>>>(a, b.block_sums, b.scan, b.totals);
  *launches += 3;
  if (a.n_reads) {
    uint32_t blocks = (a.n_reads + 7u) / 8u;
    if (blocks > 148u * 16u) blocks = 148u * 16u;
    hq_compact_copy<<<blocks, 256, 0, st>>>(a, b);
    ++*launches;
  }
  return cudaGetLastError();
}
