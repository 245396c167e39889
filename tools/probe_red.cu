// L2 reduction rates on this GPU: random 32-bit and 64-bit RED (fire-and-forget atomicAdd) and ATOM (with return) into
// a table that fits L2.   nvcc -O2 -arch=sm_100a -o tools/probe_red tools/probe_red.cu ; ./tools/probe_red
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t mix(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}
template <int MODE>
__global__ void red_kernel(void *tab, uint32_t cells, uint32_t per_thread, uint32_t *sink) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t acc = 0;
  for (uint32_t i = 0; i < per_thread; ++i) {
    const uint32_t c = mix(t * 0x9E3779B9u + i) & (cells - 1u);
    if (MODE == 0) atomicAdd(reinterpret_cast<unsigned long long *>(tab) + c, (1ull << 37) | 17ull);
    if (MODE == 1) atomicAdd(reinterpret_cast<uint32_t *>(tab) + c, (1u << 24) | 17u);
    if (MODE == 2) acc += atomicAdd(reinterpret_cast<uint32_t *>(tab) + c, (1u << 24) | 17u) >> 24;
    if (MODE == 3) acc += (uint32_t)(atomicAdd(reinterpret_cast<unsigned long long *>(tab) + c, (1ull << 37) | 17ull) >> 37);
  }
  if (acc == 0xFFFFFFFFu) *sink = acc;
}
template <int MODE>
static void run(const char *name, void *tab, uint32_t cells, uint32_t *sink) {
  const uint32_t blocks = 148 * 16, threads = 256, per = 128;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  red_kernel<MODE><<<blocks, threads>>>(tab, cells, per, sink);
  cudaEventRecord(e0);
  for (int r = 0; r < 5; ++r) red_kernel<MODE><<<blocks, threads>>>(tab, cells, per, sink);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const double n = 5.0 * blocks * threads * per;
  printf("%-28s cells %9u (%4u MB): %7.1f G ops/s\n", name, cells, (unsigned)((MODE == 0 || MODE == 3 ? 8ull : 4ull) * cells >> 20), n / (ms * 1e-3) / 1e9);
}
int main() {
  void *tab;
  uint32_t *sink;
  cudaMalloc(&tab, 256u << 20);
  cudaMalloc(&sink, 4);
  cudaMemset(tab, 0, 256u << 20);
  for (uint32_t cells : {1u << 20, 1u << 22, 1u << 24}) {
    run<0>("RED 64-bit", tab, cells, sink);
    run<1>("RED 32-bit", tab, cells, sink);
    run<2>("ATOM 32-bit (with return)", tab, cells, sink);
    run<3>("ATOM 64-bit (with return)", tab, cells, sink);
  }
  return cudaDeviceSynchronize() != cudaSuccess;
}
