// How long do the allocations of a context take on this box?  nvcc -O2 -o probe_alloc tools/probe_alloc.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <time.h>
static double now() {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return t.tv_sec + 1e-9 * t.tv_nsec;
}
int main() {
    double t = now();
    int n = 0;
    cudaGetDeviceCount(&n);
    printf("cudaGetDeviceCount        %8.2f ms\n", (now() - t) * 1e3);
    t = now();
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    printf("cudaGetDeviceProperties   %8.2f ms\n", (now() - t) * 1e3);
    t = now();
    cudaSetDevice(0);
    printf("cudaSetDevice             %8.2f ms\n", (now() - t) * 1e3);
    t = now();
    cudaFree(0);
    printf("cudaFree(0)               %8.2f ms\n", (now() - t) * 1e3);
    size_t sz[] = {64, 64, 64, 1 << 20, 16 << 20, 64 << 20, 256 << 20, 64, 16 << 20, 512 << 20};
    void *hp[10];
    for (int i = 0; i < 10; i++) {
        t = now();
        cudaError_t e = cudaHostAlloc(&hp[i], sz[i], cudaHostAllocDefault);
        printf("cudaHostAlloc %10zu  %8.2f ms %s\n", sz[i], (now() - t) * 1e3, e ? cudaGetErrorString(e) : "");
    }
    void *dp[4];
    size_t dz[] = {64, 256 << 20, 1024 << 20, 64};
    for (int i = 0; i < 4; i++) {
        t = now();
        cudaMalloc(&dp[i], dz[i]);
        printf("cudaMalloc    %10zu  %8.2f ms\n", dz[i], (now() - t) * 1e3);
    }
    t = now();
    cudaMemset(dp[1], 0, dz[1]);
    cudaDeviceSynchronize();
    printf("first cudaMemset + sync   %8.2f ms\n", (now() - t) * 1e3);
    cudaStream_t s;
    t = now();
    cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
    printf("cudaStreamCreate          %8.2f ms\n", (now() - t) * 1e3);
    for (int i = 0; i < 10; i++) {
        t = now();
        cudaFreeHost(hp[i]);
        printf("cudaFreeHost  %10zu  %8.2f ms\n", sz[i], (now() - t) * 1e3);
    }
    for (int i = 0; i < 4; i++) {
        t = now();
        cudaFree(dp[i]);
        printf("cudaFree      %10zu  %8.2f ms\n", dz[i], (now() - t) * 1e3);
    }
    t = now();
    cudaDeviceReset();
    printf("cudaDeviceReset           %8.2f ms\n", (now() - t) * 1e3);
    return 0;
}
