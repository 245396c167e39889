// Host-side probe for the text path: how fast can file bytes in the page cache reach pinned memory / the device?
//   nvcc -O2 -o probe_hostcopy tools/probe_hostcopy.cu -lpthread ; ./probe_hostcopy /dev/shm/probe.bin 2048
#include <cuda_runtime.h>
#include <fcntl.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <time.h>
#include <unistd.h>

static double now() {
    struct timespec t;
    clock_gettime(CLOCK_MONOTONIC, &t);
    return t.tv_sec + 1e-9 * t.tv_nsec;
}
#define CK(x)                                                                       \
    do {                                                                            \
        cudaError_t e_ = (x);                                                       \
        if (e_ != cudaSuccess) {                                                    \
            printf("%s -> %s\n", #x, cudaGetErrorString(e_));                       \
            exit(1);                                                                \
        }                                                                           \
    } while (0)

struct job {
    int fd;
    char *dst;
    const char *src;
    size_t off, len;
};
static void *do_pread(void *p) {
    job *j = (job *)p;
    size_t done = 0;
    while (done < j->len) {
        ssize_t r = pread(j->fd, j->dst + done, j->len - done, j->off + done);
        if (r <= 0) break;
        done += r;
    }
    return 0;
}
static void *do_memcpy(void *p) {
    job *j = (job *)p;
    memcpy(j->dst, j->src + j->off, j->len);
    return 0;
}
static void *do_pwrite(void *p) {
    job *j = (job *)p;
    size_t done = 0;
    while (done < j->len) {
        ssize_t r = pwrite(j->fd, j->src + j->off + done, j->len - done, j->off + done);
        if (r <= 0) break;
        done += r;
    }
    return 0;
}
static void *do_memcpy_out(void *p) {  // dst = mapping of the output file, src = pinned
    job *j = (job *)p;
    memcpy(j->dst, j->src + j->off, j->len);
    return 0;
}
static double run_threads(int T, void *(*fn)(void *), int fd, char *dst, const char *src, size_t n) {
    pthread_t th[64];
    job jb[64];
    size_t per = (n / T + 4095) & ~(size_t)4095;
    double t0 = now();
    for (int i = 0; i < T; i++) {
        size_t o = per * i;
        size_t l = o >= n ? 0 : (n - o < per ? n - o : per);
        jb[i] = {fd, dst + o, src, o, l};
        pthread_create(&th[i], 0, fn, &jb[i]);
    }
    for (int i = 0; i < T; i++) pthread_join(th[i], 0);
    return now() - t0;
}

int main(int argc, char **argv) {
    const char *path = argc > 1 ? argv[1] : "/dev/shm/probe.bin";
    size_t mb = argc > 2 ? atol(argv[2]) : 1024;
    size_t n = mb << 20;
    CK(cudaSetDevice(0));
    CK(cudaFree(0));
    char *pin, *dev;
    CK(cudaHostAlloc(&pin, n, cudaHostAllocDefault));
    CK(cudaMalloc(&dev, n));
    memset(pin, 'A', n);
    int fd = open(path, O_CREAT | O_RDWR | O_TRUNC, 0600);
    for (size_t o = 0; o < n; o += 64 << 20)
        if (write(fd, pin + o, 64 << 20) < 0) return 1;
    fsync(fd);
    printf("file %s, %zu MB, cores %ld\n", path, mb, sysconf(_SC_NPROCESSORS_ONLN));
    int Ts[] = {1, 2, 4, 8, 16, 24, 32, 48};
    for (int k = 0; k < 8; k++) {
        double best = 1e9;
        for (int rep = 0; rep < 3; rep++) {
            double t = run_threads(Ts[k], do_pread, fd, pin, 0, n);
            if (t < best) best = t;
        }
        printf("pread -> pinned      T=%2d : %6.2f GB/s\n", Ts[k], n / best / 1e9);
    }
    double t0 = now();
    char *map = (char *)mmap(0, n, PROT_READ, MAP_SHARED | MAP_POPULATE, fd, 0);
    printf("mmap MAP_POPULATE           : %6.3f s (%6.2f GB/s)\n", now() - t0, n / (now() - t0) / 1e9);
    for (int k = 0; k < 8; k++) {
        double best = 1e9;
        for (int rep = 0; rep < 3; rep++) {
            double t = run_threads(Ts[k], do_memcpy, fd, pin, map, n);
            if (t < best) best = t;
        }
        printf("memcpy map -> pinned T=%2d : %6.2f GB/s\n", Ts[k], n / best / 1e9);
    }
    // pageable (malloc) -> pinned, to tell the page cache apart from plain memory
    char *heap = (char *)malloc(n);
    memset(heap, 'C', n);
    for (int k = 0; k < 8; k += 2) {
        double t = run_threads(Ts[k], do_memcpy, fd, pin, heap, n);
        t = run_threads(Ts[k], do_memcpy, fd, pin, heap, n);
        printf("memcpy heap -> pinned T=%2d : %6.2f GB/s\n", Ts[k], n / t / 1e9);
    }
    // ---- output side: pinned -> a new file in the page cache
    {
        char opath[512];
        snprintf(opath, sizeof opath, "%s.out", path);
        for (int k = 0; k < 8; k++) {
            int T = Ts[k];
            if (T > 32) break;
            int ofd = open(opath, O_CREAT | O_RDWR | O_TRUNC, 0600);
            double t1 = now();
            double t = run_threads(T, do_pwrite, ofd, 0, pin, n);
            printf("pwrite pinned -> new file T=%2d : %6.2f GB/s\n", T, n / t / 1e9);
            (void)t1;
            close(ofd);
            unlink(opath);
        }
        for (int k = 0; k < 8; k++) {
            int T = Ts[k];
            if (T > 32) break;
            int ofd = open(opath, O_CREAT | O_RDWR | O_TRUNC, 0600);
            double t1 = now();
            if (ftruncate(ofd, n)) return 1;
            char *om = (char *)mmap(0, n, PROT_READ | PROT_WRITE, MAP_SHARED, ofd, 0);
            run_threads(T, do_memcpy_out, ofd, om, pin, n);
            munmap(om, n);
            double t = now() - t1;
            printf("memcpy pinned -> mapped new file T=%2d : %6.2f GB/s (ftruncate + mmap + copy + munmap)\n", T, n / t / 1e9);
            close(ofd);
            unlink(opath);
        }
        {   // the same in 64 MB windows mapped one after the other (what a streaming writer does)
            int T = 16;
            int ofd = open(opath, O_CREAT | O_RDWR | O_TRUNC, 0600);
            double t1 = now();
            const size_t win = 64 << 20;
            for (size_t o = 0; o < n; o += win) {
                if (ftruncate(ofd, o + win)) return 1;
                char *om = (char *)mmap(0, win, PROT_READ | PROT_WRITE, MAP_SHARED, ofd, o);
                run_threads(T, do_memcpy_out, ofd, om, pin + o, win);
                munmap(om, win);
            }
            double t = now() - t1;
            printf("memcpy pinned -> 64 MB windows of a new file T=16 : %6.2f GB/s\n", n / t / 1e9);
            close(ofd);
            unlink(opath);
        }
    }
    // pinned allocation rate
    {
        double t1 = now();
        char *p2;
        CK(cudaHostAlloc(&p2, 256 << 20, cudaHostAllocDefault));
        double t = now() - t1;
        printf("cudaHostAlloc 256 MB        : %6.3f s (%6.2f GB/s)\n", t, (256 << 20) / t / 1e9);
        t1 = now();
        cudaFreeHost(p2);
        printf("cudaFreeHost 256 MB         : %6.3f s\n", now() - t1);
        t1 = now();
        CK(cudaHostAlloc(&p2, 64 << 20, cudaHostAllocDefault));
        t = now() - t1;
        printf("cudaHostAlloc 64 MB         : %6.3f s (%6.2f GB/s)\n", t, (64 << 20) / t / 1e9);
        cudaFreeHost(p2);
    }
    cudaStream_t st;
    CK(cudaStreamCreate(&st));
    {
        double t1 = now();
        CK(cudaMemcpyAsync(pin, dev, n, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        printf("D2H to pinned               : %6.2f GB/s\n", n / (now() - t1) / 1e9);
    }
    t0 = now();
    CK(cudaMemcpyAsync(dev, pin, n, cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));
    printf("H2D from pinned             : %6.2f GB/s\n", n / (now() - t0) / 1e9);
    // H2D straight from the mapping (pageable path of the driver)
    t0 = now();
    CK(cudaMemcpyAsync(dev, map, n, cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));
    printf("H2D from mapping (pageable) : %6.2f GB/s\n", n / (now() - t0) / 1e9);
    // register the mapping in pieces, then copy
    size_t piece = 256 << 20;
    for (int ro = 1; ro >= 0; ro--) {
        t0 = now();
        cudaError_t e = cudaSuccess;
        for (size_t o = 0; o < n && e == cudaSuccess; o += piece)
            e = cudaHostRegister(map + o, piece, ro ? cudaHostRegisterReadOnly : cudaHostRegisterDefault);
        double tr = now() - t0;
        if (e != cudaSuccess) {
            printf("cudaHostRegister(%s) of the mapping: %s\n", ro ? "read-only" : "default", cudaGetErrorString(e));
            cudaGetLastError();
            for (size_t o = 0; o < n; o += piece) cudaHostUnregister(map + o);
            cudaGetLastError();
            continue;
        }
        printf("cudaHostRegister(%s) mapping : %6.3f s (%6.2f GB/s)\n", ro ? "read-only" : "default", tr, n / tr / 1e9);
        t0 = now();
        CK(cudaMemcpyAsync(dev, map, n, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
        printf("H2D from registered mapping : %6.2f GB/s\n", n / (now() - t0) / 1e9);
        t0 = now();
        for (size_t o = 0; o < n; o += piece) cudaHostUnregister(map + o);
        printf("cudaHostUnregister          : %6.3f s\n", now() - t0);
        break;
    }
    // private heap memory registered (the cost of pinning alone)
    t0 = now();
    if (cudaHostRegister(heap, n, cudaHostRegisterDefault) == cudaSuccess) {
        double tr = now() - t0;
        printf("cudaHostRegister heap       : %6.3f s (%6.2f GB/s)\n", tr, n / tr / 1e9);
        cudaHostUnregister(heap);
    }
    close(fd);
    unlink(path);
    return 0;
}
