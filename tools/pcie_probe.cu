// pcie_probe.cu -- what the host <-> GPU links of THIS box deliver, measured without any of the library, torch
// or Python: bare cudaHostAlloc buffers and cudaMemcpyAsync, each direction alone and both at once, on 1..N GPUs
// at the same time, driven (a) from one process with one thread, and (b) from N processes, one per GPU.
// It is the independent check of the end-to-end ceiling bench.py reports (e2e.pcie): if this probe reads well
// above what the end-to-end arm reaches the arm is the problem, if not the box is.
//
// build: nvcc -O2 -o tools/pcie_probe tools/pcie_probe.cu      run: tools/pcie_probe [max_gpus] [MiB per copy] [reps]
// Output: one line per (mode, gpus, phase): per-GPU GB/s (slowest / mean) and the aggregate over the box.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(2); } } while (0)

static double now_s(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }

struct Gpu { void *hu, *hd, *du, *dd; cudaStream_t su, sd; };

static void gpu_init(Gpu* g, int dev, size_t bytes) {
    CK(cudaSetDevice(dev));
    CK(cudaHostAlloc(&g->hu, bytes, cudaHostAllocDefault)); CK(cudaHostAlloc(&g->hd, bytes, cudaHostAllocDefault));
    memset(g->hu, 1, bytes); memset(g->hd, 2, bytes);
    CK(cudaMalloc(&g->du, bytes)); CK(cudaMalloc(&g->dd, bytes));
    CK(cudaStreamCreateWithFlags(&g->su, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&g->sd, cudaStreamNonBlocking));
}

// phase: 1 = host->device, 2 = device->host, 3 = both at once.  Returns seconds for `reps` copies per direction.
static double run_phase(Gpu* g, const int* devs, int n, size_t bytes, int reps, int phase) {
    for (int i = 0; i < n; i++) { CK(cudaSetDevice(devs[i])); CK(cudaDeviceSynchronize()); }
    const double t0 = now_s();
    for (int r = 0; r < reps; r++)
        for (int i = 0; i < n; i++) {
            CK(cudaSetDevice(devs[i]));
            if (phase & 1) CK(cudaMemcpyAsync(g[i].du, g[i].hu, bytes, cudaMemcpyHostToDevice, g[i].su));
            if (phase & 2) CK(cudaMemcpyAsync(g[i].hd, g[i].dd, bytes, cudaMemcpyDeviceToHost, g[i].sd));
        }
    for (int i = 0; i < n; i++) { CK(cudaSetDevice(devs[i])); CK(cudaStreamSynchronize(g[i].su)); CK(cudaStreamSynchronize(g[i].sd)); }
    return now_s() - t0;
}

static const char* kPhase[4] = {"", "h2d_alone", "d2h_alone", "both_at_once"};

int main(int argc, char** argv) {
    int maxg = argc > 1 ? atoi(argv[1]) : 8;
    const size_t mib = argc > 2 ? (size_t)atoi(argv[2]) : 1024;
    const int reps = argc > 3 ? atoi(argv[3]) : 4;
    const size_t bytes = mib << 20;
    // the device count must be learned without creating a CUDA context in this (parent) process: fork first
    int* shared = (int*)mmap(NULL, 4096, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    double* res = (double*)(shared + 64);            // per child: seconds of the phase
    if (fork() == 0) { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) n = 0; shared[0] = n; _exit(0); }
    wait(NULL);
    const int ndev = shared[0];
    if (ndev == 0) { fprintf(stderr, "no CUDA device\n"); return 1; }
    if (maxg > ndev) maxg = ndev;
    printf("pcie_probe: %d GPU(s) visible, %zu MiB per copy, %d copies per direction per phase, GB = 1e9 bytes\n", ndev, mib, reps);
    fflush(stdout);                                   // children inherit the stdio buffer
    for (int n = 1; n <= maxg; n *= 2) {
        // (a) one process, one thread, n GPUs
        if (fork() == 0) {
            Gpu* g = (Gpu*)calloc(n, sizeof(Gpu));
            int devs[64];
            for (int i = 0; i < n; i++) { devs[i] = i; gpu_init(&g[i], i, bytes); }
            run_phase(g, devs, n, bytes, 1, 3);                       // warm-up
            for (int phase = 1; phase <= 3; phase++) {
                const double t = run_phase(g, devs, n, bytes, reps, phase);
                const double per = (double)bytes * reps / t / 1e9;
                printf("one_process   gpus=%d %-13s per-GPU per-direction %6.1f GB/s   box per-direction %7.1f GB/s\n", n, kPhase[phase], per, per * n);
            }
            fflush(stdout);
            _exit(0);
        }
        wait(NULL);
        // (b) n processes, one GPU each, started together
        for (int phase = 1; phase <= 3; phase++) {
            shared[1] = 0; shared[2] = 0;
            for (int c = 0; c < n; c++) {
                if (fork() == 0) {
                    Gpu g; int dev = c;
                    gpu_init(&g, dev, bytes);
                    run_phase(&g, &dev, 1, bytes, 1, 3);
                    __sync_fetch_and_add(&shared[1], 1);
                    while (__atomic_load_n(&shared[1], __ATOMIC_SEQ_CST) < n) {}      // all children ready
                    res[c] = run_phase(&g, &dev, 1, bytes, reps, phase);
                    _exit(0);
                }
            }
            for (int c = 0; c < n; c++) wait(NULL);
            double worst = 0, sum = 0;
            for (int c = 0; c < n; c++) { const double gb = (double)bytes * reps / res[c] / 1e9; sum += gb; if (c == 0 || gb < worst) worst = gb; }
            printf("n_processes   gpus=%d %-13s per-GPU per-direction %6.1f GB/s (slowest; mean %6.1f)   box per-direction %7.1f GB/s\n",
                   n, kPhase[phase], worst, sum / n, sum);
            fflush(stdout);
        }
    }
    return 0;
}
