// alu_peak.cu -- measures what one B200 SM sustains for the integer instruction classes the record
// kernel is made of: LOP3 and PRMT (ALU pipe), IMAD.WIDE (FMA pipe), LDS.32 (shared-memory pipe).
// "INT32 lanes/clk/SM ... measure, do not assume".  Not part of the product.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/alu_peak tools/alu_peak.cu
// run:   tools/alu_peak            (prints one JSON object)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int kIters = 4096, kChains = 8;

template <int OP>
__global__ void __launch_bounds__(1024) probe(uint32_t* out, uint32_t seed, long long* cycles) {
    __shared__ __align__(8192) uint32_t tab[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) tab[i] = i * 2654435761u;
    __syncthreads();
    uint32_t a[kChains];
    unsigned long long w[kChains];
#pragma unroll
    for (int c = 0; c < kChains; c++) { a[c] = seed + threadIdx.x * 97 + c; w[c] = a[c]; }
    const uint32_t k1 = seed | 1, k2 = seed * 3 + 7;
    const uint32_t lane_base = (uint32_t)__cvta_generic_to_shared(tab) + 4u * (threadIdx.x & 31);   // tab is 8 KiB aligned below
    const long long t0 = clock64();
#pragma unroll 4
    for (int it = 0; it < kIters; it++) {
#pragma unroll
        for (int c = 0; c < kChains; c++) {
            if (OP == 0) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[c]) : "r"(k1), "r"(k2));
            if (OP == 1) asm volatile("prmt.b32 %0, %0, %1, 0x7604;" : "+r"(a[c]) : "r"(k1));
            if (OP == 2) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[c]) : "r"(k1), "r"(k2));
            if (OP == 4) { uint32_t m; asm volatile("match.any.sync.b32 %0, %1, 0xffffffff;" : "=r"(m) : "r"(a[c] & 0xffu)); a[c] += m; }
            if (OP == 5) { a[c] += __ballot_sync(0xffffffffu, a[c] & 1u); }
            if (OP == 3) {   // one LOP3 builds the address (row from the loaded value | base + 4 * lane), like the kernel's single PRMT
                uint32_t ad;
                asm volatile("lop3.b32 %0, %1, 0x1f80, %2, 0xea;" : "=r"(ad) : "r"(a[c]), "r"(lane_base));
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(a[c]) : "r"(ad));
            }
        }
    }
    const long long t1 = clock64();
    uint32_t r = 0;
#pragma unroll
    for (int c = 0; c < kChains; c++) r ^= a[c] ^ (uint32_t)w[c] ^ (uint32_t)(w[c] >> 32);
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
static double run(int sms, uint32_t* out, long long* cyc, float* ms_out) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<OP><<<sms, 1024>>>(out, 12345u, cyc);          // warm-up
    cudaEventRecord(e0);
    probe<OP><<<sms, 1024>>>(out, 12345u, cyc);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(ms_out, e0, e1);
    long long h[1024];
    cudaMemcpy(h, cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < sms; i++) avg += (double)h[i]; avg /= sms;
    // warp instructions issued per SM: 32 warps x kIters x kChains
    return 32.0 * kIters * kChains / avg;                  // warp instructions per clock per SM
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, sizeof(uint32_t) * sms * 1024); cudaMalloc(&cyc, sizeof(long long) * sms);
    float ms[4];
    const double lop3 = run<0>(sms, out, cyc, &ms[0]), prmt = run<1>(sms, out, cyc, &ms[1]);
    const double imadw = run<2>(sms, out, cyc, &ms[2]), lds = run<3>(sms, out, cyc, &ms[3]);
    float ms2[2];
    const double match = run<4>(sms, out, cyc, &ms2[0]), ballot = run<5>(sms, out, cyc, &ms2[1]);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    printf("{\"device\": \"%s\", \"sms\": %d, \"max_sm_clock_mhz\": %.0f, \"warp_inst_per_clk_per_sm\": {\"lop3\": %.3f, \"prmt\": %.3f, "
           "\"imad_wide\": %.3f, \"lds32\": %.3f, \"match_any_plus_iadd\": %.3f, \"ballot_plus_2alu\": %.3f}, \"kernel_ms\": {\"lop3\": %.3f, \"prmt\": %.3f, \"imad_wide\": %.3f, \"lds32\": %.3f}, "
           "\"note\": \"32 warps per SM, 8 independent chains per thread, clock64 around 4096 x 8 instructions; 1 CTA of 1024 threads per SM; lds32 = conflict-free LDS.32 with one LOP3 of address math each\"}\n",
           p.name, sms, clk_khz / 1000.0, lop3, prmt, imadw, lds, match, ballot, ms[0], ms[1], ms[2], ms[3]);
    return cudaGetLastError() == cudaSuccess ? 0 : 1;
}
