// Throughput / latency probe of the packed fp32x2 instructions of sm_100 (FFMA2) against the scalar form:
// CHAINS independent chains per thread, clock64 around the loop.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a
// Measured on B200 (round 2): 35.7 T scalar FMA/s either way at full occupancy (FFMA2 issues at half the rate, i.e.
// the same FMA throughput for half the issue slots), dependent-chain step 4.6 cycles packed vs 5.2 for two scalar FFMAs.
#include <cuda_runtime.h>
#include <cstdio>
template <int CHAINS, bool PACKED>
__global__ void __launch_bounds__(256) probe(float* out, int iters, long long* cycles) {
    float2 a[CHAINS];
    for (int c = 0; c < CHAINS; ++c) a[c] = make_float2(1.f + threadIdx.x * 1e-6f + c, 2.f + c);
    const float2 m = make_float2(1.0000001f, 0.9999999f), k = make_float2(1e-7f, -1e-7f);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            if (PACKED) a[c] = __ffma2_rn(a[c], m, k);
            else { a[c].x = fmaf(a[c].x, m.x, k.x); a[c].y = fmaf(a[c].y, m.y, k.y); }
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
    for (int c = 0; c < CHAINS; ++c) s += a[c].x + a[c].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}
template <int CHAINS, bool PACKED>
void run(const char* name, int ctas_per_sm) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float* out; long long* cyc; cudaMalloc(&out, sizeof(float) * sms * ctas_per_sm * 256); cudaMallocManaged(&cyc, 8);
    const int iters = 20000;
    probe<CHAINS, PACKED><<<sms * ctas_per_sm, 256>>>(out, iters, cyc);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    probe<CHAINS, PACKED><<<sms * ctas_per_sm, 256>>>(out, iters, cyc);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double fma = 2.0 * CHAINS * (double)iters * 256.0 * sms * ctas_per_sm;   // scalar FMAs
    printf("%-16s chains %d  CTAs/SM %d: %8.3f ms  %7.1f GFMA/s  (%.2f cycles per chain step)\n", name, CHAINS, ctas_per_sm, ms,
           fma / ms * 1e-6, (double)*cyc / iters / CHAINS);
    cudaFree(out); cudaFree(cyc);
}
int main() {
    run<1, false>("scalar FFMA x2", 1); run<1, true>("packed FFMA2", 1);
    run<4, false>("scalar FFMA x2", 1); run<4, true>("packed FFMA2", 1);
    run<8, false>("scalar FFMA x2", 4); run<8, true>("packed FFMA2", 4);
    run<8, false>("scalar FFMA x2", 8); run<8, true>("packed FFMA2", 8);
    return 0;
}
