// microbench.cu -- memory-system ceilings for the MERL eval+pdf access pattern on one B200.
//   copy40      : 6 float4 read streams + 4 float4 write streams (40 B/query), no gather
//   gather16    : idx stream (4 B) + one random 16 B gather from a 23.3 MB table + 4 B write
//   stream+gath : copy40 pattern + the gather (the kernel's memory traffic without its math)
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/microbench tools/microbench.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void copy40(const float4* a, const float4* b, const float4* c, const float4* d, const float4* e,
                       const float4* f, float4* o0, float4* o1, float4* o2, float4* o3, long n4) {
    for (long g = blockIdx.x * (long)blockDim.x + threadIdx.x; g < n4; g += (long)gridDim.x * blockDim.x) {
        float4 x = __ldcs(a + g), y = __ldcs(b + g), z = __ldcs(c + g), u = __ldcs(d + g), v = __ldcs(e + g), w = __ldcs(f + g);
        __stcs(o0 + g, make_float4(x.x + u.x, x.y + u.y, x.z + u.z, x.w + u.w));
        __stcs(o1 + g, make_float4(y.x + v.x, y.y + v.y, y.z + v.z, y.w + v.w));
        __stcs(o2 + g, make_float4(z.x + w.x, z.y + w.y, z.z + w.z, z.w + w.w));
        __stcs(o3 + g, make_float4(x.x * w.x, y.y * v.y, z.z * u.z, x.w));
    }
}
template <int MODE>  // 0: __ldg  1: ld.global.cg (L2 only)
__global__ void gather16(const float4* __restrict__ table, const int4* __restrict__ idx, float4* out, long n4) {
    for (long g = blockIdx.x * (long)blockDim.x + threadIdx.x; g < n4; g += (long)gridDim.x * blockDim.x) {
        int4 i = __ldcs(idx + g);
        float4 a, b, c, d;
        if (MODE == 0) { a = __ldg(table + i.x); b = __ldg(table + i.y); c = __ldg(table + i.z); d = __ldg(table + i.w); }
        else { a = __ldcg(table + i.x); b = __ldcg(table + i.y); c = __ldcg(table + i.z); d = __ldcg(table + i.w); }
        __stcs(out + g, make_float4(a.x + a.y, b.x + b.z, c.x + c.w, d.x + d.y));
    }
}
__global__ void stream_gather(const float4* __restrict__ table, const int4* __restrict__ idx, const float4* a,
                              const float4* b, const float4* c, const float4* d, const float4* e, float4* o0,
                              float4* o1, float4* o2, float4* o3, long n4) {
    // 5 float4 input streams + the index stream stand in for the 6 direction streams (24 B/query)
    for (long g = blockIdx.x * (long)blockDim.x + threadIdx.x; g < n4; g += (long)gridDim.x * blockDim.x) {
        int4 i = __ldcs(idx + g);
        float4 x = __ldcs(a + g), y = __ldcs(b + g), z = __ldcs(c + g), u = __ldcs(d + g), v = __ldcs(e + g);
        float4 p = __ldg(table + i.x), q = __ldg(table + i.y), r = __ldg(table + i.z), s = __ldg(table + i.w);
        __stcs(o0 + g, make_float4(p.x * x.x, q.x * x.y, r.x * x.z, s.x * x.w));
        __stcs(o1 + g, make_float4(p.y * y.x, q.y * y.y, r.y * y.z, s.y * y.w));
        __stcs(o2 + g, make_float4(p.z * z.x, q.z * z.y, r.z * z.z, s.z * z.w));
        __stcs(o3 + g, make_float4(u.x + v.x, u.y + v.y, u.z + v.z, u.w + v.w));
    }
}
// red16: idx stream (4 B) + one random red.global.add.v4.f32 into the 23.3 MB table
// stream+red: the backward kernel's memory traffic without its math (36 B/query read + the RED)
__device__ __forceinline__ void red4(float4* a, float x, float y, float z, float w) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
__global__ void red16(float4* table, const int4* __restrict__ idx, long n4) {
    for (long g = blockIdx.x * (long)blockDim.x + threadIdx.x; g < n4; g += (long)gridDim.x * blockDim.x) {
        int4 i = __ldcs(idx + g);
        red4(table + i.x, 1.f, 2.f, 3.f, 1.f); red4(table + i.y, 1.f, 2.f, 3.f, 1.f);
        red4(table + i.z, 1.f, 2.f, 3.f, 1.f); red4(table + i.w, 1.f, 2.f, 3.f, 1.f);
    }
}
__global__ void stream_red(float4* table, const int4* __restrict__ idx, const float4* a, const float4* b,
                           const float4* c, const float4* d, const float4* e, const float4* f, const float4* h,
                           const float4* k, long n4) {
    for (long g = blockIdx.x * (long)blockDim.x + threadIdx.x; g < n4; g += (long)gridDim.x * blockDim.x) {
        int4 i = __ldcs(idx + g);
        float4 x = __ldcs(a + g), y = __ldcs(b + g), z = __ldcs(c + g), u = __ldcs(d + g), v = __ldcs(e + g),
               w = __ldcs(f + g), p = __ldcs(h + g), q = __ldcs(k + g);
        red4(table + i.x, x.x + u.x, y.x + v.x, z.x + w.x, p.x + q.x); red4(table + i.y, x.y + u.y, y.y + v.y, z.y + w.y, p.y + q.y);
        red4(table + i.z, x.z + u.z, y.z + v.z, z.z + w.z, p.z + q.z); red4(table + i.w, x.w + u.w, y.w + v.w, z.w + w.w, p.w + q.w);
    }
}
__global__ void smid_gather(const float4* __restrict__ table, const int4* __restrict__ idx, const float4* a,
                            const float4* b, const float4* c, const float4* d, const float4* e, float4* o0,
                            float4* o1, float4* o2, float4* o3, long n4, int* smid_out) {
    if (threadIdx.x == 0) { unsigned s; asm("mov.u32 %0, %%smid;" : "=r"(s)); smid_out[blockIdx.x] = (int)s; }
    for (long g = blockIdx.x * (long)blockDim.x + threadIdx.x; g < n4; g += (long)gridDim.x * blockDim.x) {
        int4 i = __ldcs(idx + g);
        float4 x = __ldcs(a + g), y = __ldcs(b + g), z = __ldcs(c + g), u = __ldcs(d + g), v = __ldcs(e + g);
        float4 p = __ldg(table + i.x), q = __ldg(table + i.y), r = __ldg(table + i.z), s = __ldg(table + i.w);
        __stcs(o0 + g, make_float4(p.x * x.x, q.x * x.y, r.x * x.z, s.x * x.w));
        __stcs(o1 + g, make_float4(p.y * y.x, q.y * y.y, r.y * y.z, s.y * y.w));
        __stcs(o2 + g, make_float4(p.z * z.x, q.z * z.y, r.z * z.z, s.z * z.w));
        __stcs(o3 + g, make_float4(u.x + v.x, u.y + v.y, u.z + v.z, u.w + v.w));
    }
}
template <typename F> float timeit(F f, int iters = 10) {
    for (int i = 0; i < 3; ++i) f();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int i = 0; i < iters; ++i) f();
    cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms / iters;
}
int main(int argc, char** argv) {
    const long n = 1L << (argc > 1 ? atoi(argv[1]) : 26), n4 = n / 4;
    const int ncell = 90 * 90 * 180;
    int dev = 0, l2 = 0, pers = 0, win = 0, sms = 0;
    cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, dev);
    cudaDeviceGetAttribute(&pers, cudaDevAttrMaxPersistingL2CacheSize, dev);
    cudaDeviceGetAttribute(&win, cudaDevAttrMaxAccessPolicyWindowSize, dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    printf("L2 %d B, max persisting %d B, max window %d B, SMs %d\n", l2, pers, win, sms);
    float4 *in[6], *out[4], *table; int4* idx;
    for (auto& p : in) { CK(cudaMalloc(&p, n * 4)); CK(cudaMemset(p, 0, n * 4)); }
    for (auto& p : out) CK(cudaMalloc(&p, n * 4));
    CK(cudaMalloc(&table, ncell * 16L)); CK(cudaMemset(table, 0, ncell * 16L));
    CK(cudaMalloc(&idx, n * 4));
    { std::vector<int> h(n); unsigned s = 12345u; for (long i = 0; i < n; ++i) { s = s * 1664525u + 1013904223u; h[i] = (int)((s >> 8) % ncell); }
      CK(cudaMemcpy(idx, h.data(), n * 4, cudaMemcpyHostToDevice)); }
    for (int bps : {2, 4, 8}) {
        const int grid = sms * bps;
        float ms = timeit([&] { copy40<<<grid, 256>>>(in[0], in[1], in[2], in[3], in[4], in[5], out[0], out[1], out[2], out[3], n4); });
        printf("copy40        grid %4d: %7.3f ms  %7.1f GB/s  (%6.1f Gq/s at 40 B/q)\n", grid, ms, 40.0 * n / ms / 1e6, n / ms / 1e6);
        ms = timeit([&] { gather16<0><<<grid, 256>>>(table, idx, out[0], n4); });
        printf("gather16 ldg  grid %4d: %7.3f ms  %7.1f Ggather/s  (%6.1f GB/s of 32 B sectors)\n", grid, ms, n / ms / 1e6, 32.0 * n / ms / 1e6);
        ms = timeit([&] { gather16<1><<<grid, 256>>>(table, idx, out[0], n4); });
        printf("gather16 ldcg grid %4d: %7.3f ms  %7.1f Ggather/s\n", grid, ms, n / ms / 1e6);
        ms = timeit([&] { stream_gather<<<grid, 256>>>(table, idx, in[0], in[1], in[2], in[3], in[4], out[0], out[1], out[2], out[3], n4); });
        printf("stream+gather grid %4d: %7.3f ms  %7.1f Gq/s  (%6.1f GB/s algorithmic at 40 B/q)\n", grid, ms, n / ms / 1e6, 40.0 * n / ms / 1e6);
        ms = timeit([&] { red16<<<grid, 256>>>(table, idx, n4); });
        printf("red16 v4      grid %4d: %7.3f ms  %7.1f Gred/s\n", grid, ms, n / ms / 1e6);
        ms = timeit([&] { stream_red<<<grid, 256>>>(table, idx, in[0], in[1], in[2], in[3], in[4], in[5], out[0], out[1], n4); });
        printf("stream+red    grid %4d: %7.3f ms  %7.1f Gq/s  (%6.1f GB/s algorithmic at 36 B/q)\n", grid, ms, n / ms / 1e6, 36.0 * n / ms / 1e6);
    }
    // stream+gather read 134 / 102 / 134 Gq/s at grids 296 / 592 / 1184.  Sweep the grid in
    // steps of one CTA per SM (and two off-multiples) with more repetitions, to tell a placement effect from noise.
    printf("-- grid sweep (20 launches each) --\n");
    for (int grid : {sms * 1, sms * 2, sms * 3, 444 + 37, sms * 4, sms * 4 + 74, sms * 5, sms * 6, sms * 7, sms * 8, sms * 12, sms * 16}) {
        float a = timeit([&] { stream_gather<<<grid, 256>>>(table, idx, in[0], in[1], in[2], in[3], in[4], out[0], out[1], out[2], out[3], n4); }, 20);
        float b = timeit([&] { copy40<<<grid, 256>>>(in[0], in[1], in[2], in[3], in[4], in[5], out[0], out[1], out[2], out[3], n4); }, 20);
        float c = timeit([&] { stream_gather<<<grid, 256>>>(table, idx, in[0], in[1], in[2], in[3], in[4], out[0], out[1], out[2], out[3], n4); }, 20);
        printf("grid %5d (%5.2f CTA/SM): stream+gather %7.1f / %7.1f Gq/s (two runs)   copy40 %7.1f Gq/s\n", grid, (double)grid / sms,
               n / a / 1e6, n / c / 1e6, n / b / 1e6);
    }
    // Where do the CTAs of each grid land?  (placement imbalance vs. a memory-system effect)
    printf("-- CTAs per SM (histogram: count x SMs) and the effect of breaking the 2^28-byte stream alignment --\n");
    int* d_smid; CK(cudaMalloc(&d_smid, 4096 * 4));
    for (int grid : {sms * 2, sms * 3, 444 + 37, sms * 4, sms * 5, sms * 6, sms * 7, sms * 8}) {
        CK(cudaMemset(d_smid, 0xff, 4096 * 4));
        smid_gather<<<grid, 256>>>(table, idx, in[0], in[1], in[2], in[3], in[4], out[0], out[1], out[2], out[3], n4, d_smid);
        CK(cudaDeviceSynchronize());
        std::vector<int> h(grid); CK(cudaMemcpy(h.data(), d_smid, grid * 4, cudaMemcpyDeviceToHost));
        std::vector<int> per(256, 0); for (int v : h) if (v >= 0 && v < 256) per[v]++;
        int hist[17] = {0}; for (int sm = 0; sm < 256; ++sm) if (per[sm] > 0 || sm < sms) hist[per[sm] > 16 ? 16 : per[sm]]++;
        printf("grid %5d:", grid); for (int c = 0; c <= 16; ++c) if (hist[c]) printf("  %dx%d", c, hist[c]); printf("\n");
    }
    // same pattern, every stream shifted by a different odd multiple of 4 KB + 256 B (arrays allocated with slack)
    {
        float4 *pin[6], *pout[4];
        const long slack = 1 << 20;
        for (int k = 0; k < 6; ++k) { CK(cudaMalloc(&pin[k], n * 4 + slack)); CK(cudaMemset(pin[k], 0, n * 4 + slack)); }
        for (int k = 0; k < 4; ++k) CK(cudaMalloc(&pout[k], n * 4 + slack));
        auto sh = [&](float4* p, int k) { return (float4*)((char*)p + (long)(2 * k + 1) * 4352); };
        for (int grid : {sms * 2, sms * 3, sms * 4, sms * 5, sms * 6, sms * 8}) {
            float a = timeit([&] { stream_gather<<<grid, 256>>>(table, idx, sh(pin[0], 0), sh(pin[1], 1), sh(pin[2], 2), sh(pin[3], 3), sh(pin[4], 4),
                                                               sh(pout[0], 5), sh(pout[1], 6), sh(pout[2], 7), sh(pout[3], 8), n4); }, 20);
            float b = timeit([&] { copy40<<<grid, 256>>>(sh(pin[0], 0), sh(pin[1], 1), sh(pin[2], 2), sh(pin[3], 3), sh(pin[4], 4), sh(pin[5], 9),
                                                         sh(pout[0], 5), sh(pout[1], 6), sh(pout[2], 7), sh(pout[3], 8), n4); }, 20);
            printf("shifted streams grid %5d: stream+gather %7.1f Gq/s   copy40 %7.1f Gq/s\n", grid, n / a / 1e6, n / b / 1e6);
        }
    }
    // 200-launch hold of the no-math pattern at the kernel's own grid: does the PATTERN alone hit the power cap?
    {
        const int grid = sms * 3;
        for (int w = 0; w < 6; ++w) {
            float a = timeit([&] { stream_gather<<<grid, 256>>>(table, idx, in[0], in[1], in[2], in[3], in[4], out[0], out[1], out[2], out[3], n4); }, 200);
            printf("hold stream+gather grid %d, window %d (200 launches): %7.1f Gq/s\n", grid, w, n / a / 1e6);
        }
    }
    return 0;
}
