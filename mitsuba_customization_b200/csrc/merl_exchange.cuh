// merl_exchange.cuh -- the multi-GPU exchange step as ONE kernel over NVLink peer memory (sm_100a).
//
// the backward keeps a private fp32 gradient table per GPU and sums them once; a
// renderer sums its film the same way.  Round 1 called ncclAllReduce for that: 23.3 MB took 0.122 ms
// on 8 GPUs (about 334 GB/s of bus bandwidth), as long as the scatter it follows.  Here every rank's
// buffer is mapped into every other rank's address space (CUDA IPC over NVLink / NVSwitch) and one
// persistent kernel per rank does the whole exchange:
//
//   barrier A   all ranks have entered the kernel, i.e. their producers (scatter / render) are done
//   reduce      rank r owns slab r = elements [r n/R, (r+1) n/R): it PULLS that slab from all R buffers
//               (16-byte ld.relaxed.sys, R loads in flight per thread), sums them in rank order,
//               optionally applies the table's validity mask (backward_mask_kernel's job, fused: the
//               mask pass now touches 1/R of the table per GPU), and
//   broadcast   PUSHES the result into slab r of all R buffers (st.relaxed.sys: fire and forget)
//   barrier B   every push into this rank's buffer has landed; nobody still reads it
//
// Only the CTA with the same index on the owning rank ever touches a given element between the two
// barriers, so no barrier is needed between reduce and broadcast, and every rank ends with the SAME
// bits (one rank computes each sum).  NVLink is full duplex: pulls (inbound data) and pushes (outbound
// data) of different ranks overlap.  Bytes per GPU and direction: (R-1)/R of the buffer each way.
//
// Cross-GPU barrier: CTA i of rank r writes the epoch into word [i][r] of every peer's flag block
// (st.release.sys after a system fence) and spins on its own words [i][*] (ld.acquire.sys).  Epochs
// only grow, live in device memory (epoch[i], advanced by the kernel itself), so the kernel has no
// host-side state and can be captured in a CUDA graph.  All CTAs of the grid must be co-resident on
// every rank: the host sizes the grid from the occupancy API and the same formula on every rank.  A
// spin that exceeds ~4 s sets a sticky status word instead of hanging the GPU.
// Included exactly once, by merl_b200.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace merl {

constexpr int kMaxRanks = 16;
constexpr int kXchgThreads = 256;
constexpr int kXchgMaxCtas = 1024;
constexpr int kXchgFlagWords = kXchgMaxCtas * kMaxRanks + kXchgMaxCtas;   // arrive[cta][src rank] | epoch[cta]
constexpr long long kXchgTimeoutCycles = 8000000000ll;                          // ~4 s at 1.9 GHz

struct PeerView {
    float4* data[kMaxRanks];       // rank p's buffer in THIS process's address space (own: the local pointer)
    uint32_t* flags[kMaxRanks];    // rank p's flag block
    int nranks, rank;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_peer(const float4* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}
__device__ __forceinline__ void st_peer(float4* p, const float4 v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// CTA `cta` of every rank meets here.  The CTA barrier orders every thread's earlier peer stores before
// the flag writers' st.release.sys, and a release is cumulative over what its thread has synchronised
// with, so no thread needs a fence of its own.  (Round 2's first version had every thread execute
// __threadfence_system() first: 8 MEMBAR.SC.SYS per CTA and barrier, and a fixed cost of ~70 us per
// exchange on 8 GPUs whatever the size -- profiles/r02_exchange_n8.txt.  -DMERL_XCHG_THREAD_FENCE=1
// brings it back for an A/B.)
#ifndef MERL_XCHG_THREAD_FENCE
#define MERL_XCHG_THREAD_FENCE 0
#endif
// Returns false (to every thread of the CTA) when a wait ran into the timeout: the peers' data cannot be
// trusted then, and the callers skip their data phase instead of summing unsynchronised buffers.
__device__ __forceinline__ bool xchg_barrier(const PeerView& pv, const int cta, const uint32_t epoch, int* status) {
    __shared__ int timed_out;
#if MERL_XCHG_THREAD_FENCE
    __threadfence_system();
#endif
    if (threadIdx.x == 0) timed_out = 0;
    __syncthreads();
    if ((int)threadIdx.x < pv.nranks) {
        const int p = threadIdx.x;
        st_release_sys(pv.flags[p] + cta * kMaxRanks + pv.rank, epoch);
        const uint32_t* mine = pv.flags[pv.rank] + cta * kMaxRanks + p;
        const long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
            if (clock64() - t0 > kXchgTimeoutCycles) {
                *reinterpret_cast<volatile int*>(status) = 1;   // host-mapped word
                timed_out = 1;
                break;
            }
        }
    }
    __syncthreads();
    const bool ok = timed_out == 0;
    __syncthreads();   // everyone has read the flag before the next barrier call resets it
    return ok;
}

__device__ __forceinline__ float4 mask_cell(float4 g, const float bits) {
    const int m = (int)bits;
    if (!(m & 1)) g.x = 0.f;
    if (!(m & 2)) g.y = 0.f;
    if (!(m & 4)) g.z = 0.f;
    return g;
}

// NR > 0: number of ranks known at compile time (all R loads of an element in flight at once);
// NR == 0: any rank count, loads issued in a loop.
template <int NR, int UNROLL>
__global__ void __launch_bounds__(kXchgThreads) peer_allreduce_kernel(const PeerView pv, const int64_t n4,
                                                                      const float4* __restrict__ mask_cells,
                                                                      uint32_t* __restrict__ epoch_words,
                                                                      int* __restrict__ status) {
    __shared__ uint32_t e_s;
    const int cta = blockIdx.x;
    if (threadIdx.x == 0) e_s = epoch_words[cta];
    __syncthreads();
    const uint32_t e0 = e_s;
    const int R = NR > 0 ? NR : pv.nranks;
    const bool peers_ready = xchg_barrier(pv, cta, e0 + 1u, status);
    const int64_t per = (n4 + R - 1) / R;
    const int64_t lo = min((int64_t)pv.rank * per, n4), hi = peers_ready ? min(lo + per, n4) : lo;   // timeout: no data phase
    const int64_t step = (int64_t)gridDim.x * kXchgThreads;
    for (int64_t i0 = lo + (int64_t)cta * kXchgThreads + threadIdx.x; i0 < hi; i0 += step * UNROLL) {
        float4 acc[UNROLL];
        if (NR > 0) {
            float4 v[UNROLL][NR > 0 ? NR : 1];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const int64_t i = i0 + u * step;
#pragma unroll
                for (int p = 0; p < NR; ++p)
                    if (i < hi) v[u][p] = ld_peer(pv.data[p] + i);
            }
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                acc[u] = v[u][0];
#pragma unroll
                for (int p = 1; p < NR; ++p) {
                    acc[u].x += v[u][p].x; acc[u].y += v[u][p].y; acc[u].z += v[u][p].z; acc[u].w += v[u][p].w;
                }
            }
        } else {
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const int64_t i = i0 + u * step;
                acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (i < hi) {
                    acc[u] = ld_peer(pv.data[0] + i);
                    for (int p = 1; p < R; ++p) {
                        const float4 v = ld_peer(pv.data[p] + i);
                        acc[u].x += v.x; acc[u].y += v.y; acc[u].z += v.z; acc[u].w += v.w;
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const int64_t i = i0 + u * step;
            if (i >= hi) continue;
            if (mask_cells) acc[u] = mask_cell(acc[u], __ldg(&reinterpret_cast<const float*>(mask_cells)[4 * i + 3]));
            if (NR > 0) {
#pragma unroll
                for (int p = 0; p < NR; ++p) st_peer(pv.data[p] + i, acc[u]);
            } else {
                for (int p = 0; p < R; ++p) st_peer(pv.data[p] + i, acc[u]);
            }
        }
    }
    xchg_barrier(pv, cta, e0 + 2u, status);
    if (threadIdx.x == 0) epoch_words[cta] = e0 + 2u;
}

// ---- the same exchange through the NVSwitch (NVLS): multimem.ld_reduce + multimem.st ----------------
// The buffers of all ranks sit in one NCCL symmetric window (ncclMemAlloc + ncclCommWindowRegister), which
// also has a MULTICAST mapping: a load-reduce from the multicast address returns the sum of the element
// over all ranks, formed inside the switch, and a store to it lands in every rank's copy.  Rank r owns
// slab r as above, but per element it issues ONE 16-byte multimem.ld_reduce (instead of R peer loads)
// and ONE multimem.st (instead of R peer stores): the link carries (R-1)/R of the buffer out and in per
// GPU as before, but each GPU only issues 2/R of the requests and the adds happen in the switch.  The
// validity mask of a deferred-mask backward is applied between the two, as in the peer kernel.
// Barrier: thread 0 adds 1 to flag `cta` of EVERY rank with one multimem.red.release.sys and waits until
// its own copy has collected nranks arrivals (the flags live in the same window, after the data).
__device__ __forceinline__ float4 multimem_ld_reduce(const float4* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(float4* mc, const float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w)
                 : "memory");
}
struct NvlsView {
    float4* mc_data;          // multicast address of the window's data
    uint32_t* mc_flags;       // multicast address of the arrival counters (one per CTA)
    uint32_t* my_flags;       // this rank's copy of them
    uint32_t* epoch_words;    // this rank only: arrivals this CTA has already waited for
    int nranks, rank;
};
__device__ __forceinline__ bool nvls_barrier(const NvlsView& nv, const int cta, const uint32_t target, int* status) {
    __shared__ int timed_out;
    __syncthreads();
    if (threadIdx.x == 0) {
        timed_out = 0;
        asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(nv.mc_flags + cta), "r"(1u) : "memory");
        const long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(nv.my_flags + cta) - target) < 0) {
            if (clock64() - t0 > kXchgTimeoutCycles) {
                *reinterpret_cast<volatile int*>(status) = 1;
                timed_out = 1;
                break;
            }
        }
    }
    __syncthreads();
    return timed_out == 0;   // thread 0 alone writes the flag, and only after the next call's first __syncthreads
}
template <int UNROLL>
__global__ void __launch_bounds__(kXchgThreads) nvls_allreduce_kernel(const NvlsView nv, const int64_t n4,
                                                                      const float4* __restrict__ mask_cells,
                                                                      int* __restrict__ status) {
    __shared__ uint32_t e_s;
    const int cta = blockIdx.x;
    if (threadIdx.x == 0) e_s = nv.epoch_words[cta];
    __syncthreads();
    const uint32_t e0 = e_s;
    const uint32_t R = (uint32_t)nv.nranks;
    const bool peers_ready = nvls_barrier(nv, cta, e0 + R, status);
    const int64_t per = (n4 + R - 1) / R;
    const int64_t lo = min((int64_t)nv.rank * per, n4), hi = peers_ready ? min(lo + per, n4) : lo;   // timeout: no data phase
    const int64_t step = (int64_t)gridDim.x * kXchgThreads;
    for (int64_t i0 = lo + (int64_t)cta * kXchgThreads + threadIdx.x; i0 < hi; i0 += step * UNROLL) {
        float4 acc[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const int64_t i = i0 + u * step;
            if (i < hi) acc[u] = multimem_ld_reduce(nv.mc_data + i);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const int64_t i = i0 + u * step;
            if (i >= hi) continue;
            if (mask_cells) acc[u] = mask_cell(acc[u], __ldg(&reinterpret_cast<const float*>(mask_cells)[4 * i + 3]));
            multimem_st(nv.mc_data + i, acc[u]);
        }
    }
    nvls_barrier(nv, cta, e0 + 2u * R, status);
    if (threadIdx.x == 0) nv.epoch_words[cta] = e0 + 2u * R;
}

}  // namespace merl
