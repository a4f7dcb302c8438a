// Shared device/host helpers for liberasurehead (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../include/erasurehead.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "liberasurehead targets sm_100a (Blackwell) only"
#endif

namespace eh {

// ---------------------------------------------------------------- host error plumbing
void set_error(const char* fmt, ...);
#define EH_CUDA(call)                                                                        \
    do {                                                                                     \
        cudaError_t e__ = (call);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            eh::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            return (int)e__;                                                                 \
        }                                                                                    \
    } while (0)
#define EH_REQUIRE(cond, ...)                                                                \
    do {                                                                                     \
        if (!(cond)) {                                                                       \
            eh::set_error(__VA_ARGS__);                                                      \
            return EH_EINVAL;                                                                \
        }                                                                                    \
    } while (0)

struct DevInfo {
    int device = -1;
    int sm_count = 0;
    int coop = 0;
};
const DevInfo& dev_info();
// SMs that the gradient (worker) kernels may fill on the current device: all of them, minus the SMs
// reserved for a decoder that runs concurrently on its own stream (eh_set_reserved_sms).
int worker_sms();
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) applies per device: remember it per device
#define EH_SET_MAX_SMEM(kern, bytes)                                                              \
    do {                                                                                          \
        static bool done__[64] = {false};                                                         \
        const int dv__ = eh::dev_info().device;                                                   \
        if (dv__ < 0 || dv__ >= 64 || !done__[dv__]) {                                            \
            EH_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
            if (dv__ >= 0 && dv__ < 64) done__[dv__] = true;                                      \
        }                                                                                         \
    } while (0)
// Process-wide launch sequence number for the decision word of the cooperative gradient kernels.
uint32_t next_launch_seq();
// Launch a kernel whose CTAs wait on one another (device-wide barrier).  Cooperative by default so
// co-residency is guaranteed; EH_LAUNCH=plain uses an ordinary launch (grid <= resident capacity is
// still enforced by the callers; safe while every kernel of the process runs in one stream).
int launch_grid_sync(const void* kern, dim3 grid, dim3 block, void** args, size_t smem, cudaStream_t stream);

// ---------------------------------------------------------------- workspace control block
// Lives at lay.off_ctl in every GPU's workspace.  128-byte separated words.
struct alignas(128) CtlWord {
    uint32_t v;
    uint32_t pad[31];
};
enum CtlIndex {
    CTL_BETA_FLAG = 0,   // round+1 once beta of that round is readable in beta[round&1]
    CTL_ERROR     = 1,   // sticky error code (watchdogs)
    CTL_DECISION  = 2,   // worker kernel: (stamp<<1 | stale) decision broadcast by CTA 0
    CTL_PING      = 3,   // clock alignment: sequence number stored by the root into a peer's workspace
    CTL_PONG      = 4,   // clock alignment: peer's answer in the root's workspace (v = seq, bytes 8..15 = peer clock)
    CTL_COUNT     = 8,
};

#ifdef __CUDACC__
// ---------------------------------------------------------------- device primitives
__device__ __forceinline__ uint64_t globaltimer() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// system-scope store without its own fence: use after ONE __threadfence_system() to publish several flags
// (fence + relaxed stores form the release pattern; a release store per flag would fence every time)
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// Arrival record of a message: a 16-byte (in the workspace: 128-byte) record whose word 0 is the round stamp
// and whose bytes 8..15 hold the publishing GPU's %globaltimer at the moment the message was complete
// -- the device-side analogue of the instant before the reference worker's sleep-then-Isend
// (approximate_coding.py:198-207).  Store the time BEFORE the fence / release that publishes the stamp.
__device__ __forceinline__ void st_arrival_time(uint32_t* flag) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(flag + 2), "l"(globaltimer()) : "memory");
}
__device__ __forceinline__ uint64_t ld_arrival_time(const uint32_t* flag) {
    uint64_t t;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(t) : "l"(flag + 2) : "memory");
    return t;
}
// one store, delivered by the NVSwitch to every GPU bound to the multicast object (NVLS)
__device__ __forceinline__ void multimem_st_v4(float* mc_addr, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(v.x), "f"(v.y),
                 "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t atom_add_acq_rel_gpu(uint32_t* p, uint32_t v) {
    uint32_t old;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ void nanosleep(uint32_t ns) {
    asm volatile("nanosleep.u32 %0;" ::"r"(ns));
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Device-wide barrier for co-resident grids (cooperative launch).  `counter` is a monotonically
// increasing ticket word; `generation` counts barriers already passed by this CTA.
// Every thread of the CTA must call it.  Data written before is visible after (gpu scope).
__device__ __forceinline__ void grid_barrier(uint32_t* counter, uint32_t& generation) {
    __syncthreads();
    generation += 1;
    if (threadIdx.x == 0) {
        __threadfence();
        atom_add_acq_rel_gpu(counter, 1u);
        const uint32_t target = generation * gridDim.x;
        while (ld_acquire_gpu(counter) < target) nanosleep(32);
        __threadfence();
    }
    __syncthreads();
}

// ---- mbarrier / TMA bulk copy (sm_90+ PTX; SASS: SYNCS / UBLKCP)
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// producer-side wait (ring slot free again): back off between polls so the spinning lane does
// not take issue slots from the consumer warps that share its scheduler
__device__ __forceinline__ void mbar_wait_backoff(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) __nanosleep(32);
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// 1-D TMA bulk copy global -> shared, completion on an mbarrier (bytes % 16 == 0).
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes,
                                            uint64_t* bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
        "[%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
#endif  // __CUDACC__

}  // namespace eh
