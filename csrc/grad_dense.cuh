// Shared pieces of the dense gradient kernels (row-block kernel in grad_dense.cu, row-warp kernel
// for narrow matrices in grad_dense_small.cu).
#pragma once
#include "common.cuh"

namespace eh {

constexpr int GD_CONSUMER_WARPS = 8;
constexpr int GD_CONSUMERS = GD_CONSUMER_WARPS * 32;
constexpr int GD_THREADS = GD_CONSUMERS + 32;
constexpr int GD_MAX_STAGES = 8;
constexpr int GD_MAX_SEGS = EH_MAX_SEGS;
constexpr int GD_SMEM_BUDGET = 220 * 1024;

struct GradScratchHdr {          // first 512 bytes of the scratch buffer
    alignas(128) uint32_t barrier;     // device-wide barrier ticket
    alignas(128) uint32_t done;        // final ticket
    alignas(128) uint32_t decision;    // (seq << 2) | code, written by CTA 0
    alignas(128) uint32_t pad;
};
constexpr uint32_t DEC_GO = 1, DEC_SKIP = 2;

struct GradParams {
    const float* seg_X[EH_MAX_SEGS];
    const float* seg_y[EH_MAX_SEGS];
    const float* beta;
    float* out;
    uint32_t* flag;
    const uint32_t* beta_flag;
    uint32_t* err;
    GradScratchHdr* hdr;
    float* partials;             // [grid][ldp]
    int d, ld, ldp;
    int rows_per_seg, n_segs, model;
    int n_stages;
    int steps_per_seg;
    long long n_steps;
    uint32_t stamp, seq;
    unsigned long long timeout_ns;
    float coeff[GD_MAX_SEGS];
};

__device__ __forceinline__ float residual(int model, float p, float y) {
    if (model == EH_MODEL_LOGISTIC) {
        // -y / (exp(p*y) + 1); exp overflow -> inf -> -0, matching the float64 limit
        return __fdividef(-y, __expf(p * y) + 1.0f);
    }
    return -2.0f * (y - p);
}


// Device-wide barrier, fixed-order column-group reduction of the per-CTA partials, store into the
// destination (local or peer memory) and release of the arrival flag.  Every thread of every CTA
// calls it after this CTA's partial row has been written to p.partials[blockIdx.x].
__device__ __forceinline__ void reduce_and_ship(const GradParams& p, unsigned char* smem_raw) {
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int G = gridDim.x, b = blockIdx.x;
    uint32_t gen = 0;
    grid_barrier(&p.hdr->barrier, gen);

    if (warp < GD_CONSUMER_WARPS) {
        const int ld4 = p.ld >> 2;
        const int n_groups = (ld4 + 31) / 32;                 // 32 float4 columns per group
        float4* sred = reinterpret_cast<float4*>(smem_raw);   // [8][32] float4, ring is idle now
        for (int grp = b; grp < n_groups; grp += G) {
            const int c4 = grp * 32 + lane;
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c4 < ld4) {
                for (int row = warp; row < G; row += GD_CONSUMER_WARPS) {
                    const float4 t = __ldcg(reinterpret_cast<const float4*>(p.partials + (size_t)row * p.ldp) + c4);
                    s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
                }
            }
            sred[warp * 32 + lane] = s;
            asm volatile("bar.sync 2, %0;" ::"r"(GD_CONSUMERS) : "memory");
            if (warp == 0 && c4 < ld4) {
                float4 t = sred[lane];
#pragma unroll
                for (int w = 1; w < GD_CONSUMER_WARPS; ++w) {
                    const float4 u = sred[w * 32 + lane];
                    t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
                }
                const int c = c4 * 4;
                if (c + 3 < p.d) {
                    *reinterpret_cast<float4*>(p.out + c) = t;
                } else {
                    if (c + 0 < p.d) p.out[c + 0] = t.x;
                    if (c + 1 < p.d) p.out[c + 1] = t.y;
                    if (c + 2 < p.d) p.out[c + 2] = t.z;
                }
            }
            asm volatile("bar.sync 2, %0;" ::"r"(GD_CONSUMERS) : "memory");
        }
    }
    __syncthreads();
    if (tid == 0) {
        __threadfence_system();
        const uint32_t old = atom_add_acq_rel_gpu(&p.hdr->done, 1u);
        if (old == (uint32_t)G - 1) {
            // last CTA: every CTA is past the barrier wait -> safe to recycle the counters
            p.hdr->barrier = 0;
            p.hdr->done = 0;
            __threadfence_system();
            if (p.flag) st_release_sys(p.flag, p.stamp);
        }
    }
}

// Parameter-flag wait by CTA 0 and broadcast of the go/skip decision to the other CTAs.
// Called by one thread per CTA; returns DEC_GO or DEC_SKIP.
__device__ __forceinline__ uint32_t wait_for_parameters(const GradParams& p) {
    uint32_t code = DEC_GO;
    if (blockIdx.x == 0) {
        if (p.beta_flag) {
            const uint64_t t0 = globaltimer();
            while (true) {
                const uint32_t f = ld_acquire_sys(p.beta_flag);
                if (f == p.stamp) break;
                if (f > p.stamp) { code = DEC_SKIP; break; }                  // round already over
                if (p.err && ld_acquire_sys(p.err) != 0) { code = DEC_SKIP; break; }
                if (globaltimer() - t0 > p.timeout_ns) {
                    if (p.err) atomicCAS(p.err, 0u, (uint32_t)(-EH_ETIMEOUT));
                    code = DEC_SKIP;
                    break;
                }
                nanosleep(64);
            }
        }
        st_release_gpu(&p.hdr->decision, (p.seq << 2) | code);
    } else {
        uint32_t v;
        while (((v = ld_acquire_gpu(&p.hdr->decision)) >> 2) != p.seq) nanosleep(32);
        code = v & 3u;
    }
    return code;
}

// host-side launchers (defined next to their kernels)
int launch_rowwarp(GradParams& p, void* d_scratch, size_t scratch_bytes, cudaStream_t stream);
int launch_wide(GradParams& p, void* d_scratch, size_t scratch_bytes, cudaStream_t stream);

}  // namespace eh
