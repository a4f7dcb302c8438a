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
constexpr int GD_MAX_JOBS = 8;                 // messages batched into one launch
constexpr int GD_TOTAL_SEGS = 32;              // partitions over all batched messages
constexpr int GD_SMEM_BUDGET = 220 * 1024;

struct GradScratchHdr {          // first 512 bytes of the scratch buffer
    alignas(128) uint32_t barrier;     // device-wide barrier ticket
    alignas(128) uint32_t done;        // final ticket
    alignas(128) uint32_t decision;    // (seq << 2) | code, written by CTA 0
    alignas(128) uint32_t aborted;     // set by any CTA that abandoned the round (GradParams::abandon); reset by the last CTA
};
constexpr uint32_t DEC_GO = 1, DEC_SKIP = 2;
constexpr long long NO_STOP = 0x7fffffffffffffffll;

struct GradParams {
    const void* seg_X[GD_TOTAL_SEGS];   // fp32 rows, or bf16 rows when x_bf16
    const float* seg_y[GD_TOTAL_SEGS];
    const float* beta;
    float* out;                      // single-message paths (wide fallback)
    uint32_t* flag;
    // batched messages: job j owns CTAs [job_cta0[j], job_cta0[j+1]) and segments
    // [job_seg0[j], job_seg0[j] + job_nsegs[j]) and writes job_out[j] / releases job_flag[j]
    float* job_out[GD_MAX_JOBS];
    uint32_t* job_flag[GD_MAX_JOBS];
    int job_cta0[GD_MAX_JOBS + 1];
    int job_seg0[GD_MAX_JOBS];
    int job_nsegs[GD_MAX_JOBS];
    int n_jobs;
    int x_bf16;
    int abandon;                 // watch beta_flag while streaming and stop once the round is over
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
    float coeff[GD_TOTAL_SEGS];
};

// which batched message this CTA works for
struct JobView {
    int jb, b0, Gj, bl, seg0;
    long long n_steps;
};
__device__ __forceinline__ JobView job_of_cta(const GradParams& p) {
    JobView v;
    int jb = 0;
    while (jb + 1 < p.n_jobs && (int)blockIdx.x >= p.job_cta0[jb + 1]) ++jb;
    v.jb = jb;
    v.b0 = p.job_cta0[jb];
    v.Gj = p.job_cta0[jb + 1] - v.b0;
    v.bl = (int)blockIdx.x - v.b0;
    v.seg0 = p.job_seg0[jb];
    v.n_steps = (long long)p.steps_per_seg * p.job_nsegs[jb];
    return v;
}

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
    const int G = gridDim.x;
    const JobView jv = job_of_cta(p);
    float* out = p.job_out[jv.jb];
    uint32_t gen = 0;
    grid_barrier(&p.hdr->barrier, gen);
    // a CTA that saw the round end while streaming raised hdr->aborted before the barrier: nothing is shipped
    const bool aborted = p.abandon && ld_acquire_gpu(&p.hdr->aborted) != 0u;

    if (warp < GD_CONSUMER_WARPS && !aborted && jv.Gj <= GD_CONSUMER_WARPS) {
        // a message that runs on a handful of CTAs (small partitions): every warp reduces its own column groups --
        // the eight-warp split below would leave most warps idle and pay two block barriers per group
        const int ld4 = p.ld >> 2;
        const int n_groups = (ld4 + 31) / 32;
        for (int grp = jv.bl * GD_CONSUMER_WARPS + warp; grp < n_groups; grp += jv.Gj * GD_CONSUMER_WARPS) {
            const int c4 = grp * 32 + lane;
            if (c4 < ld4) {
                float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int row = jv.b0; row < jv.b0 + jv.Gj; ++row) {
                    const float4 u = __ldcg(reinterpret_cast<const float4*>(p.partials + (size_t)row * p.ldp) + c4);
                    t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
                }
                const int c = c4 * 4;
                if (c + 3 < p.d) {
                    *reinterpret_cast<float4*>(out + c) = t;
                } else {
                    if (c + 0 < p.d) out[c + 0] = t.x;
                    if (c + 1 < p.d) out[c + 1] = t.y;
                    if (c + 2 < p.d) out[c + 2] = t.z;
                }
            }
        }
    } else if (warp < GD_CONSUMER_WARPS && !aborted) {
        const int ld4 = p.ld >> 2;
        const int n_groups = (ld4 + 31) / 32;                 // 32 float4 columns per group
        float4* sred = reinterpret_cast<float4*>(smem_raw);   // [8][32] float4, ring is idle now
        for (int grp = jv.bl; grp < n_groups; grp += jv.Gj) {
            const int c4 = grp * 32 + lane;
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c4 < ld4) {
                for (int row = jv.b0 + warp; row < jv.b0 + jv.Gj; row += GD_CONSUMER_WARPS) {
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
                    *reinterpret_cast<float4*>(out + c) = t;
                } else {
                    if (c + 0 < p.d) out[c + 0] = t.x;
                    if (c + 1 < p.d) out[c + 1] = t.y;
                    if (c + 2 < p.d) out[c + 2] = t.z;
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
            p.hdr->aborted = 0;
            if (!aborted) {
                for (int j = 0; j < p.n_jobs; ++j)
                    if (p.job_flag[j]) st_arrival_time(p.job_flag[j]);
                __threadfence_system();
                for (int j = 0; j < p.n_jobs; ++j)
                    if (p.job_flag[j]) st_relaxed_sys(p.job_flag[j], p.stamp);      // ordered by the fence above
            }
        }
    }
}

// Parameter-flag wait by CTA 0 and broadcast of the go/skip decision to the other CTAs.
// Called by one thread per CTA; returns DEC_GO or DEC_SKIP.
template <class Params>
__device__ __forceinline__ uint32_t wait_for_parameters(const Params& p) {
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

// Abandoning a round in flight (GradParams::abandon).  The producer lane looks at the parameter flag every
// eighth tile; when the flag has passed this message's stamp the round was decided without it, so the lane
// stops issuing loads at tile k: it waits until the ring slot of tile k is free, publishes `*s_stop = k`,
// raises hdr->aborted and completes the slot's `full` barrier without data so that consumers already waiting
// for tile k wake up.  Consumers test `*s_stop` before and after every wait and leave their tile loop at k.
// All CTAs still meet in the device-wide barrier of reduce_and_ship, which then ships nothing.
__device__ __forceinline__ bool round_is_over(const GradParams& p) {
    return p.beta_flag != nullptr && ld_acquire_sys(p.beta_flag) > p.stamp;
}
__device__ __forceinline__ void producer_abandon(const GradParams& p, volatile long long* s_stop, long long k, uint64_t* full_k) {
    *s_stop = k;
    __threadfence_block();
    atomicExch(&p.hdr->aborted, 1u);
    mbar_arrive(full_k);
}

// host-side launchers (defined next to their kernels)
// CTA ranges of the batched messages, proportional to their tile counts (host)
int assign_ctas(GradParams& p, int sm_count);
int launch_rowwarp(GradParams& p, void* d_scratch, size_t scratch_bytes, cudaStream_t stream);
int launch_wide(GradParams& p, void* d_scratch, size_t scratch_bytes, cudaStream_t stream);
int launch_cluster(GradParams& p, void* d_scratch, size_t scratch_bytes, cudaStream_t stream);

}  // namespace eh
