// One-pass fused gradient for WIDE dense matrices (8192 < row stride <= 32768 floats) on
// thread-block clusters.
//
// A row no longer fits one CTA's shared-memory ring, so a cluster of CS CTAs (4 or 8) shares
// every row tile: CTA `rank` owns columns [rank*ldc, (rank+1)*ldc) of the same RB rows, TMA-streams
// only that slice, forms its partial row dots, and the CS partial dots of a row meet through
// distributed shared memory: each CTA sends its RB partials into every peer's `cdots` array with
// st.async, which completes transaction bytes on the peer's mbarrier (no release fence on the
// sender -- a release.cluster arrive per tile cost ~1.8 us); two tiles are in flight so the exchange
// of one hides behind the loads and FMAs of the next.  With the full dot in hand every CTA evaluates the residual and
// does the rank-1 accumulate for its column slice from registers -- X is still read from HBM
// exactly once (the two-pass fallback in grad_dense_wide.cu reads it twice).
//
// Same contract as eh_grad_dense: waits for the parameter flag, reduces the per-cluster partials
// in a fixed order, stores into the (possibly peer-mapped) destination and releases the flag.
// Reference math: naive.py:137-139 / 345-346, coded.py:184.
#include <stdlib.h>

#include "grad_dense.cuh"
#include "cluster_common.cuh"

namespace eh {

constexpr int GC_MAX_CS = 8;

template <int V>
struct SlotTag {
    static constexpr int value = V;
};

// DV float4 per thread over the CTA's column slice, RB rows per tile, CS CTAs per cluster (runtime)
template <int DV, int RB>
__global__ void __launch_bounds__(GD_THREADS, 1) grad_dense_cluster_kernel(const GradParams p, const int CS, const int ldc) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int rank = (int)cluster_ctarank();
    const int cluster_id = blockIdx.x / CS;
    const int n_clusters = gridDim.x / CS;
    const uint32_t row_bytes = (uint32_t)ldc * 4u;                 // this CTA's slice of one row
    const uint32_t stage_bytes = (uint32_t)RB * row_bytes;
    float* stages = reinterpret_cast<float*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)p.n_stages * stage_bytes);
    uint64_t* empty = full + GD_MAX_STAGES;
    uint64_t* cbar = empty + GD_MAX_STAGES;                        // [4] cluster exchange barriers
    float* red = reinterpret_cast<float*>(cbar + 4);               // [4][8][RB] block partials
    float* cdots = red + 4 * GD_CONSUMER_WARPS * RB;               // [4][GC_MAX_CS][RB] partial dots of all ranks
    __shared__ uint32_t s_decision;

    const long long n_steps = (long long)p.steps_per_seg * p.n_segs;
    const long long my_steps = (cluster_id < n_steps) ? (n_steps - cluster_id + n_clusters - 1) / n_clusters : 0;
    const int col0 = rank * ldc;                                   // first column of this CTA's slice

    if (tid == 0) {
        for (int s = 0; s < p.n_stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], GD_CONSUMER_WARPS);
        }
        for (int q = 0; q < 4; ++q) mbar_init(&cbar[q], 1);
        mbar_fence_init();
    }
    __syncthreads();
    cluster_sync_all();            // every CTA's barriers exist before any peer arrives on them

    uint32_t code = DEC_GO;
    if (warp == GD_CONSUMER_WARPS) {
        // ---------------- producer warp: RB row-slice copies per tile
        long long issued = 0;
        uint64_t pol = l2_policy_evict_first();
        auto issue = [&](long long k) {
            const long long step = cluster_id + k * (long long)n_clusters;
            const int sg = (int)(step / p.steps_per_seg);
            const int row0 = (int)(step % p.steps_per_seg) * RB;
            const int nr = (p.rows_per_seg - row0) < RB ? (p.rows_per_seg - row0) : RB;
            const int st = (int)(k % p.n_stages);
            mbar_arrive_expect_tx(&full[st], (uint32_t)nr * row_bytes);
            const float* src = reinterpret_cast<const float*>(p.seg_X[sg]) + (size_t)row0 * p.ld + col0;
            for (int r = 0; r < nr; ++r)
                tma_load_1d(stages + (size_t)st * RB * ldc + (size_t)r * ldc, src + (size_t)r * p.ld, row_bytes, &full[st], pol);
        };
        if (lane == 0) {
            const long long pre = my_steps < p.n_stages ? my_steps : p.n_stages;
            for (; issued < pre; ++issued) issue(issued);
            code = wait_for_parameters(p);
            s_decision = code;
        }
        __syncwarp();
        code = __shfl_sync(0xffffffffu, code, 0);
        asm volatile("bar.sync 1, %0;" ::"r"(GD_THREADS) : "memory");
        if (code == DEC_SKIP) {
            if (lane == 0)
                for (long long k = 0; k < issued; ++k) mbar_wait(&full[(int)(k % p.n_stages)], 0);
        } else if (lane == 0) {
            for (long long k = issued; k < my_steps; ++k) {
                const int st = (int)(k % p.n_stages);
                const uint32_t use = (uint32_t)(k / p.n_stages);
                mbar_wait(&empty[st], (use - 1) & 1);
                issue(k);
            }
        }
        __syncwarp();
    } else {
        asm volatile("bar.sync 1, %0;" ::"r"(GD_THREADS) : "memory");
        code = s_decision;
        if (code != DEC_SKIP) {
            // ---------------- consumers
            const int ldc4 = ldc >> 2;
            float4 bt[DV], acc[DV];
#pragma unroll
            for (int v = 0; v < DV; ++v) {
                const int c = col0 + (v * GD_CONSUMERS + tid) * 4;
                acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
                float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
                if (v * GD_CONSUMERS + tid < ldc4) {
                    if (c + 0 < p.d) t.x = p.beta[c + 0];
                    if (c + 1 < p.d) t.y = p.beta[c + 1];
                    if (c + 2 < p.d) t.z = p.beta[c + 2];
                    if (c + 3 < p.d) t.w = p.beta[c + 3];
                }
                bt[v] = t;
            }
            const uint32_t cdots_addr = smem_u32(cdots);
            const uint32_t cbar_addr = smem_u32(cbar);
            // Two tiles in flight: tile k+1 is loaded, dotted and its partial sums sent to the peers
            // BEFORE the cluster-wide sums of tile k are awaited, so the DSMEM round trip of one tile
            // hides behind the shared-memory/FMA work of the next.  Four exchange buffers (k & 3).
            float4 x[2][RB][DV];
            float yv[2][RB], cf[2];
            int nrs[2];
            auto load_send = [&](long long k, auto slot_c) {
                constexpr int SL = decltype(slot_c)::value;
                const int st = (int)(k % p.n_stages);
                const uint32_t use = (uint32_t)(k / p.n_stages);
                const int buf = (int)(k & 3);
                const long long step = cluster_id + k * (long long)n_clusters;
                const int sg = (int)(step / p.steps_per_seg);
                const int row0 = (int)(step % p.steps_per_seg) * RB;
                const int nr = (p.rows_per_seg - row0) < RB ? (p.rows_per_seg - row0) : RB;
                nrs[SL] = nr;
                cf[SL] = p.coeff[sg];
                const float* ys = p.seg_y[sg] + row0;
#pragma unroll
                for (int r = 0; r < RB; ++r) yv[SL][r] = (r < nr) ? __ldg(ys + r) : 0.f;
                mbar_wait(&full[st], use & 1);
                const float4* tile = reinterpret_cast<const float4*>(stages + (size_t)st * RB * ldc);
                float dot[RB];
#pragma unroll
                for (int r = 0; r < RB; ++r) {
                    dot[r] = 0.f;
#pragma unroll
                    for (int v = 0; v < DV; ++v) {
                        const int c4 = v * GD_CONSUMERS + tid;
                        x[SL][r][v] = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (r < nr && c4 < ldc4) x[SL][r][v] = tile[(size_t)r * ldc4 + c4];
                        dot[r] = fmaf(x[SL][r][v].x, bt[v].x, dot[r]);
                        dot[r] = fmaf(x[SL][r][v].y, bt[v].y, dot[r]);
                        dot[r] = fmaf(x[SL][r][v].z, bt[v].z, dot[r]);
                        dot[r] = fmaf(x[SL][r][v].w, bt[v].w, dot[r]);
                    }
                }
#pragma unroll
                for (int r = 0; r < RB; ++r) dot[r] = warp_sum(dot[r]);
                float* redw = red + (buf * GD_CONSUMER_WARPS + warp) * RB;
                if (lane == 0) {
#pragma unroll
                    for (int r = 0; r < RB; ++r) redw[r] = dot[r];
                    mbar_arrive(&empty[st]);
                }
                asm volatile("bar.sync 2, %0;" ::"r"(GD_CONSUMERS) : "memory");
                // thread 0 arms this CTA's exchange barrier for the CS*RB floats it is about to receive;
                // thread t < CS sends this CTA's RB block sums to CTA t (st.async completes the bytes there)
                if (tid == 0) mbar_arrive_expect_tx(&cbar[buf], (uint32_t)(CS * RB * 4));
                if (tid < CS) {
                    const float* redk = red + buf * GD_CONSUMER_WARPS * RB;
                    const uint32_t dst = mapa_shared(cdots_addr + (uint32_t)((buf * GC_MAX_CS + rank) * RB) * 4u, (uint32_t)tid);
                    const uint32_t rbar = mapa_shared(cbar_addr + (uint32_t)buf * 8u, (uint32_t)tid);
#pragma unroll
                    for (int r = 0; r < RB; ++r) {
                        float sum = 0.f;
#pragma unroll
                        for (int w = 0; w < GD_CONSUMER_WARPS; ++w) sum += redk[w * RB + r];
                        st_async_f32(dst + r * 4u, sum, rbar);
                    }
                }
            };
            auto wait_update = [&](long long k, auto slot_c) {
                constexpr int SL = decltype(slot_c)::value;
                const int buf = (int)(k & 3);
                mbar_wait(&cbar[buf], (uint32_t)((k >> 2) & 1));
                const float* cd = cdots + buf * GC_MAX_CS * RB;
#pragma unroll
                for (int r = 0; r < RB; ++r) {
                    if (r < nrs[SL]) {
                        float pr = 0.f;
                        for (int q = 0; q < CS; ++q) pr += cd[q * RB + r];
                        const float rs = cf[SL] * residual(p.model, pr, yv[SL][r]);
#pragma unroll
                        for (int v = 0; v < DV; ++v) {
                            acc[v].x = fmaf(rs, x[SL][r][v].x, acc[v].x);
                            acc[v].y = fmaf(rs, x[SL][r][v].y, acc[v].y);
                            acc[v].z = fmaf(rs, x[SL][r][v].z, acc[v].z);
                            acc[v].w = fmaf(rs, x[SL][r][v].w, acc[v].w);
                        }
                    }
                }
            };
            if (my_steps > 0) load_send(0, SlotTag<0>{});
            for (long long k = 0; k < my_steps; k += 2) {
                if (k + 1 < my_steps) load_send(k + 1, SlotTag<1>{});
                wait_update(k, SlotTag<0>{});
                if (k + 2 < my_steps) load_send(k + 2, SlotTag<0>{});
                if (k + 1 < my_steps) wait_update(k + 1, SlotTag<1>{});
            }
            // ---------------- per-cluster partial row: this CTA fills its column slice
            float4* mine = reinterpret_cast<float4*>(p.partials + (size_t)cluster_id * p.ldp + col0);
#pragma unroll
            for (int v = 0; v < DV; ++v) {
                const int c4 = v * GD_CONSUMERS + tid;
                if (c4 < ldc4) __stcg(mine + c4, acc[v]);
            }
        }
    }
    // no CTA may leave while a peer can still address its shared memory
    cluster_sync_all();
    if (s_decision == DEC_SKIP) return;

    // ---------------- device-wide barrier, ordered reduction over the clusters, ship
    const int G = gridDim.x, b = blockIdx.x;
    uint32_t gen = 0;
    grid_barrier(&p.hdr->barrier, gen);
    if (warp < GD_CONSUMER_WARPS) {
        const int ld4 = p.ld >> 2;
        const int n_groups = (ld4 + 31) / 32;
        float4* sred = reinterpret_cast<float4*>(smem_raw);
        for (int grp = b; grp < n_groups; grp += G) {
            const int c4 = grp * 32 + lane;
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c4 < ld4) {
                for (int row = warp; row < n_clusters; row += GD_CONSUMER_WARPS) {
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
            p.hdr->barrier = 0;
            p.hdr->done = 0;
            if (p.flag) st_arrival_time(p.flag);
            __threadfence_system();
            if (p.flag) st_relaxed_sys(p.flag, p.stamp);
        }
    }
}

}  // namespace eh

namespace eh {

// Host: pick the cluster size, ask the runtime how many clusters are co-resident, launch.
template <int DV, int RB>
static int launch_cluster_t(GradParams& p, void* d_scratch, size_t scratch_bytes, cudaStream_t stream, int cs_pref) {
    int CS = 0;
    for (int c : {cs_pref, 4, 8}) {
        if (p.ld % (4 * c) == 0 && p.ld / c <= 4096) { CS = c; break; }
    }
    EH_REQUIRE(CS > 0, "grad_dense(cluster): ld=%d cannot be split over a cluster of 4 or 8 CTAs", p.ld);
    const int ldc = p.ld / CS;
    const size_t stage_bytes = (size_t)RB * ldc * 4;
    int stages = (int)((GD_SMEM_BUDGET - 2048) / stage_bytes);
    if (stages > GD_MAX_STAGES) stages = GD_MAX_STAGES;
    EH_REQUIRE(stages >= 2, "grad_dense(cluster): tile too large");
    const size_t tail = (2 * GD_MAX_STAGES + 4) * sizeof(uint64_t) + 4 * GD_CONSUMER_WARPS * RB * sizeof(float) +
                        4 * GC_MAX_CS * RB * sizeof(float);
    size_t smem = stages * stage_bytes + tail;
    if (smem < 8 * 32 * sizeof(float4) + tail) smem = 8 * 32 * sizeof(float4) + tail;
    p.n_stages = stages;
    p.steps_per_seg = (p.rows_per_seg + RB - 1) / RB;
    const long long n_steps = (long long)p.steps_per_seg * p.n_segs;
    auto kern = grad_dense_cluster_kernel<DV, RB>;
    EH_SET_MAX_SMEM((kern), GD_SMEM_BUDGET);
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cudaLaunchAttribute attrs[1];
    attrs[0].id = cudaLaunchAttributeClusterDimension;
    attrs[0].val.clusterDim.x = CS;
    attrs[0].val.clusterDim.y = 1;
    attrs[0].val.clusterDim.z = 1;
    cfg.attrs = attrs;
    cfg.numAttrs = 1;
    cfg.blockDim = dim3(GD_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cfg.gridDim = dim3(worker_sms() / CS * CS);
    int max_clusters = 0;
    EH_CUDA(cudaOccupancyMaxActiveClusters(&max_clusters, kern, &cfg));
    EH_REQUIRE(max_clusters >= 1, "grad_dense(cluster): no cluster of %d CTAs fits on this device", CS);
    long long n_clusters = max_clusters;
    if (n_clusters > n_steps) n_clusters = n_steps;
    cfg.gridDim = dim3((unsigned)(n_clusters * CS));
    p.ldp = p.ld;
    const size_t need = sizeof(GradScratchHdr) + (size_t)n_clusters * p.ldp * sizeof(float);
    EH_REQUIRE(scratch_bytes >= need, "grad_dense(cluster): scratch too small (%zu < %zu)", scratch_bytes, need);
    p.hdr = reinterpret_cast<GradScratchHdr*>(d_scratch);
    p.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(d_scratch) + sizeof(GradScratchHdr));
    int cs = CS, ldc_ = ldc;
    EH_CUDA(cudaLaunchKernelEx(&cfg, kern, p, cs, ldc_));
    return EH_OK;
}

int launch_cluster(GradParams& p, void* d_scratch, size_t scratch_bytes, cudaStream_t stream) {
    const char* e = getenv("EH_CLUSTER_RB");
    const char* c = getenv("EH_CLUSTER_CS");
    const int rb = e ? atoi(e) : 2;
    const int cs = c ? atoi(c) : 4;
    if (rb == 2) return launch_cluster_t<4, 2>(p, d_scratch, scratch_bytes, stream, cs);
    if (rb == 1) return launch_cluster_t<4, 1>(p, d_scratch, scratch_bytes, stream, cs);
    return launch_cluster_t<4, 4>(p, d_scratch, scratch_bytes, stream, cs);
}

}  // namespace eh
