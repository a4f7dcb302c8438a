// Row-warp variant of the fused dense gradient kernel for narrow matrices (row stride <= 1024
// floats).  Same contract, scratch, barrier and ship tail as grad_dense.cu; only the consumer
// mapping differs: LPR lanes own one row (32/LPR rows per warp at a time), so the row dot is a
// shuffle reduction inside the warp and no block-wide barrier is needed per tile.  Each lane keeps
// its 4*DVW columns of the row in registers between the forward dot and the rank-1 accumulate;
// the warps' private accumulators are combined once at the end.
//
//   ld <=  32 / 64 / 128 / 256 : LPR = 8,  DVW = 1 / 2 / 4 / 8   (4 rows per warp instruction)
//   ld <=  512                 : LPR = 16, DVW = 8
//   ld <= 1024 / 2048          : LPR = 32, DVW = 8 / 16
// bf16 rows (E = 8 columns per 16-byte vector) use the same mapping on half as many vectors.
#include "grad_dense.cuh"

namespace eh {

// 16 bytes of a row as E floats: 4 fp32 values, or 8 bf16 values widened to fp32
template <int E>
__device__ __forceinline__ void unpack_vec(const uint4& q, float (&f)[E]) {
    if constexpr (E == 4) {
        f[0] = __uint_as_float(q.x); f[1] = __uint_as_float(q.y); f[2] = __uint_as_float(q.z); f[3] = __uint_as_float(q.w);
    } else {
        f[0] = __uint_as_float(q.x << 16); f[1] = __uint_as_float(q.x & 0xffff0000u);
        f[2] = __uint_as_float(q.y << 16); f[3] = __uint_as_float(q.y & 0xffff0000u);
        f[4] = __uint_as_float(q.z << 16); f[5] = __uint_as_float(q.z & 0xffff0000u);
        f[6] = __uint_as_float(q.w << 16); f[7] = __uint_as_float(q.w & 0xffff0000u);
    }
}

// E = columns per 16-byte vector (4 fp32 / 8 bf16), LPR lanes per row, DVW vectors per lane
template <int E, int LPR, int DVW>
__global__ void __launch_bounds__(GD_THREADS, 1) grad_dense_rowwarp_kernel(const GradParams p, const int RB) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int RPW = 32 / LPR;                       // rows a warp processes concurrently
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int b = blockIdx.x;
    constexpr int EB = 16 / E;                          // bytes per element
    const uint32_t stage_bytes = (uint32_t)RB * p.ld * EB;
    unsigned char* stages = smem_raw;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)p.n_stages * stage_bytes);
    uint64_t* empty = full + GD_MAX_STAGES;
    __shared__ uint32_t s_decision;
    __shared__ long long s_stop_store;
    volatile long long* s_stop = &s_stop_store;

    const JobView jv = job_of_cta(p);
    const long long my_steps = (jv.bl < jv.n_steps) ? (jv.n_steps - jv.bl + jv.Gj - 1) / jv.Gj : 0;

    if (tid == 0) {
        s_stop_store = NO_STOP;
        for (int s = 0; s < p.n_stages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], GD_CONSUMER_WARPS);
        }
        mbar_fence_init();
    }
    __syncthreads();

    uint32_t code = DEC_GO;
    if (warp == GD_CONSUMER_WARPS) {
        // ---------------- producer warp (identical protocol to grad_dense_kernel)
        long long issued = 0;
        uint64_t pol = l2_policy_evict_first();
        auto issue = [&](long long k) {
            const long long step = jv.bl + k * (long long)jv.Gj;
            const int sg = jv.seg0 + (int)(step / p.steps_per_seg);
            const int row0 = (int)(step % p.steps_per_seg) * RB;
            const int nr = (p.rows_per_seg - row0) < RB ? (p.rows_per_seg - row0) : RB;
            const uint32_t bytes = (uint32_t)nr * p.ld * EB;
            const int st = (int)(k % p.n_stages);
            mbar_arrive_expect_tx(&full[st], bytes);
            tma_load_1d(stages + (size_t)st * stage_bytes,
                        reinterpret_cast<const unsigned char*>(p.seg_X[sg]) + (size_t)row0 * p.ld * EB, bytes, &full[st], pol);
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
            return;
        }
        if (lane == 0) {
            for (long long k = issued; k < my_steps; ++k) {
                const int st = (int)(k % p.n_stages);
                const uint32_t use = (uint32_t)(k / p.n_stages);
                mbar_wait_backoff(&empty[st], (use - 1) & 1);
                if (p.abandon && (k & 7) == 0 && round_is_over(p)) {
                    producer_abandon(p, s_stop, k, &full[st]);
                    break;
                }
                issue(k);
            }
        }
        __syncwarp();
    } else {
        asm volatile("bar.sync 1, %0;" ::"r"(GD_THREADS) : "memory");
        code = s_decision;
        if (code == DEC_SKIP) return;

        // ---------------- consumers: LPR lanes per row
        const int ldv = p.ld / E;                       // 16-byte vectors per row
        const int sub = lane / LPR;                     // which of the warp's concurrent rows
        const int pos = lane % LPR;                     // lane position inside the row
        float bt[DVW][E], acc[DVW][E];
#pragma unroll
        for (int v = 0; v < DVW; ++v) {
            const int c = (v * LPR + pos) * E;
#pragma unroll
            for (int e = 0; e < E; ++e) {
                acc[v][e] = 0.f;
                bt[v][e] = (c + e < p.d) ? p.beta[c + e] : 0.f;
            }
        }
        // U rows in flight per lane group: narrow rows are latency-bound (shuffle + exp chain per row)
        constexpr int FL = DVW * E / 4;                 // floats per lane and row, in float4 units
        constexpr int U = (FL == 1) ? 8 : (FL == 2) ? 4 : (FL <= 8 ? 2 : 1);
        const int rows_per_group = GD_CONSUMER_WARPS * RPW;      // rows the CTA covers per sub-pass
        const int rows_per_pass = rows_per_group * U;
        for (long long k = 0; k < my_steps; ++k) {
            const int st = (int)(k % p.n_stages);
            const uint32_t use = (uint32_t)(k / p.n_stages);
            const long long step = jv.bl + k * (long long)jv.Gj;
            const int sg = jv.seg0 + (int)(step / p.steps_per_seg);
            const int row0 = (int)(step % p.steps_per_seg) * RB;
            const int nr = (p.rows_per_seg - row0) < RB ? (p.rows_per_seg - row0) : RB;
            const float cf = p.coeff[sg];
            const float* ys = p.seg_y[sg] + row0;
            if (p.abandon && k >= *s_stop) break;
            mbar_wait(&full[st], use & 1);
            if (p.abandon && k >= *s_stop) break;
            const uint4* tile = reinterpret_cast<const uint4*>(stages + (size_t)st * stage_bytes);
            for (int r0 = 0; r0 < nr; r0 += rows_per_pass) {
                float x[U][DVW][E];
                float dot[U], yv[U];
                bool valid[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int r = r0 + u * rows_per_group + warp * RPW + sub;
                    valid[u] = r < nr;
                    yv[u] = valid[u] ? __ldg(ys + r) : 0.f;
                    dot[u] = 0.f;
#pragma unroll
                    for (int v = 0; v < DVW; ++v) {
                        const int cv = v * LPR + pos;
                        uint4 q = make_uint4(0u, 0u, 0u, 0u);
                        if (valid[u] && cv < ldv) q = tile[(size_t)r * ldv + cv];
                        unpack_vec<E>(q, x[u][v]);
#pragma unroll
                        for (int e = 0; e < E; ++e) dot[u] = fmaf(x[u][v][e], bt[v][e], dot[u]);
                    }
                }
#pragma unroll
                for (int o = LPR / 2; o > 0; o >>= 1) {
#pragma unroll
                    for (int u = 0; u < U; ++u) dot[u] += __shfl_xor_sync(0xffffffffu, dot[u], o);
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const float rs = valid[u] ? cf * residual(p.model, dot[u], yv[u]) : 0.f;
#pragma unroll
                    for (int v = 0; v < DVW; ++v)
#pragma unroll
                        for (int e = 0; e < E; ++e) acc[v][e] = fmaf(rs, x[u][v][e], acc[v][e]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[st]);       // the shuffles consumed every lane's tile reads
        }
        // ---------------- combine: row groups inside the warp, then the 8 warps through smem
#pragma unroll
        for (int o = LPR; o < 32; o <<= 1) {
#pragma unroll
            for (int v = 0; v < DVW; ++v)
#pragma unroll
                for (int e = 0; e < E; ++e) acc[v][e] += __shfl_xor_sync(0xffffffffu, acc[v][e], o);
        }
        asm volatile("bar.sync 2, %0;" ::"r"(GD_CONSUMERS) : "memory");      // all tiles consumed: ring is free
        const int ld4 = p.ld >> 2;
        float4* sacc = reinterpret_cast<float4*>(smem_raw);                    // [8][ld4]
        if (sub == 0) {
#pragma unroll
            for (int v = 0; v < DVW; ++v) {
                const int cv = v * LPR + pos;
                if (cv < ldv) {
#pragma unroll
                    for (int h = 0; h < E / 4; ++h)
                        sacc[warp * ld4 + cv * (E / 4) + h] =
                            make_float4(acc[v][4 * h], acc[v][4 * h + 1], acc[v][4 * h + 2], acc[v][4 * h + 3]);
                }
            }
        }
        asm volatile("bar.sync 2, %0;" ::"r"(GD_CONSUMERS) : "memory");
        float4* mine = reinterpret_cast<float4*>(p.partials + (size_t)b * p.ldp);
        for (int c4 = tid; c4 < ld4; c4 += GD_CONSUMERS) {
            float4 t = sacc[c4];
#pragma unroll
            for (int w = 1; w < GD_CONSUMER_WARPS; ++w) {
                const float4 u = sacc[w * ld4 + c4];
                t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
            }
            __stcg(mine + c4, t);
        }
        asm volatile("bar.sync 2, %0;" ::"r"(GD_CONSUMERS) : "memory");      // sacc reads done before the tail reuses smem
    }
    reduce_and_ship(p, smem_raw);
}

template <int E, int LPR, int DVW>
static int launch_rw(GradParams& p, void* d_scratch, size_t scratch_bytes, cudaStream_t stream) {
    const DevInfo& di = dev_info();
    constexpr int RPW = 32 / LPR;
    const int row_bytes = p.ld * (16 / E);
    // rows per stage: ~32 KB, a multiple of the rows the CTA covers per pass
    int RB = (32 * 1024) / row_bytes;
    constexpr int FL = DVW * E / 4;
    constexpr int U = (FL == 1) ? 8 : (FL == 2) ? 4 : (FL <= 8 ? 2 : 1);
    const int pass = GD_CONSUMER_WARPS * RPW * U;
    if (RB < pass) RB = pass;
    RB = RB / pass * pass;
    if (RB > p.rows_per_seg) RB = (p.rows_per_seg + pass - 1) / pass * pass;
    const size_t stage_bytes = (size_t)RB * row_bytes;
    int stages = (int)((GD_SMEM_BUDGET - 1024) / stage_bytes);
    if (stages > GD_MAX_STAGES) stages = GD_MAX_STAGES;
    EH_REQUIRE(stages >= 2, "grad_dense: row tile too large for shared memory (ld=%d)", p.ld);
    const size_t tail = 2 * GD_MAX_STAGES * sizeof(uint64_t);
    size_t smem = stages * stage_bytes + tail;
    const size_t need_red = (size_t)GD_CONSUMER_WARPS * p.ld * sizeof(float);     // sacc + the tail's [8][32] float4
    if (smem < need_red + tail) smem = need_red + tail;
    if (smem < 8 * 32 * sizeof(float4) + tail) smem = 8 * 32 * sizeof(float4) + tail;
    p.n_stages = stages;
    p.steps_per_seg = (p.rows_per_seg + RB - 1) / RB;
    auto kern = grad_dense_rowwarp_kernel<E, LPR, DVW>;
    EH_SET_MAX_SMEM((kern), GD_SMEM_BUDGET);
    const int grid = assign_ctas(p, worker_sms());
    p.ldp = p.ld;
    const size_t need = sizeof(GradScratchHdr) + (size_t)grid * p.ldp * sizeof(float);
    EH_REQUIRE(scratch_bytes >= need, "grad_dense: scratch too small (%zu < %zu)", scratch_bytes, need);
    p.hdr = reinterpret_cast<GradScratchHdr*>(d_scratch);
    p.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(d_scratch) + sizeof(GradScratchHdr));
    int rb = RB;
    void* args[] = {(void*)&p, (void*)&rb};
    { int lrc = launch_grid_sync((const void*)kern, dim3((unsigned)grid), dim3(GD_THREADS), args, smem, stream); if (lrc) return lrc; }
    return EH_OK;
}

int launch_rowwarp(GradParams& p, void* d_scratch, size_t scratch_bytes, cudaStream_t stream) {
    // narrow rows: few lanes per row (several rows per warp instruction), wide rows: the whole warp per row
    if (p.x_bf16) {
        if (p.ld <= 64) return launch_rw<8, 8, 1>(p, d_scratch, scratch_bytes, stream);
        if (p.ld <= 128) return launch_rw<8, 8, 2>(p, d_scratch, scratch_bytes, stream);
        if (p.ld <= 256) return launch_rw<8, 8, 4>(p, d_scratch, scratch_bytes, stream);
        if (p.ld <= 512) return launch_rw<8, 16, 4>(p, d_scratch, scratch_bytes, stream);
        if (p.ld <= 1024) return launch_rw<8, 32, 4>(p, d_scratch, scratch_bytes, stream);
        return launch_rw<8, 32, 8>(p, d_scratch, scratch_bytes, stream);
    }
    if (p.ld <= 32) return launch_rw<4, 8, 1>(p, d_scratch, scratch_bytes, stream);
    if (p.ld <= 64) return launch_rw<4, 8, 2>(p, d_scratch, scratch_bytes, stream);
    if (p.ld <= 128) return launch_rw<4, 8, 4>(p, d_scratch, scratch_bytes, stream);
    if (p.ld <= 256) return launch_rw<4, 8, 8>(p, d_scratch, scratch_bytes, stream);
    if (p.ld <= 512) return launch_rw<4, 16, 8>(p, d_scratch, scratch_bytes, stream);
    if (p.ld <= 1024) return launch_rw<4, 32, 8>(p, d_scratch, scratch_bytes, stream);
    return launch_rw<4, 32, 16>(p, d_scratch, scratch_bytes, stream);
}

}  // namespace eh
