// Fused one-pass encoded gradient for dense fp32 features (ops K1 + K2' + K3 of SURVEY 2.4).
//
//   g = sum_r  X[r,:] * coeff[seg(r)] * resid(X[r,:].beta, y[r])
//
// Reference math: naive.py:137-139 (logistic), naive.py:345-346 (linear), encode folded into the
// residual numerator as in coded.py:92-95,184.  The reference does two BLAS GEMVs (X read
// twice, float64, CPU).  Here X is read from HBM exactly once:
//
//   * a producer warp streams RB-row tiles with 1-D TMA bulk copies (cp.async.bulk -> UBLKCP)
//     into a shared-memory ring guarded by full/empty mbarriers, L2 evict-first;
//   * 8 consumer warps pull their 4*DV columns of each row into registers, form the row dots
//     (warp shuffle + one named barrier per tile), evaluate the residual and do the rank-1
//     accumulate from the same registers;
//   * per-CTA partial gradients meet in `scratch`, a device-wide barrier (cooperative launch)
//     lets 32-column-group reducers sum them in a fixed order (deterministic), and the result is
//     stored straight into the destination slot -- local or a peer GPU's NVLink-mapped memory --
//     followed by a release/system flag (the "ship" of the message, replacing MPI Isend,
//     naive.py:150).
//
// The kernel starts prefetching X while it still waits for this round's parameters
// (`d_beta_flag`), i.e. the parameter broadcast overlaps the first tiles of the next round.
#include <stdlib.h>
#include <atomic>

#include "grad_dense.cuh"

namespace eh {

// all-tcgen05 kernel for narrow bf16 rows (grad_mc_bf16.cu)
int mc_bf16_stages(int dk, int CN);
int mc_bf16_launch(const eh_mc_job* j, void* fused_scratch, size_t fused_bytes, cudaStream_t st, bool single_vector);

// 16 bytes of a row as E floats: 4 fp32 values, or 8 bf16 values widened to fp32
template <int E>
__device__ __forceinline__ void unpack16(const uint4& q, float (&f)[E]) {
    if constexpr (E == 4) {
        f[0] = __uint_as_float(q.x); f[1] = __uint_as_float(q.y); f[2] = __uint_as_float(q.z); f[3] = __uint_as_float(q.w);
    } else {
        f[0] = __uint_as_float(q.x << 16); f[1] = __uint_as_float(q.x & 0xffff0000u);
        f[2] = __uint_as_float(q.y << 16); f[3] = __uint_as_float(q.y & 0xffff0000u);
        f[4] = __uint_as_float(q.z << 16); f[5] = __uint_as_float(q.z & 0xffff0000u);
        f[6] = __uint_as_float(q.w << 16); f[7] = __uint_as_float(q.w & 0xffff0000u);
    }
}

// E = elements per 16-byte vector: 4 (fp32 rows) or 8 (bf16 rows); DV vectors per thread; RB rows per tile
template <int E, int DV, int RB>
__global__ void __launch_bounds__(GD_THREADS, 1) grad_dense_kernel(const GradParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int EB = 16 / E;                         // bytes per element
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const uint32_t row_bytes = (uint32_t)p.ld * EB;
    const uint32_t stage_bytes = (uint32_t)RB * row_bytes;
    unsigned char* stages = smem_raw;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + (size_t)p.n_stages * stage_bytes);
    uint64_t* empty = full + GD_MAX_STAGES;
    float* red = reinterpret_cast<float*>(empty + GD_MAX_STAGES);      // [2][RB][8], 16-byte aligned
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

    // ---------------- decision: wait for beta of this round (CTA 0 decides for the grid)
    uint32_t code = DEC_GO;
    if (warp == GD_CONSUMER_WARPS) {
        // producer warp: prefetch the first tiles while the parameters are still in flight
        long long issued = 0;
        uint64_t pol = l2_policy_evict_first();
        auto issue = [&](long long k) {
            const long long step = jv.bl + k * (long long)jv.Gj;
            const int sg = jv.seg0 + (int)(step / p.steps_per_seg);
            const int row0 = (int)(step % p.steps_per_seg) * RB;
            const int nr = (p.rows_per_seg - row0) < RB ? (p.rows_per_seg - row0) : RB;
            const uint32_t bytes = (uint32_t)nr * row_bytes;
            const int st = (int)(k % p.n_stages);
            mbar_arrive_expect_tx(&full[st], bytes);
            tma_load_1d(stages + (size_t)st * stage_bytes,
                        reinterpret_cast<const unsigned char*>(p.seg_X[sg]) + (size_t)row0 * row_bytes, bytes, &full[st], pol);
        };
        if (lane == 0) {
            const long long pre = my_steps < p.n_stages ? my_steps : p.n_stages;
            for (; issued < pre; ++issued) issue(issued);
            code = wait_for_parameters(p);
            s_decision = code;
        }
        __syncwarp();
        code = __shfl_sync(0xffffffffu, code, 0);
        // named barrier 1: hand the decision to the consumers
        asm volatile("bar.sync 1, %0;" ::"r"(GD_THREADS) : "memory");
        if (code == DEC_SKIP) {
            if (lane == 0)
                for (long long k = 0; k < issued; ++k) mbar_wait(&full[(int)(k % p.n_stages)], 0);
            return;
        }
        // ---------------- producer main loop
        if (lane == 0) {
            for (long long k = issued; k < my_steps; ++k) {
                const int st = (int)(k % p.n_stages);
                const uint32_t use = (uint32_t)(k / p.n_stages);      // >= 1 here
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

        // ---------------- consumers: thread owns vectors cv = v*256 + tid, i.e. columns cv*E .. cv*E+E-1
        const int ldv = p.ld / E;
        float bt[DV][E], acc[DV][E];
#pragma unroll
        for (int v = 0; v < DV; ++v) {
            const int c = (v * GD_CONSUMERS + tid) * E;
#pragma unroll
            for (int e = 0; e < E; ++e) {
                acc[v][e] = 0.f;
                bt[v][e] = (c + e < p.d) ? p.beta[c + e] : 0.f;
            }
        }
        // tile position kept incrementally (no 64-bit divisions in the loop)
        int sg = jv.seg0 + jv.bl / p.steps_per_seg;
        int sstep = jv.bl % p.steps_per_seg;
        int st = 0;
        uint32_t phase = 0;
        const bool cols_full = (ldv == DV * GD_CONSUMERS);
        for (long long k = 0; k < my_steps; ++k) {
            const int row0 = sstep * RB;
            const int nr = (p.rows_per_seg - row0) < RB ? (p.rows_per_seg - row0) : RB;
            // labels / encode weight of this tile (independent of the TMA data)
            const float cf = p.coeff[sg];
            const float* ys = p.seg_y[sg] + row0;
            float yv[RB];
#pragma unroll
            for (int r = 0; r < RB; ++r) yv[r] = (r < nr) ? __ldg(ys + r) : 0.f;
            if (p.abandon && k >= *s_stop) break;
            mbar_wait(&full[st], phase);
            if (p.abandon && k >= *s_stop) break;
            const uint4* tile = reinterpret_cast<const uint4*>(stages + (size_t)st * stage_bytes);
            float x[RB][DV][E];
            float dot[RB];
            if (nr == RB && cols_full) {
                // full tile: no per-element predicates
#pragma unroll
                for (int r = 0; r < RB; ++r) {
                    dot[r] = 0.f;
#pragma unroll
                    for (int v = 0; v < DV; ++v) {
                        unpack16<E>(tile[(size_t)r * ldv + v * GD_CONSUMERS + tid], x[r][v]);
#pragma unroll
                        for (int e = 0; e < E; ++e) dot[r] = fmaf(x[r][v][e], bt[v][e], dot[r]);
                    }
                }
            } else {
#pragma unroll
                for (int r = 0; r < RB; ++r) {
                    dot[r] = 0.f;
#pragma unroll
                    for (int v = 0; v < DV; ++v) {
                        const int cv = v * GD_CONSUMERS + tid;
                        uint4 q = make_uint4(0u, 0u, 0u, 0u);
                        if (r < nr && cv < ldv) q = tile[(size_t)r * ldv + cv];
                        unpack16<E>(q, x[r][v]);
#pragma unroll
                        for (int e = 0; e < E; ++e) dot[r] = fmaf(x[r][v][e], bt[v][e], dot[r]);
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < RB; ++r) dot[r] = warp_sum(dot[r]);
            // all lanes' tile reads are complete (their values fed the shuffles): free the slot
            float* redk = red + (k & 1) * GD_CONSUMER_WARPS * RB;             // [RB][8 warps]
            if (lane == 0) {
#pragma unroll
                for (int r = 0; r < RB; ++r) redk[r * GD_CONSUMER_WARPS + warp] = dot[r];
                mbar_arrive(&empty[st]);
            }
            asm volatile("bar.sync 2, %0;" ::"r"(GD_CONSUMERS) : "memory");
            const float4* red4 = reinterpret_cast<const float4*>(redk);
#pragma unroll
            for (int r = 0; r < RB; ++r) {
                // rows past the end of a partial tile carry x = 0, y = 0: their residual is finite and
                // multiplies zeros, so no predicate is needed here
                const float4 a = red4[r * 2], b = red4[r * 2 + 1];
                const float pr = ((a.x + a.y) + (a.z + a.w)) + ((b.x + b.y) + (b.z + b.w));
                const float rs = cf * residual(p.model, pr, yv[r]);
#pragma unroll
                for (int v = 0; v < DV; ++v)
#pragma unroll
                    for (int e = 0; e < E; ++e) acc[v][e] = fmaf(rs, x[r][v][e], acc[v][e]);
            }
            // next tile of this CTA
            if (++st == p.n_stages) { st = 0; phase ^= 1u; }
            sstep += jv.Gj;
            while (sstep >= p.steps_per_seg) { sstep -= p.steps_per_seg; ++sg; }
        }
        // ---------------- per-CTA partial -> scratch (fp32)
        float4* mine = reinterpret_cast<float4*>(p.partials + (size_t)blockIdx.x * p.ldp);
#pragma unroll
        for (int v = 0; v < DV; ++v) {
            const int cv = v * GD_CONSUMERS + tid;
            if (cv < ldv) {
#pragma unroll
                for (int h = 0; h < E / 4; ++h)
                    __stcg(mine + (size_t)cv * (E / 4) + h,
                           make_float4(acc[v][4 * h], acc[v][4 * h + 1], acc[v][4 * h + 2], acc[v][4 * h + 3]));
            }
        }
    }

    reduce_and_ship(p, smem_raw);
}

// ------------------------------------------------------------------------------------ host

// CTA ranges of the batched messages, proportional to their tile counts; every message gets at
// least one CTA and never more CTAs than tiles.  Returns the grid size.
int assign_ctas(GradParams& p, int sm_count) {
    long long total = 0;
    long long steps[GD_MAX_JOBS];
    for (int j = 0; j < p.n_jobs; ++j) {
        steps[j] = (long long)p.steps_per_seg * p.job_nsegs[j];
        total += steps[j];
    }
    long long grid = sm_count;
    if (grid > total) grid = total;
    if (grid < p.n_jobs) grid = p.n_jobs;
    // one CTA each, the rest in proportion to the tile counts (never more CTAs than tiles)
    const long long spare = grid - p.n_jobs;
    int used = 0;
    p.job_cta0[0] = 0;
    for (int j = 0; j < p.n_jobs; ++j) {
        long long g = 1 + (total > 0 ? spare * steps[j] / total : 0);
        if (g > steps[j]) g = steps[j];
        if (g < 1) g = 1;
        used += (int)g;
        p.job_cta0[j + 1] = used;
    }
    return used;
}

template <int E, int DV, int RB>
static int launch(GradParams& p, void* d_scratch, size_t scratch_bytes, cudaStream_t stream) {
    const DevInfo& di = dev_info();
    const size_t stage_bytes = (size_t)RB * p.ld * (16 / E);
    int stages = (int)((GD_SMEM_BUDGET - 1024) / stage_bytes);
    if (stages > GD_MAX_STAGES) stages = GD_MAX_STAGES;
    EH_REQUIRE(stages >= 2, "grad_dense: row tile too large for shared memory (ld=%d)", p.ld);
    const size_t tail = 2 * GD_MAX_STAGES * sizeof(uint64_t) + 2 * GD_CONSUMER_WARPS * RB * sizeof(float);
    size_t smem = stages * stage_bytes + tail;
    if (smem < 8 * 32 * sizeof(float4) + tail) smem = 8 * 32 * sizeof(float4) + tail;
    p.n_stages = stages;
    p.steps_per_seg = (p.rows_per_seg + RB - 1) / RB;
    auto kern = grad_dense_kernel<E, DV, RB>;
    EH_SET_MAX_SMEM((kern), GD_SMEM_BUDGET);
    const int grid = assign_ctas(p, worker_sms());
    p.ldp = p.ld;
    const size_t need = sizeof(GradScratchHdr) + (size_t)grid * p.ldp * sizeof(float);
    EH_REQUIRE(scratch_bytes >= need, "grad_dense: scratch too small (%zu < %zu)", scratch_bytes, need);
    p.hdr = reinterpret_cast<GradScratchHdr*>(d_scratch);
    p.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(d_scratch) + sizeof(GradScratchHdr));
    void* args[] = {(void*)&p};
    return launch_grid_sync((const void*)kern, dim3((unsigned)grid), dim3(GD_THREADS), args, smem, stream);
}

static int fill_and_launch(const eh_grad_job* jobs, int n_jobs, void* d_scratch, size_t scratch_bytes, void* stream) {
    EH_REQUIRE(jobs && n_jobs >= 1 && n_jobs <= GD_MAX_JOBS && d_scratch, "grad_dense: bad job list (1..%d messages)",
               GD_MAX_JOBS);
    const eh_grad_job* j0 = &jobs[0];
    EH_REQUIRE(j0->d_beta, "grad_dense: null parameters");
    EH_REQUIRE(j0->rows_per_seg > 0 && j0->d > 0, "grad_dense: empty problem");
    EH_REQUIRE(j0->ld >= j0->d && (j0->ld % 4) == 0, "grad_dense: ld (%d) must be >= d and a multiple of 4", j0->ld);
    EH_REQUIRE(!j0->x_bf16 || ((j0->ld % 8) == 0 && j0->ld <= 16384),
               "grad_dense: bf16 rows need ld %% 8 == 0 and ld <= 16384 (ld=%d)", j0->ld);
    GradParams p;
    memset(&p, 0, sizeof(p));
    p.x_bf16 = j0->x_bf16 ? 1 : 0;
    p.abandon = (j0->abandon && j0->d_beta_flag) ? 1 : 0;
    int nseg = 0;
    for (int q = 0; q < n_jobs; ++q) {
        const eh_grad_job* j = &jobs[q];
        EH_REQUIRE(j->d_out, "grad_dense: null destination");
        EH_REQUIRE(j->n_segs > 0 && j->n_segs <= EH_MAX_SEGS, "grad_dense: n_segs=%d out of range (1..%d)", j->n_segs,
                   EH_MAX_SEGS);
        EH_REQUIRE(j->d == j0->d && j->ld == j0->ld && j->rows_per_seg == j0->rows_per_seg && j->model == j0->model &&
                       j->d_beta == j0->d_beta && j->d_beta_flag == j0->d_beta_flag && j->stamp == j0->stamp &&
                       (j->x_bf16 != 0) == (j0->x_bf16 != 0),
                   "grad_dense: batched messages must share shape, model, parameters and stamp");
        EH_REQUIRE(nseg + j->n_segs <= GD_TOTAL_SEGS, "grad_dense: more than %d partitions in one batch", GD_TOTAL_SEGS);
        p.job_seg0[q] = nseg;
        p.job_nsegs[q] = j->n_segs;
        p.job_out[q] = j->d_out;
        p.job_flag[q] = j->d_flag;
        for (int i = 0; i < j->n_segs; ++i, ++nseg) {
            EH_REQUIRE(j->d_X[i] && j->d_y[i], "grad_dense: null segment %d", i);
            EH_REQUIRE(((uintptr_t)j->d_X[i] % 16) == 0, "grad_dense: X must be 16-byte aligned");
            p.seg_X[nseg] = j->d_X[i];
            p.seg_y[nseg] = j->d_y[i];
            p.coeff[nseg] = j->coeff[i];
        }
    }
    p.n_jobs = n_jobs;
    p.beta = j0->d_beta; p.beta_flag = j0->d_beta_flag; p.err = j0->d_err;
    p.d = j0->d; p.ld = j0->ld;
    p.rows_per_seg = j0->rows_per_seg; p.model = j0->model;
    p.stamp = j0->stamp; p.seq = next_launch_seq();
    p.timeout_ns = j0->timeout_ns ? j0->timeout_ns : 30ull * 1000000000ull;
    cudaStream_t st = (cudaStream_t)stream;
    if (p.ld > 8192 && !p.x_bf16) {
        // two-pass fallback handles one message at a time
        for (int q = 0; q < n_jobs; ++q) {
            GradParams w = p;
            for (int i = 0; i < p.job_nsegs[q]; ++i) {
                w.seg_X[i] = p.seg_X[p.job_seg0[q] + i];
                w.seg_y[i] = p.seg_y[p.job_seg0[q] + i];
                w.coeff[i] = p.coeff[p.job_seg0[q] + i];
            }
            w.n_segs = p.job_nsegs[q]; w.out = p.job_out[q]; w.flag = p.job_flag[q];
            w.seq = next_launch_seq();
            // one pass on thread-block clusters when the row can be split over 4 or 8 CTAs,
            // else (or with EH_WIDE=twopass) the two-pass fallback
            const bool cluster_ok = p.ld <= 32768 && (p.ld % 32) == 0 && getenv("EH_WIDE_TWOPASS") == nullptr;
            int rc = cluster_ok ? launch_cluster(w, d_scratch, scratch_bytes, st)
                                : launch_wide(w, d_scratch, scratch_bytes, st);
            if (rc) return rc;
        }
        return EH_OK;
    }
    if (p.x_bf16 && n_jobs == 1 && p.ld >= 96 && p.ld <= 640 && !(p.ld > 384 && p.ld <= 512) &&
        p.ld >= (p.d + 31) / 32 * 32 &&
        mc_bf16_stages((p.d + 31) / 32 * 32, 16) > 0 && getenv("EH_BF16_CUDA_CORES") == nullptr) {
        // narrow bf16 rows: both matrix-vector products on tcgen05 from one TMA tile (grad_mc_bf16.cu; beta and
        // the residual enter as hi + lo bf16 pairs): 0.96 / 0.92 / 0.74 of HBM at 128 / 256 / 640 columns against
        // 0.75 / 0.81 / 0.52 for the CUDA-core kernels (which keep 385..512 columns: 0.80 vs 0.77)
        eh_mc_job m;
        memset(&m, 0, sizeof(m));
        for (int i = 0; i < j0->n_segs; ++i) {
            m.d_X[i] = (const float*)j0->d_X[i]; m.d_y[i] = j0->d_y[i]; m.coeff[i] = j0->coeff[i];
        }
        m.d_W = j0->d_beta; m.d_out = j0->d_out; m.d_flag = j0->d_flag; m.d_beta_flag = j0->d_beta_flag; m.d_err = j0->d_err;
        m.n_segs = j0->n_segs; m.rows_per_seg = j0->rows_per_seg; m.d = j0->d; m.ld = j0->ld;
        m.ldw = (j0->d + 31) / 32 * 32; m.C = 1; m.model = (j0->model == EH_MODEL_LOGISTIC) ? 3 : 4;
        m.stamp = j0->stamp; m.timeout_ns = j0->timeout_ns; m.x_bf16 = 1;
        return mc_bf16_launch(&m, d_scratch, scratch_bytes, st, true);
    }
    if (p.x_bf16) {
        // bf16 rows: 8 columns per 16-byte vector, 256 vectors per row pass
        // row-warp kernel up to 1024 columns (0.74-0.81 of HBM); at 2048 the row-block kernel is ahead (0.72 vs 0.63)
        if (p.ld <= 1024) return launch_rowwarp(p, d_scratch, scratch_bytes, st);
        if (p.ld <= 2048) return launch<8, 1, 8>(p, d_scratch, scratch_bytes, st);
        if (p.ld <= 4096) return launch<8, 2, 4>(p, d_scratch, scratch_bytes, st);
        if (p.ld <= 8192) return launch<8, 4, 2>(p, d_scratch, scratch_bytes, st);
        return launch<8, 8, 1>(p, d_scratch, scratch_bytes, st);
    }
    if (p.ld <= 1024 || (p.ld <= 2048 && getenv("EH_ROWBLOCK_2048") == nullptr))
        return launch_rowwarp(p, d_scratch, scratch_bytes, st);
    if (p.ld <= 2048) return launch<4, 2, 8>(p, d_scratch, scratch_bytes, st);
    if (p.ld <= 4096) return launch<4, 4, 4>(p, d_scratch, scratch_bytes, st);
    return launch<4, 8, 2>(p, d_scratch, scratch_bytes, st);
}

}  // namespace eh

extern "C" size_t eh_grad_dense_scratch_bytes(int d) {
    const int ld = (d + 3) & ~3;
    int sm = eh::dev_info().sm_count;
    if (sm <= 0) sm = 160;
    size_t b = sizeof(eh::GradScratchHdr) + (size_t)sm * (size_t)(ld + 32) * sizeof(float);
    const size_t wide = 1024 + ((size_t)65536 + ld) * sizeof(float);      // two-pass fallback (ld > 8192)
    return b > wide ? b : wide;
}

extern "C" int eh_grad_dense(const eh_grad_job* j, void* d_scratch, size_t scratch_bytes, void* stream) {
    return eh::fill_and_launch(j, 1, d_scratch, scratch_bytes, stream);
}

extern "C" int eh_grad_dense_multi(const eh_grad_job* jobs, int n_jobs, void* d_scratch, size_t scratch_bytes,
                                   void* stream) {
    return eh::fill_and_launch(jobs, n_jobs, d_scratch, scratch_bytes, stream);
}

// ------------------------------------------------------------------------------------------
// eh_grad_dense_noalloc: the same fused gradient kernel for callers that cannot allocate.
//
// Who that is: a host that drives this library from a fixed-latency loop where cudaMalloc (and
// the device-wide synchronisation it implies) is forbidden after start-up; a process whose HBM
// is already fully committed to resident partitions so that a late allocation may simply fail;
// or a plain C caller that was handed device pointers for inputs and outputs but owns no
// allocator of its own.  Such callers still pass inputs and outputs; only the scratch (the
// per-CTA partial gradients and the three barrier words) is provided here, in a file-scope
// device array that every GPU loading this library holds until the process exits.
//
// Limits: row stride ld <= EH_NOALLOC_MAX_LD floats and at most EH_NOALLOC_MAX_CTAS
// cooperative CTAs (one per SM), i.e. 512 B + 160 * 4096 * 4 B = 2.5 MiB per GPU.  Larger
// problems are rejected with EH_EINVAL and a message; use eh_grad_dense with your own scratch.
//
// Set-up: the first call on each device resolves the array's device address with
// cudaGetSymbolAddress (in host code the array's name is only a host-side stand-in, not a device
// address) and caches it per device.  Make that first call before any stream capture begins, and
// make it at start-up, while device memory is still available: it is also the moment the CUDA
// runtime loads this library's module onto that device (which is when the driver places the
// 2.5 MiB array and the kernel code) and when the kernel's shared-memory attribute is set, so a
// process that has already committed every byte of HBM can fail right there -- never later.
// Every call shares the one array, which still holds the previous call's partials: the barrier
// words that must start at zero are cleared on the caller's stream before each launch, the
// partials are fully overwritten before they are read.  Two calls that overlap on the GPU
// (different streams, or different host threads on the same device) race on the array -- keep
// calls on one device stream-ordered.
// ------------------------------------------------------------------------------------------
#define EH_NOALLOC_MAX_LD   4096
#define EH_NOALLOC_MAX_CTAS 160
#define EH_NOALLOC_BYTES    (sizeof(eh::GradScratchHdr) + (size_t)EH_NOALLOC_MAX_CTAS * EH_NOALLOC_MAX_LD * sizeof(float))

static __device__ __align__(256) unsigned char g_noalloc_scratch[EH_NOALLOC_BYTES];

extern "C" size_t eh_grad_dense_noalloc_max_ld(void) { return EH_NOALLOC_MAX_LD; }

extern "C" int eh_grad_dense_noalloc(const eh_grad_job* j, void* stream) {
    using namespace eh;
    EH_REQUIRE(j, "grad_dense_noalloc: null job");
    EH_REQUIRE(j->ld <= EH_NOALLOC_MAX_LD,
               "grad_dense_noalloc: row stride ld=%d exceeds the built-in scratch limit of %d floats; "
               "call eh_grad_dense with caller-provided scratch instead", j->ld, EH_NOALLOC_MAX_LD);
    const DevInfo& di = dev_info();
    EH_REQUIRE(di.device >= 0 && di.device < 64, "grad_dense_noalloc: no current CUDA device");
    EH_REQUIRE(di.sm_count <= EH_NOALLOC_MAX_CTAS,
               "grad_dense_noalloc: device has %d SMs, built-in scratch is sized for %d CTAs", di.sm_count,
               EH_NOALLOC_MAX_CTAS);
    // once per device, with that device current: device address of the file-scope array
    // (atomic: two host threads on different devices may both be in their first call; the value written is
    // the same for a given device, so a lost race only repeats the lookup)
    static std::atomic<void*> dev_ptr[64];
    void* scratch = dev_ptr[di.device].load(std::memory_order_acquire);
    if (scratch == nullptr) {
        void* p = nullptr;
        EH_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        EH_REQUIRE(p != nullptr, "grad_dense_noalloc: cudaGetSymbolAddress returned a null address");
        dev_ptr[di.device].store(p, std::memory_order_release);
        scratch = p;
    }
    // the array keeps the last call's contents: zero the barrier/decision words, stream-ordered
    EH_CUDA(cudaMemsetAsync(scratch, 0, sizeof(GradScratchHdr), (cudaStream_t)stream));
    return eh_grad_dense(j, scratch, EH_NOALLOC_BYTES, stream);
}
