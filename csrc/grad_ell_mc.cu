// Multi-column (softmax) gradient for one-hot ELL features -- "Covertype-shaped multi-class" at
// the one-hot shape of the reference's Covertype pipeline (arrange_real_data.py:157-178: 54
// categorical columns + bias -> 15 509 one-hot features; the reference binarises the labels,
// :154-156, so the C-class model itself has no reference counterpart; BASELINE.json config 5).
//
//   logits_r[c] = sum_k W[idx[r,k]][c]                               (gather of C-vectors)
//   G[idx[r,k]][:] += coeff[seg(r)] * (softmax(logits_r) - onehot(y_r))   (scatter of C-vectors)
//
// Parameter / gradient layout: COLUMN-major W[d][8] (the C <= 8 class weights of a feature are 32
// contiguous bytes = one L2 sector), one flat vector of 8 d floats for the decoder.  One warp owns
// a row: lanes are the row's ones (two passes for up to 64 per row), every lane gathers two
// float4 per one, the eight logits are summed over the warp with a 9-shuffle exchange (halves
// trade 4 classes, quarters 2, eighths 1, then two butterflies) so that lane group c = lane / 4
// holds class c, softmax runs across the groups, and the 8 residuals are broadcast back.  The
// scatter is `red.global.add.v4.f32` (two 16-byte L2 reductions per one); a lane keeps the last
// feature it touched and its accumulated residual vector in registers and only flushes when
// the feature changes, which removes the always-on bias feature (every row) from the atomic
// traffic altogether and shortens the runs on the popular levels; single hits go through a
// per-CTA two-way table of fixed-point accumulators in shared memory (see HotTable below) so that
// the popular levels do not serialise as same-address L2 reductions.  The accumulation vector
// lives in scratch (one per batched message), is zeroed, filled and shipped -- to local or peer
// memory, followed by the arrival record -- by this one cooperative kernel.
#include <stdlib.h>

#include "common.cuh"

namespace eh {

constexpr int GEM_THREADS = 512;
constexpr int GEM_CP = 8;                       // padded classes per feature
constexpr int GEM_MAX_SEGS = 64;
constexpr int GEM_MAX_JOBS = 16;
constexpr uint32_t GEM_GO = 1, GEM_SKIP = 2;

struct EllMcHdr {
    alignas(128) uint32_t barrier;
    alignas(128) uint32_t done;
    alignas(128) uint32_t decision;
    alignas(128) uint32_t pad;
};

struct EllMcParams;
__device__ __forceinline__ bool getenv_table_on(const EllMcParams& p);
struct EllMcParams {
    int hot_table;                   // 0 switches the per-CTA table off (EH_ELL_MC_TABLE=0, for A/B measurements)
    const int32_t* seg_idx[GEM_MAX_SEGS];
    const float* seg_y[GEM_MAX_SEGS];
    float coeff[GEM_MAX_SEGS];
    const float* W;                  // [d][8]
    const uint32_t* beta_flag;
    uint32_t* err;
    EllMcHdr* hdr;
    float* gacc;                     // [n_jobs][d][8]
    int d, nnz, rows_per_seg, C, model;
    uint32_t stamp, seq;
    unsigned long long timeout_ns;
    int n_jobs;
    int job_cta0[GEM_MAX_JOBS + 1], job_seg0[GEM_MAX_JOBS], job_nsegs[GEM_MAX_JOBS];
    float* job_out[GEM_MAX_JOBS];
    uint32_t* job_flag[GEM_MAX_JOBS];
};

__device__ __forceinline__ bool getenv_table_on(const EllMcParams& p) { return p.hot_table != 0; }

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Per-CTA table for the popular levels.  A same-address L2 reduction costs ~10 ns, so the most popular level of a
// categorical column (a fifth of all rows hit it) bounds the whole kernel when every hit goes to L2: 75 K hits
// -> 0.79 ms at the Covertype x 7 shape.  Each CTA therefore keeps a direct-mapped table of GEM_TABLE features (two-way) in
// shared memory: a feature claims a free slot with a CAS on first use, later hits are eight native integer
// shared-memory adds of the fixed-point residuals (fp32 shared-memory atomics are CAS loops on sm_100), slot
// collisions fall through to the L2 reduction, and the table is flushed with two L2 reductions per occupied slot
// at the end.  Fixed point: |coeff * residual| <= cmax < 2^e2 is scaled to 30 - R2 bits for 2^R2 rows per CTA;
// the table is used when that leaves at least 18 bits (<= 4096 rows per CTA), otherwise every hit goes to L2.
constexpr int GEM_TABLE = 4096;
constexpr int GEM_TABLE_BITS = 12;

// run-length cache of one lane: the feature it is accumulating for and the residual vector so far
struct LaneRun {
    int col, n;          // feature and number of rows accumulated
    float v[GEM_CP];
};
struct HotTable {
    int* tag;            // [GEM_TABLE], -1 = free
    int* val;            // [GEM_TABLE][8] fixed point
    float scale;         // 0 = table off
};
__device__ __forceinline__ void scatter(float* g, const HotTable& T, int col, const float (&v)[GEM_CP]) {
    if (T.scale != 0.f) {
        uint32_t slot = ((uint32_t)col * 2654435761u) >> (32 - GEM_TABLE_BITS);
#pragma unroll
        for (int probe = 0; probe < 2; ++probe, slot ^= 1u) {       // two-way: the slot and its neighbour
            int t = T.tag[slot];
            if (t == -1) {
                const int old = atomicCAS(&T.tag[slot], -1, col);
                t = (old == -1) ? col : old;
            }
            if (t == col) {
#pragma unroll
                for (int c = 0; c < GEM_CP; ++c) atomicAdd(&T.val[slot * GEM_CP + c], __float2int_rn(v[c] * T.scale));
                return;
            }
        }
    }
    red_add_v4(g + (size_t)col * GEM_CP, v[0], v[1], v[2], v[3]);
    red_add_v4(g + (size_t)col * GEM_CP + 4, v[4], v[5], v[6], v[7]);
}
// A run of n rows is one addend of magnitude <= n * bound; the sum of all addends of a CTA is still bounded by its
// row count times the bound, so runs of any length (the bias feature: every row of a warp) fit the fixed-point slot.
__device__ __forceinline__ void run_flush(float* g, const HotTable& T, LaneRun& R) {
    if (R.col >= 0) scatter(g, T, R.col, R.v);
}
__device__ __forceinline__ void run_add(float* g, const HotTable& T, LaneRun& R, int col, const float (&r)[GEM_CP]) {
    if (col != R.col) {
        run_flush(g, T, R);
        R.n = 0;
        R.col = col;
#pragma unroll
        for (int c = 0; c < GEM_CP; ++c) R.v[c] = r[c];
    } else {
#pragma unroll
        for (int c = 0; c < GEM_CP; ++c) R.v[c] += r[c];
    }
    R.n += 1;
}

__global__ void __launch_bounds__(GEM_THREADS) grad_ell_mc_kernel(const EllMcParams p) {
    __shared__ uint32_t s_code;
    extern __shared__ __align__(16) int gem_smem[];               // tags, then values
    int* s_tag = gem_smem;
    int* s_val = gem_smem + GEM_TABLE;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int G = gridDim.x, b = blockIdx.x;
    for (int i = tid; i < GEM_TABLE; i += GEM_THREADS) s_tag[i] = -1;
    for (int i = tid; i < GEM_TABLE * GEM_CP; i += GEM_THREADS) s_val[i] = 0;
    int jb = 0;
    while (jb + 1 < p.n_jobs && b >= p.job_cta0[jb + 1]) ++jb;
    const int b0 = p.job_cta0[jb], Gj = p.job_cta0[jb + 1] - b0, bl = b - b0, seg0 = p.job_seg0[jb];
    float* gacc = p.gacc + (size_t)jb * p.d * GEM_CP;
    const int n4 = p.d * GEM_CP / 4;

    // zero this message's accumulator (independent of the parameters)
    for (int i = bl * GEM_THREADS + tid; i < n4; i += Gj * GEM_THREADS)
        reinterpret_cast<float4*>(gacc)[i] = make_float4(0.f, 0.f, 0.f, 0.f);

    if (tid == 0) {
        uint32_t code = GEM_GO;
        if (b == 0) {
            if (p.beta_flag) {
                const uint64_t t0 = globaltimer();
                while (true) {
                    const uint32_t f = ld_acquire_sys(p.beta_flag);
                    if (f == p.stamp) break;
                    if (f > p.stamp) { code = GEM_SKIP; break; }
                    if (p.err && ld_acquire_sys(p.err) != 0) { code = GEM_SKIP; break; }
                    if (globaltimer() - t0 > p.timeout_ns) {
                        if (p.err) atomicCAS(p.err, 0u, (uint32_t)(-EH_ETIMEOUT));
                        code = GEM_SKIP;
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
        s_code = code;
    }
    __syncthreads();
    if (s_code == GEM_SKIP) return;

    uint32_t gen = 0;
    grid_barrier(&p.hdr->barrier, gen);               // accumulators are zero everywhere

    const int rps = p.rows_per_seg, nnz = p.nnz;
    const int job_rows = rps * p.job_nsegs[jb];
    const int gwarp = bl * (GEM_THREADS / 32) + warp, nwarps = Gj * (GEM_THREADS / 32);
    const bool l0 = lane < nnz, l1 = lane + 32 < nnz;
    const int grp = lane >> 2;                          // class owned by this lane after the exchange
    const float4* W4 = reinterpret_cast<const float4*>(p.W);
    LaneRun run0, run1;
    run0.col = run1.col = -1;
    run0.n = run1.n = 0;
    // fixed-point format of the table (see HotTable): bound on |coeff * residual| and rows this CTA adds up
    HotTable T;
    T.tag = s_tag; T.val = s_val; T.scale = 0.f;
    int e2 = 0, fbits = 0;
    {
        float cmax = 0.f;
        for (int i = 0; i < p.job_nsegs[jb]; ++i) cmax = fmaxf(cmax, fabsf(p.coeff[seg0 + i]));
        if (!(cmax > 1e-30f)) cmax = 1e-30f;
        frexpf(cmax, &e2);                                         // softmax / logistic residuals are <= 1 in magnitude
        const long long rows_cta = (long long)((job_rows + nwarps - 1) / nwarps) * (GEM_THREADS / 32);
        int R2 = 0;
        while ((1ll << R2) < rows_cta) ++R2;
        fbits = 30 - R2;
        if (fbits >= 18 && getenv_table_on(p)) T.scale = ldexpf(1.0f, fbits - e2);
    }
    int sg = seg0 + gwarp / rps, lr = gwarp % rps;
    // the next row's indices, label and weight are requested one iteration ahead (index stream latency)
    const int32_t* ri_n = nullptr;
    int c0_n = -1, c1_n = -1, sg_n = sg, lr_n = lr;
    float y_n = 0.f;
    auto fetch = [&](int row) {
        if (row < job_rows) {
            ri_n = p.seg_idx[sg_n] + (size_t)lr_n * nnz;
            c0_n = l0 ? __ldg(ri_n + lane) : -1;
            c1_n = l1 ? __ldg(ri_n + lane + 32) : -1;
            y_n = __ldg(p.seg_y[sg_n] + lr_n);
        }
    };
    fetch(gwarp);
    for (int row = gwarp; row < job_rows; row += nwarps) {
        const int32_t* ri = ri_n;
        const int c0 = c0_n, c1 = c1_n;
        const float yraw = y_n;
        const int yc = (int)yraw;
        sg = sg_n; lr = lr_n;
        const float cf = p.coeff[sg];
        lr_n += nwarps;
        while (lr_n >= rps) { lr_n -= rps; ++sg_n; }
        fetch(row + nwarps);
        float lg[GEM_CP];
#pragma unroll
        for (int c = 0; c < GEM_CP; ++c) lg[c] = 0.f;
        if (c0 >= 0) {
            const float4 a = W4[(size_t)c0 * 2], bb = W4[(size_t)c0 * 2 + 1];       // L1-cacheable: popular features stay on the SM
            lg[0] = a.x; lg[1] = a.y; lg[2] = a.z; lg[3] = a.w; lg[4] = bb.x; lg[5] = bb.y; lg[6] = bb.z; lg[7] = bb.w;
        }
        if (c1 >= 0) {
            const float4 a = W4[(size_t)c1 * 2], bb = W4[(size_t)c1 * 2 + 1];       // L1-cacheable: popular features stay on the SM
            lg[0] += a.x; lg[1] += a.y; lg[2] += a.z; lg[3] += a.w; lg[4] += bb.x; lg[5] += bb.y; lg[6] += bb.z; lg[7] += bb.w;
        }
        for (int k = lane + 64; k < nnz; k += 32) {             // rows with more than 64 ones (not the named shapes)
            const int c = __ldg(ri + k);
            const float4 a = W4[(size_t)c * 2], bb = W4[(size_t)c * 2 + 1];       // L1-cacheable: popular features stay on the SM
            lg[0] += a.x; lg[1] += a.y; lg[2] += a.z; lg[3] += a.w; lg[4] += bb.x; lg[5] += bb.y; lg[6] += bb.z; lg[7] += bb.w;
        }
        // 8 sums over 32 lanes: halves trade classes (0..3 <-> 4..7), quarters (pairs), eighths (singles)
        const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4;
        float q[4];
#pragma unroll
        for (int c = 0; c < 4; ++c)
            q[c] = (h16 ? lg[4 + c] : lg[c]) + __shfl_xor_sync(0xffffffffu, h16 ? lg[c] : lg[4 + c], 16);
        float d2[2];
#pragma unroll
        for (int c = 0; c < 2; ++c)
            d2[c] = (h8 ? q[2 + c] : q[c]) + __shfl_xor_sync(0xffffffffu, h8 ? q[c] : q[2 + c], 8);
        float m = (h4 ? d2[1] : d2[0]) + __shfl_xor_sync(0xffffffffu, h4 ? d2[0] : d2[1], 4);
        m += __shfl_xor_sync(0xffffffffu, m, 2);
        m += __shfl_xor_sync(0xffffffffu, m, 1);
        // lane group grp now holds the logit of class (bit order of the exchange): h16 -> +4, h8 -> +2, h4 -> +1
        const int cls = grp;                                   // (lane>>4)*4 + ((lane>>3)&1)*2 + ((lane>>2)&1) == lane>>2
        float r;
        if (p.model == 2) {                                    // softmax cross-entropy
            const float z = cls < p.C ? m : -INFINITY;
            float mx = z;
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
            const float e = cls < p.C ? __expf(z - mx) : 0.f;
            float se = e;
            se += __shfl_xor_sync(0xffffffffu, se, 4);
            se += __shfl_xor_sync(0xffffffffu, se, 8);
            se += __shfl_xor_sync(0xffffffffu, se, 16);
            r = cf * (__fdividef(e, se) - (cls == yc ? 1.f : 0.f));
        } else {                                               // C independent logistic columns, labels +-1 (model 3)
            const float yv = yraw;
            r = cls < p.C ? cf * __fdividef(-yv, __expf(m * yv) + 1.0f) : 0.f;
        }
        float rv[GEM_CP];
#pragma unroll
        for (int c = 0; c < GEM_CP; ++c) rv[c] = __shfl_sync(0xffffffffu, r, 4 * c);
        if (c0 >= 0) run_add(gacc, T, run0, c0, rv);
        if (c1 >= 0) run_add(gacc, T, run1, c1, rv);
        for (int k = lane + 64; k < nnz; k += 32) {
            const int c = __ldg(ri + k);
            red_add_v4(gacc + (size_t)c * GEM_CP, rv[0], rv[1], rv[2], rv[3]);
            red_add_v4(gacc + (size_t)c * GEM_CP + 4, rv[4], rv[5], rv[6], rv[7]);
        }
    }
    run_flush(gacc, T, run0);
    run_flush(gacc, T, run1);
    __syncthreads();
    if (T.scale != 0.f) {
        const float inv = ldexpf(1.0f, e2 - fbits);
        for (int i = tid; i < GEM_TABLE; i += GEM_THREADS) {
            const int t = s_tag[i];
            if (t >= 0) {
                float v[GEM_CP];
#pragma unroll
                for (int c = 0; c < GEM_CP; ++c) v[c] = (float)s_val[i * GEM_CP + c] * inv;
                red_add_v4(gacc + (size_t)t * GEM_CP, v[0], v[1], v[2], v[3]);
                red_add_v4(gacc + (size_t)t * GEM_CP + 4, v[4], v[5], v[6], v[7]);
            }
        }
    }

    grid_barrier(&p.hdr->barrier, gen);               // all reductions have landed

    float4* out4 = reinterpret_cast<float4*>(p.job_out[jb]);
    for (int i = bl * GEM_THREADS + tid; i < n4; i += Gj * GEM_THREADS)
        out4[i] = __ldcg(reinterpret_cast<const float4*>(gacc) + i);

    __syncthreads();
    if (tid == 0) {
        __threadfence_system();
        const uint32_t old = atom_add_acq_rel_gpu(&p.hdr->done, 1u);
        if (old == (uint32_t)G - 1) {
            p.hdr->barrier = 0;
            p.hdr->done = 0;
            for (int j = 0; j < p.n_jobs; ++j)
                if (p.job_flag[j]) st_arrival_time(p.job_flag[j]);
            __threadfence_system();
            for (int j = 0; j < p.n_jobs; ++j)
                if (p.job_flag[j]) st_relaxed_sys(p.job_flag[j], p.stamp);
        }
    }
}

}  // namespace eh

extern "C" int eh_grad_ell_mc_padded_classes(void) { return eh::GEM_CP; }

extern "C" size_t eh_grad_ell_mc_scratch_bytes(int d) {
    return sizeof(eh::EllMcHdr) + (size_t)eh::GEM_MAX_JOBS * (size_t)d * eh::GEM_CP * sizeof(float);
}

extern "C" int eh_grad_ell_mc_multi(const eh_grad_ell_job* jobs, int n_jobs, int classes, void* d_scratch,
                                    size_t scratch_bytes, void* stream) {
    using namespace eh;
    EH_REQUIRE(jobs && n_jobs >= 1 && n_jobs <= GEM_MAX_JOBS && d_scratch, "grad_ell_mc: bad job list (1..%d messages)",
               GEM_MAX_JOBS);
    const eh_grad_ell_job* j = &jobs[0];
    EH_REQUIRE(j->d_beta, "grad_ell_mc: null parameters");
    EH_REQUIRE(j->rows_per_seg > 0 && j->d > 0 && j->nnz > 0, "grad_ell_mc: empty problem");
    EH_REQUIRE(classes >= 1 && classes <= GEM_CP, "grad_ell_mc: 1..%d classes supported on one-hot features (got %d)",
               GEM_CP, classes);
    EH_REQUIRE(j->model == 2 || j->model == 3, "grad_ell_mc: model must be 2 (softmax) or 3 (independent logistic columns)");
    EH_REQUIRE((long long)j->rows_per_seg * GEM_MAX_SEGS < (1ll << 31), "grad_ell_mc: more than 2^31 rows in one launch");
    EH_REQUIRE(((uintptr_t)j->d_beta % 16) == 0, "grad_ell_mc: parameters must be 16-byte aligned");
    const size_t need = sizeof(EllMcHdr) + (size_t)n_jobs * j->d * GEM_CP * sizeof(float);
    EH_REQUIRE(scratch_bytes >= need, "grad_ell_mc: scratch too small (%zu < %zu)", scratch_bytes, need);
    EllMcParams p;
    memset(&p, 0, sizeof(p));
    int nseg = 0;
    for (int q = 0; q < n_jobs; ++q) {
        const eh_grad_ell_job* jq = &jobs[q];
        EH_REQUIRE(jq->d_out && ((uintptr_t)jq->d_out % 16) == 0, "grad_ell_mc: destination must be non-null and 16-byte aligned");
        EH_REQUIRE(jq->n_segs > 0 && jq->n_segs <= EH_MAX_SEGS, "grad_ell_mc: n_segs out of range");
        EH_REQUIRE(jq->d == j->d && jq->nnz == j->nnz && jq->rows_per_seg == j->rows_per_seg && jq->model == j->model &&
                       jq->d_beta == j->d_beta && jq->d_beta_flag == j->d_beta_flag && jq->stamp == j->stamp,
                   "grad_ell_mc: batched messages must share shape, model, parameters and stamp");
        EH_REQUIRE(nseg + jq->n_segs <= GEM_MAX_SEGS, "grad_ell_mc: more than %d partitions in one batch", GEM_MAX_SEGS);
        p.job_seg0[q] = nseg;
        p.job_nsegs[q] = jq->n_segs;
        p.job_out[q] = jq->d_out;
        p.job_flag[q] = jq->d_flag;
        for (int i = 0; i < jq->n_segs; ++i, ++nseg) {
            EH_REQUIRE(jq->d_idx[i] && jq->d_y[i], "grad_ell_mc: null segment %d", i);
            p.seg_idx[nseg] = jq->d_idx[i];
            p.seg_y[nseg] = jq->d_y[i];
            p.coeff[nseg] = jq->coeff[i];
        }
    }
    p.n_jobs = n_jobs;
    p.hot_table = (getenv("EH_ELL_MC_TABLE") && atoi(getenv("EH_ELL_MC_TABLE")) == 0) ? 0 : 1;
    p.W = j->d_beta; p.beta_flag = j->d_beta_flag; p.err = j->d_err;
    p.hdr = reinterpret_cast<EllMcHdr*>(d_scratch);
    p.gacc = reinterpret_cast<float*>(reinterpret_cast<char*>(d_scratch) + sizeof(EllMcHdr));
    p.d = j->d; p.nnz = j->nnz; p.rows_per_seg = j->rows_per_seg; p.C = classes; p.model = j->model;
    p.stamp = j->stamp; p.seq = next_launch_seq();
    p.timeout_ns = j->timeout_ns ? j->timeout_ns : 30ull * 1000000000ull;
    // CTAs per message in proportion to its partitions; at least one warp-row of work per warp
    const size_t smem = (size_t)GEM_TABLE * (1 + GEM_CP) * sizeof(int);
    EH_SET_MAX_SMEM(grad_ell_mc_kernel, smem);
    int occ = 0;
    EH_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, grad_ell_mc_kernel, GEM_THREADS, smem));
    if (occ > 2) occ = 2;
    if (occ < 1) occ = 1;
    const long long total_rows = (long long)j->rows_per_seg * nseg;
    long long grid = (long long)worker_sms() * occ;
    long long want = (total_rows * 32 + GEM_THREADS - 1) / GEM_THREADS;
    const long long cols = ((long long)j->d * GEM_CP / 4 + GEM_THREADS - 1) / GEM_THREADS;
    if (want < cols) want = cols;
    if (grid > want) grid = want;
    if (grid < n_jobs) grid = n_jobs;
    const long long spare = grid - n_jobs;
    int used = 0;
    p.job_cta0[0] = 0;
    for (int q = 0; q < n_jobs; ++q) {
        used += 1 + (int)(spare * p.job_nsegs[q] / nseg);
        p.job_cta0[q + 1] = used;
    }
    void* args[] = {(void*)&p};
    return launch_grid_sync((const void*)grad_ell_mc_kernel, dim3((unsigned)used), dim3(GEM_THREADS), args, smem,
                            (cudaStream_t)stream);
}
