// Encoded partial gradient for one-hot (ELL) features: every stored value is 1.0 and every row
// has the same number of ones (the reference's OneHotEncoder output, arrange_real_data.py:41-64).
//
//   p_r = sum_k beta[idx[r,k]]                      (gather; beta is L2 resident)
//   g[idx[r,k]] += coeff[seg(r)] * resid(p_r, y_r)  (scatter, red.global.add.f32)
//
// Reference: scipy CSR matvec + CSR^T matvec in float64 (approximate_coding.py:194-196 with a
// csr_matrix).  One warp owns a row; the always-on bias column (every row hits it) is summed in
// registers and added once per warp instead of `rows` colliding atomics.  The other popular
// one-hot values (a categorical column's most frequent level is hit by a large share of the
// rows) would still serialise as same-address L2 atomics -- ncu showed the Amazon-shaped kernel
// bound by exactly that -- so every CTA keeps a small direct-mapped table of accumulators in
// shared memory: a column claims a free slot with a CAS on first use, later hits are shared-memory
// atomics, slot collisions fall through to the L2 atomic, and the table is flushed with one L2
// atomic per occupied slot at the end.  The accumulation vector lives in `scratch`, zeroed and
// shipped (to local or peer memory + release flag) by the same cooperative kernel.
#include <stdlib.h>

#include "common.cuh"

namespace eh {

constexpr int GE_THREADS = 512;
constexpr int GE_TABLE = 2048;                 // per-CTA hot-column accumulators (direct-mapped)
constexpr int GE_MAX_SEGS = 64;                // partitions over all batched messages
constexpr int GE_MAX_JOBS = 16;                 // messages batched into one launch (shared-memory variant)
constexpr uint32_t GE_GO = 1, GE_SKIP = 2;

struct EllScratchHdr {
    alignas(128) uint32_t barrier;
    alignas(128) uint32_t done;
    alignas(128) uint32_t decision;
    alignas(128) uint32_t pad;
};

struct EllParams {
    const int32_t* seg_idx[GE_MAX_SEGS];
    const float* seg_y[GE_MAX_SEGS];
    const float* beta;
    float* out;
    uint32_t* flag;
    const uint32_t* beta_flag;
    uint32_t* err;
    EllScratchHdr* hdr;
    float* gacc;
    long long rows;
    int d, nnz, rows_per_seg, n_segs, model, bias_col;
    uint32_t stamp, seq;
    unsigned long long timeout_ns;
    float coeff[GE_MAX_SEGS];
    // batched messages (shared-memory variant): job j owns CTAs [job_cta0[j], job_cta0[j+1]) and segments
    // [job_seg0[j], job_seg0[j] + job_nsegs[j]) and writes job_out[j] / releases job_flag[j]
    int n_jobs;
    int job_cta0[GE_MAX_JOBS + 1], job_seg0[GE_MAX_JOBS], job_nsegs[GE_MAX_JOBS];
    float* job_out[GE_MAX_JOBS];
    uint32_t* job_flag[GE_MAX_JOBS];
};

__device__ __forceinline__ float ell_residual(int model, float p, float y) {
    if (model == EH_MODEL_LOGISTIC) return __fdividef(-y, __expf(p * y) + 1.0f);
    return -2.0f * (y - p);
}

__global__ void __launch_bounds__(GE_THREADS) grad_ell_kernel(const EllParams p) {
    __shared__ uint32_t s_code;
    __shared__ int s_tag[GE_TABLE];
    __shared__ float s_val[GE_TABLE];
    for (int i = threadIdx.x; i < GE_TABLE; i += GE_THREADS) { s_tag[i] = -1; s_val[i] = 0.f; }
    const int tid = threadIdx.x, lane = tid & 31;
    // which batched message this CTA works for: message jb owns CTAs [job_cta0[jb], job_cta0[jb+1]), its own
    // accumulation vector in scratch and its own destination / arrival record
    int jb = 0;
    while (jb + 1 < p.n_jobs && (int)blockIdx.x >= p.job_cta0[jb + 1]) ++jb;
    const int b0 = p.job_cta0[jb], Gj = p.job_cta0[jb + 1] - b0, seg0 = p.job_seg0[jb];
    const int gthreads = Gj * GE_THREADS;
    const int gtid = ((int)blockIdx.x - b0) * GE_THREADS + tid;
    float* gacc = p.gacc + (size_t)jb * (size_t)((p.d + 31) & ~31);
    const long long job_rows = (long long)p.rows_per_seg * p.job_nsegs[jb];

    // zero the accumulator (does not depend on this round's parameters)
    for (int c = gtid; c < p.d; c += gthreads) gacc[c] = 0.f;

    if (tid == 0) {
        uint32_t code = GE_GO;
        if (blockIdx.x == 0) {
            if (p.beta_flag) {
                const uint64_t t0 = globaltimer();
                while (true) {
                    const uint32_t f = ld_acquire_sys(p.beta_flag);
                    if (f == p.stamp) break;
                    if (f > p.stamp) { code = GE_SKIP; break; }
                    if (p.err && ld_acquire_sys(p.err) != 0) { code = GE_SKIP; break; }
                    if (globaltimer() - t0 > p.timeout_ns) {
                        if (p.err) atomicCAS(p.err, 0u, (uint32_t)(-EH_ETIMEOUT));
                        code = GE_SKIP;
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
    if (s_code == GE_SKIP) return;

    uint32_t gen = 0;
    grid_barrier(&p.hdr->barrier, gen);

    const long long gwarp = gtid >> 5;
    const long long nwarps = gthreads >> 5;
    float bias_acc = 0.f;
    for (long long row = gwarp; row < job_rows; row += nwarps) {
        const int sg = seg0 + (int)(row / p.rows_per_seg);
        const int lr = (int)(row % p.rows_per_seg);
        const int32_t* ri = p.seg_idx[sg] + (size_t)lr * p.nnz;
        float s = 0.f;
        for (int k = lane; k < p.nnz; k += 32) s += p.beta[__ldg(ri + k)];          // L1-cacheable: popular features stay on the SM
        s = warp_sum(s);
        const float rs = p.coeff[sg] * ell_residual(p.model, s, __ldg(p.seg_y[sg] + lr));
        for (int k = lane; k < p.nnz; k += 32) {
            const int c = __ldg(ri + k);
            if (c == p.bias_col) {
                bias_acc += rs;
            } else {
                const uint32_t slot = ((uint32_t)c * 2654435761u) >> (32 - 11);
                int t = s_tag[slot];
                if (t == -1) {
                    const int old = atomicCAS(&s_tag[slot], -1, c);
                    t = (old == -1) ? c : old;
                }
                if (t == c) atomicAdd(&s_val[slot], rs);
                else atomicAdd(gacc + c, rs);
            }
        }
    }
    if (p.bias_col >= 0) {
        bias_acc = warp_sum(bias_acc);
        if (lane == 0 && bias_acc != 0.f) atomicAdd(gacc + p.bias_col, bias_acc);
    }
    __syncthreads();
    for (int i = tid; i < GE_TABLE; i += GE_THREADS) {
        const int t = s_tag[i];
        const float v = s_val[i];
        if (t >= 0 && v != 0.f) atomicAdd(gacc + t, v);
    }

    grid_barrier(&p.hdr->barrier, gen);

    {
        float* out = p.job_out[jb];
        for (int c = gtid; c < p.d; c += gthreads) out[c] = __ldcg(gacc + c);
    }

    __syncthreads();
    if (tid == 0) {
        __threadfence_system();
        const uint32_t old = atom_add_acq_rel_gpu(&p.hdr->done, 1u);
        if (old == gridDim.x - 1) {
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


// ---------------------------------------------------------------------------------------------
// Shared-memory variant for feature spaces that fit on chip (d <= ~56K, i.e. Covertype- and
// KC-House-shaped data): every CTA keeps a private copy of the gradient (and, if it also fits, of
// the parameters) in shared memory, so the 44-55 scatter-adds per row are shared-memory atomics
// instead of L2 atomics that pile up on the popular one-hot columns.  The per-CTA gradients are then
// combined through scratch in a fixed order like the dense kernel (bitwise reproducible).
// ---------------------------------------------------------------------------------------------
constexpr int GES_THREADS = 512;

template <bool BETA_IN_SMEM>
__global__ void __launch_bounds__(GES_THREADS, 1) grad_ell_smem_kernel(const EllParams p, float* __restrict__ partials) {
    extern __shared__ __align__(16) float ges_smem[];
    float* acc = ges_smem;
    float* bsm = ges_smem + ((p.d + 3) & ~3);
    __shared__ uint32_t s_code;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int G = gridDim.x, b = blockIdx.x;

    for (int c = tid; c < p.d; c += GES_THREADS) acc[c] = 0.f;

    if (tid == 0) {
        uint32_t code = GE_GO;
        if (b == 0) {
            if (p.beta_flag) {
                const uint64_t t0 = globaltimer();
                while (true) {
                    const uint32_t f = ld_acquire_sys(p.beta_flag);
                    if (f == p.stamp) break;
                    if (f > p.stamp) { code = GE_SKIP; break; }
                    if (p.err && ld_acquire_sys(p.err) != 0) { code = GE_SKIP; break; }
                    if (globaltimer() - t0 > p.timeout_ns) {
                        if (p.err) atomicCAS(p.err, 0u, (uint32_t)(-EH_ETIMEOUT));
                        code = GE_SKIP;
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
    if (s_code == GE_SKIP) return;

    if (BETA_IN_SMEM) {
        for (int c = tid; c < p.d; c += GES_THREADS) bsm[c] = __ldcg(p.beta + c);
        __syncthreads();
    }

    // which batched message this CTA works for
    int jb = 0;
    while (jb + 1 < p.n_jobs && b >= p.job_cta0[jb + 1]) ++jb;
    const int b0 = p.job_cta0[jb], Gj = p.job_cta0[jb + 1] - b0, bl = b - b0, seg0 = p.job_seg0[jb];
    const long long job_rows = (long long)p.rows_per_seg * p.job_nsegs[jb];
    const long long gwarp = (long long)bl * (GES_THREADS / 32) + warp;
    const long long nwarps = (long long)Gj * (GES_THREADS / 32);
    float bias_acc = 0.f;
    constexpr int RW = 4;                      // rows in flight per warp: hides the index-load latency
    for (long long row0 = gwarp * RW; row0 < job_rows; row0 += nwarps * RW) {
        int cols[RW][2];
        float yv[RW], cf[RW], s[RW];
        const int32_t* rip[RW];
#pragma unroll
        for (int r = 0; r < RW; ++r) {
            const long long row = row0 + r;
            const bool valid = row < job_rows;
            const int sg = seg0 + (valid ? (int)(row / p.rows_per_seg) : 0);
            const int lr = valid ? (int)(row % p.rows_per_seg) : 0;
            rip[r] = p.seg_idx[sg] + (size_t)lr * p.nnz;
            cols[r][0] = (valid && lane < p.nnz) ? __ldg(rip[r] + lane) : -1;
            cols[r][1] = (valid && lane + 32 < p.nnz) ? __ldg(rip[r] + lane + 32) : -1;
            yv[r] = valid ? __ldg(p.seg_y[sg] + lr) : 0.f;
            cf[r] = valid ? p.coeff[sg] : 0.f;
        }
#pragma unroll
        for (int r = 0; r < RW; ++r) {
            float t = 0.f;
            if (cols[r][0] >= 0) t += BETA_IN_SMEM ? bsm[cols[r][0]] : __ldcg(p.beta + cols[r][0]);
            if (cols[r][1] >= 0) t += BETA_IN_SMEM ? bsm[cols[r][1]] : __ldcg(p.beta + cols[r][1]);
            if (p.nnz > 64 && row0 + r < job_rows)
                for (int k = lane + 64; k < p.nnz; k += 32) {
                    const int c = __ldg(rip[r] + k);
                    t += BETA_IN_SMEM ? bsm[c] : __ldcg(p.beta + c);
                }
            s[r] = t;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int r = 0; r < RW; ++r) s[r] += __shfl_xor_sync(0xffffffffu, s[r], o);
        }
#pragma unroll
        for (int r = 0; r < RW; ++r) {
            const float rs = cf[r] * ell_residual(p.model, s[r], yv[r]);
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                if (cols[r][q] >= 0) {
                    if (cols[r][q] == p.bias_col) bias_acc += rs;
                    else atomicAdd(acc + cols[r][q], rs);
                }
            }
            if (p.nnz > 64 && row0 + r < job_rows)
                for (int k = lane + 64; k < p.nnz; k += 32) {
                    const int c = __ldg(rip[r] + k);
                    if (c == p.bias_col) bias_acc += rs;
                    else atomicAdd(acc + c, rs);
                }
        }
    }
    if (p.bias_col >= 0) {
        bias_acc = warp_sum(bias_acc);
        if (lane == 0 && bias_acc != 0.f) atomicAdd(acc + p.bias_col, bias_acc);
    }
    __syncthreads();
    float* mine = partials + (size_t)b * p.d;
    for (int c = tid; c < p.d; c += GES_THREADS) __stcg(mine + c, acc[c]);

    uint32_t gen = 0;
    grid_barrier(&p.hdr->barrier, gen);

    // ordered reduction per message: CTA bl of the message owns a contiguous column range
    const int chunk = (p.d + Gj - 1) / Gj;
    const int c0 = bl * chunk, c1 = min(p.d, c0 + chunk);
    float* out = p.job_out[jb];
    for (int c = c0 + tid; c < c1; c += GES_THREADS) {
        float t = 0.f;
        for (int g = b0; g < b0 + Gj; ++g) t += __ldcg(partials + (size_t)g * p.d + c);
        out[c] = t;
    }
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
                if (p.job_flag[j]) st_relaxed_sys(p.job_flag[j], p.stamp);       // ordered by the fence above
        }
    }
}


// ---------------------------------------------------------------------------------------------
// Fixed-point shared-memory variant (d <= ~28K: Covertype- and KC-House-shaped data).
//
// fp32 atomicAdd on shared memory is not a hardware operation on sm_100 (SASS: an
// ATOMS.CAST.SPIN compare-and-swap loop), a 32-bit integer add is (ATOMS.ADD).  So the scatter
// accumulates every addend as a fixed-point integer split over two 32-bit words,
//
//     v = round(rs * 2^S),  |v| <= 2^W,      hi += v >> LB,   lo += v & (2^LB - 1)
//
// with S from a bound on |rs| (|coeff| * |y|max for the logistic residual, 2 |coeff| (|y|max +
// nnz * |beta|max) for the linear one) and W / LB from the number of rows a CTA adds up so that
// neither word can overflow (rows <= 65536: W = 30, LB = 16 -- 30 bits below the bound, finer
// than the fp32 mantissa).  Integer adds commute exactly, so the per-CTA sums -- and after the
// fixed-order combination of the CTAs of a message, the result -- are bitwise reproducible
// whatever order the warps run in.  The hot loop is branch-free (padding lanes of a row go to a
// dummy column, the bias column is an ordinary column), uses 32-bit row arithmetic only, and
// requests the next four rows of a warp before it processes the current four.
// No reference counterpart (scipy CSR^T matvec in float64, approximate_coding.py:194-196).
// ---------------------------------------------------------------------------------------------
constexpr int GFX_THREADS = 1024;            // fixed-point variant: 32 warps per SM hide the gather / shuffle / atomic chain
template <bool BETA_IN_SMEM, bool WIDE>
__global__ void __launch_bounds__(GFX_THREADS, 1) grad_ell_fx_kernel(const EllParams p, float* __restrict__ partials) {
    extern __shared__ __align__(16) float ges_smem[];
    // hi[d + 32], lo[d + 32] accumulator words (separate arrays: consecutive columns fall into consecutive banks),
    // then [d + 32] parameters; slots d .. d+31 are dummy columns, one per lane, that absorb the padding lanes (a row
    // of 55 ones occupies 64 lanes) so the hot loop has no branches -- one per lane because nine padding lanes adding
    // to ONE dummy word were a nine-way same-address conflict in every second scatter instruction
    const int dd = ((p.d + 32) + 3) & ~3;
    int* acc_hi = reinterpret_cast<int*>(ges_smem);
    int* acc_lo = acc_hi + dd;
    float* bsm = ges_smem + 2 * dd;
    __shared__ uint32_t s_code;
    __shared__ float s_red[2][GFX_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int G = gridDim.x, b = blockIdx.x;

    for (int c = tid; c < p.d + 32; c += GFX_THREADS) { acc_hi[c] = 0; acc_lo[c] = 0; }

    // which batched message this CTA works for
    int jb = 0;
    while (jb + 1 < p.n_jobs && b >= p.job_cta0[jb + 1]) ++jb;
    const int b0 = p.job_cta0[jb], Gj = p.job_cta0[jb + 1] - b0, bl = b - b0, seg0 = p.job_seg0[jb];
    // 32-bit row arithmetic throughout (the host checks rows < 2^31): a 64-bit division per row costs more
    // instructions than the rest of the row's work
    const int rps = p.rows_per_seg;
    const int job_rows = rps * p.job_nsegs[jb];
    const int gwarp = bl * (GFX_THREADS / 32) + warp;
    const int nwarps = Gj * (GFX_THREADS / 32);
    constexpr int RW = 4;                      // rows per warp and step
    const int nnz = p.nnz, dummy = p.d + lane;

    // |y|max over this CTA's rows and |coeff|max over the message's partitions (independent of the parameters)
    float ymax = 0.f, cmax = 0.f;
    for (int i = 0; i < p.job_nsegs[jb]; ++i) cmax = fmaxf(cmax, fabsf(p.coeff[seg0 + i]));
    {   // lane-parallel over this warp's rows: 32 independent label loads per step
        const int stride = nwarps * RW;
        const int n_sets = (job_rows - gwarp * RW + stride - 1) / stride;          // <= 0 if the warp has no rows
        for (int k = lane; k < n_sets * RW; k += 32) {
            const int row = gwarp * RW + (k / RW) * stride + (k % RW);
            if (row < job_rows) ymax = fmaxf(ymax, fabsf(__ldg(p.seg_y[seg0 + row / rps] + row % rps)));
        }
    }

    if (tid == 0) {
        uint32_t code = GE_GO;
        if (b == 0) {
            if (p.beta_flag) {
                const uint64_t t0 = globaltimer();
                while (true) {
                    const uint32_t f = ld_acquire_sys(p.beta_flag);
                    if (f == p.stamp) break;
                    if (f > p.stamp) { code = GE_SKIP; break; }
                    if (p.err && ld_acquire_sys(p.err) != 0) { code = GE_SKIP; break; }
                    if (globaltimer() - t0 > p.timeout_ns) {
                        if (p.err) atomicCAS(p.err, 0u, (uint32_t)(-EH_ETIMEOUT));
                        code = GE_SKIP;
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
    if (s_code == GE_SKIP) return;

    // parameters -> shared memory (if they fit); |beta|max is needed only for the linear residual's bound
    float bmax = 0.f;
    if (BETA_IN_SMEM) {
        for (int c = tid; c < p.d; c += GFX_THREADS) {
            const float t = __ldcg(p.beta + c);
            bsm[c] = t;
            bmax = fmaxf(bmax, fabsf(t));
        }
        if (tid < 32) bsm[p.d + tid] = 0.f;
    } else if (p.model != EH_MODEL_LOGISTIC) {
        for (int c = tid; c < p.d; c += GFX_THREADS) bmax = fmaxf(bmax, fabsf(__ldcg(p.beta + c)));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        ymax = fmaxf(ymax, __shfl_xor_sync(0xffffffffu, ymax, o));
        bmax = fmaxf(bmax, __shfl_xor_sync(0xffffffffu, bmax, o));
    }
    if (lane == 0) { s_red[0][warp] = ymax; s_red[1][warp] = bmax; }
    __syncthreads();
    ymax = 0.f; bmax = 0.f;
#pragma unroll
    for (int w = 0; w < GFX_THREADS / 32; ++w) { ymax = fmaxf(ymax, s_red[0][w]); bmax = fmaxf(bmax, s_red[1][w]); }
    // bound on |rs| and the fixed-point format
    float bound = (p.model == EH_MODEL_LOGISTIC) ? cmax * ymax : 2.0f * cmax * (ymax + (float)nnz * bmax);
    if (!(bound > 1e-30f)) bound = 1e-30f;
    int e2;
    frexpf(bound, &e2);                                            // bound < 2^e2
    const long long rows_cta = (long long)((job_rows + nwarps * RW - 1) / (nwarps * RW)) * (GFX_THREADS / 32) * RW;
    int R2 = 0;
    while ((1ll << R2) < rows_cta) ++R2;
    const int LB = R2 <= 16 ? 16 : 32 - R2;
    const int W = R2 <= 16 ? 30 : 63 - 2 * R2;                     // rows * 2^LB < 2^32 and rows * 2^(W-LB) < 2^31
    const float scale = ldexpf(1.0f, W - e2);
    const unsigned lo_mask = (1u << LB) - 1u;

    struct RowSet {
        int c0[RW], c1[RW];
        float yv[RW], cf[RW];
        const int32_t* rip[RW];
    };
    const bool l0 = lane < nnz, l1 = lane + 32 < nnz;
    const bool two = nnz > 32;                 // rows of at most 32 ones (KC-House: 19) need no second lane pass
    // (sg, lr) = partition and local row of the first row of a set, advanced without divisions
    auto load_rows = [&](int row0, int sg0, int lr0, RowSet& R) {
#pragma unroll
        for (int r = 0; r < RW; ++r) {
            const bool valid = row0 + r < job_rows;
            int sg = sg0, lr = lr0 + r;
            while (valid && lr >= rps) { lr -= rps; ++sg; }
            if (!valid) { sg = sg0; lr = lr0; }
            const int32_t* ri = p.seg_idx[sg] + (size_t)lr * nnz;
            R.rip[r] = ri;
            R.c0[r] = (valid && l0) ? __ldg(ri + lane) : dummy;
            R.c1[r] = (two && valid && l1) ? __ldg(ri + lane + 32) : dummy;
            R.yv[r] = __ldg(p.seg_y[sg] + lr);
            R.cf[r] = valid ? p.coeff[sg] : 0.f;
        }
    };
    auto gather = [&](int c) -> float {
        if (BETA_IN_SMEM) return bsm[c];
        return c < p.d ? p.beta[c] : 0.f;                 // L1-cacheable (nothing of beta was read before the flag)
    };
    auto process = [&](const RowSet& R, int row0) {
        float s[RW];
#pragma unroll
        for (int r = 0; r < RW; ++r) {
            float t = gather(R.c0[r]);
            if (two) t += gather(R.c1[r]);
            if (WIDE) {
                if (row0 + r < job_rows)
                    for (int k = lane + 64; k < nnz; k += 32) t += gather(__ldg(R.rip[r] + k));
            }
            s[r] = t;
        }
        // four row sums with 6 shuffles instead of 20: the halves of the warp trade rows (0,1 <-> 2,3), then the
        // quarters (0 <-> 1, 2 <-> 3), then three butterflies inside each group of 8 lanes; lane group g = lane / 8
        // ends up with the margin of row g, evaluates that row's residual once, and the four fixed-point values
        // are broadcast back
        static_assert(RW == 4, "the row-sum exchange below is written for four rows");
        const bool up = lane & 16, q1 = lane & 8;
        float a0 = (up ? s[2] : s[0]) + __shfl_xor_sync(0xffffffffu, up ? s[0] : s[2], 16);
        float a1 = (up ? s[3] : s[1]) + __shfl_xor_sync(0xffffffffu, up ? s[1] : s[3], 16);
        float m = (q1 ? a1 : a0) + __shfl_xor_sync(0xffffffffu, q1 ? a0 : a1, 8);
        m += __shfl_xor_sync(0xffffffffu, m, 4);
        m += __shfl_xor_sync(0xffffffffu, m, 2);
        m += __shfl_xor_sync(0xffffffffu, m, 1);
        const int g = lane >> 3;
        const float my_y = g == 0 ? R.yv[0] : g == 1 ? R.yv[1] : g == 2 ? R.yv[2] : R.yv[3];
        const float my_cf = g == 0 ? R.cf[0] : g == 1 ? R.cf[1] : g == 2 ? R.cf[2] : R.cf[3];
        const int my_v = __float2int_rn(my_cf * ell_residual(p.model, m, my_y) * scale);
#pragma unroll
        for (int r = 0; r < RW; ++r) {
            const int v = __shfl_sync(0xffffffffu, my_v, 8 * r);
            const int vh = v >> LB;
            const int vl = (int)((unsigned)v & lo_mask);
            atomicAdd(acc_hi + R.c0[r], vh); atomicAdd(acc_lo + R.c0[r], vl);
            if (two) { atomicAdd(acc_hi + R.c1[r], vh); atomicAdd(acc_lo + R.c1[r], vl); }
            if (WIDE) {
                if (row0 + r < job_rows)
                    for (int k = lane + 64; k < nnz; k += 32) {
                        const int c = __ldg(R.rip[r] + k);
                        atomicAdd(acc_hi + c, vh); atomicAdd(acc_lo + c, vl);
                    }
            }
        }
    };
    RowSet A, B;
    const int step = nwarps * RW;
    int row0 = gwarp * RW;
    int sg_n = seg0 + row0 / rps, lr_n = row0 % rps;             // position of the set being loaded next
    auto advance = [&]() {
        lr_n += step;
        while (lr_n >= rps) { lr_n -= rps; ++sg_n; }
    };
    if (row0 < job_rows) {
        load_rows(row0, sg_n, lr_n, A);
        advance();
        while (true) {                                           // two sets ping-pong: no register copies
            if (row0 + step < job_rows) { load_rows(row0 + step, sg_n, lr_n, B); advance(); }
            process(A, row0);
            row0 += step;
            if (row0 >= job_rows) break;
            if (row0 + step < job_rows) { load_rows(row0 + step, sg_n, lr_n, A); advance(); }
            process(B, row0);
            row0 += step;
            if (row0 >= job_rows) break;
        }
    }
    __syncthreads();
    float* mine = partials + (size_t)b * p.d;
    const double inv = (double)ldexpf(1.0f, e2 - W);
    const double hi_w = (double)(1u << LB);
    for (int c = tid; c < p.d; c += GFX_THREADS) {
        __stcg(mine + c, (float)(((double)acc_hi[c] * hi_w + (double)(unsigned)acc_lo[c]) * inv));
    }

    uint32_t gen = 0;
    grid_barrier(&p.hdr->barrier, gen);

    // ordered reduction per message: CTA bl of the message owns a contiguous column range.  Eight thread groups
    // each add every eighth partial of a column (adjacent threads = adjacent columns: coalesced), the eight
    // sub-sums meet in shared memory (the accumulators are no longer needed) and are added in a fixed order --
    // a single thread walking all Gj partials of its column was a chain of up to 148 loads
    const int chunk = (p.d + Gj - 1) / Gj;
    const int c0 = bl * chunk, c1 = min(p.d, c0 + chunk);
    float* out = p.job_out[jb];
    constexpr int SUB = 8, CW = GFX_THREADS / SUB;
    float* sred = ges_smem;
    const int sub = tid / CW, cx = tid % CW;
    if (Gj < 64) {
        // few partials per column (batched messages): one thread per column is the shorter path
        for (int c = c0 + tid; c < c1; c += GFX_THREADS) {
            float t = 0.f;
            int g = b0;
            for (; g + 8 <= b0 + Gj; g += 8) {                    // eight loads in flight, added in CTA order
                float u[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) u[q] = __ldcg(partials + (size_t)(g + q) * p.d + c);
#pragma unroll
                for (int q = 0; q < 8; ++q) t += u[q];
            }
            for (; g < b0 + Gj; ++g) t += __ldcg(partials + (size_t)g * p.d + c);
            out[c] = t;
        }
    } else
    for (int cb = c0; cb < c1; cb += CW) {
        const int c = cb + cx;
        float t = 0.f;
        if (c < c1)
            for (int g = b0 + sub; g < b0 + Gj; g += SUB) t += __ldcg(partials + (size_t)g * p.d + c);
        sred[sub * CW + cx] = t;
        __syncthreads();
        if (sub == 0 && c < c1) {
            float u = sred[cx];
#pragma unroll
            for (int q = 1; q < SUB; ++q) u += sred[q * CW + cx];
            out[c] = u;
        }
        __syncthreads();
    }
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
                if (p.job_flag[j]) st_relaxed_sys(p.job_flag[j], p.stamp);       // ordered by the fence above
        }
    }
}

constexpr size_t GES_SMEM_LIMIT = 224 * 1024;

}  // namespace eh

extern "C" size_t eh_grad_ell_scratch_bytes(int d) {
    // global-atomic variant: one accumulation vector per batched message
    size_t b = sizeof(eh::EllScratchHdr) + (size_t)eh::GE_MAX_JOBS * (size_t)((d + 31) / 32 * 32 + 32) * sizeof(float);
    if ((size_t)(d + 4) * sizeof(float) <= eh::GES_SMEM_LIMIT) {      // shared-memory variant: per-CTA partials
        int sm = eh::dev_info().sm_count;
        if (sm <= 0) sm = 160;
        b = sizeof(eh::EllScratchHdr) + (size_t)sm * (size_t)d * sizeof(float) + 256;
    }
    return b;
}

// Up to GE_MAX_JOBS messages of one shape in ONE launch (CTA ranges per message, proportional to their partition
// counts) -- in the shared-memory variants with per-CTA partial gradients, in the global-atomic variant (feature
// spaces too large for shared memory, e.g. the Amazon shape) with one accumulation vector per message in scratch.
extern "C" int eh_grad_ell_multi(const eh_grad_ell_job* jobs, int n_jobs, void* d_scratch, size_t scratch_bytes,
                                 void* stream) {
    using namespace eh;
    EH_REQUIRE(jobs && n_jobs >= 1 && n_jobs <= GE_MAX_JOBS && d_scratch, "grad_ell: bad job list (1..%d messages)", GE_MAX_JOBS);
    const eh_grad_ell_job* j = &jobs[0];
    EH_REQUIRE(j->d_beta, "grad_ell: null pointer");
    EH_REQUIRE(j->rows_per_seg > 0 && j->d > 0 && j->nnz > 0, "grad_ell: empty problem");
    EH_REQUIRE((long long)j->rows_per_seg * GE_MAX_SEGS < (1ll << 31), "grad_ell: more than 2^31 rows in one launch");
    EH_REQUIRE(scratch_bytes >= eh_grad_ell_scratch_bytes(j->d), "grad_ell: scratch too small");
    const DevInfo& di = dev_info();
    const size_t acc_bytes = (size_t)((j->d + 3) & ~3) * sizeof(float);
    const bool smem_variant = acc_bytes <= GES_SMEM_LIMIT && getenv("EH_ELL_GLOBAL") == nullptr;
    EllParams p;
    memset(&p, 0, sizeof(p));
    int nseg = 0;
    for (int q = 0; q < n_jobs; ++q) {
        const eh_grad_ell_job* jq = &jobs[q];
        EH_REQUIRE(jq->d_out, "grad_ell: null destination");
        EH_REQUIRE(jq->n_segs > 0 && jq->n_segs <= EH_MAX_SEGS, "grad_ell: n_segs out of range");
        EH_REQUIRE(jq->d == j->d && jq->nnz == j->nnz && jq->rows_per_seg == j->rows_per_seg && jq->model == j->model &&
                       jq->bias_col == j->bias_col && jq->d_beta == j->d_beta && jq->d_beta_flag == j->d_beta_flag &&
                       jq->stamp == j->stamp,
                   "grad_ell: batched messages must share shape, model, parameters and stamp");
        EH_REQUIRE(nseg + jq->n_segs <= GE_MAX_SEGS, "grad_ell: more than %d partitions in one batch", GE_MAX_SEGS);
        p.job_seg0[q] = nseg;
        p.job_nsegs[q] = jq->n_segs;
        p.job_out[q] = jq->d_out;
        p.job_flag[q] = jq->d_flag;
        for (int i = 0; i < jq->n_segs; ++i, ++nseg) {
            EH_REQUIRE(jq->d_idx[i] && jq->d_y[i], "grad_ell: null segment %d", i);
            p.seg_idx[nseg] = jq->d_idx[i];
            p.seg_y[nseg] = jq->d_y[i];
            p.coeff[nseg] = jq->coeff[i];
        }
    }
    p.n_jobs = n_jobs;
    p.beta = j->d_beta; p.out = j->d_out;
    p.flag = j->d_flag; p.beta_flag = j->d_beta_flag; p.err = j->d_err;
    p.hdr = reinterpret_cast<EllScratchHdr*>(d_scratch);
    p.gacc = reinterpret_cast<float*>(reinterpret_cast<char*>(d_scratch) + sizeof(EllScratchHdr));
    p.rows = (long long)j->rows_per_seg * j->n_segs; p.d = j->d; p.nnz = j->nnz;
    p.rows_per_seg = j->rows_per_seg; p.n_segs = j->n_segs;
    p.model = j->model; p.bias_col = j->bias_col; p.stamp = j->stamp; p.seq = next_launch_seq();
    p.timeout_ns = j->timeout_ns ? j->timeout_ns : 30ull * 1000000000ull;
    // fixed-point variant (native integer shared-memory atomics) when two words per column fit
    const size_t fx_col_bytes = (size_t)(((j->d + 32) + 3) & ~3) * sizeof(float);     // one word per column incl. 32 dummies
    const bool fx_variant = smem_variant && 2 * fx_col_bytes <= GES_SMEM_LIMIT && getenv("EH_ELL_FLOAT_ATOMICS") == nullptr;
    if (smem_variant) {
        const bool both = fx_variant ? 3 * fx_col_bytes <= GES_SMEM_LIMIT : 2 * acc_bytes <= GES_SMEM_LIMIT;
        const size_t smem = fx_variant ? (both ? 3 : 2) * fx_col_bytes : (both ? 2 * acc_bytes : acc_bytes);
        // CTAs per message in proportion to its rows (32 rows of work per warp at least), every message >= 1
        const long long total_rows = (long long)j->rows_per_seg * nseg;
        long long grid = worker_sms();
        const int threads = fx_variant ? GFX_THREADS : GES_THREADS;
        const long long want = (total_rows * 32 + threads - 1) / threads;
        if (grid > want) grid = want;
        if (grid < n_jobs) grid = n_jobs;
        const long long spare = grid - n_jobs;
        int used = 0;
        p.job_cta0[0] = 0;
        for (int q = 0; q < n_jobs; ++q) {
            used += 1 + (int)(spare * p.job_nsegs[q] / nseg);
            p.job_cta0[q + 1] = used;
        }
        grid = used;
        float* partials = p.gacc;
        EH_REQUIRE(scratch_bytes >= sizeof(EllScratchHdr) + (size_t)grid * j->d * sizeof(float),
                   "grad_ell: scratch too small for the shared-memory variant");
        const void* kern;
        if (fx_variant) {
            const bool wide = j->nnz > 64;
            if (both && !wide) { kern = (const void*)grad_ell_fx_kernel<true, false>; EH_SET_MAX_SMEM((grad_ell_fx_kernel<true, false>), (int)GES_SMEM_LIMIT); }
            else if (both) { kern = (const void*)grad_ell_fx_kernel<true, true>; EH_SET_MAX_SMEM((grad_ell_fx_kernel<true, true>), (int)GES_SMEM_LIMIT); }
            else if (!wide) { kern = (const void*)grad_ell_fx_kernel<false, false>; EH_SET_MAX_SMEM((grad_ell_fx_kernel<false, false>), (int)GES_SMEM_LIMIT); }
            else { kern = (const void*)grad_ell_fx_kernel<false, true>; EH_SET_MAX_SMEM((grad_ell_fx_kernel<false, true>), (int)GES_SMEM_LIMIT); }
        } else {
            kern = both ? (const void*)grad_ell_smem_kernel<true> : (const void*)grad_ell_smem_kernel<false>;
            if (both) { EH_SET_MAX_SMEM((grad_ell_smem_kernel<true>), (int)GES_SMEM_LIMIT); }
            else { EH_SET_MAX_SMEM((grad_ell_smem_kernel<false>), (int)GES_SMEM_LIMIT); }
        }
        void* args[] = {(void*)&p, (void*)&partials};
        return launch_grid_sync(kern, dim3((unsigned)grid), dim3(threads), args, smem, (cudaStream_t)stream);
    }
    int occ = 0;
    EH_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, grad_ell_kernel, GE_THREADS, 0));
    if (occ > 2) occ = 2;                       // fewer, larger CTAs: fewer tables to flush
    if (occ < 1) occ = 1;
    // CTAs per message in proportion to its partitions: one warp per row, and enough threads to zero / ship d columns
    const long long total_rows = (long long)j->rows_per_seg * nseg;
    long long want = (total_rows * 32 + GE_THREADS - 1) / GE_THREADS;
    const long long cols = (((long long)j->d + GE_THREADS - 1) / GE_THREADS) * n_jobs / 4;      // ~4 columns per thread
    if (cols > want) want = cols;
    long long grid = (long long)worker_sms() * occ;
    if (grid > want) grid = want;
    if (grid < n_jobs) grid = n_jobs;
    {
        const long long spare = grid - n_jobs;
        int used = 0;
        p.job_cta0[0] = 0;
        for (int q = 0; q < n_jobs; ++q) {
            used += 1 + (int)(spare * p.job_nsegs[q] / nseg);
            p.job_cta0[q + 1] = used;
        }
        grid = used;
    }
    void* args[] = {(void*)&p};
    { int lrc = launch_grid_sync((const void*)grad_ell_kernel, dim3((unsigned)grid), dim3(GE_THREADS), args, 0, (cudaStream_t)stream); if (lrc) return lrc; }
    return EH_OK;
}

extern "C" int eh_grad_ell(const eh_grad_ell_job* j, void* d_scratch, size_t scratch_bytes, void* stream) {
    EH_REQUIRE(j && j->d_out, "grad_ell: null pointer");
    return eh_grad_ell_multi(j, 1, d_scratch, scratch_bytes, stream);
}
