// Encoded partial gradient for one-hot (ELL) features: every stored value is 1.0 and every row
// has the same number of ones (the reference's OneHotEncoder output, arrange_real_data.py:41-64).
//
//   p_r = sum_k beta[idx[r,k]]                      (gather; beta is L2 resident)
//   g[idx[r,k]] += coeff[seg(r)] * resid(p_r, y_r)  (scatter, red.global.add.f32)
//
// Reference: scipy CSR matvec + CSR^T matvec in float64 (approximate_coding.py:194-196 with a
// csr_matrix).  One warp owns a row; the always-on bias column (every row hits it) is summed in
// registers and added once per warp instead of `rows` colliding atomics.  The accumulation
// vector lives in `scratch`, zeroed and shipped (to local or peer memory + release flag) by
// the same cooperative kernel.
#include "common.cuh"

namespace eh {

constexpr int GE_THREADS = 256;
constexpr int GE_MAX_SEGS = EH_MAX_SEGS;
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
};

__device__ __forceinline__ float ell_residual(int model, float p, float y) {
    if (model == EH_MODEL_LOGISTIC) return __fdividef(-y, __expf(p * y) + 1.0f);
    return -2.0f * (y - p);
}

__global__ void __launch_bounds__(GE_THREADS) grad_ell_kernel(const EllParams p) {
    __shared__ uint32_t s_code;
    const int tid = threadIdx.x, lane = tid & 31;
    const int gthreads = gridDim.x * GE_THREADS;
    const int gtid = blockIdx.x * GE_THREADS + tid;

    // zero the accumulator (does not depend on this round's parameters)
    for (int c = gtid; c < p.d; c += gthreads) p.gacc[c] = 0.f;

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
    for (long long row = gwarp; row < p.rows; row += nwarps) {
        const int sg = (int)(row / p.rows_per_seg);
        const int lr = (int)(row % p.rows_per_seg);
        const int32_t* ri = p.seg_idx[sg] + (size_t)lr * p.nnz;
        float s = 0.f;
        for (int k = lane; k < p.nnz; k += 32) s += __ldcg(p.beta + __ldg(ri + k));
        s = warp_sum(s);
        const float rs = p.coeff[sg] * ell_residual(p.model, s, __ldg(p.seg_y[sg] + lr));
        for (int k = lane; k < p.nnz; k += 32) {
            const int c = __ldg(ri + k);
            if (c == p.bias_col) bias_acc += rs;
            else atomicAdd(p.gacc + c, rs);
        }
    }
    if (p.bias_col >= 0) {
        bias_acc = warp_sum(bias_acc);
        if (lane == 0 && bias_acc != 0.f) atomicAdd(p.gacc + p.bias_col, bias_acc);
    }

    grid_barrier(&p.hdr->barrier, gen);

    for (int c = gtid; c < p.d; c += gthreads) p.out[c] = __ldcg(p.gacc + c);

    __syncthreads();
    if (tid == 0) {
        __threadfence_system();
        const uint32_t old = atom_add_acq_rel_gpu(&p.hdr->done, 1u);
        if (old == gridDim.x - 1) {
            p.hdr->barrier = 0;
            p.hdr->done = 0;
            __threadfence_system();
            if (p.flag) st_release_sys(p.flag, p.stamp);
        }
    }
}


}  // namespace eh

extern "C" size_t eh_grad_ell_scratch_bytes(int d) {
    return sizeof(eh::EllScratchHdr) + (size_t)(d + 32) * sizeof(float);
}

extern "C" int eh_grad_ell(const eh_grad_ell_job* j, void* d_scratch, size_t scratch_bytes, void* stream) {
    using namespace eh;
    EH_REQUIRE(j && j->d_beta && j->d_out && d_scratch, "grad_ell: null pointer");
    EH_REQUIRE(j->n_segs > 0 && j->n_segs <= EH_MAX_SEGS, "grad_ell: n_segs out of range");
    EH_REQUIRE(j->rows_per_seg > 0 && j->d > 0 && j->nnz > 0, "grad_ell: empty problem");
    EH_REQUIRE(scratch_bytes >= eh_grad_ell_scratch_bytes(j->d), "grad_ell: scratch too small");
    EllParams p;
    memset(&p, 0, sizeof(p));
    for (int i = 0; i < j->n_segs; ++i) {
        EH_REQUIRE(j->d_idx[i] && j->d_y[i], "grad_ell: null segment %d", i);
        p.seg_idx[i] = j->d_idx[i];
        p.seg_y[i] = j->d_y[i];
        p.coeff[i] = j->coeff[i];
    }
    p.beta = j->d_beta; p.out = j->d_out;
    p.flag = j->d_flag; p.beta_flag = j->d_beta_flag; p.err = j->d_err;
    p.hdr = reinterpret_cast<EllScratchHdr*>(d_scratch);
    p.gacc = reinterpret_cast<float*>(reinterpret_cast<char*>(d_scratch) + sizeof(EllScratchHdr));
    p.rows = (long long)j->rows_per_seg * j->n_segs; p.d = j->d; p.nnz = j->nnz;
    p.rows_per_seg = j->rows_per_seg; p.n_segs = j->n_segs;
    p.model = j->model; p.bias_col = j->bias_col; p.stamp = j->stamp; p.seq = next_launch_seq();
    p.timeout_ns = j->timeout_ns ? j->timeout_ns : 30ull * 1000000000ull;
    const DevInfo& di = dev_info();
    int occ = 0;
    EH_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, grad_ell_kernel, GE_THREADS, 0));
    if (occ > 4) occ = 4;
    if (occ < 1) occ = 1;
    long long want = (p.rows * 32 + GE_THREADS - 1) / GE_THREADS;      // one warp per row
    long long cols = ((long long)j->d + GE_THREADS - 1) / GE_THREADS;
    if (cols > want) want = cols;
    long long grid = (long long)di.sm_count * occ;
    if (grid > want) grid = want;
    if (grid < 1) grid = 1;
    void* args[] = {(void*)&p};
    { int lrc = launch_grid_sync((const void*)grad_ell_kernel, dim3((unsigned)grid), dim3(GE_THREADS), args, 0, (cudaStream_t)stream); if (lrc) return lrc; }
    return EH_OK;
}
