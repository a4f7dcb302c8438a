// Batched evaluation on the 5th-gen tensor cores: P = X . [beta_0 .. beta_{R-1}] with the loss of
// every iterate reduced in the epilogue (tcgen05.mma kind::tf32, accumulators in TMEM).
//
// Reference: the evaluation epilogue of every driver loops over the R stored iterates and does
// X_train.dot(beta) + calculate_loss per iterate on the master (naive.py:186-198,
// util.py:136-141) -- R passes over X.  Here X is streamed ONCE:
//
//   warp 0   TMA producer: 2-D tensor-map loads (cp.async.bulk.tensor, 128B swizzle) of a
//            128-row x 32-float X tile and the matching [BN x 32] slice of the iterate matrix
//            into a 6-stage shared-memory ring (mbarrier expect-tx);
//   warp 1   MMA issuer: one elected thread issues 4 x tcgen05.mma.cta_group::1.kind::tf32
//            (M=128, N=BN, K=8) per stage into one of two TMEM accumulator buffers and
//            tcgen05.commit's the stage back to the producer / the tile to the epilogue;
//   warp 2   TMEM allocator (tcgen05.alloc / dealloc);
//   warps 4-7 epilogue: tcgen05.ld the 128 x BN accumulator (one row per thread), evaluate
//            log(1+exp(-y p)) or (y-p)^2 against the row's label and keep per-column running sums
//            in registers across all tiles of the persistent CTA; optionally store P (test-set
//            margins feed the AUC kernel).  The epilogue of tile i overlaps the MMAs of tile i+1.
//
// TF32 keeps 10 mantissa bits of X and beta; margins agree with fp32 to ~1e-3 relative, which is
// below the "%5.3f" resolution of the reference's result files.  `eh_margins_dense` (fp32 CUDA
// cores) remains available as the exact path.
#include "tc_common.cuh"

namespace eh {

constexpr int TC_STAGES = 6;

struct TcParams {
    long long rows;
    int R;              // valid columns
    int num_k;          // K steps of 32 floats
    int kind;           // 0 logistic, 1 squared error
    const float* y;
    double* loss_out;   // [R] sums (atomicAdd) or NULL
    float* P;           // [rows, ldp] or NULL
    int ldp;
    int col0;           // first iterate handled by this launch
};

template <int NB>   // BN = 16 * NB accumulator columns
__global__ void __launch_bounds__(TC_THREADS, 1)
eval_tc_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapB, const TcParams p) {
    constexpr int BN = 16 * NB;
    constexpr int B_BYTES = BN * TC_BK * 4;
    constexpr int STAGE_BYTES = TC_A_BYTES + B_BYTES;
    constexpr uint32_t TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : 256;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    // 1024-byte alignment of every operand tile is required by the 128B swizzle
    unsigned char* tiles = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full = reinterpret_cast<uint64_t*>(tiles + TC_STAGES * STAGE_BYTES);
    uint64_t* empty = full + TC_STAGES;
    uint64_t* tfull = empty + TC_STAGES;       // [2] accumulator ready
    uint64_t* tempty = tfull + 2;              // [2] accumulator drained
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
    float* sred = reinterpret_cast<float*>(tiles);   // reused after the main loop: [128][BN+1]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long n_tiles = (p.rows + TC_BM - 1) / TC_BM;

    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 4); }
        mbar_fence_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            const uint64_t pol_x = l2_policy_evict_first(), pol_b = l2_policy_evict_last();
            uint32_t it = 0;
            for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                for (int k = 0; k < p.num_k; ++k, ++it) {
                    const int st = it % TC_STAGES;
                    const uint32_t use = it / TC_STAGES;
                    if (use > 0) mbar_wait(&empty[st], (use - 1) & 1);
                    unsigned char* a = tiles + (size_t)st * STAGE_BYTES;
                    mbar_arrive_expect_tx(&full[st], STAGE_BYTES);
                    tma_load_2d(a, &mapX, k * TC_BK, (int)(t * TC_BM), &full[st], pol_x);
                    tma_load_2d(a + TC_A_BYTES, &mapB, k * TC_BK, p.col0, &full[st], pol_b);
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(TC_BM, BN);
            uint32_t it = 0, ti = 0;
            for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x, ++ti) {
                const uint32_t as = ti & 1, ause = ti >> 1;
                if (ause > 0) mbar_wait(&tempty[as], (ause - 1) & 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                for (int k = 0; k < p.num_k; ++k, ++it) {
                    const int st = it % TC_STAGES;
                    mbar_wait(&full[st], (it / TC_STAGES) & 1);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(tiles + (size_t)st * STAGE_BYTES);
                    const uint64_t a_desc = make_kmajor_sw128_desc(a_addr);
                    const uint64_t b_desc = make_kmajor_sw128_desc(a_addr + TC_A_BYTES);
#pragma unroll
                    for (int kk = 0; kk < TC_BK / 8; ++kk) {
                        // advance 8 tf32 = 32 bytes inside the swizzle atom: +2 in the (addr >> 4) field
                        umma_tf32(d_tmem, a_desc + (uint64_t)(kk * 2), b_desc + (uint64_t)(kk * 2), idesc,
                                  (k > 0 || kk > 0) ? 1u : 0u);
                    }
                    umma_commit(&empty[st]);          // frees the smem stage when these MMAs retire
                }
                umma_commit(&tfull[as]);              // accumulator complete
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------ epilogue
        const int q = warp - 4;                       // TMEM lane quarter == warp % 4
        float acc[BN];
#pragma unroll
        for (int c = 0; c < BN; ++c) acc[c] = 0.f;
        uint32_t ti = 0;
        for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x, ++ti) {
            const uint32_t as = ti & 1, ause = ti >> 1;
            const long long row = t * TC_BM + q * 32 + lane;
            const bool valid = row < p.rows;
            const float yv = (valid && p.y) ? __ldg(p.y + row) : 0.f;
            mbar_wait(&tfull[as], ause & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
#pragma unroll
            for (int cb = 0; cb < NB; ++cb) {
                float v[16];
                tmem_ld16(taddr + cb * 16, v);
                if (p.P && valid) {
                    float* dst = p.P + row * p.ldp + p.col0 + cb * 16;
#pragma unroll
                    for (int e = 0; e < 16; ++e)
                        if (cb * 16 + e < p.R) dst[e] = v[e];
                }
                if (p.loss_out) {
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        float l;
                        if (p.kind == 0) {
                            const float z = -yv * v[e];
                            l = fmaxf(z, 0.f) + __logf(1.f + __expf(-fabsf(z)));
                        } else {
                            const float dlt = yv - v[e];
                            l = dlt * dlt;
                        }
                        acc[cb * 16 + e] += valid ? l : 0.f;
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[as]);
        }
        // ---- column sums over the 128 rows handled by the epilogue threads
        if (p.loss_out) {
            // all MMAs of this CTA have retired (every tfull was consumed) and no TMA is in flight
            asm volatile("bar.sync 3, 128;" ::: "memory");
            const int rloc = q * 32 + lane;
#pragma unroll
            for (int c = 0; c < BN; ++c) sred[rloc * (BN + 1) + c] = acc[c];
            asm volatile("bar.sync 3, 128;" ::: "memory");
            for (int c = rloc; c < BN; c += 128) {
                if (c < p.R) {
                    double s = 0.0;
                    for (int r = 0; r < 128; ++r) s += (double)sred[r * (BN + 1) + c];
                    atomicAdd(p.loss_out + p.col0 + c, s);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ---------------------------------------------------------------- host
template <int NB>
static int launch_tc(const CUtensorMap& mx, const CUtensorMap& mb, const TcParams& p, cudaStream_t st) {
    constexpr int BN = 16 * NB;
    constexpr size_t smem = 1024 + (size_t)TC_STAGES * (TC_A_BYTES + BN * TC_BK * 4) + 256;
    static_assert(128 * (BN + 1) * 4 <= TC_STAGES * (TC_A_BYTES + BN * TC_BK * 4), "reduction scratch fits the ring");
    auto kern = eval_tc_kernel<NB>;
    EH_SET_MAX_SMEM((kern), (int)smem);
    const long long n_tiles = (p.rows + TC_BM - 1) / TC_BM;
    long long grid = dev_info().sm_count;
    if (grid > n_tiles) grid = n_tiles;
    kern<<<(unsigned)grid, TC_THREADS, smem, st>>>(mx, mb, p);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

}  // namespace eh

extern "C" int eh_eval_tc(const float* d_X, int64_t rows, int d, int ld, const float* d_Bt, int R, int ldb,
                          const float* d_y, int kind, double* d_loss_sums, float* d_P, int ldp, void* stream) {
    using namespace eh;
    EH_REQUIRE(d_X && d_Bt && rows > 0 && d > 0 && R > 0, "eval_tc: bad argument");
    EH_REQUIRE((d_loss_sums == nullptr) || d_y, "eval_tc: labels required for the loss");
    EH_REQUIRE(d_loss_sums || d_P, "eval_tc: nothing to compute");
    EH_REQUIRE((ld % 4) == 0 && (ldb % 4) == 0 && ((uintptr_t)d_X % 16) == 0 && ((uintptr_t)d_Bt % 16) == 0,
               "eval_tc: X and Bt need 16-byte aligned rows");
    EH_REQUIRE(rows < (1ll << 31), "eval_tc: too many rows");
    const int dk = (d + TC_BK - 1) / TC_BK * TC_BK;
    EH_REQUIRE(dk <= ld && dk <= ldb, "eval_tc: rows must be zero-padded to a multiple of 32 columns (d=%d ld=%d ldb=%d)",
               d, ld, ldb);
    cudaStream_t st = (cudaStream_t)stream;
    CUtensorMap mx;
    int rc = make_map(&mx, d_X, (uint64_t)dk, (uint64_t)rows, (uint64_t)ld, TC_BM);
    if (rc) return rc;
    for (int c0 = 0; c0 < R; c0 += 128) {
        const int rc_cols = (R - c0) < 128 ? (R - c0) : 128;
        const int nb = rc_cols <= 16 ? 1 : rc_cols <= 32 ? 2 : rc_cols <= 64 ? 4 : 8;
        CUtensorMap mb;
        rc = make_map(&mb, d_Bt, (uint64_t)dk, (uint64_t)R, (uint64_t)ldb, (uint32_t)(16 * nb));
        if (rc) return rc;
        TcParams p;
        p.rows = rows; p.R = rc_cols; p.num_k = dk / TC_BK; p.kind = kind; p.y = d_y;
        p.loss_out = d_loss_sums; p.P = d_P; p.ldp = ldp; p.col0 = c0;
        switch (nb) {
            case 1: rc = launch_tc<1>(mx, mb, p, st); break;
            case 2: rc = launch_tc<2>(mx, mb, p, st); break;
            case 4: rc = launch_tc<4>(mx, mb, p, st); break;
            default: rc = launch_tc<8>(mx, mb, p, st); break;
        }
        if (rc) return rc;
    }
    return EH_OK;
}
