// Multi-column gradient on the tensor cores: softmax (multi-class) regression and batched
// independent logistic columns ("batched seeds") -- the several-column case the fused single
// vector kernel cannot serve (BASELINE.json: "tcgen05/TMEM tiles where w has several columns").
//
//   forward   P = X . W^T          (tcgen05.mma kind::tf32, M = 128 rows, N = CN classes, K = d)
//             epilogue per row:  r = coeff * (softmax(p) - onehot(y))      (model 2)
//                                r_c = coeff * -y / (exp(p_c y) + 1)        (model 3)
//             written transposed, Rt[CN, rows] (K-major operand of the second product),
//             plus the summed cross-entropy / logistic loss when asked (evaluation).
//   backward  G[j, c] += sum_rows X[row, j] * R[row, c]
//             (tcgen05.mma kind::tf32 with an MN-major A operand: the X tile is used as X^T
//             straight from the row-major TMA tile, M = 128 feature columns, N = CN, K = rows);
//             all d/128 accumulator chunks of a CTA live in TMEM (d/128 * CN <= 512 columns)
//             across its whole row range and are added to G[CN, ldw] once at the end.
//
// Parameters, gradients and messages use the class-major layout W[c][j] (ldw = d padded to 32),
// which is K-major for the forward product, so the decoder treats a C-column model as one flat
// vector of CN * ldw entries.  No counterpart in the reference (its models are single vectors).
#include <stdlib.h>
#include <string.h>

#include "tc_common.cuh"

namespace eh {

constexpr int MC_STAGES = 8;

struct McFwdParams {
    long long rows;
    int C, num_k, model;
    float coeff;
    const float* y;
    float* Rt;            // [CN, ldr]
    long long ldr;
    double* loss_out;     // [1] summed loss or NULL
};

template <int NB>
__global__ void __launch_bounds__(TC_THREADS, 1)
mc_forward_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapW, const McFwdParams p) {
    constexpr int BN = 16 * NB;
    constexpr int B_BYTES = BN * TC_BK * 4;
    constexpr int STAGE_BYTES = TC_A_BYTES + B_BYTES;
    constexpr uint32_t TMEM_COLS = (2 * BN <= 32) ? 32 : 64;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* tiles = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full = reinterpret_cast<uint64_t*>(tiles + MC_STAGES * STAGE_BYTES);
    uint64_t* empty = full + MC_STAGES;
    uint64_t* tfull = empty + MC_STAGES;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
    __shared__ double s_loss[4];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long n_tiles = (p.rows + TC_BM - 1) / TC_BM;

    if (threadIdx.x == 0) {
        for (int s = 0; s < MC_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 4); }
        mbar_fence_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            const uint64_t pol_x = l2_policy_evict_first(), pol_b = l2_policy_evict_last();
            uint32_t it = 0;
            for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                for (int k = 0; k < p.num_k; ++k, ++it) {
                    const int st = it % MC_STAGES;
                    const uint32_t use = it / MC_STAGES;
                    if (use > 0) mbar_wait(&empty[st], (use - 1) & 1);
                    unsigned char* a = tiles + (size_t)st * STAGE_BYTES;
                    mbar_arrive_expect_tx(&full[st], STAGE_BYTES);
                    tma_load_2d(a, &mapX, k * TC_BK, (int)(t * TC_BM), &full[st], pol_x);
                    tma_load_2d(a + TC_A_BYTES, &mapW, k * TC_BK, 0, &full[st], pol_b);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(TC_BM, BN);
            uint32_t it = 0, ti = 0;
            for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x, ++ti) {
                const uint32_t as = ti & 1, ause = ti >> 1;
                if (ause > 0) mbar_wait(&tempty[as], (ause - 1) & 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                for (int k = 0; k < p.num_k; ++k, ++it) {
                    const int st = it % MC_STAGES;
                    mbar_wait(&full[st], (it / MC_STAGES) & 1);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(tiles + (size_t)st * STAGE_BYTES);
                    const uint64_t a_desc = make_kmajor_sw128_desc(a_addr);
                    const uint64_t b_desc = make_kmajor_sw128_desc(a_addr + TC_A_BYTES);
#pragma unroll
                    for (int kk = 0; kk < TC_BK / 8; ++kk)
                        umma_tf32(d_tmem, a_desc + (uint64_t)(kk * 2), b_desc + (uint64_t)(kk * 2), idesc,
                                  (k > 0 || kk > 0) ? 1u : 0u);
                    umma_commit(&empty[st]);
                }
                umma_commit(&tfull[as]);
            }
        }
    } else if (warp >= 4) {
        const int q = warp - 4;
        double loss = 0.0;
        uint32_t ti = 0;
        for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x, ++ti) {
            const uint32_t as = ti & 1, ause = ti >> 1;
            const long long row = t * TC_BM + q * 32 + lane;
            const bool valid = row < p.rows;
            const float yv = valid ? __ldg(p.y + row) : 0.f;
            mbar_wait(&tfull[as], ause & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
            float v[BN];
#pragma unroll
            for (int cb = 0; cb < NB; ++cb) {
                float t16[16];
                tmem_ld16(taddr + cb * 16, t16);
#pragma unroll
                for (int e = 0; e < 16; ++e) v[cb * 16 + e] = t16[e];
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[as]);          // accumulator buffer may be overwritten
            if (p.model == 2) {
                // softmax residual + cross-entropy
                float m = -3.0e38f;
#pragma unroll
                for (int c = 0; c < BN; ++c) if (c < p.C) m = fmaxf(m, v[c]);
                float s = 0.f;
#pragma unroll
                for (int c = 0; c < BN; ++c) {
                    v[c] = (c < p.C) ? __expf(v[c] - m) : 0.f;
                    s += v[c];
                }
                const float inv = __fdividef(1.f, s);
                const int yc = (int)yv;
                float py = 0.f;
#pragma unroll
                for (int c = 0; c < BN; ++c) {
                    const float sm = v[c] * inv;
                    if (c == yc) py = sm;
                    v[c] = p.coeff * (sm - (c == yc ? 1.f : 0.f));
                }
                if (valid) loss += (double)(-__logf(fmaxf(py, 1e-30f)));
            } else {
#pragma unroll
                for (int c = 0; c < BN; ++c) {
                    const float z = v[c] * yv;
                    if (valid && c < p.C) loss += (double)(fmaxf(-z, 0.f) + __logf(1.f + __expf(-fabsf(z))));
                    v[c] = (c < p.C) ? p.coeff * __fdividef(-yv, __expf(z) + 1.0f) : 0.f;
                }
            }
            if (valid && p.Rt) {
#pragma unroll
                for (int c = 0; c < BN; ++c) p.Rt[(size_t)c * p.ldr + row] = round_tf32(v[c]);    // coalesced across lanes
            }
        }
        if (p.loss_out) {
            loss = warp_sum(loss);
            if (lane == 0) s_loss[q] = loss;
            asm volatile("bar.sync 3, 128;" ::: "memory");
            if (q == 0 && lane == 0) atomicAdd(p.loss_out, s_loss[0] + s_loss[1] + s_loss[2] + s_loss[3]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------ backward
struct McBwdParams {
    long long rows;
    int n_chunks;         // ceil(dk / 128) feature chunks, n_chunks * BN <= 512
    int d, ldw;
    float* G;             // [CN, ldw] accumulated with red.global.add
};

constexpr int MCB_KR = 32;                          // rows (K) per stage
constexpr int MCB_A_BYTES = 128 * MCB_KR * 4;       // 4 boxes of [32 rows x 32 cols]
constexpr int MCB_BOX_BYTES = 32 * MCB_KR * 4;

template <int NB>
__global__ void __launch_bounds__(TC_THREADS, 1)
mc_backward_kernel(const __grid_constant__ CUtensorMap mapX, const __grid_constant__ CUtensorMap mapR, const McBwdParams p) {
    constexpr int BN = 16 * NB;
    constexpr int B_BYTES = BN * MCB_KR * 4;        // Rt tile [BN x 32 rows], K-major
    constexpr int STAGE_BYTES = MCB_A_BYTES + B_BYTES;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* tiles = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full = reinterpret_cast<uint64_t*>(tiles + MC_STAGES * STAGE_BYTES);
    uint64_t* empty = full + MC_STAGES;
    uint64_t* tdone = empty + MC_STAGES;            // all MMAs of this CTA retired
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tdone + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long n_ksteps = (p.rows + MCB_KR - 1) / MCB_KR;
    const long long my_ksteps = (blockIdx.x < n_ksteps) ? (n_ksteps - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

    if (threadIdx.x == 0) {
        for (int s = 0; s < MC_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(tdone, 1);
        mbar_fence_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            const uint64_t pol_x = l2_policy_evict_first(), pol_b = l2_policy_evict_last();
            uint32_t it = 0;
            for (long long ks = 0; ks < my_ksteps; ++ks) {
                const int row0 = (int)((blockIdx.x + ks * gridDim.x) * MCB_KR);
                for (int ch = 0; ch < p.n_chunks; ++ch, ++it) {
                    const int st = it % MC_STAGES;
                    const uint32_t use = it / MC_STAGES;
                    if (use > 0) mbar_wait(&empty[st], (use - 1) & 1);
                    unsigned char* a = tiles + (size_t)st * STAGE_BYTES;
                    mbar_arrive_expect_tx(&full[st], STAGE_BYTES);
#pragma unroll
                    for (int bx = 0; bx < 4; ++bx)     // four 32-column boxes side by side along M
                        tma_load_2d(a + bx * MCB_BOX_BYTES, &mapX, ch * 128 + bx * 32, row0, &full[st], pol_x);
                    tma_load_2d(a + MCB_A_BYTES, &mapR, row0, 0, &full[st], pol_b);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(128, BN, /*a_mn_major=*/1, /*b_mn_major=*/0);
            uint32_t it = 0;
            for (long long ks = 0; ks < my_ksteps; ++ks) {
                for (int ch = 0; ch < p.n_chunks; ++ch, ++it) {
                    const int st = it % MC_STAGES;
                    mbar_wait(&full[st], (it / MC_STAGES) & 1);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(tiles + (size_t)st * STAGE_BYTES);
                    const uint64_t b_desc = make_kmajor_sw128_desc(a_addr + MCB_A_BYTES);
                    const uint32_t d_tmem = tmem_base + ch * BN;
#pragma unroll
                    for (int kk = 0; kk < MCB_KR / 8; ++kk) {
                        // A: X^T, MN-major (32B-atom swizzle); 8 K-rows per MMA = 1024 B, M atoms 4096 B apart
                        const uint64_t a_desc = make_mnmajor_sw128_32b_desc(a_addr + kk * 1024, MCB_BOX_BYTES);
                        umma_tf32(d_tmem, a_desc, b_desc + (uint64_t)(kk * 2), idesc, (ks > 0 || kk > 0) ? 1u : 0u);
                    }
                    umma_commit(&empty[st]);
                }
            }
            umma_commit(tdone);
        }
    } else if (warp >= 4) {
        const int q = warp - 4;
        if (my_ksteps > 0) {
            mbar_wait(tdone, 0);
            tc_fence_after();
            for (int ch = 0; ch < p.n_chunks; ++ch) {
                const int j = ch * 128 + q * 32 + lane;
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + ch * BN;
#pragma unroll
                for (int cb = 0; cb < NB; ++cb) {
                    float v[16];
                    tmem_ld16(taddr + cb * 16, v);
                    if (j < p.d) {
#pragma unroll
                        for (int e = 0; e < 16; ++e) atomicAdd(p.G + (size_t)(cb * 16 + e) * p.ldw + j, v[e]);
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// small helpers of the multi-column message path
__global__ void mc_wait_flag_kernel(const uint32_t* flag, uint32_t stamp, uint32_t* err, unsigned long long timeout_ns,
                                    uint32_t* skip) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        const uint64_t t0 = globaltimer();
        *skip = 0;
        while (true) {
            const uint32_t f = ld_acquire_sys(flag);
            if (f == stamp) break;
            if (f > stamp || (err && ld_acquire_sys(err) != 0)) { *skip = 1; break; }
            if (globaltimer() - t0 > timeout_ns) {
                if (err) atomicCAS(err, 0u, (uint32_t)(-EH_ETIMEOUT));
                *skip = 1;
                break;
            }
            nanosleep(64);
        }
    }
}
__global__ void mc_ship_kernel(const float* __restrict__ g, float* __restrict__ out, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = g[i];
    __threadfence_system();
}
__global__ void mc_flag_kernel(uint32_t* flag, uint32_t stamp, const uint32_t* skip) {
    if (threadIdx.x == 0 && blockIdx.x == 0 && !(skip && *skip)) {
        st_arrival_time(flag);
        st_release_sys(flag, stamp);
    }
}

template <int NB>
static int launch_fwd(const CUtensorMap& mx, const CUtensorMap& mw, const McFwdParams& p, cudaStream_t st) {
    constexpr size_t smem = 1024 + (size_t)MC_STAGES * (TC_A_BYTES + 16 * NB * TC_BK * 4) + 256;
    auto kern = mc_forward_kernel<NB>;
    EH_SET_MAX_SMEM((kern), (int)smem);
    const long long n_tiles = (p.rows + TC_BM - 1) / TC_BM;
    long long grid = worker_sms();
    if (grid > n_tiles) grid = n_tiles;
    kern<<<(unsigned)grid, TC_THREADS, smem, st>>>(mx, mw, p);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

template <int NB>
static int launch_bwd(const CUtensorMap& mx, const CUtensorMap& mr, const McBwdParams& p, cudaStream_t st) {
    constexpr size_t smem = 1024 + (size_t)MC_STAGES * (MCB_A_BYTES + 16 * NB * MCB_KR * 4) + 256;
    auto kern = mc_backward_kernel<NB>;
    EH_SET_MAX_SMEM((kern), (int)smem);
    const long long n_ksteps = (p.rows + MCB_KR - 1) / MCB_KR;
    long long grid = worker_sms();
    if (grid > n_ksteps) grid = n_ksteps;
    kern<<<(unsigned)grid, TC_THREADS, smem, st>>>(mx, mr, p);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

// one-pass kernel (grad_mc_fused.cu)
int mc_fused_stages(int dk, int CN);
size_t mc_fused_scratch_bytes(int d, int CN);
int mc_fused_launch(const eh_mc_job* j, void* fused_scratch, size_t fused_bytes, cudaStream_t st);
// bf16 rows, both products on tcgen05 (grad_mc_bf16.cu)
int mc_bf16_stages(int dk, int CN);
int mc_bf16_launch(const eh_mc_job* j, void* fused_scratch, size_t fused_bytes, cudaStream_t st, bool single_vector);
// one-pass kernel on thread-block clusters (grad_mc_cluster.cu)
bool mc_cluster_plan(int dk, int CN, int* cs_out, int* npc_out, int* stages_out);
int mc_cluster_launch(const eh_mc_job* j, void* fused_scratch, size_t fused_bytes, cudaStream_t st);

static size_t mc_twopass_scratch_bytes(int64_t rows_per_seg, int d, int CN) {
    const int64_t ldr = (rows_per_seg + 31) / 32 * 32;
    const int64_t ldw = (d + 31) / 32 * 32;
    return ((size_t)(CN * ldr + CN * ldw) * sizeof(float) + 1024 + 511) / 512 * 512;
}

}  // namespace eh

// Layout: [two-pass region: Rt, G, skip word][one-pass region: barrier header, per-CTA partials].
// The one-pass header must start out zero (callers allocate the scratch zero-filled).
extern "C" size_t eh_grad_mc_scratch_bytes(int64_t rows_per_seg, int d, int C) {
    const int CN = C <= 16 ? 16 : 32;
    size_t b = eh::mc_twopass_scratch_bytes(rows_per_seg, d, CN);
    const int dk = (d + 31) / 32 * 32;
    if (eh::mc_fused_stages(dk, CN) > 0 || eh::mc_cluster_plan(dk, CN, nullptr, nullptr, nullptr) ||
        eh::mc_bf16_stages(dk, CN) > 0)
        b += eh::mc_fused_scratch_bytes(d, CN);
    return b;
}

extern "C" int eh_grad_mc(const eh_mc_job* j, void* d_scratch, size_t scratch_bytes, void* stream) {
    using namespace eh;
    EH_REQUIRE(j && j->d_W && d_scratch, "grad_mc: null pointer");
    EH_REQUIRE(j->d_out || j->d_loss, "grad_mc: nothing to compute");
    EH_REQUIRE(j->C >= 1 && j->C <= 32, "grad_mc: C=%d classes not supported (1..32)", j->C);
    EH_REQUIRE(j->model == 2 || j->model == 3 || (j->model == 4 && j->x_bf16),
               "grad_mc: model must be 2 (softmax) or 3 (batched logistic); 4 (batched linear) needs bf16 rows");
    EH_REQUIRE(j->n_segs > 0 && j->n_segs <= EH_MAX_SEGS && j->rows_per_seg > 0 && j->d > 0, "grad_mc: bad shape");
    const int CN = j->C <= 16 ? 16 : 32, NB = CN / 16;
    const int dk = (j->d + 31) / 32 * 32;
    EH_REQUIRE(j->ld >= dk && (j->ld % 4) == 0 && j->ldw >= dk && (j->ldw % 4) == 0,
               "grad_mc: X rows (ld=%d) and W rows (ldw=%d) must be zero-padded to %d columns", j->ld, j->ldw, dk);
    const int n_chunks = (dk + 127) / 128;
    EH_REQUIRE(j->d_out == nullptr || n_chunks * CN <= 512,
               "grad_mc: d * classes exceeds one CTA's tensor memory (%d chunks x %d columns > 512)", n_chunks, CN);
    EH_REQUIRE(scratch_bytes >= eh_grad_mc_scratch_bytes(j->rows_per_seg, j->d, j->C), "grad_mc: scratch too small");
    cudaStream_t st = (cudaStream_t)stream;
    if (j->x_bf16) {
        EH_REQUIRE(j->d_out && !j->d_loss, "grad_mc: bf16 rows are supported for the gradient call only");
        const size_t off = mc_twopass_scratch_bytes(j->rows_per_seg, j->d, CN);
        return mc_bf16_launch(j, (char*)d_scratch + off, scratch_bytes - off, st, false);
    }
    // gradient of a narrow model: one pass over X, one launch (grad_mc_fused.cu)
    if (j->d_out && !j->d_loss && getenv("EH_MC_TWOPASS") == nullptr) {
        const size_t off = mc_twopass_scratch_bytes(j->rows_per_seg, j->d, CN);
        const char* which = getenv("EH_MC_ONEPASS");            // "cta" | "cluster" force one variant
        const bool cta_ok = mc_fused_stages(dk, CN) > 0 && !(which && strcmp(which, "cluster") == 0);
        if (cta_ok) return mc_fused_launch(j, (char*)d_scratch + off, scratch_bytes - off, st);
        // clusters pay off up to ~1024 columns (1.15-1.2x the two-pass kernels); wider rows only on request
        const bool cluster_wanted = (which && strcmp(which, "cluster") == 0) || dk <= 1024;
        if (cluster_wanted && mc_cluster_plan(dk, CN, nullptr, nullptr, nullptr))
            return mc_cluster_launch(j, (char*)d_scratch + off, scratch_bytes - off, st);
    }
    const long long ldr = ((long long)j->rows_per_seg + 31) / 32 * 32;
    float* Rt = (float*)d_scratch;
    float* G = Rt + (size_t)CN * ldr;
    uint32_t* skip = (uint32_t*)(G + (size_t)CN * j->ldw);
    if (j->d_beta_flag)
        mc_wait_flag_kernel<<<1, 32, 0, st>>>(j->d_beta_flag, j->stamp, j->d_err,
                                              j->timeout_ns ? j->timeout_ns : 30ull * 1000000000ull, skip);
    else
        EH_CUDA(cudaMemsetAsync(skip, 0, sizeof(uint32_t), st));
    if (j->d_out) EH_CUDA(cudaMemsetAsync(G, 0, (size_t)CN * j->ldw * sizeof(float), st));
    CUtensorMap mw;
    int rc = make_map(&mw, j->d_W, (uint64_t)dk, (uint64_t)CN, (uint64_t)j->ldw, (uint32_t)CN);
    if (rc) return rc;
    for (int sg = 0; sg < j->n_segs; ++sg) {
        EH_REQUIRE(j->d_X[sg] && j->d_y[sg], "grad_mc: null segment %d", sg);
        EH_REQUIRE(((uintptr_t)j->d_X[sg] % 16) == 0, "grad_mc: X must be 16-byte aligned");
        if (j->d_out) EH_CUDA(cudaMemsetAsync(Rt, 0, (size_t)CN * ldr * sizeof(float), st));
        CUtensorMap mx;
        rc = make_map(&mx, j->d_X[sg], (uint64_t)dk, (uint64_t)j->rows_per_seg, (uint64_t)j->ld, TC_BM);
        if (rc) return rc;
        McFwdParams f;
        f.rows = j->rows_per_seg; f.C = j->C; f.num_k = dk / TC_BK; f.model = j->model; f.coeff = j->coeff[sg];
        f.y = j->d_y[sg]; f.Rt = j->d_out ? Rt : nullptr; f.ldr = ldr; f.loss_out = j->d_loss;
        rc = (NB == 1) ? launch_fwd<1>(mx, mw, f, st) : launch_fwd<2>(mx, mw, f, st);
        if (rc) return rc;
        if (j->d_out) {
            CUtensorMap mxb, mr;
            rc = make_map(&mxb, j->d_X[sg], (uint64_t)dk, (uint64_t)j->rows_per_seg, (uint64_t)j->ld, MCB_KR, 32,
                          CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
            if (rc) return rc;
            rc = make_map(&mr, Rt, (uint64_t)ldr, (uint64_t)CN, (uint64_t)ldr, (uint32_t)CN, MCB_KR);
            if (rc) return rc;
            McBwdParams b;
            b.rows = j->rows_per_seg; b.n_chunks = n_chunks; b.d = j->d; b.ldw = j->ldw; b.G = G;
            rc = (NB == 1) ? launch_bwd<1>(mxb, mr, b, st) : launch_bwd<2>(mxb, mr, b, st);
            if (rc) return rc;
        }
    }
    if (j->d_out) {
        const long long n = (long long)CN * j->ldw;
        mc_ship_kernel<<<(unsigned)((n + 1023) / 1024 > 592 ? 592 : (n + 1023) / 1024), 256, 0, st>>>(G, j->d_out, n);
        if (j->d_flag) mc_flag_kernel<<<1, 32, 0, st>>>(j->d_flag, j->stamp, j->d_beta_flag ? skip : nullptr);
    }
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}
