// One-pass multi-column gradient: both products of  G = X^T . resid(X . W^T)  from ONE shared
// memory copy of each X tile (SURVEY 7.4 hard part 2), in a single launch that also waits for the
// round's parameters, reduces across CTAs in a fixed order and ships the message.
//
//   tile = 128 rows x dk columns, resident in a ring of 64 KB "super-stages" (128 rows x 128 columns,
//   four TMA boxes of 128 x 32 fp32, 32B-atom 128B swizzle = the layout the MN-major tf32 operand of
//   tcgen05.mma requires);
//   forward   P = X . W^T  on the legacy warp-level tensor-core path (mma.sync m16n8k8 tf32): the
//             K-major use of a 32B-atom-swizzled tile is not expressible as a tcgen05 descriptor
//             ("SWIZZLE_128B_BASE32B is invalid for Major-K"), while warps can read any layout; the
//             product is tiny (N = 8..32 columns) and its result lands in registers where the
//             softmax / logistic residual needs it;
//   residual  R = coeff * (softmax(P) - onehot(y))  or  coeff * -y / (exp(P y) + 1), rounded to
//             tf32 and written to shared memory in the K-major SW128 operand layout;
//   backward  G[j, c] += sum_r X[r, j] R[r, c]: tcgen05.mma kind::tf32, A = the same X tile read
//             MN-major (M = 128 feature columns, K = 8 rows per instruction), accumulators for all
//             column chunks in TMEM for the CTA's whole row range; the commit of the last MMA that
//             reads a super-stage releases it to the TMA producer;
//   tail      TMEM -> per-CTA partial in scratch, device-wide barrier, fixed-order reduction,
//             store into the (peer-mapped) destination, release flag.
//
// Handles dk <= 128 * ring slots that fit next to W and R in shared memory (dk <= 256 for 16
// columns, 384 for 8); wider problems use the two-pass tcgen05 kernels of grad_mc_tc.cu.
// No counterpart in the reference (its models are single vectors, naive.py:137-139).
#include <stdlib.h>

#include "grad_dense.cuh"
#include "mc_common.cuh"

namespace eh {

constexpr int MF_BOX_BYTES = 128 * 32 * 4;            // one TMA box: 128 rows x 32 columns
constexpr int MF_STAGE_BYTES = 4 * MF_BOX_BYTES;      // super-stage: 128 rows x 128 columns
constexpr int MF_MAX_STAGES = 3;
constexpr int MF_COMPUTE = 256;                      // 8 compute warps (16 tile rows each)
constexpr int MF_THREADS = 128 + MF_COMPUTE;          // + TMA producer, MMA issuer, TMEM allocator, spare

struct McFusedMaps {
    CUtensorMap x[EH_MAX_SEGS];
};

struct McFusedParams {
    const float* seg_y[EH_MAX_SEGS];
    float coeff[EH_MAX_SEGS];
    const float* W;                  // [CN, ldw]
    float* out;                      // [CN, ldw] destination (local or peer)
    uint32_t* flag;
    const uint32_t* beta_flag;
    uint32_t* err;
    GradScratchHdr* hdr;
    float* partials;                 // [grid][CN * ldw]
    int n_segs, rows_per_seg, tiles_per_seg;
    int d, dk, ldw, n_chunks, C, model, n_stages;
    int w_stride;                    // floats between consecutive class rows of the W image (dk + 4)
    int pf;                          // L2 prefetch distance in ring items (0 = off)
    int n_rbuf;                      // R operand images (2 when shared memory allows: backward of tile t overlaps tile t+1)
    uint32_t stamp, seq;
    unsigned long long timeout_ns;
};

// NBC = 8-column blocks of the model (BN = 8 * NBC columns in the tensor-core tiles)
template <int NBC>
__global__ void __launch_bounds__(MF_THREADS, 1)
mc_fused_kernel(const __grid_constant__ McFusedMaps maps, const McFusedParams p) {
    constexpr int BN = 8 * NBC;
    constexpr int R_BYTES = 4 * BN * 128;             // four K-major sub-tiles [BN x 32 rows]
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* ring = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    unsigned char* Rs = ring + (size_t)p.n_stages * MF_STAGE_BYTES;
    float* Ws = reinterpret_cast<float*>(Rs + p.n_rbuf * R_BYTES);          // [BN][w_stride = dk + 4]
    uint64_t* full = reinterpret_cast<uint64_t*>(Ws + (size_t)BN * p.w_stride);
    uint64_t* empty = full + MF_MAX_STAGES;
    uint64_t* rfull = empty + MF_MAX_STAGES;
    uint64_t* rempty = rfull + 2;
    uint64_t* tdone = rempty + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tdone + 1);
    volatile uint32_t& s_decision = *(tmem_slot + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int G = gridDim.x;
    const int n_tiles = p.n_segs * p.tiles_per_seg;
    const int my_tiles = ((int)blockIdx.x < n_tiles) ? (n_tiles - (int)blockIdx.x + G - 1) / G : 0;
    const int my_items = my_tiles * p.n_chunks;
    constexpr uint32_t TMEM_COLS = 128;

    if (tid == 0) {
        for (int s = 0; s < MF_MAX_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&rfull[s], MF_COMPUTE / 32); mbar_init(&rempty[s], 1); }
        mbar_init(tdone, 1);
        mbar_fence_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // box loads of one (tile, chunk) item into its ring slot
    auto issue = [&](int it, uint64_t pol) {
        const int tl = it / p.n_chunks, ch = it % p.n_chunks;
        const int T = (int)blockIdx.x + tl * G;
        const int sg = T / p.tiles_per_seg, row0 = (T % p.tiles_per_seg) * 128;
        const int st = it % p.n_stages;
        int nb = (p.dk - ch * 128 + 31) / 32;
        if (nb > 4) nb = 4;
        unsigned char* dst = ring + (size_t)st * MF_STAGE_BYTES;
        mbar_arrive_expect_tx(&full[st], (uint32_t)nb * MF_BOX_BYTES);
        for (int bx = 0; bx < nb; ++bx)
            tma_load_2d(dst + bx * MF_BOX_BYTES, &maps.x[sg], ch * 128 + bx * 32, row0, &full[st], pol);
    };

    // L2 prefetch of the boxes of a later item
    auto prefetch = [&](int it) {
        const int tl = it / p.n_chunks, ch = it % p.n_chunks;
        const int T = (int)blockIdx.x + tl * G;
        const int sg = T / p.tiles_per_seg, row0 = (T % p.tiles_per_seg) * 128;
        int nb = (p.dk - ch * 128 + 31) / 32;
        if (nb > 4) nb = 4;
        for (int bx = 0; bx < nb; ++bx) tma_prefetch_2d(&maps.x[sg], ch * 128 + bx * 32, row0);
    };

    uint32_t code = DEC_GO;
    if (warp == 0) {
        // ---------------- TMA producer; prefetches X while the parameters are still in flight
        int issued = 0;
        const uint64_t pol = l2_policy_evict_first();
        if (lane == 0) {
            const int pre = my_items < p.n_stages ? my_items : p.n_stages;
            for (; issued < pre; ++issued) issue(issued, pol);
            for (int k = issued; k < issued + p.pf && k < my_items; ++k) prefetch(k);
            code = wait_for_parameters(p);
            s_decision = code;
        }
        __syncwarp();
        asm volatile("bar.sync 1, %0;" ::"r"(MF_THREADS) : "memory");
        code = s_decision;
        if (lane == 0) {
            if (code == DEC_SKIP) {
                for (int k = 0; k < issued; ++k) mbar_wait(&full[k % p.n_stages], 0);
            } else {
                for (int it = issued; it < my_items; ++it) {
                    if (it + p.pf < my_items && p.pf > 0) prefetch(it + p.pf);
                    mbar_wait_backoff(&empty[it % p.n_stages], ((it / p.n_stages) - 1) & 1);
                    issue(it, pol);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        asm volatile("bar.sync 1, %0;" ::"r"(MF_THREADS) : "memory");
        code = s_decision;
        if (code == DEC_GO && lane == 0) {
            // ---------------- backward MMA issuer
            constexpr uint32_t idesc = make_idesc_tf32(128, BN, /*a_mn_major=*/1, /*b_mn_major=*/0);
            int it = 0;
            for (int tl = 0; tl < my_tiles; ++tl) {
                const int rb = tl % p.n_rbuf;
                mbar_wait(&rfull[rb], (tl / p.n_rbuf) & 1);
                tc_fence_after();
                // descriptors differ only in their 14-bit start-address field: build one, then add (the
                // issuing thread is a single instruction stream -- every saved instruction is MMA rate)
                const uint64_t r_desc = make_kmajor_sw128_desc(smem_u32(Rs + rb * R_BYTES));
                for (int ch = 0; ch < p.n_chunks; ++ch, ++it) {
                    const int st = it % p.n_stages;
                    mbar_wait(&full[st], (it / p.n_stages) & 1);
                    tc_fence_after();
                    // A: X^T, MN-major; 8 rows per MMA = 1024 B, 32-column atoms one box apart
                    const uint64_t a_desc0 = make_mnmajor_sw128_32b_desc(smem_u32(ring + (size_t)st * MF_STAGE_BYTES), MF_BOX_BYTES);
                    const uint32_t d_tmem = tmem_base + ch * BN;
#pragma unroll
                    for (int kk = 0; kk < 16; ++kk)
                        umma_tf32(d_tmem, a_desc0 + (uint64_t)(kk * (1024 >> 4)),
                                  r_desc + (uint64_t)((kk >> 2) * ((BN * 128) >> 4) + (kk & 3) * 2), idesc,
                                  (tl > 0 || kk > 0) ? 1u : 0u);
                    umma_commit(&empty[st]);          // slot reusable once these MMAs have read it
                }
                umma_commit(&rempty[rb]);             // this R image may be overwritten
            }
            umma_commit(tdone);
        }
        __syncwarp();
    } else if (warp >= 4) {
        asm volatile("bar.sync 1, %0;" ::"r"(MF_THREADS) : "memory");
        code = s_decision;
        if (code == DEC_GO) {
            const int cw = warp - 4;                   // compute warp 0..7 owns tile rows cw*16 .. cw*16+15
            const int ct = tid - 128;                  // 0..255
            const int g = lane >> 2, t = lane & 3;
            // ---------------- parameters -> shared memory, class-major rows of dk + 4 floats, tf32-rounded
            for (int i = ct; i < BN * p.dk; i += MF_COMPUTE) {
                const int n = i / p.dk, c = i % p.dk;
                Ws[(size_t)n * p.w_stride + c] = round_tf32(__ldcg(p.W + (size_t)n * p.ldw + c));
            }
            asm volatile("bar.sync 2, %0;" ::"r"(MF_COMPUTE) : "memory");

            // per-thread constants of the fragment addressing (shared-space byte addresses)
            const uint32_t ring_u = smem_u32(ring), ws_u = smem_u32(Ws), rs_u = smem_u32(Rs);
            // ldmatrix row addresses.  A: lane L -> tile row cw*16 + (L & 15), 16-byte half L >> 4 of the
            // k-step's 32 bytes (the 32B-chunk XOR of the swizzle folded in).  B: lane L -> class
            // (L & 7) + 8 * (L >> 4), 16-byte half (L >> 3) & 1; rows are dk*4 + 16 bytes apart (conflict-free).
            const int ra = cw * 16 + (lane & 15);
            const uint32_t xoff = (uint32_t)ra * 128u + ((uint32_t)(ra & 3) << 5) + (uint32_t)(lane >> 4) * 16u;
            const uint32_t woff = ws_u + (uint32_t)((lane & 7) + 8 * (lane >> 4)) * (uint32_t)p.w_stride * 4u +
                                  (uint32_t)((lane >> 3) & 1) * 16u;
            const int r0 = cw * 16 + g;                // accumulator rows of this thread: r0 and r0 + 8
            // R^T image: sub-tile (32 rows) cw/2, row inside it rr = (cw&1)*16 + h*8 + g
            const int rr0 = (cw & 1) * 16 + g;
            const uint32_t roff0 = rs_u + (uint32_t)(cw >> 1) * (BN * 128) + (uint32_t)(rr0 & 3) * 4u;

            int it = 0;
            int sg = (int)blockIdx.x / p.tiles_per_seg, tis = (int)blockIdx.x % p.tiles_per_seg;
            for (int tl = 0; tl < my_tiles; ++tl) {
                const int row0 = tis * 128;
                const float cf = p.coeff[sg];
                float yv[2];
                bool valid[2];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int r = row0 + r0 + h * 8;
                    valid[h] = r < p.rows_per_seg;
                    yv[h] = valid[h] ? __ldg(p.seg_y[sg] + r) : 0.f;
                }
                float acc[NBC][4];
#pragma unroll
                for (int nb = 0; nb < NBC; ++nb)
#pragma unroll
                    for (int e = 0; e < 4; ++e) acc[nb][e] = 0.f;

                // ---------------- forward: warp-level tf32 MMAs straight from the swizzled tile
                for (int ch = 0; ch < p.n_chunks; ++ch, ++it) {
                    const int st = it % p.n_stages;
                    mbar_wait(&full[st], (it / p.n_stages) & 1);
                    const uint32_t stage = ring_u + (uint32_t)st * MF_STAGE_BYTES + xoff;
                    int nbx = (p.dk - ch * 128 + 31) / 32;
                    if (nbx > 4) nbx = 4;
                    uint32_t wp = woff + (uint32_t)(ch * 128) * 4u;
#pragma unroll
                    for (int bx = 0; bx < 4; ++bx) {
                        if (bx < nbx) {
                            uint32_t a[4][4], bfr[4][NBC / 2][4];
#pragma unroll
                            for (int kk = 0; kk < 4; ++kk) {
                                mc_ldsm_x4((stage + bx * MF_BOX_BYTES) ^ ((uint32_t)kk << 5), a[kk]);
#pragma unroll
                                for (int np = 0; np < NBC / 2; ++np)      // two 8-class blocks per ldmatrix
                                    mc_ldsm_x4(wp + (uint32_t)kk * 32u + (uint32_t)np * 16u * (uint32_t)p.w_stride * 4u, bfr[kk][np]);
                            }
#pragma unroll
                            for (int kk = 0; kk < 4; ++kk)
#pragma unroll
                                for (int nb = 0; nb < NBC; ++nb)
                                    mc_mma_tf32(acc[nb], a[kk][0], a[kk][1], a[kk][2], a[kk][3], bfr[kk][nb >> 1][(nb & 1) * 2],
                                                    bfr[kk][nb >> 1][(nb & 1) * 2 + 1]);
                        }
                        wp += 32 * 4;
                    }
                }

                // ---------------- residual per row (a row's columns live in the 4 lanes of a quad)
                float res[2][NBC][2];
#pragma unroll
                for (int h = 0; h < 2; ++h) mc_quad_row_residual<NBC>(acc, h, t, p.C, p.model, yv[h], cf, valid[h], res[h]);
                // ---------------- R^T into the K-major SW128 operand image
                const int rb = tl % p.n_rbuf;
                if (tl >= p.n_rbuf) mbar_wait(&rempty[rb], ((tl / p.n_rbuf) - 1) & 1);
                const uint32_t roff = roff0 + (uint32_t)rb * R_BYTES;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int c16 = (rr0 + h * 8) >> 2;                   // 16-byte chunk of the row inside the sub-tile
#pragma unroll
                    for (int nb = 0; nb < NBC; ++nb)
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            const int n = nb * 8 + 2 * t + j;
                            mc_sts32(roff + (uint32_t)n * 128u + (uint32_t)((c16 ^ (n & 7)) << 4), res[h][nb][j]);
                        }
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(&rfull[rb]);
                tis += G;
                while (tis >= p.tiles_per_seg) { tis -= p.tiles_per_seg; ++sg; }
            }

            // ---------------- accumulators: TMEM -> this CTA's partial [CN][ldw]
            // warp w may touch TMEM lanes 32*(w%4)..+31; the two warps of a lane quarter split the columns
            float* mine = p.partials + (size_t)blockIdx.x * ((size_t)BN * p.ldw);
            if (my_tiles > 0) {
                mbar_wait(tdone, 0);
                tc_fence_after();
            }
            const int q = cw & 3, half = cw >> 2;
            for (int ch = 0; ch < p.n_chunks; ++ch) {
                const int j = ch * 128 + q * 32 + lane;
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + ch * BN;
#pragma unroll
                for (int cbi = 0; cbi < BN / 16; ++cbi) {
                    const int cb = half * (BN / 16) + cbi;
                    float v[8];
                    if (my_tiles > 0) {
                        uint32_t r[8];
                        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                                     : "r"(taddr + cb * 8)
                                     : "memory");
                        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                        for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[e]);
                    } else {
#pragma unroll
                        for (int e = 0; e < 8; ++e) v[e] = 0.f;
                    }
                    if (j < p.ldw) {
#pragma unroll
                        for (int e = 0; e < 8; ++e) __stcg(mine + (size_t)(cb * 8 + e) * p.ldw + j, (j < p.d) ? v[e] : 0.f);
                    }
                }
            }
            tc_fence_before();
        }
    } else {
        asm volatile("bar.sync 1, %0;" ::"r"(MF_THREADS) : "memory");
        code = s_decision;
    }

    // ---------------- common tail
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
    if (code == DEC_SKIP) return;

    uint32_t gen = 0;
    grid_barrier(&p.hdr->barrier, gen);
    // fixed-order reduction over the CTAs' partials, straight into the destination
    const int n4 = (BN * p.ldw) >> 2;
    for (int i4 = (int)blockIdx.x * MF_THREADS + tid; i4 < n4; i4 += G * MF_THREADS) {
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4* src = reinterpret_cast<const float4*>(p.partials) + i4;
        if ((i4 * 4) % p.ldw < p.n_chunks * 128) {         // columns past the last chunk were never written
#pragma unroll 8
            for (int b = 0; b < G; ++b) {
                const float4 v = __ldcg(src + (size_t)b * n4);
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
        }
        reinterpret_cast<float4*>(p.out)[i4] = s;
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

// ------------------------------------------------------------------------------------ host
static size_t mc_fused_smem(int dk, int BN, int stages, int n_rbuf = 1) {
    return 1024 + (size_t)stages * MF_STAGE_BYTES + (size_t)n_rbuf * 4 * BN * 128 + (size_t)BN * (dk + 4) * sizeof(float) +
           (2 * MF_MAX_STAGES + 5) * sizeof(uint64_t) + 16;
}

// ring slots for this shape, 0 when the one-pass kernel cannot hold a whole tile
int mc_fused_stages(int dk, int CN) {
    const int n_chunks = (dk + 127) / 128;
    for (int s = MF_MAX_STAGES; s >= n_chunks && s >= 1; --s)
        if (mc_fused_smem(dk, CN, s) <= (size_t)227 * 1024) return s;
    return 0;
}

size_t mc_fused_scratch_bytes(int d, int CN) {
    const size_t ldw = (size_t)(d + 31) / 32 * 32;
    int sm = dev_info().sm_count;
    if (sm <= 0) sm = 160;
    return sizeof(GradScratchHdr) + (size_t)sm * CN * ldw * sizeof(float);
}

template <int NBC>
static int launch_fused(const McFusedMaps& maps, McFusedParams& p, int grid, cudaStream_t st) {
    const size_t smem = mc_fused_smem(p.dk, 8 * NBC, p.n_stages, p.n_rbuf);
    auto kern = mc_fused_kernel<NBC>;
    EH_SET_MAX_SMEM((kern), 227 * 1024);
    void* args[] = {(void*)&maps, (void*)&p};
    return launch_grid_sync((const void*)kern, dim3((unsigned)grid), dim3(MF_THREADS), args, smem, st);
}

// `fused_scratch` points at a zero-initialised GradScratchHdr followed by the per-CTA partials.
int mc_fused_launch(const eh_mc_job* j, void* fused_scratch, size_t fused_bytes, cudaStream_t st) {
    const int CN = j->C <= 16 ? 16 : 32;
    const int dk = (j->d + 31) / 32 * 32;
    McFusedMaps maps;
    McFusedParams p;
    memset(&p, 0, sizeof(p));
    p.n_stages = mc_fused_stages(dk, CN);
    EH_REQUIRE(p.n_stages > 0, "grad_mc: one-pass kernel cannot hold %d columns", dk);
    // a second R image (backward of tile t overlapping tile t+1) measured slower (0.69 vs 0.76 at d = 128): opt-in
    p.n_rbuf = (mc_fused_smem(dk, CN, p.n_stages, 2) <= (size_t)227 * 1024 && getenv("EH_MC_RBUF2") != nullptr) ? 2 : 1;
    for (int sg = 0; sg < j->n_segs; ++sg) {
        EH_REQUIRE(j->d_X[sg] && j->d_y[sg], "grad_mc: null segment %d", sg);
        EH_REQUIRE(((uintptr_t)j->d_X[sg] % 16) == 0, "grad_mc: X must be 16-byte aligned");
        int rc = make_map(&maps.x[sg], j->d_X[sg], (uint64_t)dk, (uint64_t)j->rows_per_seg, (uint64_t)j->ld, 128, 32,
                          CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
        if (rc) return rc;
        p.seg_y[sg] = j->d_y[sg];
        p.coeff[sg] = j->coeff[sg];
    }
    for (int sg = j->n_segs; sg < EH_MAX_SEGS; ++sg) maps.x[sg] = maps.x[0];
    p.W = j->d_W; p.out = j->d_out; p.flag = j->d_flag; p.beta_flag = j->d_beta_flag; p.err = j->d_err;
    p.n_segs = j->n_segs; p.rows_per_seg = j->rows_per_seg; p.tiles_per_seg = (j->rows_per_seg + 127) / 128;
    p.d = j->d; p.dk = dk; p.ldw = j->ldw; p.n_chunks = (dk + 127) / 128; p.C = j->C; p.model = j->model;
    p.w_stride = dk + 4;
    // L2 prefetch two tiles ahead pays only when a tile spans two ring slots (d = 256: 0.84 -> 0.79 ms)
    p.pf = (getenv("EH_MC_PF") ? atoi(getenv("EH_MC_PF")) : (p.n_chunks >= 2 ? 2 : 0)) * p.n_chunks;
    p.stamp = j->stamp; p.seq = next_launch_seq();
    p.timeout_ns = j->timeout_ns ? j->timeout_ns : 30ull * 1000000000ull;
    const long long n_tiles = (long long)p.n_segs * p.tiles_per_seg;
    long long grid = worker_sms();
    if (grid > n_tiles) grid = n_tiles;
    const size_t need = sizeof(GradScratchHdr) + (size_t)grid * CN * p.ldw * sizeof(float);
    EH_REQUIRE(fused_bytes >= need, "grad_mc: scratch too small for the one-pass kernel (%zu < %zu)", fused_bytes, need);
    p.hdr = reinterpret_cast<GradScratchHdr*>(fused_scratch);
    p.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(fused_scratch) + sizeof(GradScratchHdr));
    return CN == 16 ? launch_fused<2>(maps, p, (int)grid, st) : launch_fused<4>(maps, p, (int)grid, st);
}

}  // namespace eh
