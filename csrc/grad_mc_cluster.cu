// One-pass multi-column gradient for wide models: the columns of a row tile are split over a
// thread-block cluster (CS = 1, 2, 4 or 8 CTAs), so that every CTA keeps its column slice of the
// tile in shared memory between the forward and the backward product and X is read once.
//
//   tile       32 rows; CTA `rank` of the cluster owns `npc` chunks of 128 columns; a chunk of a tile
//              is one 16 KB ring stage (four TMA boxes of 32 rows x 32 columns, 32B-atom swizzle);
//   forward    four pairs of compute warps work on four tiles at a time: mma.sync m16n8k8 tf32 over
//              the CTA's column slice -> partial margins in registers;
//   exchange   (CS > 1) the partial margins of row r go to the CTA that owns r (32/CS rows each) with
//              st.async (data + mbarrier transaction bytes in one remote store); a reducer warp of the
//              owner sums the CS partials, evaluates the softmax / logistic residual and st.async-
//              broadcasts it into every CTA's R operand image (K-major, SW128);
//              (CS = 1) the compute warps evaluate the residual themselves;
//   backward   tcgen05.mma kind::tf32, A = the resident X stage read MN-major, B = R, accumulators of
//              all owned chunks in TMEM over the CTA's whole tile range; its commit frees the stage;
//   tail       TMEM -> per-cluster partial (each rank its own columns), device-wide barrier,
//              fixed-order reduction over the clusters, store into the destination, release flag.
//
// Slot discipline (4 tiles in flight, slot = tile & 3): a compute pair sends / writes tile t only
// after the local backward of tile t-4 has retired (rempty), which itself needed the residuals of
// t-4 from every owner -- so an owner's exchange buffer and every CTA's R image of that slot are
// free again by construction, without credits travelling back.
// No counterpart in the reference (its models are single vectors, naive.py:137-139).
#include <stdlib.h>

#include "cluster_common.cuh"
#include "grad_dense.cuh"
#include "mc_common.cuh"

namespace eh {

constexpr int MC2_BOX = 32 * 32 * 4;          // TMA box: 32 rows x 32 columns fp32
constexpr int MC2_STAGE = 4 * MC2_BOX;        // ring stage: 32 rows x 128 columns
constexpr int MC2_MAX_STAGES = 12;
constexpr int MC2_SLOTS = 4;
constexpr int MC2_COMPUTE = 256;              // 8 compute warps = 4 pairs (one 16-row block per warp)
constexpr int MC2_THREADS = 128 + MC2_COMPUTE;

struct Mc2Maps {
    CUtensorMap x[EH_MAX_SEGS];
};

struct Mc2Params {
    const float* seg_y[EH_MAX_SEGS];
    float coeff[EH_MAX_SEGS];
    const float* W;
    float* out;
    uint32_t* flag;
    const uint32_t* beta_flag;
    uint32_t* err;
    GradScratchHdr* hdr;
    float* partials;                 // [clusters][CN * ldw]
    int n_segs, rows_per_seg, tiles_per_seg;
    int d, dk, ldw, n_chunks, npc, C, model, n_stages, CS;
    int npt;                         // compute-warp pairs that share one tile (1, 2 or 4): each takes every npt-th chunk
    int w_stride;
    uint32_t stamp, seq;
    unsigned long long timeout_ns;
};

template <int NBC>
__global__ void __launch_bounds__(MC2_THREADS, 1)
mc_cluster_kernel(const __grid_constant__ Mc2Maps maps, const Mc2Params p) {
    constexpr int BN = 8 * NBC;
    constexpr int R_SLOT = BN * 128;                  // [BN x 32 rows] K-major SW128
    const int X_SLOT = p.npt * 32 * BN * 4;           // [src = rank*npt + sub][rows_own][BN] partial margins
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* ring = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    unsigned char* Rs = ring + (size_t)p.n_stages * MC2_STAGE;
    unsigned char* Xb = Rs + MC2_SLOTS * R_SLOT;
    float* Ws = reinterpret_cast<float*>(Xb + MC2_SLOTS * X_SLOT);          // [npc*64][w_stride]
    uint64_t* full = reinterpret_cast<uint64_t*>(Ws + (size_t)(p.npc * 64) * p.w_stride);
    uint64_t* empty = full + MC2_MAX_STAGES;
    uint64_t* rfull = empty + MC2_MAX_STAGES;
    uint64_t* rempty = rfull + MC2_SLOTS;
    uint64_t* xfull = rempty + MC2_SLOTS;
    uint64_t* tdone = xfull + MC2_SLOTS;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tdone + 1);
    volatile uint32_t& s_decision = *(tmem_slot + 1);
    volatile uint32_t* s_issued = tmem_slot + 2;       // ring items the producer has armed and issued so far

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int CS = p.CS, NPT = p.npt;
    const bool direct = (CS == 1 && NPT == 1);         // residual in the compute warps, no exchange
    const int rank = (CS > 1) ? (int)cluster_ctarank() : 0;
    const int cluster_id = blockIdx.x / CS, n_clusters = gridDim.x / CS;
    const int n_tiles = p.n_segs * p.tiles_per_seg;
    const int my_tiles = (cluster_id < n_tiles) ? (n_tiles - cluster_id + n_clusters - 1) / n_clusters : 0;
    const int chunk0 = rank * p.npc;
    int mych = p.n_chunks - chunk0;
    mych = mych < 0 ? 0 : (mych > p.npc ? p.npc : mych);
    const int my_items = my_tiles * mych;
    const uint32_t tmem_cols = (p.npc * BN <= 32) ? 32u : (p.npc * BN <= 64) ? 64u : (p.npc * BN <= 128) ? 128u
                                                                           : (p.npc * BN <= 256) ? 256u : 512u;

    if (tid == 0) {
        for (int s = 0; s < MC2_MAX_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int s = 0; s < MC2_SLOTS; ++s) {
            mbar_init(&rfull[s], direct ? 2 : 1);      // exchange: transaction bytes; direct: the two warps of a pair
            mbar_init(&rempty[s], 1);
            mbar_init(&xfull[s], 1);
        }
        mbar_init(tdone, 1);
        *s_issued = 0;
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (CS > 1) cluster_sync_all();                    // every CTA's barriers exist before a peer targets them

    auto tile_of = [&](int tl, int& sg, int& row0) {
        const int T = cluster_id + tl * n_clusters;
        sg = T / p.tiles_per_seg;
        row0 = (T % p.tiles_per_seg) * 32;
    };
    auto issue = [&](int it, uint64_t pol) {
        const int tl = it / mych, ch = it % mych;
        int sg, row0;
        tile_of(tl, sg, row0);
        const int st = it % p.n_stages;
        const int col0 = (chunk0 + ch) * 128;
        int nb = (p.dk - col0 + 31) / 32;
        if (nb > 4) nb = 4;
        unsigned char* dst = ring + (size_t)st * MC2_STAGE;
        mbar_arrive_expect_tx(&full[st], (uint32_t)nb * MC2_BOX);
        for (int bx = 0; bx < nb; ++bx) tma_load_2d(dst + bx * MC2_BOX, &maps.x[sg], col0 + bx * 32, row0, &full[st], pol);
    };

    uint32_t code = DEC_GO;
    if (warp == 0) {
        // ---------------- TMA producer; prefetches X while the parameters are still in flight
        int issued = 0;
        const uint64_t pol = l2_policy_evict_first();
        if (lane == 0) {
            const int pre = my_items < p.n_stages ? my_items : p.n_stages;
            for (; issued < pre; ++issued) issue(issued, pol);
            __threadfence_block();
            *s_issued = (uint32_t)issued;
            code = wait_for_parameters(p);
            s_decision = code;
        }
        __syncwarp();
        asm volatile("bar.sync 1, %0;" ::"r"(MC2_THREADS) : "memory");
        code = s_decision;
        if (lane == 0) {
            if (code == DEC_SKIP) {
                for (int k = 0; k < issued; ++k) mbar_wait(&full[k % p.n_stages], 0);
            } else {
                for (int it = issued; it < my_items; ++it) {
                    mbar_wait_backoff(&empty[it % p.n_stages], ((it / p.n_stages) - 1) & 1);
                    issue(it, pol);
                    __threadfence_block();
                    *s_issued = (uint32_t)(it + 1);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        asm volatile("bar.sync 1, %0;" ::"r"(MC2_THREADS) : "memory");
        code = s_decision;
        if (code == DEC_GO && lane == 0) {
            // ---------------- backward MMA issuer (tiles in order)
            constexpr uint32_t idesc = make_idesc_tf32(128, BN, /*a_mn_major=*/1, /*b_mn_major=*/0);
            int it = 0;
            for (int tl = 0; tl < my_tiles; ++tl) {
                const int slot = tl & (MC2_SLOTS - 1);
                if (!direct) mbar_arrive_expect_tx(&rfull[slot], 32 * BN * 4);
                mbar_wait(&rfull[slot], (tl >> 2) & 1);
                if (!direct) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                tc_fence_after();
                const uint64_t r_desc = make_kmajor_sw128_desc(smem_u32(Rs + slot * R_SLOT));
                for (int ch = 0; ch < mych; ++ch, ++it) {
                    const int st = it % p.n_stages;
                    mbar_wait(&full[st], (it / p.n_stages) & 1);
                    tc_fence_after();
                    const uint64_t a_desc0 = make_mnmajor_sw128_32b_desc(smem_u32(ring + (size_t)st * MC2_STAGE), MC2_BOX);
                    const uint32_t d_tmem = tmem_base + ch * BN;
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk)
                        umma_tf32(d_tmem, a_desc0 + (uint64_t)(kk * (1024 >> 4)), r_desc + (uint64_t)(kk * 2), idesc,
                                  (tl > 0 || kk > 0) ? 1u : 0u);
                    umma_commit(&empty[st]);
                }
                umma_commit(&rempty[slot]);
            }
            umma_commit(tdone);
        }
        __syncwarp();
    }
    else if (warp == 2 || warp == 3) {
        asm volatile("bar.sync 1, %0;" ::"r"(MC2_THREADS) : "memory");
        code = s_decision;
        if (code == DEC_GO && !direct) {
            // ---------------- reducer warps: sum the CS * NPT partial margins of the rows this CTA
            // owns, evaluate the residual, broadcast it into every CTA's R image
            const int rows_own = 32 / CS;
            const int items = rows_own * (BN / 2);
            const uint32_t xb_u = smem_u32(Xb), rs_u = smem_u32(Rs), rfull_u = smem_u32(rfull);
            for (int tl = warp - 2; tl < my_tiles; tl += 2) {
                const int slot = tl & (MC2_SLOTS - 1);
                int sg, row0;
                tile_of(tl, sg, row0);
                const float cf = p.coeff[sg];
                if (lane == 0) mbar_arrive_expect_tx(&xfull[slot], (uint32_t)X_SLOT);
                __syncwarp();
                mbar_wait(&xfull[slot], (tl >> 2) & 1);
                for (int i = lane; i < items; i += 32) {
                    const int row_local = i / (BN / 2), cp = i % (BN / 2);
                    const int rt = rank * rows_own + row_local;            // row inside the tile
                    const int grow = row0 + rt;
                    const bool valid = grow < p.rows_per_seg;
                    const float yv = valid ? __ldg(p.seg_y[sg] + grow) : 0.f;
                    float v0 = 0.f, v1 = 0.f;
                    for (int src = 0; src < CS * NPT; ++src) {
                        const float2 t2 = mc_lds_f2(xb_u + (uint32_t)slot * X_SLOT +
                                                     (uint32_t)(((src * rows_own + row_local) * BN + 2 * cp) * 4));
                        v0 += t2.x; v1 += t2.y;
                    }
                    float r0v, r1v;
                    const int n0 = 2 * cp;
                    if (p.model == 2) {
                        float m = fmaxf(n0 < p.C ? v0 : -3.0e38f, n0 + 1 < p.C ? v1 : -3.0e38f);
#pragma unroll
                        for (int o = 1; o < BN / 2; o <<= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
                        const float e0 = n0 < p.C ? __expf(v0 - m) : 0.f, e1 = n0 + 1 < p.C ? __expf(v1 - m) : 0.f;
                        float sum = e0 + e1;
#pragma unroll
                        for (int o = 1; o < BN / 2; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                        const float inv = __fdividef(1.f, sum);
                        const int yc = (int)yv;
                        r0v = cf * (e0 * inv - (n0 == yc ? 1.f : 0.f));
                        r1v = cf * (e1 * inv - (n0 + 1 == yc ? 1.f : 0.f));
                    } else {
                        r0v = n0 < p.C ? mc_logistic(v0, yv, cf) : 0.f;
                        r1v = n0 + 1 < p.C ? mc_logistic(v1, yv, cf) : 0.f;
                    }
                    r0v = valid ? round_tf32(r0v) : 0.f;
                    r1v = valid ? round_tf32(r1v) : 0.f;
                    const uint32_t base = rs_u + (uint32_t)slot * R_SLOT + (uint32_t)(rt & 3) * 4u;
                    const uint32_t o0 = base + (uint32_t)n0 * 128u + (uint32_t)(((rt >> 2) ^ (n0 & 7)) << 4);
                    const uint32_t o1 = base + (uint32_t)(n0 + 1) * 128u + (uint32_t)(((rt >> 2) ^ ((n0 + 1) & 7)) << 4);
                    for (int dst = 0; dst < CS; ++dst) {
                        const uint32_t bar = mapa_shared(rfull_u + (uint32_t)slot * 8u, (uint32_t)dst);
                        st_async_f32(mapa_shared(o0, (uint32_t)dst), r0v, bar);
                        st_async_f32(mapa_shared(o1, (uint32_t)dst), r1v, bar);
                    }
                }
            }
        }
    } else {
        asm volatile("bar.sync 1, %0;" ::"r"(MC2_THREADS) : "memory");
        code = s_decision;
        if (code == DEC_GO) {
            const int cw = warp - 4;                   // compute warp 0..7
            const int pr = cw >> 1, mb = cw & 1;       // pair and 16-row block
            const int grp = pr / NPT, sub = pr % NPT;  // NPT pairs share a tile; pair `sub` takes chunks sub, sub+NPT, ...
            const int n_grp = MC2_SLOTS / NPT;
            const int ct = tid - 128;
            const int g = lane >> 2, t = lane & 3;
            // ---------------- this CTA's slice of the parameters -> shared memory as column pairs
            for (int i = ct; i < (mych * 64) * BN; i += MC2_COMPUTE) {
                const int n = i % BN, cp = i / BN;
                const int c = chunk0 * 128 + 2 * cp;
                float2 w;
                w.x = (c < p.ldw) ? round_tf32(__ldcg(p.W + (size_t)n * p.ldw + c)) : 0.f;
                w.y = (c + 1 < p.ldw) ? round_tf32(__ldcg(p.W + (size_t)n * p.ldw + c + 1)) : 0.f;
                *reinterpret_cast<float2*>(Ws + (size_t)cp * p.w_stride + 2 * n) = w;
            }
            asm volatile("bar.sync 2, %0;" ::"r"(MC2_COMPUTE) : "memory");

            const uint32_t ring_u = smem_u32(ring), ws_u = smem_u32(Ws), rs_u = smem_u32(Rs), xb_u = smem_u32(Xb);
            const uint32_t xfull_u = smem_u32(xfull);
            const int r0 = mb * 16 + g;                // tile rows r0 and r0 + 8
            const uint32_t xoff = (uint32_t)r0 * 128u + (uint32_t)t * 8u + ((uint32_t)(r0 & 3) << 5);
            const uint32_t woff = ws_u + (uint32_t)t * (uint32_t)p.w_stride * 4u + (uint32_t)g * 8u;
            const uint32_t wstep = (uint32_t)p.w_stride * 16u;
            const int rows_own = 32 / CS;

            for (int tl = grp; tl < my_tiles; tl += n_grp) {
                const int slot = tl & (MC2_SLOTS - 1);
                int sg, row0;
                tile_of(tl, sg, row0);
                float acc[NBC][4];
#pragma unroll
                for (int nb = 0; nb < NBC; ++nb)
#pragma unroll
                    for (int e = 0; e < 4; ++e) acc[nb][e] = 0.f;
                // ---------------- forward over this CTA's column slice
                for (int ch = sub; ch < mych; ch += NPT) {
                    const int it = tl * mych + ch;
                    const int st = it % p.n_stages;
                    // the pairs run several tiles ahead of one another: a parity wait is only exact once
                    // the producer has armed this very use of the stage
                    while (*s_issued <= (uint32_t)it) {
                    }
                    mbar_wait(&full[st], (it / p.n_stages) & 1);
                    const uint32_t stage = ring_u + (uint32_t)st * MC2_STAGE + xoff;
                    int nbx = (p.dk - (chunk0 + ch) * 128 + 31) / 32;
                    if (nbx > 4) nbx = 4;
                    uint32_t wp = woff + (uint32_t)(ch * 16) * wstep;
#pragma unroll
                    for (int bx = 0; bx < 4; ++bx) {
                        if (bx < nbx) {
                            uint2 a[4][2], bfr[4][NBC];
#pragma unroll
                            for (int kk = 0; kk < 4; ++kk) {
                                const uint32_t xa = (stage + bx * MC2_BOX) ^ ((uint32_t)kk << 5);
                                a[kk][0] = mc_lds64(xa);
                                a[kk][1] = mc_lds64(xa + 8 * 128);
#pragma unroll
                                for (int nb = 0; nb < NBC; ++nb) bfr[kk][nb] = mc_lds64(wp + (uint32_t)kk * wstep + nb * 64);
                            }
#pragma unroll
                            for (int kk = 0; kk < 4; ++kk)
#pragma unroll
                                for (int nb = 0; nb < NBC; ++nb)
                                    mc_mma_tf32(acc[nb], a[kk][0].x, a[kk][1].x, a[kk][0].y, a[kk][1].y, bfr[kk][nb].x, bfr[kk][nb].y);
                        }
                        wp += 4 * wstep;
                    }
                }
                // the slot's previous tile (t-4) must have left the backward product
                if (tl >= MC2_SLOTS) mbar_wait(&rempty[slot], ((tl >> 2) - 1) & 1);
                if (!direct) {
                    // ---------------- partial margins -> owners of the rows
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int r = r0 + 8 * h;
                        const int owner = r / rows_own, row_local = r % rows_own;
                        const uint32_t bar = mapa_shared(xfull_u + (uint32_t)slot * 8u, (uint32_t)owner);
#pragma unroll
                        for (int nb = 0; nb < NBC; ++nb) {
                            const uint32_t off = xb_u + (uint32_t)slot * X_SLOT +
                                                 (uint32_t)((((rank * NPT + sub) * rows_own + row_local) * BN + nb * 8 + 2 * t) * 4);
                            st_async_f32x2(mapa_shared(off, (uint32_t)owner), acc[nb][2 * h], acc[nb][2 * h + 1], bar);
                        }
                    }
                } else {
                    // ---------------- single CTA: residual in registers, R image written locally
                    const float cf = p.coeff[sg];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int rt = r0 + 8 * h;
                        const int grow = row0 + rt;
                        const bool valid = grow < p.rows_per_seg;
                        const float yv = valid ? __ldg(p.seg_y[sg] + grow) : 0.f;
                        float res[NBC][2];
                        mc_quad_row_residual<NBC>(acc, h, t, p.C, p.model, yv, cf, valid, res);
                        const uint32_t base = rs_u + (uint32_t)slot * R_SLOT + (uint32_t)(rt & 3) * 4u;
#pragma unroll
                        for (int nb = 0; nb < NBC; ++nb)
#pragma unroll
                            for (int j = 0; j < 2; ++j) {
                                const int n = nb * 8 + 2 * t + j;
                                mc_sts32(base + (uint32_t)n * 128u + (uint32_t)(((rt >> 2) ^ (n & 7)) << 4),
                                          res[nb][j]);
                            }
                    }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&rfull[slot]);
                }
            }

            // ---------------- accumulators: TMEM -> this cluster's partial [CN][ldw], own columns only
            float* mine = p.partials + (size_t)cluster_id * ((size_t)BN * p.ldw);
            if (my_tiles > 0) {
                mbar_wait(tdone, 0);
                tc_fence_after();
            }
            const int q = cw & 3, half = cw >> 2;
            for (int ch = 0; ch < mych; ++ch) {
                const int j = (chunk0 + ch) * 128 + q * 32 + lane;
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
    }

    // ---------------- common tail
    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, tmem_cols);
    }
    if (code == DEC_SKIP) return;

    uint32_t gen = 0;
    grid_barrier(&p.hdr->barrier, gen);
    const int G = gridDim.x;
    const int n4 = (BN * p.ldw) >> 2;
    for (int i4 = (int)blockIdx.x * MC2_THREADS + tid; i4 < n4; i4 += G * MC2_THREADS) {
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4* src = reinterpret_cast<const float4*>(p.partials) + i4;
        if ((i4 * 4) % p.ldw < p.n_chunks * 128) {
#pragma unroll 8
            for (int b = 0; b < n_clusters; ++b) {
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
// Pairs sharing one tile.  Splitting a tile's chunks over several pairs shortens the forward pass but
// puts every tile through the exchange; measured slower (0.35-0.45 vs 0.55 of the one-pass roofline at
// d = 512..1024), so one pair per tile is the default; EH_MC_NPT=2|4 keeps the variant reachable.
static int mc2_npt(int npc) {
    const char* e = getenv("EH_MC_NPT");
    const int want = e ? atoi(e) : 1;
    return (want >= 4 && npc >= 4) ? 4 : (want >= 2 && npc >= 2) ? 2 : 1;
}

static size_t mc2_smem(int BN, int npc, int stages) {
    const size_t w_stride = 2 * BN + 16;
    return 1024 + (size_t)stages * MC2_STAGE + (size_t)MC2_SLOTS * BN * 128 + (size_t)MC2_SLOTS * mc2_npt(npc) * 32 * BN * 4 +
           (size_t)npc * 64 * w_stride * sizeof(float) + (2 * MC2_MAX_STAGES + 3 * MC2_SLOTS + 1) * sizeof(uint64_t) + 32;
}

// smallest cluster whose CTAs hold their column slice of a tile at least twice over (so the ring
// always has a tile loading while another is in use); returns false when no cluster of <= 8 fits
bool mc_cluster_plan(int dk, int CN, int* cs_out, int* npc_out, int* stages_out) {
    const int n_chunks = (dk + 127) / 128;
    const char* force = getenv("EH_MC_CS");
    for (int CS = 1; CS <= 8; CS *= 2) {
        if (force && atoi(force) != CS) continue;
        const int npc = (n_chunks + CS - 1) / CS;
        if (npc * CN > 512) continue;
        int stages = 0;
        for (int s = MC2_MAX_STAGES; s >= 2; --s)
            if (mc2_smem(CN, npc, s) <= (size_t)227 * 1024) { stages = s; break; }
        if (stages >= 2 * npc || (CS == 8 && stages > npc)) {
            if (cs_out) *cs_out = CS;
            if (npc_out) *npc_out = npc;
            if (stages_out) *stages_out = stages;
            return true;
        }
    }
    return false;
}

template <int NBC>
static int launch_mc2(const Mc2Maps& maps, Mc2Params& p, void* fused_scratch, size_t fused_bytes, cudaStream_t st) {
    const int CN = 8 * NBC;
    const size_t smem = mc2_smem(CN, p.npc, p.n_stages);
    auto kern = mc_cluster_kernel<NBC>;
    EH_SET_MAX_SMEM((kern), 227 * 1024);
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cudaLaunchAttribute attrs[1];
    attrs[0].id = cudaLaunchAttributeClusterDimension;
    attrs[0].val.clusterDim.x = p.CS;
    attrs[0].val.clusterDim.y = 1;
    attrs[0].val.clusterDim.z = 1;
    cfg.attrs = attrs;
    cfg.numAttrs = 1;
    cfg.blockDim = dim3(MC2_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cfg.gridDim = dim3(worker_sms() / p.CS * p.CS);
    int max_clusters = 0;
    EH_CUDA(cudaOccupancyMaxActiveClusters(&max_clusters, kern, &cfg));
    EH_REQUIRE(max_clusters >= 1, "grad_mc: no cluster of %d CTAs fits on this device", p.CS);
    long long n_clusters = max_clusters;
    const long long n_tiles = (long long)p.n_segs * p.tiles_per_seg;
    if (n_clusters > n_tiles) n_clusters = n_tiles;
    cfg.gridDim = dim3((unsigned)(n_clusters * p.CS));
    const size_t need = sizeof(GradScratchHdr) + (size_t)n_clusters * CN * p.ldw * sizeof(float);
    EH_REQUIRE(fused_bytes >= need, "grad_mc: scratch too small for the one-pass kernel (%zu < %zu)", fused_bytes, need);
    p.hdr = reinterpret_cast<GradScratchHdr*>(fused_scratch);
    p.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(fused_scratch) + sizeof(GradScratchHdr));
    EH_CUDA(cudaLaunchKernelEx(&cfg, kern, maps, p));
    return EH_OK;
}

int mc_cluster_launch(const eh_mc_job* j, void* fused_scratch, size_t fused_bytes, cudaStream_t st) {
    const int CN = j->C <= 16 ? 16 : 32;
    const int dk = (j->d + 31) / 32 * 32;
    Mc2Maps maps;
    Mc2Params p;
    memset(&p, 0, sizeof(p));
    EH_REQUIRE(mc_cluster_plan(dk, CN, &p.CS, &p.npc, &p.n_stages), "grad_mc: %d columns do not fit the one-pass kernel", dk);
    p.npt = mc2_npt(p.npc);
    for (int sg = 0; sg < j->n_segs; ++sg) {
        EH_REQUIRE(j->d_X[sg] && j->d_y[sg], "grad_mc: null segment %d", sg);
        EH_REQUIRE(((uintptr_t)j->d_X[sg] % 16) == 0, "grad_mc: X must be 16-byte aligned");
        int rc = make_map(&maps.x[sg], j->d_X[sg], (uint64_t)dk, (uint64_t)j->rows_per_seg, (uint64_t)j->ld, 32, 32,
                          CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
        if (rc) return rc;
        p.seg_y[sg] = j->d_y[sg];
        p.coeff[sg] = j->coeff[sg];
    }
    for (int sg = j->n_segs; sg < EH_MAX_SEGS; ++sg) maps.x[sg] = maps.x[0];
    p.W = j->d_W; p.out = j->d_out; p.flag = j->d_flag; p.beta_flag = j->d_beta_flag; p.err = j->d_err;
    p.n_segs = j->n_segs; p.rows_per_seg = j->rows_per_seg; p.tiles_per_seg = (j->rows_per_seg + 31) / 32;
    p.d = j->d; p.dk = dk; p.ldw = j->ldw; p.n_chunks = (dk + 127) / 128; p.C = j->C; p.model = j->model;
    p.w_stride = 2 * CN + 16;
    p.stamp = j->stamp; p.seq = next_launch_seq();
    p.timeout_ns = j->timeout_ns ? j->timeout_ns : 30ull * 1000000000ull;
    return CN == 16 ? launch_mc2<2>(maps, p, fused_scratch, fused_bytes, st)
                    : launch_mc2<4>(maps, p, fused_scratch, fused_bytes, st);
}

}  // namespace eh
