// One-pass multi-column gradient for bf16 feature storage: BOTH products on tcgen05 from ONE shared
// memory tile (SURVEY 7.2 step 5 as written: "tcgen05.mma, accumulators in TMEM, residual epilogue
// from TMEM, second MMA X^T.R from the same SMEM tile").
//
// For 16-bit operands the plain 128B swizzle is a legal tensor-core layout for both majors, so the
// TMA tile [128 rows x 64 columns] (rows of 128 bytes) is
//   * the K-major A operand of the forward product  P = X . W^T   (M = 128 rows, N = classes, K = 16
//     columns per instruction), and
//   * the MN-major A operand of the backward product G = X^T . R  (M = 128 feature columns = two such
//     boxes side by side, N = classes, K = 16 rows per instruction)
// without a second copy (the fp32 / tf32 kernels cannot do this: the MN-major tf32 operand needs the
// 32-byte-atom swizzle, which is invalid for K-major -- see grad_mc_fused.cu).
//
//   warp 0   TMA producer: ring of 32 KB stages (128 rows x 128 columns bf16 = one column chunk)
//   warp 1   forward MMA issuer, warp 3 backward MMA issuer (one thread each): two independent
//            instruction streams, the forward products run ahead as far as TMEM buffers and ring
//            allow, so the TMEM round trip of the residual is off the critical path and neither stream
//            waits for the other's descriptor arithmetic; forward accumulators triple-buffered, backward
//            accumulators of all chunks resident in TMEM for the CTA's whole row range
//   warps 4-15 three epilogue groups (tile t -> group t % 3, one TMEM forward buffer and one R image each):
//            tcgen05.ld of the margins (one row per thread), softmax / logistic residual, R^T written
//            as bf16 into the K-major SW128 operand image, fence.proxy.async, arrive
//   tail     TMEM -> per-CTA partial, device-wide barrier, fixed-order reduction, ship + flag
//
// Parameters / messages stay fp32 ([CN][ldw], as for the other multi-column kernels).  For up to 16
// model columns W and R enter the MMAs as hi + lo bf16 pairs (twice the tile columns, halves added
// after the products), so only the stored features carry bf16 precision; wider models round W and R
// to bf16.  Accumulation is fp32.  dk <= 640.
// No counterpart in the reference (single-vector fp64 models, naive.py:137-139).
#include <cuda_bf16.h>
#include <stdlib.h>

#include "grad_dense.cuh"
#include "tc_common.cuh"

namespace eh {

constexpr int MB_BOX = 128 * 64 * 2;           // TMA box: 128 rows x 64 bf16 columns
constexpr int MB_STAGE = 2 * MB_BOX;           // 128 rows x 128 columns
constexpr int MB_MAX_STAGES = 6;
constexpr int MB_GROUPS = 3;             // epilogue groups = forward TMEM buffers = R images (tiles in the epilogue at once)
constexpr int MB_THREADS = 128 + MB_GROUPS * 128;   // producer, MMA issuer, TMEM allocator, spare + the epilogue groups

struct McBf16Maps {
    CUtensorMap x[EH_MAX_SEGS];
};

struct McBf16Params {
    const float* seg_y[EH_MAX_SEGS];
    float coeff[EH_MAX_SEGS];
    const float* W;
    float* out;
    uint32_t* flag;
    const uint32_t* beta_flag;
    uint32_t* err;
    GradScratchHdr* hdr;
    float* partials;
    int n_segs, rows_per_seg, tiles_per_seg;
    int d, dk, ldw, n_chunks, C, model, n_stages;
    int w_rows;                      // rows of W that exist in memory (single-vector callers: 1, d floats long)
    int out_rows;                    // rows of the result that are reduced and written (each d floats of ldw)
    uint32_t stamp, seq;
    unsigned long long timeout_ns;
};

__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N, int a_mn_major, int b_mn_major) {
    return (1u << 4)                 // accumulator F32
           | (1u << 7)               // A = BF16
           | (1u << 10)              // B = BF16
           | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) | ((uint32_t)(N >> 3) << 17) |
           ((uint32_t)(M >> 4) << 24);
}
// MN-major 16-bit operand, plain 128B swizzle: atoms of 64 elements along M/N (128 B) x 8 K-rows
// (1024 B); `lbo_bytes` between consecutive 64-element atoms along M/N, 1024 B between 8-row K groups
__device__ __forceinline__ uint64_t make_mnmajor_sw128_desc(uint32_t smem_addr, uint32_t lbo_bytes) {
    uint64_t desc = 0;
    desc |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
    desc |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    desc |= (uint64_t)(1024 >> 4) << 32;
    desc |= (uint64_t)1 << 46;
    desc |= (uint64_t)2 << 61;                                 // SWIZZLE_128B
    return desc;
}
__device__ __forceinline__ unsigned short bf16_bits(float v) { return __bfloat16_as_ushort(__float2bfloat16_rn(v)); }

// BN = 16 * NB columns in the tensor-core tiles.  SPLIT: the fp32 parameters and residuals enter the MMAs as
// hi + lo bf16 pairs (columns c and c + BN/2), i.e. with 16 mantissa bits, and the two halves are added
// after the products -- only the stored features are bf16, as in the single-column bf16 kernel.
template <int NB, bool SPLIT>
__global__ void __launch_bounds__(MB_THREADS, 1)
mc_bf16_kernel(const __grid_constant__ McBf16Maps maps, const McBf16Params p) {
    constexpr int BN = 16 * NB;
    constexpr int CO = SPLIT ? BN / 2 : BN;           // model columns (rows of the output)
    constexpr int R_BYTES = 2 * BN * 128;             // two K-major sub-tiles [BN x 64 rows] bf16
    constexpr int W_BOX = BN * 128;                   // [BN x 64 columns] bf16
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* ring = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    unsigned char* Rs = ring + (size_t)p.n_stages * MB_STAGE;
    unsigned char* Ws = Rs + MB_GROUPS * R_BYTES;     // the R images, then 2 * n_chunks W boxes
    uint64_t* full = reinterpret_cast<uint64_t*>(Ws + (size_t)2 * p.n_chunks * W_BOX);
    uint64_t* empty = full + MB_MAX_STAGES;
    uint64_t* tfull = empty + MB_MAX_STAGES;
    uint64_t* tempty = tfull + MB_GROUPS;
    uint64_t* rfull = tempty + MB_GROUPS;
    uint64_t* rempty = rfull + MB_GROUPS;
    uint64_t* tdone = rempty + MB_GROUPS;
    uint64_t* wready = tdone + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wready + 1);
    volatile uint32_t& s_decision = *(tmem_slot + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int G = gridDim.x;
    const int n_tiles = p.n_segs * p.tiles_per_seg;
    const int my_tiles = ((int)blockIdx.x < n_tiles) ? (n_tiles - (int)blockIdx.x + G - 1) / G : 0;
    const int my_items = my_tiles * p.n_chunks;
    const uint32_t need_cols = (uint32_t)(MB_GROUPS * BN + p.n_chunks * BN);
    const uint32_t tmem_cols = need_cols <= 64 ? 64u : need_cols <= 128 ? 128u : need_cols <= 256 ? 256u : 512u;

    if (tid == 0) {
        for (int s = 0; s < MB_MAX_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int s = 0; s < MB_GROUPS; ++s) {
            mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 4);
            mbar_init(&rfull[s], 4); mbar_init(&rempty[s], 1);
        }
        mbar_init(tdone, 1);
        mbar_init(wready, 4 * MB_GROUPS);
        mbar_fence_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_d2 = tmem_base + MB_GROUPS * BN;   // backward accumulators behind the forward buffers

    auto issue = [&](int it, uint64_t pol) {
        const int tl = it / p.n_chunks, ch = it % p.n_chunks;
        const int T = (int)blockIdx.x + tl * G;
        const int sg = T / p.tiles_per_seg, row0 = (T % p.tiles_per_seg) * 128;
        const int st = it % p.n_stages;
        const int nb = (p.dk - ch * 128 > 64) ? 2 : 1;
        unsigned char* dst = ring + (size_t)st * MB_STAGE;
        mbar_arrive_expect_tx(&full[st], (uint32_t)nb * MB_BOX);
        for (int bx = 0; bx < nb; ++bx) tma_load_2d(dst + bx * MB_BOX, &maps.x[sg], ch * 128 + bx * 64, row0, &full[st], pol);
    };

    uint32_t code = DEC_GO;
    if (warp == 0) {
        // ---------------- TMA producer; prefetches X while the parameters are still in flight
        int issued = 0;
        const uint64_t pol = l2_policy_evict_first();
        if (lane == 0) {
            const int pre = my_items < p.n_stages ? my_items : p.n_stages;
            for (; issued < pre; ++issued) issue(issued, pol);
            code = wait_for_parameters(p);
            s_decision = code;
        }
        __syncwarp();
        asm volatile("bar.sync 1, %0;" ::"r"(MB_THREADS) : "memory");
        code = s_decision;
        if (lane == 0) {
            if (code == DEC_SKIP) {
                for (int k = 0; k < issued; ++k) mbar_wait(&full[k % p.n_stages], 0);
            } else {
                for (int it = issued; it < my_items; ++it) {
                    mbar_wait_backoff(&empty[it % p.n_stages], ((it / p.n_stages) - 1) & 1);
                    issue(it, pol);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        asm volatile("bar.sync 1, %0;" ::"r"(MB_THREADS) : "memory");
        code = s_decision;
        if (code == DEC_GO && lane == 0) {
            // ---------------- forward MMA issuer: runs ahead of the backward issuer as far as the TMEM
            // buffers (tempty) and the ring (full) allow; the two issuers are independent threads, so no
            // issue order between forward and backward products has to be planned
            constexpr uint32_t idesc_f = make_idesc_bf16(128, BN, 0, 0);
            const uint64_t w_desc0 = make_kmajor_sw128_desc(smem_u32(Ws));
            mbar_wait(wready, 0);                                 // W operand image written
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            tc_fence_after();
            int it = 0;
            for (int tl = 0; tl < my_tiles; ++tl) {
                const uint32_t as = tl % MB_GROUPS;
                if (tl >= MB_GROUPS) mbar_wait(&tempty[as], ((tl / MB_GROUPS) - 1) & 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                for (int ch = 0; ch < p.n_chunks; ++ch, ++it) {
                    const int st = it % p.n_stages;
                    mbar_wait(&full[st], (it / p.n_stages) & 1);
                    tc_fence_after();
                    const uint64_t a_desc0 = make_kmajor_sw128_desc(smem_u32(ring + (size_t)st * MB_STAGE));
                    const int nbx = (p.dk - ch * 128 > 64) ? 2 : 1;
                    for (int bx = 0; bx < nbx; ++bx) {
                        const uint64_t a_desc = a_desc0 + (uint64_t)(bx * (MB_BOX >> 4));
                        const uint64_t b_desc = w_desc0 + (uint64_t)((ch * 2 + bx) * (W_BOX >> 4));
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk)          // 16 columns = 32 bytes per instruction
                            umma_bf16(d_tmem, a_desc + (uint64_t)(kk * 2), b_desc + (uint64_t)(kk * 2), idesc_f,
                                      (ch > 0 || bx > 0 || kk > 0) ? 1u : 0u);
                    }
                }
                umma_commit(&tfull[as]);
            }
        }
        __syncwarp();
    } else if (warp == 3) {
        asm volatile("bar.sync 1, %0;" ::"r"(MB_THREADS) : "memory");
        code = s_decision;
        if (code == DEC_GO && lane == 0) {
            // ---------------- backward MMA issuer
            constexpr uint32_t idesc_b = make_idesc_bf16(128, BN, /*a_mn_major=*/1, 0);
            const uint64_t r_desc0 = make_kmajor_sw128_desc(smem_u32(Rs));
            int it = 0;
            for (int tl = 0; tl < my_tiles; ++tl) {
                const int rb = tl % MB_GROUPS;
                mbar_wait(&rfull[rb], (tl / MB_GROUPS) & 1);
                tc_fence_after();
                const uint64_t r_desc = r_desc0 + (uint64_t)(rb * (R_BYTES >> 4));
                for (int ch = 0; ch < p.n_chunks; ++ch, ++it) {
                    const int st = it % p.n_stages;
                    const uint64_t a_desc0 = make_mnmajor_sw128_desc(smem_u32(ring + (size_t)st * MB_STAGE), MB_BOX);
                    const uint32_t d_tmem = tmem_d2 + ch * BN;
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk)              // 16 rows per instruction = 2048 bytes of the box
                        umma_bf16(d_tmem, a_desc0 + (uint64_t)(kk * (2048 >> 4)),
                                  r_desc + (uint64_t)((kk >> 2) * ((BN * 128) >> 4) + (kk & 3) * 2), idesc_b,
                                  (tl > 0 || kk > 0) ? 1u : 0u);
                    umma_commit(&empty[st]);
                }
                umma_commit(&rempty[rb]);
            }
            umma_commit(tdone);
        }
        __syncwarp();
    } else if (warp >= 4) {
        asm volatile("bar.sync 1, %0;" ::"r"(MB_THREADS) : "memory");
        code = s_decision;
        if (code == DEC_GO) {
            const int q = (warp - 4) & 3;              // TMEM lane quarter of this warp
            const int eg = (warp - 4) >> 2;            // epilogue group: tiles tl = eg, eg + MB_GROUPS, ...
            const int et = tid - 128;
            // ---------------- W^T as bf16 into the K-major SW128 operand image, one box per 64 columns
            for (int i = et; i < CO * p.n_chunks * 128; i += MB_GROUPS * 128) {
                const int n = i / (p.n_chunks * 128), c = i % (p.n_chunks * 128);
                const float w = (n < p.w_rows && c < p.d) ? __ldcg(p.W + (size_t)n * p.ldw + c) : 0.f;
                const int box = c >> 6, cc = c & 63;
                unsigned char* at = Ws + (size_t)box * W_BOX + (cc & 7) * 2;
                const unsigned short hi = bf16_bits(w);
                *reinterpret_cast<unsigned short*>(at + n * 128 + (((cc >> 3) ^ (n & 7)) << 4)) = hi;
                if constexpr (SPLIT) {
                    const int n2 = n + CO;
                    *reinterpret_cast<unsigned short*>(at + n2 * 128 + (((cc >> 3) ^ (n2 & 7)) << 4)) =
                        bf16_bits(w - __uint_as_float((uint32_t)hi << 16));
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(wready);

            const int T0 = (int)blockIdx.x + eg * G;
            int sg = T0 / p.tiles_per_seg, tis = T0 % p.tiles_per_seg;
            const int rr = q * 32 + lane;                  // this thread's tile row
            for (int tl = eg; tl < my_tiles; tl += MB_GROUPS) {
                const uint32_t as = eg;
                const int grow = tis * 128 + rr;
                const bool valid = grow < p.rows_per_seg;
                const float yv = valid ? __ldg(p.seg_y[sg] + grow) : 0.f;
                const float cf = p.coeff[sg];
                mbar_wait(&tfull[as], (tl / MB_GROUPS) & 1);
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
                if (lane == 0) mbar_arrive(&tempty[as]);
                if constexpr (SPLIT) {
#pragma unroll
                    for (int c = 0; c < CO; ++c) v[c] += v[c + CO];
                }
                if (p.model == 2) {
                    float m = -3.0e38f;
#pragma unroll
                    for (int c = 0; c < CO; ++c) if (c < p.C) m = fmaxf(m, v[c]);
                    float s = 0.f;
#pragma unroll
                    for (int c = 0; c < CO; ++c) {
                        v[c] = (c < p.C) ? __expf(v[c] - m) : 0.f;
                        s += v[c];
                    }
                    const float inv = __fdividef(1.f, s);
                    const int yc = (int)yv;
#pragma unroll
                    for (int c = 0; c < CO; ++c) v[c] = valid ? cf * (v[c] * inv - (c == yc ? 1.f : 0.f)) : 0.f;
                } else if (p.model == 3) {
#pragma unroll
                    for (int c = 0; c < CO; ++c)
                        v[c] = (valid && c < p.C) ? cf * __fdividef(-yv, __expf(v[c] * yv) + 1.0f) : 0.f;
                } else {                                   // linear regression per column: -2 (y - p)
#pragma unroll
                    for (int c = 0; c < CO; ++c) v[c] = (valid && c < p.C) ? cf * -2.0f * (yv - v[c]) : 0.f;
                }
                // R^T (bf16) into the operand image: sub-tile rr/64, K index rr%64
                if (tl >= MB_GROUPS) mbar_wait(&rempty[eg], ((tl / MB_GROUPS) - 1) & 1);
                unsigned char* rbase = Rs + eg * R_BYTES + (rr >> 6) * (BN * 128) + (rr & 7) * 2;
                const int c16 = (rr & 63) >> 3;
#pragma unroll
                for (int c = 0; c < CO; ++c) {
                    const unsigned short hi = bf16_bits(v[c]);
                    *reinterpret_cast<unsigned short*>(rbase + c * 128 + ((c16 ^ (c & 7)) << 4)) = hi;
                    if constexpr (SPLIT)
                        *reinterpret_cast<unsigned short*>(rbase + (c + CO) * 128 + ((c16 ^ ((c + CO) & 7)) << 4)) =
                            bf16_bits(v[c] - __uint_as_float((uint32_t)hi << 16));
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(&rfull[eg]);
                tis += MB_GROUPS * G;
                while (tis >= p.tiles_per_seg) { tis -= p.tiles_per_seg; ++sg; }
            }

            // ---------------- backward accumulators: TMEM -> this CTA's partial [CN][ldw]
            float* mine = p.partials + (size_t)blockIdx.x * ((size_t)p.out_rows * p.ldw);
            if (my_tiles > 0) {
                mbar_wait(tdone, 0);
                tc_fence_after();
            }
            for (int ch = eg; ch < p.n_chunks; ch += MB_GROUPS) {
                const int j = ch * 128 + q * 32 + lane;
                const uint32_t taddr = tmem_d2 + ((uint32_t)(q * 32) << 16) + ch * BN;
#pragma unroll
                for (int cb = 0; cb < CO / 16; ++cb) {
                    float t16[16];
                    if (my_tiles > 0) {
                        tmem_ld16(taddr + cb * 16, t16);
                        if constexpr (SPLIT) {                 // + the lo-residual half
                            float u16[16];
                            tmem_ld16(taddr + CO + cb * 16, u16);
#pragma unroll
                            for (int e = 0; e < 16; ++e) t16[e] += u16[e];
                        }
                    } else {
#pragma unroll
                        for (int e = 0; e < 16; ++e) t16[e] = 0.f;
                    }
                    if (j < p.ldw) {
#pragma unroll
                        for (int e = 0; e < 16; ++e)
                            if (cb * 16 + e < p.out_rows) __stcg(mine + (size_t)(cb * 16 + e) * p.ldw + j, (j < p.d) ? t16[e] : 0.f);
                    }
                }
            }
            tc_fence_before();
        }
    } else {
        asm volatile("bar.sync 1, %0;" ::"r"(MB_THREADS) : "memory");
        code = s_decision;
    }

    // ---------------- common tail
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, tmem_cols);
    }
    if (code == DEC_SKIP) return;

    uint32_t gen = 0;
    grid_barrier(&p.hdr->barrier, gen);
    const int n4 = (p.out_rows * p.ldw) >> 2;
    for (int i4 = (int)blockIdx.x * MB_THREADS + tid; i4 < n4; i4 += G * MB_THREADS) {
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4* src = reinterpret_cast<const float4*>(p.partials) + i4;
        const int col = (i4 * 4) % p.ldw;
        if (col < p.n_chunks * 128) {
#pragma unroll 8
            for (int b = 0; b < G; ++b) {
                const float4 t = __ldcg(src + (size_t)b * n4);
                s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
            }
        }
        if (p.out_rows == CO || col + 3 < p.d) {           // full [CN][ldw] result, or inside a d-float vector
            reinterpret_cast<float4*>(p.out)[i4] = s;
        } else {                                           // tail of a caller's d-float vector
            float* o = p.out + (size_t)i4 * 4;
            if (col + 0 < p.d) o[0] = s.x;
            if (col + 1 < p.d) o[1] = s.y;
            if (col + 2 < p.d) o[2] = s.z;
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

// ------------------------------------------------------------------------------------ host
static size_t mc_bf16_smem(int dk, int BN, int stages) {
    const int n_chunks = (dk + 127) / 128;
    return 1024 + (size_t)stages * MB_STAGE + (size_t)MB_GROUPS * 2 * BN * 128 + (size_t)2 * n_chunks * BN * 128 +
           (2 * MB_MAX_STAGES + 4 * MB_GROUPS + 2) * sizeof(uint64_t) + 16;
}

// hi + lo operand pairs (16 mantissa bits for parameters and residuals) for up to 16 model columns
static bool mc_bf16_split(int CN) { return CN == 16 && getenv("EH_MC_BF16_PLAIN") == nullptr; }

int mc_bf16_stages(int dk, int CN) {
    const int n_chunks = (dk + 127) / 128;
    const int BN = mc_bf16_split(CN) ? 2 * CN : CN;       // columns inside the tensor-core tiles
    if (MB_GROUPS * BN + n_chunks * BN > 512) return 0;
    for (int s = MB_MAX_STAGES; s >= n_chunks && s >= 1; --s)
        if (mc_bf16_smem(dk, BN, s) <= (size_t)227 * 1024) return s;
    return 0;
}

static int make_map_bf16(CUtensorMap* map, const void* base, uint64_t inner, uint64_t outer, uint64_t ld_elems) {
    EncodeTiledFn fn = encode_fn();
    EH_REQUIRE(fn, "grad_mc: cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {ld_elems * 2};
    cuuint32_t box[2] = {64, 128};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)base, dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    EH_REQUIRE(r == CUDA_SUCCESS, "grad_mc: cuTensorMapEncodeTiled (bf16) failed (%d)", (int)r);
    return EH_OK;
}

template <int NB, bool SPLIT>
static int launch_mc_bf16(const McBf16Maps& maps, McBf16Params& p, int grid, cudaStream_t st) {
    const size_t smem = mc_bf16_smem(p.dk, 16 * NB, p.n_stages);
    auto kern = mc_bf16_kernel<NB, SPLIT>;
    EH_SET_MAX_SMEM((kern), 227 * 1024);
    void* args[] = {(void*)&maps, (void*)&p};
    return launch_grid_sync((const void*)kern, dim3((unsigned)grid), dim3(MB_THREADS), args, smem, st);
}

// bf16 feature rows (eh_mc_job.x_bf16): `fused_scratch` = zero-initialised barrier header + per-CTA partials
// `single_vector`: W and the result are plain d-float vectors (the single-column dense entry point)
int mc_bf16_launch(const eh_mc_job* j, void* fused_scratch, size_t fused_bytes, cudaStream_t st, bool single_vector) {
    const int CN = j->C <= 16 ? 16 : 32;
    const int dk = (j->d + 31) / 32 * 32;
    EH_REQUIRE((j->ld % 8) == 0 && j->ld >= dk, "grad_mc(bf16): rows must be padded to %d columns, ld %% 8 == 0 (ld=%d)", dk, j->ld);
    McBf16Maps maps;
    McBf16Params p;
    memset(&p, 0, sizeof(p));
    p.n_stages = mc_bf16_stages(dk, CN);
    EH_REQUIRE(p.n_stages > 0, "grad_mc(bf16): %d columns x %d classes do not fit the one-pass kernel (d <= 640)", dk, CN);
    for (int sg = 0; sg < j->n_segs; ++sg) {
        EH_REQUIRE(j->d_X[sg] && j->d_y[sg], "grad_mc: null segment %d", sg);
        EH_REQUIRE(((uintptr_t)j->d_X[sg] % 16) == 0, "grad_mc: X must be 16-byte aligned");
        int rc = make_map_bf16(&maps.x[sg], j->d_X[sg], (uint64_t)dk, (uint64_t)j->rows_per_seg, (uint64_t)j->ld);
        if (rc) return rc;
        p.seg_y[sg] = j->d_y[sg];
        p.coeff[sg] = j->coeff[sg];
    }
    for (int sg = j->n_segs; sg < EH_MAX_SEGS; ++sg) maps.x[sg] = maps.x[0];
    p.W = j->d_W; p.out = j->d_out; p.flag = j->d_flag; p.beta_flag = j->d_beta_flag; p.err = j->d_err;
    p.n_segs = j->n_segs; p.rows_per_seg = j->rows_per_seg; p.tiles_per_seg = (j->rows_per_seg + 127) / 128;
    p.d = j->d; p.dk = dk; p.ldw = j->ldw; p.n_chunks = (dk + 127) / 128; p.C = j->C; p.model = j->model;
    p.w_rows = single_vector ? 1 : CN; p.out_rows = single_vector ? 1 : CN;
    p.stamp = j->stamp; p.seq = next_launch_seq();
    p.timeout_ns = j->timeout_ns ? j->timeout_ns : 30ull * 1000000000ull;
    const long long n_tiles = (long long)p.n_segs * p.tiles_per_seg;
    long long grid = worker_sms();
    if (grid > n_tiles) grid = n_tiles;
    const size_t need = sizeof(GradScratchHdr) + (size_t)grid * p.out_rows * p.ldw * sizeof(float);
    EH_REQUIRE(fused_bytes >= need, "grad_mc(bf16): scratch too small (%zu < %zu)", fused_bytes, need);
    p.hdr = reinterpret_cast<GradScratchHdr*>(fused_scratch);
    p.partials = reinterpret_cast<float*>(reinterpret_cast<char*>(fused_scratch) + sizeof(GradScratchHdr));
    if (mc_bf16_split(CN)) return launch_mc_bf16<2, true>(maps, p, (int)grid, st);
    return CN == 16 ? launch_mc_bf16<1, false>(maps, p, (int)grid, st) : launch_mc_bf16<2, false>(maps, p, (int)grid, st);
}

}  // namespace eh
