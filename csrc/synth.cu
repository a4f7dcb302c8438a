// Synthetic data on the device: two-component GMM features with logistic (or linear) targets.
//
// Same distribution as the reference generator (generate_data.py:23-45, util.py:39-47):
// beta* in {-1,+1}^d, x ~ N(+-(1.5/d) beta*, (10/sqrt(d))^2 I), y = 2 Bernoulli(sigmoid(x.beta*)) - 1.
// Counter-based Philox4x32-10 keyed by (seed) and indexed by (partition, row, column), so a
// partition replicated on several GPUs (FRC groups, cyclic supports) is bit-identical everywhere
// without ever touching a disk.  One warp generates one row, accumulates x.beta* on the fly and
// draws the label.
#include "common.cuh"

namespace eh {

__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    const uint32_t hi0 = __umulhi(M0, c[0]), lo0 = M0 * c[0];
    const uint32_t hi1 = __umulhi(M1, c[2]), lo1 = M1 * c[2];
    const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}
__device__ __forceinline__ void philox4x32(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        philox_round(c, k0, k1);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
}
__device__ __forceinline__ float u01(uint32_t x) { return ((float)(x >> 8) + 0.5f) * (1.0f / 16777216.0f); }

__device__ __forceinline__ float beta_star(uint64_t seed, int col) {
    uint32_t c[4] = {(uint32_t)col, 0x62657461u, 0u, 0u};
    philox4x32(c, (uint32_t)seed, (uint32_t)(seed >> 32) ^ 0x5bd1e995u);
    return (c[0] & 1u) ? 1.0f : -1.0f;
}

constexpr int SYNTH_MAX_CLASSES = 8;

// class c's +-1 weight of a column (multi-class labels): bit c of the second Philox word of beta_star's draw
__device__ __forceinline__ uint32_t class_sign_bits(uint64_t seed, int col) {
    uint32_t c[4] = {(uint32_t)col, 0x62657461u, 0u, 0u};
    philox4x32(c, (uint32_t)seed, (uint32_t)(seed >> 32) ^ 0x5bd1e995u);
    return c[1];
}
__device__ __forceinline__ float gumbel(uint32_t x) { return -__logf(-__logf(u01(x))); }

__global__ void __launch_bounds__(256) synth_gmm_kernel(float* __restrict__ X, float* __restrict__ y, long long rows,
                                                        int d, int ld, uint64_t seed, uint64_t partition, int task,
                                                        int classes) {
    const int lane = threadIdx.x & 31;
    const long long gwarp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const float mfac = 10.0f * rsqrtf((float)d);
    const float mu = 1.5f / (float)d;
    const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    const int ld4 = ld >> 2;
    for (long long row = gwarp; row < rows; row += nwarps) {
        // per-row draws: mixture component and the label uniform
        uint32_t rc[4] = {(uint32_t)row, (uint32_t)(row >> 32), 0xffffffffu, (uint32_t)partition};
        philox4x32(rc, k0, k1);
        const float comp = (rc[0] & 1u) ? 1.0f : -1.0f;
        float dot = 0.f;
        float cdot[SYNTH_MAX_CLASSES];
#pragma unroll
        for (int k = 0; k < SYNTH_MAX_CLASSES; ++k) cdot[k] = 0.f;
        for (int c4 = lane; c4 < ld4; c4 += 32) {
            uint32_t c[4] = {(uint32_t)row, (uint32_t)(row >> 32), (uint32_t)c4, (uint32_t)partition};
            philox4x32(c, k0, k1);
            // Box-Muller: two pairs -> four normals
            const float r0 = sqrtf(-2.0f * __logf(u01(c[0]))), r1 = sqrtf(-2.0f * __logf(u01(c[2])));
            float s0, c0, s1, c1;
            __sincosf(6.2831853f * u01(c[1]), &s0, &c0);
            __sincosf(6.2831853f * u01(c[3]), &s1, &c1);
            float v[4] = {r0 * c0, r0 * s0, r1 * c1, r1 * s1};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int col = c4 * 4 + e;
                if (col < d) {
                    const float bs = beta_star(seed, col);
                    v[e] = mfac * v[e] + comp * mu * bs;
                    dot = fmaf(v[e], bs, dot);
                    if (task == 2) {
                        const uint32_t bits = class_sign_bits(seed, col);
#pragma unroll
                        for (int k = 0; k < SYNTH_MAX_CLASSES; ++k) cdot[k] += ((bits >> k) & 1u) ? v[e] : -v[e];
                    }
                } else {
                    v[e] = 0.f;                       // zero padding up to ld
                }
            }
            *reinterpret_cast<float4*>(X + row * ld + c4 * 4) = make_float4(v[0], v[1], v[2], v[3]);
        }
        dot = warp_sum(dot);
        if (task == 2) {
            // multi-class: y = argmax_k (0.3 x.w_k + Gumbel), w_k in {-1,+1}^d  (labels 0..classes-1 as floats)
#pragma unroll
            for (int k = 0; k < SYNTH_MAX_CLASSES; ++k) cdot[k] = warp_sum(cdot[k]);
            if (lane == 0) {
                uint32_t gc[4] = {(uint32_t)row, (uint32_t)(row >> 32), 0xfffffffeu, (uint32_t)partition};
                philox4x32(gc, k0, k1);
                uint32_t gd[4] = {(uint32_t)row, (uint32_t)(row >> 32), 0xfffffffdu, (uint32_t)partition};
                philox4x32(gd, k0, k1);
                const uint32_t gu[8] = {gc[0], gc[1], gc[2], gc[3], gd[0], gd[1], gd[2], gd[3]};
                int best = 0;
                float bv = -INFINITY;
#pragma unroll
                for (int k = 0; k < SYNTH_MAX_CLASSES; ++k) {
                    const float v = 0.3f * cdot[k] + gumbel(gu[k]);
                    if (k < classes && v > bv) { bv = v; best = k; }
                }
                y[row] = (float)best;
            }
        } else if (lane == 0) {
            if (task == 1) {
                const float nz = sqrtf(-2.0f * __logf(u01(rc[1]))) * __cosf(6.2831853f * u01(rc[2]));
                y[row] = dot * rsqrtf((float)d) + 0.1f * nz;
            } else {
                const float prob = 1.0f / (1.0f + __expf(-dot));
                y[row] = (u01(rc[1]) < prob) ? 1.0f : -1.0f;
            }
        }
    }
}

}  // namespace eh

namespace eh {

// ---------------------------------------------------------------------------------------------
// One-hot data of a named dataset's SHAPE: every row has one 1 per categorical
// source column plus the always-on bias feature (arrange_real_data.py:41-64,151-178,211-221 builds
// exactly that from the real files, which cannot be fetched here).  Feature block f covers
// [edges[f], edges[f+1]); the level of a row in a block is a heavy-tailed rank (P(rank >= k) =
// (k+1)^-0.3, the tail of numpy's zipf(1.3) used by the host generator data/shapes.py) scattered
// over the block by a multiplicative hash.  Labels from a hidden N(0, 0.5^2) parameter per
// feature (logistic / linear) or per feature and class (multi-class).  One warp per row.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void hidden_params(uint64_t seed, int col, float (&w)[SYNTH_MAX_CLASSES]) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        uint32_t c[4] = {(uint32_t)col, 0x6f6e6568u + (uint32_t)h, 0u, 0u};
        philox4x32(c, (uint32_t)seed, (uint32_t)(seed >> 32) ^ 0x2545f491u);
        const float r0 = sqrtf(-2.0f * __logf(u01(c[0]))), r1 = sqrtf(-2.0f * __logf(u01(c[2])));
        float s0, c0, s1, c1;
        __sincosf(6.2831853f * u01(c[1]), &s0, &c0);
        __sincosf(6.2831853f * u01(c[3]), &s1, &c1);
        w[4 * h + 0] = 0.5f * r0 * c0; w[4 * h + 1] = 0.5f * r0 * s0;
        w[4 * h + 2] = 0.5f * r1 * c1; w[4 * h + 3] = 0.5f * r1 * s1;
    }
}

__global__ void __launch_bounds__(256) synth_onehot_kernel(int32_t* __restrict__ idx, float* __restrict__ y, long long rows,
                                                           int d, int nnz, const int32_t* __restrict__ edges, uint64_t seed,
                                                           uint64_t partition, int task, int classes) {
    const int lane = threadIdx.x & 31;
    const long long gwarp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    for (long long row = gwarp; row < rows; row += nwarps) {
        float acc[SYNTH_MAX_CLASSES];
#pragma unroll
        for (int k = 0; k < SYNTH_MAX_CLASSES; ++k) acc[k] = 0.f;
        for (int f = lane; f < nnz; f += 32) {
            int col;
            if (f == nnz - 1) {
                col = d - 1;                                      // bias feature (arrange_real_data.py:55)
            } else {
                const int lo = edges[f], size = edges[f + 1] - lo;
                uint32_t c[4] = {(uint32_t)row, (uint32_t)(row >> 32), (uint32_t)f, (uint32_t)partition};
                philox4x32(c, k0, k1);
                const float rk = fminf(__powf(u01(c[0]), -1.0f / 0.3f) - 1.0f, (float)(size - 1));
                const unsigned long long r = (unsigned long long)rk;
                col = lo + (int)((r * 2654435761ull + (unsigned long long)f * 97ull) % (unsigned long long)size);
            }
            idx[row * nnz + f] = col;
            float w[SYNTH_MAX_CLASSES];
            hidden_params(seed, col, w);
#pragma unroll
            for (int k = 0; k < SYNTH_MAX_CLASSES; ++k) acc[k] += w[k];
        }
#pragma unroll
        for (int k = 0; k < SYNTH_MAX_CLASSES; ++k) acc[k] = warp_sum(acc[k]);
        if (lane == 0) {
            uint32_t rc[4] = {(uint32_t)row, (uint32_t)(row >> 32), 0xffffffffu, (uint32_t)partition};
            philox4x32(rc, k0, k1);
            if (task == 2) {
                uint32_t gd[4] = {(uint32_t)row, (uint32_t)(row >> 32), 0xfffffffeu, (uint32_t)partition};
                philox4x32(gd, k0, k1);
                const uint32_t gu[8] = {rc[0], rc[1], rc[2], rc[3], gd[0], gd[1], gd[2], gd[3]};
                int best = 0;
                float bv = -INFINITY;
#pragma unroll
                for (int k = 0; k < SYNTH_MAX_CLASSES; ++k) {
                    const float v = acc[k] + gumbel(gu[k]);
                    if (k < classes && v > bv) { bv = v; best = k; }
                }
                y[row] = (float)best;
            } else if (task == 1) {
                const float nz = sqrtf(-2.0f * __logf(u01(rc[1]))) * __cosf(6.2831853f * u01(rc[2]));
                y[row] = acc[0] * rsqrtf((float)nnz) + 0.05f * nz;
            } else {
                y[row] = (u01(rc[1]) < 1.0f / (1.0f + __expf(-acc[0]))) ? 1.0f : -1.0f;
            }
        }
    }
}

}  // namespace eh

extern "C" int eh_synth_onehot(int32_t* d_idx, float* d_y, int64_t rows, int d, int nnz, const int32_t* d_edges,
                               uint64_t seed, uint64_t partition, int task, int classes, void* stream) {
    using namespace eh;
    EH_REQUIRE(d_idx && d_y && d_edges && rows > 0 && d > 1 && nnz >= 2 && nnz <= d, "synth_onehot: bad argument");
    EH_REQUIRE(task >= 0 && task <= 2 && (task != 2 || (classes >= 2 && classes <= SYNTH_MAX_CLASSES)),
               "synth_onehot: task 0/1/2, multi-class labels need 2..%d classes", SYNTH_MAX_CLASSES);
    const DevInfo& di = dev_info();
    long long blocks = (rows + 7) / 8;
    const long long cap = (long long)di.sm_count * 16;
    if (blocks > cap) blocks = cap;
    synth_onehot_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_idx, d_y, rows, d, nnz, d_edges, seed, partition,
                                                                            task, classes);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

extern "C" int eh_synth_gmm(float* d_X, float* d_y, int64_t rows, int d, int ld, uint64_t seed, uint64_t partition,
                            int task, int classes, void* stream) {
    using namespace eh;
    EH_REQUIRE(d_X && d_y && rows > 0 && d > 0, "synth_gmm: bad argument");
    EH_REQUIRE(task >= 0 && task <= 2 && (task != 2 || (classes >= 2 && classes <= SYNTH_MAX_CLASSES)),
               "synth_gmm: task 0 (logistic) / 1 (linear) / 2 (multi-class, 2..%d classes)", SYNTH_MAX_CLASSES);
    EH_REQUIRE(ld >= d && (ld % 4) == 0 && ((uintptr_t)d_X % 16) == 0, "synth_gmm: ld must be >= d, %%4, X 16B aligned");
    const DevInfo& di = dev_info();
    long long blocks = (rows + 7) / 8;
    const long long cap = (long long)di.sm_count * 16;
    if (blocks > cap) blocks = cap;
    synth_gmm_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_X, d_y, rows, d, ld, seed, partition, task, classes);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}
