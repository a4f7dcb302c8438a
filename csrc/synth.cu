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

__global__ void __launch_bounds__(256) synth_gmm_kernel(float* __restrict__ X, float* __restrict__ y, long long rows,
                                                        int d, int ld, uint64_t seed, uint64_t partition, int task) {
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
                } else {
                    v[e] = 0.f;                       // zero padding up to ld
                }
            }
            *reinterpret_cast<float4*>(X + row * ld + c4 * 4) = make_float4(v[0], v[1], v[2], v[3]);
        }
        dot = warp_sum(dot);
        if (lane == 0) {
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

extern "C" int eh_synth_gmm(float* d_X, float* d_y, int64_t rows, int d, int ld, uint64_t seed, uint64_t partition,
                            int task, void* stream) {
    using namespace eh;
    EH_REQUIRE(d_X && d_y && rows > 0 && d > 0, "synth_gmm: bad argument");
    EH_REQUIRE(ld >= d && (ld % 4) == 0 && ((uintptr_t)d_X % 16) == 0, "synth_gmm: ld must be >= d, %%4, X 16B aligned");
    const DevInfo& di = dev_info();
    long long blocks = (rows + 7) / 8;
    const long long cap = (long long)di.sm_count * 16;
    if (blocks > cap) blocks = cap;
    synth_gmm_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_X, d_y, rows, d, ld, seed, partition, task);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}
