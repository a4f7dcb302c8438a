// Pieces shared by the one-pass multi-column gradient kernels (grad_mc_fused.cu, grad_mc_cluster.cu):
// shared-space load/store wrappers, the warp-level tf32 MMA and the residual of a row whose class
// columns are spread over the four lanes of a quad (the m16n8k8 accumulator layout).
#pragma once
#include "tc_common.cuh"

namespace eh {

__device__ __forceinline__ uint2 mc_lds64(uint32_t addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ float2 mc_lds_f2(uint32_t addr) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ void mc_sts32(uint32_t addr, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
// four 8x(16 byte) matrices: lane L supplies the row address of matrix L/8, row L%8; thread T gets the
// 32-bit word T%4 of row T/4 of each matrix -- exactly the tf32 m16n8k8 fragment order
__device__ __forceinline__ void mc_ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
                 : "r"(addr));
}
__device__ __forceinline__ void mc_mma_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                            uint32_t b1) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float mc_logistic(float margin, float y, float cf) {
    return cf * __fdividef(-y, __expf(margin * y) + 1.0f);
}

// Residual of accumulator row h (0: row g, 1: row g + 8) of an m16n8k8 accumulator set acc[NBC][4]:
// this lane holds classes nb*8 + 2t + j; the row's other classes sit in the quad's other lanes.
// model 2: cf * (softmax(margins) - onehot(y)); model 3: cf * -y / (exp(margin y) + 1) per class.
// Result is rounded to tf32 (operand of the backward MMA) and zero for classes >= C or invalid rows.
template <int NBC>
__device__ __forceinline__ void mc_quad_row_residual(const float (&acc)[NBC][4], int h, int t, int C, int model, float y,
                                                     float cf, bool valid, float (&res)[NBC][2]) {
    if (model == 2) {
        float m = -3.0e38f;
#pragma unroll
        for (int nb = 0; nb < NBC; ++nb)
#pragma unroll
            for (int j = 0; j < 2; ++j)
                if (nb * 8 + 2 * t + j < C) m = fmaxf(m, acc[nb][2 * h + j]);
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
        m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
        float sum = 0.f;
#pragma unroll
        for (int nb = 0; nb < NBC; ++nb)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const float e = (nb * 8 + 2 * t + j < C) ? __expf(acc[nb][2 * h + j] - m) : 0.f;
                res[nb][j] = e;
                sum += e;
            }
        sum += __shfl_xor_sync(0xffffffffu, sum, 1);
        sum += __shfl_xor_sync(0xffffffffu, sum, 2);
        const float inv = __fdividef(1.f, sum);
        const int yc = (int)y;
#pragma unroll
        for (int nb = 0; nb < NBC; ++nb)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const float r = cf * (res[nb][j] * inv - ((nb * 8 + 2 * t + j) == yc ? 1.f : 0.f));
                res[nb][j] = valid ? round_tf32(r) : 0.f;
            }
    } else {
#pragma unroll
        for (int nb = 0; nb < NBC; ++nb)
#pragma unroll
            for (int j = 0; j < 2; ++j)
                res[nb][j] = (valid && nb * 8 + 2 * t + j < C) ? round_tf32(mc_logistic(acc[nb][2 * h + j], y, cf)) : 0.f;
    }
}

}  // namespace eh
