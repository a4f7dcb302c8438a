// Fallback for very wide dense matrices (row stride > 8192 floats): the row tile no longer fits
// the shared-memory ring of the one-pass kernel, so the gradient is formed in two passes over X
// (margins + residual, then X^T r), 64K rows at a time so the second pass of a chunk mostly hits
// L2.  Same contract as eh_grad_dense (wait for the parameter flag, write the destination, release
// the arrival flag); used only when ld > 8192 -- every dataset shape of the reference is either
// narrower or one-hot sparse (grad_ell.cu).
#include "grad_dense.cuh"

namespace eh {

constexpr int WIDE_CHUNK = 65536;

__global__ void __launch_bounds__(256) wide_residual_kernel(const float* __restrict__ X, const float* __restrict__ y,
                                                            const float* __restrict__ beta, int rows, int d, int ld,
                                                            float coeff, int model, float* __restrict__ r,
                                                            const uint32_t* skip) {
    if (skip && *skip) return;
    const int lane = threadIdx.x & 31;
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = (gridDim.x * blockDim.x) >> 5;
    const int d4 = d >> 2;
    for (int row = gw; row < rows; row += nw) {
        const float4* xr = reinterpret_cast<const float4*>(X + (size_t)row * ld);
        float s = 0.f;
        for (int c4 = lane; c4 < d4; c4 += 32) {
            const float4 x = __ldg(xr + c4);
            const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + c4);
            s = fmaf(x.x, b.x, fmaf(x.y, b.y, fmaf(x.z, b.z, fmaf(x.w, b.w, s))));
        }
        for (int c = (d4 << 2) + lane; c < d; c += 32) s = fmaf(X[(size_t)row * ld + c], beta[c], s);
        s = warp_sum(s);
        if (lane == 0) r[row] = coeff * residual(model, s, __ldg(y + row));
    }
}

// g[c] += sum_rows X[row, c] * r[row]; one thread owns 4 columns, a CTA 1024 columns x ROWS_PER_CTA rows
__global__ void __launch_bounds__(256) wide_xtr_kernel(const float* __restrict__ X, const float* __restrict__ r, int rows,
                                                       int ld, float* __restrict__ g, const uint32_t* skip) {
    if (skip && *skip) return;
    constexpr int ROWS_PER_CTA = 256;
    const int c4 = blockIdx.x * 256 + threadIdx.x;
    const int row0 = blockIdx.y * ROWS_PER_CTA;
    const int row1 = min(rows, row0 + ROWS_PER_CTA);
    if (c4 * 4 >= ld) return;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int row = row0; row < row1; ++row) {
        const float rv = __ldg(r + row);
        const float4 x = __ldg(reinterpret_cast<const float4*>(X + (size_t)row * ld) + c4);
        acc.x = fmaf(rv, x.x, acc.x); acc.y = fmaf(rv, x.y, acc.y);
        acc.z = fmaf(rv, x.z, acc.z); acc.w = fmaf(rv, x.w, acc.w);
    }
    float* dst = g + (size_t)c4 * 4;
    atomicAdd(dst + 0, acc.x); atomicAdd(dst + 1, acc.y); atomicAdd(dst + 2, acc.z); atomicAdd(dst + 3, acc.w);
}

__global__ void wide_wait_kernel(const uint32_t* flag, uint32_t stamp, uint32_t* err, unsigned long long timeout_ns,
                                 uint32_t* skip) {
    if (threadIdx.x || blockIdx.x) return;
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
__global__ void wide_ship_kernel(const float* __restrict__ g, float* __restrict__ out, int d, const uint32_t* skip) {
    if (skip && *skip) return;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < d; i += gridDim.x * blockDim.x) out[i] = g[i];
    __threadfence_system();
}
__global__ void wide_flag_kernel(uint32_t* flag, uint32_t stamp, const uint32_t* skip) {
    if (threadIdx.x == 0 && blockIdx.x == 0 && !(skip && *skip)) {
        st_arrival_time(flag);
        st_release_sys(flag, stamp);
    }
}

int launch_wide(GradParams& p, void* d_scratch, size_t scratch_bytes, cudaStream_t st) {
    const size_t need = 1024 + ((size_t)WIDE_CHUNK + p.ld) * sizeof(float);
    EH_REQUIRE(scratch_bytes >= need, "grad_dense(wide): scratch too small (%zu < %zu)", scratch_bytes, need);
    EH_REQUIRE(((uintptr_t)p.beta % 16) == 0, "grad_dense(wide): beta must be 16-byte aligned");
    uint32_t* skip = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(d_scratch) + 512);   // past the barrier words
    float* r = reinterpret_cast<float*>(reinterpret_cast<char*>(d_scratch) + 1024);
    float* g = r + WIDE_CHUNK;
    const uint32_t* skp = nullptr;
    if (p.beta_flag) {
        wide_wait_kernel<<<1, 32, 0, st>>>(p.beta_flag, p.stamp, p.err, p.timeout_ns, skip);
        skp = skip;
    }
    EH_CUDA(cudaMemsetAsync(g, 0, (size_t)p.ld * sizeof(float), st));
    const int sm = worker_sms();
    for (int sg = 0; sg < p.n_segs; ++sg) {
        for (int r0 = 0; r0 < p.rows_per_seg; r0 += WIDE_CHUNK) {
            const int rows = (p.rows_per_seg - r0) < WIDE_CHUNK ? (p.rows_per_seg - r0) : WIDE_CHUNK;
            const float* X = reinterpret_cast<const float*>(p.seg_X[sg]) + (size_t)r0 * p.ld;
            int blocks = (rows + 7) / 8;
            if (blocks > sm * 8) blocks = sm * 8;
            wide_residual_kernel<<<blocks, 256, 0, st>>>(X, p.seg_y[sg] + r0, p.beta, rows, p.d, p.ld, p.coeff[sg],
                                                         p.model, r, skp);
            dim3 grid((p.ld / 4 + 255) / 256, (rows + 255) / 256);
            wide_xtr_kernel<<<grid, 256, 0, st>>>(X, r, rows, p.ld, g, skp);
        }
    }
    wide_ship_kernel<<<(p.d + 1023) / 1024 > 592 ? 592 : (p.d + 1023) / 1024, 256, 0, st>>>(g, p.out, p.d, skp);
    if (p.flag) wide_flag_kernel<<<1, 32, 0, st>>>(p.flag, p.stamp, skp);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

}  // namespace eh
