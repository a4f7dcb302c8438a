// Code-matrix generation on the device (ops K8/K9 of SURVEY 2.4).
//
// eh_code_cyclic: the cyclic-MDS encode matrix B from the parity matrix H, semantics of getB
// (util.py:64-83): row i is supported on {i..i+s} mod n, B[i,i] = 1 and the other s entries are
// -H[:,tail]^{-1} H[:,i].  One warp per row solves its s x s system by pivoted elimination.
// eh_code_decode_table: one decode row per straggler pattern, semantics of getA (util.py:85-103);
// one warp per pattern reuses the decoder's elimination on the n x (n-s) consistent system.
#include "solve.cuh"

namespace eh {

constexpr int CG_MAX_S = 32;

__global__ void __launch_bounds__(32) code_cyclic_kernel(const double* __restrict__ H, int n, int s,
                                                         double* __restrict__ B) {
    __shared__ double M[CG_MAX_S][CG_MAX_S + 1];
    const int i = blockIdx.x, lane = threadIdx.x;
    for (int j = lane; j < n; j += 32) B[(size_t)i * n + j] = 0.0;
    __syncwarp();
    if (s == 0) {
        if (lane == 0) B[(size_t)i * n + i] = 1.0;
        return;
    }
    // M = [H[:, tail] | H[:, head]],  tail_c = (i + 1 + c) mod n
    for (int r = lane; r < s; r += 32) {
        for (int c = 0; c < s; ++c) M[r][c] = H[(size_t)r * n + (i + 1 + c) % n];
        M[r][s] = H[(size_t)r * n + i];
    }
    __syncwarp();
    for (int c = 0; c < s; ++c) {
        double best = -1.0;
        int bi = c;
        for (int r = c + lane; r < s; r += 32) {
            const double v = fabs(M[r][c]);
            if (v > best) { best = v; bi = r; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (bi != c)
            for (int col = lane; col <= s; col += 32) {
                const double t = M[c][col];
                M[c][col] = M[bi][col];
                M[bi][col] = t;
            }
        __syncwarp();
        const double piv = M[c][c];
        for (int r = lane; r < s; r += 32) {
            if (r == c) continue;
            const double f = M[r][c] / piv;
            for (int col = c; col <= s; ++col) M[r][col] -= f * M[c][col];
        }
        __syncwarp();
    }
    for (int c = lane; c < s; c += 32) B[(size_t)i * n + (i + 1 + c) % n] = -M[c][s] / M[c][c];
    if (lane == 0) B[(size_t)i * n + i] = 1.0;
}

__global__ void __launch_bounds__(32) decode_table_kernel(const double* __restrict__ B, int n, int s,
                                                          const int32_t* __restrict__ missing, long long n_patterns,
                                                          double* __restrict__ A) {
    __shared__ double M[EH_MAX_WORKERS][EH_MAX_WORKERS + 1];
    __shared__ double a[EH_MAX_WORKERS];
    __shared__ int S[EH_MAX_WORKERS];
    const int lane = threadIdx.x;
    for (long long pat = blockIdx.x; pat < n_patterns; pat += gridDim.x) {
        unsigned long long miss = 0;
        for (int q = 0; q < s; ++q) miss |= 1ull << missing[pat * s + q];
        if (lane == 0) {
            int k = 0;
            for (int w = 0; w < n; ++w)
                if (!((miss >> w) & 1ull)) S[k++] = w;
        }
        __syncwarp();
        const int k = n - s;
        solve_decode_row(B, n, S, k, M, a, lane);
        for (int w = lane; w < n; w += 32) A[pat * n + w] = 0.0;
        __syncwarp();
        for (int c = lane; c < k; c += 32) A[pat * n + S[c]] = a[c];
        __syncwarp();
    }
}

}  // namespace eh

extern "C" int eh_code_cyclic(const double* d_H, int n, int s, double* d_B, void* stream) {
    using namespace eh;
    EH_REQUIRE(d_B && n >= 1 && s >= 0 && s < n, "code_cyclic: bad argument");
    EH_REQUIRE(s == 0 || d_H, "code_cyclic: H required for s > 0");
    EH_REQUIRE(s <= CG_MAX_S, "code_cyclic: s=%d > %d", s, CG_MAX_S);
    code_cyclic_kernel<<<n, 32, 0, (cudaStream_t)stream>>>(d_H, n, s, d_B);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

extern "C" int eh_code_decode_table(const double* d_B, int n, int s, const int32_t* d_missing, int64_t n_patterns,
                                    double* d_A, void* stream) {
    using namespace eh;
    EH_REQUIRE(d_B && d_A && n >= 1 && n <= EH_MAX_WORKERS && s >= 0 && s < n && n_patterns > 0,
               "decode_table: bad argument");
    EH_REQUIRE(s == 0 || d_missing, "decode_table: missing-worker list required");
    long long grid = n_patterns < 4096 ? n_patterns : 4096;
    decode_table_kernel<<<(unsigned)grid, 32, 0, (cudaStream_t)stream>>>(d_B, n, s, d_missing, n_patterns, d_A);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}
