// Small dense solvers shared by the decoder kernel and the code-generation kernels.
#pragma once
#include "common.cuh"

namespace eh {

// Solve  sum_c a_c * B[S_c, :] = 1  for the k = |S| selected workers by Gauss-Jordan elimination
// with row pivoting on the n x (k+1) augmented system (consistent by construction of B).
__device__ inline void solve_decode_row(const double* __restrict__ B, int n, const int* S, int k,
                                 double (*M)[EH_MAX_WORKERS + 1], double* a_out, int lane) {
    for (int j = lane; j < n; j += 32) {
        for (int c = 0; c < k; ++c) M[j][c] = B[(size_t)S[c] * n + j];
        M[j][k] = 1.0;
    }
    __syncwarp();
    for (int c = 0; c < k; ++c) {
        // pivot search over rows >= c
        double best = -1.0;
        int bi = c;
        for (int j = c + lane; j < n; j += 32) {
            const double v = fabs(M[j][c]);
            if (v > best) { best = v; bi = j; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        if (bi != c) {
            for (int col = lane; col <= k; col += 32) {
                const double t = M[c][col];
                M[c][col] = M[bi][col];
                M[bi][col] = t;
            }
        }
        __syncwarp();
        const double piv = M[c][c];
        if (fabs(piv) > 0.0) {
            for (int j = lane; j < n; j += 32) {
                if (j == c) continue;
                const double f = M[j][c] / piv;
                if (f != 0.0)
                    for (int col = c; col <= k; ++col) M[j][col] -= f * M[c][col];
            }
        }
        __syncwarp();
    }
    for (int c = lane; c < k; c += 32) {
        const double piv = M[c][c];
        a_out[c] = (fabs(piv) > 0.0) ? M[c][k] / piv : 0.0;
    }
    __syncwarp();
}

__device__ inline long long combinadic_rank(unsigned long long done, int n, const int64_t* __restrict__ binom) {
    // calculate_indexA (util.py:125-134): scan j = n-1 .. 0, ctr counts set bits seen so far
    long long ind = 0;
    int ctr = 0;
    for (int j = n - 1; j >= 0; --j) {
        if ((done >> j) & 1ull) {
            ctr += 1;
            ind += binom[(size_t)(n - 1 - j) * (n + 2) + ctr];
        }
    }
    return ind;
}


}  // namespace eh
