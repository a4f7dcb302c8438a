// One-hot feature construction on the device (op K15 of SURVEY 2.4).
//
// Reference: sklearn OneHotEncoder -> scipy CSR (arrange_real_data.py:61-64) and the hashed
// pair features of util.py:49-55.  On the GPU a one-hot row with one active entry per source
// column is just an index vector: idx[r, c] = offset[c] + code[r, c]; pair features are the
// mixed-radix combination of two label-encoded columns.
#include "common.cuh"

namespace eh {

__global__ void onehot_assemble_kernel(const long long* __restrict__ codes, const long long* __restrict__ offsets,
                                       long long rows, int cols, int32_t* __restrict__ idx) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * cols) return;
    const int c = (int)(i % cols);
    idx[i] = (int32_t)(codes[i] + offsets[c]);
}

__global__ void pair_codes_kernel(const long long* __restrict__ codes, long long rows, int cols,
                                  const int32_t* __restrict__ pairs, int n_pairs,
                                  const long long* __restrict__ radix, long long* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * n_pairs) return;
    const long long r = i / n_pairs;
    const int q = (int)(i % n_pairs);
    const int a = pairs[2 * q], b = pairs[2 * q + 1];
    out[i] = codes[r * cols + a] * radix[b] + codes[r * cols + b];
}

}  // namespace eh

extern "C" int eh_onehot_assemble(const long long* d_codes, const long long* d_offsets, int64_t rows, int cols,
                                  int32_t* d_idx, void* stream) {
    using namespace eh;
    EH_REQUIRE(d_codes && d_offsets && d_idx && rows > 0 && cols > 0, "onehot_assemble: bad argument");
    const long long n = rows * cols;
    onehot_assemble_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_codes, d_offsets, rows, cols,
                                                                                         d_idx);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

extern "C" int eh_pair_codes(const long long* d_codes, int64_t rows, int cols, const int32_t* d_pairs, int n_pairs,
                             const long long* d_radix, long long* d_out, void* stream) {
    using namespace eh;
    EH_REQUIRE(d_codes && d_pairs && d_radix && d_out && rows > 0 && cols > 0 && n_pairs > 0, "pair_codes: bad argument");
    const long long n = rows * n_pairs;
    pair_codes_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_codes, rows, cols, d_pairs, n_pairs,
                                                                                    d_radix, d_out);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}
