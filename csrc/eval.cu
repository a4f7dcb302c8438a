// Evaluation kernels: batched margins for all stored iterates, column-wise loss, tie-aware AUC.
//
// Reference: the evaluation epilogue of every driver (naive.py:154-209) loops over the R stored
// iterates and does two matvecs + loss + sklearn roc_curve/auc per iterate on the master.  Here
// all iterates are evaluated at once: P = X . [beta_0 .. beta_{R-1}] is one GEMM-shaped pass
// over X (CUDA-core tiled version in this file; the tcgen05/TMEM version lives in
// eval_tc.cu), the loss of every column is reduced in one kernel (util.py:136-141), and AUC is
// a device radix sort + scans reproducing sklearn's trapezoid over distinct thresholds.
#include <cub/cub.cuh>

#include "common.cuh"

namespace eh {

// ---------------------------------------------------------------- P = X . Bt^T (fp32, tiled)
constexpr int MT = 64, NT = 64, KT = 16;

__global__ void __launch_bounds__(256) margins_dense_kernel(const float* __restrict__ X, long long rows, int d,
                                                            int ld, const float* __restrict__ Bt, int R, int ldb,
                                                            float* __restrict__ P) {
    __shared__ float As[KT][MT + 4];
    __shared__ float Bs[KT][NT + 4];
    const int tid = threadIdx.x;
    const long long m0 = (long long)blockIdx.x * MT;
    const int n0 = blockIdx.y * NT;
    const int tx = tid & 15, ty = tid >> 4;          // 16 x 16 threads, 4 x 4 outputs each
    float acc[4][4] = {};
    for (int k0 = 0; k0 < d; k0 += KT) {
        // load A tile (64 rows x 16 k): 1024 elements, 4 per thread
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int e = tid + i * 256;
            const int rr = e >> 4, kk = e & 15;
            const long long gr = m0 + rr;
            As[kk][rr] = (gr < rows && k0 + kk < d) ? X[gr * ld + k0 + kk] : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int e = tid + i * 256;
            const int cc = e >> 4, kk = e & 15;
            const int gc = n0 + cc;
            Bs[kk][cc] = (gc < R && k0 + kk < d) ? Bt[(size_t)gc * ldb + k0 + kk] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < KT; ++kk) {
            float a[4], bb[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) bb[j] = Bs[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const long long gr = m0 + ty * 4 + i;
        if (gr >= rows) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gc = n0 + tx * 4 + j;
            if (gc < R) P[gr * R + gc] = acc[i][j];
        }
    }
}

// one-hot rows: P[r, c] = sum_k Bt[c, idx[r,k]]
__global__ void __launch_bounds__(256) margins_ell_kernel(const int32_t* __restrict__ idx, long long rows, int nnz,
                                                          const float* __restrict__ Bt, int R, int ldb,
                                                          float* __restrict__ P) {
    const long long row = blockIdx.x;
    if (row >= rows) return;
    const int32_t* ri = idx + row * nnz;
    for (int c = threadIdx.x; c < R; c += blockDim.x) {
        float s = 0.f;
        for (int k = 0; k < nnz; ++k) s += __ldg(Bt + (size_t)c * ldb + __ldg(ri + k));
        P[row * R + c] = s;
    }
}

// ---------------------------------------------------------------- column-wise loss
__global__ void __launch_bounds__(256) loss_columns_kernel(const float* __restrict__ P, const float* __restrict__ y,
                                                           long long rows, int R, int kind, double* __restrict__ out) {
    const long long chunk = 512;
    const long long r0 = (long long)blockIdx.x * chunk;
    const long long r1 = (r0 + chunk < rows) ? r0 + chunk : rows;
    for (int c = threadIdx.x; c < R; c += blockDim.x) {
        double s = 0.0;
        for (long long r = r0; r < r1; ++r) {
            const float p = P[r * R + c], yy = __ldg(y + r);
            if (kind == 0) {
                // log(1 + exp(-y p)) evaluated stably (util.py:136-137)
                const float z = -yy * p;
                s += (double)(fmaxf(z, 0.f) + log1pf(__expf(-fabsf(z))));
            } else {
                const float e = yy - p;
                s += (double)e * (double)e;
            }
        }
        atomicAdd(out + c, s);
    }
}

// ---------------------------------------------------------------- ROC-AUC
__global__ void auc_prepare_kernel(const float* __restrict__ s, const float* __restrict__ lab, long long n,
                                   int* __restrict__ is_pos, long long* __restrict__ gstart) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    is_pos[i] = lab[i] > 0.f ? 1 : 0;
    gstart[i] = (i == 0 || s[i - 1] != s[i]) ? i : 0;      // tie-group start marker for a max-scan
}

__global__ void __launch_bounds__(256) auc_area_kernel(const float* __restrict__ s, const long long* __restrict__ cumpos,
                                                       const long long* __restrict__ gstart, long long n,
                                                       double* __restrict__ area) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    double a = 0.0;
    if (i < n && (i == n - 1 || s[i] != s[i + 1])) {     // i closes a tie group = one ROC vertex
        const long long gs = gstart[i];
        const long long tp_prev = gs > 0 ? cumpos[gs - 1] : 0;
        const long long fp_prev = gs - tp_prev;
        const long long tp = cumpos[i];
        const long long fp = (i + 1) - tp;
        a = (double)(fp - fp_prev) * 0.5 * (double)(tp + tp_prev);
    }
    a = warp_sum(a);
    __shared__ double sh[8];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += sh[w];
        if (t != 0.0) atomicAdd(area, t);
    }
}

__global__ void auc_finish_kernel(const long long* __restrict__ cumpos, long long n, const double* __restrict__ area,
                                  double* __restrict__ out) {
    const double Pn = (double)cumpos[n - 1];
    const double Nn = (double)n - Pn;
    *out = (Pn > 0.0 && Nn > 0.0) ? (*area) / (Pn * Nn) : nan("");
}

struct AucLayout {
    size_t off_keys, off_vals, off_pos, off_cum, off_gs, off_area, off_tmp, tmp_bytes, total;
};

static AucLayout auc_layout(long long n) {
    AucLayout L;
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    size_t off = 0;
    L.off_keys = off; off += up(n * sizeof(float));
    L.off_vals = off; off += up(n * sizeof(float));
    L.off_pos = off;  off += up(n * sizeof(int));
    L.off_cum = off;  off += up(n * sizeof(long long));
    L.off_gs = off;   off += up(n * sizeof(long long));
    L.off_area = off; off += 256;
    size_t t1 = 0, t2 = 0, t3 = 0;
    cub::DeviceRadixSort::SortPairsDescending(nullptr, t1, (const float*)nullptr, (float*)nullptr, (const float*)nullptr,
                                              (float*)nullptr, (int)n);
    cub::DeviceScan::InclusiveSum(nullptr, t2, (const int*)nullptr, (long long*)nullptr, (int)n);
    cub::DeviceScan::InclusiveScan(nullptr, t3, (const long long*)nullptr, (long long*)nullptr, cub::Max(), (int)n);
    L.tmp_bytes = t1 > t2 ? t1 : t2;
    if (t3 > L.tmp_bytes) L.tmp_bytes = t3;
    L.off_tmp = off;
    off += up(L.tmp_bytes);
    L.total = off;
    return L;
}

// ---------------------------------------------------------------- ROC-AUC of R score columns in one pass
// One device-wide radix sort of 64-bit keys (column id << 32 | descending-order bits of the score) replaces R
// separate sorts; the positive counts and tie-group starts are two scans over the whole array, and every column
// subtracts the running total at its own start.  Same tie-aware trapezoid as above (sklearn roc_curve + auc,
// naive.py:196-197), one launch sequence for all iterates of the evaluation epilogue.
__device__ __forceinline__ uint32_t desc_bits(float f) {
    if (f == 0.f) f = 0.f;                                   // -0 and +0 tie
    uint32_t u = __float_as_uint(f);
    u ^= (u >> 31) ? 0xffffffffu : 0x80000000u;              // ascending order-preserving
    return ~u;                                               // descending
}
__global__ void aucb_keys_kernel(const float* __restrict__ S, const float* __restrict__ lab, long long n, int R,
                                 unsigned long long* __restrict__ keys, unsigned char* __restrict__ vals) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * R) return;
    const long long seg = i / n, j = i - seg * n;
    keys[i] = ((unsigned long long)seg << 32) | desc_bits(S[i]);
    vals[i] = lab[j] > 0.f ? 1 : 0;
}
__global__ void aucb_prepare_kernel(const unsigned long long* __restrict__ keys, const unsigned char* __restrict__ vals,
                                    long long n, long long total, int* __restrict__ is_pos, long long* __restrict__ gstart) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    is_pos[i] = vals[i];
    gstart[i] = (i == 0 || keys[i - 1] != keys[i]) ? i : 0;   // a new column is a new key: also a group start
}
__global__ void __launch_bounds__(256) aucb_area_kernel(const unsigned long long* __restrict__ keys,
                                                        const long long* __restrict__ cumpos,
                                                        const long long* __restrict__ gstart, long long n,
                                                        double* __restrict__ area) {
    const long long seg = blockIdx.y, base = seg * n;
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    double a = 0.0;
    if (j < n) {
        const long long i = base + j;
        if (j == n - 1 || keys[i] != keys[i + 1]) {          // i closes a tie group = one ROC vertex
            const long long cb = base > 0 ? cumpos[base - 1] : 0;
            const long long gs = gstart[i];
            const long long tp_prev = gs > base ? cumpos[gs - 1] - cb : 0;
            const long long fp_prev = (gs - base) - tp_prev;
            const long long tp = cumpos[i] - cb;
            const long long fp = (j + 1) - tp;
            a = (double)(fp - fp_prev) * 0.5 * (double)(tp + tp_prev);
        }
    }
    a = warp_sum(a);
    __shared__ double sh[8];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = a;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += sh[w];
        if (t != 0.0) atomicAdd(area + seg, t);
    }
}
__global__ void aucb_finish_kernel(const long long* __restrict__ cumpos, long long n, int R, const double* __restrict__ area,
                                   double* __restrict__ out) {
    const int seg = blockIdx.x * blockDim.x + threadIdx.x;
    if (seg >= R) return;
    const long long base = (long long)seg * n;
    const double Pn = (double)(cumpos[base + n - 1] - (base > 0 ? cumpos[base - 1] : 0));
    const double Nn = (double)n - Pn;
    out[seg] = (Pn > 0.0 && Nn > 0.0) ? area[seg] / (Pn * Nn) : nan("");
}

struct AucBatchLayout {
    size_t off_k0, off_k1, off_v0, off_v1, off_pos, off_cum, off_gs, off_area, off_tmp, tmp_bytes, total;
};
static AucBatchLayout aucb_layout(long long n, int R) {
    AucBatchLayout L;
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    const size_t T = (size_t)n * R;
    size_t off = 0;
    L.off_k0 = off; off += up(T * 8);
    L.off_k1 = off; off += up(T * 8);
    L.off_v0 = off; off += up(T);
    L.off_v1 = off; off += up(T);
    L.off_pos = off; off += up(T * sizeof(int));
    L.off_cum = off; off += up(T * sizeof(long long));
    L.off_gs = off;  off += up(T * sizeof(long long));
    L.off_area = off; off += up((size_t)R * sizeof(double));
    size_t t1 = 0, t2 = 0, t3 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, t1, (const unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                    (const unsigned char*)nullptr, (unsigned char*)nullptr, (long long)T);
    cub::DeviceScan::InclusiveSum(nullptr, t2, (const int*)nullptr, (long long*)nullptr, (long long)T);
    cub::DeviceScan::InclusiveScan(nullptr, t3, (const long long*)nullptr, (long long*)nullptr, cub::Max(), (long long)T);
    L.tmp_bytes = t1 > t2 ? t1 : t2;
    if (t3 > L.tmp_bytes) L.tmp_bytes = t3;
    L.off_tmp = off;
    off += up(L.tmp_bytes);
    L.total = off;
    return L;
}

}  // namespace eh

extern "C" size_t eh_auc_batch_scratch_bytes(int64_t n, int R) { return (n > 0 && R > 0) ? eh::aucb_layout(n, R).total : 0; }

extern "C" int eh_auc_batch(const float* d_scores, const float* d_labels, int64_t n, int R, double* d_auc, void* d_scratch,
                            size_t scratch_bytes, void* stream) {
    using namespace eh;
    EH_REQUIRE(d_scores && d_labels && d_auc && d_scratch && n > 0 && R > 0, "auc_batch: bad argument");
    EH_REQUIRE(n < (1ll << 31) && (long long)n * R < (1ll << 40), "auc_batch: too many scores");
    const AucBatchLayout L = aucb_layout(n, R);
    EH_REQUIRE(scratch_bytes >= L.total, "auc_batch: scratch too small (%zu < %zu)", scratch_bytes, L.total);
    char* base = (char*)d_scratch;
    unsigned long long* k0 = (unsigned long long*)(base + L.off_k0);
    unsigned long long* k1 = (unsigned long long*)(base + L.off_k1);
    unsigned char* v0 = (unsigned char*)(base + L.off_v0);
    unsigned char* v1 = (unsigned char*)(base + L.off_v1);
    int* pos = (int*)(base + L.off_pos);
    long long* cum = (long long*)(base + L.off_cum);
    long long* gs = (long long*)(base + L.off_gs);
    double* area = (double*)(base + L.off_area);
    void* tmp = base + L.off_tmp;
    cudaStream_t st = (cudaStream_t)stream;
    const long long T = (long long)n * R;
    const unsigned gridT = (unsigned)((T + 255) / 256);
    aucb_keys_kernel<<<gridT, 256, 0, st>>>(d_scores, d_labels, n, R, k0, v0);
    int seg_bits = 1;
    while ((1ll << seg_bits) < R) ++seg_bits;
    size_t tb = L.tmp_bytes;
    EH_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tb, k0, k1, v0, v1, T, 0, 32 + seg_bits, st));
    aucb_prepare_kernel<<<gridT, 256, 0, st>>>(k1, v1, n, T, pos, gs);
    tb = L.tmp_bytes;
    EH_CUDA(cub::DeviceScan::InclusiveSum(tmp, tb, pos, cum, T, st));
    tb = L.tmp_bytes;
    EH_CUDA(cub::DeviceScan::InclusiveScan(tmp, tb, gs, gs, cub::Max(), T, st));
    EH_CUDA(cudaMemsetAsync(area, 0, (size_t)R * sizeof(double), st));
    dim3 grid((unsigned)((n + 255) / 256), (unsigned)R);
    aucb_area_kernel<<<grid, 256, 0, st>>>(k1, cum, gs, n, area);
    aucb_finish_kernel<<<(R + 127) / 128, 128, 0, st>>>(cum, n, R, area, d_auc);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

extern "C" int eh_margins_dense(const float* d_X, int64_t rows, int d, int ld, const float* d_Bt, int R, int ldb,
                                float* d_P, void* stream) {
    using namespace eh;
    EH_REQUIRE(d_X && d_Bt && d_P && rows > 0 && d > 0 && R > 0, "margins_dense: bad argument");
    dim3 grid((unsigned)((rows + MT - 1) / MT), (unsigned)((R + NT - 1) / NT));
    margins_dense_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d_X, rows, d, ld, d_Bt, R, ldb, d_P);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

extern "C" int eh_margins_ell(const int32_t* d_idx, int64_t rows, int nnz, const float* d_Bt, int R, int ldb,
                              float* d_P, void* stream) {
    using namespace eh;
    EH_REQUIRE(d_idx && d_Bt && d_P && rows > 0 && nnz > 0 && R > 0, "margins_ell: bad argument");
    EH_REQUIRE(rows < (1ll << 31), "margins_ell: too many rows for one launch");
    margins_ell_kernel<<<(unsigned)rows, 128, 0, (cudaStream_t)stream>>>(d_idx, rows, nnz, d_Bt, R, ldb, d_P);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

extern "C" int eh_loss_columns(const float* d_P, const float* d_y, int64_t rows, int R, int kind, double* d_out,
                               void* stream) {
    using namespace eh;
    EH_REQUIRE(d_P && d_y && d_out && rows > 0 && R > 0, "loss_columns: bad argument");
    const unsigned grid = (unsigned)((rows + 511) / 512);
    loss_columns_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d_P, d_y, rows, R, kind, d_out);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}

extern "C" size_t eh_auc_scratch_bytes(int64_t n) { return n > 0 ? eh::auc_layout(n).total : 0; }

extern "C" int eh_auc(const float* d_scores, const float* d_labels, int64_t n, double* d_auc, void* d_scratch,
                      size_t scratch_bytes, void* stream) {
    using namespace eh;
    EH_REQUIRE(d_scores && d_labels && d_auc && d_scratch && n > 0, "auc: bad argument");
    EH_REQUIRE(n < (1ll << 31), "auc: n too large");
    const AucLayout L = auc_layout(n);
    EH_REQUIRE(scratch_bytes >= L.total, "auc: scratch too small (%zu < %zu)", scratch_bytes, L.total);
    char* base = (char*)d_scratch;
    float* keys = (float*)(base + L.off_keys);
    float* vals = (float*)(base + L.off_vals);
    int* pos = (int*)(base + L.off_pos);
    long long* cum = (long long*)(base + L.off_cum);
    long long* gs = (long long*)(base + L.off_gs);
    double* area = (double*)(base + L.off_area);
    void* tmp = base + L.off_tmp;
    size_t tb = L.tmp_bytes;
    cudaStream_t st = (cudaStream_t)stream;
    EH_CUDA(cub::DeviceRadixSort::SortPairsDescending(tmp, tb, d_scores, keys, d_labels, vals, (int)n, 0, 32, st));
    const unsigned grid = (unsigned)((n + 255) / 256);
    auc_prepare_kernel<<<grid, 256, 0, st>>>(keys, vals, n, pos, gs);
    tb = L.tmp_bytes;
    EH_CUDA(cub::DeviceScan::InclusiveSum(tmp, tb, pos, cum, (int)n, st));
    tb = L.tmp_bytes;
    EH_CUDA(cub::DeviceScan::InclusiveScan(tmp, tb, gs, gs, cub::Max(), (int)n, st));
    EH_CUDA(cudaMemsetAsync(area, 0, sizeof(double), st));
    auc_area_kernel<<<grid, 256, 0, st>>>(keys, cum, gs, n, area);
    auc_finish_kernel<<<1, 1, 0, st>>>(cum, n, area, d_auc);
    EH_CUDA(cudaGetLastError());
    return EH_OK;
}
