// csr.cuh -- sparse (CSR) primitives of the materialized-Jacobian mode, as CUDA kernels.
//
// Replaces the cuSPARSE wrappers of the reference's CUDA backend (API/src/backend_cuda.t:541-654: cusparseXcsrgemmNnz,
// cusparseScsrgemm, cusparseScsr2csc, cusparseScsrmv -- entry points that no longer exist in current cuSPARSE) and follows
// the contract of its CPU library (API/src/linalg_cpu.t:203-617), whose known-answer test (API/src/linalg_cpu_test.t:20-190)
// is the golden of tests/test_csr_gpu.py:
//   computeNnzPatternAT / computeAT   -> csr_transpose_pattern (stable radix sort of the column indices; the permutation is
//                                        kept, so refreshing the VALUES of A^T every nonlinear iteration is one gather)
//   computeNnzPatternATA              -> csr_ata_pattern (expand every row's column pairs, sort, unique: rows and columns of
//                                        A^T A come out sorted, like the reference's)
//   computeATA                        -> csr_ata_values (one thread per entry (i, j): sum over the rows r of column i of
//                                        A[r][i] * A[r][j], r ascending -- deterministic)
//   applyAtoVector                    -> csr_spmv (CSR-vector: 2^k lanes per row chosen from the mean row length; the rows of
//                                        J have 2-6 entries, those of J^T J a few dozen)
// Every kernel is bandwidth-bound integer / float streaming; no tensor cores.
#pragma once
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include "common.cuh"

namespace optk {

template <class T>
struct CsrMat {
    int nrows = 0, ncols = 0;
    long nnz = 0;
    int* rowPtr = nullptr;      // nrows + 1
    int* colInd = nullptr;      // nnz, ascending inside a row
    T* val = nullptr;           // nnz
    void release() { cudaFree(rowPtr); cudaFree(colInd); cudaFree(val); rowPtr = colInd = nullptr; val = nullptr; nnz = 0; }
};

// temporary device storage that grows on demand (cub temp storage and sort buffers)
struct Scratch {
    void* p = nullptr; size_t cap = 0;
    void* get(size_t bytes) {
        if (bytes > cap) { cudaFree(p); OPTK_CUDA(cudaMalloc(&p, bytes)); cap = bytes; }
        return p;
    }
    void release() { cudaFree(p); p = nullptr; cap = 0; }
};

static __global__ void k_csr_row_of(int nrows, const int* __restrict__ rowPtr, int* __restrict__ rowOf) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrows) return;
    for (int k = rowPtr[r]; k < rowPtr[r + 1]; ++k) rowOf[k] = r;
}
static __global__ void k_iota(long n, int* v) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i < n) v[i] = (int)i;
}
// ptr[k] = first position whose (sorted) key is >= k, for k in [0, nkeys]
template <class K>
static __global__ void k_offsets_from_sorted(const K* __restrict__ keys, long n, long nkeys, int shift, int* ptr) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i > n) return;
    const long lo = i == 0 ? 0 : (long)(keys[i - 1] >> shift) + 1;
    const long hi = i == n ? nkeys : (long)(keys[i] >> shift);
    for (long k = lo; k <= hi; ++k) ptr[k] = (int)i;
}
static __global__ void k_gather_int(long n, const int* __restrict__ src, const int* __restrict__ perm, int* __restrict__ dst) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[perm[i]];
}
template <class T>
static __global__ void k_gather_val(long n, const T* __restrict__ src, const int* __restrict__ perm, T* __restrict__ dst) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[perm[i]];
}

inline unsigned blocks_for(long n, int t = 256) { return (unsigned)((n + t - 1) / t > 0 ? (n + t - 1) / t : 1); }
inline int bits_for(long n) { int b = 1; while ((1L << b) < n) ++b; return b; }

// Pattern of A^T (rowPtr, colInd allocated here) and the permutation perm[k] = position in A of the k-th entry of A^T.
template <class T>
void csr_transpose_pattern(const CsrMat<T>& A, CsrMat<T>& AT, int** perm_out, Scratch& tmp, cudaStream_t st) {
    AT.nrows = A.ncols; AT.ncols = A.nrows; AT.nnz = A.nnz;
    const long nnz = A.nnz;
    OPTK_CUDA(cudaMalloc(&AT.rowPtr, (size_t)(AT.nrows + 1) * sizeof(int)));
    OPTK_CUDA(cudaMalloc(&AT.colInd, (size_t)std::max<long>(nnz, 1) * sizeof(int)));
    OPTK_CUDA(cudaMalloc(&AT.val, (size_t)std::max<long>(nnz, 1) * sizeof(T)));
    int* perm = nullptr;
    OPTK_CUDA(cudaMalloc(&perm, (size_t)std::max<long>(nnz, 1) * sizeof(int)));
    int *rowOf = nullptr, *iota = nullptr, *keys = nullptr;
    OPTK_CUDA(cudaMalloc(&rowOf, (size_t)std::max<long>(nnz, 1) * sizeof(int)));
    OPTK_CUDA(cudaMalloc(&iota, (size_t)std::max<long>(nnz, 1) * sizeof(int)));
    OPTK_CUDA(cudaMalloc(&keys, (size_t)std::max<long>(nnz, 1) * sizeof(int)));
    if (nnz > 0) {
        k_csr_row_of<<<blocks_for(A.nrows), 256, 0, st>>>(A.nrows, A.rowPtr, rowOf);
        k_iota<<<blocks_for(nnz), 256, 0, st>>>(nnz, iota);
        size_t bytes = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, bytes, A.colInd, keys, iota, perm, (int)nnz, 0, bits_for(A.ncols), st);
        cub::DeviceRadixSort::SortPairs(tmp.get(bytes), bytes, A.colInd, keys, iota, perm, (int)nnz, 0, bits_for(A.ncols), st);   // stable: rows ascending inside a column
        k_gather_int<<<blocks_for(nnz), 256, 0, st>>>(nnz, rowOf, perm, AT.colInd);
    }
    k_offsets_from_sorted<int><<<blocks_for(nnz + 1), 256, 0, st>>>(keys, nnz, (long)AT.nrows, 0, AT.rowPtr);
    OPTK_CUDA(cudaStreamSynchronize(st));
    cudaFree(rowOf); cudaFree(iota); cudaFree(keys);
    *perm_out = perm;
}
template <class T>
void csr_transpose_values(const CsrMat<T>& A, CsrMat<T>& AT, const int* perm, cudaStream_t st) {
    if (A.nnz > 0) k_gather_val<T><<<blocks_for(A.nnz), 256, 0, st>>>(A.nnz, A.val, perm, AT.val);
}

// ---- A^T A: pattern by expand / sort / unique ---------------------------------------------------------------
static __global__ void k_ata_pair_counts(int nrows, const int* __restrict__ rowPtr, unsigned long long* __restrict__ cnt) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrows) return;
    const unsigned long long n = (unsigned long long)(rowPtr[r + 1] - rowPtr[r]);
    cnt[r] = n * n;
}
static __global__ void k_ata_expand(int nrows, const int* __restrict__ rowPtr, const int* __restrict__ colInd,
                                    const unsigned long long* __restrict__ base, unsigned long long* __restrict__ keys) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrows) return;
    const int b = rowPtr[r], e = rowPtr[r + 1];
    unsigned long long o = base[r];
    for (int a = b; a < e; ++a)
        for (int c = b; c < e; ++c) keys[o++] = ((unsigned long long)(unsigned)colInd[a] << 32) | (unsigned)colInd[c];
}
static __global__ void k_ata_split(long n, const unsigned long long* __restrict__ keys, int* __restrict__ colInd, int* __restrict__ rowOf) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    colInd[i] = (int)(keys[i] & 0xffffffffull);
    rowOf[i] = (int)(keys[i] >> 32);
}
// rowOf_out: row index of every entry of A^T A (used by csr_ata_values)
template <class T>
void csr_ata_pattern(const CsrMat<T>& A, CsrMat<T>& ATA, int** rowOf_out, Scratch& tmp, cudaStream_t st) {
    ATA.nrows = ATA.ncols = A.ncols;
    OPTK_CUDA(cudaMalloc(&ATA.rowPtr, (size_t)(ATA.nrows + 1) * sizeof(int)));
    unsigned long long *cnt = nullptr, *base = nullptr;
    OPTK_CUDA(cudaMalloc(&cnt, (size_t)(A.nrows + 1) * sizeof(unsigned long long)));
    OPTK_CUDA(cudaMalloc(&base, (size_t)(A.nrows + 1) * sizeof(unsigned long long)));
    OPTK_CUDA(cudaMemsetAsync(cnt, 0, (size_t)(A.nrows + 1) * sizeof(unsigned long long), st));
    k_ata_pair_counts<<<blocks_for(A.nrows), 256, 0, st>>>(A.nrows, A.rowPtr, cnt);
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, cnt, base, A.nrows + 1, st);
    cub::DeviceScan::ExclusiveSum(tmp.get(bytes), bytes, cnt, base, A.nrows + 1, st);
    unsigned long long total = 0;
    OPTK_CUDA(cudaMemcpyAsync(&total, base + A.nrows, sizeof(total), cudaMemcpyDeviceToHost, st));
    OPTK_CUDA(cudaStreamSynchronize(st));
    if (total >= (1ull << 31)) optk_fatal(1, "Opt: J^T J pattern needs %llu candidate pairs (more than 2^31): use the unfused materialized mode\n", total);
    unsigned long long *ka = nullptr, *kb = nullptr;
    OPTK_CUDA(cudaMalloc(&ka, (size_t)std::max<unsigned long long>(total, 1) * sizeof(unsigned long long)));
    OPTK_CUDA(cudaMalloc(&kb, (size_t)std::max<unsigned long long>(total, 1) * sizeof(unsigned long long)));
    int* nuniq = nullptr;
    OPTK_CUDA(cudaMalloc(&nuniq, sizeof(int)));
    int h_n = 0;
    if (total > 0) {
        k_ata_expand<<<blocks_for(A.nrows), 256, 0, st>>>(A.nrows, A.rowPtr, A.colInd, base, ka);
        const int bits = 32 + bits_for(A.ncols);
        cub::DeviceRadixSort::SortKeys(nullptr, bytes, ka, kb, (int)total, 0, bits, st);
        cub::DeviceRadixSort::SortKeys(tmp.get(bytes), bytes, ka, kb, (int)total, 0, bits, st);
        cub::DeviceSelect::Unique(nullptr, bytes, kb, ka, nuniq, (int)total, st);
        cub::DeviceSelect::Unique(tmp.get(bytes), bytes, kb, ka, nuniq, (int)total, st);
        OPTK_CUDA(cudaMemcpyAsync(&h_n, nuniq, sizeof(int), cudaMemcpyDeviceToHost, st));
        OPTK_CUDA(cudaStreamSynchronize(st));
    }
    ATA.nnz = h_n;
    int* rowOf = nullptr;
    OPTK_CUDA(cudaMalloc(&ATA.colInd, (size_t)std::max(h_n, 1) * sizeof(int)));
    OPTK_CUDA(cudaMalloc(&ATA.val, (size_t)std::max(h_n, 1) * sizeof(T)));
    OPTK_CUDA(cudaMalloc(&rowOf, (size_t)std::max(h_n, 1) * sizeof(int)));
    if (h_n > 0) k_ata_split<<<blocks_for(h_n), 256, 0, st>>>(h_n, ka, ATA.colInd, rowOf);
    k_offsets_from_sorted<unsigned long long><<<blocks_for((long)h_n + 1), 256, 0, st>>>(ka, (long)h_n, (long)ATA.nrows, 32, ATA.rowPtr);
    OPTK_CUDA(cudaStreamSynchronize(st));
    cudaFree(cnt); cudaFree(base); cudaFree(ka); cudaFree(kb); cudaFree(nuniq);
    *rowOf_out = rowOf;
}
// (A^T A)[i][j] = sum over the rows r that hold column i (= row i of A^T, ascending) of A[r][i] * A[r][j]
template <class T>
static __global__ void k_ata_values(long nnzATA, const int* __restrict__ rowOf, const int* __restrict__ colATA,
                                    const int* __restrict__ rowPtrAT, const int* __restrict__ colAT, const T* __restrict__ valAT,
                                    const int* __restrict__ rowPtrA, const int* __restrict__ colA, const T* __restrict__ valA,
                                    T* __restrict__ valATA) {
    const long k = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (k >= nnzATA) return;
    const int i = rowOf[k], j = colATA[k];
    T acc = T(0);
    for (int t = rowPtrAT[i]; t < rowPtrAT[i + 1]; ++t) {
        const int r = colAT[t];
        const T ari = valAT[t];
        for (int q = rowPtrA[r]; q < rowPtrA[r + 1]; ++q) {
            const int c = colA[q];
            if (c == j) { acc += ari * valA[q]; break; }
            if (c > j) break;
        }
    }
    valATA[k] = acc;
}
template <class T>
void csr_ata_values(const CsrMat<T>& A, const CsrMat<T>& AT, CsrMat<T>& ATA, const int* rowOf, cudaStream_t st) {
    if (ATA.nnz > 0)
        k_ata_values<T><<<blocks_for(ATA.nnz), 256, 0, st>>>(ATA.nnz, rowOf, ATA.colInd, AT.rowPtr, AT.colInd, AT.val, A.rowPtr, A.colInd, A.val, ATA.val);
}

// ---- y = A x, CSR-vector with L lanes per row ------------------------------------------------------------------
template <class T, int L>
static __global__ void __launch_bounds__(256) k_spmv(int nrows, const int* __restrict__ rowPtr, const int* __restrict__ colInd,
                                                     const T* __restrict__ val, const T* __restrict__ x, T* __restrict__ y) {
    constexpr int RPW = 32 / L;                                  // rows per warp
    const int lane = threadIdx.x & 31, sub = lane & (L - 1);
    const long warp = (blockIdx.x * (long)blockDim.x + threadIdx.x) >> 5, nwarps = ((long)gridDim.x * blockDim.x) >> 5;
    for (long r0 = warp * RPW; r0 < nrows; r0 += nwarps * RPW) {      // warp-uniform trip count: the shuffles below need every lane
        const long row = r0 + lane / L;
        T acc = T(0);
        if (row < nrows) {
            const int b = rowPtr[row], e = rowPtr[row + 1];
            for (int k = b + sub; k < e; k += L) acc = fma(val[k], x[colInd[k]], acc);
        }
#pragma unroll
        for (int o = L >> 1; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o, L);
        if (sub == 0 && row < nrows) y[row] = acc;
    }
}
template <class T>
void csr_spmv(const CsrMat<T>& A, const T* x, T* y, int n_sm, cudaStream_t st) {
    if (A.nrows == 0) return;
    const double mean = A.nrows > 0 ? (double)A.nnz / A.nrows : 0.0;
    int L = 1;
    while (L < 32 && L < mean) L <<= 1;
    const long threads = (long)A.nrows * L;
    const unsigned grid = (unsigned)std::min<long>((threads + 255) / 256, (long)n_sm * 16);
    switch (L) {
        case 1:  k_spmv<T, 1><<<grid, 256, 0, st>>>(A.nrows, A.rowPtr, A.colInd, A.val, x, y); break;
        case 2:  k_spmv<T, 2><<<grid, 256, 0, st>>>(A.nrows, A.rowPtr, A.colInd, A.val, x, y); break;
        case 4:  k_spmv<T, 4><<<grid, 256, 0, st>>>(A.nrows, A.rowPtr, A.colInd, A.val, x, y); break;
        case 8:  k_spmv<T, 8><<<grid, 256, 0, st>>>(A.nrows, A.rowPtr, A.colInd, A.val, x, y); break;
        case 16: k_spmv<T, 16><<<grid, 256, 0, st>>>(A.nrows, A.rowPtr, A.colInd, A.val, x, y); break;
        default: k_spmv<T, 32><<<grid, 256, 0, st>>>(A.nrows, A.rowPtr, A.colInd, A.val, x, y); break;
    }
}

// ---- Jacobian rows from a functor ------------------------------------------------------------------------------
// A functor with HAS_ROWS_J emits, for a site (rowsJ) or an edge (graphRowsJ), its residual rows in a fixed order
// through this buffer: begin(k) opens local row k, add(site, channel, value) appends dR/dX(site)[channel].  The STRUCTURE a
// functor emits depends on the grid bounds only (entries of residuals that a mask or a validity test switches off are
// emitted with value 0), so the pattern is built once per solve and only the values are refreshed (generateDumpJ,
// solver:385-442, writes a fixed number of entries per residual for the same reason).  Entries are sorted by column and
// duplicates merged when the row is closed.  COUNT: row lengths into cnt[]; otherwise columns / values at rowPtr[].
template <class T, class VecT, bool COUNT>
struct RowSink {
    static constexpr int CAP = 8;
    int* cnt_or_ptr; int* colInd; T* val; long row0; long plane;
    int k = -1, n = 0;
    int c[CAP]; T v[CAP];
    __device__ void close() {
        if (k < 0) return;
        for (int a = 1; a < n; ++a) {                       // insertion sort by column (n <= 8)
            const int ca = c[a]; const T va = v[a];
            int b = a - 1;
            while (b >= 0 && c[b] > ca) { c[b + 1] = c[b]; v[b + 1] = v[b]; --b; }
            c[b + 1] = ca; v[b + 1] = va;
        }
        int m = 0;
        for (int a = 0; a < n; ++a) {
            if (m > 0 && c[m - 1] == c[a]) v[m - 1] += v[a];
            else { c[m] = c[a]; v[m] = v[a]; ++m; }
        }
        if (COUNT) cnt_or_ptr[row0 + k] = m;
        else {
            const int o = cnt_or_ptr[row0 + k];
            for (int a = 0; a < m; ++a) { colInd[o + a] = c[a]; val[o + a] = v[a]; }
        }
        k = -1;
    }
    __device__ void begin(int row) { close(); k = row; n = 0; }
    __device__ void add(long site, int channel, T value) {
        if (n < CAP) { c[n] = (int)VecT::index(channel, site, plane); v[n] = value; ++n; }
    }
};

}  // namespace optk
