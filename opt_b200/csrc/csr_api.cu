// csr_api.cu -- C ABI over the CSR kernels of csr.cuh (include/OptB200Sparse.h): what the reference's matrix library
// exposes to its solver (backend_cuda.t:541-654 / linalg_cpu.t:203-617), for callers and for the known-answer test.
// All pointers are DEVICE pointers; outputs are allocated by the caller.
#include "csr.cuh"
#include "../../include/OptB200Sparse.h"

using namespace optk;

extern "C" void optk_report_fatal(int code);   // registry.cu: runs the installed error handler

namespace {
CsrMat<float> view(int nrows, int ncols, int nnz, const int* rowPtr, const int* colInd, const float* val) {
    CsrMat<float> A;
    A.nrows = nrows; A.ncols = ncols; A.nnz = nnz;
    A.rowPtr = const_cast<int*>(rowPtr); A.colInd = const_cast<int*>(colInd); A.val = const_cast<float*>(val);
    return A;
}
// optk_fatal throws when the application installed an error handler (common.cuh): end the call with the error code, after the
// handler ran; temporaries are released by the Cleanup objects on the way out
template <class Fn> int guarded(Fn&& fn) {
    try { return fn(); } catch (const optk::Fatal& f) { optk_report_fatal(f.code); return f.code ? f.code : -1; }
}
template <class Fn> struct Cleanup { Fn fn; ~Cleanup() { fn(); } };
template <class Fn> Cleanup<Fn> at_exit(Fn fn) { return Cleanup<Fn>{fn}; }
}  // namespace

extern "C" {

int OptB200_CsrTranspose(int nrows, int ncols, int nnz, const int* rowPtr, const int* colInd, const float* val,
                         int* rowPtrT, int* colIndT, float* valT) {
    CsrMat<float> A = view(nrows, ncols, nnz, rowPtr, colInd, val), AT;
    Scratch tmp; int* perm = nullptr;
    auto cleanup = at_exit([&] { AT.release(); cudaFree(perm); tmp.release(); });
    return guarded([&] {
        csr_transpose_pattern(A, AT, &perm, tmp, 0);
        csr_transpose_values(A, AT, perm, 0);
        OPTK_CUDA(cudaMemcpy(rowPtrT, AT.rowPtr, (size_t)(ncols + 1) * sizeof(int), cudaMemcpyDeviceToDevice));
        OPTK_CUDA(cudaMemcpy(colIndT, AT.colInd, (size_t)nnz * sizeof(int), cudaMemcpyDeviceToDevice));
        OPTK_CUDA(cudaMemcpy(valT, AT.val, (size_t)nnz * sizeof(float), cudaMemcpyDeviceToDevice));
        return 0;
    });
}

int OptB200_CsrAta(int nrows, int ncols, int nnz, const int* rowPtr, const int* colInd, const float* val,
                   int* rowPtrATA, int* colIndATA, float* valATA, int capacity, int* nnzATA) {
    CsrMat<float> A = view(nrows, ncols, nnz, rowPtr, colInd, val), AT, ATA;
    Scratch tmp; int *perm = nullptr, *rowOf = nullptr;
    auto cleanup = at_exit([&] { AT.release(); ATA.release(); cudaFree(perm); cudaFree(rowOf); tmp.release(); });
    return guarded([&] {
        csr_transpose_pattern(A, AT, &perm, tmp, 0);
        csr_transpose_values(A, AT, perm, 0);
        csr_ata_pattern(A, ATA, &rowOf, tmp, 0);
        *nnzATA = (int)ATA.nnz;
        if (ATA.nnz > capacity) return -1;
        csr_ata_values(A, AT, ATA, rowOf, 0);
        OPTK_CUDA(cudaMemcpy(rowPtrATA, ATA.rowPtr, (size_t)(ncols + 1) * sizeof(int), cudaMemcpyDeviceToDevice));
        OPTK_CUDA(cudaMemcpy(colIndATA, ATA.colInd, (size_t)ATA.nnz * sizeof(int), cudaMemcpyDeviceToDevice));
        OPTK_CUDA(cudaMemcpy(valATA, ATA.val, (size_t)ATA.nnz * sizeof(float), cudaMemcpyDeviceToDevice));
        return 0;
    });
}

int OptB200_CsrSpmv(int nrows, int ncols, int nnz, const int* rowPtr, const int* colInd, const float* val, const float* x, float* y) {
    CsrMat<float> A = view(nrows, ncols, nnz, rowPtr, colInd, val);
    int dev = 0, nsm = 148;
    cudaGetDevice(&dev); cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
    return guarded([&] {
        csr_spmv(A, x, y, nsm, 0);
        OPTK_CUDA(cudaGetLastError());
        OPTK_CUDA(cudaStreamSynchronize(0));
        return 0;
    });
}

}  // extern "C"
