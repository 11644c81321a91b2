/* OptB200Sparse.h -- the sparse-matrix (CSR) primitives behind the materialized-Jacobian mode, as a C ABI.
 *
 * The reference's solver reaches its matrix library through five functions (API/src/backend_cuda.t:541-654, cuSPARSE
 * wrappers; API/src/linalg_cpu.t:203-617, the CPU library with the same contract; known-answer test
 * API/src/linalg_cpu_test.t:20-190).  Here they are CUDA kernels (opt_b200/csrc/csr.cuh):
 *
 *   reference                                        here
 *   computeNnzPatternAT + computeAT     (csr2csc)    OptB200_CsrTranspose
 *   computeNnzPatternATA + computeATA   (csrgemm)    OptB200_CsrAta
 *   applyAtoVector                      (csrmv)      OptB200_CsrSpmv
 *
 * All arrays are DEVICE pointers owned by the caller, 0-based CSR with ascending columns inside a row; single precision
 * (the reference's matrix values are float, solverGPUGaussNewton.t:1585).  Return 0 on success. */
#ifndef OPTB200_SPARSE_H
#define OPTB200_SPARSE_H
#ifdef __cplusplus
extern "C" {
#endif

/* A is nrows x ncols with nnz entries; writes A^T (ncols + 1 row offsets, nnz columns / values; rows of A^T sorted). */
int OptB200_CsrTranspose(int nrows, int ncols, int nnz, const int* rowPtr, const int* colInd, const float* val,
                         int* rowPtrT, int* colIndT, float* valT);
/* A^T A (ncols x ncols): row offsets (ncols + 1), columns and values into arrays of `capacity` entries; *nnzATA receives
 * the entry count.  Returns -1 (with *nnzATA set) if capacity is too small: call again with larger arrays. */
int OptB200_CsrAta(int nrows, int ncols, int nnz, const int* rowPtr, const int* colInd, const float* val,
                   int* rowPtrATA, int* colIndATA, float* valATA, int capacity, int* nnzATA);
/* y = A x */
int OptB200_CsrSpmv(int nrows, int ncols, int nnz, const int* rowPtr, const int* colInd, const float* val,
                    const float* x, float* y);

#ifdef __cplusplus
}
#endif
#endif
