/* oracle/shim/cblas_shim.h -- TEST INFRASTRUCTURE (oracle build only).
 * Force-included declarations of the only CBLAS symbols the reference uses
 * (libsanm/tensor_linalg.cpp:66-68,98-102,202-207).  Implemented naively in
 * shim_impl.cpp.  The 3x3 element path never reaches them (fixed-size Eigen
 * branches, tensor_linalg.cpp:192-195). */
#pragma once
#ifdef __cplusplus
extern "C" {
#endif
typedef enum { CblasRowMajor = 101, CblasColMajor = 102 } CBLAS_LAYOUT;
typedef enum { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 } CBLAS_TRANSPOSE;
void cblas_dgemm(CBLAS_LAYOUT layout, CBLAS_TRANSPOSE ta, CBLAS_TRANSPOSE tb,
                 int m, int n, int k, double alpha, const double* a, int lda,
                 const double* b, int ldb, double beta, double* c, int ldc);
void cblas_dgemm_batch_strided(CBLAS_LAYOUT layout, CBLAS_TRANSPOSE ta,
                               CBLAS_TRANSPOSE tb, int m, int n, int k,
                               double alpha, const double* a, int lda,
                               int stridea, const double* b, int ldb,
                               int strideb, double beta, double* c, int ldc,
                               int stridec, int batch);
#ifdef __cplusplus
}
#endif
