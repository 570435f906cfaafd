/* Minimal CBLAS declaration used only to build the reference oracle with USE_BLAS.
 * TEST INFRASTRUCTURE: nothing under oracle/ is linked into the product library. */
#pragma once
#ifdef __cplusplus
extern "C" {
#endif
enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 };
enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 };
void cblas_sgemm(enum CBLAS_ORDER order, enum CBLAS_TRANSPOSE ta, enum CBLAS_TRANSPOSE tb,
                 int m, int n, int k, float alpha, const float* a, int lda,
                 const float* b, int ldb, float beta, float* c, int ldc);
void openblas_set_num_threads(int n);
#ifdef __cplusplus
}
#endif
