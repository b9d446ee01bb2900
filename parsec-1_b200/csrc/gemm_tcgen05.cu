// BF16-in / FP32-accumulate tile GEMM (tcgen05 + TMEM + TMA).  Placeholder until the kernel lands.
#include "pb2_internal.h"
extern "C" int parsec_b200_gemm_bf16_sm100(char transa, char transb, int m, int n, int k, float alpha, const void* A,
                                           int lda, const void* B, int ldb, float beta, float* C, int ldc,
                                           void* stream) {
    (void)transa; (void)transb; (void)m; (void)n; (void)k; (void)alpha; (void)A; (void)lda; (void)B; (void)ldb;
    (void)beta; (void)C; (void)ldc; (void)stream;
    return PB2_ERR_NOT_IMPLEMENTED;
}
