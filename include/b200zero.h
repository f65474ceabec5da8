/*
 * b200zero — C ABI of the sm_100a kernel library (libb200zero.so).
 *
 * Every entry point takes raw device pointers plus a CUDA stream handle (as
 * void*; NULL = legacy default stream), launches asynchronously on that stream
 * and returns 0 on success. On failure it returns non-zero and
 * bz_last_error() describes why (thread-local string). Unless stated otherwise
 * an entry point allocates nothing and synchronises nothing.
 *
 * Python binds these with ctypes (ops/lib.py); C/C++ callers can link the .so
 * directly. Citations are to the reference repo the framework replaces
 * (SURVEY.md); the reference itself contains no native code.
 */
#ifndef B200ZERO_H_
#define B200ZERO_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ---------------------------------------------------------- */
const char* bz_last_error(void);
int bz_version(void);
int bz_sm_count(void);

/* ---- K4: bf16 GEMM on tcgen05/TMEM/TMA --------------------------------
 * C[M,N] (+)= A*B with fp32 accumulation.
 *   a_mn_major=0: A stored [M,K] row-major (lda = row stride in elements)
 *   a_mn_major=1: A stored [K,M] row-major
 *   b_mn_major=0: B stored [N,K] row-major (nn.Linear weight layout)
 *   b_mn_major=1: B stored [K,N] row-major
 *   out_mode: 0 bf16 store, 1 bf16 accumulate (C += AB), 2 fp32 store, 3 fp32 accumulate
 *   cta_group: 1 = one CTA per tile (128x256), 2 = CTA pair (256x256, cta_group::2), 0 = auto
 *   max_ctas:  0 = all SMs; otherwise cap the persistent grid (leave SMs for a
 *              concurrent communication kernel)
 * Requirements: pointers 16-byte aligned; lda, ldb multiples of 8.
 * Replaces the cuBLAS calls a DeepSpeed user script would make.
 */
int bz_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                 int a_mn_major, int b_mn_major, int out_mode, int cta_group, int max_ctas, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200ZERO_H_ */
