/* C ABI of libdpfs_b200.so — the sm_100a kernel library behind distributed_pytorch_from_scratch_b200.
 *
 * Conventions
 *   - Every entry point enqueues work on `stream` and returns immediately; the return value is 0 on success,
 *     a positive cudaError_t, or a negative library error (bad alignment / unsupported shape).
 *   - Pointers are device pointers unless stated otherwise. bf16 tensors are `__nv_bfloat16` (passed as void*).
 *   - No entry point allocates device memory unless its comment says so; scratch is passed in by the caller.
 *   - "ref:" comments cite the file:line in ldh127/distributed_pytorch_from_scratch whose behaviour the kernel
 *     provides on B200.
 */
#pragma once

#include <cuda_runtime.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- GEMM (tcgen05 / TMEM / TMA) -------------------------------------------------------------------------
 * out[M,N] = A[M,K] * B[N,K]^T (+ bias[N]) (+ residual[M,N]),  bf16 inputs, fp32 accumulation in TMEM.
 *   a_mn_major = 0: A is stored [M][lda] (K contiguous).   1: A is stored [K][lda] (M contiguous).
 *   b_mn_major = 0: B is stored [N][ldb] (K contiguous).   1: B is stored [K][ldb] (N contiguous).
 *   out_fp32   = 0: out is bf16 [M][ldo];                  1: out is fp32 [M][ldo].
 *   bias (fp32, nullable), residual (fp32 [M][ldr], nullable, only with splits == 1).
 *   splits > 1 : split-K over `splits` CTAs groups; `workspace` must hold splits*M*N floats.
 *   bn         : N tile (128 or 256; 0 = 128).
 * Requirements: N % 8 == 0, lda/ldb % 8 == 0, 16-byte aligned bases.
 * ref: F.linear in models/layers.py:49,93 and its autograd dgrad/wgrad (cuBLAS in the reference). */
int dpfs_gemm_bf16(const void* A, long long lda, int a_mn_major, const void* B, long long ldb, int b_mn_major,
                   void* out, long long ldo, int out_fp32, const float* bias, const float* residual,
                   long long ldr, int M, int N, int K, int splits, float* workspace, int bn,
                   cudaStream_t stream);

#ifdef __cplusplus
}
#endif
