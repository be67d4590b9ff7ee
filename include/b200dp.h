/* libb200dp - C ABI of the hand-written sm_100a kernels and the native runtime of b200dp.
 *
 * Conventions
 *   - every function returns 0 (B200DP_OK) on success, non-zero on failure; the text of the last
 *     failure on the calling thread is returned by b200dp_last_error().
 *   - pointers are device pointers unless the name says "host"; `stream` is a cudaStream_t passed
 *     as void* (NULL = legacy default stream).  Nothing here synchronises the host unless stated.
 *   - matrices are row-major; `ld*` are leading dimensions in ELEMENTS.
 *   - bf16 = __nv_bfloat16 (2 bytes), fp32 = float.
 *
 * Each declaration cites the part of the reference (ref.py = dataParallelTraining_NN_MPI.py) whose
 * work it performs.
 */
#ifndef B200DP_H_
#define B200DP_H_

#ifdef __cplusplus
extern "C" {
#endif

#define B200DP_OK 0
#define B200DP_ERR 1
#define B200DP_MAX_RANKS 8
#define B200DP_MAX_LAYERS 64

const char* b200dp_last_error(void);
int b200dp_version(void);

/* ------------------------------------------------------------------------------------------------
 * Linear-layer GEMM on tcgen05 / TMEM / TMA:  D[M,N] = epilogue(A[M,K] * B[N,K]^T), bf16 inputs,
 * fp32 accumulation.  Covers aten::addmm of the forward (ref.py:42-44,170) and both mm's of
 * AddmmBackward0 (ref.py:176):
 *   forward  : A = X  (K-major, lda=K)    B = W  (K-major,  [N,K], ldb=K)   bias, relu
 *   dgrad    : A = dY (K-major)           B = W  (MN-major: stored [K,N], ldb=N)   mask_src = X
 *   wgrad    : A = dY (MN-major: stored [K,M], lda=M)   B = X (MN-major: stored [K,N], ldb=N)
 * Epilogue order: + bias[n] -> relu -> zero where mask_src[m,n] <= 0 -> store (bf16 or fp32).
 * Requirements: pointers 16-byte aligned, lda/ldb/ldd/ldmask multiples of 8 elements (4 for an
 * fp32 D), N % 8 == 0.  M/N/K tails are handled (TMA zero fill, guarded stores).
 */
typedef struct b200dp_gemm_args {
  int M, N, K;
  const void* A;        /* bf16 */
  long long lda;
  int a_mn_major;       /* 0: A stored [M,K];  1: A stored [K,M] */
  const void* B;        /* bf16 */
  long long ldb;
  int b_mn_major;       /* 0: B stored [N,K];  1: B stored [K,N] */
  void* D;              /* bf16 or fp32 [M,N] */
  long long ldd;
  int d_fp32;
  const float* bias;    /* fp32 [N] or NULL */
  int relu;
  const void* mask_src; /* bf16 [M,N] or NULL */
  long long ldmask;
  int tile_n;           /* 0 = auto, 128 or 256 */
  int max_ctas;         /* 0 = one persistent CTA per SM; >0 caps the grid (SM budget when overlapping) */
} b200dp_gemm_args;

int b200dp_gemm_bf16(const b200dp_gemm_args* args, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DP_H_ */
