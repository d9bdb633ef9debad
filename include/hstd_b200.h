/* C ABI of libhstd_b200.so — the sm_100a kernel library of
 * huggingface_sagemaker_tensorflow_distributed_b200.
 *
 * Every entry point takes raw device pointers plus a `stream` (a cudaStream_t passed as
 * void*), launches asynchronously on that stream and returns 0 on success or a negative
 * error code; `hstd_last_error()` returns a human-readable message for the calling thread.
 * Unless a comment says otherwise, tensors are row-major bf16 and no entry point allocates
 * device memory (workspaces are passed by the caller).
 *
 * The reference (philschmid/huggingface_sagemaker_tensorflow_distributed) has no native
 * code; each entry point cites the reference line whose *implied* library call it replaces
 * (SURVEY.md §2.4, K1..K13).
 */
#ifndef HSTD_B200_H_
#define HSTD_B200_H_

#include <stdint.h>

#if defined(HSTD_BUILDING_LIB)
#define HSTD_API __attribute__((visibility("default")))
#else
#define HSTD_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

const char* hstd_last_error(void);
int hstd_version(void);

/* ------------------------------------------------------------------ GEMM (K2,K4,K5,K6 + attention GEMMs)
 * tcgen05/TMEM GEMM fed by TMA.  Replaces the cuBLAS calls behind model.fit()
 * (reference scripts/train.py:145).  For every batch index z = (z0, z1):
 *     acc[m,n] = sum_k A_z(m,k) * B_z(n,k)
 *     C_z[m,n] = epilogue(alpha * acc[m,n] + bias[n])
 * Operand X (A or B) in K-major form is read at  X + z0*bs0 + z1*bs1 + row*ld + k   (row = m or n),
 * in MN-major form ("transposed") at            X + z0*bs0 + z1*bs1 + k*ld   + row.
 * All strides are in elements; ld, bs0, bs1 must be multiples of 8 and pointers 16-byte aligned.
 */
enum {
  HSTD_EPI_NONE = 0,     /* C = v                                                        */
  HSTD_EPI_GELU = 1,     /* aux (if non-null) = v ; C = gelu_erf(v)                       */
  HSTD_EPI_ADD = 2,      /* C = v + aux        (aux: bf16 [M,N] read with ldaux)          */
  HSTD_EPI_GELU_BWD = 3, /* C = v * gelu_erf'(aux)  (aux: pre-activation, ldaux)          */
};

typedef struct hstd_gemm_args {
  const void* A;
  const void* B;
  void* C;
  const void* bias; /* bf16 [N] or NULL */
  void* aux;        /* see epilogue */
  int64_t lda, a_bs0, a_bs1;
  int64_t ldb, b_bs0, b_bs1;
  int64_t ldc, c_bs0, c_bs1;
  int64_t ldaux; /* aux uses the same batch strides as C */
  int32_t M, N, K;
  int32_t nb0, nb1; /* batch extents (>=1) */
  int32_t a_mn_major, b_mn_major;
  int32_t epilogue;
  float alpha;
  int32_t block_n; /* 0 = auto, else 64/128/256 */
  int32_t max_ctas; /* 0 = one persistent CTA per SM */
} hstd_gemm_args;

int hstd_gemm_bf16(const hstd_gemm_args* args, void* stream);

/* Split-K variant for weight gradients (K = tokens is long, M x N is small):
 * dW[M,N] = sum_k A(m,k) B(n,k) with `splits` partial sums written as fp32 to
 * `workspace` ([splits, M, N] floats, caller-allocated) and reduced to bf16 C by a second
 * kernel.  Only alpha and EPI_NONE are honoured.  No batch dimensions. */
int hstd_gemm_bf16_splitk(const hstd_gemm_args* args, int splits, void* workspace, int64_t workspace_bytes,
                          void* stream);
int64_t hstd_gemm_splitk_workspace_bytes(int M, int N, int splits);

#ifdef __cplusplus
}
#endif
#endif /* HSTD_B200_H_ */
