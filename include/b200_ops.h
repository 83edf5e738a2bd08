/* C ABI of libb200ops.so — the sm_100a kernel library of llm_in_practise_b200.
 *
 * Every entry point takes raw device pointers plus a cudaStream_t (passed as void*), launches
 * asynchronously on that stream and returns 0 on success, a cudaError_t (< 1000) or a
 * B200_ERR_* code (>= 10001). Nothing here synchronises the device. Entry points that need
 * temporary global memory take it from the caller (`workspace` arguments).
 *
 * Each declaration cites the reference call site (path:line under /root/reference) whose
 * library kernel it replaces; the reference itself contains no native code.
 */
#ifndef B200_OPS_H_
#define B200_OPS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
  B200_DT_BF16 = 0,
  B200_DT_FP16 = 1,
  B200_DT_FP32 = 2, /* as GEMM input: multiplied on the tensor cores as TF32 */
};

/* ------------------------------------------------------------------------------------------
 * K1 / K3: tcgen05 GEMM.  D[b] = alpha * A[b] x B[b / b_div]^T (+ A2 x B2^T) (+ bias) + beta * D[b]
 *
 *   A is logically [M, K], B is logically [N, K].  "K-major" means the K index is contiguous
 *   (an nn.Linear weight, an activation matrix); "MN-major" (a_mn / b_mn = 1) means the operand
 *   is stored transposed, [K, M] or [K, N] row-major, so dgrad (dY x W) and wgrad (dY^T x X)
 *   read their operands in place with no transpose copy.
 *   *_rs is the stride in elements between consecutive rows of the stored matrix;
 *   *_bs0 / *_bs1 are batch strides (batch index b -> (b % nb0, b / nb0)); B uses b / b_div
 *   (grouped-query attention shares one K/V head among b_div query heads).
 *   A2 [M, K2], B2 [N, K2] (both K-major, K2 <= 64) add a low-rank term on the tensor cores as
 *   one more k-block: the LoRA update  y += (s * x A^T) B^T  of peft's lora.Linear.
 *   Replaces: cuBLAS behind every nn.Linear (LLM_Distributed_Trainning/PyTorch/ddp_basics/
 *   ddp_gpt_wikitext2.py:86-165) and peft LoRA (Fine-Tuning/qwen3-8b-lora-dist.py:118-125).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const void* A;
  const void* B;
  void* D;
  const void* bias; /* [N], dtype bias_dtype, may be NULL */
  const void* A2;   /* optional low-rank term, may be NULL */
  const void* B2;
  int32_t M, N, K, K2;
  int32_t batch, nb0, b_div, bias_dtype;
  int64_t a_rs, a_bs0, a_bs1;
  int64_t b_rs, b_bs0, b_bs1;
  int64_t d_rs, d_bs0, d_bs1;
  int64_t a2_rs, b2_rs;
  int32_t a_mn, b_mn, in_dtype, out_dtype;
  float alpha, beta;
} b200_gemm_args;

int b200_gemm(const b200_gemm_args* args, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_OPS_H_ */
