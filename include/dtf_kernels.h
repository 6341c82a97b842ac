/* C ABI of libdtf_kernels.so -- the hand-written sm_100a kernels of distributed_tensorflow_b200.
 *
 * Every function returns 0 on success and a negative value on failure; dtf_last_error() then
 * holds a description (thread local).  Pointers are device pointers unless stated otherwise;
 * `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  All launches are
 * asynchronous with respect to the host.
 *
 * The reference (yaokeepmoving/distributed_tensorflow) has no native code of its own; each entry
 * cites the reference call site whose TF-runtime kernel it stands in for.
 */
#ifndef DTF_KERNELS_H_
#define DTF_KERNELS_H_

#ifdef __cplusplus
extern "C" {
#endif

#define DTF_KERNELS_ABI_VERSION 1

/* ---- bookkeeping ---------------------------------------------------------------------------- */
const char* dtf_last_error(void);
long long dtf_launch_count(void);      /* kernels launched by this library since the last reset */
void dtf_reset_launch_count(void);
int dtf_kernels_abi_version(void);

/* ---- epilogues of dtf_gemm_bf16 -------------------------------------------------------------- */
enum {
  DTF_EPI_BF16 = 0,           /* D = bf16(acc)                                  */
  DTF_EPI_BIAS_BF16 = 1,      /* D = bf16(acc + bias[n])                        */
  DTF_EPI_BIAS_RELU_BF16 = 2, /* D = bf16(max(acc + bias[n], 0))                */
  DTF_EPI_DRELU_BF16 = 3,     /* D = bf16(mask bit(m,n) ? acc : 0)              */
  DTF_EPI_F32 = 4,            /* D = acc (fp32)                                 */
  DTF_EPI_F32_ATOMIC = 5      /* D += acc (fp32 red.global.add; for split-K)    */
};

/* D[M,N] = epilogue(A * B^T), bf16 operands, fp32 accumulation on tcgen05 tensor cores.
 *   a_mn == 0: A is stored [M,K] (row stride lda elements);  a_mn == 1: A is stored [K,M].
 *   b_mn == 0: B is stored [N,K] (row stride ldb elements);  b_mn == 1: B is stored [K,N].
 * Supported layouts: (0,0) forward, (0,1) data gradient, (1,1) weight gradient.
 * Row strides must be multiples of 8 elements and bases 16-byte aligned (TMA).
 * splits > 1 partitions K across CTAs and requires DTF_EPI_F32_ATOMIC (D pre-zeroed by caller).
 * ReLU masks are bit matrices: uint32 word [m, n/32] bit (n%32); ldmask = words per row, even,
 *   >= ceil(N/32).  DTF_EPI_BIAS_RELU_BF16 writes one to mask_out when it is non-NULL (bit =
 *   output > 0); DTF_EPI_DRELU_BF16 reads mask_in.  D must satisfy the same alignment as A/B
 *   (it is written by TMA stores).
 * bn: N tile (64/128/256) or 0 to let the library choose.
 * Stands in for the Mul/Add/Sum Eigen kernels behind reference trainer/task.py:137-139. */
int dtf_gemm_bf16(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn,
                  void* D, long long ldd, int M, int N, int K, int epilogue, const float* bias,
                  const void* mask_in, void* mask_out, long long ldmask, int splits, int bn,
                  void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DTF_KERNELS_H_ */
