/* libmuscle_b200_diag.so — a TEST fixture, not part of the product library (libmuscle_b200.so does not contain it):
 * one-tile checks of the tcgen05 building blocks the pair passes are made of (TMA 128-byte-swizzled loads, K-major and
 * MN-major shared-memory descriptors, tcgen05.ld / tcgen05.st, an MMA with its A operand in tensor memory).  Built from
 * muscle_b200/csrc/debug_tile.cu by `python -m muscle_b200.build`; used by tests/test_gpu_infonce.py and
 * scripts/gpu_diag.py.  No counterpart in the reference. */
#ifndef MUSCLE_B200_DIAG_H_
#define MUSCLE_B200_DIAG_H_

#ifdef __cplusplus
extern "C" {
#endif

/* S[128,128] = A[128,dim] B[128,dim]^T ;  G[128,dim] = bf16(S) B  (A from TMEM).  a_bf16, b_bf16: device, bf16,
 * row-major, 128-byte aligned; dim in {64,128,256}.  Returns 0 or an error code of muscle_b200.h. */
int muscle_debug_tile(const void* a_bf16, const void* b_bf16, int dim, float* s_out,
                      float* g_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif  /* MUSCLE_B200_DIAG_H_ */
