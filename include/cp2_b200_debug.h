/* cp2_b200 — TEST-ONLY reference kernels (CUDA cores, fp32 math, no tensor cores).
 *
 * Same contracts as cp2_dense_lse / cp2_dense_grad / cp2_global_nce in cp2_b200.h, but
 * written as plain shared-memory tiled SIMT loops.  They exist so that tests/ can
 * cross-check the tcgen05 kernels ON THE GPU at sizes the CPU oracle cannot reach in
 * seconds (P = 3136), and to localise a failure to one kernel.  cp2_b200/ (the product
 * path) never calls them.
 */
#ifndef CP2_B200_DEBUG_H_
#define CP2_B200_DEBUG_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

int cp2_dbg_dense_lse_simt(const void *qhat_t, const void *khat_t, float *lse, float inv_tau,
                           int B, int C, int P, void *stream);
int cp2_dbg_dense_grad_simt(const void *qhat_t, const void *khat_t, const float *lse,
                            const float *mask_a, const float *mask_b, const float *inv_norm,
                            const float *alpha, const float *v1, const float *v2, float *dq,
                            float inv_tau, int B, int C, int P, void *stream);
int cp2_dbg_global_nce_simt(const float *q_pos, const float *queue, float *z, float *g,
                            float inv_tau, int B, int C, int K, void *stream);

/* Wait-time accounting of the next cp2_dense_grad launches: buf = device array [n_sms*16] of int64 cycle
 * counters per CTA (0/1 producer waits on q_empty/kempty; 2-5 MMA issuer waits on q_full/d_empty/kfull/
 * p_full; 6/7 first softmax warp waits on kfull/s_full; 8 epilogue wait on d_full; 10-12 role lifetimes),
 * or NULL to switch it off. */
int cp2_dbg_dense_trace(long long *buf);

#ifdef __cplusplus
}
#endif
#endif
