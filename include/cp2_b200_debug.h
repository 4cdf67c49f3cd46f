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

#ifdef __cplusplus
}
#endif
#endif
