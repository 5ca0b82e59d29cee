// Placeholder until the classifier kernel lands (next commit).
#include "elbo.cuh"
int smi_model_cls_logp_grad(const SmiModelDesc& md, const float* Z, const float* x, const int32_t* y, int ndraws,
                            int B, float lik_scale, float* DZ, float* LOGP, cudaStream_t st) {
  (void)md; (void)Z; (void)x; (void)y; (void)ndraws; (void)B; (void)lik_scale; (void)DZ; (void)LOGP; (void)st;
  SMI_FAIL(SMI_ELIMIT, "bnn_cls kernel not built");
}
