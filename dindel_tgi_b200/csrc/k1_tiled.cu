// k1_tiled.cu — K1, register-tiled variant (placeholder until the tiled kernel lands).
#include "k1_common.cuh"
extern "C" int k1_tiled_supported(int maxS, int maxR, int J) { (void)maxS; (void)maxR; (void)J; return 0; }
extern "C" cudaError_t k1_tiled_launch(const K1Batch *B, int num_sms, cudaStream_t stream, int *launches)
{
    (void)B; (void)num_sms; (void)stream; (void)launches;
    return cudaErrorNotSupported;
}
