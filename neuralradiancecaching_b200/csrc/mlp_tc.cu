// mlp_tc.cu -- tcgen05 (bf16, fp32 accumulate in TMEM) fused encode + MLP.  Placeholder until the kernel lands:
// selecting NRC_MLP_BF16_TC fails loudly instead of silently running the fp32 path.
#include "nrc_internal.cuh"

nrc_status launch_infer_tc(nrc_mlp*, nrc_mrhe*, const float*, const float*, const uint32_t*, size_t, float*, cudaStream_t)
{
    nrc_set_error("NRC_MLP_BF16_TC: the tcgen05 kernel is not part of this build");
    return NRC_ERR_INVALID;
}
