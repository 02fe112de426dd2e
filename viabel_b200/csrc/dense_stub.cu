// Placeholder until gemm_tf32x3.cu lands: the fp32 dense entry points report "unsupported".
#include "dense.cuh"
namespace vb {
size_t gemm_tf32x3_workspace_bytes(long long, long long, long long) { return 0; }
int gemm_tf32x3(int, int, long long, long long, long long, float, const float*, long long, const float*, long long,
                float, float*, long long, void*, size_t, cudaStream_t) { return VB_ERR_UNSUPPORTED; }
int fr_sample_tf32(const double*, const float*, double, long long, int, const float*, const float*, float*, void*,
                   size_t, cudaStream_t) { return VB_ERR_UNSUPPORTED; }
int gauss_logp_grad_f32(const double*, const double*, int, const float*, long long, double*, float*, void*, size_t,
                        cudaStream_t) { return VB_ERR_UNSUPPORTED; }
size_t regression_tf32_workspace_bytes(long long, long long, int) { return 0; }
int regression_logp_grad_tf32(const ModelHost*, const float*, long long, double*, float*, void*, size_t,
                              cudaStream_t) { return VB_ERR_UNSUPPORTED; }
}  // namespace vb
