// Entry points of the fp32-storage / 3xTF32 tcgen05 dense path (gemm_tf32x3.cu), used by the
// full-rank family, the correlated-Gaussian plugin and the Student-t regression plugin.
#pragma once
#include "common.cuh"

namespace vb {

struct ModelHost;

size_t gemm_tf32x3_workspace_bytes(long long m, long long n, long long k);
// C (m, n) = alpha * op(A) op(B) + beta * C, row-major fp32, 3xTF32 on tcgen05 (fp32-accurate).
int gemm_tf32x3(int transa, int transb, long long m, long long n, long long k, float alpha,
                const float* A, long long lda, const float* B, long long ldb, float beta,
                float* C, long long ldc, void* ws, size_t ws_bytes, cudaStream_t st);

size_t fr_sample_tf32_workspace_bytes(long long n, int dim);
int fr_sample_tf32(const double* mu, const float* scale, double df, long long n, int dim, const float* z,
                   const float* chi2, float* out_x, void* ws, size_t ws_bytes, cudaStream_t st);

size_t gauss_f32_workspace_bytes(long long n, int dim);
int gauss_logp_grad_f32(const double* mean, const double* prec, int dim, const float* x, long long n,
                        double* out_logp, float* out_grad_or_null, void* ws, size_t ws_bytes, cudaStream_t st,
                        const float* cache = nullptr);
size_t gauss_cache_bytes(int dim);
int gauss_fill_cache(const double* mean, const double* prec, int dim, float* cache, cudaStream_t st);

size_t fr_weighted_tf32_workspace_bytes(long long n, int dim);
// G_S (dim, dim) fp64 = sum_i (w_i / sqrt(chi2_i/df)) z_i' g_i on the tensor path (split-K, fp64 cross-split sum)
int fr_weighted_GS_tf32(double df, long long n, int dim, const float* z, const float* chi2, const double* w,
                        const float* grad, double* out_GS, void* ws, size_t ws_bytes, cudaStream_t st,
                        double* out_gmu_or_null = nullptr, const float* z_planes_or_null = nullptr);
// true when fr_weighted_GS_tf32 also returns g_mu = sum_i w_i g_i (its row-scaling pass carries the column sums)
bool fr_weighted_colsum_fused(long long n, int dim);

size_t regression_tf32_workspace_bytes(long long n, long long n_rows, int dim, bool have_planes);
int regression_logp_grad_tf32(const ModelHost* m, const float* b, long long n, double* out_logp,
                              float* out_grad_or_null, void* ws, size_t ws_bytes, cudaStream_t st);

// Fused fp16x3 kernel (reg_fused.cu): needs the scaled fp16 planes of X attached to the model and dim <= 128
bool regression_fused_ok(const ModelHost* m);
size_t regression_fused_workspace_bytes(long long n, long long n_rows);
int regression_logp_grad_fused(const ModelHost* m, const float* b, long long n, double* out_logp,
                               float* out_grad_or_null, void* ws, size_t ws_bytes, cudaStream_t st);

}  // namespace vb
