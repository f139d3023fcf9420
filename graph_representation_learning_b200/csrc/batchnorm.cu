// BatchNorm1d over the node rows of a [M, C] activation + leaky ReLU, forward and backward, for the DeepRP sibling
// models (reference: GCNBlock / EmbeddingBlock, gnn/models/networks/deep_rp_robust_gcn.py:30-62 -- nn.BatchNorm1d on
// the permuted [B, C, N] tensor = per-channel statistics over all B*N rows, then F.leaky_relu(slope 0.2)).
//
// HBM-bound column reductions: every reduction is two-stage with a FIXED summation order (row blocks -> per-block
// partials -> one thread per column over the partials), so results do not depend on scheduling.  Variance is taken
// around the mean in a second pass (no E[x^2]-E[x]^2 cancellation).  fp32 only.
#include "common.cuh"

namespace grl {

constexpr int kBnRowsPerBlock = 256;   // rows reduced by one block (32 x 8 threads, 32 columns)

// MODE 0: sum (x - X[0, c])  MODE 1: sum (x - mean)^2
// MODE 2: two sums over g = dY * leaky'(Y): sum g and sum g * xhat, xhat = (x - mean) * rstd
template <int MODE>
__global__ void __launch_bounds__(256)
bn_partial_kernel(const float* __restrict__ X, int64_t ldx, const float* __restrict__ dY, int64_t lddy,
                  const float* __restrict__ Y, int64_t ldy, const float* __restrict__ mean,
                  const float* __restrict__ rstd, float slope, int64_t M, int C, float* __restrict__ part0,
                  float* __restrict__ part1) {
    __shared__ float s0[8][33], s1[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    const int64_t r0 = static_cast<int64_t>(blockIdx.y) * kBnRowsPerBlock;
    const int64_t r1 = r0 + kBnRowsPerBlock < M ? r0 + kBnRowsPerBlock : M;
    float a0 = 0.0f, a1 = 0.0f;
    if (c < C) {
        // MODE 0 sums x - X[0, c] (the finalize step adds the shift back): columns with |mean| >> sigma keep their digits
        const float mu = MODE >= 1 ? mean[c] : X[c];
        const float rs = MODE == 2 ? rstd[c] : 0.0f;
        for (int64_t m = r0 + threadIdx.y; m < r1; m += 8) {
            const float x = X[m * ldx + c];
            if (MODE == 0) a0 += x - mu;
            else if (MODE == 1) a0 += (x - mu) * (x - mu);
            else {
                const float y = Y[m * ldy + c];
                const float g = dY[m * lddy + c] * (y > 0.0f ? 1.0f : slope);
                a0 += g;
                a1 += g * ((x - mu) * rs);
            }
        }
    }
    s0[threadIdx.y][threadIdx.x] = a0;
    s1[threadIdx.y][threadIdx.x] = a1;
    __syncthreads();
    if (threadIdx.y == 0 && c < C) {
        float t0 = 0.0f, t1 = 0.0f;
#pragma unroll
        for (int k = 0; k < 8; ++k) { t0 += s0[k][threadIdx.x]; t1 += s1[k][threadIdx.x]; }
        part0[static_cast<int64_t>(blockIdx.y) * C + c] = t0;
        if (MODE == 2) part1[static_cast<int64_t>(blockIdx.y) * C + c] = t1;
    }
}

// One thread per column walks the partials in block order.
// STEP 0: mean = shift + sum / M (shift = `mean` argument = first row of X) -> out0
// STEP 1: var = sum / M; rstd -> out0; running statistics updated (momentum, unbiased variance) when given
// STEP 2: out0 = sum0 (d beta), out1 = sum1 (d gamma)
template <int STEP>
__global__ void bn_finalize_kernel(const float* __restrict__ part0, const float* __restrict__ part1, int nblk, int C,
                                   int64_t M, float eps, float momentum, const float* __restrict__ mean,
                                   float* __restrict__ out0, float* __restrict__ out1, float* running_mean,
                                   float* running_var) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    float t0 = 0.0f, t1 = 0.0f;
    for (int b = 0; b < nblk; ++b) {
        t0 += part0[static_cast<int64_t>(b) * C + c];
        if (STEP == 2) t1 += part1[static_cast<int64_t>(b) * C + c];
    }
    if (STEP == 0) {
        out0[c] = mean[c] + t0 / static_cast<float>(M);          // mean = shift row X[0, :] + mean of the shifted values
    } else if (STEP == 1) {
        const float var = t0 / static_cast<float>(M);
        out0[c] = rsqrtf(var + eps);
        if (running_mean) running_mean[c] = (1.0f - momentum) * running_mean[c] + momentum * mean[c];
        if (running_var) {
            const float unbiased = M > 1 ? t0 / static_cast<float>(M - 1) : var;
            running_var[c] = (1.0f - momentum) * running_var[c] + momentum * unbiased;
        }
    } else {
        out0[c] = t0;
        out1[c] = t1;
    }
}

__global__ void bn_eval_stats_kernel(const float* __restrict__ running_mean, const float* __restrict__ running_var,
                                     float eps, int C, float* __restrict__ mean, float* __restrict__ rstd) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < C) {
        mean[c] = running_mean[c];
        rstd[c] = rsqrtf(running_var[c] + eps);
    }
}

// y = leaky((x - mean) * rstd * gamma + beta)
__global__ void __launch_bounds__(256)
bn_apply_fwd_kernel(const float* __restrict__ X, int64_t ldx, const float* __restrict__ mean,
                    const float* __restrict__ rstd, const float* __restrict__ gamma, const float* __restrict__ beta,
                    float slope, float* __restrict__ Y, int64_t ldy, int64_t M, int C) {
    const int64_t total = M * C;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t m = t / C;
        const int c = static_cast<int>(t - m * C);
        const float z = (X[m * ldx + c] - mean[c]) * rstd[c] * gamma[c] + beta[c];
        Y[m * ldy + c] = z > 0.0f ? z : slope * z;
    }
}

// training: dx = gamma * rstd * (g - sum_g / M - xhat * sum_gx / M);   eval (batch_stats == 0): dx = gamma * rstd * g
__global__ void __launch_bounds__(256)
bn_apply_bwd_kernel(const float* __restrict__ dY, int64_t lddy, const float* __restrict__ Y, int64_t ldy,
                    const float* __restrict__ X, int64_t ldx, const float* __restrict__ mean,
                    const float* __restrict__ rstd, const float* __restrict__ gamma, const float* __restrict__ sum_g,
                    const float* __restrict__ sum_gx, float slope, int batch_stats, float* __restrict__ dX, int64_t lddx,
                    int64_t M, int C) {
    const int64_t total = M * C;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    const float inv_m = 1.0f / static_cast<float>(M);
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t m = t / C;
        const int c = static_cast<int>(t - m * C);
        const float y = Y[m * ldy + c];
        float g = dY[m * lddy + c] * (y > 0.0f ? 1.0f : slope);
        if (batch_stats) {
            const float xhat = (X[m * ldx + c] - mean[c]) * rstd[c];
            g = g - sum_g[c] * inv_m - xhat * sum_gx[c] * inv_m;
        }
        dX[m * lddx + c] = gamma[c] * rstd[c] * g;
    }
}

static inline int bn_blocks(int64_t M) { return static_cast<int>((M + kBnRowsPerBlock - 1) / kBnRowsPerBlock); }

static inline int bn_ew_grid(int64_t total) {
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = static_cast<int64_t>(kNumSMs) * 16;
    return static_cast<int>(blocks > cap ? cap : (blocks < 1 ? 1 : blocks));
}

}  // namespace grl

using namespace grl;

extern "C" size_t grl_bn_workspace_bytes(int64_t M, int C) {
    if (M <= 0 || C <= 0) return 0;
    return static_cast<size_t>(bn_blocks(M)) * C * 2 * sizeof(float);
}

extern "C" int grl_bn_act_fwd(const float* X, int64_t ldx, const float* gamma, const float* beta, float* running_mean,
                              float* running_var, float momentum, float eps, int training, float slope, float* Y,
                              int64_t ldy, float* save_mean, float* save_rstd, int64_t M, int C, void* ws,
                              size_t ws_bytes, void* stream) {
    GRL_CHECK_ARG(X && Y && gamma && beta && save_mean && save_rstd, GRL_E_BADARG, "grl_bn_act_fwd: null pointer");
    GRL_CHECK_ARG(M > 0 && C > 0 && ldx >= C && ldy >= C, GRL_E_BADARG, "grl_bn_act_fwd: bad shape M=%lld C=%d",
                  (long long)M, C);
    GRL_CHECK_ARG(training || (running_mean && running_var), GRL_E_BADARG,
                  "grl_bn_act_fwd: eval mode needs the running statistics");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int cb = (C + 127) / 128;
    if (training) {
        GRL_CHECK_ARG(ws && ws_bytes >= grl_bn_workspace_bytes(M, C), GRL_E_BADARG, "grl_bn_act_fwd: workspace too small");
        const int nblk = bn_blocks(M);
        float* part = static_cast<float*>(ws);
        dim3 grid((C + 31) / 32, nblk), block(32, 8);
        bn_partial_kernel<0><<<grid, block, 0, st>>>(X, ldx, nullptr, 0, nullptr, 0, nullptr, nullptr, 0.0f, M, C, part, nullptr);
        GRL_LAUNCH_CHECK("bn_partial_kernel<0>");
        bn_finalize_kernel<0><<<cb, 128, 0, st>>>(part, nullptr, nblk, C, M, eps, momentum, X, save_mean, nullptr,
                                                  nullptr, nullptr);
        GRL_LAUNCH_CHECK("bn_finalize_kernel<0>");
        bn_partial_kernel<1><<<grid, block, 0, st>>>(X, ldx, nullptr, 0, nullptr, 0, save_mean, nullptr, 0.0f, M, C, part, nullptr);
        GRL_LAUNCH_CHECK("bn_partial_kernel<1>");
        bn_finalize_kernel<1><<<cb, 128, 0, st>>>(part, nullptr, nblk, C, M, eps, momentum, save_mean, save_rstd, nullptr,
                                                  running_mean, running_var);
        GRL_LAUNCH_CHECK("bn_finalize_kernel<1>");
    } else {
        bn_eval_stats_kernel<<<cb, 128, 0, st>>>(running_mean, running_var, eps, C, save_mean, save_rstd);
        GRL_LAUNCH_CHECK("bn_eval_stats_kernel");
    }
    bn_apply_fwd_kernel<<<bn_ew_grid(M * C), 256, 0, st>>>(X, ldx, save_mean, save_rstd, gamma, beta, slope, Y, ldy, M, C);
    GRL_LAUNCH_CHECK("bn_apply_fwd_kernel");
    return 0;
}

extern "C" int grl_bn_act_bwd(const float* dY, int64_t lddy, const float* Y, int64_t ldy, const float* X, int64_t ldx,
                              const float* gamma, const float* save_mean, const float* save_rstd, float slope,
                              int training, float* dX, int64_t lddx, float* dgamma, float* dbeta, int64_t M, int C,
                              void* ws, size_t ws_bytes, void* stream) {
    GRL_CHECK_ARG(dY && Y && X && gamma && save_mean && save_rstd && dX && dgamma && dbeta, GRL_E_BADARG,
                  "grl_bn_act_bwd: null pointer");
    GRL_CHECK_ARG(M > 0 && C > 0 && lddy >= C && ldy >= C && ldx >= C && lddx >= C, GRL_E_BADARG,
                  "grl_bn_act_bwd: bad shape M=%lld C=%d", (long long)M, C);
    GRL_CHECK_ARG(ws && ws_bytes >= grl_bn_workspace_bytes(M, C), GRL_E_BADARG, "grl_bn_act_bwd: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nblk = bn_blocks(M);
    float* part0 = static_cast<float*>(ws);
    float* part1 = part0 + static_cast<size_t>(nblk) * C;
    dim3 grid((C + 31) / 32, nblk), block(32, 8);
    bn_partial_kernel<2><<<grid, block, 0, st>>>(X, ldx, dY, lddy, Y, ldy, save_mean, save_rstd, slope, M, C, part0, part1);
    GRL_LAUNCH_CHECK("bn_partial_kernel<2>");
    bn_finalize_kernel<2><<<(C + 127) / 128, 128, 0, st>>>(part0, part1, nblk, C, M, 0.0f, 0.0f, nullptr, dbeta, dgamma,
                                                          nullptr, nullptr);
    GRL_LAUNCH_CHECK("bn_finalize_kernel<2>");
    bn_apply_bwd_kernel<<<bn_ew_grid(M * C), 256, 0, st>>>(dY, lddy, Y, ldy, X, ldx, save_mean, save_rstd, gamma, dbeta,
                                                          dgamma, slope, training ? 1 : 0, dX, lddx, M, C);
    GRL_LAUNCH_CHECK("bn_apply_bwd_kernel");
    return 0;
}
