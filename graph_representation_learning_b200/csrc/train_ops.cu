// Loss and optimizer kernels of the KV train step (all deterministic: fixed-order reductions, no atomics).
//
//   grl_ce_fwd / grl_ce_bwd   nn.CrossEntropyLoss()(pred.transpose(1, 2), target) with ignore_index and optional class
//                             weights -- gnn/trainer/losses/cross_entropy_loss.py:19-35 (mean over the non-ignored
//                             nodes, weighted mean when weights are given) and its gradient.
//   grl_l2_norm               total gradient norm of clip_grad_norm_ -- gnn/trainer/training_procedures/kv_procedure.py:161.
//   grl_adam_step             torch.optim.Adam.step (the optimizer BuitlinOptimizer builds by name,
//                             gnn/trainer/optimizers/builtin_optimizer.py:22-26) over ONE flat parameter / gradient /
//                             moment buffer, with the clip coefficient min(1, max_norm / (norm + 1e-6)) applied on the
//                             fly from the device-resident norm (no host sync between clip and step).
#include "common.cuh"
#include <mutex>

namespace grl {

constexpr int kTrainThreads = 256;

// one warp per row; per-block partial sums (loss * w, w, #labels outside [0, C) that are not ignore_index) in block order
// -> partial[3 * blockIdx.x + {0, 1, 2}]
__global__ void __launch_bounds__(kTrainThreads)
ce_fwd_kernel(const float* __restrict__ logits, int64_t ld, const int64_t* __restrict__ target,
              const float* __restrict__ weight, int64_t ignore_index, int64_t M, int C, float* __restrict__ lse,
              float* __restrict__ partial) {
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    float loss_sum = 0.f, w_sum = 0.f, bad = 0.f;
    if (C <= 64) {
        // a row fits two registers per lane: loaded once, the next row (and its label) prefetched while this one is reduced,
        // x[y] taken from the registers by a shuffle.  Same operations in the same order as the general loop below; that one
        // paid four dependent memory latencies per row (49 us for 128 000 rows of 53 classes).
        const int c1 = lane + 32;
        float n0 = -INFINITY, n1 = -INFINITY;
        int64_t ny = 0;
        if (warp < M) {
            const float* x = logits + warp * ld;
            if (lane < C) n0 = x[lane];
            if (c1 < C) n1 = x[c1];
            ny = target[warp];
        }
        for (int64_t r = warp; r < M; r += n_warps) {
            const float v0 = n0, v1 = n1;
            const int64_t y = ny;
            const int64_t rn = r + n_warps;
            if (rn < M) {
                const float* x = logits + rn * ld;
                n0 = lane < C ? x[lane] : -INFINITY;
                n1 = c1 < C ? x[c1] : -INFINITY;
                ny = target[rn];
            }
            float m = -INFINITY;
            if (lane < C) m = fmaxf(m, v0);
            if (c1 < C) m = fmaxf(m, v1);
            for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
            float s = 0.f;
            if (lane < C) s += expf(v0 - m);
            if (c1 < C) s += expf(v1 - m);
            for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            const float l = m + logf(s);
            if (lane == 0) lse[r] = l;
            const bool ok = y != ignore_index && y >= 0 && y < C;
            const int yy = ok ? static_cast<int>(y) : 0;
            const float xy = __shfl_sync(0xffffffffu, yy < 32 ? v0 : v1, yy & 31);
            if (ok) {
                const float w = weight ? weight[y] : 1.0f;
                loss_sum += w * (l - xy);
                w_sum += w;
            } else if (y != ignore_index) {
                bad += 1.0f;
            }
        }
    } else
    for (int64_t r = warp; r < M; r += n_warps) {
        const float* x = logits + r * ld;
        float m = -INFINITY;
        for (int c = lane; c < C; c += 32) m = fmaxf(m, x[c]);
        for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        float s = 0.f;
        for (int c = lane; c < C; c += 32) s += expf(x[c] - m);
        for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        const float l = m + logf(s);
        if (lane == 0) lse[r] = l;
        const int64_t y = target[r];
        if (y != ignore_index && y >= 0 && y < C) {
            const float w = weight ? weight[y] : 1.0f;
            loss_sum += w * (l - x[y]);
            w_sum += w;
        } else if (y != ignore_index) {
            bad += 1.0f;                     // nn.CrossEntropyLoss raises on such a target; here the loss turns NaN
        }
    }
    __shared__ float sl[kTrainThreads / 32], sw[kTrainThreads / 32], sb[kTrainThreads / 32];
    if (lane == 0) { sl[wib] = loss_sum; sw[wib] = w_sum; sb[wib] = bad; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.f, b = 0.f, c = 0.f;
        for (int k = 0; k < kTrainThreads / 32; ++k) { a += sl[k]; b += sw[k]; c += sb[k]; }
        partial[3 * blockIdx.x] = a;
        partial[3 * blockIdx.x + 1] = b;
        partial[3 * blockIdx.x + 2] = c;
    }
}

// out[0] = sum(loss) / sum(w), out[1] = sum(w), out[2] = number of out-of-range labels, out[3] = sum(loss)
// (one block; fixed order).  All rows ignored (sum(w) == 0): loss 0 and, through out[1] = 0, a zero gradient.
// Any label outside [0, C) that is not ignore_index: loss = NaN (the reference's nn.CrossEntropyLoss raises).
__global__ void __launch_bounds__(kTrainThreads)
ce_finish_kernel(const float* __restrict__ partial, int n_blocks, float* __restrict__ out) {
    __shared__ float sa[kTrainThreads], sb[kTrainThreads], sc[kTrainThreads];
    float a = 0.f, b = 0.f, c = 0.f;
    for (int k = threadIdx.x; k < n_blocks; k += kTrainThreads) {
        a += partial[3 * k]; b += partial[3 * k + 1]; c += partial[3 * k + 2];
    }
    sa[threadIdx.x] = a; sb[threadIdx.x] = b; sc[threadIdx.x] = c;
    __syncthreads();
    for (int o = kTrainThreads / 2; o >= 1; o >>= 1) {
        if (threadIdx.x < o) {
            sa[threadIdx.x] += sa[threadIdx.x + o]; sb[threadIdx.x] += sb[threadIdx.x + o]; sc[threadIdx.x] += sc[threadIdx.x + o];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const float mean = sb[0] > 0.f ? sa[0] / sb[0] : 0.f;
        out[0] = sc[0] > 0.f ? __int_as_float(0x7fc00000) : mean;
        out[1] = sb[0];
        out[2] = sc[0];
        out[3] = sa[0];
    }
}

// dlogits[r][c] = g * w_y / sum(w) * (softmax(x)[c] - [c == y]);  0 for ignored rows
__global__ void __launch_bounds__(kTrainThreads)
ce_bwd_kernel(const float* __restrict__ logits, int64_t ld, const int64_t* __restrict__ target,
              const float* __restrict__ weight, int64_t ignore_index, int64_t M, int C, const float* __restrict__ lse,
              const float* __restrict__ loss_out, const float* __restrict__ grad_out, float* __restrict__ dlogits,
              int64_t ldd) {
    const int64_t total = M * C;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    const float g = loss_out[1] > 0.f ? grad_out[0] / loss_out[1] : 0.f;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t r = t / C;
        const int c = static_cast<int>(t - r * C);
        const int64_t y = target[r];
        float v = 0.f;
        if (y != ignore_index && y >= 0 && y < C) {
            const float w = weight ? weight[y] : 1.0f;
            v = g * w * (expf(logits[r * ld + c] - lse[r]) - (c == y ? 1.0f : 0.0f));
        }
        dlogits[r * ldd + c] = v;
    }
}

// partial[blockIdx.x] = sum of squares of this block's grid-stride share (fixed order)
__global__ void __launch_bounds__(kTrainThreads)
sumsq_kernel(const float* __restrict__ x, int64_t n, float* __restrict__ partial) {
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    float s = 0.f;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) s = fmaf(x[i], x[i], s);
    __shared__ float sh[kTrainThreads];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = kTrainThreads / 2; o >= 1; o >>= 1) {
        if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}

__global__ void __launch_bounds__(kTrainThreads)
norm_finish_kernel(const float* __restrict__ partial, int n_blocks, float* __restrict__ out) {
    __shared__ double sh[kTrainThreads];
    double s = 0.0;
    for (int k = threadIdx.x; k < n_blocks; k += kTrainThreads) s += static_cast<double>(partial[k]);
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = kTrainThreads / 2; o >= 1; o >>= 1) {
        if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = static_cast<float>(sqrt(sh[0]));
}

// torch.optim.Adam (amsgrad = False, maximize = False), single-tensor formulas:
//   g += wd * p;  m.lerp_(g, 1 - b1);  v = b2 * v + (1 - b2) * g * g;
//   p -= (lr / bc1) * m / (sqrt(v) / sqrt(bc2) + eps)
// g is first scaled by the clip coefficient min(1, max_norm / (norm + 1e-6)) (clip_grad_norm_) and written back.
__global__ void __launch_bounds__(kTrainThreads)
adam_step_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, int64_t n,
                 float lr, float b1, float b2, float eps, float wd, float bc1, float bc2_sqrt,
                 const float* __restrict__ norm, float max_norm) {
    float coef = 1.0f;
    if (norm && max_norm > 0.0f) coef = fminf(max_norm / (norm[0] + 1e-6f), 1.0f);
    const float step_size = lr / bc1;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        float gi = g[i] * coef;
        g[i] = gi;
        const float pi = p[i];
        if (wd != 0.0f) gi = fmaf(wd, pi, gi);
        const float mi = m[i] + (gi - m[i]) * (1.0f - b1);
        const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
        m[i] = mi;
        v[i] = vi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        p[i] = pi - step_size * (mi / denom);
    }
}

static int grid_for(int64_t n_items, int per_block) {
    int64_t b = (n_items + per_block - 1) / per_block;
    if (b > kNumSMs * 8) b = kNumSMs * 8;
    if (b < 1) b = 1;
    return static_cast<int>(b);
}

}  // namespace grl

using namespace grl;

extern "C" size_t grl_reduce_workspace_bytes(void) { return sizeof(float) * 3 * kNumSMs * 8; }

extern "C" int grl_ce_fwd(const float* logits, int64_t ld, const int64_t* target, const float* weight,
                          int64_t ignore_index, int64_t M, int C, float* lse, float* loss_out, void* workspace,
                          void* stream) {
    GRL_CHECK_ARG(logits && target && lse && loss_out && workspace, GRL_E_BADARG, "grl_ce_fwd: null pointer");
    GRL_CHECK_ARG(M > 0 && C > 0 && ld >= C, GRL_E_BADARG, "grl_ce_fwd: bad shape");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int blocks = grid_for(M, kTrainThreads / 32);
    float* partial = static_cast<float*>(workspace);
    ce_fwd_kernel<<<blocks, kTrainThreads, 0, st>>>(logits, ld, target, weight, ignore_index, M, C, lse, partial);
    GRL_LAUNCH_CHECK("ce_fwd_kernel");
    ce_finish_kernel<<<1, kTrainThreads, 0, st>>>(partial, blocks, loss_out);
    GRL_LAUNCH_CHECK("ce_finish_kernel");
    return 0;
}

extern "C" int grl_ce_bwd(const float* logits, int64_t ld, const int64_t* target, const float* weight,
                          int64_t ignore_index, int64_t M, int C, const float* lse, const float* loss_out,
                          const float* grad_out, float* dlogits, int64_t ldd, void* stream) {
    GRL_CHECK_ARG(logits && target && lse && loss_out && grad_out && dlogits, GRL_E_BADARG, "grl_ce_bwd: null pointer");
    GRL_CHECK_ARG(M > 0 && C > 0 && ld >= C && ldd >= C, GRL_E_BADARG, "grl_ce_bwd: bad shape");
    ce_bwd_kernel<<<grid_for(M * C, kTrainThreads), kTrainThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        logits, ld, target, weight, ignore_index, M, C, lse, loss_out, grad_out, dlogits, ldd);
    GRL_LAUNCH_CHECK("ce_bwd_kernel");
    return 0;
}

extern "C" int grl_l2_norm(const float* x, int64_t n, float* out, void* workspace, void* stream) {
    GRL_CHECK_ARG(x && out && workspace, GRL_E_BADARG, "grl_l2_norm: null pointer");
    GRL_CHECK_ARG(n > 0, GRL_E_BADARG, "grl_l2_norm: n <= 0");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int blocks = grid_for(n, kTrainThreads * 4);
    float* partial = static_cast<float*>(workspace);
    sumsq_kernel<<<blocks, kTrainThreads, 0, st>>>(x, n, partial);
    GRL_LAUNCH_CHECK("sumsq_kernel");
    norm_finish_kernel<<<1, kTrainThreads, 0, st>>>(partial, blocks, out);
    GRL_LAUNCH_CHECK("norm_finish_kernel");
    return 0;
}

// ---------------------------------------------------------------- grl_l2_norm without any allocation by anybody
// The block partial sums of sumsq_kernel live in a file-scope device array instead of a caller workspace, for callers that
// cannot allocate (CUDA graph capture, serving loops).  Every GPU that loads the library holds the 16 KB until exit.  The
// array's name is only a host-side handle here: its device address is looked up once per device, with that device
// current, and that pointer is what the kernels get.  Nothing in the array has to start at zero (sumsq_kernel overwrites
// the `blocks` partials norm_finish_kernel reads); calls that overlap on one GPU (two streams) race on it.
namespace grl {
constexpr int kL2StaticFloats = 4096;                                  // 16 KB of block partials
constexpr int64_t kL2StaticMaxN = int64_t(1) << 28;                    // 268 435 456 floats (1 GiB of gradients)
static __device__ __align__(256) float g_l2_static_scratch[kL2StaticFloats];

static int l2_static_scratch_ptr(float** out) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static float* ptrs[kMaxDevices] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("grl_l2_norm_static: device ordinal %d outside [0, %d)", dev, kMaxDevices);
        return GRL_E_UNSUP;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        float* p = nullptr;
        e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_l2_static_scratch);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_l2_static_scratch)");
        if (!p) {
            set_error("grl_l2_norm_static: cudaGetSymbolAddress returned a null device address");
            return static_cast<int>(cudaErrorInvalidSymbol);
        }
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return 0;
}
}  // namespace grl

extern "C" int grl_l2_norm_static(const float* x, int64_t n, float* out, void* stream) {
    GRL_CHECK_ARG(x && out, GRL_E_BADARG, "grl_l2_norm_static: null pointer");
    GRL_CHECK_ARG(n > 0, GRL_E_BADARG, "grl_l2_norm_static: n <= 0");
    GRL_CHECK_ARG(n <= kL2StaticMaxN, GRL_E_UNSUP,
                  "grl_l2_norm_static: n = %lld exceeds the %lld elements this entry point serves (use grl_l2_norm with a "
                  "caller workspace)", static_cast<long long>(n), static_cast<long long>(kL2StaticMaxN));
    const int blocks = grid_for(n, kTrainThreads * 4);                 // the same launch shape as grl_l2_norm: same bits
    GRL_CHECK_ARG(blocks <= kL2StaticFloats, GRL_E_UNSUP, "grl_l2_norm_static: %d block partials exceed the static scratch (%d)",
                  blocks, kL2StaticFloats);
    float* partial = nullptr;
    const int rc = l2_static_scratch_ptr(&partial);
    if (rc != 0) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    sumsq_kernel<<<blocks, kTrainThreads, 0, st>>>(x, n, partial);
    GRL_LAUNCH_CHECK("sumsq_kernel");
    norm_finish_kernel<<<1, kTrainThreads, 0, st>>>(partial, blocks, out);
    GRL_LAUNCH_CHECK("norm_finish_kernel");
    return 0;
}

extern "C" int grl_adam_step(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                             float eps, float weight_decay, int64_t step, const float* grad_norm, float max_norm,
                             void* stream) {
    GRL_CHECK_ARG(p && g && m && v, GRL_E_BADARG, "grl_adam_step: null pointer");
    GRL_CHECK_ARG(n > 0 && step >= 1, GRL_E_BADARG, "grl_adam_step: n <= 0 or step < 1");
    const double bc1 = 1.0 - pow(static_cast<double>(beta1), static_cast<double>(step));
    const double bc2 = 1.0 - pow(static_cast<double>(beta2), static_cast<double>(step));
    adam_step_kernel<<<grid_for(n, kTrainThreads * 4), kTrainThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        p, g, m, v, n, lr, beta1, beta2, eps, weight_decay, static_cast<float>(bc1), static_cast<float>(sqrt(bc2)),
        grad_norm, max_norm);
    GRL_LAUNCH_CHECK("adam_step_kernel");
    return 0;
}
