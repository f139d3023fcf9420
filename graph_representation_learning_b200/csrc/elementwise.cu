// Fused bias + ReLU + dropout epilogue (and its gradient), column sums for bias gradients.
//
// Reference call sites: F.relu / nn.Dropout(p=0.5) / torch.cat in
// gnn/models/networks/drop_robust_gcn.py:64,76-77,80-81,84-89,99-100 and the bias add of
// gnn/models/networks/robust_gcn.py:50.  Writing with ldy > C places the result directly in a
// column slice of the concat buffer, so torch.cat never runs.
#include "common.cuh"

namespace grl {

// 4 consecutive elements; `vec` (warp-uniform) says the address is 4*sizeof(T)-aligned.
template <typename T> __device__ __forceinline__ void load4(const T* p, bool vec, float (&v)[4]);
template <> __device__ __forceinline__ void load4<float>(const float* p, bool vec, float (&v)[4]) {
    if (vec) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = p[k];
    }
}
template <> __device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16* p, bool vec, float (&v)[4]) {
    if (vec) {
        const uint2 t = *reinterpret_cast<const uint2*>(p);
        v[0] = __uint_as_float(t.x << 16); v[1] = __uint_as_float(t.x & 0xFFFF0000u);
        v[2] = __uint_as_float(t.y << 16); v[3] = __uint_as_float(t.y & 0xFFFF0000u);
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = __bfloat162float(p[k]);
    }
}
template <typename T> __device__ __forceinline__ void store4(T* p, bool vec, const float (&v)[4]);
template <> __device__ __forceinline__ void store4<float>(float* p, bool vec, const float (&v)[4]) {
    if (vec) *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    else {
#pragma unroll
        for (int k = 0; k < 4; ++k) p[k] = v[k];
    }
}
template <> __device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, bool vec, const float (&v)[4]) {
    if (vec) {
        __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
        *reinterpret_cast<uint2*>(p) = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) p[k] = __float2bfloat16_rn(v[k]);
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
bias_act_dropout_fwd_kernel(const T* X, int64_t ldx, const float* __restrict__ bias, T* Y,
                            int64_t ldy, int64_t M, int C, uint32_t flags, float scale, uint32_t thresh,
                            const uint8_t* __restrict__ mask, uint64_t seed, uint64_t stream_id, bool vec,
                            const uint64_t* __restrict__ rng_off) {
    // one thread = 4 consecutive columns of one row (C % 4 == 0): one Philox block per thread
    stream_id = rng_stream(stream_id, rng_off);
    const int c4 = C >> 2;
    const int64_t total = M * c4;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t m = t / c4;
        const int c = static_cast<int>(t - m * c4) << 2;
        float v[4];
        load4<T>(X + m * ldx + c, vec, v);
        if (flags & GRL_EPI_BIAS) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(bias + c));
            v[0] += b.x; v[1] += b.y; v[2] += b.z; v[3] += b.w;
        }
        if (flags & GRL_EPI_RELU) {
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = fmaxf(v[k], 0.0f);
        } else if (flags & kEpiAnyLeaky) {
            const float slope = leaky_slope(flags);
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = v[k] > 0.0f ? v[k] : slope * v[k];
        }
        if (flags & GRL_EPI_DROPOUT) {
            const int64_t e0 = m * C + c;  // logical element index, independent of ld
            bool keep[4];
            if (mask) {
#pragma unroll
                for (int k = 0; k < 4; ++k) keep[k] = mask[e0 + k] != 0;
            } else {
                const uint4 d = Philox::draw(seed, stream_id, static_cast<uint64_t>(e0) >> 2);
                keep[0] = d.x >= thresh; keep[1] = d.y >= thresh; keep[2] = d.z >= thresh; keep[3] = d.w >= thresh;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = keep[k] ? v[k] * scale : 0.0f;
        }
        store4<T>(Y + m * ldy + c, vec, v);
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
bias_act_dropout_bwd_kernel(const T* dY, int64_t lddy, const T* Y, int64_t ldy,
                            T* dX, int64_t lddx, int64_t M, int C, bool relu, float scale, float neg, float zero,
                            bool vec) {
    const int c4 = C >> 2;
    const int64_t total = M * c4;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t m = t / c4;
        const int c = static_cast<int>(t - m * c4) << 2;
        float g[4], y[4] = {1.0f, 1.0f, 1.0f, 1.0f};
        load4<T>(dY + m * lddy + c, vec, g);
        if (relu) load4<T>(Y + m * ldy + c, vec, y);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float f = y[k] > 0.0f ? scale : (y[k] < 0.0f ? neg : zero);
            g[k] = f != 0.0f ? g[k] * f : 0.0f;          // +0 (not -0) where the derivative is zero
        }
        store4<T>(dX + m * lddx + c, vec, g);
    }
}

// Any-width variants (C % 4 != 0, e.g. the 2-wide attention projections of tiny models or the 53-wide head):
// one thread per element; Philox::word(idx) picks the same bit the 4-wide kernel would.
template <typename T>
__global__ void __launch_bounds__(256)
bias_act_dropout_fwd_scalar_kernel(const T* X, int64_t ldx, const float* __restrict__ bias, T* Y, int64_t ldy, int64_t M,
                                   int C, uint32_t flags, float scale, uint32_t thresh,
                                   const uint8_t* __restrict__ mask, uint64_t seed, uint64_t stream_id,
                                   const uint64_t* __restrict__ rng_off) {
    stream_id = rng_stream(stream_id, rng_off);
    const int64_t total = M * C;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t m = t / C;
        const int c = static_cast<int>(t - m * C);
        float v = to_float(X[m * ldx + c]);
        if (flags & GRL_EPI_BIAS) v += bias[c];
        if (flags & GRL_EPI_RELU) v = fmaxf(v, 0.0f);
        else if (flags & kEpiAnyLeaky) v = v > 0.0f ? v : leaky_slope(flags) * v;
        if (flags & GRL_EPI_DROPOUT) {
            const bool keep = mask ? (mask[t] != 0) : (Philox::word(seed, stream_id, static_cast<uint64_t>(t)) >= thresh);
            v = keep ? v * scale : 0.0f;
        }
        Y[m * ldy + c] = from_float<T>(v);
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
bias_act_dropout_bwd_scalar_kernel(const T* dY, int64_t lddy, const T* Y, int64_t ldy, T* dX, int64_t lddx, int64_t M,
                                   int C, bool relu, float scale, float neg, float zero) {
    const int64_t total = M * C;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t m = t / C;
        const int c = static_cast<int>(t - m * C);
        const float y = relu ? to_float(Y[m * ldy + c]) : 1.0f;
        const float f = y > 0.0f ? scale : (y < 0.0f ? neg : zero);
        dX[m * lddx + c] = from_float<T>(f != 0.0f ? to_float(dY[m * lddy + c]) * f : 0.0f);
    }
}

// out[c] += sum over a slab of rows; one block = 32 x 8 threads covering 32 columns, rows strided by gridDim.y.
template <typename T>
__global__ void colsum_kernel(const T* __restrict__ X, int64_t ldx, int64_t M, int C, float* __restrict__ out) {
    __shared__ float part[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    float acc = 0.0f;
    if (c < C) {
        for (int64_t m = static_cast<int64_t>(blockIdx.y) * 8 + threadIdx.y; m < M; m += static_cast<int64_t>(gridDim.y) * 8)
            acc += to_float(X[m * ldx + c]);
    }
    part[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && c < C) {
        float s = 0.0f;
#pragma unroll
        for (int k = 0; k < 8; ++k) s += part[k][threadIdx.x];
        atomicAdd(out + c, s);
    }
}

static inline int ew_grid(int64_t total) {
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = static_cast<int64_t>(kNumSMs) * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

}  // namespace grl

using namespace grl;

extern "C" int grl_bias_act_dropout_fwd(const void* X, int64_t ldx, const float* bias, void* Y, int64_t ldy, int dtype,
                                        int64_t M, int C, uint32_t flags, float p, const uint8_t* mask, uint64_t seed,
                                        uint64_t stream_id, void* stream) {
    GRL_CHECK_ARG(X && Y, GRL_E_BADARG, "grl_bias_act_dropout_fwd: null pointer");
    GRL_CHECK_ARG(!(flags & GRL_EPI_BIAS) || bias, GRL_E_BADARG, "grl_bias_act_dropout_fwd: BIAS flag without bias");
    GRL_CHECK_ARG(M >= 0 && C > 0, GRL_E_BADARG, "grl_bias_act_dropout_fwd: bad shape M=%lld C=%d", (long long)M, C);
    GRL_CHECK_ARG(ldx >= C && ldy >= C, GRL_E_BADARG, "grl_bias_act_dropout_fwd: leading dimension < C");
    GRL_CHECK_ARG(p >= 0.0f && p < 1.0f, GRL_E_BADARG, "grl_bias_act_dropout_fwd: p outside [0,1)");
    if (M == 0) return 0;
    if (p == 0.0f) flags &= ~GRL_EPI_DROPOUT;
    const float scale = 1.0f / (1.0f - p);
    const uint32_t thresh = keep_threshold(p);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (C % 4 != 0 || ((flags & GRL_EPI_BIAS) && !aligned_to(bias, 16))) {
        const int sgrid = ew_grid(M * C);
        if (dtype == GRL_F32)
            bias_act_dropout_fwd_scalar_kernel<float><<<sgrid, 256, 0, st>>>(
                static_cast<const float*>(X), ldx, bias, static_cast<float*>(Y), ldy, M, C, flags, scale, thresh, mask, seed,
                stream_id, rng_offset_ptr());
        else if (dtype == GRL_BF16)
            bias_act_dropout_fwd_scalar_kernel<__nv_bfloat16><<<sgrid, 256, 0, st>>>(
                static_cast<const __nv_bfloat16*>(X), ldx, bias, static_cast<__nv_bfloat16*>(Y), ldy, M, C, flags, scale,
                thresh, mask, seed, stream_id, rng_offset_ptr());
        else {
            set_error("grl_bias_act_dropout_fwd: dtype %d unsupported", dtype);
            return GRL_E_UNSUP;
        }
        GRL_LAUNCH_CHECK("bias_act_dropout_fwd_scalar_kernel");
        return 0;
    }
    const int grid = ew_grid(M * (C / 4));
    const size_t a4 = 4 * (dtype == GRL_F32 ? 4 : 2);
    const bool vec = aligned_to(X, a4) && aligned_to(Y, a4) && ldx % 4 == 0 && ldy % 4 == 0;
    if (dtype == GRL_F32)
        bias_act_dropout_fwd_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(X), ldx, bias,
                                                                static_cast<float*>(Y), ldy, M, C, flags, scale, thresh,
                                                                mask, seed, stream_id, vec, rng_offset_ptr());
    else if (dtype == GRL_BF16)
        bias_act_dropout_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(X), ldx, bias,
                                                                        static_cast<__nv_bfloat16*>(Y), ldy, M, C, flags,
                                                                        scale, thresh, mask, seed, stream_id, vec, rng_offset_ptr());
    else {
        set_error("grl_bias_act_dropout_fwd: dtype %d unsupported", dtype);
        return GRL_E_UNSUP;
    }
    GRL_LAUNCH_CHECK("bias_act_dropout_fwd_kernel");
    return 0;
}

extern "C" int grl_bias_act_dropout_bwd(const void* dY, int64_t lddy, const void* Y, int64_t ldy, void* dX, int64_t lddx,
                                        int dtype, int64_t M, int C, uint32_t flags, float p, void* stream) {
    GRL_CHECK_ARG(dY && dX, GRL_E_BADARG, "grl_bias_act_dropout_bwd: null pointer");
    GRL_CHECK_ARG(M >= 0 && C > 0, GRL_E_BADARG, "grl_bias_act_dropout_bwd: bad shape M=%lld C=%d", (long long)M, C);
    // "relu" below = the derivative is read off the sign of the forward output Y (ReLU or leaky ReLU)
    const bool leaky = (flags & kEpiAnyLeaky) != 0 && !(flags & GRL_EPI_RELU);
    const float slope = leaky_slope(flags);
    const bool relu = (flags & GRL_EPI_RELU) != 0 || leaky;
    GRL_CHECK_ARG(relu || !(flags & GRL_EPI_DROPOUT) || p == 0.0f, GRL_E_UNSUP,
                  "grl_bias_act_dropout_bwd: dropout without an activation is not supported");
    GRL_CHECK_ARG(!relu || Y, GRL_E_BADARG, "grl_bias_act_dropout_bwd: RELU / LEAKY needs the forward output Y");
    if (M == 0) return 0;
    const bool drop = (flags & GRL_EPI_DROPOUT) && p > 0.0f;
    const float scale = drop ? 1.0f / (1.0f - p) : 1.0f;
    // y < 0: leaky slope; y == 0: a dropped element (0) or, without dropout, leaky ReLU at z == 0 (slope, as torch)
    const float neg = leaky ? slope * scale : 0.0f;
    const float zero = (leaky && !drop) ? slope : 0.0f;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (C % 4 != 0) {
        const int sgrid = ew_grid(M * C);
        if (dtype == GRL_F32)
            bias_act_dropout_bwd_scalar_kernel<float><<<sgrid, 256, 0, st>>>(
                static_cast<const float*>(dY), lddy, static_cast<const float*>(Y), ldy, static_cast<float*>(dX), lddx, M, C,
                relu, scale, neg, zero);
        else if (dtype == GRL_BF16)
            bias_act_dropout_bwd_scalar_kernel<__nv_bfloat16><<<sgrid, 256, 0, st>>>(
                static_cast<const __nv_bfloat16*>(dY), lddy, static_cast<const __nv_bfloat16*>(Y), ldy,
                static_cast<__nv_bfloat16*>(dX), lddx, M, C, relu, scale, neg, zero);
        else {
            set_error("grl_bias_act_dropout_bwd: dtype %d unsupported", dtype);
            return GRL_E_UNSUP;
        }
        GRL_LAUNCH_CHECK("bias_act_dropout_bwd_scalar_kernel");
        return 0;
    }
    const int grid = ew_grid(M * (C / 4));
    const size_t a4 = 4 * (dtype == GRL_F32 ? 4 : 2);
    const bool vec = aligned_to(dY, a4) && aligned_to(dX, a4) && (!relu || aligned_to(Y, a4)) && lddy % 4 == 0 &&
                     lddx % 4 == 0 && (!relu || ldy % 4 == 0);
    if (dtype == GRL_F32)
        bias_act_dropout_bwd_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(dY), lddy,
                                                                static_cast<const float*>(Y), ldy,
                                                                static_cast<float*>(dX), lddx, M, C, relu, scale, neg, zero,
                                                                vec);
    else if (dtype == GRL_BF16)
        bias_act_dropout_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(
            static_cast<const __nv_bfloat16*>(dY), lddy, static_cast<const __nv_bfloat16*>(Y), ldy,
            static_cast<__nv_bfloat16*>(dX), lddx, M, C, relu, scale, neg, zero, vec);
    else {
        set_error("grl_bias_act_dropout_bwd: dtype %d unsupported", dtype);
        return GRL_E_UNSUP;
    }
    GRL_LAUNCH_CHECK("bias_act_dropout_bwd_kernel");
    return 0;
}

extern "C" int grl_colsum(const void* X, int64_t ldx, int dtype, int64_t M, int C, float* out, void* stream) {
    GRL_CHECK_ARG(X && out, GRL_E_BADARG, "grl_colsum: null pointer");
    GRL_CHECK_ARG(M >= 0 && C > 0 && ldx >= C, GRL_E_BADARG, "grl_colsum: bad shape");
    if (M == 0) return 0;
    dim3 block(32, 8);
    int64_t gy = (M + 8 * 64 - 1) / (8 * 64);  // >= 64 rows per thread
    if (gy > 1024) gy = 1024;
    if (gy < 1) gy = 1;
    dim3 grid((C + 31) / 32, static_cast<unsigned>(gy));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == GRL_F32) colsum_kernel<float><<<grid, block, 0, st>>>(static_cast<const float*>(X), ldx, M, C, out);
    else if (dtype == GRL_BF16)
        colsum_kernel<__nv_bfloat16><<<grid, block, 0, st>>>(static_cast<const __nv_bfloat16*>(X), ldx, M, C, out);
    else {
        set_error("grl_colsum: dtype %d unsupported", dtype);
        return GRL_E_UNSUP;
    }
    GRL_LAUNCH_CHECK("colsum_kernel");
    return 0;
}
