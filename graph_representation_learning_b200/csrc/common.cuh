// Shared device/host helpers for the sm_100a GraphCNNDropEdge kernels.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/grl_b200.h"

namespace grl {

// ---------------------------------------------------------------- error plumbing (host)
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define GRL_CHECK_ARG(cond, code, ...)                  \
    do {                                                \
        if (!(cond)) {                                  \
            ::grl::set_error(__VA_ARGS__);              \
            return (code);                              \
        }                                               \
    } while (0)

#define GRL_LAUNCH_CHECK(what)                          \
    do {                                                \
        cudaError_t e__ = cudaGetLastError();           \
        if (e__ != cudaSuccess) return ::grl::cuda_fail(e__, what); \
    } while (0)

// Optional device-side offset added to every Philox stream id (grl_rng_set_offset_ptr): a CUDA graph that captured the
// train step replays with fresh dropout / DropEdge masks by bumping ONE device counter between replays.
const uint64_t* rng_offset_ptr();
__device__ __forceinline__ uint64_t rng_stream(uint64_t stream_id, const uint64_t* off) { return off ? stream_id + __ldg(off) : stream_id; }

static inline bool aligned_to(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

// ---------------------------------------------------------------- Philox4x32-10 (counter based)
// The same (seed, stream_id, index) always yields the same bit whichever kernel asks, so the
// forward (CSR order) and the transposed backward (CSC order through perm) agree on DropEdge.
struct Philox {
    static constexpr uint32_t kM0 = 0xD2511F53u, kM1 = 0xCD9E8D57u;
    static constexpr uint32_t kW0 = 0x9E3779B9u, kW1 = 0xBB67AE85u;

    __host__ __device__ static inline void mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
#ifdef __CUDA_ARCH__
        hi = __umulhi(a, b);
        lo = a * b;
#else
        uint64_t p = static_cast<uint64_t>(a) * b;
        hi = static_cast<uint32_t>(p >> 32);
        lo = static_cast<uint32_t>(p);
#endif
    }

    // counter = (idx4 lo, idx4 hi, stream lo, stream hi); key = seed
    __host__ __device__ static inline uint4 draw(uint64_t seed, uint64_t stream_id, uint64_t idx4) {
        uint32_t c0 = static_cast<uint32_t>(idx4), c1 = static_cast<uint32_t>(idx4 >> 32);
        uint32_t c2 = static_cast<uint32_t>(stream_id), c3 = static_cast<uint32_t>(stream_id >> 32);
        uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            uint32_t hi0, lo0, hi1, lo1;
            mulhilo(kM0, c0, hi0, lo0);
            mulhilo(kM1, c2, hi1, lo1);
            uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
            c0 = n0; c1 = n1; c2 = n2; c3 = n3;
            k0 += kW0; k1 += kW1;
        }
        return make_uint4(c0, c1, c2, c3);
    }

    // one 32-bit word for element `idx` (4 consecutive elements share one Philox block)
    __host__ __device__ static inline uint32_t word(uint64_t seed, uint64_t stream_id, uint64_t idx) {
        uint4 r = draw(seed, stream_id, idx >> 2);
        switch (idx & 3u) {
            case 0: return r.x;
            case 1: return r.y;
            case 2: return r.z;
            default: return r.w;
        }
    }
};

// keep iff u >= p with u = word * 2^-32 in [0,1)  <=>  word >= thresh,  thresh = ceil(p * 2^32)
__host__ __device__ static inline uint32_t keep_threshold(float p) {
    double t = static_cast<double>(p) * 4294967296.0;
    if (t <= 0.0) return 0u;
    if (t >= 4294967295.0) return 0xFFFFFFFFu;
    return static_cast<uint32_t>(t);
}

// ---------------------------------------------------------------- 128-bit vector helpers
template <typename T> struct Vec16;  // 16-byte vector of T, unpacked to floats
template <> struct Vec16<float> {
    static constexpr int kElems = 4;
    __device__ static inline void load(const float* p, float (&v)[4]) {
        float4 t = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    __device__ static inline void store(float* p, const float (&v)[4]) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
    __device__ static inline void store_streaming(float* p, const float (&v)[4]) {
        __stcs(reinterpret_cast<float4*>(p), make_float4(v[0], v[1], v[2], v[3]));
    }
};
template <> struct Vec16<__nv_bfloat16> {
    static constexpr int kElems = 8;
    __device__ static inline void load(const __nv_bfloat16* p, float (&v)[8]) {
        uint4 t = __ldg(reinterpret_cast<const uint4*>(p));
        const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[2 * i] = __uint_as_float(w[i] << 16);
            v[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
        }
    }
    __device__ static inline uint4 pack(const float (&v)[8]) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
            w[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
    __device__ static inline void store(__nv_bfloat16* p, const float (&v)[8]) {
        *reinterpret_cast<uint4*>(p) = pack(v);
    }
    __device__ static inline void store_streaming(__nv_bfloat16* p, const float (&v)[8]) {
        __stcs(reinterpret_cast<uint4*>(p), pack(v));
    }
};

__device__ static inline float to_float(float x) { return x; }
__device__ static inline float to_float(__nv_bfloat16 x) { return __bfloat162float(x); }
__device__ static inline float to_float(__half x) { return __half2float(x); }
template <typename T> __device__ static inline T from_float(float x);
template <> __device__ inline float from_float<float>(float x) { return x; }
template <> __device__ inline __nv_bfloat16 from_float<__nv_bfloat16>(float x) { return __float2bfloat16_rn(x); }

constexpr int kNumSMs = 148;  // B200
// negative slope of the leaky-ReLU epilogues: F.leaky_relu's default 0.01 (GRL_EPI_LEAKY) or 0.2 (GRL_EPI_LEAKY02)
constexpr uint32_t kEpiAnyLeaky = GRL_EPI_LEAKY | GRL_EPI_LEAKY02;
__host__ __device__ static inline float leaky_slope(uint32_t flags) { return (flags & GRL_EPI_LEAKY02) ? 0.2f : 0.01f; }

}  // namespace grl
