// Pieces shared by the tcgen05 GEMM kernels: parameters, the fused epilogue, TMA tensor-map construction.
#pragma once
#include "common.cuh"
#include "tcgen05.cuh"

#include <mutex>

namespace grl {

enum GemmMode { kBF16 = 0, kTF32 = 1, kTF32X3 = 2, kTF32BC = 3 };

// fp32 bits -> nearest tf32 (10 explicit mantissa bits), ties away from zero.  Every operand split of the fp32-grade
// kernels rounds BOTH parts this way (x_hi = rn(x), x_lo = rn(x - x_hi)): handing the tensor core raw fp32 bits makes it
// TRUNCATE the 13 low bits, a bias of the same sign as x that adds up coherently over K (measured: 4-5e-7 norm-wise per
// product against 1-2e-7 with rounded parts).
__host__ __device__ __forceinline__ uint32_t tf32_rn_bits(uint32_t u) { return (u + 0x1000u) & 0xFFFFE000u; }

struct GemmParams {
    int M, N;
    int kblocks0, kblocks1;     // K blocks of segment 0 / 1 (segment 1 may be 0)
    void* C;
    int64_t ldc;
    int out_bf16;               // 0: fp32 output, 1: bf16 output
    int atomic_add;             // 1: red.global.add.f32 into C (split-K partial tiles; fp32 output only)
    const float* bias;
    uint32_t flags;
    float scale;                // 1/(1-p)
    uint32_t thresh;            // keep iff philox word >= thresh
    const uint8_t* mask;        // injected keep bits [M*N] or null
    uint64_t seed, stream_id;
    const uint64_t* rng_off;    // device-side offset added to stream_id (null: none); see common.cuh rng_stream
};

// Fused epilogue for 4 consecutive columns [col, col+4) of one output row held in registers (bias already added).
// Same Philox indexing (row*N + col) as grl_bias_act_dropout_fwd, so fused and unfused paths agree bit for bit.
__device__ __forceinline__ void epilogue_store4(const GemmParams& p, int row, int col, float (&v)[4], bool vec_ok) {
    if (p.flags & GRL_EPI_RELU) {
#pragma unroll
        for (int t = 0; t < 4; ++t) v[t] = fmaxf(v[t], 0.0f);
    } else if (p.flags & kEpiAnyLeaky) {
        const float slope = leaky_slope(p.flags);
#pragma unroll
        for (int t = 0; t < 4; ++t) v[t] = v[t] > 0.0f ? v[t] : slope * v[t];
    }
    if (p.flags & GRL_EPI_DROPOUT) {
        const int64_t e0 = static_cast<int64_t>(row) * p.N + col;
        bool keep[4];
        if (p.mask) {
#pragma unroll
            for (int t = 0; t < 4; ++t) keep[t] = (col + t < p.N) && p.mask[e0 + t] != 0;
        } else if ((e0 & 3) == 0) {       // the 4 elements share one Philox block (always when N % 4 == 0)
            const uint4 d = Philox::draw(p.seed, rng_stream(p.stream_id, p.rng_off), static_cast<uint64_t>(e0) >> 2);
            keep[0] = d.x >= p.thresh; keep[1] = d.y >= p.thresh; keep[2] = d.z >= p.thresh; keep[3] = d.w >= p.thresh;
        } else {
#pragma unroll
            for (int t = 0; t < 4; ++t)
                keep[t] = Philox::word(p.seed, rng_stream(p.stream_id, p.rng_off), static_cast<uint64_t>(e0 + t)) >= p.thresh;
        }
#pragma unroll
        for (int t = 0; t < 4; ++t) v[t] = keep[t] ? v[t] * p.scale : 0.0f;
    }
    if (p.out_bf16) {
        __nv_bfloat16* out = static_cast<__nv_bfloat16*>(p.C) + static_cast<int64_t>(row) * p.ldc + col;
        if (vec_ok && col + 3 < p.N) {
            __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
            *reinterpret_cast<uint2*>(out) = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
        } else {
#pragma unroll
            for (int t = 0; t < 4; ++t)
                if (col + t < p.N) out[t] = __float2bfloat16_rn(v[t]);
        }
    } else {
        float* out = static_cast<float*>(p.C) + static_cast<int64_t>(row) * p.ldc + col;
        if (p.atomic_add) {
#pragma unroll
            for (int t = 0; t < 4; ++t)
                if (col + t < p.N) atomicAdd(out + t, v[t]);
        } else if (vec_ok && col + 3 < p.N) {
            *reinterpret_cast<float4*>(out) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
#pragma unroll
            for (int t = 0; t < 4; ++t)
                if (col + t < p.N) out[t] = v[t];
        }
    }
}

__device__ __forceinline__ bool epilogue_vec_ok(const GemmParams& p) {
    const int eb = p.out_bf16 ? 2 : 4;
    return (p.ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.C) % (4 * eb)) == 0);
}

// Epilogue of one warp over ITS 32 accumulator rows, staged in shared memory as fp32 [128][BLOCK_N + 4] (the pad keeps
// the row-per-thread staging stores and the row-contiguous reads below bank-conflict free).  A compact loop -- the
// fully unrolled register version was 10k straight-line instructions per warp and ran out of the instruction cache --
// and every global store is one contiguous 512-byte (fp32) / 256-byte (bf16) row segment per warp.
template <int BLOCK_N>
__device__ __forceinline__ void epilogue_warp_rows(const GemmParams& p, const float* stage, int warp_row0, int m0, int n0,
                                                   int lane) {
    constexpr int PITCH = BLOCK_N + 4;
    const bool vec_ok = epilogue_vec_ok(p);
#pragma unroll 1
    for (int cpass = 0; cpass < BLOCK_N; cpass += 128) {
        const int cl = cpass + lane * 4;
        const int col = n0 + cl;
        if (cl >= BLOCK_N || col >= p.N) continue;
        float b4[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        if (p.flags & GRL_EPI_BIAS) {
#pragma unroll
            for (int t = 0; t < 4; ++t)
                if (col + t < p.N) b4[t] = __ldg(p.bias + col + t);
        }
        const int rows = min(32, p.M - (m0 + warp_row0));
#pragma unroll 2
        for (int rr = 0; rr < rows; ++rr) {
            const float4 t4 = *reinterpret_cast<const float4*>(stage + (warp_row0 + rr) * PITCH + cl);
            float v[4] = {t4.x + b4[0], t4.y + b4[1], t4.z + b4[2], t4.w + b4[3]};
            epilogue_store4(p, m0 + warp_row0 + rr, col, v, vec_ok);
        }
    }
}

// ------------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    });
    return fn;
}

// Row-major matrix [rows, cols] with leading dimension ld (elements); box = one 128-byte swizzle atom along the
// contiguous dimension x box_rows rows.  K-major operands: cols = K.  MN-major operands: cols = the M/N extent.
// atom32 = true selects CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B (32-byte chunks XOR (row % 4)), the only shared-memory
// layout the tensor core accepts for MN-major tf32 operands (UMMA layout type SWIZZLE_128B_BASE32B).
inline int make_tmap_sw128(CUtensorMap* out, const void* ptr, int elem_bytes, int64_t rows, int64_t cols, int64_t ld,
                           int box_rows, bool atom32 = false) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return GRL_E_UNSUP;
    }
    const cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    const cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * elem_bytes};
    const cuuint32_t box[2] = {static_cast<cuuint32_t>(128 / elem_bytes), static_cast<cuuint32_t>(box_rows)};
    const cuuint32_t estr[2] = {1, 1};
    const CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    CUresult r = enc(out, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d): ptr=%p rows=%lld cols=%lld ld=%lld elem=%d box_rows=%d",
                  static_cast<int>(r), ptr, static_cast<long long>(rows), static_cast<long long>(cols),
                  static_cast<long long>(ld), elem_bytes, box_rows);
        return GRL_E_BADARG;
    }
    return 0;
}

// Plain (un-swizzled) 2-D tile map over a row-major [rows, cols] matrix: box = box_cols x box_rows elements, rows and
// columns past the tensor are zero-filled.  Used by the SpMM kernels to stage one graph's feature slab in shared memory.
inline int make_tmap_plain(CUtensorMap* out, const void* ptr, int elem_bytes, int64_t rows, int64_t cols, int64_t ld,
                           int box_cols, int box_rows) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return GRL_E_UNSUP;
    }
    const cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    const cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * elem_bytes};
    const cuuint32_t box[2] = {static_cast<cuuint32_t>(box_cols), static_cast<cuuint32_t>(box_rows)};
    const cuuint32_t estr[2] = {1, 1};
    const CUtensorMapDataType dt = elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    CUresult r = enc(out, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(plain) failed (%d): ptr=%p rows=%lld cols=%lld ld=%lld elem=%d box=%dx%d",
                  static_cast<int>(r), ptr, static_cast<long long>(rows), static_cast<long long>(cols),
                  static_cast<long long>(ld), elem_bytes, box_cols, box_rows);
        return GRL_E_BADARG;
    }
    return 0;
}

struct GemmArgs {
    const void *A0, *B0, *B0lo, *A1, *B1, *B1lo;
    int64_t lda0, ldb0, lda1, ldb1;
    int K0, K1;
};

int launch_gemm_x3(const GemmArgs& a, GemmParams p, cudaStream_t st);   // gemm_x3_tcgen05.cu
int launch_gemm_bc(const GemmArgs& a, GemmParams p, cudaStream_t st, bool x3, int promo_x3);   // gemm_bc_tcgen05.cu
// CTA-pair (cta_group::2) variant of the NT product; returns 1 when it is switched off / the product is too small for it
int try_launch_gemm_bc_2sm(const GemmArgs& a, GemmParams p, cudaStream_t st, bool x3, int promo);   // gemm_bc_2sm.cu
int try_launch_gemm_nt_bf16_2sm(const GemmArgs& a, GemmParams p, cudaStream_t st);   // gemm_bc_2sm.cu
int try_launch_gemm_bc_tn_2sm(GemmParams p, int promo, int transpose_out, const CUtensorMap& tA, const CUtensorMap& tB,
                              const CUtensorMap& tBp, int* splits_out, size_t workspace_bytes, cudaStream_t st);   // gemm_bc_2sm.cu

}  // namespace grl
