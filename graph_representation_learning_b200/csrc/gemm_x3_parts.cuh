// Pieces shared by the fp32-grade GEMM kernels (gemm_x3_tcgen05.cu: three tf32 products; gemm_bc_tcgen05.cu: tf32 main
// product + bf16 correction product with the activation operand in TMEM): tile configuration, the register epilogue of
// the drain warps, the split-K plan and reduction of the weight-gradient kernels.
#pragma once
#include "gemm_common.cuh"

namespace grl {
using namespace sm100;

struct X3Cfg {
    static constexpr int BLOCK_M = 128, BLOCK_N = 128, BLOCK_K = 32, UMMA_K = 8;
    static constexpr int TILE_BYTES = 128 * 128;                    // every operand tile is 16 KB
    static constexpr int STAGE_BYTES = 4 * TILE_BYTES;              // A | B(_hi) | B_lo | A_lo
    static constexpr int STAGES = 3;
    static constexpr int DRAIN_WARPS = 8;                           // two per TMEM lane quarter: 64 columns each
    static constexpr int DRAIN_COLS = BLOCK_N / (DRAIN_WARPS / 4);
    static constexpr int EPI_COLS = 16;                             // columns transposed per pass
    static constexpr int EPI_PITCH = EPI_COLS + 4;                  // floats (conflict-free STS.128 rows)
    static constexpr int EPI_BYTES = DRAIN_WARPS * 32 * EPI_PITCH * 4;   // one 32 x 16 transposition buffer per drain warp
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 + 256;
    static constexpr int TMEM_COLS = 512;                           // main0 | main1 | small0 | small1
    static constexpr int THREADS = 64 + 128 + 32 * DRAIN_WARPS;
};

// Epilogue of one drain warp over ITS 32 accumulator rows x DRAIN_COLS columns held in registers (one row per lane):
// 16-column chunks are transposed through a private 32 x 20 fp32 buffer so that every global store instruction writes
// eight 64-byte row segments.  Private buffer + registers only: the operand stages keep streaming the NEXT tile.
__device__ __forceinline__ void epilogue_from_regs(const GemmParams& p, const float (&acc)[X3Cfg::DRAIN_COLS], float* buf,
                                                   int row0, int n0, int lane) {
    constexpr int PITCH = X3Cfg::EPI_PITCH, CH = X3Cfg::EPI_COLS;
    const bool vec_ok = epilogue_vec_ok(p);
    const int rsub = lane >> 2, c4 = (lane & 3) * 4;
    const int rows = min(32, p.M - row0);
    // Fast path (warp-uniform): no dropout, no atomics, aligned output, all 64 columns inside N.  The general path below
    // re-tests every flag and rebuilds a 64-bit address per 4 elements (~50 instructions per element group); short-K
    // products (GraphConv dX: 8 K-blocks per tile) were bound by exactly that instruction stream in the drain warps.
    const bool drop = (p.flags & GRL_EPI_DROPOUT) != 0;      // fast path: Philox dropout only (one draw per 4 columns)
    if ((!drop || (p.mask == nullptr && (p.N & 3) == 0)) && !p.atomic_add && vec_ok && n0 + X3Cfg::DRAIN_COLS <= p.N) {
        const bool relu = (p.flags & GRL_EPI_RELU) != 0;
        const bool leaky = !relu && (p.flags & kEpiAnyLeaky) != 0;
        const float slope = leaky_slope(p.flags);
        const int64_t ldc = p.ldc;
        const uint64_t sid = drop ? rng_stream(p.stream_id, p.rng_off) : 0ull;
#pragma unroll
        for (int c0 = 0; c0 < X3Cfg::DRAIN_COLS; c0 += CH) {
            __syncwarp();
#pragma unroll
            for (int j = 0; j < CH; j += 4)
                *reinterpret_cast<float4*>(buf + lane * PITCH + j) = make_float4(acc[c0 + j], acc[c0 + j + 1], acc[c0 + j + 2], acc[c0 + j + 3]);
            __syncwarp();
            const int col = n0 + c0 + c4;
            float4 b4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            if (p.flags & GRL_EPI_BIAS) b4 = make_float4(__ldg(p.bias + col), __ldg(p.bias + col + 1), __ldg(p.bias + col + 2), __ldg(p.bias + col + 3));
            const int64_t off = static_cast<int64_t>(row0 + rsub) * ldc + col;
            if (p.out_bf16) {
                __nv_bfloat16* out = static_cast<__nv_bfloat16*>(p.C) + off;
#pragma unroll
                for (int rr = rsub; rr < 32; rr += 8, out += 8 * ldc) {
                    if (rr < rows) {
                        float4 v = *reinterpret_cast<const float4*>(buf + rr * PITCH + c4);
                        v.x += b4.x; v.y += b4.y; v.z += b4.z; v.w += b4.w;
                        if (relu) {
                            v.x = fmaxf(v.x, 0.0f); v.y = fmaxf(v.y, 0.0f); v.z = fmaxf(v.z, 0.0f); v.w = fmaxf(v.w, 0.0f);
                        } else if (leaky) {
                            v.x = v.x > 0.0f ? v.x : slope * v.x; v.y = v.y > 0.0f ? v.y : slope * v.y;
                            v.z = v.z > 0.0f ? v.z : slope * v.z; v.w = v.w > 0.0f ? v.w : slope * v.w;
                        }
                        if (drop) {     // same (row * N + col) >> 2 Philox block as epilogue_store4 / grl_bias_act_dropout_fwd
                            const uint4 d = Philox::draw(p.seed, sid, static_cast<uint64_t>(static_cast<int64_t>(row0 + rr) * p.N + col) >> 2);
                            v.x = d.x >= p.thresh ? v.x * p.scale : 0.0f; v.y = d.y >= p.thresh ? v.y * p.scale : 0.0f;
                            v.z = d.z >= p.thresh ? v.z * p.scale : 0.0f; v.w = d.w >= p.thresh ? v.w * p.scale : 0.0f;
                        }
                        const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
                        *reinterpret_cast<uint2*>(out) = make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
                    }
                }
            } else {
                float* out = static_cast<float*>(p.C) + off;
#pragma unroll
                for (int rr = rsub; rr < 32; rr += 8, out += 8 * ldc) {
                    if (rr < rows) {
                        float4 v = *reinterpret_cast<const float4*>(buf + rr * PITCH + c4);
                        v.x += b4.x; v.y += b4.y; v.z += b4.z; v.w += b4.w;
                        if (relu) {
                            v.x = fmaxf(v.x, 0.0f); v.y = fmaxf(v.y, 0.0f); v.z = fmaxf(v.z, 0.0f); v.w = fmaxf(v.w, 0.0f);
                        } else if (leaky) {
                            v.x = v.x > 0.0f ? v.x : slope * v.x; v.y = v.y > 0.0f ? v.y : slope * v.y;
                            v.z = v.z > 0.0f ? v.z : slope * v.z; v.w = v.w > 0.0f ? v.w : slope * v.w;
                        }
                        if (drop) {
                            const uint4 d = Philox::draw(p.seed, sid, static_cast<uint64_t>(static_cast<int64_t>(row0 + rr) * p.N + col) >> 2);
                            v.x = d.x >= p.thresh ? v.x * p.scale : 0.0f; v.y = d.y >= p.thresh ? v.y * p.scale : 0.0f;
                            v.z = d.z >= p.thresh ? v.z * p.scale : 0.0f; v.w = d.w >= p.thresh ? v.w * p.scale : 0.0f;
                        }
                        *reinterpret_cast<float4*>(out) = v;
                    }
                }
            }
        }
        return;
    }
#pragma unroll
    for (int c0 = 0; c0 < X3Cfg::DRAIN_COLS; c0 += CH) {
        if (n0 + c0 >= p.N) break;                                  // warp-uniform
        __syncwarp();                                               // readers of the previous chunk are done
#pragma unroll
        for (int j = 0; j < CH; j += 4)
            *reinterpret_cast<float4*>(buf + lane * PITCH + j) = make_float4(acc[c0 + j], acc[c0 + j + 1], acc[c0 + j + 2], acc[c0 + j + 3]);
        __syncwarp();
        const int col = n0 + c0 + c4;
        if (col < p.N) {
            float b4[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            if (p.flags & GRL_EPI_BIAS) {
#pragma unroll
                for (int t = 0; t < 4; ++t)
                    if (col + t < p.N) b4[t] = __ldg(p.bias + col + t);
            }
#pragma unroll
            for (int rr = rsub; rr < 32; rr += 8) {
                if (rr < rows) {
                    const float4 t4 = *reinterpret_cast<const float4*>(buf + rr * PITCH + c4);
                    float v[4] = {t4.x + b4[0], t4.y + b4[1], t4.z + b4[2], t4.w + b4[3]};
                    epilogue_store4(p, row0 + rr, col, v, vec_ok);
                }
            }
        }
    }
}

int tn_plan_splits(int tiles, int kblocks);                                                        // gemm_x3_tcgen05.cu
int launch_reduce_splits(const float* ws, float* out, int64_t n, int splits, cudaStream_t st);      // gemm_x3_tcgen05.cu

}  // namespace grl
