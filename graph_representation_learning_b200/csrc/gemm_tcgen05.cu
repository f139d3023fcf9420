// tcgen05 / TMEM / TMA GEMM with fused bias + ReLU + dropout epilogue for the dense projections of the
// GraphCNNDropEdge stack (reference: nn.Linear / torch.matmul(new_V, h_weights) --
// gnn/models/networks/robust_gcn.py:10-11,50; drop_robust_gcn.py:59,64,89,99-103 -- and their autograd).
//
//   C[M, N] = epilogue( sum_seg  A_seg[M, K_seg] * B_seg[N, K_seg]^T )        ("NT": both operands K-major)
//
// * up to two K segments = the reference's torch.cat([g1, g2]) / cat([g1, g3]) consumed without a concat buffer;
// * this file: single-pass kernels, BF16 (kind::f16) and TF32 (kind::tf32, fp32 operands, ~1e-3);
//   the fp32-grade error-compensated TF32X3 kernel lives in gemm_x3_tcgen05.cu;
// * one CTA = one 128 x BLOCK_N output tile; warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner,
//   warps 2..5 = epilogue: tcgen05.ld -> bias/ReLU/dropout -> global.  fp32 accumulators live in TMEM.
#include "gemm_common.cuh"
#include "gemm_x3_parts.cuh"

namespace grl {
using namespace sm100;

template <int kMode, int BLOCK_N>
struct GemmCfg {
    static constexpr int BLOCK_M = 128;
    static constexpr int ELEM = (kMode == kBF16) ? 2 : 4;
    static constexpr int BLOCK_K = 128 / ELEM;          // one 128-byte swizzle atom along K
    static constexpr int UMMA_K = 32 / ELEM;            // 16 (bf16) / 8 (tf32)
    static constexpr int A_BYTES = BLOCK_M * 128;
    static constexpr int B_BYTES = BLOCK_N * 128;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int kMaxSmem = 227 * 1024 - 1024 /*align slack*/ - 256 /*barriers*/;
    static constexpr int STAGES_RAW = kMaxSmem / STAGE_BYTES;
    static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
    static constexpr int TMEM_COLS = BLOCK_N < 32 ? 32 : BLOCK_N;
    static_assert(STAGES >= 2, "pipeline needs at least two stages");
    static_assert(BLOCK_N % 16 == 0 && BLOCK_N >= 16 && BLOCK_N <= 256, "UMMA N for M=128");
};

constexpr int kGemmThreads = 192;

template <int kMode, int BLOCK_N>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_nt_kernel(const GemmParams p, const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmB0,
               const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1) {
    static_assert(kMode == kBF16 || kMode == kTF32, "TF32X3 has its own kernel (gemm_nt_x3_kernel)");
    using Cfg = GemmCfg<kMode, BLOCK_N>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    uint64_t* bars = reinterpret_cast<uint64_t*>(tiles + STAGES * Cfg::STAGE_BYTES);
    uint64_t* full_bar = bars;                 // [STAGES] TMA bytes landed
    uint64_t* empty_bar = bars + STAGES;       // [STAGES] MMAs reading the stage retired
    uint64_t* accum_bar = bars + 2 * STAGES;   // accumulator complete
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 1);

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);   // provably warp-uniform
    const int lane = threadIdx.x & 31;
    const int tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
    const int m0 = (blockIdx.x / tiles_n) * Cfg::BLOCK_M;
    const int n0 = (blockIdx.x % tiles_n) * BLOCK_N;
    const int total_kb = p.kblocks0 + p.kblocks1;

    auto stage_a = [&](int s) { return tiles + s * Cfg::STAGE_BYTES; };
    auto stage_b = [&](int s) { return tiles + s * Cfg::STAGE_BYTES + Cfg::A_BYTES; };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA0);
        tma_prefetch_desc(&tmB0);
        if (p.kblocks1) {
            tma_prefetch_desc(&tmA1);
            tma_prefetch_desc(&tmB1);
        }
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(accum_bar, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================================================================== TMA producer (one lane)
        if (lane == 0) {
            for (int kb = 0; kb < total_kb; ++kb) {
                const int s = kb % STAGES;
                const uint32_t ph = (kb / STAGES) & 1;
                mbar_wait(&empty_bar[s], ph ^ 1);
                mbar_expect_tx(&full_bar[s], Cfg::STAGE_BYTES);
                const bool seg1 = kb >= p.kblocks0;
                const int k_elem = (seg1 ? kb - p.kblocks0 : kb) * Cfg::BLOCK_K;
                tma_load_2d(stage_a(s), seg1 ? &tmA1 : &tmA0, &full_bar[s], k_elem, m0);
                tma_load_2d(stage_b(s), seg1 ? &tmB1 : &tmB0, &full_bar[s], k_elem, n0);
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer (warp-wide loop, one elected lane issues)
        {
            constexpr uint32_t idesc = umma_idesc(kMode == kBF16 ? 1 : 2, Cfg::BLOCK_M, BLOCK_N, 0, 0);
            for (int kb = 0; kb < total_kb; ++kb) {
                const int s = kb % STAGES;
                const uint32_t ph = (kb / STAGES) & 1;
                mbar_wait(&full_bar[s], ph);
                tc_fence_after_sync();
                const uint64_t a_desc = umma_desc_kmajor_sw128(smem_u32(stage_a(s)));
                const uint64_t b_desc = umma_desc_kmajor_sw128(smem_u32(stage_b(s)));
                if (elect_one_sync()) {
#pragma unroll
                    for (int k = 0; k < Cfg::BLOCK_K / Cfg::UMMA_K; ++k) {
                        const uint64_t koff = static_cast<uint64_t>(k * 2);   // 32 bytes per UMMA_K step, >> 4
                        umma<kMode != kBF16>(tmem_base, a_desc + koff, b_desc + koff, idesc, (kb | k) != 0);
                    }
                    umma_commit(&empty_bar[s]);       // frees the stage once these MMAs have read it
                }
                __syncwarp();
            }
            if (elect_one_sync()) umma_commit(accum_bar);               // accumulator final
            __syncwarp();
        }
    } else {
        // ===================================================================== epilogue (warps 2..5)
        mbar_wait(accum_bar, 0);
        tc_fence_after_sync();
        const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
        // accumulator complete => every TMA load was consumed: the operand stages are free to stage the tile
        float* stage = reinterpret_cast<float*>(tiles);
        constexpr int PITCH = BLOCK_N + 4;
        float* my_row = stage + (quarter * 32 + lane) * PITCH;
#pragma unroll 1
        for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
            if (n0 + c0 >= p.N) break;                      // warp-uniform
            uint32_t r[32];
            tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<uint4*>(my_row + c0 + j) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
        }
        __syncwarp();
        epilogue_warp_rows<BLOCK_N>(p, stage, quarter * 32, m0, n0, lane);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

template <int kMode, int BLOCK_N>
static int launch_gemm(const GemmArgs& a, GemmParams p, cudaStream_t st) {
    using Cfg = GemmCfg<kMode, BLOCK_N>;
    const int eb = Cfg::ELEM;
    CUtensorMap tA0, tB0, tA1, tB1;
    int rc;
    if ((rc = make_tmap_sw128(&tA0, a.A0, eb, p.M, a.K0, a.lda0, Cfg::BLOCK_M))) return rc;
    if ((rc = make_tmap_sw128(&tB0, a.B0, eb, p.N, a.K0, a.ldb0, BLOCK_N))) return rc;
    tA1 = tA0;
    tB1 = tB0;
    if (a.K1 > 0) {
        if ((rc = make_tmap_sw128(&tA1, a.A1, eb, p.M, a.K1, a.lda1, Cfg::BLOCK_M))) return rc;
        if ((rc = make_tmap_sw128(&tB1, a.B1, eb, p.N, a.K1, a.ldb1, BLOCK_N))) return rc;
    }
    p.kblocks0 = (a.K0 + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K;
    p.kblocks1 = a.K1 > 0 ? (a.K1 + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K : 0;
    auto kern = gemm_nt_kernel<kMode, BLOCK_N>;
    static bool attr_set = false;   // per template instantiation
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)");
        attr_set = true;
    }
    const int tiles_m = (p.M + Cfg::BLOCK_M - 1) / Cfg::BLOCK_M, tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
    kern<<<tiles_m * tiles_n, kGemmThreads, Cfg::SMEM_BYTES, st>>>(p, tA0, tB0, tA1, tB1);
    GRL_LAUNCH_CHECK("gemm_nt_kernel");
    return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// Persistent bf16 NT kernel for wide outputs (N > 128): one CTA per SM walks 128 x 256 tiles (n fastest, so the CTAs of
// one wave share their A tiles through L2), TWO 256-column accumulators in TMEM so the MMAs of tile t+1 run while eight
// drain warps read tile t out (two warps per TMEM lane quarter, 64 columns per pass) and run the fused epilogue from
// registers through private transposition buffers -- the operand ring never stops for an epilogue.  The one-tile-per-CTA
// kernel above paid a prologue + a serial epilogue per tile: at K = 512 (GraphConv dX at S = 512) that was half the time.
struct NtPCfg {
    static constexpr int BLOCK_M = 128, BLOCK_N = 256, BLOCK_K = 64, UMMA_K = 16;
    static constexpr int A_BYTES = BLOCK_M * 128, B_BYTES = BLOCK_N * 128, STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = 4;
    static constexpr int DRAIN_WARPS = 8;
    static constexpr int EPI_BYTES = DRAIN_WARPS * 32 * X3Cfg::EPI_PITCH * 4;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 + 256;
    static constexpr int TMEM_COLS = 512;
    static constexpr int THREADS = 64 + 32 * DRAIN_WARPS;
};
static_assert(NtPCfg::SMEM_BYTES <= 227 * 1024, "persistent bf16 NT kernel: shared memory budget");

__global__ void __launch_bounds__(NtPCfg::THREADS, 1)
gemm_nt_bf16_persistent_kernel(const GemmParams p, int n_tiles, int tiles_n, const __grid_constant__ CUtensorMap tmA0,
                               const __grid_constant__ CUtensorMap tmB0, const __grid_constant__ CUtensorMap tmA1,
                               const __grid_constant__ CUtensorMap tmB1) {
    using Cfg = NtPCfg;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    float* epi_buf = reinterpret_cast<float*>(tiles + STAGES * Cfg::STAGE_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(tiles + STAGES * Cfg::STAGE_BYTES + Cfg::EPI_BYTES);
    uint64_t* full_bar = bars;                     // [STAGES] TMA bytes landed
    uint64_t* empty_bar = bars + STAGES;           // [STAGES] MMAs reading the stage retired
    uint64_t* acc_full = bars + 2 * STAGES;        // [2] accumulator of a tile complete
    uint64_t* acc_empty = acc_full + 2;            // [2] accumulator read out by all drain warps
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 4);

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);   // provably warp-uniform
    const int lane = threadIdx.x & 31;
    const int total_kb = p.kblocks0 + p.kblocks1;
    auto stage_a = [&](int s) { return tiles + s * Cfg::STAGE_BYTES; };
    auto stage_b = [&](int s) { return tiles + s * Cfg::STAGE_BYTES + Cfg::A_BYTES; };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA0);
        tma_prefetch_desc(&tmB0);
        if (p.kblocks1) {
            tma_prefetch_desc(&tmA1);
            tma_prefetch_desc(&tmB1);
        }
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 32 * Cfg::DRAIN_WARPS);
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================================================================== TMA producer (one lane)
        if (lane == 0) {
            int it = 0;
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                const int m0 = (t / tiles_n) * Cfg::BLOCK_M, n0 = (t % tiles_n) * Cfg::BLOCK_N;
                for (int kb = 0; kb < total_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
                    mbar_expect_tx(&full_bar[s], Cfg::STAGE_BYTES);
                    const bool seg1 = kb >= p.kblocks0;
                    const int k_elem = (seg1 ? kb - p.kblocks0 : kb) * Cfg::BLOCK_K;
                    tma_load_2d(stage_a(s), seg1 ? &tmA1 : &tmA0, &full_bar[s], k_elem, m0);
                    tma_load_2d(stage_b(s), seg1 ? &tmB1 : &tmB0, &full_bar[s], k_elem, n0);
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer (warp-wide loop, one elected lane issues)
        {
            constexpr uint32_t idesc = umma_idesc(1, Cfg::BLOCK_M, Cfg::BLOCK_N, 0, 0);
            int it = 0, ct = 0;
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++ct) {
                const int buf = ct & 1;
                mbar_wait(&acc_empty[buf], ((ct >> 1) & 1) ^ 1);
                tc_fence_after_sync();
                const uint32_t d = tmem_base + buf * Cfg::BLOCK_N;
                for (int kb = 0; kb < total_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    mbar_wait(&full_bar[s], (it / STAGES) & 1);
                    tc_fence_after_sync();
                    const uint64_t a_desc = umma_desc_kmajor_sw128(smem_u32(stage_a(s)));
                    const uint64_t b_desc = umma_desc_kmajor_sw128(smem_u32(stage_b(s)));
                    if (elect_one_sync()) {
#pragma unroll
                        for (int k = 0; k < Cfg::BLOCK_K / Cfg::UMMA_K; ++k) {
                            const uint64_t koff = static_cast<uint64_t>(k * 2);   // 32 bytes per UMMA_K step, >> 4
                            umma<false>(d, a_desc + koff, b_desc + koff, idesc, (kb | k) != 0);
                        }
                        umma_commit(&empty_bar[s]);
                    }
                    __syncwarp();
                }
                if (elect_one_sync()) umma_commit(&acc_full[buf]);
                __syncwarp();
            }
        }
    } else {
        // ===================================================================== drain + epilogue (warps 2..9)
        const int quarter = warp & 3, half = (warp - 2) >> 2;
        const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + half * 128;
        float* my_buf = epi_buf + (warp - 2) * 32 * X3Cfg::EPI_PITCH;
        int ct = 0;
        for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++ct) {
            const int m0 = (t / tiles_n) * Cfg::BLOCK_M, n0 = (t % tiles_n) * Cfg::BLOCK_N;
            const int buf = ct & 1;
            mbar_wait(&acc_full[buf], (ct >> 1) & 1);
            tc_fence_after_sync();
#pragma unroll 1
            for (int pass = 0; pass < 2; ++pass) {
                uint32_t r0[32], r1[32];
                tmem_ld_32x32(lane_base + buf * Cfg::BLOCK_N + pass * 64, r0);
                tmem_ld_32x32(lane_base + buf * Cfg::BLOCK_N + pass * 64 + 32, r1);
                tmem_ld_wait();
                if (pass == 1) {
                    tc_fence_before_sync();
                    mbar_arrive(&acc_empty[buf]);               // the MMAs of tile t+2 may overwrite this accumulator
                }
                float acc[X3Cfg::DRAIN_COLS];
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[q] = __uint_as_float(r0[q]);
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[32 + q] = __uint_as_float(r1[q]);
                if (m0 + quarter * 32 < p.M)
                    epilogue_from_regs(p, acc, my_buf, m0 + quarter * 32, n0 + half * 128 + pass * 64, lane);
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

static bool g_nt_bf16_persistent = true;      // grl_gemm_nt_bf16_set_persistent(0): the one-tile-per-CTA kernel (A/B runs)

static int launch_gemm_bf16_persistent(const GemmArgs& a, GemmParams p, cudaStream_t st) {
    using Cfg = NtPCfg;
    {
        const int rc2 = try_launch_gemm_nt_bf16_2sm(a, p, st);      // CTA pairs (cta_group::2) when switched on and large enough
        if (rc2 != 1) return rc2;
    }
    CUtensorMap tA0, tB0, tA1, tB1;
    int rc;
    if ((rc = make_tmap_sw128(&tA0, a.A0, 2, p.M, a.K0, a.lda0, Cfg::BLOCK_M))) return rc;
    if ((rc = make_tmap_sw128(&tB0, a.B0, 2, p.N, a.K0, a.ldb0, Cfg::BLOCK_N))) return rc;
    tA1 = tA0;
    tB1 = tB0;
    if (a.K1 > 0) {
        if ((rc = make_tmap_sw128(&tA1, a.A1, 2, p.M, a.K1, a.lda1, Cfg::BLOCK_M))) return rc;
        if ((rc = make_tmap_sw128(&tB1, a.B1, 2, p.N, a.K1, a.ldb1, Cfg::BLOCK_N))) return rc;
    }
    p.kblocks0 = (a.K0 + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K;
    p.kblocks1 = a.K1 > 0 ? (a.K1 + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K : 0;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm_nt_bf16_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)");
        attr_set = true;
    }
    const int tiles_m = (p.M + Cfg::BLOCK_M - 1) / Cfg::BLOCK_M, tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    const int n_tiles = tiles_m * tiles_n;
    const int grid = n_tiles < kNumSMs ? n_tiles : kNumSMs;
    gemm_nt_bf16_persistent_kernel<<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(p, n_tiles, tiles_n, tA0, tB0, tA1, tB1);
    GRL_LAUNCH_CHECK("gemm_nt_bf16_persistent_kernel");
    return 0;
}

// bf16 weight gradients: C[Ma, Nb] (fp32) = A[R, Ma]^T * B[R, Nb], both operands MN-major bf16 activations read in place
// by TMA (two boxes of {64 columns, 64 reduction rows} per operand and K-block, SWIZZLE_128B = the tensor core's MN-major
// 16-bit layout: 64-column atoms 8192 B apart, 8 reduction rows per 1024-byte swizzle pattern).  Split-K over R
// (blockIdx.y) into a workspace, deterministic reduction by launch_reduce_splits.  Same warp roles as gemm_nt_kernel.
struct TnBf16Cfg {
    static constexpr int BLOCK_M = 128, BLOCK_N = 128, BLOCK_K = 64, UMMA_K = 16;
    static constexpr int TILE_BYTES = 128 * 128;
    static constexpr int STAGE_BYTES = 2 * TILE_BYTES;
    static constexpr int STAGES = 6;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
    static constexpr int TMEM_COLS = 128;
};

__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tn_bf16_kernel(const GemmParams p, const int kb_per_split, const __grid_constant__ CUtensorMap tmA,
                    const __grid_constant__ CUtensorMap tmB) {
    using Cfg = TnBf16Cfg;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    uint64_t* bars = reinterpret_cast<uint64_t*>(tiles + STAGES * Cfg::STAGE_BYTES);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + STAGES;
    uint64_t* accum_bar = bars + 2 * STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 1);

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);   // provably warp-uniform
    const int lane = threadIdx.x & 31;
    const int tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    const int m0 = (blockIdx.x / tiles_n) * Cfg::BLOCK_M;
    const int n0 = (blockIdx.x % tiles_n) * Cfg::BLOCK_N;
    const int kb_begin = static_cast<int>(blockIdx.y) * kb_per_split;
    const int total_kb = max(min(p.kblocks0, kb_begin + kb_per_split) - kb_begin, 0);

    auto stage_a = [&](int s) { return tiles + s * Cfg::STAGE_BYTES; };
    auto stage_b = [&](int s) { return tiles + s * Cfg::STAGE_BYTES + Cfg::TILE_BYTES; };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(accum_bar, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int kk = 0; kk < total_kb; ++kk) {
                const int s = kk % STAGES;
                mbar_wait(&empty_bar[s], ((kk / STAGES) & 1) ^ 1);
                mbar_expect_tx(&full_bar[s], Cfg::STAGE_BYTES);
                const int r0 = (kb_begin + kk) * Cfg::BLOCK_K;
#pragma unroll
                for (int a = 0; a < 2; ++a) {
                    tma_load_2d(stage_a(s) + a * 8192, &tmA, &full_bar[s], m0 + a * 64, r0);
                    tma_load_2d(stage_b(s) + a * 8192, &tmB, &full_bar[s], n0 + a * 64, r0);
                }
            }
        }
    } else if (warp == 1) {
        {   // warp-wide loop, one elected lane issues (elect_one_sync in tcgen05.cuh)
            constexpr uint32_t idesc = umma_idesc(1, Cfg::BLOCK_M, Cfg::BLOCK_N, 1, 1);
            for (int kk = 0; kk < total_kb; ++kk) {
                const int s = kk % STAGES;
                mbar_wait(&full_bar[s], (kk / STAGES) & 1);
                tc_fence_after_sync();
                const uint64_t a_desc = umma_desc_mnmajor_sw128(smem_u32(stage_a(s)), 8192, 1024, 2);
                const uint64_t b_desc = umma_desc_mnmajor_sw128(smem_u32(stage_b(s)), 8192, 1024, 2);
                if (elect_one_sync()) {
#pragma unroll
                    for (int k = 0; k < Cfg::BLOCK_K / Cfg::UMMA_K; ++k) {
                        const uint64_t koff = static_cast<uint64_t>(k * 128);       // 16 reduction rows = 2048 B, >> 4
                        umma<false>(tmem_base, a_desc + koff, b_desc + koff, idesc, (kk | k) != 0);
                    }
                    umma_commit(&empty_bar[s]);
                }
                __syncwarp();
            }
            if (elect_one_sync()) umma_commit(accum_bar);
            __syncwarp();
        }
    } else {
        mbar_wait(accum_bar, 0);
        tc_fence_after_sync();
        const int quarter = warp & 3;
        float* stage = reinterpret_cast<float*>(tiles);
        constexpr int PITCH = Cfg::BLOCK_N + 4;
        float* my_row = stage + (quarter * 32 + lane) * PITCH;
        const bool any = total_kb > 0;
#pragma unroll 1
        for (int c0 = 0; c0 < Cfg::BLOCK_N; c0 += 32) {
            if (n0 + c0 >= p.N) break;
            uint32_t r[32];
            if (any) {
                tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + c0, r);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) r[j] = 0u;
            }
#pragma unroll
            for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<uint4*>(my_row + c0 + j) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
        }
        __syncwarp();
        GemmParams q = p;
        q.C = static_cast<float*>(p.C) + static_cast<int64_t>(blockIdx.y) * p.M * p.ldc;
        epilogue_warp_rows<Cfg::BLOCK_N>(q, stage, quarter * 32, m0, n0, lane);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

// cs: element stride between the columns of `src` (1 = row-major; a transposed weight view has cs = its pitch, ld_src = 1)
__global__ void split_tf32_kernel(const float* __restrict__ src, int64_t ld_src, float* __restrict__ hi,
                                  float* __restrict__ lo, int64_t ld_dst, int64_t rows, int cols, int64_t cs) {
    const int64_t total = rows * cols;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t r = t / cols;
        const int c = static_cast<int>(t - r * cols);
        const float v = src[r * ld_src + c * cs];
        const float h = __uint_as_float(tf32_rn_bits(__float_as_uint(v)));
        hi[r * ld_dst + c] = h;
        lo[r * ld_dst + c] = __uint_as_float(tf32_rn_bits(__float_as_uint(v - h)));
    }
}

}  // namespace grl

using namespace grl;

extern "C" int grl_split_tf32(const float* src, int64_t ld_src, float* hi, float* lo, int64_t ld_dst, int64_t rows,
                              int cols, void* stream) {
    GRL_CHECK_ARG(src && hi && lo, GRL_E_BADARG, "grl_split_tf32: null pointer");
    GRL_CHECK_ARG(rows >= 0 && cols > 0 && ld_src >= cols && ld_dst >= cols, GRL_E_BADARG, "grl_split_tf32: bad shape");
    if (rows == 0) return 0;
    int64_t blocks = (rows * cols + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    split_tf32_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(src, ld_src, hi, lo, ld_dst,
                                                                                             rows, cols, 1);
    GRL_LAUNCH_CHECK("split_tf32_kernel");
    return 0;
}

extern "C" int grl_split_tf32_strided(const float* src, int64_t stride_row, int64_t stride_col, float* hi, float* lo,
                                      int64_t ld_dst, int64_t rows, int cols, void* stream) {
    GRL_CHECK_ARG(src && hi && lo, GRL_E_BADARG, "grl_split_tf32_strided: null pointer");
    GRL_CHECK_ARG(rows >= 0 && cols > 0 && stride_row > 0 && stride_col > 0 && ld_dst >= cols, GRL_E_BADARG,
                  "grl_split_tf32_strided: bad shape");
    if (rows == 0) return 0;
    int64_t blocks = (rows * cols + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    split_tf32_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(src, stride_row, hi, lo, ld_dst,
                                                                                             rows, cols, stride_col);
    GRL_LAUNCH_CHECK("split_tf32_kernel");
    return 0;
}

extern "C" int grl_gemm_nt(const void* A0, int64_t lda0, const void* B0, const void* B0_lo, int64_t ldb0, int K0,
                           const void* A1, int64_t lda1, const void* B1, const void* B1_lo, int64_t ldb1, int K1,
                           void* C, int64_t ldc, int out_dtype, int M, int N, int mode, const float* bias,
                           uint32_t flags, float p, const uint8_t* mask, uint64_t seed, uint64_t stream_id,
                           void* stream) {
    GRL_CHECK_ARG(A0 && B0 && C, GRL_E_BADARG, "grl_gemm_nt: null pointer");
    GRL_CHECK_ARG(M > 0 && N > 0 && K0 > 0 && K1 >= 0, GRL_E_BADARG, "grl_gemm_nt: bad shape M=%d N=%d K0=%d K1=%d", M, N, K0, K1);
    GRL_CHECK_ARG(mode >= kBF16 && mode <= kTF32BC, GRL_E_BADARG, "grl_gemm_nt: bad mode %d", mode);
    GRL_CHECK_ARG(K1 == 0 || (A1 && B1), GRL_E_BADARG, "grl_gemm_nt: second K segment without operands");
    GRL_CHECK_ARG(mode < kTF32X3 || (B0_lo && (K1 == 0 || B1_lo)), GRL_E_BADARG, "grl_gemm_nt: TF32X3 / TF32_BF16C need the low / packed parts of B");
    GRL_CHECK_ARG(out_dtype == GRL_F32 || out_dtype == GRL_BF16, GRL_E_UNSUP, "grl_gemm_nt: output dtype %d", out_dtype);
    GRL_CHECK_ARG(!(flags & GRL_EPI_BIAS) || bias, GRL_E_BADARG, "grl_gemm_nt: BIAS flag without bias");
    GRL_CHECK_ARG(p >= 0.0f && p < 1.0f, GRL_E_BADARG, "grl_gemm_nt: p outside [0,1)");
    GRL_CHECK_ARG(ldc >= N, GRL_E_BADARG, "grl_gemm_nt: ldc < N");
    const int eb = mode == kBF16 ? 2 : 4;
    const int64_t ld_mult = 16 / eb;
    // TMA: 16-byte aligned base pointers and row pitches
    GRL_CHECK_ARG(lda0 % ld_mult == 0 && ldb0 % ld_mult == 0 && (K1 == 0 || (lda1 % ld_mult == 0 && ldb1 % ld_mult == 0)),
                  GRL_E_ALIGN, "grl_gemm_nt: leading dimensions must be multiples of %lld elements (16 bytes)", (long long)ld_mult);
    GRL_CHECK_ARG(aligned_to(A0, 16) && aligned_to(B0, 16) && (K1 == 0 || (aligned_to(A1, 16) && aligned_to(B1, 16))) &&
                      (mode < kTF32X3 || (aligned_to(B0_lo, 16) && (K1 == 0 || aligned_to(B1_lo, 16)))),
                  GRL_E_ALIGN, "grl_gemm_nt: operand base pointers must be 16-byte aligned");
    GRL_CHECK_ARG(lda0 >= K0 && ldb0 >= K0 && (K1 == 0 || (lda1 >= K1 && ldb1 >= K1)), GRL_E_BADARG, "grl_gemm_nt: ld < K");

    if (p == 0.0f) flags &= ~GRL_EPI_DROPOUT;
    GemmParams gp{};
    gp.M = M; gp.N = N; gp.C = C; gp.ldc = ldc; gp.out_bf16 = out_dtype == GRL_BF16; gp.bias = bias; gp.flags = flags;
    gp.scale = 1.0f / (1.0f - p); gp.thresh = keep_threshold(p); gp.mask = mask; gp.seed = seed; gp.stream_id = stream_id; gp.rng_off = rng_offset_ptr();
    GemmArgs ga{A0, B0, B0_lo, A1, B1, B1_lo, lda0, ldb0, lda1, ldb1, K0, K1};
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool wide = N > 128;
    switch (mode) {
        case kBF16: return wide ? (g_nt_bf16_persistent ? launch_gemm_bf16_persistent(ga, gp, st) : launch_gemm<kBF16, 256>(ga, gp, st)) : (N > 64 ? launch_gemm<kBF16, 128>(ga, gp, st) : launch_gemm<kBF16, 64>(ga, gp, st));
        case kTF32: return wide ? launch_gemm<kTF32, 256>(ga, gp, st) : (N > 64 ? launch_gemm<kTF32, 128>(ga, gp, st) : launch_gemm<kTF32, 64>(ga, gp, st));
        case kTF32BC: return launch_gemm_bc(ga, gp, st, false, 0);
        default:    return launch_gemm_x3(ga, gp, st);
    }
}

extern "C" int grl_gemm_nt_bf16_set_persistent(int on) {
    g_nt_bf16_persistent = on != 0;
    return 0;
}

namespace grl {
int tn_plan_splits(int tiles, int kblocks);                                                        // gemm_x3_tcgen05.cu
int launch_reduce_splits(const float* ws, float* out, int64_t n, int splits, cudaStream_t st);
}

extern "C" int grl_gemm_tn_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, float* C, int64_t ldc, int Ma,
                                int Nb, int64_t R, void* workspace, size_t workspace_bytes, void* stream) {
    using Cfg = TnBf16Cfg;
    GRL_CHECK_ARG(A && B && C && workspace, GRL_E_BADARG, "grl_gemm_tn_bf16: null pointer");
    GRL_CHECK_ARG(Ma > 0 && Nb > 0 && R > 0 && R < (1LL << 31), GRL_E_BADARG, "grl_gemm_tn_bf16: bad shape");
    GRL_CHECK_ARG(lda >= Ma && ldb >= Nb && ldc == Nb, GRL_E_BADARG, "grl_gemm_tn_bf16: bad leading dimension (C must be dense)");
    GRL_CHECK_ARG(lda % 8 == 0 && ldb % 8 == 0 && aligned_to(A, 16) && aligned_to(B, 16), GRL_E_ALIGN,
                  "grl_gemm_tn_bf16: operands need 16-byte aligned base pointers and row pitches");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CUtensorMap tA, tB;
    int rc;
    if ((rc = make_tmap_sw128(&tA, A, 2, R, Ma, lda, Cfg::BLOCK_K))) return rc;      // box {64 columns, 64 rows}
    if ((rc = make_tmap_sw128(&tB, B, 2, R, Nb, ldb, Cfg::BLOCK_K))) return rc;
    GemmParams p{};
    p.M = Ma; p.N = Nb; p.ldc = Nb; p.C = workspace;
    p.kblocks0 = static_cast<int>((R + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K);
    const int tiles = ((Ma + Cfg::BLOCK_M - 1) / Cfg::BLOCK_M) * ((Nb + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N);
    int splits = tn_plan_splits(tiles, p.kblocks0);
    const int kb_per_split = (p.kblocks0 + splits - 1) / splits;
    splits = (p.kblocks0 + kb_per_split - 1) / kb_per_split;
    const size_t need = sizeof(float) * static_cast<size_t>(Ma) * Nb * splits;
    GRL_CHECK_ARG(workspace_bytes >= need, GRL_E_WORKSPACE, "grl_gemm_tn_bf16: workspace %zu < %zu", workspace_bytes, need);
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm_tn_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)");
        attr_set = true;
    }
    gemm_tn_bf16_kernel<<<dim3(tiles, splits), kGemmThreads, Cfg::SMEM_BYTES, st>>>(p, kb_per_split, tA, tB);
    GRL_LAUNCH_CHECK("gemm_tn_bf16_kernel");
    return launch_reduce_splits(static_cast<const float*>(workspace), C, static_cast<int64_t>(Ma) * Nb, splits, st);
}
