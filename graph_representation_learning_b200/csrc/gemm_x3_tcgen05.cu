// fp32-grade GEMM on the tf32 tensor cores ("TF32X3"), the fp32 parity mode of every dense projection:
//
//   C = A_hi*B_hi  +  (A_lo*B_hi + A_hi*B_lo)          a = a_hi + a_lo,  a_hi = a with 13 low mantissa bits cleared
//
// Two measured facts of the sm_100a tensor core shape this kernel (scripts/gemm_accuracy.py, profiles/):
//  (1) each tcgen05.mma accumulation into TMEM truncates (~1.5e-8 relative per step, biased), so a single
//      accumulator over K = 4369 (1638 steps) ends at 3e-5 -- 30x worse than an fp32 FFMA loop;
//  (2) the error is proportional to the number of accumulation steps into ONE accumulator.
// Hence: the dominant A_hi*B_hi chain is accumulated in TMEM only over a short chunk of `promo` K-blocks
// (4*promo steps), then drained by four dedicated warps into fp32 REGISTERS (round-to-nearest FADD) while the
// tensor core fills the other of two TMEM accumulators; the two correction products (2^-11 smaller, so their
// truncation is irrelevant) share a third TMEM accumulator that is read once at the end.
//
// Warp roles (320 threads): 0 = TMA producer, 1 = MMA issuer + TMEM owner, 2..5 = A splitter (hi/lo in smem),
// 6..9 = accumulator drain + fused epilogue.  B (weights) arrives pre-split from grl_split_tf32.
#include "gemm_common.cuh"

namespace grl {
using namespace sm100;

struct X3Cfg {
    static constexpr int BLOCK_M = 128, BLOCK_N = 128, BLOCK_K = 32, UMMA_K = 8;
    static constexpr int TILE_BYTES = 128 * 128;                    // every operand tile: 128 rows x 128 bytes
    static constexpr int STAGE_BYTES = 4 * TILE_BYTES;              // A | B_hi | B_lo | A_lo
    static constexpr int TX_BYTES = 3 * TILE_BYTES;                 // TMA delivers A, B_hi, B_lo
    static constexpr int STAGES = 3;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
    static constexpr int TMEM_COLS = 512;                           // main0 | main1 | small | (unused)
    static constexpr int THREADS = 320;
};

#define GRL_X3_RAW_HI (1u << 8)   // experiment: trust the tensor core to ignore the low 13 bits of A (no in-place hi)

__global__ void __launch_bounds__(X3Cfg::THREADS, 1)
gemm_nt_x3_kernel(const GemmParams p, const int promo, const __grid_constant__ CUtensorMap tmA0,
                  const __grid_constant__ CUtensorMap tmB0, const __grid_constant__ CUtensorMap tmBlo0,
                  const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
                  const __grid_constant__ CUtensorMap tmBlo1) {
    using Cfg = X3Cfg;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    uint64_t* bars = reinterpret_cast<uint64_t*>(tiles + STAGES * Cfg::STAGE_BYTES);
    uint64_t* full_bar = bars;                     // [STAGES] TMA bytes landed
    uint64_t* empty_bar = bars + STAGES;           // [STAGES] MMAs reading the stage retired
    uint64_t* split_bar = bars + 2 * STAGES;       // [STAGES] A_hi / A_lo written by the splitter warps
    uint64_t* acc_full = bars + 3 * STAGES;        // [2] chunk accumulator complete
    uint64_t* acc_empty = bars + 3 * STAGES + 2;   // [2] chunk accumulator drained
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * STAGES + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    const int m0 = (blockIdx.x / tiles_n) * Cfg::BLOCK_M;
    const int n0 = (blockIdx.x % tiles_n) * Cfg::BLOCK_N;
    const int total_kb = p.kblocks0 + p.kblocks1;
    const int n_chunks = (total_kb + promo - 1) / promo;

    auto tile = [&](int s, int which) { return tiles + s * Cfg::STAGE_BYTES + which * Cfg::TILE_BYTES; };
    enum { T_A = 0, T_B = 1, T_BLO = 2, T_ALO = 3 };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA0);
        tma_prefetch_desc(&tmB0);
        tma_prefetch_desc(&tmBlo0);
        if (p.kblocks1) {
            tma_prefetch_desc(&tmA1);
            tma_prefetch_desc(&tmB1);
            tma_prefetch_desc(&tmBlo1);
        }
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
            mbar_init(&split_bar[s], 128);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 128);
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
            for (int kb = 0; kb < total_kb; ++kb) {
                const int s = kb % STAGES;
                const uint32_t ph = (kb / STAGES) & 1;
                mbar_wait(&empty_bar[s], ph ^ 1);
                mbar_expect_tx(&full_bar[s], Cfg::TX_BYTES);
                const bool seg1 = kb >= p.kblocks0;
                const int k_elem = (seg1 ? kb - p.kblocks0 : kb) * Cfg::BLOCK_K;
                tma_load_2d(tile(s, T_A), seg1 ? &tmA1 : &tmA0, &full_bar[s], k_elem, m0);
                tma_load_2d(tile(s, T_B), seg1 ? &tmB1 : &tmB0, &full_bar[s], k_elem, n0);
                tma_load_2d(tile(s, T_BLO), seg1 ? &tmBlo1 : &tmBlo0, &full_bar[s], k_elem, n0);
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer (one lane)
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc(2, Cfg::BLOCK_M, Cfg::BLOCK_N, 0, 0);
            const uint32_t d_small = tmem_base + 256;
            int kb = 0;
            for (int c = 0; c < n_chunks; ++c) {
                const int buf = c & 1;
                mbar_wait(&acc_empty[buf], ((c >> 1) & 1) ^ 1);      // drained (first two uses pass immediately)
                tc_fence_after_sync();
                const uint32_t d_main = tmem_base + buf * 128;
                for (int i = 0; i < promo && kb < total_kb; ++i, ++kb) {
                    const int s = kb % STAGES;
                    const uint32_t ph = (kb / STAGES) & 1;
                    mbar_wait(&split_bar[s], ph);
                    tc_fence_after_sync();
                    const uint64_t a_desc = umma_desc_kmajor_sw128(smem_u32(tile(s, T_A)));
                    const uint64_t b_desc = umma_desc_kmajor_sw128(smem_u32(tile(s, T_B)));
                    const uint64_t alo_desc = umma_desc_kmajor_sw128(smem_u32(tile(s, T_ALO)));
                    const uint64_t blo_desc = umma_desc_kmajor_sw128(smem_u32(tile(s, T_BLO)));
#pragma unroll
                    for (int k = 0; k < Cfg::BLOCK_K / Cfg::UMMA_K; ++k) {
                        const uint64_t koff = static_cast<uint64_t>(k * 2);
                        umma<true>(d_main, a_desc + koff, b_desc + koff, idesc, (i | k) != 0);
                        umma<true>(d_small, alo_desc + koff, b_desc + koff, idesc, (kb | k) != 0);
                        umma<true>(d_small, a_desc + koff, blo_desc + koff, idesc, 1u);
                    }
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(&acc_full[buf]);
            }
        }
    } else if (warp < 6) {
        // ===================================================================== A splitter (warps 2..5)
        // elementwise on the raw tile bytes, so the 128B swizzle written by TMA is irrelevant here
        const int tid = threadIdx.x - 64;
        const bool raw_hi = (p.flags & GRL_X3_RAW_HI) != 0;
        for (int kb = 0; kb < total_kb; ++kb) {
            const int s = kb % STAGES;
            const uint32_t ph = (kb / STAGES) & 1;
            mbar_wait(&full_bar[s], ph);
            uint4* a = reinterpret_cast<uint4*>(tile(s, T_A));
            uint4* alo = reinterpret_cast<uint4*>(tile(s, T_ALO));
#pragma unroll
            for (int i = 0; i < Cfg::TILE_BYTES / 16 / 128; ++i) {
                const int idx = i * 128 + tid;
                const uint4 v = a[idx];
                const uint4 hi = make_uint4(v.x & 0xFFFFE000u, v.y & 0xFFFFE000u, v.z & 0xFFFFE000u, v.w & 0xFFFFE000u);
                uint4 lo;
                lo.x = __float_as_uint(__uint_as_float(v.x) - __uint_as_float(hi.x));
                lo.y = __float_as_uint(__uint_as_float(v.y) - __uint_as_float(hi.y));
                lo.z = __float_as_uint(__uint_as_float(v.z) - __uint_as_float(hi.z));
                lo.w = __float_as_uint(__uint_as_float(v.w) - __uint_as_float(hi.w));
                if (!raw_hi) a[idx] = hi;
                alo[idx] = lo;
            }
            fence_proxy_async_smem();
            mbar_arrive(&split_bar[s]);
        }
    } else {
        // ===================================================================== drain + epilogue (warps 6..9)
        const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
        const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
        float acc[Cfg::BLOCK_N];
#pragma unroll
        for (int j = 0; j < Cfg::BLOCK_N; ++j) acc[j] = 0.0f;
        for (int c = 0; c < n_chunks; ++c) {
            const int buf = c & 1;
            mbar_wait(&acc_full[buf], (c >> 1) & 1);
            tc_fence_after_sync();
#pragma unroll
            for (int j = 0; j < Cfg::BLOCK_N; j += 32) {
                uint32_t r[32];
                tmem_ld_32x32(lane_base + buf * 128 + j, r);
                tmem_ld_wait();
#pragma unroll
                for (int t = 0; t < 32; ++t) acc[j + t] += __uint_as_float(r[t]);
            }
            tc_fence_before_sync();
            mbar_arrive(&acc_empty[buf]);
        }
        // correction terms (complete: the last acc_full commit covers every earlier MMA)
#pragma unroll
        for (int j = 0; j < Cfg::BLOCK_N; j += 32) {
            uint32_t r[32];
            tmem_ld_32x32(lane_base + 256 + j, r);
            tmem_ld_wait();
#pragma unroll
            for (int t = 0; t < 32; ++t) acc[j + t] += __uint_as_float(r[t]);
        }
        const int row = m0 + quarter * 32 + lane;
        if (row < p.M) {
            const bool vec_ok = epilogue_vec_ok(p);
#pragma unroll
            for (int j = 0; j < Cfg::BLOCK_N; j += 4) {
                const int col = n0 + j;
                if (col < p.N) {
                    float v[4] = {acc[j], acc[j + 1], acc[j + 2], acc[j + 3]};
                    epilogue_store4(p, row, col, v, vec_ok);
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

static int g_x3_promo = 2;

int launch_gemm_x3(const GemmArgs& a, GemmParams p, cudaStream_t st) {
    using Cfg = X3Cfg;
    CUtensorMap tA0, tB0, tBlo0, tA1, tB1, tBlo1;
    int rc;
    if ((rc = make_tmap_sw128(&tA0, a.A0, 4, p.M, a.K0, a.lda0, Cfg::BLOCK_M))) return rc;
    if ((rc = make_tmap_sw128(&tB0, a.B0, 4, p.N, a.K0, a.ldb0, Cfg::BLOCK_N))) return rc;
    if ((rc = make_tmap_sw128(&tBlo0, a.B0lo, 4, p.N, a.K0, a.ldb0, Cfg::BLOCK_N))) return rc;
    tA1 = tA0; tB1 = tB0; tBlo1 = tBlo0;
    if (a.K1 > 0) {
        if ((rc = make_tmap_sw128(&tA1, a.A1, 4, p.M, a.K1, a.lda1, Cfg::BLOCK_M))) return rc;
        if ((rc = make_tmap_sw128(&tB1, a.B1, 4, p.N, a.K1, a.ldb1, Cfg::BLOCK_N))) return rc;
        if ((rc = make_tmap_sw128(&tBlo1, a.B1lo, 4, p.N, a.K1, a.ldb1, Cfg::BLOCK_N))) return rc;
    }
    p.kblocks0 = (a.K0 + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K;
    p.kblocks1 = a.K1 > 0 ? (a.K1 + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K : 0;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm_nt_x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)");
        attr_set = true;
    }
    const int tiles_m = (p.M + Cfg::BLOCK_M - 1) / Cfg::BLOCK_M, tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    gemm_nt_x3_kernel<<<tiles_m * tiles_n, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(p, g_x3_promo, tA0, tB0, tBlo0, tA1, tB1,
                                                                              tBlo1);
    GRL_LAUNCH_CHECK("gemm_nt_x3_kernel");
    return 0;
}

}  // namespace grl

// Tuning / experiment knobs of the TF32X3 kernel (host call): promo = K-blocks (of 32) accumulated in TMEM before
// the partial sum is promoted to fp32 registers (>= 1).
extern "C" int grl_gemm_x3_set_promo(int promo) {
    if (promo < 1 || promo > 1 << 20) {
        grl::set_error("grl_gemm_x3_set_promo: promo=%d must be >= 1", promo);
        return GRL_E_BADARG;
    }
    grl::g_x3_promo = promo;
    return 0;
}
