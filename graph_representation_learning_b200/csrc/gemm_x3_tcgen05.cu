// fp32-grade GEMMs on the tf32 tensor cores ("TF32X3"), the fp32 parity mode of every dense contraction:
//
//   C = A_hi*B_hi  +  (A_lo*B_hi + A_hi*B_lo)          a = a_hi + a_lo,  a_hi = a with 13 low mantissa bits cleared
//
// Two measured facts of the sm_100a tensor core shape these kernels (scripts/gemm_accuracy.py, profiles/):
//  (1) each tcgen05.mma accumulation into TMEM truncates (~1.5e-8 relative per step, biased), so a single
//      accumulator over K = 4369 (1638 steps) ends at 3e-5 -- 30x worse than an fp32 FFMA loop;
//  (2) the error is proportional to the number of accumulation steps into ONE accumulator.
// Hence: the dominant A_hi*B_hi chain is accumulated in TMEM only over a short chunk of `promo` K-blocks
// (4*promo steps), then drained by four dedicated warps into fp32 REGISTERS (round-to-nearest FADD) while the
// tensor core fills the other of two TMEM accumulators; the two correction products (2^-11 smaller, so their
// truncation is irrelevant) share a third TMEM accumulator that is read once at the end.
// Measured: 4-7e-7 norm-wise for K = 32 .. 8192 (cuBLAS fp32 SIMT: 1e-7 .. 1.2e-6) at 2.8x its speed.
//
// Two operand arrangements share the kernel body:
//   kTN = false  "NT":  C[M,N] = A[M,K] * B[N,K]^T      forward projections and dX.  Both operands K-major.
//                       B = weights, pre-split (hi/lo) by grl_split_tf32; A split in shared memory.
//   kTN = true   "TN":  C[Ma,Nb] = A[R,Ma]^T * B[R,Nb]  weight gradients: reduction over the node dimension R.
//                       Both operands MN-major activations, both split in shared memory; split-K over R
//                       (blockIdx.y), partial tiles to a workspace that grl_gemm_tn reduces deterministically.
//
// Warp roles (448 threads): 0 = TMA producer, 1 = MMA issuer + TMEM owner, 2..5 = operand splitter,
// 6..13 = accumulator drain + fused epilogue (two warps per TMEM lane quarter, 64 columns each).  The kernel is
// persistent: the TMA / split / MMA pipeline runs into the next tile while the drain warps finish the current one.
#include "gemm_x3_parts.cuh"

namespace grl {
using namespace sm100;

#define GRL_X3_RAW_HI (1u << 8)   // experiment: trust the tensor core to ignore the 13 low bits (no in-place hi write)

// split one 16 KB fp32 tile held in shared memory into hi (in place) and lo; elementwise on raw bytes, so the
// 128B swizzle written by TMA does not matter
__device__ __forceinline__ void split_tile(uint8_t* raw, uint8_t* lo_tile, int tid, bool raw_hi) {
    uint4* a = reinterpret_cast<uint4*>(raw);
    uint4* alo = reinterpret_cast<uint4*>(lo_tile);
#pragma unroll
    for (int i = 0; i < X3Cfg::TILE_BYTES / 16 / 128; ++i) {
        const int idx = i * 128 + tid;
        const uint4 v = a[idx];
        const uint4 hi = make_uint4(tf32_rn_bits(v.x), tf32_rn_bits(v.y), tf32_rn_bits(v.z), tf32_rn_bits(v.w));
        uint4 lo;
        lo.x = tf32_rn_bits(__float_as_uint(__uint_as_float(v.x) - __uint_as_float(hi.x)));
        lo.y = tf32_rn_bits(__float_as_uint(__uint_as_float(v.y) - __uint_as_float(hi.y)));
        lo.z = tf32_rn_bits(__float_as_uint(__uint_as_float(v.z) - __uint_as_float(hi.z)));
        lo.w = tf32_rn_bits(__float_as_uint(__uint_as_float(v.w) - __uint_as_float(hi.w)));
        (void)raw_hi;                  // hi is ROUNDED now: the raw tile can no longer stand in for it
        a[idx] = hi;
        alo[idx] = lo;
    }
}

// Persistent: a CTA walks output tiles blockIdx.x, blockIdx.x + gridDim.x, ...; the TMA / split / MMA pipeline runs
// ahead into the next tile while the drain warps finish the epilogue of the current one.
template <bool kTN>
__global__ void __launch_bounds__(X3Cfg::THREADS, 1)
gemm_x3_kernel(const GemmParams p, const int promo, const int kb_per_split, const int n_tiles,
               const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmB0,
               const __grid_constant__ CUtensorMap tmBlo0, const __grid_constant__ CUtensorMap tmA1,
               const __grid_constant__ CUtensorMap tmB1, const __grid_constant__ CUtensorMap tmBlo1) {
    using Cfg = X3Cfg;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    float* epi_buf = reinterpret_cast<float*>(tiles + STAGES * Cfg::STAGE_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(tiles + STAGES * Cfg::STAGE_BYTES + Cfg::EPI_BYTES);
    uint64_t* full_bar = bars;                     // [STAGES] TMA bytes landed
    uint64_t* empty_bar = bars + STAGES;           // [STAGES] MMAs reading the stage retired
    uint64_t* split_bar = bars + 2 * STAGES;       // [STAGES] hi / lo tiles written by the splitter warps
    uint64_t* acc_full = bars + 3 * STAGES;        // [2] chunk accumulator complete
    uint64_t* acc_empty = bars + 3 * STAGES + 2;   // [2] chunk accumulator drained
    uint64_t* small_empty = bars + 3 * STAGES + 4; // [2] correction accumulator (one per tile parity) read out
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * STAGES + 6);

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);   // provably warp-uniform
    const int lane = threadIdx.x & 31;
    const int tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    // K-block range of this CTA: everything for NT; one slice of the reduction for TN (split-K over blockIdx.y)
    const int all_kb = p.kblocks0 + p.kblocks1;
    const int kb_begin = kTN ? static_cast<int>(blockIdx.y) * kb_per_split : 0;
    const int kb_end = kTN ? min(all_kb, kb_begin + kb_per_split) : all_kb;
    const int total_kb = max(kb_end - kb_begin, 0);
    const int n_chunks = (total_kb + promo - 1) / promo;

    auto tile = [&](int s, int which) { return tiles + s * Cfg::STAGE_BYTES + which * Cfg::TILE_BYTES; };
    enum { T_A = 0, T_B = 1, T_BLO = 2, T_ALO = 3 };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA0);
        tma_prefetch_desc(&tmB0);
        if (!kTN) tma_prefetch_desc(&tmBlo0);
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
            mbar_init(&acc_empty[b], 32 * Cfg::DRAIN_WARPS);
        }
        mbar_init(&small_empty[0], 32 * Cfg::DRAIN_WARPS);
        mbar_init(&small_empty[1], 32 * Cfg::DRAIN_WARPS);
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
            int it = 0;                                            // K-blocks issued so far, over all tiles of this CTA
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
                const int m0 = (t / tiles_n) * Cfg::BLOCK_M, n0 = (t % tiles_n) * Cfg::BLOCK_N;
                for (int kk = 0; kk < total_kb; ++kk, ++it) {
                    const int kb = kb_begin + kk;
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    if constexpr (kTN) {
                        // operands are [R, Ma] / [R, Nb] row-major: 4 boxes of {32 columns (128 B), 32 reduction rows}
                        mbar_expect_tx(&full_bar[s], 2 * Cfg::TILE_BYTES);
                        const int r0 = kb * Cfg::BLOCK_K;
#pragma unroll
                        for (int a = 0; a < 4; ++a) {
                            tma_load_2d(tile(s, T_A) + a * 4096, &tmA0, &full_bar[s], m0 + a * 32, r0);
                            tma_load_2d(tile(s, T_B) + a * 4096, &tmB0, &full_bar[s], n0 + a * 32, r0);
                        }
                    } else {
                        mbar_expect_tx(&full_bar[s], 3 * Cfg::TILE_BYTES);
                        const bool seg1 = kb >= p.kblocks0;
                        const int k_elem = (seg1 ? kb - p.kblocks0 : kb) * Cfg::BLOCK_K;
                        tma_load_2d(tile(s, T_A), seg1 ? &tmA1 : &tmA0, &full_bar[s], k_elem, m0);
                        tma_load_2d(tile(s, T_B), seg1 ? &tmB1 : &tmB0, &full_bar[s], k_elem, n0);
                        tma_load_2d(tile(s, T_BLO), seg1 ? &tmBlo1 : &tmBlo0, &full_bar[s], k_elem, n0);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer (warp-wide loop, one elected lane issues:
        // see elect_one_sync in tcgen05.cuh)
        {
            constexpr uint32_t idesc = umma_idesc(2, Cfg::BLOCK_M, Cfg::BLOCK_N, kTN ? 1 : 0, kTN ? 1 : 0);
            // two correction accumulators (TMEM columns 256.. and 384..), alternating per tile: the MMAs of tile i+1 start
            // while the drain warps still read tile i's corrections (short-K products spent a third of every tile there)
            int it = 0, cg = 0, tl = 0;                            // K-blocks, chunks, tiles issued by this CTA
            for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++tl) {
                const uint32_t d_small = tmem_base + 256 + (tl & 1) * 128;
                mbar_wait(&small_empty[tl & 1], ((tl >> 1) & 1) ^ 1);   // this accumulator's previous tile was read out
                tc_fence_after_sync();
                int kk = 0;
                for (int c = 0; c < n_chunks; ++c, ++cg) {
                    const int buf = cg & 1;
                    mbar_wait(&acc_empty[buf], ((cg >> 1) & 1) ^ 1);   // drained (first two uses pass immediately)
                    tc_fence_after_sync();
                    const uint32_t d_main = tmem_base + buf * 128;
                    for (int i = 0; i < promo && kk < total_kb; ++i, ++kk, ++it) {
                        const int s = it % STAGES;
                        const uint32_t ph = (it / STAGES) & 1;
                        mbar_wait(&split_bar[s], ph);
                        tc_fence_after_sync();
                        uint64_t a_desc, b_desc, alo_desc, blo_desc;
                        if constexpr (kTN) {
                            // MN-major tf32: 128-byte atoms of 32 columns 4096 B apart (LBO); swizzle pattern of 4
                            // reduction rows = 512 B (SBO); layout SWIZZLE_128B_BASE32B as written by TMA's ATOM_32B
                            constexpr uint32_t lbo = 4096, sbo = 512;
                            a_desc = umma_desc_mnmajor_sw128(smem_u32(tile(s, T_A)), lbo, sbo, 1);
                            b_desc = umma_desc_mnmajor_sw128(smem_u32(tile(s, T_B)), lbo, sbo, 1);
                            alo_desc = umma_desc_mnmajor_sw128(smem_u32(tile(s, T_ALO)), lbo, sbo, 1);
                            blo_desc = umma_desc_mnmajor_sw128(smem_u32(tile(s, T_BLO)), lbo, sbo, 1);
                        } else {
                            a_desc = umma_desc_kmajor_sw128(smem_u32(tile(s, T_A)));
                            b_desc = umma_desc_kmajor_sw128(smem_u32(tile(s, T_B)));
                            alo_desc = umma_desc_kmajor_sw128(smem_u32(tile(s, T_ALO)));
                            blo_desc = umma_desc_kmajor_sw128(smem_u32(tile(s, T_BLO)));
                        }
                        if (elect_one_sync()) {
#pragma unroll
                            for (int k = 0; k < Cfg::BLOCK_K / Cfg::UMMA_K; ++k) {
                                // one UMMA_K step: 32 bytes along a K-major row, or 8 rows (1024 B) of an MN-major tile
                                const uint64_t koff = static_cast<uint64_t>(kTN ? k * 64 : k * 2);
                                umma<true>(d_main, a_desc + koff, b_desc + koff, idesc, (i | k) != 0);
                                umma<true>(d_small, alo_desc + koff, b_desc + koff, idesc, (kk | k) != 0);
                                umma<true>(d_small, a_desc + koff, blo_desc + koff, idesc, 1u);
                            }
                            umma_commit(&empty_bar[s]);
                        }
                        __syncwarp();
                    }
                    if (elect_one_sync()) umma_commit(&acc_full[buf]);
                    __syncwarp();
                }
            }
        }
    } else if (warp < 6) {
        // ===================================================================== operand splitter (warps 2..5)
        const int tid = threadIdx.x - 64;
        const bool raw_hi = (p.flags & GRL_X3_RAW_HI) != 0;
        int it = 0;
        for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
            for (int kk = 0; kk < total_kb; ++kk, ++it) {
                const int s = it % STAGES;
                const uint32_t ph = (it / STAGES) & 1;
                mbar_wait(&full_bar[s], ph);
                split_tile(tile(s, T_A), tile(s, T_ALO), tid, raw_hi);
                if constexpr (kTN) split_tile(tile(s, T_B), tile(s, T_BLO), tid, raw_hi);
                fence_proxy_async_smem();
                mbar_arrive(&split_bar[s]);
            }
        }
    } else {
        // ===================================================================== drain + epilogue (warps 6..13)
        // TMEM lane quarter = warp % 4 (hardware rule); the two warps of a quarter take 64 accumulator columns each
        const int quarter = warp & 3, half = (warp - 6) >> 2;
        const int colbase = half * Cfg::DRAIN_COLS;
        const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + colbase;
        float* my_buf = epi_buf + (warp - 6) * 32 * Cfg::EPI_PITCH;
        int cg = 0, tl = 0;
        for (int t = blockIdx.x; t < n_tiles; t += gridDim.x, ++tl) {
            const int m0 = (t / tiles_n) * Cfg::BLOCK_M, n0 = (t % tiles_n) * Cfg::BLOCK_N;
            const uint32_t small_col = 256 + (tl & 1) * 128;
            float acc[Cfg::DRAIN_COLS];
#pragma unroll
            for (int j = 0; j < Cfg::DRAIN_COLS; ++j) acc[j] = 0.0f;
            for (int c = 0; c < n_chunks; ++c, ++cg) {
                const int buf = cg & 1;
                mbar_wait(&acc_full[buf], (cg >> 1) & 1);
                tc_fence_after_sync();
                uint32_t r0[32], r1[32];
                tmem_ld_32x32(lane_base + buf * 128, r0);
                tmem_ld_32x32(lane_base + buf * 128 + 32, r1);
                tmem_ld_wait();
                tc_fence_before_sync();
                mbar_arrive(&acc_empty[buf]);                // the data is in registers: release the accumulator first
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[q] += __uint_as_float(r0[q]);
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[32 + q] += __uint_as_float(r1[q]);
            }
            if (n_chunks > 0) {
                // correction terms (complete: the last acc_full commit covers every earlier MMA of the tile)
                uint32_t r0[32], r1[32];
                tmem_ld_32x32(lane_base + small_col, r0);
                tmem_ld_32x32(lane_base + small_col + 32, r1);
                tmem_ld_wait();
                tc_fence_before_sync();
                mbar_arrive(&small_empty[tl & 1]);           // this correction accumulator may be reused two tiles on
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[q] += __uint_as_float(r0[q]);
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[32 + q] += __uint_as_float(r1[q]);
            } else {
                mbar_arrive(&small_empty[tl & 1]);
            }
            GemmParams q = p;
            if constexpr (kTN)   // partial tile of split blockIdx.y -> its slab of the workspace
                q.C = static_cast<float*>(p.C) + static_cast<int64_t>(blockIdx.y) * p.M * p.ldc;
            if (m0 + quarter * 32 < p.M) epilogue_from_regs(q, acc, my_buf, m0 + quarter * 32, n0 + colbase, lane);
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

static int g_x3_promo = 2;
static int g_x3_persistent = 1;   // 0: one CTA per tile (A/B experiments via grl_gemm_x3_set_promo(-1 / -2))
static int g_x3_tmem_operand = 1; // 1 (-3, default): NT products run on the TMEM-operand kernel of gemm_bc_tcgen05.cu, same
                                  // arithmetic, half the shared-memory traffic: 5-9 % faster once the MMA issue loop stopped
                                  // being the bound (scripts/gemm_x3_ab.py); 0 (-4): the shared-memory-operand kernel here

int gemm_x3_current_promo() { return g_x3_promo; }      // graphconv_fused.cu promotes its main product like the forward GEMMs

template <bool kTN>
static int ensure_smem_attr() {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm_x3_kernel<kTN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             X3Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)");
        attr_set = true;
    }
    return 0;
}

int launch_gemm_x3(const GemmArgs& a, GemmParams p, cudaStream_t st) {
    using Cfg = X3Cfg;
    if (g_x3_tmem_operand) return launch_gemm_bc(a, p, st, true, g_x3_promo);
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
    if ((rc = ensure_smem_attr<false>())) return rc;
    const int tiles_m = (p.M + Cfg::BLOCK_M - 1) / Cfg::BLOCK_M, tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    const int n_tiles = tiles_m * tiles_n;
    const int grid = g_x3_persistent ? (n_tiles < kNumSMs ? n_tiles : kNumSMs) : n_tiles;   // one CTA per SM, tiles strided
    gemm_x3_kernel<false><<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(p, g_x3_promo, 0, n_tiles, tA0, tB0, tBlo0, tA1, tB1,
                                                                      tBlo1);
    GRL_LAUNCH_CHECK("gemm_x3_kernel<NT>");
    return 0;
}

// out[i] = sum_s ws[s][i]   (deterministic split-K reduction; float4 lanes)
__global__ void reduce_splits_kernel(const float* __restrict__ ws, float* __restrict__ out, int64_t n, int splits) {
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        float s = 0.0f;
        for (int k = 0; k < splits; ++k) s += ws[static_cast<int64_t>(k) * n + i];
        out[i] = s;
    }
}

}  // namespace grl

using namespace grl;

// Tuning / experiment knob of the TF32X3 kernels (host call): promo = K-blocks (of 32) accumulated in TMEM before
// the partial sum is promoted to fp32 registers (>= 1).
extern "C" int grl_gemm_x3_set_promo(int promo) {
    if (promo == -1 || promo == -2) {             // experiment switch: -1 = one CTA per tile, -2 = persistent (default)
        g_x3_persistent = promo == -2;
        return 0;
    }
    if (promo == -3 || promo == -4) {             // -3 = activation operand through TMEM (default), -4 = shared memory
        g_x3_tmem_operand = promo == -3;
        return 0;
    }
    if (promo < 1 || promo > 1 << 20) {
        set_error("grl_gemm_x3_set_promo: promo=%d must be >= 1", promo);
        return GRL_E_BADARG;
    }
    g_x3_promo = promo;
    return 0;
}

namespace grl {
// Split-K plan of the TN kernel: all tiles * splits CTAs run in waves of kNumSMs, a CTA costs its K-blocks plus a fixed
// prologue + drain + epilogue (~12 K-blocks of time).  One wave with fewer CTAs than SMs can lose to two fuller waves
// (gcn3 dW: 56 tiles -> 2 splits = 112 CTAs = 2000 K-blocks each, against 5 splits = 280 CTAs = 2 waves of 800).
int tn_plan_splits(int tiles, int kblocks) {
    constexpr int kOverheadKb = 12, kMaxSplits = 16;
    int cap = kNumSMs / tiles;
    if (cap < kMaxSplits) cap = kMaxSplits;
    if (cap > kblocks) cap = kblocks;
    if (cap < 1) cap = 1;
    int64_t best_cost = -1;
    int best = 1;
    for (int sp = 1; sp <= cap; ++sp) {
        const int64_t waves = (static_cast<int64_t>(tiles) * sp + kNumSMs - 1) / kNumSMs;
        const int64_t cost = waves * ((kblocks + sp - 1) / sp + kOverheadKb);
        if (best_cost < 0 || cost * 100 < best_cost * 98) {      // a larger split count must win by 2 %
            best_cost = cost;
            best = sp;
        }
    }
    return best;
}

int launch_reduce_splits(const float* ws, float* out, int64_t n, int splits, cudaStream_t st) {
    int blocks = static_cast<int>((n + 255) / 256);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    reduce_splits_kernel<<<blocks, 256, 0, st>>>(ws, out, n, splits);
    GRL_LAUNCH_CHECK("reduce_splits_kernel");
    return 0;
}
}  // namespace grl

extern "C" size_t grl_gemm_tn_workspace_bytes(int Ma, int Nb) {
    // one [Ma, Nb] fp32 slab per split; splits <= max(16, kNumSMs / tiles)
    const int tiles = ((Ma + 127) / 128) * ((Nb + 127) / 128);
    int cap = kNumSMs / (tiles > 0 ? tiles : 1);
    if (cap < 16) cap = 16;
    return sizeof(float) * static_cast<size_t>(Ma) * Nb * cap;
}

extern "C" int grl_gemm_tn(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc, int Ma,
                           int Nb, int64_t R, void* workspace, size_t workspace_bytes, void* stream) {
    using Cfg = X3Cfg;
    GRL_CHECK_ARG(A && B && C && workspace, GRL_E_BADARG, "grl_gemm_tn: null pointer");
    GRL_CHECK_ARG(Ma > 0 && Nb > 0 && R > 0 && R < (1LL << 31), GRL_E_BADARG, "grl_gemm_tn: bad shape");
    GRL_CHECK_ARG(lda >= Ma && ldb >= Nb && ldc == Nb, GRL_E_BADARG, "grl_gemm_tn: bad leading dimension (C must be dense)");
    GRL_CHECK_ARG(lda % 4 == 0 && ldb % 4 == 0 && aligned_to(A, 16) && aligned_to(B, 16), GRL_E_ALIGN,
                  "grl_gemm_tn: operands need 16-byte aligned base pointers and row pitches");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CUtensorMap tA, tB;
    int rc;
    // [R, cols] row-major; box = {32 columns, 32 rows}
    if ((rc = make_tmap_sw128(&tA, A, 4, R, Ma, lda, Cfg::BLOCK_K, true))) return rc;
    if ((rc = make_tmap_sw128(&tB, B, 4, R, Nb, ldb, Cfg::BLOCK_K, true))) return rc;
    GemmParams p{};
    p.M = Ma; p.N = Nb; p.ldc = Nb; p.C = workspace;
    p.kblocks0 = static_cast<int>((R + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K);
    p.kblocks1 = 0;
    const int tiles = ((Ma + Cfg::BLOCK_M - 1) / Cfg::BLOCK_M) * ((Nb + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N);
    int splits = tn_plan_splits(tiles, p.kblocks0);
    const int kb_per_split = (p.kblocks0 + splits - 1) / splits;
    splits = (p.kblocks0 + kb_per_split - 1) / kb_per_split;          // no empty split
    const size_t need = sizeof(float) * static_cast<size_t>(Ma) * Nb * splits;
    GRL_CHECK_ARG(workspace_bytes >= need, GRL_E_WORKSPACE, "grl_gemm_tn: workspace %zu < %zu", workspace_bytes, need);
    if ((rc = ensure_smem_attr<true>())) return rc;
    gemm_x3_kernel<true><<<dim3(tiles, splits), Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(p, g_x3_promo, kb_per_split, tiles, tA,
                                                                                    tB, tB, tA, tB, tB);
    GRL_LAUNCH_CHECK("gemm_x3_kernel<TN>");
    return launch_reduce_splits(static_cast<const float*>(workspace), C, static_cast<int64_t>(Ma) * Nb, splits, st);
}
