// fp32-grade GEMMs, second generation ("TF32_BF16C"): one tf32 product + ONE bf16 correction product, with the
// activation operand fed to the tensor core from TMEM.
//
//   C = A_hi*B_hi (kind::tf32)  +  [bf16(A_lo) | bf16(A_hi)] * [bf16(B_hi) | bf16(B_lo)]^T (kind::f16 over a doubled K)
//   x_hi = x rounded to nearest tf32, x_lo = x - x_hi (exact, |x_lo| <= 2^-11 |x|)
//
// Why (measured on the three-product kernel, gemm_x3_tcgen05.cu, profiles/r01_s31_*): that kernel is bound by the
// 128 B/clk shared-memory port, not by the tensor pipe.  Per 128x128x32 K-block it moves 192 KB through shared memory
// (TMA writes 48 KB, the operand splitter reads 16 KB and writes 32 KB, twelve SS-mode MMAs read 8 KB each) = 1536
// cycles against the 1550 measured, while the MMAs alone need 768.  Replacing the two tf32 correction products by one
// bf16 product (8 instead of 12 MMA steps, 160 KB) gained only 5-12 %.  Here the activation tile never returns to
// shared memory: the splitter warps read the raw fp32 tile once (16 KB), split it in registers and write a_hi and the
// packed bf16 [a_lo | a_hi] with tcgen05.st into a two-deep TMEM ring; the MMAs take A from TMEM ("TS" form) and read
// only the weight tiles from shared memory: 48 + 16 + 8*4 = 96 KB per K-block.
// Accuracy: rounding a_lo / b_hi to bf16 costs 2^-9 relative on terms that are 2^-11 of the product: 6-7e-7 norm-wise
// measured for K = 32 .. 4369 (three tf32 products: 4-5e-7; cuBLAS fp32 SIMT: 1e-7 .. 1.2e-6).
//
//   kTN = false "NT":  C[M,N] = A[M,K] * B[N,K]^T.  B = weights, prepared once per step by grl_split_tf32_mix.
//   kTN = true  "TN":  C[Ma,Nb] = A[R,Ma]^T * B[R,Nb] (weight gradients).  A = the WIDE activation operand, transposed
//                      by the splitter's shared-memory reads (lane m takes column m); B = the narrow operand (dZ), split
//                      ONCE in HBM by grl_split_act_mn instead of once per output-tile row.  Split-K over R as before.
//
// The main product is still promoted to fp32 registers every `promo` K-blocks (TMEM accumulation truncates, see
// gemm_x3_tcgen05.cu); warp roles and the persistent tile loop are those of the three-product kernel.
#include "gemm_x3_parts.cuh"

namespace grl {

struct BcCfg {
    static constexpr int BLOCK_M = 128, BLOCK_N = 128, BLOCK_K = 32;
    static constexpr int TILE_BYTES = 128 * 128;
    static constexpr int STAGE_BYTES = 3 * TILE_BYTES;              // A raw | B_hi | B packed
    static constexpr int STAGES = 4;
    static constexpr int TA_STAGES = 2;                             // TMEM ring of split activation tiles
    static constexpr int TA_COL0 = 384;                             // main0 [0,128) | main1 [128,256) | small [256,384)
    static constexpr int TA_COLS = 64;                              // a_hi: 32 columns, packed bf16: 32 columns
    static constexpr int EPI_BYTES = X3Cfg::EPI_BYTES;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 + 256;
    static constexpr int TMEM_COLS = 512;
    static constexpr int THREADS = X3Cfg::THREADS;
};
static_assert(BcCfg::SMEM_BYTES <= 227 * 1024, "shared memory budget");

__device__ __forceinline__ uint32_t pack_bf16x2(float first, float second) {   // `first` at the lower address
    const __nv_bfloat162 v = __floats2bfloat162_rn(first, second);
    return *reinterpret_cast<const uint32_t*>(&v);
}

// kX3 (NT only): the three-tf32-product arithmetic of gemm_x3_tcgen05.cu (a_hi = a truncated to tf32, corrections
// a_lo*b_hi + a_hi*b_lo in tf32 against the fp32 low parts of grl_split_tf32) on this kernel's TMEM operand path.
// kMC (NT only): launched as thread-block clusters of 2 CTAs that work on two adjacent row tiles of the SAME column tile;
// each CTA fetches half of every weight tile (64 of the 128 rows of B_hi and of the packed B) and TMA-multicasts it into
// both CTAs, so a K-block costs 32 KB of L2 -> SM traffic per CTA instead of 48 KB.  A stage is reusable when BOTH
// MMA issuers have retired it: empty barriers count 2 and every commit is multicast to the pair.
template <bool kTN, bool kX3, bool kMC = false>
__global__ void __launch_bounds__(BcCfg::THREADS, 1)
gemm_bc_kernel(const GemmParams p, const int promo, const int kb_per_split, const int n_tiles, const int transpose_out,
               const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmB0,
               const __grid_constant__ CUtensorMap tmBp0, const __grid_constant__ CUtensorMap tmA1,
               const __grid_constant__ CUtensorMap tmB1, const __grid_constant__ CUtensorMap tmBp1) {
    static_assert(!(kTN && kX3), "the three-product arithmetic on the TMEM path exists for NT only");
    static_assert(!(kMC && (kTN || kX3)), "weight-tile multicast is built for the NT second-generation product");
    using Cfg = BcCfg;
    constexpr int STAGES = Cfg::STAGES, TA = Cfg::TA_STAGES, DRAIN_WARPS = X3Cfg::DRAIN_WARPS;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    float* epi_buf = reinterpret_cast<float*>(tiles + STAGES * Cfg::STAGE_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(tiles + STAGES * Cfg::STAGE_BYTES + Cfg::EPI_BYTES);
    uint64_t* full_bar = bars;                          // [STAGES] TMA bytes landed
    uint64_t* empty_bar = bars + STAGES;                // [STAGES] MMAs reading the stage retired
    uint64_t* split_bar = bars + 2 * STAGES;            // [TA] split activation tile written to TMEM
    uint64_t* ta_empty = bars + 2 * STAGES + TA;        // [TA] MMAs reading the TMEM tile retired
    uint64_t* acc_full = bars + 2 * STAGES + 2 * TA;    // [2] chunk accumulator complete
    uint64_t* acc_empty = acc_full + 2;                 // [2] chunk accumulator drained
    uint64_t* small_empty = acc_full + 4;               // correction accumulator of the previous tile read out
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 5);

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);   // provably warp-uniform
    const int lane = threadIdx.x & 31;
    const int tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    const int all_kb = p.kblocks0 + p.kblocks1;
    const int kb_begin = kTN ? static_cast<int>(blockIdx.y) * kb_per_split : 0;
    const int kb_end = kTN ? min(all_kb, kb_begin + kb_per_split) : all_kb;
    const int total_kb = max(kb_end - kb_begin, 0);
    const int n_chunks = (total_kb + promo - 1) / promo;
    // tile walk: one tile per step, or (kMC) one PAIR of row tiles per step, this CTA taking row tile 2*pair + rank
    const int cta_rank = kMC ? static_cast<int>(cluster_ctarank()) : 0;
    const int t_first = kMC ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
    const int t_step = kMC ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
    auto tile_m0 = [&](int t) { return (kMC ? 2 * (t / tiles_n) + cta_rank : t / tiles_n) * Cfg::BLOCK_M; };

    auto tile = [&](int s, int which) { return tiles + s * Cfg::STAGE_BYTES + which * Cfg::TILE_BYTES; };
    enum { T_A = 0, T_B = 1, T_BP = 2 };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA0);
        tma_prefetch_desc(&tmB0);
        tma_prefetch_desc(&tmBp0);
        if (p.kblocks1) {
            tma_prefetch_desc(&tmA1);
            tma_prefetch_desc(&tmB1);
            tma_prefetch_desc(&tmBp1);
        }
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kMC ? 2 : 1);
        }
        for (int j = 0; j < TA; ++j) {
            mbar_init(&split_bar[j], 128);
            mbar_init(&ta_empty[j], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 32 * DRAIN_WARPS);
        }
        mbar_init(small_empty, 32 * DRAIN_WARPS);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    tc_fence_before_sync();
    if constexpr (kMC) cluster_sync_all();      // the peer's barriers are initialised before anything is multicast
    else __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================================================================== TMA producer (one lane)
        if (lane == 0) {
            int it = 0;
            for (int t = t_first; t < n_tiles; t += t_step) {
                const int m0 = tile_m0(t), n0 = (t % tiles_n) * Cfg::BLOCK_N;
                for (int kk = 0; kk < total_kb; ++kk, ++it) {
                    const int kb = kb_begin + kk;
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    mbar_expect_tx(&full_bar[s], 3 * Cfg::TILE_BYTES);
                    if constexpr (kTN) {
                        const int r0 = kb * Cfg::BLOCK_K;
                        // A: one plain box {128 columns, 32 reduction rows} (read by the splitter only)
                        tma_load_2d(tile(s, T_A), &tmA0, &full_bar[s], m0, r0);
#pragma unroll
                        for (int a = 0; a < 4; ++a)   // B_hi: MN-major tf32, 4 boxes of {32 columns, 32 rows}
                            tma_load_2d(tile(s, T_B) + a * 4096, &tmB0, &full_bar[s], n0 + a * 32, r0);
#pragma unroll
                        for (int a = 0; a < 2; ++a)   // packed bf16: MN-major, 2 boxes of {64 columns, 64 rows}
                            tma_load_2d(tile(s, T_BP) + a * 8192, &tmBp0, &full_bar[s], n0 + a * 64, 2 * r0);
                    } else {
                        const bool seg1 = kb >= p.kblocks0;
                        const int k_elem = (seg1 ? kb - p.kblocks0 : kb) * Cfg::BLOCK_K;
                        tma_load_2d(tile(s, T_A), seg1 ? &tmA1 : &tmA0, &full_bar[s], k_elem, m0);
                        if constexpr (kMC) {
                            // my half (64 rows = 8 KB) of both weight tiles, into both CTAs of the pair
                            const int nh = n0 + cta_rank * 64, off = cta_rank * 8192;
                            tma_load_2d_multicast(tile(s, T_B) + off, seg1 ? &tmB1 : &tmB0, &full_bar[s], k_elem, nh, 3);
                            tma_load_2d_multicast(tile(s, T_BP) + off, seg1 ? &tmBp1 : &tmBp0, &full_bar[s], 2 * k_elem, nh, 3);
                        } else {
                            tma_load_2d(tile(s, T_B), seg1 ? &tmB1 : &tmB0, &full_bar[s], k_elem, n0);
                            tma_load_2d(tile(s, T_BP), seg1 ? &tmBp1 : &tmBp0, &full_bar[s], kX3 ? k_elem : 2 * k_elem, n0);
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer (warp-wide loop, one elected lane issues)
        {
            constexpr uint32_t idesc = umma_idesc(2, Cfg::BLOCK_M, Cfg::BLOCK_N, 0, kTN ? 1 : 0);     // tf32, A K-major (TMEM)
            constexpr uint32_t idesc_c = umma_idesc(1, Cfg::BLOCK_M, Cfg::BLOCK_N, 0, kTN ? 1 : 0);   // bf16
            const uint32_t d_small = tmem_base + 256;
            int it = 0, cg = 0, tl = 0;
            for (int t = t_first; t < n_tiles; t += t_step, ++tl) {
                mbar_wait(small_empty, (tl & 1) ^ 1);
                tc_fence_after_sync();
                int kk = 0;
                for (int c = 0; c < n_chunks; ++c, ++cg) {
                    const int buf = cg & 1;
                    mbar_wait(&acc_empty[buf], ((cg >> 1) & 1) ^ 1);
                    tc_fence_after_sync();
                    const uint32_t d_main = tmem_base + buf * 128;
                    for (int i = 0; i < promo && kk < total_kb; ++i, ++kk, ++it) {
                        const int s = it % STAGES, j = it % TA;
                        mbar_wait(&full_bar[s], (it / STAGES) & 1);       // weight tiles landed
                        mbar_wait(&split_bar[j], (it / TA) & 1);          // activation tile split into TMEM
                        tc_fence_after_sync();
                        const uint32_t ta_hi = tmem_base + Cfg::TA_COL0 + j * Cfg::TA_COLS, ta_pk = ta_hi + 32;
                        uint64_t b_desc, bp_desc;
                        if constexpr (kTN) {
                            b_desc = umma_desc_mnmajor_sw128(smem_u32(tile(s, T_B)), 4096, 512, 1);
                            bp_desc = umma_desc_mnmajor_sw128(smem_u32(tile(s, T_BP)), 8192, 1024, 2);
                        } else {
                            b_desc = umma_desc_kmajor_sw128(smem_u32(tile(s, T_B)));
                            bp_desc = umma_desc_kmajor_sw128(smem_u32(tile(s, T_BP)));
                        }
                        if (elect_one_sync()) {
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                // 8 tf32 / 16 packed bf16 per step = 8 TMEM columns of A; B: 32 bytes of a K-major row, or
                                // 8 rows (tf32, 1024 B) / 16 rows (bf16, 2048 B) of an MN-major tile
                                const uint64_t koff = static_cast<uint64_t>(kTN ? k * 64 : k * 2);
                                const uint64_t koff_c = static_cast<uint64_t>(kTN ? k * 128 : k * 2);
                                umma_ts<true>(d_main, ta_hi + 8 * k, b_desc + koff, idesc, (i | k) != 0);
                                if constexpr (kX3) {
                                    umma_ts<true>(d_small, ta_pk + 8 * k, b_desc + koff, idesc, (kk | k) != 0);    // a_lo * b_hi
                                    umma_ts<true>(d_small, ta_hi + 8 * k, bp_desc + koff, idesc, 1u);              // a_hi * b_lo
                                } else {
                                    umma_ts<false>(d_small, ta_pk + 8 * k, bp_desc + koff_c, idesc_c, (kk | k) != 0);
                                }
                            }
                            if constexpr (kMC) umma_commit_multicast(&empty_bar[s], 3);
                            else umma_commit(&empty_bar[s]);
                            umma_commit(&ta_empty[j]);
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
        // TMEM lane quarter = warp % 4: this thread owns row (quarter*32 + lane) of the 128-row activation tile
        const int quarter = warp & 3, row = quarter * 32 + lane;
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + Cfg::TA_COL0;
        int it = 0;
        for (int t = t_first; t < n_tiles; t += t_step) {
            for (int kk = 0; kk < total_kb; ++kk, ++it) {
                const int s = it % STAGES, j = it % TA;
                mbar_wait(&full_bar[s], (it / STAGES) & 1);
                uint32_t hi[32], pk[32];
                if constexpr (kTN) {
                    // [32 reduction rows][128 columns] plain: lane m reads column m (conflict-free LDS.32)
                    const uint32_t* colp = reinterpret_cast<const uint32_t*>(tile(s, T_A)) + row;
                    float lo[32];
#pragma unroll
                    for (int r = 0; r < 32; ++r) {
                        const uint32_t v = colp[r * 128];
                        hi[r] = tf32_rn_bits(v);
                        lo[r] = __uint_as_float(v) - __uint_as_float(hi[r]);
                    }
                    // packed K' order: [lo(0..31) | hi(0..31)] against B' rows [hi(0..31) | lo(0..31)]
#pragma unroll
                    for (int q = 0; q < 16; ++q) {
                        pk[q] = pack_bf16x2(lo[2 * q], lo[2 * q + 1]);
                        pk[16 + q] = pack_bf16x2(__uint_as_float(hi[2 * q]), __uint_as_float(hi[2 * q + 1]));
                    }
                } else {
                    // K-major row of 128 bytes under the 128B swizzle: 16-byte chunk c sits at c ^ (row & 7)
                    const uint8_t* rowp = tile(s, T_A) + row * 128;
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const uint4 v = *reinterpret_cast<const uint4*>(rowp + ((c ^ (row & 7)) << 4));
                        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
                        float lo[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            hi[4 * c + e] = tf32_rn_bits(w[e]);
                            lo[e] = __uint_as_float(w[e]) - __uint_as_float(hi[4 * c + e]);
                        }
                        if constexpr (kX3) {
#pragma unroll
                            for (int e = 0; e < 4; ++e) pk[4 * c + e] = tf32_rn_bits(__float_as_uint(lo[e]));
                            continue;
                        }
                        // packed chunk c = {lo[4c..4c+3], hi[4c..4c+3]} against B' chunk c = {b_hi[4c..], b_lo[4c..]}
                        pk[4 * c + 0] = pack_bf16x2(lo[0], lo[1]);
                        pk[4 * c + 1] = pack_bf16x2(lo[2], lo[3]);
                        pk[4 * c + 2] = pack_bf16x2(__uint_as_float(hi[4 * c]), __uint_as_float(hi[4 * c + 1]));
                        pk[4 * c + 3] = pack_bf16x2(__uint_as_float(hi[4 * c + 2]), __uint_as_float(hi[4 * c + 3]));
                    }
                }
                mbar_wait(&ta_empty[j], ((it / TA) & 1) ^ 1);            // the MMAs that read this TMEM slot retired
                tc_fence_after_sync();
                tmem_st_32x32(lane_addr + j * Cfg::TA_COLS, hi);
                tmem_st_32x32(lane_addr + j * Cfg::TA_COLS + 32, pk);
                tmem_st_wait();
                tc_fence_before_sync();
                mbar_arrive(&split_bar[j]);
            }
        }
    } else {
        // ===================================================================== drain + epilogue (warps 6..13)
        const int quarter = warp & 3, half = (warp - 6) >> 2;
        const int colbase = half * X3Cfg::DRAIN_COLS;
        const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + colbase;
        float* my_buf = epi_buf + (warp - 6) * 32 * X3Cfg::EPI_PITCH;
        int cg = 0;
        for (int t = t_first; t < n_tiles; t += t_step) {
            const int m0 = tile_m0(t), n0 = (t % tiles_n) * Cfg::BLOCK_N;
            float acc[X3Cfg::DRAIN_COLS];
#pragma unroll
            for (int q = 0; q < X3Cfg::DRAIN_COLS; ++q) acc[q] = 0.0f;
            for (int c = 0; c < n_chunks; ++c, ++cg) {
                const int buf = cg & 1;
                mbar_wait(&acc_full[buf], (cg >> 1) & 1);
                tc_fence_after_sync();
                uint32_t r0[32], r1[32];
                tmem_ld_32x32(lane_base + buf * 128, r0);
                tmem_ld_32x32(lane_base + buf * 128 + 32, r1);
                tmem_ld_wait();
                tc_fence_before_sync();
                mbar_arrive(&acc_empty[buf]);
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[q] += __uint_as_float(r0[q]);
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[32 + q] += __uint_as_float(r1[q]);
            }
            if (n_chunks > 0) {
                uint32_t r0[32], r1[32];
                tmem_ld_32x32(lane_base + 256, r0);
                tmem_ld_32x32(lane_base + 256 + 32, r1);
                tmem_ld_wait();
                tc_fence_before_sync();
                mbar_arrive(small_empty);
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[q] += __uint_as_float(r0[q]);
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[32 + q] += __uint_as_float(r1[q]);
            } else {
                mbar_arrive(small_empty);
            }
            if constexpr (kTN) {
                // partial tile of split blockIdx.y -> its slab of the workspace; transposed: slab is [N, M] and a warp
                // store writes 32 consecutive m of one column n (coalesced straight from the row-per-lane registers)
                float* slab = static_cast<float*>(p.C) + static_cast<int64_t>(blockIdx.y) * p.M * p.N;
                const int m = m0 + quarter * 32 + lane;
                if (transpose_out) {
                    if (m < p.M) {
#pragma unroll
                        for (int q = 0; q < X3Cfg::DRAIN_COLS; ++q) {
                            const int n = n0 + colbase + q;
                            if (n < p.N) slab[static_cast<int64_t>(n) * p.M + m] = acc[q];
                        }
                    }
                } else {
                    GemmParams q = p;
                    q.C = slab;
                    if (m0 + quarter * 32 < p.M) epilogue_from_regs(q, acc, my_buf, m0 + quarter * 32, n0 + colbase, lane);
                }
            } else {
                if (m0 + quarter * 32 < p.M) epilogue_from_regs(p, acc, my_buf, m0 + quarter * 32, n0 + colbase, lane);
            }
        }
    }
    tc_fence_before_sync();
    if constexpr (kMC) cluster_sync_all();      // the peer may still multicast into / arrive on this CTA's shared memory
    else __syncthreads();
    if (warp == 1) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

// K-blocks accumulated in TMEM before the partial sum is promoted to fp32 registers.  Measured (scripts/gemm_promo.py):
// the drain of a 128x128 accumulator costs about as much as 3 K-blocks of MMAs here, so promo = 2 stalls the tensor
// pipe (0.59 ms for [128000x1792]x[1792x256]) and promo = 4 does not (0.50 ms) at 7.4e-7 instead of 6.9e-7.
static int g_bc_promo = 4;

static bool g_bc_multicast = false;     // grl_gemm_bc_set_multicast(1): CTA pairs sharing weight tiles; measured EQUAL

template <bool kTN, bool kX3, bool kMC = false>
static int ensure_bc_smem_attr() {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm_bc_kernel<kTN, kX3, kMC>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             BcCfg::SMEM_BYTES);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)");
        attr_set = true;
    }
    return 0;
}

// x3 = false: B*lo of GemmArgs = the packed bf16 weight operands [N, 2 * ceil4(K)] of grl_split_tf32_mix, row pitch
// 2 * ldb.  x3 = true: B*lo = the fp32 low parts of grl_split_tf32 (pitch ldb), three-product arithmetic.
int launch_gemm_bc(const GemmArgs& a, GemmParams p, cudaStream_t st, bool x3, int promo_x3) {
    using Cfg = BcCfg;
    {
        const int rc2 = try_launch_gemm_bc_2sm(a, p, st, x3, x3 ? promo_x3 : g_bc_promo);
        if (rc2 != 1) return rc2;
    }
    CUtensorMap tA0, tB0, tBp0, tA1, tB1, tBp1;
    int rc;
    auto lo_map = [&](CUtensorMap* out, const void* ptr, int K, int64_t ldb) {
        return x3 ? make_tmap_sw128(out, ptr, 4, p.N, K, ldb, Cfg::BLOCK_N)
                  : make_tmap_sw128(out, ptr, 2, p.N, 2 * ((K + 3) / 4 * 4), 2 * ldb, Cfg::BLOCK_N);
    };
    if ((rc = make_tmap_sw128(&tA0, a.A0, 4, p.M, a.K0, a.lda0, Cfg::BLOCK_M))) return rc;
    if ((rc = make_tmap_sw128(&tB0, a.B0, 4, p.N, a.K0, a.ldb0, Cfg::BLOCK_N))) return rc;
    if ((rc = lo_map(&tBp0, a.B0lo, a.K0, a.ldb0))) return rc;
    tA1 = tA0; tB1 = tB0; tBp1 = tBp0;
    if (a.K1 > 0) {
        if ((rc = make_tmap_sw128(&tA1, a.A1, 4, p.M, a.K1, a.lda1, Cfg::BLOCK_M))) return rc;
        if ((rc = make_tmap_sw128(&tB1, a.B1, 4, p.N, a.K1, a.ldb1, Cfg::BLOCK_N))) return rc;
        if ((rc = lo_map(&tBp1, a.B1lo, a.K1, a.ldb1))) return rc;
    }
    p.kblocks0 = (a.K0 + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K;
    p.kblocks1 = a.K1 > 0 ? (a.K1 + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K : 0;
    if ((rc = x3 ? ensure_bc_smem_attr<false, true>() : ensure_bc_smem_attr<false, false>())) return rc;
    const int tiles_m = (p.M + Cfg::BLOCK_M - 1) / Cfg::BLOCK_M, tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    const int n_tiles = tiles_m * tiles_n;
    const int n_pairs = (tiles_m + 1) / 2 * tiles_n;
    if (!x3 && g_bc_multicast && n_pairs >= kNumSMs / 2) {
        // clusters of 2 CTAs on adjacent row tiles sharing their weight tiles by TMA multicast (64-row weight boxes)
        auto half_map = [&](CUtensorMap* hi, CUtensorMap* pk, const void* b, const void* blo, int K, int64_t ldb) {
            int r = make_tmap_sw128(hi, b, 4, p.N, K, ldb, Cfg::BLOCK_N / 2);
            return r ? r : make_tmap_sw128(pk, blo, 2, p.N, 2 * ((K + 3) / 4 * 4), 2 * ldb, Cfg::BLOCK_N / 2);
        };
        if ((rc = half_map(&tB0, &tBp0, a.B0, a.B0lo, a.K0, a.ldb0))) return rc;
        tB1 = tB0; tBp1 = tBp0;
        if (a.K1 > 0 && (rc = half_map(&tB1, &tBp1, a.B1, a.B1lo, a.K1, a.ldb1))) return rc;
        if ((rc = ensure_bc_smem_attr<false, false, true>())) return rc;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(kNumSMs);
        cfg.blockDim = dim3(Cfg::THREADS);
        cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        // persistent grid = the clusters that are co-resident (GPCs with an odd SM count leave an SM without a partner)
        static int max_clusters = 0;
        if (max_clusters == 0) {
            int n = 0;
            cudaError_t eo = cudaOccupancyMaxActiveClusters(&n, gemm_bc_kernel<false, false, true>, &cfg);
            if (eo != cudaSuccess) return cuda_fail(eo, "cudaOccupancyMaxActiveClusters(gemm_bc_kernel<NT, multicast>)");
            max_clusters = n < 1 ? 1 : (n > kNumSMs / 2 ? kNumSMs / 2 : n);
        }
        cfg.gridDim = dim3(2 * (n_pairs < max_clusters ? n_pairs : max_clusters));
        cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_bc_kernel<false, false, true>, p, g_bc_promo, 0, n_pairs, 0, tA0, tB0,
                                           tBp0, tA1, tB1, tBp1);
        if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(gemm_bc_kernel<NT, multicast>)");
        return 0;
    }
    const int grid = n_tiles < kNumSMs ? n_tiles : kNumSMs;
    if (x3)
        gemm_bc_kernel<false, true><<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(p, promo_x3, 0, n_tiles, 0, tA0, tB0, tBp0,
                                                                                tA1, tB1, tBp1);
    else
        gemm_bc_kernel<false, false><<<grid, Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(p, g_bc_promo, 0, n_tiles, 0, tA0, tB0, tBp0,
                                                                                 tA1, tB1, tBp1);
    GRL_LAUNCH_CHECK("gemm_bc_kernel<NT>");
    return 0;
}

// Weight operand: hi = src rounded to nearest tf32 (fp32 storage), packed = bf16 [rows, 2 * ceil4(cols)] with 16-byte
// chunk c of a row = {hi[4c..4c+3], lo[4c..4c+3]} (lo = src - hi), zeros past cols.
__global__ void split_tf32_mix_kernel(const float* __restrict__ src, int64_t ld_src, float* __restrict__ hi, int64_t ld_hi,
                                      uint4* __restrict__ packed, int64_t ld_packed_chunks, int64_t rows, int cols,
                                      int64_t cs) {          // cs: element stride between the columns of `src`
    const int chunks = (cols + 3) / 4;
    const int64_t total = rows * chunks;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t r = t / chunks;
        const int c = static_cast<int>(t - r * chunks);
        float h[4], l[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int k = 4 * c + e;
            const float v = k < cols ? src[r * ld_src + k * cs] : 0.0f;
            h[e] = __uint_as_float(tf32_rn_bits(__float_as_uint(v)));
            l[e] = v - h[e];
            if (k < cols) hi[r * ld_hi + k] = h[e];
        }
        packed[r * ld_packed_chunks + c] = make_uint4(pack_bf16x2(h[0], h[1]), pack_bf16x2(h[2], h[3]),
                                                      pack_bf16x2(l[0], l[1]), pack_bf16x2(l[2], l[3]));
    }
}

// Narrow activation operand of the weight-gradient kernel, split once in HBM: hi [R, cols] fp32 (pitch ld_hi) and
// packed bf16 [ceil(R/32)*64, ld_packed]: K-block kb holds rows 64kb + r%32 = bf16(hi[r]) and 64kb + 32 + r%32 =
// bf16(lo[r]); rows past R and columns in [cols, ld_*) are zero.  One thread per (row, 8-column group).
__global__ void split_act_mn_kernel(const float* __restrict__ src, int64_t ld_src, float* __restrict__ hi, int64_t ld_hi,
                                    __nv_bfloat16* __restrict__ packed, int64_t ld_packed, int64_t R, int cols) {
    const int groups = static_cast<int>(ld_packed / 8);
    const int64_t Rp = (R + 31) / 32 * 32;
    const int64_t total = Rp * groups;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t r = t / groups;
        const int c0 = static_cast<int>(t - r * groups) * 8;
        float h[8], l[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const float v = (r < R && c0 + e < cols) ? src[r * ld_src + c0 + e] : 0.0f;
            h[e] = __uint_as_float(tf32_rn_bits(__float_as_uint(v)));
            l[e] = v - h[e];
            if (r < R && c0 + e < ld_hi) hi[r * ld_hi + c0 + e] = h[e];
        }
        const int64_t row_hi = (r >> 5) * 64 + (r & 31);
        *reinterpret_cast<uint4*>(packed + row_hi * ld_packed + c0) =
            make_uint4(pack_bf16x2(h[0], h[1]), pack_bf16x2(h[2], h[3]), pack_bf16x2(h[4], h[5]), pack_bf16x2(h[6], h[7]));
        *reinterpret_cast<uint4*>(packed + (row_hi + 32) * ld_packed + c0) =
            make_uint4(pack_bf16x2(l[0], l[1]), pack_bf16x2(l[2], l[3]), pack_bf16x2(l[4], l[5]), pack_bf16x2(l[6], l[7]));
    }
}

// Producer of the backward operands of one layer in ONE pass over dY / Y (fp32):
//   dZ = dY * (Y > 0 ? 1/(1-p) : 0)                      (autograd of F.relu + nn.Dropout, drop_robust_gcn.py:64-100)
//   hi / packed = the grl_split_act_mn split of dZ        (B operand of the weight-gradient kernel)
//   partial[block][c] = column sums of dZ over the block's 32 rows   (bias gradient; reduced in block order afterwards)
// Block = 32 rows (one K-block of the packed layout) x C columns, 8 columns per thread; needs C % 8 == 0 and
// 256 % (C / 8) == 0.  Replaces bias_act_dropout_bwd + colsum (atomics) + split_act_mn: 0.66 GB instead of 0.92 GB of
// traffic per 128000 x 256 layer, and a bias gradient that is summed in a fixed order.
__global__ void __launch_bounds__(256)
dz_split_kernel(const float* __restrict__ dY, int64_t lddy, const float* __restrict__ Y, int64_t ldy,
                float* __restrict__ dZ, int64_t lddz, float* __restrict__ hi, __nv_bfloat16* __restrict__ packed,
                int64_t ld_split, int64_t M, int C, bool relu, float scale, float* __restrict__ partial) {
    __shared__ float red[256 * 8];
    const int groups = C >> 3, rows_per_pass = 256 / groups;
    const int cg = threadIdx.x % groups, r_first = threadIdx.x / groups;
    const int c0 = cg * 8;
    const int64_t row0 = static_cast<int64_t>(blockIdx.x) * 32;
    float sum[8] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    for (int r = r_first; r < 32; r += rows_per_pass) {
        const int64_t m = row0 + r;
        float z[8] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
        if (m < M) {
            const float4 g0 = *reinterpret_cast<const float4*>(dY + m * lddy + c0);
            const float4 g1 = *reinterpret_cast<const float4*>(dY + m * lddy + c0 + 4);
            float g[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
            if (relu) {
                const float4 y0 = *reinterpret_cast<const float4*>(Y + m * ldy + c0);
                const float4 y1 = *reinterpret_cast<const float4*>(Y + m * ldy + c0 + 4);
                const float y[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
#pragma unroll
                for (int k = 0; k < 8; ++k) z[k] = y[k] > 0.0f ? g[k] * scale : 0.0f;
            } else {
#pragma unroll
                for (int k = 0; k < 8; ++k) z[k] = g[k];
            }
            *reinterpret_cast<float4*>(dZ + m * lddz + c0) = make_float4(z[0], z[1], z[2], z[3]);
            *reinterpret_cast<float4*>(dZ + m * lddz + c0 + 4) = make_float4(z[4], z[5], z[6], z[7]);
#pragma unroll
            for (int k = 0; k < 8; ++k) sum[k] += z[k];
        }
        if (packed) {
            float h[8], l[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                h[k] = __uint_as_float(tf32_rn_bits(__float_as_uint(z[k])));
                l[k] = z[k] - h[k];
            }
            if (m < M) {
                *reinterpret_cast<float4*>(hi + m * ld_split + c0) = make_float4(h[0], h[1], h[2], h[3]);
                *reinterpret_cast<float4*>(hi + m * ld_split + c0 + 4) = make_float4(h[4], h[5], h[6], h[7]);
            }
            const int64_t prow = static_cast<int64_t>(blockIdx.x) * 64 + r;          // rows past M are zero-filled
            *reinterpret_cast<uint4*>(packed + prow * ld_split + c0) =
                make_uint4(pack_bf16x2(h[0], h[1]), pack_bf16x2(h[2], h[3]), pack_bf16x2(h[4], h[5]), pack_bf16x2(h[6], h[7]));
            *reinterpret_cast<uint4*>(packed + (prow + 32) * ld_split + c0) =
                make_uint4(pack_bf16x2(l[0], l[1]), pack_bf16x2(l[2], l[3]), pack_bf16x2(l[4], l[5]), pack_bf16x2(l[6], l[7]));
        }
    }
    if (partial) {
#pragma unroll
        for (int k = 0; k < 8; ++k) red[r_first * C + c0 + k] = sum[k];
        __syncthreads();
        for (int c = threadIdx.x; c < C; c += 256) {
            float s = 0.0f;
            for (int q = 0; q < rows_per_pass; ++q) s += red[q * C + c];
            partial[static_cast<int64_t>(blockIdx.x) * C + c] = s;
        }
    }
}

// The same producer for bf16 activations (no operand split: the bf16 weight-gradient kernel reads dZ in place):
//   dZ = bf16(dY * (Y > 0 ? 1/(1-p) : 0)),  partial[block][c] = column sums of the STORED dZ over the block's 32 rows.
// Replaces bias_act_dropout_bwd + colsum (atomics) of the bf16 backward: one read of dY / Y, a bias gradient summed in a
// fixed order.
__global__ void __launch_bounds__(256)
dz_colsum_bf16_kernel(const __nv_bfloat16* __restrict__ dY, int64_t lddy, const __nv_bfloat16* __restrict__ Y, int64_t ldy,
                      __nv_bfloat16* __restrict__ dZ, int64_t lddz, int64_t M, int C, bool relu, float scale,
                      float* __restrict__ partial) {
    __shared__ float red[256 * 8];
    const int groups = C >> 3, rows_per_pass = 256 / groups;
    const int cg = threadIdx.x % groups, r_first = threadIdx.x / groups;
    const int c0 = cg * 8;
    const int64_t row0 = static_cast<int64_t>(blockIdx.x) * 32;
    float sum[8] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    for (int r = r_first; r < 32; r += rows_per_pass) {
        const int64_t m = row0 + r;
        if (m >= M) continue;
        float g[8], y[8];
        Vec16<__nv_bfloat16>::load(dY + m * lddy + c0, g);
        if (relu) {
            Vec16<__nv_bfloat16>::load(Y + m * ldy + c0, y);
#pragma unroll
            for (int k = 0; k < 8; ++k) g[k] = y[k] > 0.0f ? g[k] * scale : 0.0f;
        }
        const uint4 packed = Vec16<__nv_bfloat16>::pack(g);
        *reinterpret_cast<uint4*>(dZ + m * lddz + c0) = packed;
        const uint32_t w[4] = {packed.x, packed.y, packed.z, packed.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            sum[2 * k] += __uint_as_float(w[k] << 16);
            sum[2 * k + 1] += __uint_as_float(w[k] & 0xFFFF0000u);
        }
    }
    if (partial) {
#pragma unroll
        for (int k = 0; k < 8; ++k) red[r_first * C + c0 + k] = sum[k];
        __syncthreads();
        for (int c = threadIdx.x; c < C; c += 256) {
            float sacc = 0.0f;
            for (int q = 0; q < rows_per_pass; ++q) sacc += red[q * C + c];
            partial[static_cast<int64_t>(blockIdx.x) * C + c] = sacc;
        }
    }
}

// out[c] = sum_b partial[b][c] in a fixed order: thread (x, y) adds rows y, y + 32, ... of column blockIdx.x*32 + x, then the
// 32 row-lane sums are added in lane order.  (A single block walking all partial rows serially costs 0.2 ms per layer.)
__global__ void __launch_bounds__(1024) colsum_partials_kernel(const float* __restrict__ partial, int n_rows, int C,
                                                               float* __restrict__ out) {
    __shared__ float part[32][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    float acc = 0.0f;
    if (c < C) {
        // eight blocks for C = 256: the pass is bound by load latency, so sixteen loads are in flight per thread (four gave
        // 21 us for 4000 partial rows); the additions keep their order
        int r = threadIdx.y;
        for (; r + 15 * 32 < n_rows; r += 16 * 32) {
            float v[16];
#pragma unroll
            for (int u = 0; u < 16; ++u) v[u] = __ldg(partial + static_cast<int64_t>(r + 32 * u) * C + c);
#pragma unroll
            for (int u = 0; u < 16; ++u) acc += v[u];
        }
        for (; r < n_rows; r += 32) acc += __ldg(partial + static_cast<int64_t>(r) * C + c);
    }
    part[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && c < C) {
        float s = 0.0f;
#pragma unroll
        for (int k = 0; k < 32; ++k) s += part[k][threadIdx.x];
        out[c] = s;
    }
}

}  // namespace grl

using namespace grl;

extern "C" int grl_gemm_bc_set_multicast(int on) {
    g_bc_multicast = on != 0;
    return 0;
}

extern "C" int grl_gemm_bc_set_promo(int promo) {
    if (promo < 1 || promo > 1 << 20) {
        set_error("grl_gemm_bc_set_promo: promo=%d must be >= 1", promo);
        return GRL_E_BADARG;
    }
    g_bc_promo = promo;
    return 0;
}

extern "C" int grl_split_tf32_mix(const float* src, int64_t ld_src, float* hi, int64_t ld_hi, void* packed,
                                  int64_t ld_packed, int64_t rows, int cols, void* stream) {
    GRL_CHECK_ARG(src && hi && packed, GRL_E_BADARG, "grl_split_tf32_mix: null pointer");
    const int64_t c4 = (static_cast<int64_t>(cols) + 3) / 4 * 4;
    GRL_CHECK_ARG(rows >= 0 && cols > 0 && ld_src >= cols && ld_hi >= cols && ld_packed >= 2 * c4, GRL_E_BADARG,
                  "grl_split_tf32_mix: bad shape");
    GRL_CHECK_ARG(ld_packed % 8 == 0 && aligned_to(packed, 16), GRL_E_ALIGN,
                  "grl_split_tf32_mix: packed rows need a 16-byte aligned base and pitch");
    if (rows == 0) return 0;
    int64_t blocks = (rows * (c4 / 4) + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    split_tf32_mix_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        src, ld_src, hi, ld_hi, static_cast<uint4*>(packed), ld_packed / 8, rows, cols, 1);
    GRL_LAUNCH_CHECK("split_tf32_mix_kernel");
    return 0;
}

extern "C" int grl_split_tf32_mix_strided(const float* src, int64_t stride_row, int64_t stride_col, float* hi, int64_t ld_hi,
                                          void* packed, int64_t ld_packed, int64_t rows, int cols, void* stream) {
    GRL_CHECK_ARG(src && hi && packed, GRL_E_BADARG, "grl_split_tf32_mix_strided: null pointer");
    const int64_t c4 = (static_cast<int64_t>(cols) + 3) / 4 * 4;
    GRL_CHECK_ARG(rows >= 0 && cols > 0 && stride_row > 0 && stride_col > 0 && ld_hi >= cols && ld_packed >= 2 * c4, GRL_E_BADARG,
                  "grl_split_tf32_mix_strided: bad shape");
    GRL_CHECK_ARG(ld_packed % 8 == 0 && aligned_to(packed, 16), GRL_E_ALIGN,
                  "grl_split_tf32_mix_strided: packed rows need a 16-byte aligned base and pitch");
    if (rows == 0) return 0;
    int64_t blocks = (rows * (c4 / 4) + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    split_tf32_mix_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        src, stride_row, hi, ld_hi, static_cast<uint4*>(packed), ld_packed / 8, rows, cols, stride_col);
    GRL_LAUNCH_CHECK("split_tf32_mix_kernel");
    return 0;
}

extern "C" int grl_split_act_mn(const float* src, int64_t ld_src, float* hi, int64_t ld_hi, void* packed,
                                int64_t ld_packed, int64_t R, int cols, void* stream) {
    GRL_CHECK_ARG(src && hi && packed, GRL_E_BADARG, "grl_split_act_mn: null pointer");
    GRL_CHECK_ARG(R > 0 && cols > 0 && ld_src >= cols && ld_hi >= cols && ld_packed >= cols, GRL_E_BADARG,
                  "grl_split_act_mn: bad shape");
    GRL_CHECK_ARG(ld_packed % 8 == 0 && ld_hi % 4 == 0 && ld_hi <= ld_packed && aligned_to(packed, 16) && aligned_to(hi, 16),
                  GRL_E_ALIGN, "grl_split_act_mn: hi / packed need 16-byte aligned bases and pitches (ld_hi <= ld_packed)");
    const int64_t Rp = (R + 31) / 32 * 32;
    int64_t blocks = (Rp * (ld_packed / 8) + 255) / 256;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    split_act_mn_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        src, ld_src, hi, ld_hi, static_cast<__nv_bfloat16*>(packed), ld_packed, R, cols);
    GRL_LAUNCH_CHECK("split_act_mn_kernel");
    return 0;
}

extern "C" int grl_gemm_tn_bc(const float* A, int64_t lda, int Ma, const float* B_hi, int64_t ldb_hi, const void* B_packed,
                              int64_t ldb_packed, int Nb, int64_t R, float* C, int transpose_out, void* workspace,
                              size_t workspace_bytes, void* stream) {
    using Cfg = BcCfg;
    GRL_CHECK_ARG(A && B_hi && B_packed && C && workspace, GRL_E_BADARG, "grl_gemm_tn_bc: null pointer");
    GRL_CHECK_ARG(Ma > 0 && Nb > 0 && R > 0 && R < (1LL << 30), GRL_E_BADARG, "grl_gemm_tn_bc: bad shape");
    GRL_CHECK_ARG(lda >= Ma && ldb_hi >= Nb && ldb_packed >= Nb, GRL_E_BADARG, "grl_gemm_tn_bc: bad leading dimension");
    GRL_CHECK_ARG(lda % 4 == 0 && ldb_hi % 4 == 0 && ldb_packed % 8 == 0 && aligned_to(A, 16) && aligned_to(B_hi, 16) &&
                      aligned_to(B_packed, 16),
                  GRL_E_ALIGN, "grl_gemm_tn_bc: operands need 16-byte aligned base pointers and row pitches");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CUtensorMap tA, tB, tBp;
    int rc;
    if ((rc = make_tmap_plain(&tA, A, 4, R, Ma, lda, Cfg::BLOCK_M, Cfg::BLOCK_K))) return rc;
    if ((rc = make_tmap_sw128(&tB, B_hi, 4, R, Nb, ldb_hi, Cfg::BLOCK_K, true))) return rc;
    const int64_t kblocks = (R + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K;
    if ((rc = make_tmap_sw128(&tBp, B_packed, 2, kblocks * 64, Nb, ldb_packed, 64))) return rc;
    GemmParams p{};
    p.M = Ma; p.N = Nb; p.ldc = Nb; p.C = workspace;
    p.kblocks0 = static_cast<int>(kblocks);
    p.kblocks1 = 0;
    {
        int sp = 0;
        const int rc2 = try_launch_gemm_bc_tn_2sm(p, g_bc_promo, transpose_out, tA, tB, tBp, &sp, workspace_bytes, st);
        if (rc2 == 0) return launch_reduce_splits(static_cast<const float*>(workspace), C, static_cast<int64_t>(Ma) * Nb, sp, st);
        if (rc2 != 1) return rc2;
    }
    const int tiles = ((Ma + Cfg::BLOCK_M - 1) / Cfg::BLOCK_M) * ((Nb + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N);
    int splits = tn_plan_splits(tiles, p.kblocks0);
    const int kb_per_split = (p.kblocks0 + splits - 1) / splits;
    splits = (p.kblocks0 + kb_per_split - 1) / kb_per_split;
    const size_t need = sizeof(float) * static_cast<size_t>(Ma) * Nb * splits;
    GRL_CHECK_ARG(workspace_bytes >= need, GRL_E_WORKSPACE, "grl_gemm_tn_bc: workspace %zu < %zu", workspace_bytes, need);
    if ((rc = ensure_bc_smem_attr<true, false>())) return rc;
    gemm_bc_kernel<true, false><<<dim3(tiles, splits), Cfg::THREADS, Cfg::SMEM_BYTES, st>>>(p, g_bc_promo, kb_per_split, tiles,
                                                                                    transpose_out, tA, tB, tBp, tA, tB, tBp);
    GRL_LAUNCH_CHECK("gemm_bc_kernel<TN>");
    return launch_reduce_splits(static_cast<const float*>(workspace), C, static_cast<int64_t>(Ma) * Nb, splits, st);
}

extern "C" size_t grl_dz_split_workspace_bytes(int64_t M, int C);

extern "C" int grl_dz_colsum_bf16(const void* dY, int64_t lddy, const void* Y, int64_t ldy, void* dZ, int64_t lddz, int64_t M,
                                  int C, uint32_t flags, float p, float* colsum_out, void* workspace, size_t workspace_bytes,
                                  void* stream) {
    GRL_CHECK_ARG(dY && dZ, GRL_E_BADARG, "grl_dz_colsum_bf16: null pointer");
    GRL_CHECK_ARG(M > 0 && C > 0 && C % 8 == 0 && C <= 2048 && 256 % (C / 8) == 0, GRL_E_UNSUP,
                  "grl_dz_colsum_bf16: C=%d must be a multiple of 8 with 256 %% (C/8) == 0", C);
    const bool relu = (flags & GRL_EPI_RELU) != 0;
    GRL_CHECK_ARG(relu || !(flags & GRL_EPI_DROPOUT) || p == 0.0f, GRL_E_UNSUP, "grl_dz_colsum_bf16: dropout without ReLU is not supported");
    GRL_CHECK_ARG(!relu || Y, GRL_E_BADARG, "grl_dz_colsum_bf16: RELU needs the forward output Y");
    GRL_CHECK_ARG(lddy % 8 == 0 && lddz % 8 == 0 && (!relu || ldy % 8 == 0) && aligned_to(dY, 16) && aligned_to(dZ, 16) &&
                      (!relu || aligned_to(Y, 16)),
                  GRL_E_ALIGN, "grl_dz_colsum_bf16: operands need 16-byte aligned bases and pitches");
    GRL_CHECK_ARG(!colsum_out || (workspace && workspace_bytes >= grl_dz_split_workspace_bytes(M, C)), GRL_E_WORKSPACE,
                  "grl_dz_colsum_bf16: workspace too small for the column sums");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const float scale = (flags & GRL_EPI_DROPOUT) ? 1.0f / (1.0f - p) : 1.0f;
    const int blocks = static_cast<int>((M + 31) / 32);
    dz_colsum_bf16_kernel<<<blocks, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(dY), lddy, static_cast<const __nv_bfloat16*>(Y), ldy,
                                                  static_cast<__nv_bfloat16*>(dZ), lddz, M, C, relu, scale,
                                                  colsum_out ? static_cast<float*>(workspace) : nullptr);
    GRL_LAUNCH_CHECK("dz_colsum_bf16_kernel");
    if (colsum_out) {
        colsum_partials_kernel<<<(C + 31) / 32, dim3(32, 32), 0, st>>>(static_cast<const float*>(workspace), blocks, C, colsum_out);
        GRL_LAUNCH_CHECK("colsum_partials_kernel");
    }
    return 0;
}

extern "C" size_t grl_dz_split_workspace_bytes(int64_t M, int C) {
    return sizeof(float) * static_cast<size_t>((M + 31) / 32) * C;
}

extern "C" int grl_dz_split(const float* dY, int64_t lddy, const float* Y, int64_t ldy, float* dZ, int64_t lddz, float* hi,
                            void* packed, int64_t ld_split, int64_t M, int C, uint32_t flags, float p, float* colsum_out,
                            void* workspace, size_t workspace_bytes, void* stream) {
    GRL_CHECK_ARG(dY && dZ, GRL_E_BADARG, "grl_dz_split: null pointer");
    GRL_CHECK_ARG(M > 0 && C > 0 && C % 8 == 0 && C <= 2048 && 256 % (C / 8) == 0, GRL_E_UNSUP,
                  "grl_dz_split: C=%d must be a multiple of 8 with 256 %% (C/8) == 0", C);
    const bool relu = (flags & GRL_EPI_RELU) != 0;
    GRL_CHECK_ARG(relu || !(flags & GRL_EPI_DROPOUT) || p == 0.0f, GRL_E_UNSUP, "grl_dz_split: dropout without ReLU is not supported");
    GRL_CHECK_ARG(!relu || Y, GRL_E_BADARG, "grl_dz_split: RELU needs the forward output Y");
    GRL_CHECK_ARG((hi == nullptr) == (packed == nullptr), GRL_E_BADARG, "grl_dz_split: hi and packed go together");
    GRL_CHECK_ARG(lddy % 4 == 0 && lddz % 4 == 0 && (!relu || ldy % 4 == 0) && aligned_to(dY, 16) && aligned_to(dZ, 16) &&
                      (!relu || aligned_to(Y, 16)),
                  GRL_E_ALIGN, "grl_dz_split: operands need 16-byte aligned bases and pitches");
    GRL_CHECK_ARG(!packed || (ld_split >= C && ld_split % 8 == 0 && aligned_to(hi, 16) && aligned_to(packed, 16)), GRL_E_ALIGN,
                  "grl_dz_split: hi / packed need 16-byte aligned bases and a pitch that is a multiple of 8");
    GRL_CHECK_ARG(!colsum_out || (workspace && workspace_bytes >= grl_dz_split_workspace_bytes(M, C)), GRL_E_WORKSPACE,
                  "grl_dz_split: workspace too small for the column sums");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const float scale = (flags & GRL_EPI_DROPOUT) ? 1.0f / (1.0f - p) : 1.0f;
    const int blocks = static_cast<int>((M + 31) / 32);
    dz_split_kernel<<<blocks, 256, 0, st>>>(dY, lddy, Y, ldy, dZ, lddz, hi, static_cast<__nv_bfloat16*>(packed), ld_split, M, C,
                                            relu, scale, colsum_out ? static_cast<float*>(workspace) : nullptr);
    GRL_LAUNCH_CHECK("dz_split_kernel");
    if (colsum_out) {
        colsum_partials_kernel<<<(C + 31) / 32, dim3(32, 32), 0, st>>>(static_cast<const float*>(workspace), blocks, C, colsum_out);
        GRL_LAUNCH_CHECK("colsum_partials_kernel");
    }
    return 0;
}
