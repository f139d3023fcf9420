// fp32-grade NT GEMMs on CTA PAIRS (tcgen05 cta_group::2): the arithmetic and the warp roles of gemm_bc_kernel<NT>
// (gemm_bc_tcgen05.cu: activation operand split in registers and fed from TMEM, weights from shared memory, promoted fp32
// accumulation, register epilogue), with two CTAs of a cluster on adjacent row tiles of the SAME column tile running ONE
// MMA of M = 256.
//
// Why: the single-CTA kernel moves, per 128x128x32 K-block, 16 KB (TMA, A) + 16 KB (splitter read) + 32 KB (TMA, weight
// tiles hi / lo-or-packed) + 32-48 KB (MMA reads of the weight tiles) through the SM's 128 B/clk shared-memory port --
// ~900 of the 960-1180 cycles a K-block takes.  Multicast (gemm_bc_kernel<NT, -, kMC>) halves only where the weight bytes
// COME FROM, and measured nothing.  In a pair each CTA holds HALF of every weight tile (64 of its 128 rows): the TMA
// writes and the MMA reads of the weight tiles halve per SM.
//
// Protocol (rank 0 = leader issues every MMA and commit):
//   a_full[s]    local   TMA bytes of my raw activation tile           -> my splitter warps
//   b_full[s]    leader  TMA bytes of BOTH CTAs' weight halves (cp.async.bulk.tensor .cta_group::2 signals the leader's
//                        barrier from the peer)                         -> the MMA warp
//   split[j]     leader  128 local + 128 remote arrivals: both CTAs' splitters have written their A rows to their TMEM
//   empty[s], ta_empty[j], acc_full[b]   both   tcgen05.commit.cta_group::2 ... multicast::cluster
//   acc_empty[b], small_empty            leader 256 local + 256 remote arrivals from the drain warps
#include "gemm_x3_parts.cuh"

namespace grl {

int gemm_x3_current_promo();    // gemm_x3_tcgen05.cu

struct Bc2Cfg {
    static constexpr int BLOCK_M = 128, BLOCK_N = 128, BLOCK_K = 32;
    static constexpr int TILE_A = 128 * 128, HALF_B = 64 * 128;
    static constexpr int STAGE_BYTES = TILE_A + 2 * HALF_B;         // A raw | B_hi half | B_lo / packed half
    static constexpr int STAGES = 6;
    static constexpr int TA_STAGES = 2, TA_COL0 = 384, TA_COLS = 64;
    static constexpr int EPI_BYTES = X3Cfg::EPI_BYTES;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 + 256;
    static constexpr int TMEM_COLS = 512;
    static constexpr int THREADS = X3Cfg::THREADS;
};
static_assert(Bc2Cfg::SMEM_BYTES <= 227 * 1024, "shared memory budget");

__device__ __forceinline__ uint32_t pack_bf16x2_2sm(float first, float second) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(first, second);
    return *reinterpret_cast<const uint32_t*>(&v);
}

template <bool kX3>
__global__ void __launch_bounds__(Bc2Cfg::THREADS, 1)
gemm_bc_2sm_kernel(const GemmParams p, const int promo, const int n_pairs, const __grid_constant__ CUtensorMap tmA0,
                   const __grid_constant__ CUtensorMap tmB0, const __grid_constant__ CUtensorMap tmBp0,
                   const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
                   const __grid_constant__ CUtensorMap tmBp1) {
    using Cfg = Bc2Cfg;
    constexpr int STAGES = Cfg::STAGES, TA = Cfg::TA_STAGES, DRAIN_WARPS = X3Cfg::DRAIN_WARPS;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    float* epi_buf = reinterpret_cast<float*>(tiles + STAGES * Cfg::STAGE_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(tiles + STAGES * Cfg::STAGE_BYTES + Cfg::EPI_BYTES);
    uint64_t* a_full = bars;                            // [STAGES]
    uint64_t* b_full = bars + STAGES;                   // [STAGES] (used in the leader)
    uint64_t* empty_bar = bars + 2 * STAGES;            // [STAGES]
    uint64_t* split_bar = bars + 3 * STAGES;            // [TA]     (leader)
    uint64_t* ta_empty = split_bar + TA;                // [TA]
    uint64_t* acc_full = ta_empty + TA;                 // [2]
    uint64_t* acc_empty = acc_full + 2;                 // [2]      (leader)
    uint64_t* small_empty = acc_empty + 2;              //          (leader)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(small_empty + 1);

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const int rank = static_cast<int>(cluster_ctarank());
    const bool leader = rank == 0;
    const int tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    const int total_kb = p.kblocks0 + p.kblocks1;
    const int n_chunks = (total_kb + promo - 1) / promo;
    const int t_first = static_cast<int>(blockIdx.x >> 1), t_step = static_cast<int>(gridDim.x >> 1);
    auto tile_m0 = [&](int t) { return (2 * (t / tiles_n) + rank) * Cfg::BLOCK_M; };
    auto tile_a = [&](int s) { return tiles + s * Cfg::STAGE_BYTES; };
    auto tile_b = [&](int s, int which) { return tiles + s * Cfg::STAGE_BYTES + Cfg::TILE_A + which * Cfg::HALF_B; };
    // barriers the leader owns, as shared::cluster addresses valid from either CTA
    auto leader_addr = [&](uint64_t* bar) { return mapa_cluster(smem_u32(bar), 0u); };

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
            mbar_init(&a_full[s], 1);
            mbar_init(&b_full[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int j = 0; j < TA; ++j) {
            mbar_init(&split_bar[j], 256);
            mbar_init(&ta_empty[j], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 2 * 32 * DRAIN_WARPS);
        }
        mbar_init(small_empty, 2 * 32 * DRAIN_WARPS);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc_2sm<Cfg::TMEM_COLS>(tmem_slot);
    tc_fence_before_sync();
    cluster_sync_all();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================================================================== TMA producer (one lane, both CTAs)
        if (lane == 0) {
            int it = 0;
            for (int t = t_first; t < n_pairs; t += t_step) {
                const int m0 = tile_m0(t), nh = (t % tiles_n) * Cfg::BLOCK_N + rank * 64;
                for (int kb = 0; kb < total_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    const bool seg1 = kb >= p.kblocks0;
                    const int k_elem = (seg1 ? kb - p.kblocks0 : kb) * Cfg::BLOCK_K;
                    mbar_expect_tx(&a_full[s], Cfg::TILE_A);
                    tma_load_2d(tile_a(s), seg1 ? &tmA1 : &tmA0, &a_full[s], k_elem, m0);
                    if (leader) mbar_expect_tx(&b_full[s], 4 * Cfg::HALF_B);          // two halves from each CTA
                    const uint32_t bf = leader_addr(&b_full[s]);
                    tma_load_2d_2sm(tile_b(s, 0), seg1 ? &tmB1 : &tmB0, bf, k_elem, nh);
                    tma_load_2d_2sm(tile_b(s, 1), seg1 ? &tmBp1 : &tmBp0, bf, kX3 ? k_elem : 2 * k_elem, nh);
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================================== MMA issuer (leader only; warp-wide loop)
        if (leader) {
            constexpr uint32_t idesc = umma_idesc(2, 2 * Cfg::BLOCK_M, Cfg::BLOCK_N, 0, 0);      // tf32, M = 256 over the pair
            constexpr uint32_t idesc_c = umma_idesc(1, 2 * Cfg::BLOCK_M, Cfg::BLOCK_N, 0, 0);    // bf16
            const uint32_t d_small = tmem_base + 256;
            int it = 0, cg = 0, tl = 0;
            for (int t = t_first; t < n_pairs; t += t_step, ++tl) {
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
                        mbar_wait(&b_full[s], (it / STAGES) & 1);
                        mbar_wait(&split_bar[j], (it / TA) & 1);
                        tc_fence_after_sync();
                        const uint32_t ta_hi = tmem_base + Cfg::TA_COL0 + j * Cfg::TA_COLS, ta_pk = ta_hi + 32;
                        const uint64_t b_desc = umma_desc_kmajor_sw128(smem_u32(tile_b(s, 0)));
                        const uint64_t bp_desc = umma_desc_kmajor_sw128(smem_u32(tile_b(s, 1)));
                        if (elect_one_sync()) {
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const uint64_t koff = static_cast<uint64_t>(k * 2);
                                umma_ts_2sm<true>(d_main, ta_hi + 8 * k, b_desc + koff, idesc, (i | k) != 0);
                                if constexpr (kX3) {
                                    umma_ts_2sm<true>(d_small, ta_pk + 8 * k, b_desc + koff, idesc, (kk | k) != 0);   // a_lo * b_hi
                                    umma_ts_2sm<true>(d_small, ta_hi + 8 * k, bp_desc + koff, idesc, 1u);             // a_hi * b_lo
                                } else {
                                    umma_ts_2sm<false>(d_small, ta_pk + 8 * k, bp_desc + koff, idesc_c, (kk | k) != 0);
                                }
                            }
                            umma_commit_2sm_multicast(&empty_bar[s], 3);
                            umma_commit_2sm_multicast(&ta_empty[j], 3);
                        }
                        __syncwarp();
                    }
                    if (elect_one_sync()) umma_commit_2sm_multicast(&acc_full[buf], 3);
                    __syncwarp();
                }
            }
        }
    } else if (warp < 6) {
        // ===================================================================== operand splitter (warps 2..5, both CTAs)
        const int quarter = warp & 3, row = quarter * 32 + lane;
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + Cfg::TA_COL0;
        const uint32_t split_remote0 = leader_addr(&split_bar[0]), split_remote1 = leader_addr(&split_bar[1]);
        int it = 0;
        for (int t = t_first; t < n_pairs; t += t_step) {
            for (int kk = 0; kk < total_kb; ++kk, ++it) {
                const int s = it % STAGES, j = it % TA;
                mbar_wait(&a_full[s], (it / STAGES) & 1);
                uint32_t hi[32], pk[32];
                const uint8_t* rowp = tile_a(s) + row * 128;
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
                    pk[4 * c + 0] = pack_bf16x2_2sm(lo[0], lo[1]);
                    pk[4 * c + 1] = pack_bf16x2_2sm(lo[2], lo[3]);
                    pk[4 * c + 2] = pack_bf16x2_2sm(__uint_as_float(hi[4 * c]), __uint_as_float(hi[4 * c + 1]));
                    pk[4 * c + 3] = pack_bf16x2_2sm(__uint_as_float(hi[4 * c + 2]), __uint_as_float(hi[4 * c + 3]));
                }
                mbar_wait(&ta_empty[j], ((it / TA) & 1) ^ 1);
                tc_fence_after_sync();
                tmem_st_32x32(lane_addr + j * Cfg::TA_COLS, hi);
                tmem_st_32x32(lane_addr + j * Cfg::TA_COLS + 32, pk);
                tmem_st_wait();
                tc_fence_before_sync();
                mbar_arrive_cluster_relaxed(j ? split_remote1 : split_remote0);
            }
        }
    } else {
        // ===================================================================== drain + epilogue (warps 6..13, both CTAs)
        const int quarter = warp & 3, half = (warp - 6) >> 2;
        const int colbase = half * X3Cfg::DRAIN_COLS;
        const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + colbase;
        float* my_buf = epi_buf + (warp - 6) * 32 * X3Cfg::EPI_PITCH;
        const uint32_t acc_empty_remote0 = leader_addr(&acc_empty[0]), acc_empty_remote1 = leader_addr(&acc_empty[1]);
        const uint32_t small_empty_remote = leader_addr(small_empty);
        int cg = 0;
        for (int t = t_first; t < n_pairs; t += t_step) {
            const int m0 = tile_m0(t), n0 = (t % tiles_n) * Cfg::BLOCK_N;
            float acc[X3Cfg::DRAIN_COLS];
#pragma unroll
            for (int q = 0; q < X3Cfg::DRAIN_COLS; ++q) acc[q] = 0.0f;
            for (int c = 0; c < n_chunks; ++c, ++cg) {
                const int buf = cg & 1;
                mbar_wait_parked(&acc_full[buf], (cg >> 1) & 1);
                tc_fence_after_sync();
                uint32_t r0[32], r1[32];
                tmem_ld_32x32(lane_base + buf * 128, r0);
                tmem_ld_32x32(lane_base + buf * 128 + 32, r1);
                tmem_ld_wait();
                tc_fence_before_sync();
                mbar_arrive_cluster_relaxed(buf ? acc_empty_remote1 : acc_empty_remote0);
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[q] += __uint_as_float(r0[q]);
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[32 + q] += __uint_as_float(r1[q]);
            }
            {
                uint32_t r0[32], r1[32];
                tmem_ld_32x32(lane_base + 256, r0);
                tmem_ld_32x32(lane_base + 256 + 32, r1);
                tmem_ld_wait();
                tc_fence_before_sync();
                mbar_arrive_cluster_relaxed(small_empty_remote);
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[q] += __uint_as_float(r0[q]);
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[32 + q] += __uint_as_float(r1[q]);
            }
            if (m0 + quarter * 32 < p.M) epilogue_from_regs(p, acc, my_buf, m0 + quarter * 32, n0 + colbase, lane);
        }
    }
    tc_fence_before_sync();
    cluster_sync_all();
    if (warp == 1) tmem_dealloc_2sm<Cfg::TMEM_COLS>(tmem_base);
}

// Weight gradients C[Ma, Nb] = A[R, Ma]^T B[R, Nb] on CTA pairs: gemm_bc_kernel<TN> (wide operand A transposed by the
// splitter's shared-memory reads into TMEM, narrow operand B = dZ pre-split in HBM and read MN-major, split-K over R with
// a deterministic reduce) with two CTAs on adjacent Ma tiles of the same Nb tile and the same K split; each CTA holds 64 of
// the 128 B columns.  One tile per CTA (grid = pairs x splits).
__global__ void __launch_bounds__(Bc2Cfg::THREADS, 1)
gemm_bc_tn_2sm_kernel(const GemmParams p, const int promo, const int kb_per_split, const int transpose_out,
                      const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmBp) {
    using Cfg = Bc2Cfg;
    constexpr int STAGES = Cfg::STAGES, TA = Cfg::TA_STAGES, DRAIN_WARPS = X3Cfg::DRAIN_WARPS;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    float* epi_buf = reinterpret_cast<float*>(tiles + STAGES * Cfg::STAGE_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(tiles + STAGES * Cfg::STAGE_BYTES + Cfg::EPI_BYTES);
    uint64_t* a_full = bars;
    uint64_t* b_full = bars + STAGES;
    uint64_t* empty_bar = bars + 2 * STAGES;
    uint64_t* split_bar = bars + 3 * STAGES;
    uint64_t* ta_empty = split_bar + TA;
    uint64_t* acc_full = ta_empty + TA;
    uint64_t* acc_empty = acc_full + 2;
    uint64_t* small_empty = acc_empty + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(small_empty + 1);

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const int rank = static_cast<int>(cluster_ctarank());
    const bool leader = rank == 0;
    const int tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    const int pair = static_cast<int>(blockIdx.x >> 1);
    const int m0 = (2 * (pair / tiles_n) + rank) * Cfg::BLOCK_M, n0 = (pair % tiles_n) * Cfg::BLOCK_N;
    const int kb_begin = static_cast<int>(blockIdx.y) * kb_per_split;
    const int total_kb = max(min(p.kblocks0, kb_begin + kb_per_split) - kb_begin, 0);
    const int n_chunks = (total_kb + promo - 1) / promo;
    auto tile_a = [&](int s) { return tiles + s * Cfg::STAGE_BYTES; };
    auto tile_b = [&](int s, int which) { return tiles + s * Cfg::STAGE_BYTES + Cfg::TILE_A + which * Cfg::HALF_B; };
    auto leader_addr = [&](uint64_t* bar) { return mapa_cluster(smem_u32(bar), 0u); };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmBp);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&a_full[s], 1);
            mbar_init(&b_full[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int j = 0; j < TA; ++j) {
            mbar_init(&split_bar[j], 256);
            mbar_init(&ta_empty[j], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 2 * 32 * DRAIN_WARPS);
        }
        mbar_init(small_empty, 2 * 32 * DRAIN_WARPS);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc_2sm<Cfg::TMEM_COLS>(tmem_slot);
    tc_fence_before_sync();
    cluster_sync_all();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int it = 0; it < total_kb; ++it) {
                const int s = it % STAGES;
                const uint32_t ph = (it / STAGES) & 1;
                const int r0 = (kb_begin + it) * Cfg::BLOCK_K;
                mbar_wait(&empty_bar[s], ph ^ 1);
                mbar_expect_tx(&a_full[s], Cfg::TILE_A);
                tma_load_2d(tile_a(s), &tmA, &a_full[s], m0, r0);                      // plain box {128 columns, 32 reduction rows}
                if (leader) mbar_expect_tx(&b_full[s], 4 * Cfg::HALF_B);
                const uint32_t bf = leader_addr(&b_full[s]);
                const int nh = n0 + rank * 64;
                tma_load_2d_2sm(tile_b(s, 0), &tmB, bf, nh, r0);                       // B_hi: my two {32 columns, 32 rows} atoms
                tma_load_2d_2sm(tile_b(s, 0) + 4096, &tmB, bf, nh + 32, r0);
                tma_load_2d_2sm(tile_b(s, 1), &tmBp, bf, nh, 2 * r0);                  // packed bf16: my {64 columns, 64 rows} atom
            }
        }
    } else if (warp == 1) {
        if (leader) {
            constexpr uint32_t idesc = umma_idesc(2, 2 * Cfg::BLOCK_M, Cfg::BLOCK_N, 0, 1);      // tf32, B MN-major
            constexpr uint32_t idesc_c = umma_idesc(1, 2 * Cfg::BLOCK_M, Cfg::BLOCK_N, 0, 1);    // bf16
            const uint32_t d_small = tmem_base + 256;
            int it = 0, kk = 0;
            for (int c = 0; c < n_chunks; ++c) {
                const int buf = c & 1;
                mbar_wait(&acc_empty[buf], ((c >> 1) & 1) ^ 1);
                tc_fence_after_sync();
                const uint32_t d_main = tmem_base + buf * 128;
                for (int i = 0; i < promo && kk < total_kb; ++i, ++kk, ++it) {
                    const int s = it % STAGES, j = it % TA;
                    mbar_wait(&b_full[s], (it / STAGES) & 1);
                    mbar_wait(&split_bar[j], (it / TA) & 1);
                    tc_fence_after_sync();
                    const uint32_t ta_hi = tmem_base + Cfg::TA_COL0 + j * Cfg::TA_COLS, ta_pk = ta_hi + 32;
                    const uint64_t b_desc = umma_desc_mnmajor_sw128(smem_u32(tile_b(s, 0)), 4096, 512, 1);
                    const uint64_t bp_desc = umma_desc_mnmajor_sw128(smem_u32(tile_b(s, 1)), 8192, 1024, 2);
                    if (elect_one_sync()) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            umma_ts_2sm<true>(d_main, ta_hi + 8 * k, b_desc + static_cast<uint64_t>(k * 64), idesc, (i | k) != 0);
                            umma_ts_2sm<false>(d_small, ta_pk + 8 * k, bp_desc + static_cast<uint64_t>(k * 128), idesc_c, (kk | k) != 0);
                        }
                        umma_commit_2sm_multicast(&empty_bar[s], 3);
                        umma_commit_2sm_multicast(&ta_empty[j], 3);
                    }
                    __syncwarp();
                }
                if (elect_one_sync()) umma_commit_2sm_multicast(&acc_full[buf], 3);
                __syncwarp();
            }
        }
    } else if (warp < 6) {
        const int quarter = warp & 3, row = quarter * 32 + lane;
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + Cfg::TA_COL0;
        const uint32_t split_remote0 = leader_addr(&split_bar[0]), split_remote1 = leader_addr(&split_bar[1]);
        for (int it = 0; it < total_kb; ++it) {
            const int s = it % STAGES, j = it % TA;
            mbar_wait(&a_full[s], (it / STAGES) & 1);
            uint32_t hi[32], pk[32];
            // [32 reduction rows][128 columns] plain: lane m reads column m (conflict-free LDS.32)
            const uint32_t* colp = reinterpret_cast<const uint32_t*>(tile_a(s)) + row;
            float lo[32];
#pragma unroll
            for (int r = 0; r < 32; ++r) {
                const uint32_t v = colp[r * 128];
                hi[r] = tf32_rn_bits(v);
                lo[r] = __uint_as_float(v) - __uint_as_float(hi[r]);
            }
#pragma unroll
            for (int q = 0; q < 16; ++q) {
                pk[q] = pack_bf16x2_2sm(lo[2 * q], lo[2 * q + 1]);
                pk[16 + q] = pack_bf16x2_2sm(__uint_as_float(hi[2 * q]), __uint_as_float(hi[2 * q + 1]));
            }
            mbar_wait(&ta_empty[j], ((it / TA) & 1) ^ 1);
            tc_fence_after_sync();
            tmem_st_32x32(lane_addr + j * Cfg::TA_COLS, hi);
            tmem_st_32x32(lane_addr + j * Cfg::TA_COLS + 32, pk);
            tmem_st_wait();
            tc_fence_before_sync();
            mbar_arrive_cluster_relaxed(j ? split_remote1 : split_remote0);
        }
    } else {
        const int quarter = warp & 3, half = (warp - 6) >> 2;
        const int colbase = half * X3Cfg::DRAIN_COLS;
        const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + colbase;
        float* my_buf = epi_buf + (warp - 6) * 32 * X3Cfg::EPI_PITCH;
        const uint32_t acc_empty_remote0 = leader_addr(&acc_empty[0]), acc_empty_remote1 = leader_addr(&acc_empty[1]);
        float acc[X3Cfg::DRAIN_COLS];
#pragma unroll
        for (int q = 0; q < X3Cfg::DRAIN_COLS; ++q) acc[q] = 0.0f;
        for (int c = 0; c < n_chunks; ++c) {
            const int buf = c & 1;
            mbar_wait_parked(&acc_full[buf], (c >> 1) & 1);
            tc_fence_after_sync();
            uint32_t r0[32], r1[32];
            tmem_ld_32x32(lane_base + buf * 128, r0);
            tmem_ld_32x32(lane_base + buf * 128 + 32, r1);
            tmem_ld_wait();
            tc_fence_before_sync();
            mbar_arrive_cluster_relaxed(buf ? acc_empty_remote1 : acc_empty_remote0);
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
#pragma unroll
            for (int q = 0; q < 32; ++q) acc[q] += __uint_as_float(r0[q]);
#pragma unroll
            for (int q = 0; q < 32; ++q) acc[32 + q] += __uint_as_float(r1[q]);
        }
        // partial tile of split blockIdx.y -> its slab of the workspace (transposed: slab is [N, M], a warp store writes 32
        // consecutive m of one column n)
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
    }
    tc_fence_before_sync();
    cluster_sync_all();
    if (warp == 1) tmem_dealloc_2sm<Cfg::TMEM_COLS>(tmem_base);
}

// bf16 NT products with wide outputs (N > 128) on CTA pairs: gemm_nt_bf16_persistent_kernel (gemm_tcgen05.cu) with one
// MMA of 256 x 256 x 16 over two CTAs.  Each CTA loads ITS 128 rows of A and HALF (128 rows) of the B tile: 32 KB instead
// of 48 KB per K-block through the shared-memory port, and the tensor core reads 32 instead of 48 KB -- the single-CTA
// kernel needs 768 port cycles per K-block for 512 cycles of MMA.  Two 256-column accumulators per CTA as before.
struct NtP2Cfg {
    static constexpr int BLOCK_M = 128, BLOCK_N = 256, BLOCK_K = 64, UMMA_K = 16;
    static constexpr int A_BYTES = 128 * 128, BH_BYTES = 128 * 128, STAGE_BYTES = A_BYTES + BH_BYTES;
    static constexpr int STAGES = 6;
    static constexpr int DRAIN_WARPS = 8;
    static constexpr int EPI_BYTES = DRAIN_WARPS * 32 * X3Cfg::EPI_PITCH * 4;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 1024 + 256;
    static constexpr int TMEM_COLS = 512;
    static constexpr int THREADS = 64 + 32 * DRAIN_WARPS;
};
static_assert(NtP2Cfg::SMEM_BYTES <= 227 * 1024, "bf16 NT pair kernel: shared memory budget");

__global__ void __launch_bounds__(NtP2Cfg::THREADS, 1)
gemm_nt_bf16_2sm_kernel(const GemmParams p, int n_pairs, int tiles_n, const __grid_constant__ CUtensorMap tmA0,
                        const __grid_constant__ CUtensorMap tmB0, const __grid_constant__ CUtensorMap tmA1,
                        const __grid_constant__ CUtensorMap tmB1) {
    using Cfg = NtP2Cfg;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    float* epi_buf = reinterpret_cast<float*>(tiles + STAGES * Cfg::STAGE_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(tiles + STAGES * Cfg::STAGE_BYTES + Cfg::EPI_BYTES);
    uint64_t* full_bar = bars;                     // [STAGES] (leader) TMA bytes of both CTAs landed
    uint64_t* empty_bar = bars + STAGES;           // [STAGES] (both, multicast commit)
    uint64_t* acc_full = bars + 2 * STAGES;        // [2] (both, multicast commit)
    uint64_t* acc_empty = acc_full + 2;            // [2] (leader) 256 local + 256 remote arrivals
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 4);

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const int rank = static_cast<int>(cluster_ctarank());
    const bool leader = rank == 0;
    const int total_kb = p.kblocks0 + p.kblocks1;
    const int t_first = static_cast<int>(blockIdx.x >> 1), t_step = static_cast<int>(gridDim.x >> 1);
    auto stage_a = [&](int s) { return tiles + s * Cfg::STAGE_BYTES; };
    auto stage_b = [&](int s) { return tiles + s * Cfg::STAGE_BYTES + Cfg::A_BYTES; };
    auto leader_addr = [&](uint64_t* bar) { return mapa_cluster(smem_u32(bar), 0u); };

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
            mbar_init(&acc_empty[b], 2 * 32 * Cfg::DRAIN_WARPS);
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc_2sm<Cfg::TMEM_COLS>(tmem_slot);
    tc_fence_before_sync();
    cluster_sync_all();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            int it = 0;
            for (int t = t_first; t < n_pairs; t += t_step) {
                const int m0 = (2 * (t / tiles_n) + rank) * Cfg::BLOCK_M, nh = (t % tiles_n) * Cfg::BLOCK_N + rank * 128;
                for (int kb = 0; kb < total_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
                    if (leader) mbar_expect_tx(&full_bar[s], 2 * Cfg::STAGE_BYTES);      // A rows + B half of each CTA
                    const uint32_t fb = leader_addr(&full_bar[s]);
                    const bool seg1 = kb >= p.kblocks0;
                    const int k_elem = (seg1 ? kb - p.kblocks0 : kb) * Cfg::BLOCK_K;
                    tma_load_2d_2sm(stage_a(s), seg1 ? &tmA1 : &tmA0, fb, k_elem, m0);
                    tma_load_2d_2sm(stage_b(s), seg1 ? &tmB1 : &tmB0, fb, k_elem, nh);
                }
            }
        }
    } else if (warp == 1) {
        if (leader) {
            constexpr uint32_t idesc = umma_idesc(1, 2 * Cfg::BLOCK_M, Cfg::BLOCK_N, 0, 0);
            int it = 0, ct = 0;
            for (int t = t_first; t < n_pairs; t += t_step, ++ct) {
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
                            const uint64_t koff = static_cast<uint64_t>(k * 2);
                            umma_2sm<false>(d, a_desc + koff, b_desc + koff, idesc, (kb | k) != 0);
                        }
                        umma_commit_2sm_multicast(&empty_bar[s], 3);
                    }
                    __syncwarp();
                }
                if (elect_one_sync()) umma_commit_2sm_multicast(&acc_full[buf], 3);
                __syncwarp();
            }
        }
    } else {
        const int quarter = warp & 3, half = (warp - 2) >> 2;
        const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + half * 128;
        float* my_buf = epi_buf + (warp - 2) * 32 * X3Cfg::EPI_PITCH;
        const uint32_t acc_empty_remote0 = leader_addr(&acc_empty[0]), acc_empty_remote1 = leader_addr(&acc_empty[1]);
        int ct = 0;
        for (int t = t_first; t < n_pairs; t += t_step, ++ct) {
            const int m0 = (2 * (t / tiles_n) + rank) * Cfg::BLOCK_M, n0 = (t % tiles_n) * Cfg::BLOCK_N;
            const int buf = ct & 1;
            mbar_wait_parked(&acc_full[buf], (ct >> 1) & 1);
            tc_fence_after_sync();
#pragma unroll 1
            for (int pass = 0; pass < 2; ++pass) {
                uint32_t r0[32], r1[32];
                tmem_ld_32x32(lane_base + buf * Cfg::BLOCK_N + pass * 64, r0);
                tmem_ld_32x32(lane_base + buf * Cfg::BLOCK_N + pass * 64 + 32, r1);
                tmem_ld_wait();
                if (pass == 1) {
                    tc_fence_before_sync();
                    mbar_arrive_cluster_relaxed(buf ? acc_empty_remote1 : acc_empty_remote0);
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
    cluster_sync_all();
    if (warp == 1) tmem_dealloc_2sm<Cfg::TMEM_COLS>(tmem_base);
}

static int g_bc_pairs = 1;      // grl_gemm_bc_set_pairs(0 / 1): NT fp32-grade products on CTA pairs (cta_group::2); measured
                                // 1.07-1.20x the single-CTA kernel at identical bits (profiles/r02_s52_gemm_2sm_ab.jsonl)

template <bool kX3>
static int launch_2sm(const GemmParams& p, int promo, int n_pairs, const CUtensorMap& tA0, const CUtensorMap& tB0,
                      const CUtensorMap& tBp0, const CUtensorMap& tA1, const CUtensorMap& tB1, const CUtensorMap& tBp1,
                      cudaStream_t st) {
    auto kern = gemm_bc_2sm_kernel<kX3>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Bc2Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(gemm_bc_2sm_kernel)");
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(Bc2Cfg::THREADS);
    cfg.dynamicSmemBytes = Bc2Cfg::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    static int max_clusters = 0;
    if (max_clusters == 0) {
        int n = 0;
        cfg.gridDim = dim3(kNumSMs);
        cudaError_t eo = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
        if (eo != cudaSuccess) return cuda_fail(eo, "cudaOccupancyMaxActiveClusters(gemm_bc_2sm_kernel)");
        max_clusters = n < 1 ? 1 : (n > kNumSMs / 2 ? kNumSMs / 2 : n);
    }
    cfg.gridDim = dim3(2 * (n_pairs < max_clusters ? n_pairs : max_clusters));
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p, promo, n_pairs, tA0, tB0, tBp0, tA1, tB1, tBp1);
    if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(gemm_bc_2sm_kernel)");
    return 0;
}

// returns 1 when the pair kernel is switched off or the product is too small for it (the caller runs gemm_bc_kernel)
int try_launch_gemm_bc_2sm(const GemmArgs& a, GemmParams p, cudaStream_t st, bool x3, int promo) {
    using Cfg = Bc2Cfg;
    if (!g_bc_pairs) return 1;
    const int tiles_m = (p.M + Cfg::BLOCK_M - 1) / Cfg::BLOCK_M, tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    const int n_pairs = (tiles_m + 1) / 2 * tiles_n;
    if (n_pairs < kNumSMs / 2) return 1;
    CUtensorMap tA0, tB0, tBp0, tA1, tB1, tBp1;
    int rc;
    auto lo_map = [&](CUtensorMap* out, const void* ptr, int K, int64_t ldb) {
        return x3 ? make_tmap_sw128(out, ptr, 4, p.N, K, ldb, 64) : make_tmap_sw128(out, ptr, 2, p.N, 2 * ((K + 3) / 4 * 4), 2 * ldb, 64);
    };
    if ((rc = make_tmap_sw128(&tA0, a.A0, 4, p.M, a.K0, a.lda0, Cfg::BLOCK_M))) return rc;
    if ((rc = make_tmap_sw128(&tB0, a.B0, 4, p.N, a.K0, a.ldb0, 64))) return rc;
    if ((rc = lo_map(&tBp0, a.B0lo, a.K0, a.ldb0))) return rc;
    tA1 = tA0; tB1 = tB0; tBp1 = tBp0;
    if (a.K1 > 0) {
        if ((rc = make_tmap_sw128(&tA1, a.A1, 4, p.M, a.K1, a.lda1, Cfg::BLOCK_M))) return rc;
        if ((rc = make_tmap_sw128(&tB1, a.B1, 4, p.N, a.K1, a.ldb1, 64))) return rc;
        if ((rc = lo_map(&tBp1, a.B1lo, a.K1, a.ldb1))) return rc;
    }
    p.kblocks0 = (a.K0 + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K;
    p.kblocks1 = a.K1 > 0 ? (a.K1 + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K : 0;
    return x3 ? launch_2sm<true>(p, promo, n_pairs, tA0, tB0, tBp0, tA1, tB1, tBp1, st)
              : launch_2sm<false>(p, promo, n_pairs, tA0, tB0, tBp0, tA1, tB1, tBp1, st);
}

// returns 1 when the pair kernel is switched off or the product is too small (the caller runs the single-CTA kernel)
int try_launch_gemm_nt_bf16_2sm(const GemmArgs& a, GemmParams p, cudaStream_t st) {
    using Cfg = NtP2Cfg;
    if (!g_bc_pairs) return 1;
    const int tiles_m = (p.M + Cfg::BLOCK_M - 1) / Cfg::BLOCK_M, tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    const int n_pairs = (tiles_m + 1) / 2 * tiles_n;
    if (n_pairs < kNumSMs / 2) return 1;
    CUtensorMap tA0, tB0, tA1, tB1;
    int rc;
    if ((rc = make_tmap_sw128(&tA0, a.A0, 2, p.M, a.K0, a.lda0, Cfg::BLOCK_M))) return rc;
    if ((rc = make_tmap_sw128(&tB0, a.B0, 2, p.N, a.K0, a.ldb0, 128))) return rc;
    tA1 = tA0;
    tB1 = tB0;
    if (a.K1 > 0) {
        if ((rc = make_tmap_sw128(&tA1, a.A1, 2, p.M, a.K1, a.lda1, Cfg::BLOCK_M))) return rc;
        if ((rc = make_tmap_sw128(&tB1, a.B1, 2, p.N, a.K1, a.ldb1, 128))) return rc;
    }
    p.kblocks0 = (a.K0 + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K;
    p.kblocks1 = a.K1 > 0 ? (a.K1 + Cfg::BLOCK_K - 1) / Cfg::BLOCK_K : 0;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm_nt_bf16_2sm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(gemm_nt_bf16_2sm_kernel)");
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
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
    static int max_clusters = 0;
    if (max_clusters == 0) {
        int n = 0;
        cfg.gridDim = dim3(kNumSMs);
        cudaError_t eo = cudaOccupancyMaxActiveClusters(&n, gemm_nt_bf16_2sm_kernel, &cfg);
        if (eo != cudaSuccess) return cuda_fail(eo, "cudaOccupancyMaxActiveClusters(gemm_nt_bf16_2sm_kernel)");
        max_clusters = n < 1 ? 1 : (n > kNumSMs / 2 ? kNumSMs / 2 : n);
    }
    cfg.gridDim = dim3(2 * (n_pairs < max_clusters ? n_pairs : max_clusters));
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_nt_bf16_2sm_kernel, p, n_pairs, tiles_n, tA0, tB0, tA1, tB1);
    if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(gemm_nt_bf16_2sm_kernel)");
    return 0;
}

// returns 1 when the pair kernel is switched off (the caller launches gemm_bc_kernel<TN>); otherwise launches the split-K
// partial products into `p.C` (the workspace) and reports the split count it used
int try_launch_gemm_bc_tn_2sm(GemmParams p, int promo, int transpose_out, const CUtensorMap& tA, const CUtensorMap& tB,
                              const CUtensorMap& tBp, int* splits_out, size_t workspace_bytes, cudaStream_t st) {
    using Cfg = Bc2Cfg;
    if (!g_bc_pairs) return 1;
    const int tiles_m = (p.M + Cfg::BLOCK_M - 1) / Cfg::BLOCK_M, tiles_n = (p.N + Cfg::BLOCK_N - 1) / Cfg::BLOCK_N;
    const int n_pairs = (tiles_m + 1) / 2 * tiles_n;
    int splits = tn_plan_splits(2 * n_pairs, p.kblocks0);
    const int kb_per_split = (p.kblocks0 + splits - 1) / splits;
    splits = (p.kblocks0 + kb_per_split - 1) / kb_per_split;
    const size_t need = sizeof(float) * static_cast<size_t>(p.M) * p.N * splits;
    GRL_CHECK_ARG(workspace_bytes >= need, GRL_E_WORKSPACE, "grl_gemm_tn_bc: workspace %zu < %zu", workspace_bytes, need);
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gemm_bc_tn_2sm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(gemm_bc_tn_2sm_kernel)");
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * n_pairs, splits);
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
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_bc_tn_2sm_kernel, p, promo, kb_per_split, transpose_out, tA, tB, tBp);
    if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(gemm_bc_tn_2sm_kernel)");
    *splits_out = splits;
    return 0;
}

}  // namespace grl

using namespace grl;

extern "C" int grl_gemm_bc_set_pairs(int on) {
    g_bc_pairs = on != 0;
    return 0;
}
