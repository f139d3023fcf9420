// Fused GraphConv forward: aggregation producer -> tcgen05 projection, new_V never reaches HBM.
//
// Reference: GraphConv.forward, gnn/models/networks/robust_gcn.py:45-51
//     new_V = matmul(A', V)                  [B, (L+1)N, F]   (A' DropEdge-masked in drop_robust_gcn.py:76,80,85)
//     out   = matmul(new_V.view(B, N, (L+1)F), h_weights) + bias        K = (L+1) F
// The unfused path (grl_spmm_fwd_stream + grl_gemm_nt) writes new_V ((L+1) x the input, 0.92 GB per layer at cfg2) and
// reads it back.  Here one CTA cluster owns a 128-row node tile of ONE graph; its K loop runs over
// (input, 32-column chunk kc, edge type t), and for every K-block
//   * 8 gather warps aggregate the 128 x 32 fp32 tile  sum_e w[e] X[col[e], kc*32 .. +32]  of type t in CSR order (the
//     order of every SpMM kernel here) -- 8 lanes x 16 bytes per row, rows taken from a shared-memory slab
//     [N, 32 columns] of the graph staged by TMA once per (tile, kc) (kSlab), or straight from global / L2 when the
//     graph is too large for a slab -- and write it into the K-major SWIZZLE_128B operand tile;
//   * the splitter / MMA / drain pipeline of gemm_bc_kernel<NT, x3> (gemm_bc_tcgen05.cu) takes it from there: hi / lo
//     split in registers -> TMEM operand ring -> three tf32 products against the weight tiles (TMA) -> promoted fp32
//     accumulation -> bias / ReLU / dropout epilogue from registers.
// Output columns: a cluster of cs = ceil(C / 128) CTAs shares the node tile, CTA `rank` owns columns [128 rank, +128).
// The aggregation is done ONCE per cluster: CTA rank aggregates rows [rank 128/cs, +128/cs) and stores them into the
// operand tile of every CTA of the cluster with st.async: each 16-byte store signals complete_tx on the destination's
// `agg` barrier when it lands (no fence, no cluster-scope release / acquire -- a first version with st.shared::cluster +
// fence.acq_rel.cluster + release-arrives spent its time in MEMBAR / CCTL.IVALL, which also flushed the cached edge
// lists out of L1 every K-block).  An operand slot is reused when the splitters of ALL CTAs have read it (remote
// arrives on `a_free`).
// Register budget by role (setmaxnreg per warpgroup, 896 threads launched at 72): TMA / MMA / slab warps 24, splitter 88,
// drain 200 (FOUR warps holding 128 promoted columns each: at the pace the producer sets, half the drain warps of the
// plain GEMM are enough, and their registers go to the producer), SIXTEEN gather warps at 48.  Measured steps of the
// producer (cfg2 gcn1 layer): 8 warps, one row per lane group 1.27 ms -> two rows in flight (72 registers) 1.23 ms ->
// 16 warps, one row 1.15 ms (direct producer 2.58 -> 1.91 -> 1.47 ms).
//
// Arithmetic = the unfused path's (CSR-order fp32 aggregation, three-product tf32 GEMM with promotion), but the K-blocks
// are visited in (kc, t) instead of (t, kc) order: results agree to fp32 rounding (~1e-7), not bit for bit.
#include "gemm_x3_parts.cuh"

namespace grl {

int gemm_x3_current_promo();    // gemm_x3_tcgen05.cu

struct FusedCfg {
    static constexpr int BLOCK_M = 128, BLOCK_N = 128, BLOCK_K = 32;
    static constexpr int TILE_BYTES = 128 * 128;
    static constexpr int B_STAGE_BYTES = 2 * TILE_BYTES;            // B_hi | B_lo
    static constexpr int MAX_A = 3, MAX_B = 4;
    static constexpr int TA_STAGES = 2, TA_COL0 = 384, TA_COLS = 64;
    static constexpr int EPI_BYTES = 4 * 32 * X3Cfg::EPI_PITCH * 4;   // one 32 x 16 transposition buffer per drain warp
    static constexpr int BAR_BYTES = 256;
    static constexpr int GATHER_WARPS = 16;
    static constexpr int THREADS = 896;                             // 7 warpgroups: control | splitter | drain | gather x4
    static constexpr int TMEM_COLS = 512;
    static constexpr int DRAIN_WARPS = 4;                           // one per TMEM lane quarter, all 128 columns
    static constexpr int GATHER_WARP0 = 12;
};

struct FusedParams {
    GemmParams g;                       // epilogue: M = B*N, N = C, output, bias, flags, dropout
    const int32_t* rowptr;              // [B*N*(L+1) + 1]
    const int32_t* col;                 // [nnz] local node index
    const float* w;                     // [nnz] edge weights (val, or the DropEdge weights of the layer)
    const float* X0;                    // [B*N, F0]
    const float* X1;                    // [B*N, F1] or null
    int64_t ldx0, ldx1;
    int F0, F1;                         // multiples of 32
    int B, N, Lp1;
    int tiles_per_graph, n_mtiles;
    int promo, n_a, n_b;
    int box_rows, n_boxes;              // slab: n_boxes TMA boxes of box_rows rows x 32 columns
    uint32_t off_b, off_epi, off_rp, off_slab, off_bars;
};

struct Ring {
    int s = 0, n;
    uint32_t ph = 0;
    __device__ explicit Ring(int n_) : n(n_) {}
    __device__ __forceinline__ void next() {
        if (++s == n) { s = 0; ph ^= 1u; }
    }
};

template <bool kSlab>
__global__ void __launch_bounds__(FusedCfg::THREADS, 1)
graphconv_fused_kernel(const FusedParams p, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmBlo,
                       const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1) {
    using Cfg = FusedCfg;
    constexpr int TA = Cfg::TA_STAGES, DRAIN_WARPS = Cfg::DRAIN_WARPS;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    // [A ring: n_a x 16 KB][B ring: n_b x 32 KB][epilogue buffers][row pointers][slab][barriers]
    float* epi_buf = reinterpret_cast<float*>(tiles + p.off_epi);
    int32_t* rp_s = reinterpret_cast<int32_t*>(tiles + p.off_rp);
    uint64_t* bars = reinterpret_cast<uint64_t*>(tiles + p.off_bars);
    uint64_t* full_bar = bars;                      // [MAX_B] weight tiles landed
    uint64_t* empty_bar = bars + 4;                 // [MAX_B] MMAs reading the weight stage retired
    uint64_t* agg_bar = bars + 8;                   // [MAX_A] aggregated tile written by every CTA of the cluster
    uint64_t* a_free = bars + 11;                   // [MAX_A] the splitters of every CTA have read the tile
    uint64_t* split_bar = bars + 14;                // [TA]
    uint64_t* ta_empty = bars + 16;                 // [TA]
    uint64_t* acc_full = bars + 18;                 // [2]
    uint64_t* acc_empty = bars + 20;                // [2]
    uint64_t* small_empty = bars + 22;
    uint64_t* slab_full = bars + 23;
    uint64_t* slab_empty = bars + 24;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 25);

    const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
    const int lane = threadIdx.x & 31;
    const int cs = static_cast<int>(cluster_nctarank());
    const int rank = static_cast<int>(cluster_ctarank());
    const int t_first = static_cast<int>(blockIdx.x) / cs, t_step = static_cast<int>(gridDim.x) / cs;
    const int n0 = rank * Cfg::BLOCK_N;
    const int rows_share = Cfg::BLOCK_M / cs;
    const int Lp1 = p.Lp1, rp_pitch = Lp1 + 1;
    const int F_total = p.F0 + p.F1;
    const int total_kb = Lp1 * (F_total / Cfg::BLOCK_K);
    const int n_chunks = (total_kb + p.promo - 1) / p.promo;

    auto tile_a = [&](int s) { return tiles + s * Cfg::TILE_BYTES; };
    auto tile_b = [&](int s, int which) { return tiles + p.off_b + s * Cfg::B_STAGE_BYTES + which * Cfg::TILE_BYTES; };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmBlo);
        if (kSlab) {
            tma_prefetch_desc(&tmX0);
            if (p.F1) tma_prefetch_desc(&tmX1);
        }
        for (int s = 0; s < Cfg::MAX_B; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < Cfg::MAX_A; ++s) {
            mbar_init(&agg_bar[s], 1);                      // one local expect_tx(16 KB); the bytes come from every CTA
            mbar_init(&a_free[s], 4 * cs);
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
        mbar_init(slab_full, 1);
        mbar_init(slab_empty, Cfg::GATHER_WARPS);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    tc_fence_before_sync();
    cluster_sync_all();                     // every CTA's barriers exist before anything arrives on them remotely
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < 4) {
        setmaxnreg_dec<24>();
        if (warp == 0) {
            // ================================================================= weight tiles (one lane)
            if (lane == 0) {
                Ring rb(p.n_b);
                for (int t = t_first; t < p.n_mtiles; t += t_step) {
                    for (int seg = 0; seg < 2; ++seg) {
                        const int Fk = seg ? p.F1 : p.F0, off_k = seg ? p.F0 : 0;
                        for (int kc = 0; kc < Fk / Cfg::BLOCK_K; ++kc) {
                            for (int ty = 0; ty < Lp1; ++ty) {
                                const int k_elem = ty * F_total + off_k + kc * Cfg::BLOCK_K;
                                mbar_wait(&empty_bar[rb.s], rb.ph ^ 1u);
                                mbar_expect_tx(&full_bar[rb.s], 2 * Cfg::TILE_BYTES);
                                tma_load_2d(tile_b(rb.s, 0), &tmB, &full_bar[rb.s], k_elem, n0);
                                tma_load_2d(tile_b(rb.s, 1), &tmBlo, &full_bar[rb.s], k_elem, n0);
                                rb.next();
                            }
                        }
                    }
                }
            }
        } else if (warp == 1) {
            // ================================================================= MMA issuer (warp-wide loop, elected lane issues)
            constexpr uint32_t idesc = umma_idesc(2, Cfg::BLOCK_M, Cfg::BLOCK_N, 0, 0);     // tf32, A from TMEM, B K-major
            const uint32_t d_small = tmem_base + 256;
            Ring rb(p.n_b);
            int it = 0, cg = 0, tl = 0;
            for (int t = t_first; t < p.n_mtiles; t += t_step, ++tl) {
                mbar_wait(small_empty, (tl & 1) ^ 1);
                tc_fence_after_sync();
                int kk = 0;
                for (int c = 0; c < n_chunks; ++c, ++cg) {
                    const int buf = cg & 1;
                    mbar_wait(&acc_empty[buf], ((cg >> 1) & 1) ^ 1);
                    tc_fence_after_sync();
                    const uint32_t d_main = tmem_base + buf * 128;
                    for (int i = 0; i < p.promo && kk < total_kb; ++i, ++kk, ++it) {
                        const int j = it % TA;
                        mbar_wait(&full_bar[rb.s], rb.ph);
                        mbar_wait(&split_bar[j], (it / TA) & 1);
                        tc_fence_after_sync();
                        const uint32_t ta_hi = tmem_base + Cfg::TA_COL0 + j * Cfg::TA_COLS, ta_lo = ta_hi + 32;
                        const uint64_t b_desc = umma_desc_kmajor_sw128(smem_u32(tile_b(rb.s, 0)));
                        const uint64_t bl_desc = umma_desc_kmajor_sw128(smem_u32(tile_b(rb.s, 1)));
                        if (elect_one_sync()) {
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const uint64_t koff = static_cast<uint64_t>(k * 2);
                                umma_ts<true>(d_main, ta_hi + 8 * k, b_desc + koff, idesc, (i | k) != 0);
                                umma_ts<true>(d_small, ta_lo + 8 * k, b_desc + koff, idesc, (kk | k) != 0);    // a_lo * b_hi
                                umma_ts<true>(d_small, ta_hi + 8 * k, bl_desc + koff, idesc, 1u);              // a_hi * b_lo
                            }
                            umma_commit(&empty_bar[rb.s]);
                            umma_commit(&ta_empty[j]);
                        }
                        __syncwarp();
                        rb.next();
                    }
                    if (elect_one_sync()) umma_commit(&acc_full[buf]);
                    __syncwarp();
                }
            }
        } else if (warp == 2) {
            // ================================================================= feature slab of the tile's graph (one lane)
            if (kSlab && lane == 0) {
                uint8_t* slab = tiles + p.off_slab;
                uint32_t ph = 0;
                for (int t = t_first; t < p.n_mtiles; t += t_step) {
                    const int b = t / p.tiles_per_graph;
                    for (int seg = 0; seg < 2; ++seg) {
                        const int Fk = seg ? p.F1 : p.F0;
                        for (int kc = 0; kc < Fk / Cfg::BLOCK_K; ++kc) {
                            mbar_wait(slab_empty, ph ^ 1u);
                            mbar_expect_tx(slab_full, static_cast<uint32_t>(p.n_boxes) * p.box_rows * 128u);
                            for (int q = 0; q < p.n_boxes; ++q)
                                tma_load_2d(slab + static_cast<size_t>(q) * p.box_rows * 128, seg ? &tmX1 : &tmX0, slab_full,
                                            kc * Cfg::BLOCK_K, b * p.N + q * p.box_rows);
                            ph ^= 1u;
                        }
                    }
                }
            }
        }
    } else if (warp < 8) {
        // ===================================================================== operand splitter (warps 4..7)
        setmaxnreg_inc<88>();
        const int quarter = warp & 3, row = quarter * 32 + lane;
        const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + Cfg::TA_COL0;
        Ring ra(p.n_a);
        int it = 0;
        for (int t = t_first; t < p.n_mtiles; t += t_step) {
            for (int kk = 0; kk < total_kb; ++kk, ++it) {
                const int j = it % TA;
                mbar_wait(&agg_bar[ra.s], ra.ph);
                uint32_t hi[32], lo[32];
                const uint8_t* rowp = tile_a(ra.s) + row * 128;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const uint4 v = *reinterpret_cast<const uint4*>(rowp + ((c ^ (row & 7)) << 4));
                    const uint32_t wv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        hi[4 * c + e] = tf32_rn_bits(wv[e]);
                        lo[4 * c + e] = tf32_rn_bits(__float_as_uint(__uint_as_float(wv[e]) - __uint_as_float(hi[4 * c + e])));
                    }
                }
                mbar_wait(&ta_empty[j], ((it / TA) & 1) ^ 1);
                tc_fence_after_sync();
                tmem_st_32x32(lane_addr + j * Cfg::TA_COLS, hi);
                tmem_st_32x32(lane_addr + j * Cfg::TA_COLS + 32, lo);
                tmem_st_wait();
                tc_fence_before_sync();
                mbar_arrive(&split_bar[j]);
                // the shared-memory tile is free once every splitter warp of every CTA has read it
                __syncwarp();
                if (lane == 0) {
                    const uint32_t fb = smem_u32(&a_free[ra.s]);
                    for (int r = 0; r < cs; ++r) mbar_arrive_cluster_relaxed(mapa_cluster(fb, r));
                }
                ra.next();
            }
        }
    } else if (warp < 12) {
        // ===================================================================== drain + epilogue (warps 8..11)
        setmaxnreg_inc<200>();
        const int quarter = warp & 3;
        const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
        float* my_buf = epi_buf + (warp - 8) * 32 * X3Cfg::EPI_PITCH;
        int cg = 0;
        for (int t = t_first; t < p.n_mtiles; t += t_step) {
            const int b = t / p.tiles_per_graph, q = t - b * p.tiles_per_graph;
            const int m0 = b * p.N + q * Cfg::BLOCK_M;
            float acc0[64], acc1[64];               // columns [0, 64) and [64, 128) of this lane's row
#pragma unroll
            for (int x = 0; x < 64; ++x) { acc0[x] = 0.0f; acc1[x] = 0.0f; }
            auto add_from = [&](uint32_t col0, bool last_of_buffer, uint64_t* release) {
                uint32_t r0[32], r1[32];
                tmem_ld_32x32(lane_base + col0, r0);
                tmem_ld_32x32(lane_base + col0 + 32, r1);
                tmem_ld_wait();
#pragma unroll
                for (int x = 0; x < 32; ++x) acc0[x] += __uint_as_float(r0[x]);
#pragma unroll
                for (int x = 0; x < 32; ++x) acc0[32 + x] += __uint_as_float(r1[x]);
                tmem_ld_32x32(lane_base + col0 + 64, r0);
                tmem_ld_32x32(lane_base + col0 + 96, r1);
                tmem_ld_wait();
                if (last_of_buffer) {
                    tc_fence_before_sync();
                    mbar_arrive(release);
                }
#pragma unroll
                for (int x = 0; x < 32; ++x) acc1[x] += __uint_as_float(r0[x]);
#pragma unroll
                for (int x = 0; x < 32; ++x) acc1[32 + x] += __uint_as_float(r1[x]);
            };
            for (int c = 0; c < n_chunks; ++c, ++cg) {
                const int buf = cg & 1;
                mbar_wait(&acc_full[buf], (cg >> 1) & 1);
                tc_fence_after_sync();
                add_from(buf * 128, true, &acc_empty[buf]);
            }
            add_from(256, true, small_empty);
            GemmParams gq = p.g;
            gq.M = b * p.N + min(p.N, (q + 1) * Cfg::BLOCK_M);       // rows of THIS graph only
            if (m0 + quarter * 32 < gq.M) {
                if (n0 < gq.N) epilogue_from_regs(gq, acc0, my_buf, m0 + quarter * 32, n0, lane);
                if (n0 + 64 < gq.N) epilogue_from_regs(gq, acc1, my_buf, m0 + quarter * 32, n0 + 64, lane);
            }
        }
    } else {
        // ===================================================================== aggregation producer (warps 12..19)
        setmaxnreg_dec<48>();
        const int gw = warp - Cfg::GATHER_WARP0, grp = lane >> 3, l8 = lane & 7;
        const int gtid = static_cast<int>(threadIdx.x) - Cfg::GATHER_WARP0 * 32;
        const uint32_t slab_lane = base_u32 + p.off_slab + l8 * 16;
        Ring ra(p.n_a);
        uint32_t slab_ph = 0;
        uint32_t peer_base[4];              // shared::cluster address of `tiles` in every CTA of the cluster (cs is 1, 2 or 4)
#pragma unroll
        for (int r = 0; r < 4; ++r) peer_base[r] = mapa_cluster(base_u32, static_cast<uint32_t>(r < cs ? r : 0));
        // Work units of one (input, 32-column chunk): (edge type ty, pass) -- a lane group of 8 aggregates row
        // lr = gw*4 + grp + 64*pass of my share for type ty; a K-block is n_pass consecutive units (cs = 4: half of the
        // gather warps have no row).  The edge list of a unit (<= 16 entries: lane l8 holds entries l8 and l8 + 8) is
        // fetched ONE UNIT AHEAD and broadcast inside the group with shuffles, so the latency of col / w never sits in
        // front of the gather; longer typed rows finish in a plain loop.
        constexpr int ROWS_PER_PASS = Cfg::GATHER_WARPS * 4;
        const int n_pass = rows_share > ROWS_PER_PASS ? rows_share / ROWS_PER_PASS : 1, pass_shift = n_pass == 2 ? 1 : 0;
        const int units = Lp1 * n_pass;
        struct Pre { int e0, cnt, c0, c1; float w0, w1; };
        auto prefetch = [&](int u) -> Pre {
            const int ty = u >> pass_shift, lr = gw * 4 + grp + ROWS_PER_PASS * (u & (n_pass - 1));
            Pre x;
            x.e0 = 0; x.cnt = 0; x.c0 = 0; x.c1 = 0; x.w0 = 0.0f; x.w1 = 0.0f;
            if (lr < rows_share) {
                x.e0 = rp_s[lr * rp_pitch + ty];
                x.cnt = rp_s[lr * rp_pitch + ty + 1] - x.e0;
                if (l8 < x.cnt) { x.c0 = __ldg(p.col + x.e0 + l8); x.w0 = __ldg(p.w + x.e0 + l8); }
                if (l8 + 8 < x.cnt) { x.c1 = __ldg(p.col + x.e0 + l8 + 8); x.w1 = __ldg(p.w + x.e0 + l8 + 8); }
            }
            return x;
        };
        for (int t = t_first; t < p.n_mtiles; t += t_step) {
            const int b = t / p.tiles_per_graph, q = t - b * p.tiles_per_graph;
            const int n_valid = min(Cfg::BLOCK_M, p.N - q * Cfg::BLOCK_M);
            const int64_t node0 = static_cast<int64_t>(b) * p.N + q * Cfg::BLOCK_M;
            // row pointers of my share of the tile's rows (rows past the graph: empty)
            named_bar_sync(1, Cfg::GATHER_WARPS * 32);
            for (int idx = gtid; idx < rows_share * rp_pitch; idx += Cfg::GATHER_WARPS * 32) {
                const int lr = idx / rp_pitch, k = idx - lr * rp_pitch;
                const int row = rank * rows_share + lr;
                rp_s[idx] = row < n_valid ? __ldg(p.rowptr + (node0 + row) * Lp1 + k) : 0;
            }
            named_bar_sync(1, Cfg::GATHER_WARPS * 32);
            Pre nxt = prefetch(0);
            for (int seg = 0; seg < 2; ++seg) {
                const int Fk = seg ? p.F1 : p.F0;
                const float* Xg = (seg ? p.X1 : p.X0) + static_cast<int64_t>(b) * p.N * (seg ? p.ldx1 : p.ldx0) + l8 * 4;
                const int64_t ldx = seg ? p.ldx1 : p.ldx0;
                for (int kc = 0; kc < Fk / Cfg::BLOCK_K; ++kc) {
                    if (kSlab) {
                        mbar_wait(slab_full, slab_ph);
                        slab_ph ^= 1u;
                    }
                    auto fetch = [&](int c) -> float4 {
                        if (kSlab) {
                            float4 v;
                            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                                         : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                                         : "r"(slab_lane + static_cast<uint32_t>(c) * 128u));
                            return v;
                        }
                        return __ldg(reinterpret_cast<const float4*>(Xg + static_cast<int64_t>(c) * ldx + kc * Cfg::BLOCK_K));
                    };
                    auto fma4 = [](float4& a, float w, const float4& v) {
                        a.x = fmaf(w, v.x, a.x); a.y = fmaf(w, v.y, a.y); a.z = fmaf(w, v.z, a.z); a.w = fmaf(w, v.w, a.w);
                    };
                    for (int u = 0; u < units; ++u) {
                        const Pre cur = nxt;
                        nxt = prefetch(u + 1 == units ? 0 : u + 1);        // wraps to the next chunk's first unit (same rows)
                        const int pass = u & (n_pass - 1);
                        if (pass == 0) {
                            mbar_wait(&a_free[ra.s], ra.ph ^ 1u);
                            if (gw == 0 && lane == 0) mbar_expect_tx(&agg_bar[ra.s], Cfg::TILE_BYTES);
                        }
                        float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                        const int kmax = __reduce_max_sync(0xffffffffu, min(cur.cnt, 16));
                        const int src0 = lane & 24;
#pragma unroll 1
                        for (int k = 0; k < kmax; k += 4) {       // entries past the row's end carry w = 0 (col = 0: a valid row)
                            int cc[4];
                            float ww[4];
                            float4 vv[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const int kk = k + j, src = src0 | (kk & 7);
                                cc[j] = __shfl_sync(0xffffffffu, kk < 8 ? cur.c0 : cur.c1, src);
                                ww[j] = __shfl_sync(0xffffffffu, kk < 8 ? cur.w0 : cur.w1, src);
                            }
#pragma unroll
                            for (int j = 0; j < 4; ++j) vv[j] = fetch(cc[j]);
#pragma unroll
                            for (int j = 0; j < 4; ++j) fma4(acc, ww[j], vv[j]);
                        }
                        for (int e = cur.e0 + 16; e < cur.e0 + cur.cnt; ++e)       // typed rows longer than 16 entries
                            fma4(acc, __ldg(p.w + e), fetch(__ldg(p.col + e)));
                        const int lr = gw * 4 + grp + ROWS_PER_PASS * pass;
                        if (lr < rows_share) {
                            // K-major SWIZZLE_128B tile: 16-byte chunk c of row r sits at c ^ (r & 7)
                            const int row = rank * rows_share + lr;
                            const uint32_t dst = ra.s * Cfg::TILE_BYTES + row * 128 + ((l8 ^ (row & 7)) << 4);
                            const uint32_t ab = p.off_bars + (8 + ra.s) * 8;                 // agg_bar[ra.s]
                            st_async_v4(peer_base[0] + dst, acc, peer_base[0] + ab);
                            if (cs > 1) st_async_v4(peer_base[1] + dst, acc, peer_base[1] + ab);
                            if (cs > 2) {
                                st_async_v4(peer_base[2] + dst, acc, peer_base[2] + ab);
                                st_async_v4(peer_base[3] + dst, acc, peer_base[3] + ab);
                            }
                        }
                        if (pass == n_pass - 1) ra.next();
                    }
                    if (kSlab) {
                        __syncwarp();                           // every lane's slab reads of this chunk are in registers
                        if (lane == 0) mbar_arrive(slab_empty);
                    }
                }
            }
        }
    }
    tc_fence_before_sync();
    cluster_sync_all();             // peers may still store into / arrive on this CTA's shared memory
    if (warp == 1) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

// 0 = choose, 1 = force the slab producer (GRL_E_UNSUP when it does not fit), 2 = force the direct (global / L2) producer
static int g_fused_mode = 0;

template <bool kSlab>
static int launch_fused(const FusedParams& p, const CUtensorMap& tB, const CUtensorMap& tBlo, const CUtensorMap& tX0,
                        const CUtensorMap& tX1, int cs, size_t smem, cudaStream_t st) {
    auto kern = graphconv_fused_kernel<kSlab>;
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(graphconv_fused_kernel, smem)");
        attr_smem = smem;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(FusedCfg::THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cfg.gridDim = dim3(cs);
    // persistent grid = the clusters that are co-resident (one CTA per SM)
    static int max_clusters[5] = {0, 0, 0, 0, 0};
    static size_t max_clusters_smem[5] = {0, 0, 0, 0, 0};
    if (max_clusters[cs] == 0 || max_clusters_smem[cs] != smem) {
        int n = 0;
        cfg.gridDim = dim3(kNumSMs / cs * cs);
        cudaError_t eo = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
        if (eo != cudaSuccess) return cuda_fail(eo, "cudaOccupancyMaxActiveClusters(graphconv_fused_kernel)");
        max_clusters[cs] = n < 1 ? 1 : (n > kNumSMs / cs ? kNumSMs / cs : n);
        max_clusters_smem[cs] = smem;
    }
    const int n_clusters = p.n_mtiles < max_clusters[cs] ? p.n_mtiles : max_clusters[cs];
    cfg.gridDim = dim3(n_clusters * cs);
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, p, tB, tBlo, tX0, tX1);
    if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx(graphconv_fused_kernel)");
    return 0;
}

}  // namespace grl

using namespace grl;

extern "C" int grl_graphconv_fused_set_mode(int mode) {
    if (mode < 0 || mode > 2) {
        set_error("grl_graphconv_fused_set_mode: mode=%d not in {0 auto, 1 slab, 2 direct}", mode);
        return GRL_E_BADARG;
    }
    g_fused_mode = mode;
    return 0;
}

extern "C" int grl_graphconv_fused_fwd(const int32_t* rowptr, const int32_t* col, const float* w, const float* X0, int64_t ldx0,
                                       int F0, const float* X1, int64_t ldx1, int F1, const float* W_hi, const float* W_lo,
                                       int64_t ldw, const float* bias, float* Y, int64_t ldy, int B, int N, int L, int C,
                                       uint32_t flags, float p_drop, const uint8_t* mask, uint64_t seed, uint64_t stream_id,
                                       void* stream) {
    using Cfg = FusedCfg;
    GRL_CHECK_ARG(rowptr && col && w && X0 && W_hi && W_lo && Y, GRL_E_BADARG, "grl_graphconv_fused_fwd: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0 && L > 0 && C > 0 && F0 > 0 && F1 >= 0, GRL_E_BADARG, "grl_graphconv_fused_fwd: bad shape");
    GRL_CHECK_ARG(F1 == 0 || X1, GRL_E_BADARG, "grl_graphconv_fused_fwd: second input without a pointer");
    GRL_CHECK_ARG(static_cast<int64_t>(B) * N < (1LL << 31), GRL_E_BADARG, "grl_graphconv_fused_fwd: B*N overflows int32");
    GRL_CHECK_ARG(!(flags & GRL_EPI_BIAS) || bias, GRL_E_BADARG, "grl_graphconv_fused_fwd: BIAS flag without bias");
    GRL_CHECK_ARG(p_drop >= 0.0f && p_drop < 1.0f, GRL_E_BADARG, "grl_graphconv_fused_fwd: p outside [0,1)");
    const int F_total = F0 + F1, K = (L + 1) * F_total;
    // shapes the kernel is built for; anything else -> the caller's unfused path
    GRL_CHECK_ARG(F0 % 32 == 0 && F1 % 32 == 0 && C <= 4 * Cfg::BLOCK_N && L + 2 <= 32, GRL_E_UNSUP,
                  "grl_graphconv_fused_fwd: needs F0, F1 multiples of 32, C <= 512, L <= 30 (F0=%d F1=%d C=%d L=%d)", F0, F1, C, L);
    GRL_CHECK_ARG(ldx0 >= F0 && (F1 == 0 || ldx1 >= F1) && ldw >= K && ldy >= C, GRL_E_BADARG,
                  "grl_graphconv_fused_fwd: leading dimension too small");
    GRL_CHECK_ARG(ldx0 % 4 == 0 && (F1 == 0 || ldx1 % 4 == 0) && ldw % 4 == 0 && aligned_to(X0, 16) &&
                      (F1 == 0 || aligned_to(X1, 16)) && aligned_to(W_hi, 16) && aligned_to(W_lo, 16),
                  GRL_E_ALIGN, "grl_graphconv_fused_fwd: operands need 16-byte aligned bases and row pitches");
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    FusedParams fp{};
    if (p_drop == 0.0f) flags &= ~GRL_EPI_DROPOUT;
    fp.g.M = B * N; fp.g.N = C; fp.g.C = Y; fp.g.ldc = ldy; fp.g.out_bf16 = 0; fp.g.bias = bias; fp.g.flags = flags;
    fp.g.scale = 1.0f / (1.0f - p_drop); fp.g.thresh = keep_threshold(p_drop); fp.g.mask = mask; fp.g.seed = seed;
    fp.g.stream_id = stream_id; fp.g.rng_off = rng_offset_ptr();
    fp.rowptr = rowptr; fp.col = col; fp.w = w;
    fp.X0 = X0; fp.X1 = X1; fp.ldx0 = ldx0; fp.ldx1 = ldx1; fp.F0 = F0; fp.F1 = F1;
    fp.B = B; fp.N = N; fp.Lp1 = L + 1;
    fp.tiles_per_graph = (N + Cfg::BLOCK_M - 1) / Cfg::BLOCK_M;
    fp.n_mtiles = B * fp.tiles_per_graph;
    fp.promo = gemm_x3_current_promo();
    const int cs = C <= 128 ? 1 : (C <= 256 ? 2 : 4);

    // shared-memory plan: the slab producer when one graph's [N, 32] fp32 slab fits next to >= 2 weight stages
    const int n_boxes = (N + 255) / 256, box_rows = (N + n_boxes - 1) / n_boxes;
    const size_t slab = static_cast<size_t>(n_boxes) * box_rows * 128;
    const size_t rp = (static_cast<size_t>(Cfg::BLOCK_M) * (L + 2) * 4 + 127) / 128 * 128;
    const size_t fixed = Cfg::EPI_BYTES + rp + Cfg::BAR_BYTES;
    const size_t budget = 227 * 1024 - 1024;
    auto total = [&](int na, int nb, size_t s) { return static_cast<size_t>(na) * Cfg::TILE_BYTES + static_cast<size_t>(nb) * Cfg::B_STAGE_BYTES + fixed + s; };
    bool use_slab = false;
    int na = 3, nb = 4;
    if (g_fused_mode != 2) {
        if (total(3, 3, slab) <= budget) { use_slab = true; na = 3; nb = 3; }
        else if (total(3, 2, slab) <= budget) { use_slab = true; na = 3; nb = 2; }
        else if (total(2, 2, slab) <= budget) { use_slab = true; na = 2; nb = 2; }
    }
    if (g_fused_mode == 1 && !use_slab) {
        set_error("grl_graphconv_fused_fwd: the [N=%d, 32] slab (%zu bytes) does not fit", N, slab);
        return GRL_E_UNSUP;
    }
    fp.n_a = na; fp.n_b = nb; fp.box_rows = box_rows; fp.n_boxes = n_boxes;
    fp.off_b = static_cast<uint32_t>(na) * Cfg::TILE_BYTES;
    fp.off_epi = fp.off_b + static_cast<uint32_t>(nb) * Cfg::B_STAGE_BYTES;
    fp.off_rp = fp.off_epi + Cfg::EPI_BYTES;
    fp.off_slab = fp.off_rp + static_cast<uint32_t>(rp);
    fp.off_bars = fp.off_slab + static_cast<uint32_t>(use_slab ? slab : 0);
    const size_t smem = fp.off_bars + Cfg::BAR_BYTES + 1024;

    CUtensorMap tB, tBlo, tX0, tX1;
    int rc;
    if ((rc = make_tmap_sw128(&tB, W_hi, 4, C, K, ldw, Cfg::BLOCK_N))) return rc;
    if ((rc = make_tmap_sw128(&tBlo, W_lo, 4, C, K, ldw, Cfg::BLOCK_N))) return rc;
    tX0 = tB; tX1 = tB;
    if (use_slab) {
        if ((rc = make_tmap_plain(&tX0, X0, 4, static_cast<int64_t>(B) * N, F0, ldx0, Cfg::BLOCK_K, box_rows))) return rc;
        tX1 = tX0;
        if (F1 > 0 && (rc = make_tmap_plain(&tX1, X1, 4, static_cast<int64_t>(B) * N, F1, ldx1, Cfg::BLOCK_K, box_rows))) return rc;
    }
    return use_slab ? launch_fused<true>(fp, tB, tBlo, tX0, tX1, cs, smem, st)
                    : launch_fused<false>(fp, tB, tBlo, tX0, tX1, cs, smem, st);
}
