// NodeSelfAtten on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), flash style: no N x N tensor in HBM.
//
// Reference: NodeSelfAtten.forward, gnn/models/networks/robust_gcn.py:90-96
//     s   = softmax(bmm(f_out, g_out^T), dim=-1)        f_out, g_out: [B, N, d]   (d = H/8, already ReLU'd; UNSCALED,
//     o   = bmm(s, h_out)                               h_out: [B, N, H]           UNMASKED: padded nodes are live)
//     out = gamma * o + V
// and its autograd (dh = P^T do, dP = do h^T, dS = P (dP - D), df = dS g, dg = dS^T f, do = gamma dOut, D = do . o).
//
// Every contraction is a kind::f16 tcgen05.mma on bf16 operands with fp32 accumulation in TMEM.  fp32 models keep their
// accuracy by operand SPLITTING: x = x0 + x1 (+ x2) with bf16 pieces (x0 = bf16(x), x1 = bf16(x - x0), ...), and
// a.b ~ sum over piece pairs (i, j), i + j < NP, of a_i.b_j  -- NP = 3 (six products) carries 24 mantissa bits like an
// fp32 FFMA, NP = 2 (three products) 16 bits, NP = 1 is plain bf16.  All-bf16 operands also mean ONE shared-memory image
// of a tile serves both as a K-major operand (rows x 64 columns of 128 B, SWIZZLE_128B) and as an MN-major operand (the
// same bytes read with the transposed descriptor) -- h, do, f and g are each used both ways.
//
//   attn_tc_fwd_kernel    CTA = (graph, 128 query rows), 128 threads, thread t = TMEM lane t = query row t.
//                         pass 1 over the key blocks: S = f g^T -> row max (K = d is tiny: redoing S is cheap);
//                         pass 2: S again, p = exp(s - max), row sum, P (bf16 pieces) -> TMEM, O += P h with P read from
//                         TMEM (TS form) and h MN-major from shared memory; epilogue o = O / sum, out = gamma o + V.
//                         Training saves o (fp32), max and sum per row -- 3 * N floats per graph instead of N^2.
//   attn_tc_bwd_q_kernel  CTA = (graph, 128 query rows): per key block S and dP = do h^T (K = H), dS in registers,
//                         written over the S columns of TMEM as the A operand of df += dS g.
//   attn_tc_bwd_k_kernel  CTA = (graph, 128 key rows): per query block S^T = g f^T and dP^T = h do^T, then P^T / dS^T in
//                         registers (per-column statistics from shared memory), written over the dP^T / S^T columns as
//                         the A operands of dh += P^T do and dg += dS^T f.  Queries in order: no atomics.
//
// The kernels are serialised per block (TMA -> MMA -> softmax warps -> MMA): the products here are 2-6 % of the model's
// flops; the previous fp32 FFMA kernels (attention.cu, kept for the widths this file does not instantiate) spent
// 12 % / 34 % of the cfg2 / cfg3 step on them.
#include "gemm_common.cuh"

namespace grl {
using namespace sm100;

namespace {

// piece pairs (a_i, b_j), i + j < NP, in the order (0,0) (0,1) (1,0) (0,2) (1,1) (2,0): NP = 1 takes the first one,
// NP = 2 the first three, NP = 3 all six
__host__ __device__ constexpr int kPA(int i) { return (i == 2 || i == 4) ? 1 : (i == 5 ? 2 : 0); }
__host__ __device__ constexpr int kPB(int i) { return (i == 1 || i == 4) ? 1 : (i == 3 ? 2 : 0); }
__host__ __device__ constexpr int n_combos(int np) { return np * (np + 1) / 2; }

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}

__device__ __forceinline__ uint32_t bf16x2_bits(float lo_addr, float hi_addr) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo_addr, hi_addr);
    return *reinterpret_cast<const uint32_t*>(&v);
}

// 32 fp32 values of one row (consecutive K positions) -> NP bf16 pieces, 16 packed 32-bit columns each
template <int NP>
__device__ __forceinline__ void split_row32(const float (&v)[32], uint32_t (&pk)[NP][16]) {
#pragma unroll
    for (int q = 0; q < 16; ++q) {
        float a = v[2 * q], b = v[2 * q + 1];
#pragma unroll
        for (int pc = 0; pc < NP; ++pc) {
            const __nv_bfloat16 ha = __float2bfloat16_rn(a), hb = __float2bfloat16_rn(b);
            pk[pc][q] = static_cast<uint32_t>(__bfloat16_as_ushort(ha)) | (static_cast<uint32_t>(__bfloat16_as_ushort(hb)) << 16);
            a -= __bfloat162float(ha);
            b -= __bfloat162float(hb);
        }
    }
}

struct AttnTcParams {
    int B, N;
    int io_bf16;               // dtype of V / out / df / dg / dh: 0 fp32, 1 bf16
    const float* gamma;        // [H]
    const void* V;             // [B*N, H]
    void* out;                 // [B*N, H]
    float* o_save;             // [B*N, H] fp32 (training) or null
    float* m_save;             // [B*N] row max (training) or null
    float* l_save;             // [B*N] row sum of exp(s - max)
    float* Dvec;               // [B*N] D_i = sum_j P_ij dP_ij (written by the query-major backward kernel, read by the key-major one)
    void* df;                  // [B*N, D]
    void* dg;                  // [B*N, D]
    void* dh;                  // [B*N, H]
};

template <int NPS, int NP, int H>
struct AttnSmem {
    static constexpr int SLAB = 128 * 128;            // 128 rows x 128 bytes (64 bf16 columns), SWIZZLE_128B
    static constexpr int HS = H / 64;                 // slabs of one piece of an H-wide tile
};

__device__ __forceinline__ void store_row_chunk(void* base, int io_bf16, int64_t row, int ld, int col, const float (&v)[32]) {
    if (io_bf16) {
        __nv_bfloat16* p = static_cast<__nv_bfloat16*>(base) + row * ld + col;
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
            const uint4 w = make_uint4(bf16x2_bits(v[j], v[j + 1]), bf16x2_bits(v[j + 2], v[j + 3]),
                                       bf16x2_bits(v[j + 4], v[j + 5]), bf16x2_bits(v[j + 6], v[j + 7]));
            *reinterpret_cast<uint4*>(p + j) = w;
        }
    } else {
        float* p = static_cast<float*>(base) + row * ld + col;
#pragma unroll
        for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(p + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    }
}

__device__ __forceinline__ void load_row_chunk(const void* base, int io_bf16, int64_t row, int ld, int col, float (&v)[32]) {
    if (io_bf16) {
        const __nv_bfloat16* p = static_cast<const __nv_bfloat16*>(base) + row * ld + col;
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
            const uint4 w = __ldg(reinterpret_cast<const uint4*>(p + j));
            const uint32_t u[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                v[j + 2 * t] = __uint_as_float(u[t] << 16);
                v[j + 2 * t + 1] = __uint_as_float(u[t] & 0xFFFF0000u);
            }
        }
    } else {
        const float* p = static_cast<const float*>(base) + row * ld + col;
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(p + j));
            v[j] = t.x; v[j + 1] = t.y; v[j + 2] = t.z; v[j + 3] = t.w;
        }
    }
}

// ================================================================================================ forward
// NP pieces for every operand.  TMEM columns: S [0,128) | O [128, 128+H) | P pieces [128+H, 128+H+64*NP).
// kFwdGroups warps share every TMEM lane quarter (a warp may touch lanes 32 (warp % 4) ...): thread (quarter, lane) of
// group g owns query row quarter*32 + lane and the 32-column score chunks c with c % kFwdGroups == g.  With one group
// (128 threads) the exp / piece-split / TMEM traffic of 128 x 128 scores per key block ran on 4 warps while the tensor
// core idled: ~6.8 us per key block.
constexpr int kFwdGroups = 4;
template <int NP, int H, int D>
__global__ void __launch_bounds__(128 * kFwdGroups, 1)
attn_tc_fwd_kernel(const AttnTcParams p, const __grid_constant__ CUtensorMap tmF, const __grid_constant__ CUtensorMap tmG,
                   const __grid_constant__ CUtensorMap tmH) {
    constexpr int NC = n_combos(NP), HS = H / 64, SLAB = 128 * 128;
    constexpr int COL_S = 0, COL_O = 128, COL_P = 128 + H;
    static_assert(COL_P + 64 * NP <= 512, "TMEM budget");
    static_assert(D == 16 || D == 32, "score width");
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    uint8_t* sQ = tiles;                          // NP slabs
    uint8_t* sG = sQ + NP * SLAB;                 // NP slabs
    uint8_t* sH = sG + NP * SLAB;                 // NP * HS slabs
    uint64_t* bars = reinterpret_cast<uint64_t*>(sH + NP * HS * SLAB);
    uint64_t* bar_tma = bars;                     // the query tile, then the G tile of the next key block
    uint64_t* bar_mma = bars + 1;
    uint64_t* bar_h = bars + 2;                   // the h tile of the next key block
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);
    float* red = reinterpret_cast<float*>(bars + 8);           // [kFwdGroups][128] row max / row sum partials

    const int tid = threadIdx.x;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int grp = warp >> 2, rit = (warp & 3) * 32 + (tid & 31);      // column group, query row inside the tile
    const int b = blockIdx.y, q0 = blockIdx.x * 128;
    const int N = p.N, nkb = (N + 127) / 128;

    if (tid == 0) {
        tma_prefetch_desc(&tmF);
        tma_prefetch_desc(&tmG);
        tma_prefetch_desc(&tmH);
        mbar_init(bar_tma, 1);
        mbar_init(bar_mma, 1);
        mbar_init(bar_h, 1);
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc<512>(tmem_slot);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t lane_sel = static_cast<uint32_t>((warp & 3) * 32) << 16;
    uint32_t ph_tma = 0, ph_mma = 0;

    if (tid == 0) {
        mbar_expect_tx(bar_tma, NP * SLAB);
#pragma unroll
        for (int pc = 0; pc < NP; ++pc) tma_load_3d(sQ + pc * SLAB, &tmF, bar_tma, 0, q0, pc * p.B + b);
    }
    mbar_wait(bar_tma, ph_tma);
    ph_tma ^= 1;

    constexpr uint32_t idescS = umma_idesc(1, 128, 128, 0, 0);
    constexpr uint32_t idescO = umma_idesc(1, 128, H, 0, 1);

    auto issue_S = [&]() {            // warp 0, converged
        tc_fence_after_sync();
        if (elect_one_sync()) {
#pragma unroll
            for (int i = 0; i < NC; ++i) {
                const uint64_t a = umma_desc_kmajor_sw128(smem_u32(sQ + kPA(i) * SLAB));
                const uint64_t bd = umma_desc_kmajor_sw128(smem_u32(sG + kPB(i) * SLAB));
#pragma unroll
                for (int k = 0; k < D / 16; ++k)
                    umma<false>(tmem_base + COL_S, a + 2 * k, bd + 2 * k, idescS, (i | k) != 0);
            }
            umma_commit(bar_mma);
        }
        __syncwarp();
    };

    // The operand tiles of the NEXT key block are fetched as soon as the MMAs that read the current ones have retired (one
    // buffer per operand): the G tile right after the score product, the h tile right after O += P h -- so a TMA round
    // trip no longer sits in front of every score product.  The first h tile travels during pass 1.
    auto load_G = [&](int kb) {            // tid 0
        mbar_expect_tx(bar_tma, NP * SLAB);
#pragma unroll
        for (int pc = 0; pc < NP; ++pc) tma_load_3d(sG + pc * SLAB, &tmG, bar_tma, 0, kb * 128, pc * p.B + b);
    };
    auto load_H = [&](int kb) {            // tid 0
        mbar_expect_tx(bar_h, NP * HS * SLAB);
#pragma unroll
        for (int pc = 0; pc < NP; ++pc)
#pragma unroll
            for (int s = 0; s < HS; ++s)
                tma_load_3d(sH + (pc * HS + s) * SLAB, &tmH, bar_h, 64 * s, kb * 128, pc * p.B + b);
    };
    uint32_t ph_h = 0;
    if (tid == 0) {
        load_G(0);
        load_H(0);
    }

    // ------------------------------------------------------------ pass 1: row max
    float m = -INFINITY;
    for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(bar_tma, ph_tma);
        ph_tma ^= 1;
        if (warp == 0) issue_S();
        mbar_wait(bar_mma, ph_mma);
        ph_mma ^= 1;
        tc_fence_after_sync();
        if (tid == 0) load_G(kb + 1 < nkb ? kb + 1 : 0);        // the last one is pass 2's first tile
#pragma unroll
        for (int c = grp; c < 4; c += kFwdGroups) {
            uint32_t r[32];
            tmem_ld_32x32(tmem_base + lane_sel + COL_S + 32 * c, r);
            tmem_ld_wait();
            const int valid = N - (kb * 128 + 32 * c);
#pragma unroll
            for (int e = 0; e < 32; ++e)
                if (e < valid) m = fmaxf(m, __uint_as_float(r[e]));
        }
        tc_fence_before_sync();
        __syncthreads();
    }
    red[grp * 128 + rit] = m;                   // the row max over all column groups
    __syncthreads();
#pragma unroll
    for (int g2 = 0; g2 < kFwdGroups; ++g2) m = fmaxf(m, red[g2 * 128 + rit]);
    __syncthreads();

    // ------------------------------------------------------------ pass 2: P, row sum, O += P h
    float l = 0.0f;
    for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(bar_tma, ph_tma);
        ph_tma ^= 1;
        if (warp == 0) issue_S();
        mbar_wait(bar_mma, ph_mma);
        ph_mma ^= 1;
        tc_fence_after_sync();
        if (tid == 0 && kb + 1 < nkb) load_G(kb + 1);
#pragma unroll
        for (int c = grp; c < 4; c += kFwdGroups) {
            uint32_t r[32];
            tmem_ld_32x32(tmem_base + lane_sel + COL_S + 32 * c, r);
            tmem_ld_wait();
            const int valid = N - (kb * 128 + 32 * c);
            float pv[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) {
                pv[e] = e < valid ? __expf(__uint_as_float(r[e]) - m) : 0.0f;
                l += pv[e];
            }
            uint32_t pk[NP][16];
            split_row32<NP>(pv, pk);
#pragma unroll
            for (int pc = 0; pc < NP; ++pc) tmem_st_32x16(tmem_base + lane_sel + COL_P + 64 * pc + 16 * c, pk[pc]);
        }
        tmem_st_wait();
        tc_fence_before_sync();
        __syncthreads();
        mbar_wait(bar_h, ph_h);
        ph_h ^= 1;
        if (warp == 0) {
            tc_fence_after_sync();
            if (elect_one_sync()) {
#pragma unroll
                for (int i = 0; i < NC; ++i) {
                    const uint64_t hd = umma_desc_mnmajor_sw128(smem_u32(sH + kPB(i) * HS * SLAB), SLAB, 1024, 2);
#pragma unroll
                    for (int k = 0; k < 8; ++k)
                        umma_ts<false>(tmem_base + COL_O, tmem_base + COL_P + 64 * kPA(i) + 8 * k, hd + 128 * k, idescO,
                                       (kb | i | k) != 0);
                }
                umma_commit(bar_mma);
            }
            __syncwarp();
        }
        mbar_wait(bar_mma, ph_mma);
        ph_mma ^= 1;
        tc_fence_after_sync();
        if (tid == 0 && kb + 1 < nkb) load_H(kb + 1);
        __syncthreads();
    }

    // ------------------------------------------------------------ epilogue
    red[grp * 128 + rit] = l;                   // the row sum over all column groups, added in group order
    __syncthreads();
    l = 0.0f;
#pragma unroll
    for (int g2 = 0; g2 < kFwdGroups; ++g2) l += red[g2 * 128 + rit];
    const int row = q0 + rit;
    const bool live = row < N;
    const int64_t grow = static_cast<int64_t>(b) * N + row;
    const float inv_l = 1.0f / l;
    if (live && p.m_save && grp == 0) {
        p.m_save[grow] = m;
        p.l_save[grow] = l;
    }
#pragma unroll 1
    for (int c = grp; c < H / 32; c += kFwdGroups) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + lane_sel + COL_O + 32 * c, r);
        tmem_ld_wait();
        if (live) {
            float o[32], v[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) o[e] = __uint_as_float(r[e]) * inv_l;
            if (p.o_save) store_row_chunk(p.o_save, 0, grow, H, 32 * c, o);
            load_row_chunk(p.V, p.io_bf16, grow, H, 32 * c, v);
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] = fmaf(__ldg(p.gamma + 32 * c + e), o[e], v[e]);
            store_row_chunk(p.out, p.io_bf16, grow, H, 32 * c, v);
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem_base);
}

// ================================================================================================ backward, query-major
// df_i = sum_j dS_ij g_j.  NPS pieces for the score product, NP (<= 2) for dP = do h^T and for dS g.
// Shared memory: F (NPS slabs), DO (NP*HS) of this query block; G (NPS), Hh (NP*HS) of the current key block.
// TMEM: S [0,128) (overwritten chunk by chunk with the dS pieces) | dP [128,256) | dF [256,320).
// kBwdGroups warps per TMEM lane quarter (as in the forward): group g owns the score chunks c with c % kBwdGroups == g.
constexpr int kBwdGroups = 2;
template <int NPS, int NP, int H, int D>
__global__ void __launch_bounds__(128 * kBwdGroups, 1)
attn_tc_bwd_q_kernel(const AttnTcParams p, const __grid_constant__ CUtensorMap tmF, const __grid_constant__ CUtensorMap tmG,
                     const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmDO) {
    constexpr int NCS = n_combos(NPS), NC = n_combos(NP), HS = H / 64, SLAB = 128 * 128;
    constexpr int COL_S = 0, COL_DP = 128, COL_DF = 256;
    static_assert(NP <= 2, "dS pieces are written over the 32-column S chunk they come from");
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    uint8_t* sF = tiles;
    uint8_t* sDO = sF + NPS * SLAB;
    uint8_t* sG = sDO + NP * HS * SLAB;
    uint8_t* sH = sG + NPS * SLAB;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sH + NP * HS * SLAB);
    uint64_t* bar_tma = bars;
    uint64_t* bar_mma = bars + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
    float* red = reinterpret_cast<float*>(bars + 8);           // [kBwdGroups][128] partial D_i

    const int tid = threadIdx.x;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int grp = warp >> 2, rit = (warp & 3) * 32 + (tid & 31);
    const int b = blockIdx.y, q0 = blockIdx.x * 128;
    const int N = p.N, nkb = (N + 127) / 128;

    if (tid == 0) {
        tma_prefetch_desc(&tmF);
        tma_prefetch_desc(&tmG);
        tma_prefetch_desc(&tmH);
        tma_prefetch_desc(&tmDO);
        mbar_init(bar_tma, 1);
        mbar_init(bar_mma, 1);
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc<512>(tmem_slot);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t lane_sel = static_cast<uint32_t>((warp & 3) * 32) << 16;
    uint32_t ph_tma = 0, ph_mma = 0;

    if (tid == 0) {
        mbar_expect_tx(bar_tma, (NPS + NP * HS) * SLAB);
#pragma unroll
        for (int pc = 0; pc < NPS; ++pc) tma_load_3d(sF + pc * SLAB, &tmF, bar_tma, 0, q0, pc * p.B + b);
#pragma unroll
        for (int pc = 0; pc < NP; ++pc)
#pragma unroll
            for (int s = 0; s < HS; ++s) tma_load_3d(sDO + (pc * HS + s) * SLAB, &tmDO, bar_tma, 64 * s, q0, pc * p.B + b);
    }
    const int row = q0 + rit;
    const bool live = row < N;
    const int64_t grow = static_cast<int64_t>(b) * N + row;
    const float m = live ? p.m_save[grow] : 0.0f;
    const float inv_l = live ? 1.0f / p.l_save[grow] : 0.0f;
    mbar_wait(bar_tma, ph_tma);
    ph_tma ^= 1;

    constexpr uint32_t idescS = umma_idesc(1, 128, 128, 0, 0);
    constexpr uint32_t idescDF = umma_idesc(1, 128, 64, 0, 1);

    // Two passes over the key blocks.  Pass 0 accumulates D_i = sum_j P_ij dP_ij from the SAME P and dP the second pass
    // uses (not do_i . o_i from the forward's rounded products): sum_j dS_ij is then zero to fp32 rounding, and the
    // cancellation in dS = P (dP - D) does not amplify the rounding of the operand pieces into df / dg.
    float Dv = 0.0f;
    for (int pass = 0; pass < 2; ++pass) {
        for (int kb = 0; kb < nkb; ++kb) {
            if (tid == 0) {
                mbar_expect_tx(bar_tma, (NPS + NP * HS) * SLAB);
#pragma unroll
                for (int pc = 0; pc < NPS; ++pc) tma_load_3d(sG + pc * SLAB, &tmG, bar_tma, 0, kb * 128, pc * p.B + b);
#pragma unroll
                for (int pc = 0; pc < NP; ++pc)
#pragma unroll
                    for (int s = 0; s < HS; ++s)
                        tma_load_3d(sH + (pc * HS + s) * SLAB, &tmH, bar_tma, 64 * s, kb * 128, pc * p.B + b);
            }
            mbar_wait(bar_tma, ph_tma);
            ph_tma ^= 1;
            if (warp == 0) {
                tc_fence_after_sync();
                if (elect_one_sync()) {
#pragma unroll
                    for (int i = 0; i < NCS; ++i) {     // S = f g^T
                        const uint64_t a = umma_desc_kmajor_sw128(smem_u32(sF + kPA(i) * SLAB));
                        const uint64_t bd = umma_desc_kmajor_sw128(smem_u32(sG + kPB(i) * SLAB));
#pragma unroll
                        for (int k = 0; k < D / 16; ++k)
                            umma<false>(tmem_base + COL_S, a + 2 * k, bd + 2 * k, idescS, (i | k) != 0);
                    }
#pragma unroll
                    for (int i = 0; i < NC; ++i) {      // dP = do h^T, K = H
#pragma unroll
                        for (int kk = 0; kk < H / 16; ++kk) {
                            const uint64_t a = umma_desc_kmajor_sw128(smem_u32(sDO + (kPA(i) * HS + kk / 4) * SLAB)) + 2 * (kk % 4);
                            const uint64_t bd = umma_desc_kmajor_sw128(smem_u32(sH + (kPB(i) * HS + kk / 4) * SLAB)) + 2 * (kk % 4);
                            umma<false>(tmem_base + COL_DP, a, bd, idescS, (i | kk) != 0);
                        }
                    }
                    umma_commit(bar_mma);
                }
                __syncwarp();
            }
            mbar_wait(bar_mma, ph_mma);
            ph_mma ^= 1;
            tc_fence_after_sync();
            if (pass == 0) {
                float dacc = 0.0f;
#pragma unroll
                for (int c = grp; c < 4; c += kBwdGroups) {
                    uint32_t rs[32], rp[32];
                    tmem_ld_32x32(tmem_base + lane_sel + COL_S + 32 * c, rs);
                    tmem_ld_32x32(tmem_base + lane_sel + COL_DP + 32 * c, rp);
                    tmem_ld_wait();
                    const int valid = N - (kb * 128 + 32 * c);
#pragma unroll
                    for (int e = 0; e < 32; ++e) {
                        const float pe = e < valid ? __expf(__uint_as_float(rs[e]) - m) * inv_l : 0.0f;
                        dacc = fmaf(pe, __uint_as_float(rp[e]), dacc);
                    }
                }
                Dv += dacc;
                tc_fence_before_sync();
                __syncthreads();
                continue;
            }
#pragma unroll
            for (int c = grp; c < 4; c += kBwdGroups) {
                uint32_t rs[32], rp[32];
                tmem_ld_32x32(tmem_base + lane_sel + COL_S + 32 * c, rs);
                tmem_ld_32x32(tmem_base + lane_sel + COL_DP + 32 * c, rp);
                tmem_ld_wait();
                const int valid = N - (kb * 128 + 32 * c);
                float ds[32];
#pragma unroll
                for (int e = 0; e < 32; ++e) {
                    const float pe = e < valid ? __expf(__uint_as_float(rs[e]) - m) * inv_l : 0.0f;
                    ds[e] = pe * (__uint_as_float(rp[e]) - Dv);
                }
                uint32_t pk[NP][16];
                split_row32<NP>(ds, pk);
#pragma unroll
                for (int pc = 0; pc < NP; ++pc) tmem_st_32x16(tmem_base + lane_sel + COL_S + 32 * c + 16 * pc, pk[pc]);
            }
            tmem_st_wait();
            tc_fence_before_sync();
            __syncthreads();
            if (warp == 0) {
                tc_fence_after_sync();
                if (elect_one_sync()) {
#pragma unroll
                    for (int i = 0; i < NC; ++i) {      // df += dS g: A = dS pieces in TMEM, B = g MN-major (N = 64, zero padded)
                        const uint64_t gd = umma_desc_mnmajor_sw128(smem_u32(sG + kPB(i) * SLAB), SLAB, 1024, 2);
#pragma unroll
                        for (int k = 0; k < 8; ++k)
                            umma_ts<false>(tmem_base + COL_DF, tmem_base + COL_S + 32 * (k / 2) + 16 * kPA(i) + 8 * (k % 2),
                                           gd + 128 * k, idescDF, (kb | i | k) != 0);
                    }
                    umma_commit(bar_mma);
                }
                __syncwarp();
            }
            mbar_wait(bar_mma, ph_mma);
            ph_mma ^= 1;
            tc_fence_after_sync();
            __syncthreads();
        }
        if (pass == 0) {                        // D_i over all column groups, added in group order
            red[grp * 128 + rit] = Dv;
            __syncthreads();
            Dv = 0.0f;
#pragma unroll
            for (int g2 = 0; g2 < kBwdGroups; ++g2) Dv += red[g2 * 128 + rit];
            if (live && grp == 0) p.Dvec[grow] = Dv;
        }
    }

    if (grp == 0) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + lane_sel + COL_DF, r);
        tmem_ld_wait();
        if (live) {
            if (p.io_bf16) {
                __nv_bfloat16* o = static_cast<__nv_bfloat16*>(p.df) + grow * D;
#pragma unroll
                for (int j = 0; j < D; j += 8)
                    *reinterpret_cast<uint4*>(o + j) = make_uint4(
                        bf16x2_bits(__uint_as_float(r[j]), __uint_as_float(r[j + 1])), bf16x2_bits(__uint_as_float(r[j + 2]), __uint_as_float(r[j + 3])),
                        bf16x2_bits(__uint_as_float(r[j + 4]), __uint_as_float(r[j + 5])), bf16x2_bits(__uint_as_float(r[j + 6]), __uint_as_float(r[j + 7])));
            } else {
                float* o = static_cast<float*>(p.df) + grow * D;
#pragma unroll
                for (int j = 0; j < D; j += 4) *reinterpret_cast<uint4*>(o + j) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem_base);
}

// ================================================================================================ backward, key-major
// dh_j = sum_i P_ij do_i,  dg_j = sum_i dS_ij f_i.  CTA = 128 key rows; query blocks of BQ rows.
// Shared memory: G (NPS slabs), Hh (NP*HS) of this key block; F (NPS slabs of BQ rows), DO (NP*HS) of the current query
// block; row statistics of the query block.  TMEM: S^T [0,BQ) | dP^T [BQ,2BQ) | dG [2BQ, 2BQ+64) | dH [2BQ+64, 2BQ+64+H).
template <int NPS, int NP, int H, int D, int BQ>
__global__ void __launch_bounds__(128 * kBwdGroups, 1)
attn_tc_bwd_k_kernel(const AttnTcParams p, const __grid_constant__ CUtensorMap tmF, const __grid_constant__ CUtensorMap tmG,
                     const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmDO) {
    constexpr int NCS = n_combos(NPS), NC = n_combos(NP), HS = H / 64, SLAB = 128 * 128, SLABQ = BQ * 128;
    constexpr int COL_S = 0, COL_DP = BQ, COL_DG = 2 * BQ, COL_DH = 2 * BQ + 64;
    static_assert(NP <= 2 && (BQ == 64 || BQ == 128) && COL_DH + H <= 512, "TMEM budget");
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* tiles = smem_raw + (base_u32 - smem_u32(smem_raw));
    uint8_t* sG = tiles;
    uint8_t* sH = sG + NPS * SLAB;
    uint8_t* sF = sH + NP * HS * SLAB;
    uint8_t* sDO = sF + NPS * SLABQ;
    float* stat = reinterpret_cast<float*>(sDO + NP * HS * SLABQ);      // m | 1/l | D, BQ each
    uint64_t* bars = reinterpret_cast<uint64_t*>(stat + 3 * BQ);
    uint64_t* bar_tma = bars;
    uint64_t* bar_mma = bars + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);

    const int tid = threadIdx.x;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int grp = warp >> 2, rit = (warp & 3) * 32 + (tid & 31);
    const int b = blockIdx.y, k0 = blockIdx.x * 128;
    const int N = p.N, nqb = (N + BQ - 1) / BQ;

    if (tid == 0) {
        tma_prefetch_desc(&tmF);
        tma_prefetch_desc(&tmG);
        tma_prefetch_desc(&tmH);
        tma_prefetch_desc(&tmDO);
        mbar_init(bar_tma, 1);
        mbar_init(bar_mma, 1);
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc<512>(tmem_slot);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t lane_sel = static_cast<uint32_t>((warp & 3) * 32) << 16;
    uint32_t ph_tma = 0, ph_mma = 0;

    if (tid == 0) {
        mbar_expect_tx(bar_tma, (NPS + NP * HS) * SLAB);
#pragma unroll
        for (int pc = 0; pc < NPS; ++pc) tma_load_3d(sG + pc * SLAB, &tmG, bar_tma, 0, k0, pc * p.B + b);
#pragma unroll
        for (int pc = 0; pc < NP; ++pc)
#pragma unroll
            for (int s = 0; s < HS; ++s) tma_load_3d(sH + (pc * HS + s) * SLAB, &tmH, bar_tma, 64 * s, k0, pc * p.B + b);
    }
    const int krow = k0 + rit;
    const bool live = krow < N;
    const int64_t gkrow = static_cast<int64_t>(b) * N + krow;
    mbar_wait(bar_tma, ph_tma);
    ph_tma ^= 1;

    constexpr uint32_t idescS = umma_idesc(1, 128, BQ, 0, 0);
    constexpr uint32_t idescDH = umma_idesc(1, 128, H, 0, 1);
    constexpr uint32_t idescDG = umma_idesc(1, 128, 64, 0, 1);

    for (int qb = 0; qb < nqb; ++qb) {
        const int q0 = qb * BQ;
        if (tid == 0) {
            mbar_expect_tx(bar_tma, (NPS + NP * HS) * SLABQ);
#pragma unroll
            for (int pc = 0; pc < NPS; ++pc) tma_load_3d(sF + pc * SLABQ, &tmF, bar_tma, 0, q0, pc * p.B + b);
#pragma unroll
            for (int pc = 0; pc < NP; ++pc)
#pragma unroll
                for (int s = 0; s < HS; ++s)
                    tma_load_3d(sDO + (pc * HS + s) * SLABQ, &tmDO, bar_tma, 64 * s, q0, pc * p.B + b);
        }
        if (tid < BQ) {
            const int q = q0 + tid;
            const int64_t gq = static_cast<int64_t>(b) * N + q;
            const bool ql = q < N;
            stat[tid] = ql ? p.m_save[gq] : 0.0f;
            stat[BQ + tid] = ql ? 1.0f / p.l_save[gq] : 0.0f;
            stat[2 * BQ + tid] = ql ? p.Dvec[gq] : 0.0f;
        }
        mbar_wait(bar_tma, ph_tma);
        ph_tma ^= 1;
        if (warp == 0) {
            tc_fence_after_sync();
            if (elect_one_sync()) {
#pragma unroll
                for (int i = 0; i < NCS; ++i) {     // S^T = g f^T   (M = keys, N = queries, K = D)
                    const uint64_t a = umma_desc_kmajor_sw128(smem_u32(sG + kPA(i) * SLAB));
                    const uint64_t bd = umma_desc_kmajor_sw128(smem_u32(sF + kPB(i) * SLABQ));
#pragma unroll
                    for (int k = 0; k < D / 16; ++k)
                        umma<false>(tmem_base + COL_S, a + 2 * k, bd + 2 * k, idescS, (i | k) != 0);
                }
#pragma unroll
                for (int i = 0; i < NC; ++i) {      // dP^T = h do^T  (K = H)
#pragma unroll
                    for (int kk = 0; kk < H / 16; ++kk) {
                        const uint64_t a = umma_desc_kmajor_sw128(smem_u32(sH + (kPA(i) * HS + kk / 4) * SLAB)) + 2 * (kk % 4);
                        const uint64_t bd = umma_desc_kmajor_sw128(smem_u32(sDO + (kPB(i) * HS + kk / 4) * SLABQ)) + 2 * (kk % 4);
                        umma<false>(tmem_base + COL_DP, a, bd, idescS, (i | kk) != 0);
                    }
                }
                umma_commit(bar_mma);
            }
            __syncwarp();
        }
        __syncthreads();                  // query-block statistics are in shared memory
        mbar_wait(bar_mma, ph_mma);
        ph_mma ^= 1;
        tc_fence_after_sync();
#pragma unroll
        for (int c = grp; c < BQ / 32; c += kBwdGroups) {
            uint32_t rs[32], rp[32];
            tmem_ld_32x32(tmem_base + lane_sel + COL_S + 32 * c, rs);
            tmem_ld_32x32(tmem_base + lane_sel + COL_DP + 32 * c, rp);
            tmem_ld_wait();
            float pv[32], ds[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) {
                const int qi = 32 * c + e;
                // padded queries carry 1/l = 0, padded keys are masked by `live`
                const float pe = live ? __expf(__uint_as_float(rs[e]) - stat[qi]) * stat[BQ + qi] : 0.0f;
                pv[e] = pe;
                ds[e] = pe * (__uint_as_float(rp[e]) - stat[2 * BQ + qi]);
            }
            uint32_t pk[NP][16];
            split_row32<NP>(pv, pk);
#pragma unroll
            for (int pc = 0; pc < NP; ++pc) tmem_st_32x16(tmem_base + lane_sel + COL_DP + 32 * c + 16 * pc, pk[pc]);
            split_row32<NP>(ds, pk);
#pragma unroll
            for (int pc = 0; pc < NP; ++pc) tmem_st_32x16(tmem_base + lane_sel + COL_S + 32 * c + 16 * pc, pk[pc]);
        }
        tmem_st_wait();
        tc_fence_before_sync();
        __syncthreads();
        if (warp == 0) {
            tc_fence_after_sync();
            if (elect_one_sync()) {
#pragma unroll
                for (int i = 0; i < NC; ++i) {
                    const uint64_t dod = umma_desc_mnmajor_sw128(smem_u32(sDO + kPB(i) * HS * SLABQ), SLABQ, 1024, 2);
                    const uint64_t fd = umma_desc_mnmajor_sw128(smem_u32(sF + kPB(i) * SLABQ), SLABQ, 1024, 2);
#pragma unroll
                    for (int k = 0; k < BQ / 16; ++k) {
                        const uint32_t acol = 32 * (k / 2) + 16 * kPA(i) + 8 * (k % 2);
                        umma_ts<false>(tmem_base + COL_DH, tmem_base + COL_DP + acol, dod + 128 * k, idescDH, (qb | i | k) != 0);
                        umma_ts<false>(tmem_base + COL_DG, tmem_base + COL_S + acol, fd + 128 * k, idescDG, (qb | i | k) != 0);
                    }
                }
                umma_commit(bar_mma);
            }
            __syncwarp();
        }
        mbar_wait(bar_mma, ph_mma);
        ph_mma ^= 1;
        tc_fence_after_sync();
        __syncthreads();
    }

#pragma unroll 1
    for (int c = grp; c < H / 32; c += kBwdGroups) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + lane_sel + COL_DH + 32 * c, r);
        tmem_ld_wait();
        if (live) {
            float v[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] = __uint_as_float(r[e]);
            store_row_chunk(p.dh, p.io_bf16, gkrow, H, 32 * c, v);
        }
    }
    if (grp == 0) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + lane_sel + COL_DG, r);
        tmem_ld_wait();
        if (live) {
            if (p.io_bf16) {
                __nv_bfloat16* o = static_cast<__nv_bfloat16*>(p.dg) + gkrow * D;
#pragma unroll
                for (int j = 0; j < D; j += 8)
                    *reinterpret_cast<uint4*>(o + j) = make_uint4(
                        bf16x2_bits(__uint_as_float(r[j]), __uint_as_float(r[j + 1])), bf16x2_bits(__uint_as_float(r[j + 2]), __uint_as_float(r[j + 3])),
                        bf16x2_bits(__uint_as_float(r[j + 4]), __uint_as_float(r[j + 5])), bf16x2_bits(__uint_as_float(r[j + 6]), __uint_as_float(r[j + 7])));
            } else {
                float* o = static_cast<float*>(p.dg) + gkrow * D;
#pragma unroll
                for (int j = 0; j < D; j += 4) *reinterpret_cast<uint4*>(o + j) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem_base);
}

// ================================================================================================ operand preparation
// x [rows, C] (fp32 or bf16, leading dimension ld) times an optional per-column scale -> np bf16 pieces, each a dense
// [rows, C] matrix, piece pc at out + pc * rows * C:  x0 = bf16(x), x1 = bf16(x - x0), x2 = bf16(x - x0 - x1).
template <typename T>
__global__ void attn_split_kernel(const T* __restrict__ x, int64_t ld, int64_t rows, int C, const float* __restrict__ scale,
                                  int np, __nv_bfloat16* __restrict__ out) {
    const int64_t total = rows * C;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
        const int64_t r = t / C;
        const int c = static_cast<int>(t - r * C);
        float v = to_float(x[r * ld + c]);
        if (scale) v *= scale[c];
        for (int pc = 0; pc < np; ++pc) {
            const __nv_bfloat16 h = __float2bfloat16_rn(v);
            out[pc * total + t] = h;
            v -= __bfloat162float(h);
        }
    }
}

// The same, eight consecutive columns per thread: 16-byte loads and one 16-byte store per piece (the element-per-thread
// form above -- a 64-bit division and 2-byte stores per element -- took 73 us for the 128 000 x 128 h operand of cfg2).
// Needs C % 8 == 0 and 16-byte aligned rows of x / out.
__device__ __forceinline__ void load8(const float* p, float (&v)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 a = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        v[2 * k] = __uint_as_float(w[k] << 16);
        v[2 * k + 1] = __uint_as_float(w[k] & 0xFFFF0000u);
    }
}
// v -> np bf16 pieces of eight values each, stored as one uint4 per piece at out + pc * piece_stride
__device__ __forceinline__ void split_store8(float (&v)[8], int np, __nv_bfloat16* out, int64_t piece_stride) {
    for (int pc = 0; pc < np; ++pc) {
        uint32_t w[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const __nv_bfloat16 a = __float2bfloat16_rn(v[2 * k]), b = __float2bfloat16_rn(v[2 * k + 1]);
            w[k] = static_cast<uint32_t>(__bfloat16_as_ushort(a)) | (static_cast<uint32_t>(__bfloat16_as_ushort(b)) << 16);
            v[2 * k] -= __bfloat162float(a);
            v[2 * k + 1] -= __bfloat162float(b);
        }
        *reinterpret_cast<uint4*>(out + pc * piece_stride) = make_uint4(w[0], w[1], w[2], w[3]);
    }
}
template <typename T>
__global__ void attn_split_vec8_kernel(const T* __restrict__ x, int64_t ld, int64_t rows, int C, const float* __restrict__ scale,
                                       int np, __nv_bfloat16* __restrict__ out) {
    const int c8 = C >> 3;
    const int64_t total8 = rows * c8, total = rows * C;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total8; t += stride) {
        const int64_t r = t / c8;
        const int c = static_cast<int>(t - r * c8) << 3;
        float v[8];
        load8(x + r * ld + c, v);
        if (scale) {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] *= __ldg(scale + c + e);
        }
        split_store8(v, np, out + r * C + c, total);
    }
}

// Backward preparation, one pass over dOut and o: do = gamma * dOut as np bf16 pieces and the per-block partial sums of
// d(gamma)_c = sum_i dOut_ic o_ic (rows of a block in order; the caller adds the blocks up).  D_i is computed by the
// query-major kernel itself.
// Block = 256 threads = 8 warps over kPrepRows rows.
constexpr int kPrepRows = 64;
__device__ __forceinline__ void load4(const float* p, float (&v)[4]) {
    const float4 a = *reinterpret_cast<const float4*>(p);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
__device__ __forceinline__ void load4(const __nv_bfloat16* p, float (&v)[4]) {
    const uint2 a = *reinterpret_cast<const uint2*>(p);
    v[0] = __uint_as_float(a.x << 16); v[1] = __uint_as_float(a.x & 0xFFFF0000u);
    v[2] = __uint_as_float(a.y << 16); v[3] = __uint_as_float(a.y & 0xFFFF0000u);
}
// One pass: a lane owns four consecutive columns (16-byte loads of dOut and o, one 8-byte store per piece) and carries
// their dOut * o sums over its warp's rows; the eight warps' sums are added in warp order.  H % 4 == 0, H <= 256.
// (The first version stored 2 bytes per thread and then re-read dOut and o column by column with half the block idle:
// 97 us at cfg2.)
template <typename T>
__global__ void __launch_bounds__(256)
attn_bwd_prep_kernel(const T* __restrict__ dOut, const float* __restrict__ o, const float* __restrict__ gamma, int64_t rows,
                     int H, int np, __nv_bfloat16* __restrict__ do_p, float* __restrict__ part) {
    __shared__ float red[8][256];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t r0 = static_cast<int64_t>(blockIdx.x) * kPrepRows;
    const int64_t total = rows * H;
    for (int c = lane * 4; c < H; c += 128) {
        const float4 gm = make_float4(__ldg(gamma + c), __ldg(gamma + c + 1), __ldg(gamma + c + 2), __ldg(gamma + c + 3));   // a parameter view: 4-byte aligned
        float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        for (int rr = wib; rr < kPrepRows; rr += 8) {
            const int64_t r = r0 + rr;
            if (r >= rows) break;
            float d[4];
            load4(dOut + r * H + c, d);
            const float4 ov = *reinterpret_cast<const float4*>(o + r * H + c);
            acc[0] = fmaf(d[0], ov.x, acc[0]); acc[1] = fmaf(d[1], ov.y, acc[1]);
            acc[2] = fmaf(d[2], ov.z, acc[2]); acc[3] = fmaf(d[3], ov.w, acc[3]);
            float v[4] = {d[0] * gm.x, d[1] * gm.y, d[2] * gm.z, d[3] * gm.w};
            for (int pc = 0; pc < np; ++pc) {
                uint32_t w[2];
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    const __nv_bfloat16 a = __float2bfloat16_rn(v[2 * k]), b = __float2bfloat16_rn(v[2 * k + 1]);
                    w[k] = static_cast<uint32_t>(__bfloat16_as_ushort(a)) | (static_cast<uint32_t>(__bfloat16_as_ushort(b)) << 16);
                    v[2 * k] -= __bfloat162float(a);
                    v[2 * k + 1] -= __bfloat162float(b);
                }
                *reinterpret_cast<uint2*>(do_p + pc * total + r * H + c) = make_uint2(w[0], w[1]);
            }
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) red[wib][c + e] = acc[e];
    }
    __syncthreads();
    for (int c = threadIdx.x; c < H; c += 256) {
        float sum = 0.0f;
#pragma unroll
        for (int w8 = 0; w8 < 8; ++w8) sum += red[w8][c];
        part[static_cast<int64_t>(blockIdx.x) * H + c] = sum;
    }
}

// bf16 3-D map over pieces: dims {cols, N, n_batches}, box {64, box_rows, 1}, SWIZZLE_128B, zero fill out of bounds
// (columns past `cols` -- the score operands are 16 / 32 wide -- and rows past the graph's N).
int make_tmap_graph_rows(CUtensorMap* out, const void* ptr, int64_t n_batches, int64_t N, int64_t cols, int box_rows) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return GRL_E_UNSUP;
    }
    const cuuint64_t dims[3] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(N), static_cast<cuuint64_t>(n_batches)};
    const cuuint64_t strides[2] = {static_cast<cuuint64_t>(cols) * 2, static_cast<cuuint64_t>(N) * cols * 2};
    const cuuint32_t box[3] = {64, static_cast<cuuint32_t>(box_rows), 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(3d) failed (%d): ptr=%p batches=%lld N=%lld cols=%lld box_rows=%d", static_cast<int>(r),
                  ptr, static_cast<long long>(n_batches), static_cast<long long>(N), static_cast<long long>(cols), box_rows);
        return GRL_E_BADARG;
    }
    return 0;
}

template <typename K>
int set_smem(K kern, int bytes, const char* what) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return cuda_fail(e, what);
    return 0;
}

template <int NP, int H, int D>
int launch_fwd(const AttnTcParams& p, const void* f_p, const void* g_p, const void* h_p, cudaStream_t st) {
    CUtensorMap tF, tG, tH;
    int rc;
    if ((rc = make_tmap_graph_rows(&tF, f_p, static_cast<int64_t>(NP) * p.B, p.N, D, 128))) return rc;
    if ((rc = make_tmap_graph_rows(&tG, g_p, static_cast<int64_t>(NP) * p.B, p.N, D, 128))) return rc;
    if ((rc = make_tmap_graph_rows(&tH, h_p, static_cast<int64_t>(NP) * p.B, p.N, H, 128))) return rc;
    constexpr int SMEM = (2 * NP + NP * (H / 64)) * 16384 + 1024 + 64 + kFwdGroups * 128 * 4;
    static_assert(SMEM <= 227 * 1024, "forward shared memory");
    auto kern = attn_tc_fwd_kernel<NP, H, D>;
    static bool attr = false;
    if (!attr) {
        if ((rc = set_smem(kern, SMEM, "cudaFuncSetAttribute(attn_tc_fwd_kernel)"))) return rc;
        attr = true;
    }
    kern<<<dim3((p.N + 127) / 128, p.B), 128 * kFwdGroups, SMEM, st>>>(p, tF, tG, tH);
    GRL_LAUNCH_CHECK("attn_tc_fwd_kernel");
    return 0;
}

template <int NPS, int NP, int H, int D, int BQ>
int launch_bwd(const AttnTcParams& p, const void* f_p, const void* g_p, const void* h_p, const void* do_p, cudaStream_t st) {
    CUtensorMap tF, tG, tH, tDO, tFq, tDOq;
    int rc;
    const int64_t nbS = static_cast<int64_t>(NPS) * p.B, nb = static_cast<int64_t>(NP) * p.B;
    if ((rc = make_tmap_graph_rows(&tF, f_p, nbS, p.N, D, 128))) return rc;
    if ((rc = make_tmap_graph_rows(&tG, g_p, nbS, p.N, D, 128))) return rc;
    if ((rc = make_tmap_graph_rows(&tH, h_p, nb, p.N, H, 128))) return rc;
    if ((rc = make_tmap_graph_rows(&tDO, do_p, nb, p.N, H, 128))) return rc;
    if ((rc = make_tmap_graph_rows(&tFq, f_p, nbS, p.N, D, BQ))) return rc;
    if ((rc = make_tmap_graph_rows(&tDOq, do_p, nb, p.N, H, BQ))) return rc;
    constexpr int HS = H / 64;
    constexpr int SMEM_Q = 2 * (NPS + NP * HS) * 16384 + 1024 + 64 + kBwdGroups * 128 * 4;
    constexpr int SMEM_K = (NPS + NP * HS) * (16384 + BQ * 128) + 3 * BQ * 4 + 1024 + 64;
    static_assert(SMEM_Q <= 227 * 1024 && SMEM_K <= 227 * 1024, "backward shared memory");
    auto kq = attn_tc_bwd_q_kernel<NPS, NP, H, D>;
    auto kk = attn_tc_bwd_k_kernel<NPS, NP, H, D, BQ>;
    static bool attr = false;
    if (!attr) {
        if ((rc = set_smem(kq, SMEM_Q, "cudaFuncSetAttribute(attn_tc_bwd_q_kernel)"))) return rc;
        if ((rc = set_smem(kk, SMEM_K, "cudaFuncSetAttribute(attn_tc_bwd_k_kernel)"))) return rc;
        attr = true;
    }
    const dim3 grid((p.N + 127) / 128, p.B);
    kq<<<grid, 128 * kBwdGroups, SMEM_Q, st>>>(p, tF, tG, tH, tDO);
    GRL_LAUNCH_CHECK("attn_tc_bwd_q_kernel");
    kk<<<grid, 128 * kBwdGroups, SMEM_K, st>>>(p, tFq, tG, tH, tDOq);
    GRL_LAUNCH_CHECK("attn_tc_bwd_k_kernel");
    return 0;
}

}  // namespace
}  // namespace grl

using namespace grl;

// pieces used by the forward (all operands) and by the backward (score product / the H-wide products)
static bool tc_config(int io_dtype, int H, int D, int* np_fwd, int* nps_bwd, int* np_bwd) {
    if (io_dtype == GRL_BF16 && ((H == 128 && D == 16) || (H == 256 && D == 32))) {
        *np_fwd = 1; *nps_bwd = 1; *np_bwd = 1;
        return true;
    }
    if (io_dtype == GRL_F32 && H == 128 && D == 16) {
        *np_fwd = 3; *nps_bwd = 3; *np_bwd = 2;
        return true;
    }
    return false;
}

extern "C" int grl_selfatten_tc_pieces(int io_dtype, int H, int D, int* np_fwd, int* nps_bwd, int* np_bwd) {
    int a = 0, b = 0, c = 0;
    const bool ok = tc_config(io_dtype, H, D, &a, &b, &c);
    if (np_fwd) *np_fwd = a;
    if (nps_bwd) *nps_bwd = b;
    if (np_bwd) *np_bwd = c;
    return ok ? 1 : 0;
}

extern "C" int grl_attn_split(const void* x, int64_t ld, int dtype, int64_t rows, int C, const float* scale, int np,
                              void* pieces, void* stream) {
    GRL_CHECK_ARG(x && pieces, GRL_E_BADARG, "grl_attn_split: null pointer");
    GRL_CHECK_ARG(rows > 0 && C > 0 && ld >= C && np >= 1 && np <= 3, GRL_E_BADARG, "grl_attn_split: bad shape");
    int64_t blocks = (rows * C + 255) / 256;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    __nv_bfloat16* out = static_cast<__nv_bfloat16*>(pieces);
    const int es = dtype == GRL_F32 ? 4 : 2;
    if ((dtype == GRL_F32 || dtype == GRL_BF16) && C % 8 == 0 && (ld * es) % 16 == 0 && aligned_to(x, 16) && aligned_to(out, 16)) {
        int64_t b8 = (rows * (C / 8) + 255) / 256;
        if (b8 > kNumSMs * 16) b8 = kNumSMs * 16;
        if (dtype == GRL_F32)
            attn_split_vec8_kernel<float><<<static_cast<int>(b8), 256, 0, st>>>(static_cast<const float*>(x), ld, rows, C, scale, np, out);
        else
            attn_split_vec8_kernel<__nv_bfloat16><<<static_cast<int>(b8), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), ld, rows, C,
                                                                                      scale, np, out);
        GRL_LAUNCH_CHECK("attn_split_vec8_kernel");
        return 0;
    }
    if (dtype == GRL_F32)
        attn_split_kernel<float><<<static_cast<int>(blocks), 256, 0, st>>>(static_cast<const float*>(x), ld, rows, C, scale, np, out);
    else if (dtype == GRL_BF16)
        attn_split_kernel<__nv_bfloat16><<<static_cast<int>(blocks), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), ld, rows, C, scale, np, out);
    else {
        set_error("grl_attn_split: dtype %d unsupported", dtype);
        return GRL_E_UNSUP;
    }
    GRL_LAUNCH_CHECK("attn_split_kernel");
    return 0;
}

extern "C" int grl_selfatten_tc_fwd(const void* f_p, const void* g_p, const void* h_p, const void* V, const float* gamma,
                                    void* out, float* o_save, float* m_save, float* l_save, int io_dtype, int B, int N,
                                    int H, int D, void* stream) {
    GRL_CHECK_ARG(f_p && g_p && h_p && V && gamma && out, GRL_E_BADARG, "grl_selfatten_tc_fwd: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0, GRL_E_BADARG, "grl_selfatten_tc_fwd: bad shape");
    GRL_CHECK_ARG((o_save != nullptr) == (m_save != nullptr) && (m_save != nullptr) == (l_save != nullptr), GRL_E_BADARG,
                  "grl_selfatten_tc_fwd: o_save / m_save / l_save go together");
    GRL_CHECK_ARG(aligned_to(f_p, 16) && aligned_to(g_p, 16) && aligned_to(h_p, 16) && aligned_to(V, 16) && aligned_to(out, 16) &&
                      aligned_to(o_save, 16), GRL_E_ALIGN, "grl_selfatten_tc_fwd: 16-byte aligned pointers");
    int npf, nps, npb;
    GRL_CHECK_ARG(tc_config(io_dtype, H, D, &npf, &nps, &npb), GRL_E_UNSUP,
                  "grl_selfatten_tc_fwd: (dtype %d, H %d, D %d) has no tensor-core instantiation", io_dtype, H, D);
    AttnTcParams p{};
    p.B = B; p.N = N; p.io_bf16 = io_dtype == GRL_BF16; p.gamma = gamma; p.V = V; p.out = out;
    p.o_save = o_save; p.m_save = m_save; p.l_save = l_save;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (io_dtype == GRL_BF16) return H == 128 ? launch_fwd<1, 128, 16>(p, f_p, g_p, h_p, st) : launch_fwd<1, 256, 32>(p, f_p, g_p, h_p, st);
    return launch_fwd<3, 128, 16>(p, f_p, g_p, h_p, st);
}

extern "C" size_t grl_selfatten_tc_prep_blocks(int64_t rows) { return static_cast<size_t>((rows + kPrepRows - 1) / kPrepRows); }

extern "C" int grl_selfatten_tc_bwd(const void* dOut, const float* o_save, const float* m_save, const float* l_save,
                                    const void* f_p, const void* g_p, const void* h_p, const float* gamma, void* do_p,
                                    float* Dvec, float* part, void* df, void* dg, void* dh, int io_dtype, int B, int N,
                                    int H, int D, void* stream) {
    GRL_CHECK_ARG(dOut && o_save && m_save && l_save && f_p && g_p && h_p && gamma && do_p && Dvec && part && df && dg && dh,
                  GRL_E_BADARG, "grl_selfatten_tc_bwd: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0, GRL_E_BADARG, "grl_selfatten_tc_bwd: bad shape");
    int npf, nps, npb;
    GRL_CHECK_ARG(tc_config(io_dtype, H, D, &npf, &nps, &npb), GRL_E_UNSUP,
                  "grl_selfatten_tc_bwd: (dtype %d, H %d, D %d) has no tensor-core instantiation", io_dtype, H, D);
    GRL_CHECK_ARG(aligned_to(do_p, 16) && aligned_to(df, 16) && aligned_to(dg, 16) && aligned_to(dh, 16) && aligned_to(dOut, 16) &&
                      aligned_to(o_save, 16), GRL_E_ALIGN, "grl_selfatten_tc_bwd: 16-byte aligned pointers");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t rows = static_cast<int64_t>(B) * N;
    const int blocks = static_cast<int>((rows + kPrepRows - 1) / kPrepRows);
    if (io_dtype == GRL_BF16)
        attn_bwd_prep_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(dOut), o_save, gamma, rows, H,
                                                                    npb, static_cast<__nv_bfloat16*>(do_p), part);
    else
        attn_bwd_prep_kernel<float><<<blocks, 256, 0, st>>>(static_cast<const float*>(dOut), o_save, gamma, rows, H, npb,
                                                            static_cast<__nv_bfloat16*>(do_p), part);
    GRL_LAUNCH_CHECK("attn_bwd_prep_kernel");
    AttnTcParams p{};
    p.B = B; p.N = N; p.io_bf16 = io_dtype == GRL_BF16; p.gamma = gamma;
    p.m_save = const_cast<float*>(m_save); p.l_save = const_cast<float*>(l_save); p.Dvec = Dvec;
    p.df = df; p.dg = dg; p.dh = dh;
    if (io_dtype == GRL_BF16)
        return H == 128 ? launch_bwd<1, 1, 128, 16, 128>(p, f_p, g_p, h_p, do_p, st)
                        : launch_bwd<1, 1, 256, 32, 64>(p, f_p, g_p, h_p, do_p, st);
    return launch_bwd<3, 2, 128, 16, 128>(p, f_p, g_p, h_p, do_p, st);
}
