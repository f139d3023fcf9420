// Shared-memory slab SpMM, second generation (forward aggregation; the default kernel behind grl_spmm_fwd).
//
// Reference semantics: new_V = torch.matmul(A', V) (gnn/models/networks/robust_gcn.py:45-47), A' already DropEdge-masked
// (gnn/models/networks/drop_robust_gcn.py:76,80,85).
//
// The first slab kernel (spmm_tile.cu) was issue-bound: ~95 instructions of bookkeeping per typed row and 11.5 per
// entry step (ncu: 333 M warp instructions, 72 % issue-slot utilisation, 0.42 ms for 1.10 GB of compulsory traffic).
// This kernel keeps its data movement -- one CTA stages the [N, FC] feature slab of one graph with TMA and every gather
// is a conflict-free LDS.128 -- and restructures the walk:
//
//   * NODE-level entry stream.  All (L+1) typed rows of a node are contiguous in CSR, so a lane group loads the node's
//     entries in one go (coalesced, prefetched one node ahead), tags every entry with its row type, and the walk runs
//     over the node's stream with ONE compare per entry to detect "this entry opens the next typed row".
//   * Dropped entries are compacted away before the walk (ballot + popc rank), so the 30 % of the entries DropEdge
//     removed cost one predicate in the prologue instead of a walk step.  (Entries are summed in CSR order with the
//     dropped ones skipped -- exactly what the gather kernels do, so all forward kernels stay bit-identical.)
//   * The compacted stream lives in a per-warp shared-memory staging area and is read back with broadcast LDS.64:
//     no shuffles in the walk.  An entry is (t << 24 | col * ROW_BYTES, weight): "new row" is `x >= (t_open+1) << 24`
//     and the gather address is `S + x` with S = lane's slab base - (t_open << 24), maintained when a row opens.
//     Walk step: LDS.64, ISETP, BRA, IADD, LDS.128, 4 FFMA.
//   * Persistent CTAs over (graph, column chunk) items; when two slabs fit, the next item's slab is fetched by TMA while
//     the current one is walked (the fill of a 125 KB slab is ~15 % of an item's time otherwise).
#include "gemm_common.cuh"

namespace grl {
using namespace sm100;

struct Slab2Params {
    const int32_t* rowptr;
    const int32_t* col;
    const float* w;
    uint8_t* Y;            // bytes
    int64_t ldy_bytes;
    int N, Lp1, n_chunks, box_rows, n_boxes, n_items, n_buf;
    uint32_t slab_bytes;   // one slab buffer (multiple of 128)
};

template <typename T> struct Acc16v2;   // one 16-byte slab word as fp32 accumulators
template <> struct Acc16v2<float> {
    static constexpr int E = 4;
    __device__ static __forceinline__ void fma(float (&acc)[4], float wt, const uint4& q) {
        acc[0] = fmaf(wt, __uint_as_float(q.x), acc[0]); acc[1] = fmaf(wt, __uint_as_float(q.y), acc[1]);
        acc[2] = fmaf(wt, __uint_as_float(q.z), acc[2]); acc[3] = fmaf(wt, __uint_as_float(q.w), acc[3]);
    }
    __device__ static __forceinline__ uint4 pack(const float (&acc)[4]) {
        return make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
    }
};
template <> struct Acc16v2<__nv_bfloat16> {
    static constexpr int E = 8;
    __device__ static __forceinline__ void fma(float (&acc)[8], float wt, const uint4& q) {
        const uint32_t ww[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            acc[2 * i] = fmaf(wt, __uint_as_float(ww[i] << 16), acc[2 * i]);
            acc[2 * i + 1] = fmaf(wt, __uint_as_float(ww[i] & 0xFFFF0000u), acc[2 * i + 1]);
        }
    }
    __device__ static __forceinline__ uint4 pack(const float (&acc)[8]) {
        uint32_t o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(acc[2 * i], acc[2 * i + 1]);
            o[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        return make_uint4(o[0], o[1], o[2], o[3]);
    }
};

__device__ __forceinline__ uint2 lds64(uint32_t addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts64(uint32_t addr, uint32_t x, uint32_t y) {
    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(x), "r"(y) : "memory");
}

// GL lanes per node group (16 bytes per lane: ROW_BYTES = GL * 16 per slab row); J entry rounds per batch (CAP = GL*J).
template <int GL> struct Slab2Cfg {
    static constexpr int J = GL == 32 ? 3 : (GL == 16 ? 6 : 8);
    static constexpr int CAP = GL * J;
    static constexpr int G = 32 / GL;
    static constexpr int STAGE_SLOTS = CAP + 4;                 // + terminator, + padding of the walk length to even
    static constexpr int STAGE_BYTES_PER_WARP = G * STAGE_SLOTS * 8;
};

template <typename T, int GL>
__global__ void __launch_bounds__(1024, 1)
spmm_slab2_kernel(const Slab2Params p, const __grid_constant__ CUtensorMap tmX) {
    using A16 = Acc16v2<T>;
    using Cfg = Slab2Cfg<GL>;
    constexpr int E = A16::E;
    constexpr int G = Cfg::G, J = Cfg::J, CAP = Cfg::CAP;
    constexpr uint32_t ROW_BYTES = GL * 16;            // one slab row = the chunk's FC columns
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t bar[2];

    // buffer b: [zero row][slab rows...] at base + b * slab_bytes; staging areas behind the buffers
    const uint32_t base_u32 = (smem_u32(smem_raw) + 127u) & ~127u;
    uint8_t* const base_ptr = smem_raw + (base_u32 - smem_u32(smem_raw));
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const int grp = lane / GL, gl = lane % GL;
    const uint32_t stage_g = base_u32 + p.n_buf * p.slab_bytes + (warp * G + grp) * (Cfg::STAGE_SLOTS * 8);
    const uint32_t gshift = grp * GL;
    const uint32_t gfull = GL == 32 ? 0xFFFFFFFFu : ((1u << GL) - 1u);
    const uint32_t glt = (1u << gl) - 1u;

    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_mbar_init();
    }
    for (int q = threadIdx.x; q < p.n_buf * (int)(ROW_BYTES / 16); q += blockDim.x) {   // the zero rows (fillers point here)
        const int bsel = q / (int)(ROW_BYTES / 16), r16 = q % (int)(ROW_BYTES / 16);
        *reinterpret_cast<uint4*>(base_ptr + bsel * p.slab_bytes + r16 * 16) = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();

    auto issue_fill = [&](int item, int bsel) {
        const int chunk = item % p.n_chunks, b = item / p.n_chunks;
        mbar_expect_tx(&bar[bsel], static_cast<uint32_t>(p.n_boxes) * p.box_rows * ROW_BYTES);
        uint8_t* dst = base_ptr + bsel * p.slab_bytes + ROW_BYTES;
        for (int k = 0; k < p.n_boxes; ++k)
            tma_load_2d(dst + static_cast<size_t>(k) * p.box_rows * ROW_BYTES, &tmX, &bar[bsel],
                        chunk * static_cast<int>(ROW_BYTES / sizeof(T)), b * p.N + k * p.box_rows);
    };

    const int Lp1 = p.Lp1;
    const int step = n_warps * G;
    int buf = 0;
    uint32_t phases = 0;
    int item = blockIdx.x;
    if (threadIdx.x == 0 && item < p.n_items) issue_fill(item, 0);

    int32_t rc[J];
    float rw[J];
    auto load_raw = [&](int32_t ebeg, int32_t eend) {
#pragma unroll
        for (int j = 0; j < J; ++j) {
            const int32_t e = ebeg + gl + GL * j;
            rc[j] = 0;
            rw[j] = 0.0f;
            if (e < eend) {
                rc[j] = __ldg(p.col + e);
                rw[j] = __ldg(p.w + e);
            }
        }
    };

    for (; item < p.n_items; item += gridDim.x) {
        const int nxt = item + gridDim.x;
        if (p.n_buf == 2 && threadIdx.x == 0 && nxt < p.n_items) issue_fill(nxt, buf ^ 1);
        const int chunk = item % p.n_chunks, b = item / p.n_chunks;
        const int64_t node0 = static_cast<int64_t>(b) * p.N;
        const uint32_t slab_lane = base_u32 + buf * p.slab_bytes + gl * 16;     // the zero row; node j is row j + 1
        uint8_t* const ychunk = p.Y + static_cast<size_t>(chunk) * ROW_BYTES + gl * 16;
        auto load_rp = [&](int ii) -> int32_t {
            return (ii < p.N && gl <= Lp1) ? __ldg(p.rowptr + (node0 + ii) * Lp1 + gl) : 0;
        };

        int i = warp * G + grp;
        int32_t rp = load_rp(i);
        int32_t e0 = __shfl_sync(0xffffffffu, rp, 0, GL);
        int32_t eN = __shfl_sync(0xffffffffu, rp, Lp1, GL);
        load_raw(e0, eN);
        int32_t rp_next = load_rp(i + step);
        mbar_wait(&bar[buf], (phases >> buf) & 1u);      // slab landed (the index loads above overlap the fill)
        phases ^= 1u << buf;

        for (; (i - grp) < p.N; i += step) {
            const bool valid = i < p.N;
            uint8_t* const ynode = ychunk + (node0 + i) * Lp1 * p.ldy_bytes;
            const int maxtotal = __reduce_max_sync(0xffffffffu, eN - e0);
            int t_open = -1;
            uint32_t thr = 0, S = slab_lane;
            float acc[E];
#pragma unroll
            for (int k = 0; k < E; ++k) acc[k] = 0.0f;
            int32_t ebase = e0;

            auto open_row = [&](int t) {
                if (valid) {
                    if (t_open >= 0)
                        __stcs(reinterpret_cast<uint4*>(ynode + t_open * p.ldy_bytes), A16::pack(acc));
                    for (int tt = t_open + 1; tt < t; ++tt)       // typed rows without a kept entry
                        __stcs(reinterpret_cast<uint4*>(ynode + tt * p.ldy_bytes), make_uint4(0, 0, 0, 0));
                }
#pragma unroll
                for (int k = 0; k < E; ++k) acc[k] = 0.0f;
                t_open = t;
                thr = static_cast<uint32_t>(t + 1) << 24;
                S = slab_lane - (static_cast<uint32_t>(t) << 24);
            };

            for (int eb = 0;; eb += CAP) {
                // ---------------- prologue: tag, compact and stage the batch held in (rc, rw)
                const bool final_warp = eb + CAP >= maxtotal;
                const bool final_g = ebase + CAP >= eN;
                const int rem = min(maxtotal - eb, CAP);
                uint32_t tj[J];
#pragma unroll
                for (int j = 0; j < J; ++j) tj[j] = 0;
                for (int t = 1; t < Lp1; ++t) {
                    const int32_t bt = __shfl_sync(0xffffffffu, rp, t, GL) - ebase - gl;
#pragma unroll
                    for (int j = 0; j < J; ++j)
                        if (GL * j < rem) tj[j] += (GL * j >= bt) ? 1u : 0u;
                }
                uint32_t kept = 0;
#pragma unroll
                for (int j = 0; j < J; ++j) {
                    if (GL * j < rem) {
                        const bool keep = rw[j] != 0.0f;
                        const uint32_t m = (__ballot_sync(0xffffffffu, keep) >> gshift) & gfull;
                        if (keep)
                            sts64(stage_g + (kept + __popc(m & glt)) * 8,
                                  (tj[j] << 24) | ((static_cast<uint32_t>(rc[j]) + 1u) * ROW_BYTES), __float_as_uint(rw[j]));
                        kept += __popc(m);
                    }
                }
                const uint32_t t_last = __shfl_sync(0xffffffffu, tj[J - 1], GL - 1, GL);
                const uint32_t term = (final_g ? static_cast<uint32_t>(Lp1) : t_last) << 24;
                int n = __reduce_max_sync(0xffffffffu, static_cast<int>(kept) + 1);
                n = (n + 1) & ~1;
                for (int s = kept + gl; s < n; s += GL) sts64(stage_g + s * 8, term, 0u);
                __syncwarp();

                // ---------------- refill the registers: next batch of this node, or the next node's first batch
                if (final_warp) {
                    rp = rp_next;
                    e0 = __shfl_sync(0xffffffffu, rp, 0, GL);
                    eN = __shfl_sync(0xffffffffu, rp, Lp1, GL);
                    load_raw(e0, eN);
                    rp_next = load_rp(i + 2 * step);
                } else {
                    ebase += CAP;
                    load_raw(ebase, eN);
                }

                // ---------------- walk
                uint32_t sa = stage_g;
                uint2 ew = lds64(sa);
#pragma unroll 1
                for (int k = 0; k < n; k += 2) {
                    const uint2 ew1 = lds64(sa + 8);
                    const uint2 ew2 = lds64(sa + 16);
                    sa += 16;
                    if (ew.x >= thr) open_row(static_cast<int>(ew.x >> 24));
                    A16::fma(acc, __uint_as_float(ew.y), lds128(S + ew.x));
                    if (ew1.x >= thr) open_row(static_cast<int>(ew1.x >> 24));
                    A16::fma(acc, __uint_as_float(ew1.y), lds128(S + ew1.x));
                    ew = ew2;
                }
                __syncwarp();                                // staging is rewritten by the next prologue
                if (final_warp) break;
            }
        }
        __syncthreads();                                     // every warp is done with slab[buf]
        if (p.n_buf == 1 && threadIdx.x == 0 && nxt < p.n_items) issue_fill(nxt, 0);
        if (p.n_buf == 2) buf ^= 1;
    }
}

template <typename T, int GL>
static int launch_slab2(const Slab2Params& p, const CUtensorMap& tm, int grid, size_t smem, int threads, cudaStream_t st) {
    auto kern = spmm_slab2_kernel<T, GL>;
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(spmm_slab2_kernel)");
        attr_smem = smem;
    }
    kern<<<grid, threads, smem, st>>>(p, tm);
    GRL_LAUNCH_CHECK("spmm_slab2_kernel");
    return 0;
}

// returns 0 = launched, 1 = not applicable (caller tries the next kernel), other = error.
// gl_req: 0 = choose the group width, 8 / 16 / 32 = force it.
int try_launch_spmm_slab2(const int32_t* rowptr, const int32_t* col, const float* w, const void* X, int64_t ldx, void* Y,
                          int64_t ldy, int dtype, int B, int N, int L, int F, int gl_req, cudaStream_t st) {
    const int es = dtype == GRL_F32 ? 4 : 2;
    if (!aligned_to(X, 16) || (ldx * es) % 16 != 0 || !aligned_to(Y, 16) || (ldy * es) % 16 != 0) return 1;
    if (L + 1 >= 127) return 1;                                   // row type lives in 7 bits of the staged entry
    const int n_boxes = (N + 255) / 256;
    const int box_rows = (N + n_boxes - 1) / n_boxes;
    const int rows = n_boxes * box_rows;
    constexpr size_t kMaxSmem = 227 * 1024 - 1024;

    struct Plan { int gl, threads, n_buf; size_t smem; uint32_t slab_bytes; };
    auto plan_for = [&](int gl, Plan* out) -> bool {
        if (L + 2 > gl || F % (gl * 16 / es) != 0) return false;
        const int G = 32 / gl;
        int threads = 1024;
        while (threads > 32 && (threads / 32) * G > N) threads >>= 1;              // tiny graphs: no idle warps
        const int stage_slots = (gl == 32 ? 96 : (gl == 16 ? 96 : 64)) + 4;
        const size_t stage = static_cast<size_t>(threads / 32) * G * stage_slots * 8;
        const size_t slab = static_cast<size_t>(rows + 1) * gl * 16;                // + the zero row
        if (static_cast<uint64_t>(rows + 1) * gl * 16 >= (1u << 24)) return false;
        const size_t one = slab + stage + 128, two = 2 * slab + stage + 128;
        if (one > kMaxSmem) return false;
        out->gl = gl; out->threads = threads; out->slab_bytes = static_cast<uint32_t>(slab);
        out->n_buf = two <= kMaxSmem ? 2 : 1;
        out->smem = out->n_buf == 2 ? two : one;
        return true;
    };
    Plan pl{};
    bool have = false;
    if (gl_req) {
        have = plan_for(gl_req, &pl);
    } else {
        // widest group whose slab can be double-buffered, else the widest that fits at all
        Plan cand{};
        for (int gl : {32, 16, 8})
            if (plan_for(gl, &cand) && cand.n_buf == 2) { pl = cand; have = true; break; }
        if (!have)
            for (int gl : {32, 16, 8})
                if (plan_for(gl, &cand)) { pl = cand; have = true; break; }
    }
    if (!have) return 1;
    const int n_chunks = F / (pl.gl * 16 / es);
    const int n_items = B * n_chunks;
    int per_sm = static_cast<int>((228 * 1024) / (pl.smem + 1024));
    if (per_sm > 2048 / pl.threads) per_sm = 2048 / pl.threads;
    if (per_sm < 1) per_sm = 1;
    const int grid = n_items < kNumSMs * per_sm ? n_items : kNumSMs * per_sm;

    CUtensorMap tm;
    int rc = make_tmap_plain(&tm, X, es, static_cast<int64_t>(B) * N, F, ldx, pl.gl * 16 / es, box_rows);
    if (rc) return rc;
    Slab2Params p{rowptr, col, w, static_cast<uint8_t*>(Y), ldy * es, N, L + 1, n_chunks, box_rows, n_boxes, n_items,
                  pl.n_buf, pl.slab_bytes};
#define GRL_SLAB2_CASE(TT, V) \
    if (pl.gl == V) return launch_slab2<TT, V>(p, tm, grid, pl.smem, pl.threads, st)
    if (dtype == GRL_F32) {
        GRL_SLAB2_CASE(float, 8); GRL_SLAB2_CASE(float, 16); GRL_SLAB2_CASE(float, 32);
    } else {
        GRL_SLAB2_CASE(__nv_bfloat16, 8); GRL_SLAB2_CASE(__nv_bfloat16, 16); GRL_SLAB2_CASE(__nv_bfloat16, 32);
    }
#undef GRL_SLAB2_CASE
    return 1;
}

}  // namespace grl
