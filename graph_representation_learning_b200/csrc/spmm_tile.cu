// Shared-memory-staged segmented SpMM (forward aggregation) -- the fast path of grl_spmm_fwd.
//
// Reference semantics: new_V = torch.matmul(A', V) (gnn/models/networks/robust_gcn.py:45-47), A' already DropEdge-masked.
//
// Why: every stored entry gathers one neighbour row of F features.  From global memory that is nnz * F * s bytes of
// L2 -> SM traffic (cfg2: 6.4 GB for a kernel whose compulsory HBM traffic is 1.1 GB), and the L2 fabric, not HBM,
// becomes the bound.  A document graph is small (N = 200 .. 1000 nodes), so ONE CTA stages the [N, FC] feature slab of
// ONE graph (FC = a column chunk of the features) in shared memory with TMA and all gathers hit shared memory; global
// traffic drops to the compulsory bytes: X read once, indices once per chunk (L2-resident across the chunk CTAs of a
// graph, which are adjacent in the grid), new_V written once with full 128 B+ segments.
//
// Work split: CTA = (graph b, feature chunk of FC columns).  The issue slots, not shared-memory bandwidth, bound the
// walk over the entries (two shuffles + address + load + FMAs per entry), so every lane moves the widest word there
// is, 16 bytes, and a warp is cut into G = 32/GL groups of GL lanes (GL*16 bytes = one slab row) that each own one
// node: one issued instruction advances G entries.  A group walks the (L+1) typed rows of its node in lock step with
// the other groups of the warp (trip count = the longest of the G rows); the (column, weight) pairs of the next
// row are prefetched while the current one is accumulated; a finished row leaves with one 16-byte streaming store per
// lane (GL*16 contiguous bytes per group).
#include "gemm_common.cuh"

namespace grl {
using namespace sm100;

extern int g_spmm_mode;   // spmm.cu (grl_spmm_set_tile_mode)

struct SpmmSlabParams {
    const int32_t* rowptr;
    const int32_t* col;
    const float* w;
    uint8_t* Y;            // bytes
    int64_t ldy_bytes;
    int N, Lp1, n_chunks, box_rows, n_boxes;
};

template <typename T> struct Acc16;   // one 16-byte slab word as fp32 accumulators
template <> struct Acc16<float> {
    static constexpr int E = 4;
    __device__ static inline void fma(float (&acc)[4], float wt, const uint4& q) {
        acc[0] = fmaf(wt, __uint_as_float(q.x), acc[0]); acc[1] = fmaf(wt, __uint_as_float(q.y), acc[1]);
        acc[2] = fmaf(wt, __uint_as_float(q.z), acc[2]); acc[3] = fmaf(wt, __uint_as_float(q.w), acc[3]);
    }
    __device__ static inline uint4 pack(const float (&acc)[4]) {
        return make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]), __float_as_uint(acc[3]));
    }
};
template <> struct Acc16<__nv_bfloat16> {
    static constexpr int E = 8;
    __device__ static inline void fma(float (&acc)[8], float wt, const uint4& q) {
        const uint32_t ww[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            acc[2 * i] = fmaf(wt, __uint_as_float(ww[i] << 16), acc[2 * i]);
            acc[2 * i + 1] = fmaf(wt, __uint_as_float(ww[i] & 0xFFFF0000u), acc[2 * i + 1]);
        }
    }
    __device__ static inline uint4 pack(const float (&acc)[8]) {
        uint32_t o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(acc[2 * i], acc[2 * i + 1]);
            o[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        return make_uint4(o[0], o[1], o[2], o[3]);
    }
};

template <typename T, int GL>
__global__ void __launch_bounds__(1024, 1)
spmm_slab_kernel(const SpmmSlabParams p, const __grid_constant__ CUtensorMap tmX) {
    using A16 = Acc16<T>;
    constexpr int E = A16::E;
    constexpr int G = 32 / GL;
    constexpr int ROW_BYTES = GL * 16;                  // one slab row = the chunk's FC columns
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 127u) & ~127u;
    uint8_t* slab = smem_raw + (base_u32 - smem_u32(smem_raw));
    __shared__ uint64_t bar;

    const int chunk = blockIdx.x % p.n_chunks;
    const int b = blockIdx.x / p.n_chunks;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const int grp = lane / GL, gl = lane % GL;

    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
        mbar_expect_tx(&bar, static_cast<uint32_t>(p.n_boxes) * p.box_rows * ROW_BYTES);
        for (int k = 0; k < p.n_boxes; ++k)
            tma_load_2d(slab + static_cast<size_t>(k) * p.box_rows * ROW_BYTES, &tmX, &bar, chunk * (ROW_BYTES / (int)sizeof(T)),
                        b * p.N + k * p.box_rows);
    }
    __syncthreads();   // barrier initialised before anyone polls it

    const int64_t node0 = static_cast<int64_t>(b) * p.N;
    const int Lp1 = p.Lp1;
    uint8_t* ylane = p.Y + static_cast<size_t>(chunk) * ROW_BYTES + gl * 16;
    const uint8_t* slab_lane = slab + gl * 16;
    const int node_step = n_warps * G;

    // (column, weight) pairs of one typed row, two per lane (covers 2*GL entries; longer rows spill to a slow tail)
    auto load_entries = [&](int32_t beg, int cnt, uint32_t& o0, float& w0, uint32_t& o1, float& w1) {
        o0 = 0; o1 = 0; w0 = 0.0f; w1 = 0.0f;
        if (gl < cnt) {
            o0 = static_cast<uint32_t>(__ldg(p.col + beg + gl)) * ROW_BYTES;
            w0 = __ldg(p.w + beg + gl);
        }
        if (GL + gl < cnt) {
            o1 = static_cast<uint32_t>(__ldg(p.col + beg + GL + gl)) * ROW_BYTES;
            w1 = __ldg(p.w + beg + GL + gl);
        }
    };

    int i = warp * G + grp;                              // this group's node; all groups of a warp iterate together
    bool valid = i < p.N;
    int32_t rp = (valid && gl <= Lp1) ? __ldg(p.rowptr + (node0 + i) * Lp1 + gl) : 0;
    mbar_wait(&bar, 0);                                  // slab landed (the row-pointer load above overlaps the fill)

    for (; (i - grp) < p.N; i += node_step) {
        const int64_t r0 = (node0 + i) * Lp1;
        // prefetch the next node's row pointers
        const int inext = i + node_step;
        const bool vnext = inext < p.N;
        const int32_t rp_next = (vnext && gl <= Lp1) ? __ldg(p.rowptr + (node0 + inext) * Lp1 + gl) : 0;

        int32_t beg = __shfl_sync(0xffffffffu, rp, 0, GL);
        int32_t end = __shfl_sync(0xffffffffu, rp, 1, GL);
        uint32_t o0, o1;
        float w0, w1;
        load_entries(beg, end - beg, o0, w0, o1, w1);
        uint8_t* yrow = ylane + r0 * p.ldy_bytes;
        for (int t = 0; t < Lp1; ++t, yrow += p.ldy_bytes) {
            const int cnt = end - beg;
            const int32_t cur_beg = beg;
            const uint32_t co0 = o0, co1 = o1;
            const float cw0 = w0, cw1 = w1;
            if (t + 1 < Lp1) {                            // prefetch the next typed row's entries
                beg = end;
                end = __shfl_sync(0xffffffffu, rp, t + 2, GL);
                load_entries(beg, end - beg, o0, w0, o1, w1);
            }
            const int cmax = min(__reduce_max_sync(0xffffffffu, cnt), 2 * GL);
            float acc[E];
#pragma unroll
            for (int k = 0; k < E; ++k) acc[k] = 0.0f;
            const int lim0 = min(cmax, GL);
#pragma unroll 2
            for (int kk = 0; kk < lim0; ++kk) {
                const uint32_t off = __shfl_sync(0xffffffffu, co0, kk, GL);
                const float wt = __shfl_sync(0xffffffffu, cw0, kk, GL);
                if (wt != 0.0f) A16::fma(acc, wt, *reinterpret_cast<const uint4*>(slab_lane + off));
            }
#pragma unroll 2
            for (int kk = GL; kk < cmax; ++kk) {
                const uint32_t off = __shfl_sync(0xffffffffu, co1, kk - GL, GL);
                const float wt = __shfl_sync(0xffffffffu, cw1, kk - GL, GL);
                if (wt != 0.0f) A16::fma(acc, wt, *reinterpret_cast<const uint4*>(slab_lane + off));
            }
            for (int k = 2 * GL; k < cnt; ++k) {          // rare: a typed row with more than 2*GL entries
                const float wt = __ldg(p.w + cur_beg + k);
                if (wt != 0.0f)
                    A16::fma(acc, wt, *reinterpret_cast<const uint4*>(
                                          slab_lane + static_cast<uint32_t>(__ldg(p.col + cur_beg + k)) * ROW_BYTES));
            }
            if (valid) __stcs(reinterpret_cast<uint4*>(yrow), A16::pack(acc));
        }
        rp = rp_next;
        valid = vnext;
    }
}

template <typename T, int GL>
static int launch_slab(const SpmmSlabParams& p, const CUtensorMap& tm, int B, size_t smem, int threads, cudaStream_t st) {
    auto kern = spmm_slab_kernel<T, GL>;
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(spmm_slab_kernel)");
        attr_smem = smem;
    }
    kern<<<B * p.n_chunks, threads, smem, st>>>(p, tm);
    GRL_LAUNCH_CHECK("spmm_slab_kernel");
    return 0;
}

// returns 0 = launched, 1 = not applicable (caller falls back to the gather kernel), other = error
// mode: 0 = pick the group width automatically; 8 / 16 / 32 = lanes per node group
int try_launch_spmm_tile(const int32_t* rowptr, const int32_t* col, const float* w, const void* X, int64_t ldx, void* Y,
                         int64_t ldy, int dtype, int B, int N, int L, int F, cudaStream_t st) {
    const int mode = g_spmm_mode;
    if (mode < 0) return 1;
    const int es = dtype == GRL_F32 ? 4 : 2;
    if (!aligned_to(X, 16) || (ldx * es) % 16 != 0 || !aligned_to(Y, 16) || (ldy * es) % 16 != 0) return 1;
    const int n_boxes = (N + 255) / 256;
    const int box_rows = (N + n_boxes - 1) / n_boxes;
    const int rows = n_boxes * box_rows;
    constexpr size_t kMaxSmem = 227 * 1024 - 1024;
    auto smem_of = [&](int gl) { return static_cast<size_t>(rows) * gl * 16 + 128; };
    auto ok = [&](int gl) { return L + 2 <= gl && F % (gl * 16 / es) == 0 && smem_of(gl) <= kMaxSmem; };
    int gl = 0;
    if (mode > 0) {
        if (ok(mode)) gl = mode;
    } else {
        // measured (scripts/spmm_bench.py, cfg2): GL = 16 beats the L2 gather kernel by 10 %, GL = 8 loses to it; a batch
        // with fewer than two CTAs per SM cannot hide the slab fill at all -> gather kernel
        if (ok(16) && static_cast<int64_t>(B) * (F / (16 * 16 / es)) >= 2 * kNumSMs) gl = 16;
    }
    if (!gl) return 1;
    const size_t smem = smem_of(gl);
    const int per_sm = static_cast<int>((228 * 1024) / (smem + 1024)) < 1 ? 1 : static_cast<int>((228 * 1024) / (smem + 1024));
    int threads = per_sm >= 3 ? 256 : (per_sm == 2 ? 512 : 1024);
    const int G = 32 / gl;
    while (threads > 32 && (threads / 32) * G > N) threads >>= 1;     // tiny graphs: no idle warps

    CUtensorMap tm;
    int rc = make_tmap_plain(&tm, X, es, static_cast<int64_t>(B) * N, F, ldx, gl * 16 / es, box_rows);
    if (rc) return rc;
    SpmmSlabParams p{rowptr, col, w, static_cast<uint8_t*>(Y), ldy * es, N, L + 1, F / (gl * 16 / es), box_rows, n_boxes};
#define GRL_SLAB_CASE(TT, V) \
    if (gl == V) return launch_slab<TT, V>(p, tm, B, smem, threads, st)
    if (dtype == GRL_F32) {
        GRL_SLAB_CASE(float, 8); GRL_SLAB_CASE(float, 16); GRL_SLAB_CASE(float, 32);
    } else {
        GRL_SLAB_CASE(__nv_bfloat16, 8); GRL_SLAB_CASE(__nv_bfloat16, 16); GRL_SLAB_CASE(__nv_bfloat16, 32);
    }
#undef GRL_SLAB_CASE
    return 1;
}

}  // namespace grl
