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
// Work split: CTA = (graph b, feature chunk); warp = one node at a time, all (L+1) typed rows of it; lane = VB bytes
// of the chunk (VB = 4, 8 or 16 -> FC = 32*VB/sizeof(T) columns).  The node's entries are contiguous in CSR, so the
// warp fetches its (L+2) row pointers with one load and up to 64 (col, weight) pairs with two coalesced loads, then
// walks them with shuffles; a typed row is flushed with one streaming store when the walk crosses its end.
#include "gemm_common.cuh"

namespace grl {
using namespace sm100;

static int g_spmm_tile_mode = 0;   // 0 auto, -1 never (gather kernel), 4/8/16 force that many bytes per lane

template <typename T, int VB> struct LaneVec;
template <> struct LaneVec<float, 4> {
    static constexpr int VE = 1;
    __device__ static inline void load(const uint8_t* p, float (&v)[1]) { v[0] = *reinterpret_cast<const float*>(p); }
    __device__ static inline void store(uint8_t* p, const float (&v)[1]) { __stcs(reinterpret_cast<float*>(p), v[0]); }
};
template <> struct LaneVec<float, 8> {
    static constexpr int VE = 2;
    __device__ static inline void load(const uint8_t* p, float (&v)[2]) {
        const float2 t = *reinterpret_cast<const float2*>(p);
        v[0] = t.x; v[1] = t.y;
    }
    __device__ static inline void store(uint8_t* p, const float (&v)[2]) {
        __stcs(reinterpret_cast<float2*>(p), make_float2(v[0], v[1]));
    }
};
template <> struct LaneVec<float, 16> {
    static constexpr int VE = 4;
    __device__ static inline void load(const uint8_t* p, float (&v)[4]) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    __device__ static inline void store(uint8_t* p, const float (&v)[4]) {
        __stcs(reinterpret_cast<float4*>(p), make_float4(v[0], v[1], v[2], v[3]));
    }
};
__device__ static inline uint32_t pack_bf16x2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
template <> struct LaneVec<__nv_bfloat16, 4> {
    static constexpr int VE = 2;
    __device__ static inline void load(const uint8_t* p, float (&v)[2]) {
        const uint32_t t = *reinterpret_cast<const uint32_t*>(p);
        v[0] = __uint_as_float(t << 16); v[1] = __uint_as_float(t & 0xFFFF0000u);
    }
    __device__ static inline void store(uint8_t* p, const float (&v)[2]) {
        __stcs(reinterpret_cast<uint32_t*>(p), pack_bf16x2(v[0], v[1]));
    }
};
template <> struct LaneVec<__nv_bfloat16, 8> {
    static constexpr int VE = 4;
    __device__ static inline void load(const uint8_t* p, float (&v)[4]) {
        const uint2 t = *reinterpret_cast<const uint2*>(p);
        v[0] = __uint_as_float(t.x << 16); v[1] = __uint_as_float(t.x & 0xFFFF0000u);
        v[2] = __uint_as_float(t.y << 16); v[3] = __uint_as_float(t.y & 0xFFFF0000u);
    }
    __device__ static inline void store(uint8_t* p, const float (&v)[4]) {
        __stcs(reinterpret_cast<uint2*>(p), make_uint2(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3])));
    }
};
template <> struct LaneVec<__nv_bfloat16, 16> {
    static constexpr int VE = 8;
    __device__ static inline void load(const uint8_t* p, float (&v)[8]) {
        const uint4 t = *reinterpret_cast<const uint4*>(p);
        const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[2 * i] = __uint_as_float(w[i] << 16);
            v[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
        }
    }
    __device__ static inline void store(uint8_t* p, const float (&v)[8]) {
        __stcs(reinterpret_cast<uint4*>(p), make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                                       pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7])));
    }
};

struct SpmmTileParams {
    const int32_t* rowptr;
    const int32_t* col;
    const float* w;
    uint8_t* Y;            // bytes
    int64_t ldy_bytes;
    int N, Lp1, n_chunks, box_rows, n_boxes;
};

template <typename T, int VB>
__global__ void __launch_bounds__(1024, 1)
spmm_tile_kernel(const SpmmTileParams p, const __grid_constant__ CUtensorMap tmX) {
    using LV = LaneVec<T, VB>;
    constexpr int VE = LV::VE;
    constexpr int ROW_BYTES = 32 * VB;                  // one slab row = the chunk's FC columns
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base_u32 = (smem_u32(smem_raw) + 127u) & ~127u;
    uint8_t* slab = smem_raw + (base_u32 - smem_u32(smem_raw));
    __shared__ uint64_t bar;

    const int chunk = blockIdx.x % p.n_chunks;
    const int b = blockIdx.x / p.n_chunks;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;

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
    uint8_t* ybase = p.Y + static_cast<size_t>(chunk) * ROW_BYTES + lane * VB;
    bool slab_ready = false;

    for (int i = warp; i < p.N; i += n_warps) {
        const int64_t r0 = (node0 + i) * p.Lp1;
        const int32_t rp = (lane <= p.Lp1) ? __ldg(p.rowptr + r0 + lane) : 0;
        const int32_t e0 = __shfl_sync(0xffffffffu, rp, 0);
        const int32_t total = __shfl_sync(0xffffffffu, rp, p.Lp1) - e0;
        int t = 0;
        int32_t next_end = __shfl_sync(0xffffffffu, rp, 1);
        float acc[VE];
#pragma unroll
        for (int k = 0; k < VE; ++k) acc[k] = 0.0f;

        for (int32_t basee = 0; basee < total; basee += 64) {
            const int n = min(64, total - basee);
            int32_t c0 = 0, c1 = 0;
            float w0 = 0.0f, w1 = 0.0f;
            if (lane < n) {
                c0 = __ldg(p.col + e0 + basee + lane);
                w0 = __ldg(p.w + e0 + basee + lane);
            }
            if (32 + lane < n) {
                c1 = __ldg(p.col + e0 + basee + 32 + lane);
                w1 = __ldg(p.w + e0 + basee + 32 + lane);
            }
            if (!slab_ready) {            // first use: the index loads above overlap the TMA fill
                mbar_wait(&bar, 0);
                slab_ready = true;
            }
#pragma unroll 1
            for (int half = 0; half < 2; ++half) {
                const int32_t cc = half ? c1 : c0;
                const float ww = half ? w1 : w0;
                const int cnt = min(32, n - half * 32);
#pragma unroll 4
                for (int k = 0; k < cnt; ++k) {
                    const int32_t e = e0 + basee + half * 32 + k;
                    while (e >= next_end) {      // warp-uniform: close the typed row(s) that end here
                        LV::store(ybase + (r0 + t) * p.ldy_bytes, acc);
#pragma unroll
                        for (int q = 0; q < VE; ++q) acc[q] = 0.0f;
                        ++t;
                        next_end = __shfl_sync(0xffffffffu, rp, t + 1);
                    }
                    const int32_t src = __shfl_sync(0xffffffffu, cc, k);
                    const float wt = __shfl_sync(0xffffffffu, ww, k);
                    if (wt != 0.0f) {            // dropped edges are never read
                        float v[VE];
                        LV::load(slab + static_cast<size_t>(src) * ROW_BYTES + lane * VB, v);
#pragma unroll
                        for (int q = 0; q < VE; ++q) acc[q] = fmaf(wt, v[q], acc[q]);
                    }
                }
            }
        }
        for (; t < p.Lp1; ++t) {                 // last row with entries + trailing empty rows
            LV::store(ybase + (r0 + t) * p.ldy_bytes, acc);
#pragma unroll
            for (int q = 0; q < VE; ++q) acc[q] = 0.0f;
        }
    }
    if (!slab_ready) mbar_wait(&bar, 0);         // never leave with a TMA write still in flight
}

template <typename T, int VB>
static int launch_tile(const SpmmTileParams& p, const CUtensorMap& tm, int B, size_t smem, int threads, cudaStream_t st) {
    auto kern = spmm_tile_kernel<T, VB>;
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(spmm_tile_kernel)");
        attr_smem = smem;
    }
    kern<<<B * p.n_chunks, threads, smem, st>>>(p, tm);
    GRL_LAUNCH_CHECK("spmm_tile_kernel");
    return 0;
}

// returns 0 = launched, 1 = not applicable (caller falls back to the gather kernel), other = error
int try_launch_spmm_tile(const int32_t* rowptr, const int32_t* col, const float* w, const void* X, int64_t ldx, void* Y,
                         int64_t ldy, int dtype, int B, int N, int L, int F, cudaStream_t st) {
    if (g_spmm_tile_mode < 0) return 1;
    const int es = dtype == GRL_F32 ? 4 : 2;
    if (L + 1 >= 32) return 1;
    if (!aligned_to(X, 16) || (ldx * es) % 16 != 0) return 1;
    const int n_boxes = (N + 255) / 256;
    const int box_rows = (N + n_boxes - 1) / n_boxes;
    const int rows = n_boxes * box_rows;
    constexpr size_t kMaxSmem = 227 * 1024 - 256;
    // bytes per lane: the widest chunk that still lets two CTAs share an SM, else the widest that fits at all
    int vb = 0;
    const int cand[3] = {16, 8, 4};
    auto ok = [&](int v) {
        const int fc = 32 * v / es;
        return F % fc == 0 && (ldy * es) % v == 0 && (reinterpret_cast<uintptr_t>(Y) % v) == 0;
    };
    if (g_spmm_tile_mode > 0) {
        if (ok(g_spmm_tile_mode) && static_cast<size_t>(rows) * 32 * g_spmm_tile_mode + 128 <= kMaxSmem) vb = g_spmm_tile_mode;
    } else {
        for (int v : cand)
            if (!vb && ok(v) && static_cast<size_t>(rows) * 32 * v + 128 <= kMaxSmem / 2) vb = v;
        for (int v : cand)
            if (!vb && ok(v) && static_cast<size_t>(rows) * 32 * v + 128 <= kMaxSmem) vb = v;
    }
    if (!vb) return 1;
    const int fc = 32 * vb / es;
    const size_t smem = static_cast<size_t>(rows) * 32 * vb + 128;
    const int per_sm = static_cast<int>(kMaxSmem / smem) < 1 ? 1 : static_cast<int>(kMaxSmem / smem);
    int threads = per_sm >= 4 ? 256 : (per_sm >= 2 ? 512 : 1024);
    while (threads > 32 && threads / 32 > N) threads >>= 1;     // tiny graphs: no idle warps

    CUtensorMap tm;
    int rc = make_tmap_plain(&tm, X, es, static_cast<int64_t>(B) * N, F, ldx, fc, box_rows);
    if (rc) return rc;
    SpmmTileParams p{rowptr, col, w, static_cast<uint8_t*>(Y), ldy * es, N, L + 1, F / fc, box_rows, n_boxes};
#define GRL_TILE_CASE(TT, V) \
    if (vb == V) return launch_tile<TT, V>(p, tm, B, smem, threads, st)
    if (dtype == GRL_F32) {
        GRL_TILE_CASE(float, 4); GRL_TILE_CASE(float, 8); GRL_TILE_CASE(float, 16);
    } else {
        GRL_TILE_CASE(__nv_bfloat16, 4); GRL_TILE_CASE(__nv_bfloat16, 8); GRL_TILE_CASE(__nv_bfloat16, 16);
    }
#undef GRL_TILE_CASE
    return 1;
}

}  // namespace grl

extern "C" int grl_spmm_set_tile_mode(int mode) {
    if (mode != 0 && mode != -1 && mode != 4 && mode != 8 && mode != 16) {
        grl::set_error("grl_spmm_set_tile_mode: mode %d not in {0 auto, -1 off, 4, 8, 16}", mode);
        return GRL_E_BADARG;
    }
    grl::g_spmm_tile_mode = mode;
    return 0;
}
