// Segmented SpMM over the (L+1) typed adjacencies and its transpose (HBM-bound gather/stream kernels).
//
// Reference semantics: new_V = torch.matmul(A', V) (gnn/models/networks/robust_gcn.py:45-47) with A'
// already multiplied by the DropEdge mask (gnn/models/networks/drop_robust_gcn.py:76,80,85); the
// backward is autograd's A'^T * d(new_V).  Both are the same "row gather" kernel here:
//   forward : rows = CSR rows r,     sources = X[base(r) + col[e]],   weight = w[e]
//   backward: rows = CSC columns c,  sources = dY[rowidx[k]],         weight = w[perm[k]]
#include "common.cuh"

namespace grl {

// One warp per output row.  A row of F elements is CHUNKS * 32 lanes * 16 bytes.
// Entries are consumed four at a time so that 4*CHUNKS independent 128-bit gathers are in flight
// per lane before the first FMA (MLP; rows have only ~8 entries, so the latency has to be covered
// inside the row rather than across loop iterations).
template <typename T, int CHUNKS, bool kTransposed, bool kAccumulate>
__global__ void __launch_bounds__(256)
spmm_rows_kernel(const int32_t* __restrict__ ptr, const int32_t* __restrict__ idx, const int32_t* __restrict__ widx,
                 const float* __restrict__ w, const T* __restrict__ X, int64_t ldx, T* __restrict__ Y, int64_t ldy,
                 int64_t n_rows, int rows_per_graph, int cols_per_graph) {
    constexpr int E = Vec16<T>::kElems;
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;

    for (int64_t r = warp; r < n_rows; r += n_warps) {
        const int32_t beg = __ldg(ptr + r), end = __ldg(ptr + r + 1);
        const int64_t base = kTransposed ? 0 : (r / rows_per_graph) * static_cast<int64_t>(cols_per_graph);
        float acc[CHUNKS][E];
#pragma unroll
        for (int c = 0; c < CHUNKS; ++c)
#pragma unroll
            for (int k = 0; k < E; ++k) acc[c][k] = 0.0f;

        for (int32_t e0 = beg; e0 < end; e0 += 32) {
            // lanes fetch (source row, weight) of up to 32 entries, then broadcast them
            int32_t my_src = 0;
            float my_w = 0.0f;
            if (e0 + lane < end) {
                my_src = __ldg(idx + e0 + lane);
                const int32_t we = kTransposed ? __ldg(widx + e0 + lane) : (e0 + lane);
                my_w = __ldg(w + we);
            }
            const int n_here = min(32, end - e0);
            for (int t0 = 0; t0 < n_here; t0 += 4) {
                float v[4][CHUNKS][E];
                float wt[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int t = t0 + u;
                    const int32_t src = __shfl_sync(0xffffffffu, my_src, t & 31);
                    wt[u] = (t < n_here) ? __shfl_sync(0xffffffffu, my_w, t & 31) : 0.0f;
                    if (wt[u] != 0.0f) {  // dropped edges (weight 0) are never fetched
                        const T* row = X + (base + src) * ldx + lane * E;
#pragma unroll
                        for (int c = 0; c < CHUNKS; ++c) Vec16<T>::load(row + c * 32 * E, v[u][c]);
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (wt[u] != 0.0f) {
#pragma unroll
                        for (int c = 0; c < CHUNKS; ++c)
#pragma unroll
                            for (int k = 0; k < E; ++k) acc[c][k] = fmaf(wt[u], v[u][c][k], acc[c][k]);
                    }
                }
            }
        }
        T* out = Y + r * ldy + lane * E;
#pragma unroll
        for (int c = 0; c < CHUNKS; ++c) {
            if (kAccumulate) {
                float old[E];
                Vec16<T>::load(out + c * 32 * E, old);
#pragma unroll
                for (int k = 0; k < E; ++k) acc[c][k] += old[k];
                Vec16<T>::store(out + c * 32 * E, acc[c]);
            } else if (kTransposed) {
                Vec16<T>::store(out + c * 32 * E, acc[c]);  // dX is re-read by the next GEMM: keep in L2
            } else {
                Vec16<T>::store_streaming(out + c * 32 * E, acc[c]);  // new_V is (L+1)x larger than X: stream it
            }
        }
    }
}

// Forward aggregation, one warp per NODE: all (L+1) typed rows of a node are contiguous in CSR, so the warp fetches
// the node's (L+2) row pointers with one load and up to 64 (column, weight) pairs with two coalesced loads, turns
// each column into a ready-made 16-byte-unit row offset in its own lane, and then walks the entries with two
// shuffles + one address add per entry.  The per-row bookkeeping of spmm_rows_kernel (53 issued instructions per
// entry, 63 % issue-slot utilisation in ncu: the kernel was issue-bound, not HBM-bound) is amortised over ~50 entries.
template <typename T, int CHUNKS>
__global__ void __launch_bounds__(256)
spmm_nodes_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, const float* __restrict__ w,
                  const T* __restrict__ X, uint32_t ldx16, T* __restrict__ Y, int64_t ldy, int64_t n_nodes,
                  int nodes_per_graph, int Lp1) {
    constexpr int E = Vec16<T>::kElems;
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    const uint4* xlane = reinterpret_cast<const uint4*>(X) + lane;      // 16-byte units

    for (int64_t node = warp; node < n_nodes; node += n_warps) {
        const int64_t r0 = node * Lp1;
        const int32_t rp = (lane <= Lp1) ? __ldg(rowptr + r0 + lane) : 0;
        const int32_t e0 = __shfl_sync(0xffffffffu, rp, 0);
        const int32_t total = __shfl_sync(0xffffffffu, rp, Lp1) - e0;
        const uint32_t base = static_cast<uint32_t>((node / nodes_per_graph) * nodes_per_graph);
        // this lane's two entries: (row offset in 16-byte units, weight); weight 0 = dropped or out of range
        uint32_t o0 = 0, o1 = 0;
        float w0 = 0.0f, w1 = 0.0f;
        if (lane < total) {
            o0 = (base + static_cast<uint32_t>(__ldg(col + e0 + lane))) * ldx16;
            w0 = __ldg(w + e0 + lane);
        }
        if (lane + 32 < total) {
            o1 = (base + static_cast<uint32_t>(__ldg(col + e0 + 32 + lane))) * ldx16;
            w1 = __ldg(w + e0 + 32 + lane);
        }
        T* out = Y + r0 * ldy + lane * E;
        for (int t = 0; t < Lp1; ++t, out += ldy) {
            const int beg = __shfl_sync(0xffffffffu, rp, t) - e0;
            const int end = __shfl_sync(0xffffffffu, rp, t + 1) - e0;
            float acc[CHUNKS][E];
#pragma unroll
            for (int c = 0; c < CHUNKS; ++c)
#pragma unroll
                for (int k = 0; k < E; ++k) acc[c][k] = 0.0f;
            for (int k0 = beg; k0 < end; k0 += 4) {
                float v[4][CHUNKS][E];
                float wt[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int k = k0 + u;
                    uint32_t off;
                    if (k < 64) {                                   // warp-uniform
                        off = __shfl_sync(0xffffffffu, k < 32 ? o0 : o1, k & 31);
                        wt[u] = __shfl_sync(0xffffffffu, k < 32 ? w0 : w1, k & 31);
                    } else if (k < end) {                           // rare: > 64 entries on one node
                        off = (base + static_cast<uint32_t>(__ldg(col + e0 + k))) * ldx16;
                        wt[u] = __ldg(w + e0 + k);
                    } else {
                        off = 0;
                        wt[u] = 0.0f;
                    }
                    if (k >= end) wt[u] = 0.0f;
                    if (wt[u] != 0.0f) {                            // dropped edges (weight 0) are never fetched
#pragma unroll
                        for (int c = 0; c < CHUNKS; ++c) {
                            const uint4 q = __ldg(xlane + off + c * 32);
                            if constexpr (E == 4) {
                                v[u][c][0] = __uint_as_float(q.x); v[u][c][1] = __uint_as_float(q.y);
                                v[u][c][2] = __uint_as_float(q.z); v[u][c][3] = __uint_as_float(q.w);
                            } else {
                                const uint32_t ww[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                                for (int i = 0; i < 4; ++i) {
                                    v[u][c][2 * i] = __uint_as_float(ww[i] << 16);
                                    v[u][c][2 * i + 1] = __uint_as_float(ww[i] & 0xFFFF0000u);
                                }
                            }
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (wt[u] != 0.0f) {
#pragma unroll
                        for (int c = 0; c < CHUNKS; ++c)
#pragma unroll
                            for (int k = 0; k < E; ++k) acc[c][k] = fmaf(wt[u], v[u][c][k], acc[c][k]);
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < CHUNKS; ++c) Vec16<T>::store_streaming(out + c * 32 * E, acc[c]);
        }
    }
}

// Generic width fallback (F not a multiple of 32*16 bytes): lanes stride over columns.
template <typename T, bool kTransposed, bool kAccumulate>
__global__ void __launch_bounds__(256)
spmm_rows_generic_kernel(const int32_t* __restrict__ ptr, const int32_t* __restrict__ idx,
                         const int32_t* __restrict__ widx, const float* __restrict__ w, const T* __restrict__ X,
                         int64_t ldx, T* __restrict__ Y, int64_t ldy, int64_t n_rows, int rows_per_graph,
                         int cols_per_graph, int F) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n_rows; r += n_warps) {
        const int32_t beg = ptr[r], end = ptr[r + 1];
        const int64_t base = kTransposed ? 0 : (r / rows_per_graph) * static_cast<int64_t>(cols_per_graph);
        for (int f0 = 0; f0 < F; f0 += 32) {
            const int f = f0 + lane;
            float acc = 0.0f;
            for (int32_t e = beg; e < end; ++e) {
                const float wt = w[kTransposed ? widx[e] : e];
                if (wt != 0.0f && f < F) acc = fmaf(wt, to_float(X[(base + idx[e]) * ldx + f]), acc);
            }
            if (f < F) {
                T* o = Y + r * ldy + f;
                if (kAccumulate) acc += to_float(*o);
                *o = from_float<T>(acc);
            }
        }
    }
}

// grl_spmm_set_tile_mode: 0 auto, -1 node-per-warp gather kernel, -2 legacy row-per-warp kernel (both: grl_spmm_fwd);
// 108/116/132 (208/216) force 8/16/32 lanes per node group (two 16-byte words per lane) in grl_spmm_fwd_stream
int g_spmm_mode = 0;
extern int g_slab2_gl;    // spmm_slab2.cu
extern int g_slab2_vw;    // spmm_slab2.cu

template <typename T, bool kTransposed>
static int launch_spmm(const int32_t* ptr, const int32_t* idx, const int32_t* widx, const float* w, const void* Xv,
                       int64_t ldx, void* Yv, int64_t ldy, int64_t n_rows, int rows_per_graph, int cols_per_graph, int F,
                       bool accumulate, cudaStream_t st) {
    const T* X = static_cast<const T*>(Xv);
    T* Y = static_cast<T*>(Yv);
    constexpr int E = Vec16<T>::kElems;
    const int chunk = 32 * E;
    const bool vec_ok = (F % chunk == 0) && (F / chunk <= 4) && (F / chunk != 3) && (ldx % E == 0) && (ldy % E == 0) &&
                        aligned_to(X, 16) && aligned_to(Y, 16);
    // 8 warps per block; grid capped at a multiple of the SM count, warps loop over rows.
    int64_t blocks = (n_rows + 7) / 8;
    const int64_t cap = static_cast<int64_t>(kNumSMs) * 32;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    const int grid = static_cast<int>(blocks);
#define GRL_SPMM_CASE(C, ACC)                                                                             \
    spmm_rows_kernel<T, C, kTransposed, ACC><<<grid, 256, 0, st>>>(ptr, idx, widx, w, X, ldx, Y, ldy, n_rows, \
                                                                  rows_per_graph, cols_per_graph)
    if (vec_ok && !kTransposed && !accumulate && g_spmm_mode != -2 && rows_per_graph / cols_per_graph < 32 &&
        static_cast<uint64_t>(n_rows / (rows_per_graph / cols_per_graph)) * static_cast<uint64_t>(ldx) * sizeof(T) / 16 <
            (1ull << 32)) {
        const int Lp1 = rows_per_graph / cols_per_graph;
        const int64_t n_nodes = n_rows / Lp1;
        int64_t nb = (n_nodes + 7) / 8;
        if (nb > cap) nb = cap;
        if (nb < 1) nb = 1;
        const uint32_t ldx16 = static_cast<uint32_t>(ldx * sizeof(T) / 16);
        const int chunks = F / chunk;
#define GRL_NODE_CASE(C) \
    spmm_nodes_kernel<T, C><<<static_cast<int>(nb), 256, 0, st>>>(ptr, idx, w, X, ldx16, Y, ldy, n_nodes, cols_per_graph, Lp1)
        if (chunks == 1) GRL_NODE_CASE(1);
        else if (chunks == 2) GRL_NODE_CASE(2);
        else GRL_NODE_CASE(4);
#undef GRL_NODE_CASE
        GRL_LAUNCH_CHECK("spmm_nodes_kernel");
        return 0;
    }
    if (vec_ok) {
        const int chunks = F / chunk;
        if (accumulate) {
            if (chunks == 1) GRL_SPMM_CASE(1, true);
            else if (chunks == 2) GRL_SPMM_CASE(2, true);
            else GRL_SPMM_CASE(4, true);
        } else {
            if (chunks == 1) GRL_SPMM_CASE(1, false);
            else if (chunks == 2) GRL_SPMM_CASE(2, false);
            else GRL_SPMM_CASE(4, false);
        }
    } else {
        if (accumulate)
            spmm_rows_generic_kernel<T, kTransposed, true><<<grid, 256, 0, st>>>(ptr, idx, widx, w, X, ldx, Y, ldy, n_rows,
                                                                                rows_per_graph, cols_per_graph, F);
        else
            spmm_rows_generic_kernel<T, kTransposed, false><<<grid, 256, 0, st>>>(ptr, idx, widx, w, X, ldx, Y, ldy, n_rows,
                                                                                 rows_per_graph, cols_per_graph, F);
    }
#undef GRL_SPMM_CASE
    GRL_LAUNCH_CHECK("spmm_rows_kernel");
    return 0;
}

__global__ void dropedge_weights_kernel(const float* __restrict__ val, const uint8_t* __restrict__ mask, int64_t nnz,
                                        float scale, uint32_t thresh, uint64_t seed, uint64_t stream_id,
                                        float* __restrict__ w, const uint64_t* __restrict__ rng_off) {
    stream_id = rng_stream(stream_id, rng_off);
    const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    // each thread owns 4 consecutive entries = one Philox block
    for (int64_t q = tid; q * 4 < nnz; q += stride) {
        uint32_t bits[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
        if (!mask && thresh) {
            const uint4 d = Philox::draw(seed, stream_id, static_cast<uint64_t>(q));
            bits[0] = d.x; bits[1] = d.y; bits[2] = d.z; bits[3] = d.w;
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int64_t e = q * 4 + k;
            if (e < nnz) {
                const bool keep = mask ? (mask[e] != 0) : (bits[k] >= thresh);
                w[e] = keep ? val[e] * scale : 0.0f;
            }
        }
    }
}

}  // namespace grl

using namespace grl;

extern "C" int grl_dropedge_weights(const float* val, const uint8_t* mask, int64_t nnz, float p, uint64_t seed,
                                    uint64_t stream_id, float* w, void* stream) {
    GRL_CHECK_ARG(val && w, GRL_E_BADARG, "grl_dropedge_weights: null pointer");
    GRL_CHECK_ARG(p >= 0.0f && p < 1.0f, GRL_E_BADARG, "grl_dropedge_weights: p=%f outside [0,1)", p);
    if (nnz == 0) return 0;
    const float scale = 1.0f / (1.0f - p);
    int64_t blocks = (nnz / 4 + 255) / 256 + 1;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    dropedge_weights_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        val, mask, nnz, scale, p > 0.0f ? keep_threshold(p) : 0u, seed, stream_id, w, rng_offset_ptr());
    GRL_LAUNCH_CHECK("dropedge_weights_kernel");
    return 0;
}

extern "C" int grl_spmm_set_tile_mode(int mode) {
    if (mode != 0 && mode != -1 && mode != -2 && mode != 108 && mode != 116 && mode != 132 && mode != 208 && mode != 216) {
        set_error("grl_spmm_set_tile_mode: mode %d not in {0 auto, -1 gather, -2 legacy rows, 108/116/132/208/216 slab "
                  "group width}", mode);
        return GRL_E_BADARG;
    }
    g_spmm_mode = mode > 100 ? 0 : mode;             // 1xx / 2xx only steer grl_spmm_fwd_stream
    g_slab2_gl = mode > 200 ? mode - 200 : (mode > 100 ? mode - 100 : 0);
    g_slab2_vw = mode > 200 ? 2 : (mode > 100 ? 1 : 0);    // 208 / 216: two 16-byte words per lane
    return 0;
}

extern "C" int grl_spmm_fwd(const int32_t* rowptr, const int32_t* col, const float* w, const void* X, int64_t ldx,
                            void* Y, int64_t ldy, int dtype, int B, int N, int L, int F, void* stream) {
    GRL_CHECK_ARG(rowptr && col && w && X && Y, GRL_E_BADARG, "grl_spmm_fwd: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0 && L > 0 && F > 0, GRL_E_BADARG, "grl_spmm_fwd: bad shape");
    GRL_CHECK_ARG(ldx >= F && ldy >= F, GRL_E_BADARG, "grl_spmm_fwd: leading dimension < F");
    const int64_t R = static_cast<int64_t>(B) * N * (L + 1);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == GRL_F32)
        return launch_spmm<float, false>(rowptr, col, nullptr, w, X, ldx, Y, ldy, R, N * (L + 1), N, F, false, st);
    if (dtype == GRL_BF16)
        return launch_spmm<__nv_bfloat16, false>(rowptr, col, nullptr, w, X, ldx, Y, ldy, R, N * (L + 1), N, F, false, st);
    set_error("grl_spmm_fwd: dtype %d unsupported", dtype);
    return GRL_E_UNSUP;
}

extern "C" int grl_spmm_bwd(const int32_t* colptr, const int32_t* rowidx, const int32_t* perm, const float* w,
                            const void* dY, int64_t lddy, void* dX, int64_t lddx, int dtype, int B, int N, int L, int F,
                            int accumulate, void* stream) {
    GRL_CHECK_ARG(colptr && rowidx && perm && w && dY && dX, GRL_E_BADARG, "grl_spmm_bwd: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0 && L > 0 && F > 0, GRL_E_BADARG, "grl_spmm_bwd: bad shape");
    GRL_CHECK_ARG(lddy >= F && lddx >= F, GRL_E_BADARG, "grl_spmm_bwd: leading dimension < F");
    const int64_t n_nodes = static_cast<int64_t>(B) * N;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == GRL_F32)
        return launch_spmm<float, true>(colptr, rowidx, perm, w, dY, lddy, dX, lddx, n_nodes, 1, 0, F, accumulate != 0, st);
    if (dtype == GRL_BF16)
        return launch_spmm<__nv_bfloat16, true>(colptr, rowidx, perm, w, dY, lddy, dX, lddx, n_nodes, 1, 0, F,
                                                accumulate != 0, st);
    set_error("grl_spmm_bwd: dtype %d unsupported", dtype);
    return GRL_E_UNSUP;
}
