// CSR / CSC packing of the typed document-graph adjacency (integer, bit-exact work; HBM-bound).
//
// Reference semantics: GraphConv.preprocess_adj, gnn/models/networks/robust_gcn.py:53-72 --
// A'[b, i*(L+1)+t, j] with t=0 the identity block.  Here A' is never materialised: the dense
// [B,N,L,N] batch tensor is read twice (count, fill) and only rowptr/col/val are written.
#include "common.cuh"
#include <cub/device/device_scan.cuh>

namespace grl {

// One warp per (b, i, l) adjacency row of N contiguous elements.  VEC elements per lane per step.
template <typename T, int VEC>
__device__ __forceinline__ unsigned load_nz_bits(const T* row, int j0, int N) {
    // returns a VEC-bit mask of non-zero elements row[j0 .. j0+VEC)
    unsigned bits = 0;
    if (VEC > 1 && j0 + VEC <= N) {
        if constexpr (sizeof(T) * VEC == 16) {
            uint4 t = __ldcs(reinterpret_cast<const uint4*>(row + j0));
            const uint32_t w[4] = {t.x, t.y, t.z, t.w};
            if constexpr (sizeof(T) == 4) {
#pragma unroll
                for (int k = 0; k < 4; ++k) bits |= ((w[k] << 1) != 0u) << k;  // ignore sign bit: -0.0 is zero
            } else {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    bits |= ((w[k] & 0x7FFFu) != 0u) << (2 * k);
                    bits |= ((w[k] & 0x7FFF0000u) != 0u) << (2 * k + 1);
                }
            }
        } else if constexpr (sizeof(T) * VEC == 8) {
            uint2 t = __ldcs(reinterpret_cast<const uint2*>(row + j0));
            const uint32_t w[2] = {t.x, t.y};
            if constexpr (sizeof(T) == 4) {
#pragma unroll
                for (int k = 0; k < 2; ++k) bits |= ((w[k] << 1) != 0u) << k;
            } else {
#pragma unroll
                for (int k = 0; k < 2; ++k) {
                    bits |= ((w[k] & 0x7FFFu) != 0u) << (2 * k);
                    bits |= ((w[k] & 0x7FFF0000u) != 0u) << (2 * k + 1);
                }
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            if (j0 + k < N) bits |= (to_float(row[j0 + k]) != 0.0f) << k;
    }
    return bits;
}

template <typename T, int VEC>
__global__ void __launch_bounds__(256) csr_count_kernel(const T* __restrict__ adj, int64_t n_adj_rows, int N, int L,
                                                        int self_loop, int32_t* __restrict__ rowptr) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    if (warp == 0 && lane == 0) rowptr[0] = 0;
    for (int64_t ar = warp; ar < n_adj_rows; ar += n_warps) {  // ar = (b*N+i)*L + l
        const T* row = adj + ar * N;
        int cnt = 0;
        for (int j0 = lane * VEC; j0 < N; j0 += 32 * VEC) cnt += __popc(load_nz_bits<T, VEC>(row, j0, N));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if (lane == 0) {
            const int64_t node = ar / L;
            const int l = static_cast<int>(ar - node * L);
            const int64_t r = node * (L + self_loop) + l + self_loop;
            rowptr[r + 1] = cnt;
            if (self_loop && l == 0) rowptr[r] = 1;  // self-loop row (t = 0) holds exactly the diagonal entry
        }
    }
}

template <typename T, int VEC>
__global__ void __launch_bounds__(256) csr_fill_kernel(const T* __restrict__ adj, int64_t n_adj_rows, int N, int L,
                                                       int self_loop, const int32_t* __restrict__ rowptr,
                                                       int32_t* __restrict__ col, float* __restrict__ val) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t ar = warp; ar < n_adj_rows; ar += n_warps) {
        const T* row = adj + ar * N;
        const int64_t node = ar / L;
        const int l = static_cast<int>(ar - node * L);
        const int64_t r = node * (L + self_loop) + l + self_loop;
        if (self_loop && l == 0 && lane == 0) {
            const int32_t e = rowptr[r - 1];
            col[e] = static_cast<int32_t>(node % N);
            val[e] = 1.0f;
        }
        int32_t out = rowptr[r];
        const int32_t end = rowptr[r + 1];
        if (out == end) continue;  // empty row: nothing to re-read
        for (int jb = 0; jb < N && out < end; jb += 32 * VEC) {
            const int j0 = jb + lane * VEC;
            unsigned bits = (j0 < N) ? load_nz_bits<T, VEC>(row, j0, N) : 0u;
            const int mine = __popc(bits);
            int incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            int pos = out + incl - mine;
            while (bits) {
                const int k = __ffs(bits) - 1;
                bits &= bits - 1;
                col[pos] = j0 + k;
                val[pos] = to_float(row[j0 + k]);
                ++pos;
            }
            out += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
}

// ---------------------------------------------------------------- dense -> CSR through a bitmap (one read of A)
// The count pass already sees every element: it also leaves one bit per element (row-major, ceil(N/32) words per
// adjacency row, 1/16 of an fp16 A), so the fill pass never re-reads the dense tensor -- it walks the bitmap and fetches
// only the stored values.  Loads of a row are issued four deep (the plain kernels keep one 8/16-byte load per lane in
// flight: 0.24 + 0.36 ms for the 768 MB fp16 adjacency of cfg2, against 0.12 ms per pass at HBM speed).
// The same test on a register copy: the bitmap pass issues every load of its rows first (plain predicated loads, no
// tail path: its vector widths divide N) and only then looks at the bits -- with the test inlined next to each load the
// compiler serialised them (one DRAM latency per 8-byte load: 2.0 TB/s on cfg2's adjacency).
template <typename T, int VEC>
__device__ __forceinline__ uint4 load_raw_vec(const T* p, bool ok) {
    static_assert(sizeof(T) * VEC == 16 || sizeof(T) * VEC == 8, "vector loads of 8 or 16 bytes");
    uint4 r = make_uint4(0u, 0u, 0u, 0u);
    if constexpr (sizeof(T) * VEC == 16) {
        if (ok) r = __ldcs(reinterpret_cast<const uint4*>(p));
    } else {
        if (ok) {
            const uint2 t = __ldcs(reinterpret_cast<const uint2*>(p));
            r.x = t.x;
            r.y = t.y;
        }
    }
    return r;
}

template <typename T, int VEC>
__device__ __forceinline__ unsigned nz_bits_of(const uint4& r) {
    constexpr int WORDS = static_cast<int>(sizeof(T)) * VEC / 4;
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
    unsigned bits = 0;
#pragma unroll
    for (int k = 0; k < WORDS; ++k) {
        if constexpr (sizeof(T) == 4) {
            bits |= ((w[k] << 1) != 0u) << k;                     // ignore the sign bit: -0.0 is zero
        } else {
            bits |= ((w[k] & 0x7FFFu) != 0u) << (2 * k);
            bits |= ((w[k] & 0x7FFF0000u) != 0u) << (2 * k + 1);
        }
    }
    return bits;
}

// Non-zero when some element of the vector is neither +0 nor exactly +1 (1.0 = 0x3C00 / 0x3F800000: a run of exponent
// bits; anything else has a bit outside the run or two unequal neighbours inside it).  Three instructions per word.
// -0.0 counts as "other": the batch then only takes the slower, still exact, value-fetching fill.
template <typename T, int VEC>
__device__ __forceinline__ uint32_t not_zero_or_one(const uint4& r) {
    constexpr int WORDS = static_cast<int>(sizeof(T)) * VEC / 4;
    constexpr uint32_t ONE = sizeof(T) == 4 ? 0x3F800000u : 0x3C003C00u;
    constexpr uint32_t INNER = sizeof(T) == 4 ? 0x1F800000u : 0x1C001C00u;    // the run without its top bit
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
    uint32_t bad = 0;
#pragma unroll
    for (int k = 0; k < WORDS; ++k) bad |= (w[k] & ~ONE) | ((w[k] ^ (w[k] >> 1)) & INNER);
    return bad;
}

template <typename T, int VEC>
__global__ void __launch_bounds__(256) csr_count_bitmap_kernel(const T* __restrict__ adj, int64_t n_adj_rows, int N, int L,
                                                               int self_loop, int32_t* __restrict__ rowptr,
                                                               uint32_t* __restrict__ bitmap, int words,
                                                               uint32_t* __restrict__ weighted) {
    constexpr int LPW = 32 / VEC;                        // lanes that share one 32-bit bitmap word
    uint32_t other = 0;                                  // non-zero: this lane saw an element that is neither 0 nor 1
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    if (warp == 0 && lane == 0) rowptr[0] = 0;
    // two adjacency rows per warp iteration, all loads of both issued before the first use: the pass is bound by bytes in
    // flight per SM (one 1000-byte row per warp gave 2.9 TB/s)
    constexpr int ROWS = 2;
    // (node, l) of row ar0 is stepped along with it: a 64-bit division per row was a third of this kernel's instructions
    const uint32_t uL = static_cast<uint32_t>(L);
    const uint32_t step_node = static_cast<uint32_t>((n_warps * ROWS) / uL), step_l = static_cast<uint32_t>((n_warps * ROWS) % uL);
    uint32_t node0 = static_cast<uint32_t>((warp * ROWS) / uL), l0 = static_cast<uint32_t>((warp * ROWS) % uL);
    for (int64_t ar0 = warp * ROWS; ar0 < n_adj_rows; ar0 += n_warps * ROWS) {
        int cnt[ROWS];
#pragma unroll
        for (int q = 0; q < ROWS; ++q) cnt[q] = 0;
        for (int jb = 0; jb < N; jb += 4 * 32 * VEC) {
            unsigned b[ROWS][4];
            if constexpr (VEC > 1) {
                uint4 raw[ROWS][4];
#pragma unroll
                for (int q = 0; q < ROWS; ++q) {
                    const T* row = adj + (ar0 + q) * N;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int j0 = jb + u * 32 * VEC + lane * VEC;
                        raw[q][u] = load_raw_vec<T, VEC>(row + j0, ar0 + q < n_adj_rows && j0 < N);
                    }
                }
#pragma unroll
                for (int q = 0; q < ROWS; ++q)
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        b[q][u] = nz_bits_of<T, VEC>(raw[q][u]);
                        other |= not_zero_or_one<T, VEC>(raw[q][u]);
                    }
            } else {
#pragma unroll
                for (int q = 0; q < ROWS; ++q) {
                    const T* row = adj + (ar0 + q) * N;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int j0 = jb + u * 32 * VEC + lane * VEC;
                        b[q][u] = (ar0 + q < n_adj_rows && j0 < N) ? load_nz_bits<T, VEC>(row, j0, N) : 0u;
                        if (b[q][u]) other |= to_float(row[j0]) != 1.0f;      // VEC == 1: the one element of this lane
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < ROWS; ++q) {
                uint32_t* brow = bitmap + (ar0 + q) * words;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    cnt[q] += __popc(b[q][u]);
                    unsigned v = b[q][u] << (VEC * (lane % LPW));
#pragma unroll
                    for (int o = 1; o < LPW; o <<= 1) v |= __shfl_xor_sync(0xffffffffu, v, o);
                    const int w = (jb + u * 32 * VEC + lane * VEC) >> 5;
                    if (lane % LPW == 0 && w < words && ar0 + q < n_adj_rows) brow[w] = v;
                }
            }
        }
#pragma unroll
        for (int q = 0; q < ROWS; ++q) {
            int c = cnt[q];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
            const int64_t ar = ar0 + q;
            if (lane == 0 && ar < n_adj_rows) {
                uint32_t node = node0, l = l0 + q;
                if (l >= uL) {
                    l -= uL;
                    ++node;
                }
                const int64_t r = static_cast<int64_t>(node) * (L + self_loop) + l + self_loop;
                rowptr[r + 1] = c;
                if (self_loop && l == 0) rowptr[r] = 1;
            }
        }
        l0 += step_l;
        node0 += step_node;
        if (l0 >= uL) {
            l0 -= uL;
            ++node0;
        }
    }
    if (other) *weighted = 1u;                           // cleared by the launcher; every writer stores the same word
}

// One THREAD per (b, i, l) adjacency row: a row of cfg2 holds 8 entries in 16 bitmap words, so a warp per row spent ~100
// warp instructions (scan shuffles, a divergent bit loop) on 8 stores; a lane walking its own row's words needs no
// shuffle at all, the 32 rows of a warp are contiguous in the bitmap (L1 serves the re-touched sectors) and their entries
// are contiguous in col / val.  The count pass left a word behind the bitmap: 0 = every stored value is 1 (a 0/1
// adjacency) and nothing of A is read here -- a 2-byte value per entry costs a 32-byte DRAM sector at a random address.
// Measured at cfg2 (6.3 M entries): 0.197 ms (warp per row) -> 0.147 ms with values fetched, 0.060 ms without.
template <typename T>
__global__ void __launch_bounds__(256) csr_fill_bitmap_kernel(const T* __restrict__ adj, const uint32_t* __restrict__ bitmap,
                                                              int words, int64_t n_adj_rows, int N, int L, int self_loop,
                                                              const int32_t* __restrict__ rowptr, int32_t* __restrict__ col,
                                                              float* __restrict__ val) {
    const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    const bool weighted = __ldg(bitmap + n_adj_rows * words) != 0u;
    const uint32_t uL = static_cast<uint32_t>(L);
    const bool vec4 = (words & 3) == 0 && (reinterpret_cast<uintptr_t>(bitmap) & 15) == 0;   // 16-byte aligned bitmap rows
    for (int64_t ar = tid; ar < n_adj_rows; ar += stride) {
        const uint32_t node = static_cast<uint32_t>(ar) / uL;          // n_adj_rows < 2^31 (checked by the launcher)
        const uint32_t l = static_cast<uint32_t>(ar) - node * uL;
        const int64_t r = static_cast<int64_t>(node) * (L + self_loop) + l + self_loop;
        int32_t pos = __ldg(rowptr + r);
        const int32_t end = __ldg(rowptr + r + 1);                     // may be the clamped capacity (grl_csr_clamp_rowptr)
        if (self_loop && l == 0) {
            // the self-loop row (t = 0) holds exactly the diagonal entry; empty when the capacity clamp cut it off
            const int32_t e = __ldg(rowptr + r - 1);
            if (e < pos) {
                col[e] = static_cast<int32_t>(node % static_cast<uint32_t>(N));
                val[e] = 1.0f;
            }
        }
        const T* row = adj + ar * N;
        const uint32_t* brow = bitmap + ar * words;
        auto emit = [&](uint32_t bits, int j_base) {
            while (bits && pos < end) {
                const int k = __ffs(bits) - 1;
                bits &= bits - 1;
                col[pos] = j_base + k;
                val[pos] = weighted ? to_float(row[j_base + k]) : 1.0f;
                ++pos;
            }
        };
        if (vec4) {
            for (int w = 0; w < words && pos < end; w += 4) {
                const uint4 q = __ldg(reinterpret_cast<const uint4*>(brow + w));
                emit(q.x, w << 5);
                emit(q.y, (w + 1) << 5);
                emit(q.z, (w + 2) << 5);
                emit(q.w, (w + 3) << 5);
            }
        } else {
            for (int w = 0; w < words && pos < end; ++w) emit(__ldg(brow + w), w << 5);
        }
    }
}

// ---------------------------------------------------------------- COO (edge list) -> CSR
// An edge whose (b, i, l, j) lies outside the batch must not become an out-of-bounds atomicAdd: it is re-pointed at entry
// (0, 0, 0, 0) with value 0 (no contribution; the entry count E + B*N the caller sized its buffers for still holds).
// collate.EdgeListCollate range-checks on the host and raises, so this only guards raw device edge lists.
__device__ __forceinline__ bool coo_edge_ok(const int4& q, int B, int N, int L) {
    return q.x >= 0 && q.x < B && q.y >= 0 && q.y < N && q.z >= 0 && q.z < L && q.w >= 0 && q.w < N;
}

__global__ void coo_count_kernel(const int32_t* __restrict__ edges, int64_t E, int N, int L,
                                 int32_t* __restrict__ rowptr, int64_t n_nodes) {
    const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t e = tid; e < E; e += stride) {
        int4 q = __ldg(reinterpret_cast<const int4*>(edges) + e);  // (b, i, l, j)
        if (!coo_edge_ok(q, static_cast<int>(n_nodes / N), N, L)) q = make_int4(0, 0, 0, 0);
        const int64_t r = (static_cast<int64_t>(q.x) * N + q.y) * (L + 1) + q.z + 1;
        atomicAdd(&rowptr[r + 1], 1);
    }
    for (int64_t v = tid; v < n_nodes; v += stride) atomicAdd(&rowptr[v * (L + 1) + 1], 1);  // self loop
}

__global__ void coo_scatter_kernel(const int32_t* __restrict__ edges, const float* __restrict__ weights, int64_t E,
                                   int N, int L, const int32_t* __restrict__ rowptr, int32_t* __restrict__ cursor,
                                   int32_t* __restrict__ col, float* __restrict__ val, int64_t n_nodes) {
    const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t e = tid; e < E; e += stride) {
        int4 q = __ldg(reinterpret_cast<const int4*>(edges) + e);
        const bool ok = coo_edge_ok(q, static_cast<int>(n_nodes / N), N, L);
        if (!ok) q = make_int4(0, 0, 0, 0);
        const int64_t r = (static_cast<int64_t>(q.x) * N + q.y) * (L + 1) + q.z + 1;
        const int32_t pos = rowptr[r] + atomicAdd(&cursor[r], 1);
        col[pos] = q.w;
        val[pos] = ok ? (weights ? weights[e] : 1.0f) : 0.0f;
    }
    for (int64_t v = tid; v < n_nodes; v += stride) {
        const int32_t pos = rowptr[v * (L + 1)];
        col[pos] = static_cast<int32_t>(v % N);
        val[pos] = 1.0f;
    }
}

// Rows are tiny (a handful of entries): one thread insertion-sorts its row by (col ascending, val descending) so the
// result is independent of the atomic arrival order above.  Repeated (b, i, l, j) triples then sit next to each other: the
// reference's dense build `adj[start, label, end] = w` (graph_utils.py:821-829) is idempotent, so all but the first
// (largest) keep their slot with value 0 -- same aggregation as the dense collate, no second compaction pass.
__global__ void csr_sort_rows_kernel(const int32_t* __restrict__ rowptr, int64_t R, int32_t* __restrict__ col,
                                     float* __restrict__ val) {
    const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t r = tid; r < R; r += stride) {
        const int32_t s = rowptr[r], e = rowptr[r + 1];
        for (int32_t a = s + 1; a < e; ++a) {
            const int32_t c = col[a];
            const float v = val[a];
            int32_t b = a - 1;
            while (b >= s && (col[b] > c || (col[b] == c && val[b] < v))) {
                col[b + 1] = col[b];
                val[b + 1] = val[b];
                --b;
            }
            col[b + 1] = c;
            val[b + 1] = v;
        }
        for (int32_t a = s + 1; a < e; ++a)
            if (col[a] == col[a - 1]) val[a] = 0.0f;
    }
}

// ---------------------------------------------------------------- CSR -> CSC (+perm)
__global__ void csc_count_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, int64_t R,
                                 int rows_per_graph, int N, int32_t* __restrict__ colptr) {
    const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    if (tid == 0) colptr[0] = 0;
    for (int64_t r = tid; r < R; r += stride) {
        const int64_t base = (r / rows_per_graph) * N;
        for (int32_t e = rowptr[r]; e < rowptr[r + 1]; ++e) atomicAdd(&colptr[base + col[e] + 1], 1);
    }
}

__global__ void csc_scatter_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, int64_t R,
                                   int rows_per_graph, int N, const int32_t* __restrict__ colptr,
                                   int32_t* __restrict__ cursor, int32_t* __restrict__ tmp_e,
                                   int32_t* __restrict__ tmp_r) {
    const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t r = tid; r < R; r += stride) {
        const int64_t base = (r / rows_per_graph) * N;
        for (int32_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
            const int64_t c = base + col[e];
            const int32_t pos = colptr[c] + atomicAdd(&cursor[c], 1);
            tmp_e[pos] = e;
            tmp_r[pos] = static_cast<int32_t>(r);
        }
    }
}

// One warp per column: rank every entry by its CSR position (unique) -> deterministic order.
__global__ void __launch_bounds__(256) csc_rank_kernel(const int32_t* __restrict__ colptr, int64_t n_cols,
                                                       const int32_t* __restrict__ tmp_e,
                                                       const int32_t* __restrict__ tmp_r,
                                                       int32_t* __restrict__ rowidx, int32_t* __restrict__ perm) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t c = warp; c < n_cols; c += n_warps) {
        const int32_t s = colptr[c], d = colptr[c + 1] - s;
        for (int32_t a = lane; a < d; a += 32) {
            const int32_t mine = tmp_e[s + a];
            int32_t rank = 0;
            for (int32_t t = 0; t < d; ++t) rank += (tmp_e[s + t] < mine);
            perm[s + rank] = mine;
            rowidx[s + rank] = tmp_r[s + a];
        }
    }
}

// Typed transpose: regroup every CSC column (entries sorted by CSR row r = (i, t)) by edge type, keeping the i order
// inside a type -- the CSR of the transposed typed adjacency A_t^T, rows (j, t), over the same node numbering.  One
// warp per column, two passes over its entries in chunks of 32: per-type totals (lane t keeps the count of type t; an
// exclusive warp scan gives the typed row pointers), then a stable placement by ballot ranks.  (A first version ranked
// every entry against all others of the column, O(d^2) with a modulo each: 0.45 ms at cfg2 against 0.03 ms.)
__global__ void __launch_bounds__(256) csc_typed_kernel(const int32_t* __restrict__ colptr, int64_t n_cols,
                                                        const int32_t* __restrict__ rowidx, const int32_t* __restrict__ perm,
                                                        int Lp1, int N, int32_t* __restrict__ t_rowptr,
                                                        int32_t* __restrict__ t_col, int32_t* __restrict__ t_perm) {
    const int lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t c = warp; c < n_cols; c += n_warps) {
        const int32_t s = colptr[c], d = colptr[c + 1] - s;
        int32_t total = 0;                                  // lane t: entries of type t in this column
        for (int32_t a0 = 0; a0 < d; a0 += 32) {
            const int32_t a = a0 + lane;
            const int ta = a < d ? rowidx[s + a] % Lp1 : -1;
            for (int t = 0; t < Lp1; ++t) {
                const uint32_t m = __ballot_sync(0xffffffffu, ta == t);
                if (lane == t) total += __popc(m);
            }
        }
        int32_t off = total;                                // inclusive scan over the types, then made exclusive
#pragma unroll
        for (int k = 1; k < 32; k <<= 1) {
            const int32_t v = __shfl_up_sync(0xffffffffu, off, k);
            if (lane >= k) off += v;
        }
        off -= total;
        if (lane < Lp1) t_rowptr[c * Lp1 + lane] = s + off;
        if (c == n_cols - 1 && lane == 0) t_rowptr[n_cols * Lp1] = s + d;
        const int32_t node_base = static_cast<int32_t>((c / N) * N);
        int32_t placed = 0;                                 // lane t: entries of type t already placed
        for (int32_t a0 = 0; a0 < d; a0 += 32) {
            const int32_t a = a0 + lane;
            const int32_t r = a < d ? rowidx[s + a] : 0;
            const int ta = a < d ? r % Lp1 : -1;
            int32_t pos = 0;
            for (int t = 0; t < Lp1; ++t) {
                const uint32_t m = __ballot_sync(0xffffffffu, ta == t);
                const int32_t base = __shfl_sync(0xffffffffu, off + placed, t);
                if (ta == t) pos = s + base + __popc(m & lt);
                if (lane == t) placed += __popc(m);
            }
            if (a < d) {
                t_col[pos] = r / Lp1 - node_base;
                t_perm[pos] = perm[s + a];
            }
        }
    }
}

static inline int grid_for(int64_t work_items, int threads, int items_per_thread = 1) {
    int64_t blocks = (work_items + static_cast<int64_t>(threads) * items_per_thread - 1) /
                     (static_cast<int64_t>(threads) * items_per_thread);
    const int64_t cap = static_cast<int64_t>(kNumSMs) * 64;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return static_cast<int>(blocks);
}

template <typename T>
static int launch_count(const T* adj, int B, int N, int L, int self_loop, int32_t* rowptr, cudaStream_t st) {
    const int64_t rows = static_cast<int64_t>(B) * N * L;
    const int grid = grid_for(rows * 32, 256);
    constexpr int V16 = 16 / sizeof(T), V8 = 8 / sizeof(T);
    if (N % V16 == 0 && aligned_to(adj, 16)) csr_count_kernel<T, V16><<<grid, 256, 0, st>>>(adj, rows, N, L, self_loop, rowptr);
    else if (N % V8 == 0 && aligned_to(adj, 8)) csr_count_kernel<T, V8><<<grid, 256, 0, st>>>(adj, rows, N, L, self_loop, rowptr);
    else csr_count_kernel<T, 1><<<grid, 256, 0, st>>>(adj, rows, N, L, self_loop, rowptr);
    GRL_LAUNCH_CHECK("csr_count_kernel");
    return 0;
}

template <typename T>
static int launch_fill(const T* adj, int B, int N, int L, int self_loop, const int32_t* rowptr, int32_t* col, float* val,
                       cudaStream_t st) {
    const int64_t rows = static_cast<int64_t>(B) * N * L;
    const int grid = grid_for(rows * 32, 256);
    constexpr int V16 = 16 / sizeof(T), V8 = 8 / sizeof(T);
    if (N % V16 == 0 && aligned_to(adj, 16))
        csr_fill_kernel<T, V16><<<grid, 256, 0, st>>>(adj, rows, N, L, self_loop, rowptr, col, val);
    else if (N % V8 == 0 && aligned_to(adj, 8))
        csr_fill_kernel<T, V8><<<grid, 256, 0, st>>>(adj, rows, N, L, self_loop, rowptr, col, val);
    else
        csr_fill_kernel<T, 1><<<grid, 256, 0, st>>>(adj, rows, N, L, self_loop, rowptr, col, val);
    GRL_LAUNCH_CHECK("csr_fill_kernel");
    return 0;
}

template <typename T>
static int launch_count_bitmap(const T* adj, int B, int N, int L, int self_loop, int32_t* rowptr, uint32_t* bitmap,
                               cudaStream_t st) {
    const int64_t rows = static_cast<int64_t>(B) * N * L;
    const int grid = grid_for(rows * 32, 256);
    const int words = (N + 31) / 32;
    constexpr int V16 = 16 / sizeof(T), V8 = 8 / sizeof(T);
    uint32_t* weighted = bitmap + rows * words;          // trailer word of the bitmap buffer (grl_csr_bitmap_bytes)
    cudaError_t e = cudaMemsetAsync(weighted, 0, sizeof(uint32_t), st);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(bitmap trailer)");
    if (N % V16 == 0 && aligned_to(adj, 16))
        csr_count_bitmap_kernel<T, V16><<<grid, 256, 0, st>>>(adj, rows, N, L, self_loop, rowptr, bitmap, words, weighted);
    else if (N % V8 == 0 && aligned_to(adj, 8))
        csr_count_bitmap_kernel<T, V8><<<grid, 256, 0, st>>>(adj, rows, N, L, self_loop, rowptr, bitmap, words, weighted);
    else
        csr_count_bitmap_kernel<T, 1><<<grid, 256, 0, st>>>(adj, rows, N, L, self_loop, rowptr, bitmap, words, weighted);
    GRL_LAUNCH_CHECK("csr_count_bitmap_kernel");
    return 0;
}

}  // namespace grl

using namespace grl;

extern "C" size_t grl_csr_bitmap_bytes(int B, int N, int L) {
    // one bit per element, row-major, plus a trailer word: 0 = every stored value is exactly 1 (the fill pass then skips A)
    return sizeof(uint32_t) * (static_cast<size_t>(B) * N * L * ((N + 31) / 32) + 4);
}

extern "C" int grl_csr_count_bitmap(const void* adj, int adj_dtype, int B, int N, int L, int self_loop, int32_t* rowptr,
                                    void* bitmap, void* stream) {
    GRL_CHECK_ARG(adj && rowptr && bitmap, GRL_E_BADARG, "grl_csr_count_bitmap: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0 && L > 0, GRL_E_BADARG, "grl_csr_count_bitmap: B=%d N=%d L=%d must be positive", B, N, L);
    self_loop = self_loop ? 1 : 0;
    GRL_CHECK_ARG(static_cast<int64_t>(B) * N * (L + 1) < INT32_MAX, GRL_E_UNSUP, "grl_csr_count_bitmap: too many rows for int32");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    uint32_t* bm = static_cast<uint32_t*>(bitmap);
    if (adj_dtype == GRL_F32) return launch_count_bitmap(static_cast<const float*>(adj), B, N, L, self_loop, rowptr, bm, st);
    if (adj_dtype == GRL_F16) return launch_count_bitmap(static_cast<const __half*>(adj), B, N, L, self_loop, rowptr, bm, st);
    set_error("grl_csr_count_bitmap: adjacency dtype %d unsupported (f32 or f16)", adj_dtype);
    return GRL_E_UNSUP;
}

extern "C" int grl_csr_fill_bitmap(const void* adj, int adj_dtype, const void* bitmap, int B, int N, int L, int self_loop,
                                   const int32_t* rowptr, int32_t* col, float* val, void* stream) {
    GRL_CHECK_ARG(adj && bitmap && rowptr && col && val, GRL_E_BADARG, "grl_csr_fill_bitmap: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0 && L > 0, GRL_E_BADARG, "grl_csr_fill_bitmap: B=%d N=%d L=%d must be positive", B, N, L);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    self_loop = self_loop ? 1 : 0;
    const int64_t rows = static_cast<int64_t>(B) * N * L;
    GRL_CHECK_ARG(rows < INT32_MAX, GRL_E_UNSUP, "grl_csr_fill_bitmap: too many rows for int32");
    const int grid = grid_for(rows, 256);                // one thread per row
    const int words = (N + 31) / 32;
    const uint32_t* bm = static_cast<const uint32_t*>(bitmap);
    if (adj_dtype == GRL_F32)
        csr_fill_bitmap_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(adj), bm, words, rows, N, L, self_loop,
                                                            rowptr, col, val);
    else if (adj_dtype == GRL_F16)
        csr_fill_bitmap_kernel<__half><<<grid, 256, 0, st>>>(static_cast<const __half*>(adj), bm, words, rows, N, L, self_loop,
                                                             rowptr, col, val);
    else {
        set_error("grl_csr_fill_bitmap: adjacency dtype %d unsupported (f32 or f16)", adj_dtype);
        return GRL_E_UNSUP;
    }
    GRL_LAUNCH_CHECK("csr_fill_bitmap_kernel");
    return 0;
}

extern "C" int grl_csr_count(const void* adj, int adj_dtype, int B, int N, int L, int self_loop, int32_t* rowptr,
                             void* stream) {
    GRL_CHECK_ARG(adj && rowptr, GRL_E_BADARG, "grl_csr_count: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0 && L > 0, GRL_E_BADARG, "grl_csr_count: B=%d N=%d L=%d must be positive", B, N, L);
    self_loop = self_loop ? 1 : 0;
    GRL_CHECK_ARG(static_cast<int64_t>(B) * N * (L + 1) < INT32_MAX, GRL_E_UNSUP, "grl_csr_count: too many rows for int32");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (adj_dtype == GRL_F32) return launch_count(static_cast<const float*>(adj), B, N, L, self_loop, rowptr, st);
    if (adj_dtype == GRL_F16) return launch_count(static_cast<const __half*>(adj), B, N, L, self_loop, rowptr, st);
    set_error("grl_csr_count: adjacency dtype %d unsupported (f32 or f16)", adj_dtype);
    return GRL_E_UNSUP;
}

extern "C" int grl_csr_fill(const void* adj, int adj_dtype, int B, int N, int L, int self_loop, const int32_t* rowptr,
                            int32_t* col, float* val, void* stream) {
    GRL_CHECK_ARG(adj && rowptr && col && val, GRL_E_BADARG, "grl_csr_fill: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0 && L > 0, GRL_E_BADARG, "grl_csr_fill: B=%d N=%d L=%d must be positive", B, N, L);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    self_loop = self_loop ? 1 : 0;
    if (adj_dtype == GRL_F32) return launch_fill(static_cast<const float*>(adj), B, N, L, self_loop, rowptr, col, val, st);
    if (adj_dtype == GRL_F16) return launch_fill(static_cast<const __half*>(adj), B, N, L, self_loop, rowptr, col, val, st);
    set_error("grl_csr_fill: adjacency dtype %d unsupported (f32 or f16)", adj_dtype);
    return GRL_E_UNSUP;
}

extern "C" size_t grl_scan_workspace_bytes(int64_t n) {
    size_t bytes = 0;
    cub::DeviceScan::InclusiveSum(nullptr, bytes, static_cast<int32_t*>(nullptr), static_cast<int32_t*>(nullptr),
                                  static_cast<int>(n));
    return bytes + 256;
}

extern "C" int grl_scan_inclusive_i32(int32_t* data, int64_t n, void* workspace, size_t workspace_bytes, void* stream) {
    GRL_CHECK_ARG(data && workspace, GRL_E_BADARG, "grl_scan_inclusive_i32: null pointer");
    GRL_CHECK_ARG(n > 0 && n < INT32_MAX, GRL_E_BADARG, "grl_scan_inclusive_i32: bad n");
    size_t need = 0;
    cub::DeviceScan::InclusiveSum(nullptr, need, data, data, static_cast<int>(n));
    GRL_CHECK_ARG(workspace_bytes >= need, GRL_E_WORKSPACE, "grl_scan_inclusive_i32: workspace %zu < %zu", workspace_bytes, need);
    cudaError_t e = cub::DeviceScan::InclusiveSum(workspace, need, data, data, static_cast<int>(n),
                                                  static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return cuda_fail(e, "cub::DeviceScan::InclusiveSum");
    return 0;
}

// rowptr[r] = min(rowptr[r], capacity) for r in [0, R]; nnz_out[0] = the un-clamped entry count.  After this every
// consumer of (rowptr, col, val) stays inside `capacity` slots whatever the batch holds; rows past the capacity come out
// empty and nnz_out > capacity tells the host (graph.PackedGraphBatch reads it one step later, without a stall).
namespace grl {
__global__ void csr_clamp_rowptr_kernel(int32_t* __restrict__ rowptr, int64_t R, int32_t capacity, int32_t* __restrict__ nnz_out) {
    const int64_t tid = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    for (int64_t r = tid; r <= R; r += stride) {
        const int32_t v = rowptr[r];
        if (r == R) nnz_out[0] = v;
        if (v > capacity) rowptr[r] = capacity;
    }
}
}  // namespace grl

extern "C" int grl_csr_clamp_rowptr(int32_t* rowptr, int64_t R, int64_t capacity, int32_t* nnz_out, void* stream) {
    GRL_CHECK_ARG(rowptr && nnz_out, GRL_E_BADARG, "grl_csr_clamp_rowptr: null pointer");
    GRL_CHECK_ARG(R > 0 && capacity >= 0 && capacity < INT32_MAX, GRL_E_BADARG, "grl_csr_clamp_rowptr: bad shape");
    csr_clamp_rowptr_kernel<<<grid_for(R + 1, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        rowptr, R, static_cast<int32_t>(capacity), nnz_out);
    GRL_LAUNCH_CHECK("csr_clamp_rowptr_kernel");
    return 0;
}

extern "C" int grl_coo_count(const int32_t* edges, int64_t E, int B, int N, int L, int32_t* rowptr, void* stream) {
    GRL_CHECK_ARG(rowptr && (edges || E == 0), GRL_E_BADARG, "grl_coo_count: null pointer");
    GRL_CHECK_ARG(E == 0 || aligned_to(edges, 16), GRL_E_ALIGN, "grl_coo_count: edges must be 16-byte aligned");
    const int64_t n_nodes = static_cast<int64_t>(B) * N;
    coo_count_kernel<<<grid_for(E > n_nodes ? E : n_nodes, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        edges, E, N, L, rowptr, n_nodes);
    GRL_LAUNCH_CHECK("coo_count_kernel");
    return 0;
}

extern "C" int grl_coo_fill(const int32_t* edges, const float* weights, int64_t E, int B, int N, int L,
                            const int32_t* rowptr, int32_t* cursor, int32_t* col, float* val, void* stream) {
    GRL_CHECK_ARG(rowptr && cursor && col && val && (edges || E == 0), GRL_E_BADARG, "grl_coo_fill: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t n_nodes = static_cast<int64_t>(B) * N, R = n_nodes * (L + 1);
    cudaError_t e = cudaMemsetAsync(cursor, 0, sizeof(int32_t) * R, st);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(cursor)");
    coo_scatter_kernel<<<grid_for(E > n_nodes ? E : n_nodes, 256), 256, 0, st>>>(edges, weights, E, N, L, rowptr, cursor,
                                                                                   col, val, n_nodes);
    GRL_LAUNCH_CHECK("coo_scatter_kernel");
    csr_sort_rows_kernel<<<grid_for(R, 256), 256, 0, st>>>(rowptr, R, col, val);
    GRL_LAUNCH_CHECK("csr_sort_rows_kernel");
    return 0;
}

extern "C" size_t grl_csr_transpose_workspace_bytes(int64_t n_nodes, int64_t nnz) {
    return sizeof(int32_t) * static_cast<size_t>(n_nodes + 2 * nnz) + 256 + grl_scan_workspace_bytes(n_nodes);
}

// shared body: CSC of an [R x n_cols] CSR matrix whose stored column index is local to a block of rows_per_block rows
// (column of entry e in row r = (r / rows_per_block) * cols_per_block + col[e]); one block = the whole matrix otherwise
static int csr_transpose_impl(const int32_t* rowptr, const int32_t* col, int64_t R, int64_t rows_per_block,
                              int cols_per_block, int64_t n_cols, int64_t nnz, int32_t* colptr, int32_t* rowidx,
                              int32_t* perm, void* workspace, size_t workspace_bytes, void* stream, const char* who) {
    GRL_CHECK_ARG(rowptr && col && colptr && rowidx && perm && workspace, GRL_E_BADARG, "%s: null pointer", who);
    GRL_CHECK_ARG(workspace_bytes >= grl_csr_transpose_workspace_bytes(n_cols, nnz), GRL_E_WORKSPACE,
                  "%s: workspace too small", who);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int32_t* cursor = static_cast<int32_t*>(workspace);
    int32_t* tmp_e = cursor + n_cols;
    int32_t* tmp_r = tmp_e + nnz;
    uintptr_t scan_ws = reinterpret_cast<uintptr_t>(tmp_r + nnz);
    scan_ws = (scan_ws + 255) & ~static_cast<uintptr_t>(255);
    const size_t scan_bytes = grl_scan_workspace_bytes(n_cols) - 256;

    cudaError_t e = cudaMemsetAsync(colptr, 0, sizeof(int32_t) * (n_cols + 1), st);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(colptr)");
    e = cudaMemsetAsync(cursor, 0, sizeof(int32_t) * n_cols, st);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(cursor)");
    const int rpb = static_cast<int>(rows_per_block);
    csc_count_kernel<<<grid_for(R, 256), 256, 0, st>>>(rowptr, col, R, rpb, cols_per_block, colptr);
    GRL_LAUNCH_CHECK("csc_count_kernel");
    int rc = grl_scan_inclusive_i32(colptr + 1, n_cols, reinterpret_cast<void*>(scan_ws), scan_bytes, stream);
    if (rc) return rc;
    csc_scatter_kernel<<<grid_for(R, 256), 256, 0, st>>>(rowptr, col, R, rpb, cols_per_block, colptr, cursor, tmp_e, tmp_r);
    GRL_LAUNCH_CHECK("csc_scatter_kernel");
    csc_rank_kernel<<<grid_for(n_cols * 32, 256), 256, 0, st>>>(colptr, n_cols, tmp_e, tmp_r, rowidx, perm);
    GRL_LAUNCH_CHECK("csc_rank_kernel");
    return 0;
}

extern "C" int grl_csc_typed_regroup(const int32_t* colptr, const int32_t* rowidx, const int32_t* perm, int B, int N, int L,
                                    int32_t* t_rowptr, int32_t* t_col, int32_t* t_perm, void* stream) {
    GRL_CHECK_ARG(colptr && rowidx && perm && t_rowptr && t_col && t_perm, GRL_E_BADARG, "grl_csc_typed_regroup: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0 && L > 0 && L + 1 <= 31, GRL_E_BADARG, "grl_csc_typed_regroup: bad shape (L + 1 must be <= 31)");
    const int64_t n_cols = static_cast<int64_t>(B) * N;
    GRL_CHECK_ARG(n_cols * (L + 1) < (1LL << 31), GRL_E_BADARG, "grl_csc_typed_regroup: row count overflows int32");
    csc_typed_kernel<<<grid_for(n_cols * 32, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(colptr, n_cols, rowidx, perm, L + 1,
                                                                                              N, t_rowptr, t_col, t_perm);
    GRL_LAUNCH_CHECK("csc_typed_kernel");
    return 0;
}

extern "C" int grl_csr_transpose(const int32_t* rowptr, const int32_t* col, int B, int N, int L, int64_t nnz,
                                 int32_t* colptr, int32_t* rowidx, int32_t* perm, void* workspace,
                                 size_t workspace_bytes, void* stream) {
    const int64_t n_nodes = static_cast<int64_t>(B) * N, R = n_nodes * (L + 1);
    return csr_transpose_impl(rowptr, col, R, static_cast<int64_t>(N) * (L + 1), N, n_nodes, nnz, colptr, rowidx, perm,
                              workspace, workspace_bytes, stream, "grl_csr_transpose");
}

extern "C" int grl_csr_transpose_general(const int32_t* rowptr, const int32_t* col, int64_t R, int n_cols, int64_t nnz,
                                         int32_t* colptr, int32_t* rowidx, int32_t* perm, void* workspace,
                                         size_t workspace_bytes, void* stream) {
    GRL_CHECK_ARG(R > 0 && R < (1LL << 31) && n_cols > 0, GRL_E_BADARG, "grl_csr_transpose_general: bad shape");
    return csr_transpose_impl(rowptr, col, R, R, n_cols, n_cols, nnz, colptr, rowidx, perm, workspace, workspace_bytes,
                              stream, "grl_csr_transpose_general");
}
