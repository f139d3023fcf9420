// Typed transpose of the packed adjacency, built directly: one CTA per graph, counters in shared memory.
//
// The backward aggregation U = A_t^T dZ (autograd of GraphConv.forward's bmm, gnn/models/networks/robust_gcn.py:45-50) is a
// forward aggregation over the CSR of the transposed typed adjacency: rows (j, t) list the source nodes i of the entries
// (i, t) -> j in CSR order of the original packing, with t_perm giving that position.  csr_pack.cu builds it through a CSC
// (count / scatter / rank over columns, then a regroup by type: 4 kernels + a device scan + 1, 0.34 ms at cfg2 -- atomics on
// global counters, and every column of ~56 entries ranked O(d^2)).  A graph's adjacency only ever references its own N
// nodes, so here a CTA owns a graph: its N (L+1) typed-row counters (and, when they fit, its columns and the scattered
// keys) live in shared memory, the scan is a block scan, and the entries of a typed row (~8) are put in order in the
// registers of the thread that owns the row.  Same t_rowptr / t_col / t_perm, bit for
// bit (tests/test_gpu_bwd_agg.py).
#include "common.cuh"

namespace grl {

constexpr int kTtThreads = 1024;
constexpr int kTtWarps = kTtThreads / 32;
constexpr int kTtSmall = 16;                         // rows up to this length: sorted in registers by the owning thread
constexpr int kTtWarpStage = 256;                    // rows up to this length: rank sort by a warp through shared memory
constexpr int kTtStageBytes = kTtWarps * kTtWarpStage * 4;   // 32 KB of packed keys
constexpr int kTtMidList = 1024;
constexpr int kTtBigList = 64;
constexpr int kTtSmemBytes = 200 * 1024;             // always requested: 64 registers x 1024 threads leave one CTA per SM anyway

struct TtShared {
    int32_t warp_sums[kTtWarps];
    int32_t total;
    int32_t n_mid;
    int32_t n_big;
    int32_t wide;                                    // a source row longer than 65535 entries: packed keys cannot hold it
    int32_t mid_rows[kTtMidList];
    int32_t big_rows[kTtBigList];
};

__host__ __device__ constexpr size_t tt_align16(size_t x) { return (x + 15) & ~static_cast<size_t>(15); }

// exclusive prefix of v over the CTA's threads (thread order); *total = the sum.  Two barriers.
__device__ __forceinline__ int32_t tt_block_exclusive(int32_t v, TtShared& sh, int32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int32_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    __syncthreads();                                 // the previous use of warp_sums is over
    if (lane == 31) sh.warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const int32_t w = sh.warp_sums[lane];
        int32_t wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int32_t u = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += u;
        }
        sh.warp_sums[lane] = wi - w;
        if (lane == 31) sh.total = wi;
    }
    __syncthreads();
    *total = sh.total;
    return sh.warp_sums[warp] + incl - v;
}

// Entries are scattered as ONE packed word, (source node i) << 16 | (position inside source row (i, t)): ascending order of
// the word is ascending CSR position, so a short row is sorted by a min / max network in registers (two instructions per
// compare-exchange, no dynamic register indexing) and decoded on the way out.
template <int W>
__device__ __forceinline__ void tt_sort_network(uint32_t (&v)[W]) {
#pragma unroll
    for (int k = 2; k <= W; k <<= 1)
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1)
#pragma unroll
            for (int a = 0; a < W; ++a) {
                const int b = a ^ j;
                if (b > a) {
                    const uint32_t lo = min(v[a], v[b]), hi = max(v[a], v[b]);
                    const bool up = (a & k) == 0;
                    v[a] = up ? lo : hi;
                    v[b] = up ? hi : lo;
                }
            }
}

template <int W>
__device__ __forceinline__ void tt_small_row(const uint32_t* tmp, int d, int t, int Lp1, const int32_t* rp_s, int32_t e0,
                                             int32_t* out_col, int32_t* out_perm) {
    uint32_t v[W];
#pragma unroll
    for (int a = 0; a < W; ++a) v[a] = a < d ? tmp[a] : 0xFFFFFFFFu;
    tt_sort_network<W>(v);
#pragma unroll
    for (int a = 0; a < W; ++a)
        if (a < d) {
            const int i = static_cast<int>(v[a] >> 16);
            out_col[a] = i;
            out_perm[a] = e0 + rp_s[i * Lp1 + t] + static_cast<int32_t>(v[a] & 0xFFFFu);
        }
}

__global__ void __launch_bounds__(kTtThreads, 1)
typed_transpose_graph_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, int B, int N, int Lp1,
                             int32_t* __restrict__ t_rowptr, int32_t* t_col, int32_t* t_perm) {
    extern __shared__ __align__(16) unsigned char tt_smem[];
    const int rpg = N * Lp1;
    int32_t* cnt = reinterpret_cast<int32_t*>(tt_smem);                        // [N (L+1)] counters / cursors
    int32_t* rp_s = reinterpret_cast<int32_t*>(tt_smem + tt_align16(static_cast<size_t>(rpg) * 4));   // [N (L+1) + 1] local rowptr
    unsigned char* rest = reinterpret_cast<unsigned char*>(rp_s) + tt_align16(static_cast<size_t>(rpg + 1) * 4);
    const size_t rest_bytes = kTtSmemBytes - (rest - tt_smem);                 // >= kTtStageBytes (launcher)
    uint16_t* col_s = reinterpret_cast<uint16_t*>(rest);                       // staged columns, steps 1-3 ...
    uint32_t* stage = reinterpret_cast<uint32_t*>(rest);                       // ... then [kTtWarps][kTtWarpStage] keys, step 5
    __shared__ TtShared sh;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int b = blockIdx.x; b < B; b += gridDim.x) {
        const int64_t base_r = static_cast<int64_t>(b) * rpg;
        const int32_t e0 = rowptr[base_r];
        const int32_t n_e = rowptr[base_r + rpg] - e0;
        // What fits is staged in shared memory: the graph's columns (coalesced, every load independent -- the passes below
        // chase rowptr -> col -> counter per entry) and the packed keys between the scatter and the sort (else t_perm holds them).
        const size_t col_b = tt_align16(static_cast<size_t>(n_e) * 2);
        const bool col_in = col_b <= rest_bytes;
        const size_t u_b = col_in && col_b > kTtStageBytes ? col_b : kTtStageBytes;
        const bool tmp_in = u_b + static_cast<size_t>(n_e) * 4 <= rest_bytes;
        uint32_t* tmp = tmp_in ? reinterpret_cast<uint32_t*>(rest + u_b) : reinterpret_cast<uint32_t*>(t_perm + e0);
        for (int k = tid; k < rpg; k += kTtThreads) cnt[k] = 0;
        for (int k = tid; k <= rpg; k += kTtThreads) rp_s[k] = rowptr[base_r + k] - e0;
        if (col_in)
            for (int k = tid; k < n_e; k += kTtThreads) col_s[k] = static_cast<uint16_t>(col[e0 + k]);
        if (tid == 0) {
            sh.n_mid = 0;
            sh.n_big = 0;
            sh.wide = 0;
        }
        __syncthreads();
        // 1. entries per transposed typed row (j, t); a thread walks source rows (i, t)
        for (int r = tid; r < rpg; r += kTtThreads) {
            const int t = r % Lp1;
            const int32_t s = rp_s[r], e = rp_s[r + 1];
            if (e - s > 65535) sh.wide = 1;
            if (col_in)
                for (int32_t k = s; k < e; ++k) atomicAdd(&cnt[col_s[k] * Lp1 + t], 1);
            else
                for (int32_t k = s; k < e; ++k) atomicAdd(&cnt[col[e0 + k] * Lp1 + t], 1);
        }
        __syncthreads();
        // 2. exclusive scan in place -> t_rowptr
        {
            const int per = (rpg + kTtThreads - 1) / kTtThreads;
            const int k0 = min(tid * per, rpg), k1 = min(k0 + per, rpg);
            int32_t sum = 0;
            for (int k = k0; k < k1; ++k) sum += cnt[k];
            int32_t total;
            int32_t run = tt_block_exclusive(sum, sh, &total);
            for (int k = k0; k < k1; ++k) {
                const int32_t c = cnt[k];
                cnt[k] = run;
                t_rowptr[base_r + k] = e0 + run;
                run += c;
            }
            if (b == B - 1 && tid == 0) t_rowptr[base_r + rpg] = e0 + total;
        }
        __syncthreads();
        // 3. scatter of packed keys (order inside a row is whatever the atomics gave); cnt[k] ends as the END of row k
        for (int r = tid; r < rpg; r += kTtThreads) {
            const int i = r / Lp1, t = r - i * Lp1;
            const int32_t s = rp_s[r], e = rp_s[r + 1];
            const uint32_t hi = static_cast<uint32_t>(i) << 16;
            if (col_in)
                for (int32_t k = s; k < e; ++k) tmp[atomicAdd(&cnt[col_s[k] * Lp1 + t], 1)] = hi | static_cast<uint32_t>(k - s);
            else
                for (int32_t k = s; k < e; ++k) tmp[atomicAdd(&cnt[col[e0 + k] * Lp1 + t], 1)] = hi | static_cast<uint32_t>(k - s);
        }
        __syncthreads();
        // 4. every row into CSR order of the source packing, decoded into t_col / t_perm.  A thread reads its whole row
        // before it writes (rows are disjoint, so this is safe where the keys sit in t_perm itself).
        const bool wide = sh.wide != 0;
        for (int k = tid; k < rpg; k += kTtThreads) {
            const int32_t s = k ? cnt[k - 1] : 0, d = cnt[k] - s;
            if (d == 0) continue;
            const int t = k % Lp1;
            if (d > kTtWarpStage || wide) {
                const int slot = atomicAdd(&sh.n_big, 1);
                if (slot < kTtBigList) sh.big_rows[slot] = k;
            } else if (d > kTtSmall) {
                const int slot = atomicAdd(&sh.n_mid, 1);
                if (slot < kTtMidList) sh.mid_rows[slot] = k;
            } else if (d > 8) {
                tt_small_row<16>(tmp + s, d, t, Lp1, rp_s, e0, t_col + e0 + s, t_perm + e0 + s);
            } else {
                tt_small_row<8>(tmp + s, d, t, Lp1, rp_s, e0, t_col + e0 + s, t_perm + e0 + s);
            }
        }
        __syncthreads();
        // 5. longer rows: rank sort by a warp (its keys staged where the columns were)
        const int n_mid = sh.n_mid;
        if (n_mid) {
            uint32_t* ws = stage + warp * kTtWarpStage;
            const bool listed = n_mid <= kTtMidList;
            const int n_iter = listed ? n_mid : rpg;
            for (int it = warp; it < n_iter; it += kTtWarps) {
                const int k = listed ? sh.mid_rows[it] : it;
                const int32_t s = k ? cnt[k - 1] : 0, d = cnt[k] - s;
                if (d <= kTtSmall || d > kTtWarpStage) continue;
                const int t = k % Lp1;
                for (int a = lane; a < d; a += 32) ws[a] = tmp[s + a];
                __syncwarp();
                for (int a = lane; a < d; a += 32) {
                    const uint32_t mine = ws[a];
                    int rank = 0;
                    for (int q = 0; q < d; ++q) rank += ws[q] < mine;
                    const int i = static_cast<int>(mine >> 16);
                    t_col[e0 + s + rank] = i;
                    t_perm[e0 + s + rank] = e0 + rp_s[i * Lp1 + t] + static_cast<int32_t>(mine & 0xFFFFu);
                }
                __syncwarp();
            }
        }
        // 6. hub rows (and every row of a graph with a source row too long for the packed keys): rebuilt in order by the
        // whole CTA instead of sorted -- thread i looks for j in source row (i, t), a block scan places the hits.
        const int n_big = sh.n_big;
        if (n_big) {
            const bool listed = n_big <= kTtBigList;
            const int n_iter = listed ? n_big : rpg;
            for (int it = 0; it < n_iter; ++it) {
                const int k = listed ? sh.big_rows[it] : it;
                const int32_t s = k ? cnt[k - 1] : 0, d = cnt[k] - s;
                if (d == 0 || !(d > kTtWarpStage || wide)) continue;           // uniform: every thread reads the same words
                const int j = k / Lp1, t = k - j * Lp1;
                int32_t out = e0 + s;
                for (int i0 = 0; i0 < N; i0 += kTtThreads) {
                    const int i = i0 + tid;
                    int32_t m = 0, rs = 0, re = 0;
                    if (i < N) {
                        rs = e0 + rp_s[i * Lp1 + t];
                        re = e0 + rp_s[i * Lp1 + t + 1];
                        for (int32_t q = rs; q < re; ++q) m += col[q] == j;
                    }
                    int32_t total;
                    int32_t p = out + tt_block_exclusive(m, sh, &total);
                    if (m)
                        for (int32_t q = rs; q < re; ++q)
                            if (col[q] == j) {
                                t_perm[p] = q;
                                t_col[p] = i;
                                ++p;
                            }
                    out += total;
                }
            }
        }
        __syncthreads();
    }
}

}  // namespace grl

using namespace grl;

extern "C" int grl_typed_transpose(const int32_t* rowptr, const int32_t* col, int B, int N, int L, int32_t* t_rowptr,
                                   int32_t* t_col, int32_t* t_perm, void* stream) {
    GRL_CHECK_ARG(rowptr && col && t_rowptr && t_col && t_perm, GRL_E_BADARG, "grl_typed_transpose: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0 && L > 0, GRL_E_BADARG, "grl_typed_transpose: B=%d N=%d L=%d must be positive", B, N, L);
    const int64_t rpg = static_cast<int64_t>(N) * (L + 1);
    GRL_CHECK_ARG(static_cast<int64_t>(B) * rpg < INT32_MAX, GRL_E_BADARG, "grl_typed_transpose: row count overflows int32");
    if (tt_align16(4 * rpg) + tt_align16(4 * (rpg + 1)) + kTtStageBytes > kTtSmemBytes) {
        set_error("grl_typed_transpose: %lld typed rows per graph do not fit in shared memory (use grl_csr_transpose + "
                  "grl_csc_typed_regroup)", static_cast<long long>(rpg));
        return GRL_E_UNSUP;
    }
    cudaError_t e = cudaFuncSetAttribute(typed_transpose_graph_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTtSmemBytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(typed_transpose_graph_kernel)");
    const int grid = B < kNumSMs ? B : kNumSMs;          // 63 registers x 1024 threads: one CTA per SM
    typed_transpose_graph_kernel<<<grid, kTtThreads, kTtSmemBytes, static_cast<cudaStream_t>(stream)>>>(
        rowptr, col, B, N, L + 1, t_rowptr, t_col, t_perm);
    GRL_LAUNCH_CHECK("typed_transpose_graph_kernel");
    return 0;
}
