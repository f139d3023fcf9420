// Shared-memory slab SpMM, second generation (forward aggregation; the default kernel behind the forward pass).
//
// Reference semantics: new_V = torch.matmul(A', V) (gnn/models/networks/robust_gcn.py:45-47), A' already DropEdge-masked
// (gnn/models/networks/drop_robust_gcn.py:76,80,85).
//
// The first slab kernel (removed in round 2; profiles/r01_s10_*) was issue-bound: ~95 instructions of bookkeeping per typed row and 11.5 per
// entry step (ncu: 333 M warp instructions, 72 % issue-slot utilisation, 0.42 ms for 1.10 GB of compulsory traffic).
// A first node-level rewrite that tagged / compacted the entries inside the SpMM kernel executed 390 M instructions:
// the tagging prologue (~600 per warp iteration) ran once per (node, column chunk).  Hence the split used here:
//
//   grl_dropedge_stream  (ONE pass per layer; replaces grl_dropedge_weights on the forward path)
//       draws the DropEdge keep bits, writes w[e] (the transposed kernel still reads it through perm) and a
//       NODE-level entry stream: the kept entries of all (L+1) typed rows of a node, in CSR order, each tagged with
//       its row type:  ew[start + k] = (t << 24 | col, weight),  nodeinfo[node] = (start, kept).
//   spmm_slab2_kernel    (one launch per input; persistent CTAs over (graph, column chunk) items)
//       * one CTA stages the [N, FC] feature slab of one graph with TMA; when two slabs fit, the next item's slab is
//         fetched while the current one is walked;
//       * a lane group copies its node's stream (coalesced, prefetched one node ahead) into a per-warp staging area,
//         turning col into a slab byte offset, and appends a terminator;
//       * the walk reads the stream back with broadcast LDS (no shuffles): "new typed row" is ONE compare
//         `x >= (t_open+1) << 24`, the gather address is `S + x` with S = lane slab base - (t_open << 24), the
//         gather is a conflict-free LDS.128, the arithmetic is two packed fma.rn.f32x2 (fp32) per 16 bytes.
//   Dropped entries never reach the walk (30 % fewer steps), and every kernel still sums a typed row in CSR order with
//   the dropped entries skipped, so all forward kernels stay bit-identical.
#include "gemm_common.cuh"

namespace grl {
using namespace sm100;

struct Slab2Params {
    const int2* nodeinfo;  // [B*N] (start, kept)
    const uint2* ew;       // node streams (t << 24 | col, weight bits)
    uint8_t* Y;            // bytes
    int64_t ldy_bytes;
    int N, Lp1, n_chunks, box_rows, n_boxes, n_items, n_buf;
    uint32_t slab_bytes;   // one slab buffer (multiple of 128)
};

// Accumulators of one 16-byte slab word per lane.  fp32: two packed f32x2 registers (FFMA2 / FMUL2, IEEE per half, so
// bit-identical to fmaf); bf16: eight fp32 scalars.  gather() is split from the arithmetic so that the walk can keep
// the next entries' LDS.128 in flight while it accumulates the current ones.
template <typename T> struct SlabAcc;
template <> struct SlabAcc<float> {
    struct Q { uint64_t q01, q23; };
    uint64_t a01, a23;
    __device__ static __forceinline__ Q gather(uint32_t addr) {
        Q q;
        asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(q.q01), "=l"(q.q23) : "r"(addr));
        return q;
    }
    __device__ __forceinline__ void zero() { a01 = 0ull; a23 = 0ull; }
    __device__ __forceinline__ void fma(const Q& q, uint32_t wbits) {
        uint64_t ww;
        asm("mov.b64 %0, {%1, %1};" : "=l"(ww) : "r"(wbits));
        asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(a01) : "l"(q.q01), "l"(ww));
        asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(a23) : "l"(q.q23), "l"(ww));
    }
    __device__ __forceinline__ void mul(const Q& q, uint32_t wbits) {          // first entry of a typed row: acc = w * x
        uint64_t ww;
        asm("mov.b64 %0, {%1, %1};" : "=l"(ww) : "r"(wbits));
        asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(a01) : "l"(q.q01), "l"(ww));
        asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(a23) : "l"(q.q23), "l"(ww));
    }
    __device__ __forceinline__ void store_streaming(uint8_t* dst) const {     // straight from the packed registers
        asm volatile("st.global.cs.v2.b64 [%0], {%1, %2};" ::"l"(dst), "l"(a01), "l"(a23) : "memory");
    }
};
template <> struct SlabAcc<__nv_bfloat16> {
    struct Q { uint32_t ww[4]; };
    float acc[8];
    __device__ static __forceinline__ Q gather(uint32_t addr) {
        Q q;
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(q.ww[0]), "=r"(q.ww[1]), "=r"(q.ww[2]), "=r"(q.ww[3]) : "r"(addr));
        return q;
    }
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = 0.0f;
    }
    __device__ __forceinline__ void fma(const Q& q, uint32_t wbits) {
        const float wt = __uint_as_float(wbits);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            acc[2 * i] = fmaf(wt, __uint_as_float(q.ww[i] << 16), acc[2 * i]);
            acc[2 * i + 1] = fmaf(wt, __uint_as_float(q.ww[i] & 0xFFFF0000u), acc[2 * i + 1]);
        }
    }
    __device__ __forceinline__ void mul(const Q& q, uint32_t wbits) {
        const float wt = __uint_as_float(wbits);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            acc[2 * i] = wt * __uint_as_float(q.ww[i] << 16);
            acc[2 * i + 1] = wt * __uint_as_float(q.ww[i] & 0xFFFF0000u);
        }
    }
    __device__ __forceinline__ void store_streaming(uint8_t* dst) const {
        uint32_t o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(acc[2 * i], acc[2 * i + 1]);
            o[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        __stcs(reinterpret_cast<uint4*>(dst), make_uint4(o[0], o[1], o[2], o[3]));
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
    static constexpr int CAP = 64;                              // kept entries staged per batch (cfg2: 34 +- 6 per node)
    static constexpr int J = CAP / GL;
    static constexpr int G = 32 / GL;
    static constexpr int STAGE_SLOTS = CAP + 8;                 // + terminator, padding of the walk to 4, read-ahead
    static constexpr int STAGE_BYTES_PER_WARP = G * STAGE_SLOTS * 8;
};


// ------------------------------------------------------------------------------------------------ stream producer
// One warp per node.  Philox indexing (block e >> 2, word e & 3) is the one grl_dropedge_weights uses, so both entry
// points produce the same w for the same (seed, stream_id).
__global__ void __launch_bounds__(256)
dropedge_stream_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, const float* __restrict__ val,
                       const uint8_t* __restrict__ mask, int64_t n_nodes, int Lp1, float scale, uint32_t thresh,
                       uint64_t seed, uint64_t stream_id, float* __restrict__ w, uint2* __restrict__ ew,
                       int2* __restrict__ nodeinfo, const int32_t* __restrict__ perm = nullptr,
                       const uint64_t* __restrict__ rng_off = nullptr) {
    stream_id = rng_stream(stream_id, rng_off);
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    for (int64_t node = warp; node < n_nodes; node += n_warps) {
        const int32_t rp = lane <= Lp1 ? __ldg(rowptr + node * Lp1 + lane) : 0;
        const int32_t e0 = __shfl_sync(0xffffffffu, rp, 0), eN = __shfl_sync(0xffffffffu, rp, Lp1);
        int kept = 0;
        for (int32_t eb = e0; eb < eN; eb += 32) {
            const int32_t e = eb + lane;
            const bool inb = e < eN;
            int32_t c = 0;
            float wv = 0.0f;
            if (inb) {
                c = __ldg(col + e);
                bool keep = true;
                if (mask) keep = mask[e] != 0;
                else if (thresh) keep = Philox::word(seed, stream_id, static_cast<uint64_t>(e)) >= thresh;
                wv = keep ? __ldg(val + (perm ? perm[e] : e)) * scale : 0.0f;    // perm: weights of another entry order
                w[e] = wv;
            }
            uint32_t t = 0;
            for (int tt = 1; tt < Lp1; ++tt) t += (e >= __shfl_sync(0xffffffffu, rp, tt)) ? 1u : 0u;
            const bool live = inb && wv != 0.0f;
            const uint32_t m = __ballot_sync(0xffffffffu, live);
            if (live) ew[e0 + kept + __popc(m & lt)] = make_uint2((t << 24) | static_cast<uint32_t>(c), __float_as_uint(wv));
            kept += __popc(m);
        }
        if (lane == 0) nodeinfo[node] = make_int2(e0, kept);
    }
}

// Typed rows (t_from, t_to) of a node that kept no entry: written as zeros.  Out of line: rare on dense synthetic
// graphs, common on real documents (padding nodes carry only the self loop).
__device__ __noinline__ void slab2_zero_rows(uint8_t* yfirst, uint32_t ldy, uint32_t n_rows, uint32_t word_pitch, int words) {
    for (uint32_t tt = 0; tt < n_rows; ++tt)
        for (int v = 0; v < words; ++v)
            __stcs(reinterpret_cast<uint4*>(yfirst + static_cast<size_t>(tt) * ldy + v * word_pitch), make_uint4(0, 0, 0, 0));
}

// VW = 16-byte words per lane (1 or 2): word v of lane gl sits at byte gl*16 + v*GL*16 of the slab / output row, so every
// LDS.128 / st.global of a group stays one contiguous GL*16-byte segment.  VW = 2 halves the bookkeeping instructions per
// byte (entry decode, compare, address) at 512 threads per CTA (the second accumulator pair needs the registers).
template <typename T, int GL, int VW>
__global__ void __launch_bounds__(VW == 2 ? 512 : 1024, 1)
spmm_slab2_kernel(const Slab2Params p, const __grid_constant__ CUtensorMap tmX) {
    using Cfg = Slab2Cfg<GL>;
    constexpr int G = Cfg::G, J = Cfg::J, CAP = Cfg::CAP;
    constexpr uint32_t WORD_PITCH = GL * 16;           // distance between the words of one lane
    constexpr uint32_t ROW_BYTES = GL * 16 * VW;       // one slab row = the chunk's FC columns
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t bar[2];

    // buffer b: [zero row][slab rows...] at base + b * slab_bytes; staging areas behind the buffers
    const uint32_t base_u32 = (smem_u32(smem_raw) + 127u) & ~127u;
    uint8_t* const base_ptr = smem_raw + (base_u32 - smem_u32(smem_raw));
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const int grp = lane / GL, gl = lane % GL;
    const uint32_t stage_g = base_u32 + p.n_buf * p.slab_bytes + (warp * G + grp) * (Cfg::STAGE_SLOTS * 8);

    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_mbar_init();
    }
    for (int q = threadIdx.x; q < p.n_buf * (int)(ROW_BYTES / 16); q += blockDim.x) {   // the zero rows (terminators point here)
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
    const uint32_t ldy = static_cast<uint32_t>(p.ldy_bytes);
    const int step = n_warps * G;
    int buf = 0;
    uint32_t phases = 0;
    int item = blockIdx.x;
    if (threadIdx.x == 0 && item < p.n_items) issue_fill(item, 0);

    uint2 re[J];                                        // this lane's share of the node stream (raw entries)
    auto load_raw = [&](int32_t start, int32_t count) { // entries start + gl + GL*j, j < J, of `count` remaining
#pragma unroll
        for (int j = 0; j < J; ++j) {
            const int k = gl + GL * j;
            re[j] = make_uint2(0u, 0u);
            if (k < count) re[j] = __ldg(p.ew + start + k);
        }
    };

    for (; item < p.n_items; item += gridDim.x) {
        const int nxt = item + gridDim.x;
        if (p.n_buf == 2 && threadIdx.x == 0 && nxt < p.n_items) issue_fill(nxt, buf ^ 1);
        const int chunk = item % p.n_chunks, b = item / p.n_chunks;
        const int64_t node0 = static_cast<int64_t>(b) * p.N;
        const uint32_t slab_lane = base_u32 + buf * p.slab_bytes + gl * 16;     // the zero row; node j is row j + 1
        uint8_t* const ychunk = p.Y + static_cast<size_t>(chunk) * ROW_BYTES + gl * 16;
        // A group whose node index runs past the graph (last round of some warps) repeats the graph's last node: the
        // duplicate writes the same bytes to the same addresses, and the walk needs no "is this node real" predicate.
        auto load_ni = [&](int ii) -> int2 { return __ldg(p.nodeinfo + node0 + min(ii, p.N - 1)); };
        using Q = typename SlabAcc<T>::Q;

        int i = warp * G + grp;
        int2 ni = load_ni(i);
        load_raw(ni.x, ni.y);
        int2 ni_next = load_ni(i + step);
        mbar_wait(&bar[buf], (phases >> buf) & 1u);      // slab landed (the stream loads above overlap the fill)
        phases ^= 1u << buf;

        for (; (i - grp) < p.N; i += step) {
            uint8_t* const ynode = ychunk + (node0 + min(i, p.N - 1)) * Lp1 * p.ldy_bytes;
            const int maxkept = __reduce_max_sync(0xffffffffu, ni.y);
            // walk state: thr = (open row type + 1) << 24 (0: nothing open yet); S + x = gather address of entry x of
            // the open row; yrow = output row of the open typed row (one row before the node's first while none is)
            uint32_t thr = 0;
            uint8_t* yrow = ynode - ldy;
            SlabAcc<T> acc[VW];
#pragma unroll
            for (int v = 0; v < VW; ++v) acc[v].zero();

            // One entry (x = t << 24 | slab byte offset).  The first entry of a typed row stores the previous row,
            // zero-fills skipped rows, re-bases the state and starts the accumulator with a multiply
            // (w*x == fma(w, x, +0) up to the sign of zero).
            // The gather address never depends on the walk state: S + x = slab_lane - (t_open << 24) + (t << 24 | offset) is
            // slab_lane + offset for every entry of the open row, and a row opening re-bases S first.  So the gather of an
            // entry is issued ONE STEP AHEAD (before the previous entry's compare / FMAs), which takes the LDS latency off
            // the walk's dependency chain at the price of one mask per entry.
            auto gather_of = [&](uint32_t x, Q (&q)[VW]) {
                const uint32_t a = slab_lane + (x & 0x00FFFFFFu);
#pragma unroll
                for (int v = 0; v < VW; ++v) q[v] = SlabAcc<T>::gather(a + v * WORD_PITCH);
            };
            auto walk_step = [&](uint32_t x, uint32_t wbits, const Q (&q)[VW]) {
                if (x >= thr) {
                    const uint32_t d = x - thr;                   // (typed rows skipped) << 24 | column offset
                    if (thr != 0u) {
#pragma unroll
                        for (int v = 0; v < VW; ++v) acc[v].store_streaming(yrow + v * WORD_PITCH);
                    }
                    if (d >= 0x01000000u) slab2_zero_rows(yrow + ldy, ldy, d >> 24, WORD_PITCH, VW);
                    uint64_t yn;
                    asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(yn) : "r"((d >> 24) + 1u), "r"(ldy), "l"(reinterpret_cast<uint64_t>(yrow)));
                    yrow = reinterpret_cast<uint8_t*>(yn);
                    const uint32_t tb = x & 0xFF000000u;
                    thr = tb + 0x01000000u;
#pragma unroll
                    for (int v = 0; v < VW; ++v) acc[v].mul(q[v], wbits);
                } else {
#pragma unroll
                    for (int v = 0; v < VW; ++v) acc[v].fma(q[v], wbits);
                }
            };

            for (int eb = 0;; eb += CAP) {
                // ---------------- stage the batch held in `re`: col -> slab byte offset, then the terminator(s)
                const bool final_warp = eb + CAP >= maxkept;
                const bool final_g = eb + CAP >= ni.y;
                const int rem = min(maxkept - eb, CAP);
                const int cnt = max(0, min(ni.y - eb, CAP));
#pragma unroll
                for (int j = 0; j < J; ++j) {
                    if (GL * j < rem) {
                        const int k = gl + GL * j;
                        if (k < cnt)
                            sts64(stage_g + k * 8, (re[j].x & 0xFF000000u) | (((re[j].x & 0x00FFFFFFu) + 1u) * ROW_BYTES), re[j].y);
                    }
                }
                // terminator: end of node -> row type Lp1 (closes the last row, zero-fills the rest); end of a batch of
                // a longer node -> the row type of the batch's last entry with weight 0 (state carries over)
                const uint32_t t_last = __shfl_sync(0xffffffffu, re[J - 1].x, GL - 1, GL) & 0xFF000000u;
                const uint32_t term = final_g ? static_cast<uint32_t>(Lp1) << 24 : t_last;
                int n = __reduce_max_sync(0xffffffffu, cnt + 1);
                n = (n + 3) & ~3;
                for (int k = cnt + gl; k < n + 2; k += GL) sts64(stage_g + k * 8, term, 0u);
                __syncwarp();

                // ---------------- refill the registers: next batch of this node, or the next node's first batch
                if (final_warp) {
                    ni = ni_next;
                    load_raw(ni.x, ni.y);
                    ni_next = load_ni(i + 2 * step);
                } else {
                    load_raw(ni.x + eb + CAP, ni.y - eb - CAP);
                }

                // ---------------- walk: 4 steps per iteration, entries fetched two at a time
                uint32_t sa = stage_g;
                uint4 ea = lds128(sa);
                Q qa[VW], qb[VW];
                gather_of(ea.x, qa);
#pragma unroll 1
                for (int k = 0; k < n; k += 4) {             // every gather is issued one step ahead of its use (two ahead
                    const uint4 eb2 = lds128(sa + 16);       // measured equal in fp32 and slower in bf16: registers)
                    gather_of(ea.z, qb);
                    walk_step(ea.x, ea.y, qa);
                    gather_of(eb2.x, qa);
                    walk_step(ea.z, ea.w, qb);
                    ea = lds128(sa + 32);
                    sa += 32;
                    gather_of(eb2.z, qb);
                    walk_step(eb2.x, eb2.y, qa);
                    gather_of(ea.x, qa);                     // first entry of the next iteration (a terminator past the end)
                    walk_step(eb2.z, eb2.w, qb);
                }
                __syncwarp();                                // staging is rewritten by the next batch
                if (final_warp) break;
            }
        }
        __syncthreads();                                     // every warp is done with slab[buf]
        if (p.n_buf == 1 && threadIdx.x == 0 && nxt < p.n_items) issue_fill(nxt, 0);
        if (p.n_buf == 2) buf ^= 1;
    }
}

template <typename T, int GL, int VW>
static int launch_slab2(const Slab2Params& p, const CUtensorMap& tm, int grid, size_t smem, int threads, cudaStream_t st) {
    auto kern = spmm_slab2_kernel<T, GL, VW>;
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

int g_slab2_gl = 0;   // 0 = choose the group width, 8 / 16 / 32 = force it (grl_spmm_set_tile_mode 108 / 116 / 132)
constexpr bool kSlab2PreferTwoWords = false;   // flipped once the A/B (scripts/spmm_bench.py --modes=116,208) says so
int g_slab2_vw = 0;   // 0 = choose, 1 / 2 = force the 16-byte words per lane (grl_spmm_set_tile_mode 2xx = two words)

// returns 0 = launched, 1 = the slab of one graph does not fit / layout not TMA-able, other = error
static int try_launch_spmm_slab2(const int2* nodeinfo, const uint2* ew, const void* X, int64_t ldx, void* Y, int64_t ldy,
                                 int dtype, int B, int N, int L, int F, cudaStream_t st) {
    const int es = dtype == GRL_F32 ? 4 : 2;
    if (!aligned_to(X, 16) || (ldx * es) % 16 != 0 || !aligned_to(Y, 16) || (ldy * es) % 16 != 0) return 1;
    if (L + 1 > 31 || N >= (1 << 24) - 2) return 1;               // row type: 8 bits, column: 24 bits of a stream entry
    const int n_boxes = (N + 255) / 256;
    const int box_rows = (N + n_boxes - 1) / n_boxes;
    const int rows = n_boxes * box_rows;
    constexpr size_t kMaxSmem = 227 * 1024 - 1024;

    struct Plan { int gl, vw, threads, n_buf; size_t smem; uint32_t slab_bytes; };
    auto plan_for = [&](int gl, int vw, Plan* out) -> bool {
        if (F % (gl * vw * 16 / es) != 0) return false;
        const int G = 32 / gl;
        int threads = vw == 2 ? 512 : 1024;
        while (threads > 32 && (threads / 32) * G > N) threads >>= 1;              // tiny graphs: no idle warps
        const size_t stage = static_cast<size_t>(threads / 32) * G * (64 + 8) * 8;  // Slab2Cfg::STAGE_SLOTS
        const size_t slab = static_cast<size_t>(rows + 1) * gl * vw * 16;           // + the zero row
        if (slab >= (1u << 24)) return false;
        const size_t one = slab + stage + 128, two = 2 * slab + stage + 128;
        if (one > kMaxSmem) return false;
        out->gl = gl; out->vw = vw; out->threads = threads; out->slab_bytes = static_cast<uint32_t>(slab);
        out->n_buf = two <= kMaxSmem ? 2 : 1;
        out->smem = out->n_buf == 2 ? two : one;
        return true;
    };
    Plan pl{};
    bool have = false;
    if (g_slab2_gl) {
        have = plan_for(g_slab2_gl, g_slab2_vw ? g_slab2_vw : 1, &pl);
    } else {
        for (int gl : {32, 16, 8})                       // the widest group that fits (fewest column chunks per graph)
            if (plan_for(gl, 1, &pl)) { have = true; break; }
        if (have && kSlab2PreferTwoWords && pl.gl >= 16) {   // same row width with half the lanes and two words per lane
            Plan p2{};
            if (plan_for(pl.gl / 2, 2, &p2)) pl = p2;
        }
    }
    if (!have) return 1;
    const int n_chunks = F / (pl.gl * pl.vw * 16 / es);
    const int n_items = B * n_chunks;
    int per_sm = static_cast<int>((228 * 1024) / (pl.smem + 1024));
    if (per_sm > 2048 / pl.threads) per_sm = 2048 / pl.threads;
    if (per_sm < 1) per_sm = 1;
    const int grid = n_items < kNumSMs * per_sm ? n_items : kNumSMs * per_sm;

    CUtensorMap tm;
    int rc = make_tmap_plain(&tm, X, es, static_cast<int64_t>(B) * N, F, ldx, pl.gl * pl.vw * 16 / es, box_rows);
    if (rc) return rc;
    Slab2Params p{nodeinfo, ew, static_cast<uint8_t*>(Y), ldy * es, N, L + 1, n_chunks, box_rows, n_boxes, n_items,
                  pl.n_buf, pl.slab_bytes};
#define GRL_SLAB2_CASE(TT, V) \
    if (pl.gl == V && pl.vw == 1) return launch_slab2<TT, V, 1>(p, tm, grid, pl.smem, pl.threads, st); \
    if (pl.gl == V && pl.vw == 2) return launch_slab2<TT, V, 2>(p, tm, grid, pl.smem, pl.threads, st)
    if (dtype == GRL_F32) {
        GRL_SLAB2_CASE(float, 8); GRL_SLAB2_CASE(float, 16); GRL_SLAB2_CASE(float, 32);
    } else {
        GRL_SLAB2_CASE(__nv_bfloat16, 8); GRL_SLAB2_CASE(__nv_bfloat16, 16); GRL_SLAB2_CASE(__nv_bfloat16, 32);
    }
#undef GRL_SLAB2_CASE
    return 1;
}

}  // namespace grl

using namespace grl;

extern "C" int grl_dropedge_stream(const int32_t* rowptr, const int32_t* col, const float* val, const uint8_t* mask,
                                   int64_t n_nodes, int L, float p, uint64_t seed, uint64_t stream_id, float* w,
                                   void* ew, int32_t* nodeinfo, void* stream) {
    GRL_CHECK_ARG(rowptr && col && val && w && ew && nodeinfo, GRL_E_BADARG, "grl_dropedge_stream: null pointer");
    GRL_CHECK_ARG(p >= 0.0f && p < 1.0f, GRL_E_BADARG, "grl_dropedge_stream: p=%f outside [0,1)", p);
    GRL_CHECK_ARG(n_nodes > 0 && L > 0 && L + 1 <= 31, GRL_E_BADARG,
                  "grl_dropedge_stream: bad shape (a node's L + 2 row pointers live one per lane: L + 1 must be <= 31)");
    GRL_CHECK_ARG(aligned_to(ew, 8) && aligned_to(nodeinfo, 8), GRL_E_ALIGN, "grl_dropedge_stream: ew / nodeinfo need 8-byte alignment");
    int64_t blocks = (n_nodes + 7) / 8;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    dropedge_stream_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        rowptr, col, val, mask, n_nodes, L + 1, 1.0f / (1.0f - p), p > 0.0f ? keep_threshold(p) : 0u, seed, stream_id, w,
        static_cast<uint2*>(ew), reinterpret_cast<int2*>(nodeinfo), nullptr, rng_offset_ptr());
    GRL_LAUNCH_CHECK("dropedge_stream_kernel");
    return 0;
}

extern "C" int grl_stream_from_weights(const int32_t* rowptr, const int32_t* col, const int32_t* perm, const float* w_src,
                                      int64_t n_nodes, int L, float* w, void* ew, int32_t* nodeinfo, void* stream) {
    GRL_CHECK_ARG(rowptr && col && perm && w_src && w && ew && nodeinfo, GRL_E_BADARG, "grl_stream_from_weights: null pointer");
    GRL_CHECK_ARG(n_nodes > 0 && L > 0 && L + 1 <= 31, GRL_E_BADARG, "grl_stream_from_weights: bad shape (L + 1 must be <= 31)");
    GRL_CHECK_ARG(aligned_to(ew, 8) && aligned_to(nodeinfo, 8), GRL_E_ALIGN, "grl_stream_from_weights: ew / nodeinfo need 8-byte alignment");
    int64_t blocks = (n_nodes + 7) / 8;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    dropedge_stream_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        rowptr, col, w_src, nullptr, n_nodes, L + 1, 1.0f, 0u, 0ull, 0ull, w, static_cast<uint2*>(ew),
        reinterpret_cast<int2*>(nodeinfo), perm);
    GRL_LAUNCH_CHECK("dropedge_stream_kernel<perm>");
    return 0;
}

extern "C" int grl_spmm_fwd_stream(const int32_t* nodeinfo, const void* ew, const void* X, int64_t ldx, void* Y,
                                   int64_t ldy, int dtype, int B, int N, int L, int F, void* stream) {
    GRL_CHECK_ARG(nodeinfo && ew && X && Y, GRL_E_BADARG, "grl_spmm_fwd_stream: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0 && L > 0 && F > 0, GRL_E_BADARG, "grl_spmm_fwd_stream: bad shape");
    GRL_CHECK_ARG(ldx >= F && ldy >= F, GRL_E_BADARG, "grl_spmm_fwd_stream: leading dimension < F");
    GRL_CHECK_ARG(dtype == GRL_F32 || dtype == GRL_BF16, GRL_E_UNSUP, "grl_spmm_fwd_stream: dtype %d unsupported", dtype);
    const int rc = try_launch_spmm_slab2(reinterpret_cast<const int2*>(nodeinfo), static_cast<const uint2*>(ew), X, ldx, Y,
                                         ldy, dtype, B, N, L, F, static_cast<cudaStream_t>(stream));
    if (rc == 1) {
        set_error("grl_spmm_fwd_stream: one graph's feature slab (N=%d, F=%d) does not fit in shared memory or the "
                  "layout is not 16-byte aligned; use grl_spmm_fwd", N, F);
        return GRL_E_UNSUP;
    }
    return rc;
}
