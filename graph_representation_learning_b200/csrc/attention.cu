// NodeSelfAtten core: per-graph dense self attention, unscaled and unmasked, exact two-pass softmax.
//
// Reference: NodeSelfAtten.forward, gnn/models/networks/robust_gcn.py:90-96
//     s   = softmax(bmm(f_out, g_out^T), dim=-1)        f_out, g_out: [B, N, d]   (d = H/8, already ReLU'd)
//     o   = bmm(s, h_out)                               h_out: [B, N, H]
//     out = gamma * o + V
// and its autograd.  The reference materialises the B x N x N scores three times (bmm, softmax, bmm); here one kernel
// per direction streams 64 x 64 score tiles through shared memory:
//
//   attn_fwd_kernel      CTA = (graph, 64 query rows).  Pass 1 over the key blocks: row max and row sum (the score GEMM
//                        has K = d = 16 / 32, it is cheap to redo).  Pass 2: p = exp(s - max) / sum, O += P h.  In
//                        training the normalised P (and o, before gamma) are saved for the backward; in eval nothing
//                        of size N^2 is ever written, so N = 10 000 (cfg5) needs no 25 GB score tensor.
//   attn_bwd_q_kernel    CTA = (graph, 64 query rows): D_i = do_i . o_i, dP = do h^T, dS = P (dP - D), df = dS g;
//                        dS is written once for the key-major kernel; per-CTA partial sums of d(gamma).
//   attn_bwd_k_kernel    CTA = (graph, 64 key rows): dh = P^T do, dg = dS^T f (reduction over the queries, in order:
//                        no atomics, deterministic).
//
// All products are fp32 FFMA on register tiles (4 x 4 ... 4 x 16 per thread) -- the parity mode of the model is fp32 at
// 1e-5 and these are 2 % of the step's FLOPs; operands are staged in shared memory with pitches chosen so that every
// LDS.128 is conflict-free or a broadcast.  bf16 inputs / outputs are converted at the shared-memory boundary.
#include "common.cuh"

namespace grl {

constexpr int kAttnTile = 64;        // query rows per CTA and key rows per block
constexpr int kAttnThreads = 256;    // 16 x 16 thread grid: ty = 4 rows, tx = interleaved columns

template <typename T> __device__ __forceinline__ float4 attn_ld4(const T* p);
template <> __device__ __forceinline__ float4 attn_ld4<float>(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
template <> __device__ __forceinline__ float4 attn_ld4<__nv_bfloat16>(const __nv_bfloat16* p) {
    const uint2 t = __ldg(reinterpret_cast<const uint2*>(p));
    return make_float4(__uint_as_float(t.x << 16), __uint_as_float(t.x & 0xFFFF0000u), __uint_as_float(t.y << 16),
                       __uint_as_float(t.y & 0xFFFF0000u));
}
template <typename T> __device__ __forceinline__ void attn_st4(T* p, const float4& v);
template <> __device__ __forceinline__ void attn_st4<float>(float* p, const float4& v) { *reinterpret_cast<float4*>(p) = v; }
template <> __device__ __forceinline__ void attn_st4<__nv_bfloat16>(__nv_bfloat16* p, const float4& v) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    *reinterpret_cast<uint2*>(p) = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
}

// 64 rows x COLS of a row-major [*, ld] matrix -> shared memory (pitch PITCH floats) as fp32; rows >= n_rows are zero.
// scale (optional, [COLS]) multiplies column-wise (do = gamma * dOut).
template <typename T, int COLS, int PITCH>
__device__ __forceinline__ void attn_load_rows(float* dst, const T* src, int64_t ld, int row0, int n_rows,
                                               const float* __restrict__ scale, int tid) {
    constexpr int UNITS = kAttnTile * COLS / 4;
#pragma unroll
    for (int u0 = 0; u0 < UNITS; u0 += kAttnThreads) {
        const int u = u0 + tid;
        if (UNITS % kAttnThreads != 0 && u >= UNITS) break;
        const int r = u / (COLS / 4), c4 = (u % (COLS / 4)) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row0 + r < n_rows) {
            v = attn_ld4<T>(src + static_cast<int64_t>(row0 + r) * ld + c4);
            if (scale) {
                const float4 s = __ldg(reinterpret_cast<const float4*>(scale + c4));
                v.x *= s.x; v.y *= s.y; v.z *= s.z; v.w *= s.w;
            }
        }
        *reinterpret_cast<float4*>(dst + r * PITCH + c4) = v;
    }
}

// reduce over the 16 tx lanes of a row group (lanes with equal ty are 16 consecutive lanes)
__device__ __forceinline__ float attn_row_max(float v) {
#pragma unroll
    for (int o = 8; o >= 1; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float attn_row_sum(float v) {
#pragma unroll
    for (int o = 8; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// s[i][j] = sum_k A[ty*4+i][k] * Bm[tx+16j][k]   (both operands row-major in shared memory, K contiguous)
template <int K, int PA, int PB>
__device__ __forceinline__ void attn_tile_nt(const float* A, const float* Bm, int ty, int tx, float (&s)[4][4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) s[i][j] = 0.f;
#pragma unroll 4
    for (int k = 0; k < K; k += 4) {
        float4 a[4], b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(A + (ty * 4 + i) * PA + k);
#pragma unroll
        for (int j = 0; j < 4; ++j) b[j] = *reinterpret_cast<const float4*>(Bm + (tx + 16 * j) * PB + k);
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                s[i][j] = fmaf(a[i].x, b[j].x, s[i][j]);
                s[i][j] = fmaf(a[i].y, b[j].y, s[i][j]);
                s[i][j] = fmaf(a[i].z, b[j].z, s[i][j]);
                s[i][j] = fmaf(a[i].w, b[j].w, s[i][j]);
            }
    }
}

// acc[i][c*4+e] += sum_k A[ty*4+i][k] * Bk[k][c*64 + tx*4 + e]   (A row-major [m][k]; Bk row-major [k][n]; k < 64)
template <int NC, int PA, int PB>
__device__ __forceinline__ void attn_mac_nn(const float* A, const float* Bk, int ty, int tx, float (&acc)[4][NC * 4]) {
#pragma unroll 2
    for (int k = 0; k < kAttnTile; k += 4) {
        float4 a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(A + (ty * 4 + i) * PA + k);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const float4 b = *reinterpret_cast<const float4*>(Bk + (k + kk) * PB + c * 64 + tx * 4);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float av = kk == 0 ? a[i].x : (kk == 1 ? a[i].y : (kk == 2 ? a[i].z : a[i].w));
                    acc[i][c * 4 + 0] = fmaf(av, b.x, acc[i][c * 4 + 0]);
                    acc[i][c * 4 + 1] = fmaf(av, b.y, acc[i][c * 4 + 1]);
                    acc[i][c * 4 + 2] = fmaf(av, b.z, acc[i][c * 4 + 2]);
                    acc[i][c * 4 + 3] = fmaf(av, b.w, acc[i][c * 4 + 3]);
                }
            }
        }
    }
}

// acc[i][c*4+e] += sum_k Ak[k][ty*4+i] * Bk[k][c*64 + tx*4 + e]   (both operands row-major [k][.]; k < 64)
template <int NC, int PA, int PB>
__device__ __forceinline__ void attn_mac_tn(const float* Ak, const float* Bk, int ty, int tx, float (&acc)[4][NC * 4]) {
#pragma unroll 4
    for (int k = 0; k < kAttnTile; ++k) {
        const float4 a = *reinterpret_cast<const float4*>(Ak + k * PA + ty * 4);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const float4 b = *reinterpret_cast<const float4*>(Bk + k * PB + c * 64 + tx * 4);
            const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                acc[i][c * 4 + 0] = fmaf(av[i], b.x, acc[i][c * 4 + 0]);
                acc[i][c * 4 + 1] = fmaf(av[i], b.y, acc[i][c * 4 + 1]);
                acc[i][c * 4 + 2] = fmaf(av[i], b.z, acc[i][c * 4 + 2]);
                acc[i][c * 4 + 3] = fmaf(av[i], b.w, acc[i][c * 4 + 3]);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ forward
template <typename T, int H, int D>
__global__ void __launch_bounds__(kAttnThreads)
attn_fwd_kernel(const T* __restrict__ f, const T* __restrict__ g, const T* __restrict__ h, const T* __restrict__ V,
                const float* __restrict__ gamma, T* __restrict__ out, float* __restrict__ o_save,
                float* __restrict__ P_save, int N, int Np) {
    constexpr int PF = D + 4, PH = H + 4, PP = kAttnTile + 4, NC = H / 64;
    extern __shared__ __align__(16) float attn_smem[];
    float* fs = attn_smem;                       // [64][PF] queries
    float* gs = fs + kAttnTile * PF;             // [64][PF] keys
    float* Ps = gs + kAttnTile * PF;             // [64][PP] probabilities of the current tile
    float* hs = Ps + kAttnTile * PP;             // [64][PH] values

    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int b = blockIdx.y, q0 = blockIdx.x * kAttnTile;
    const int64_t node0 = static_cast<int64_t>(b) * N;
    f += node0 * D; g += node0 * D; h += node0 * H; V += node0 * H; out += node0 * H;

    attn_load_rows<T, D, PF>(fs, f, D, q0, N, nullptr, tid);
    const int n_kb = (N + kAttnTile - 1) / kAttnTile;

    // pass 1: row max and row sum of exp
    float m[4], l[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { m[i] = -INFINITY; l[i] = 0.f; }
    for (int kb = 0; kb < n_kb; ++kb) {
        __syncthreads();                         // previous tile's readers of gs are done (and fs is visible)
        attn_load_rows<T, D, PF>(gs, g, D, kb * kAttnTile, N, nullptr, tid);
        __syncthreads();
        float s[4][4];
        attn_tile_nt<D, PF, PF>(fs, gs, ty, tx, s);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float tm = -INFINITY;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (kb * kAttnTile + tx + 16 * j >= N) s[i][j] = -INFINITY;
                tm = fmaxf(tm, s[i][j]);
            }
            const float m_new = fmaxf(m[i], attn_row_max(tm));
            float ps = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) ps += expf(s[i][j] - m_new);
            l[i] = l[i] * expf(m[i] - m_new) + attn_row_sum(ps);
            m[i] = m_new;
        }
    }
    float inv_l[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) inv_l[i] = 1.0f / l[i];

    // pass 2: P tiles and O += P h
    float acc[4][NC * 4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int c = 0; c < NC * 4; ++c) acc[i][c] = 0.f;
    for (int kb = 0; kb < n_kb; ++kb) {
        __syncthreads();                         // readers of gs / hs / Ps of the previous tile are done
        attn_load_rows<T, D, PF>(gs, g, D, kb * kAttnTile, N, nullptr, tid);
        attn_load_rows<T, H, PH>(hs, h, H, kb * kAttnTile, N, nullptr, tid);
        __syncthreads();
        float s[4][4];
        attn_tile_nt<D, PF, PF>(fs, gs, ty, tx, s);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = q0 + ty * 4 + i;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int col = kb * kAttnTile + tx + 16 * j;
                const float pv = col < N ? expf(s[i][j] - m[i]) * inv_l[i] : 0.f;
                Ps[(ty * 4 + i) * PP + tx + 16 * j] = pv;
                if (P_save && row < N) P_save[(node0 + row) * Np + col] = pv;
            }
        }
        __syncthreads();
        attn_mac_nn<NC, PP, PH>(Ps, hs, ty, tx, acc);
    }

    // epilogue: o (saved for the backward, before gamma) and out = gamma * o + V
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = q0 + ty * 4 + i;
        if (row >= N) continue;
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const int col = c * 64 + tx * 4;
            const float4 o4 = make_float4(acc[i][c * 4], acc[i][c * 4 + 1], acc[i][c * 4 + 2], acc[i][c * 4 + 3]);
            if (o_save) *reinterpret_cast<float4*>(o_save + (node0 + row) * H + col) = o4;
            const float4 gm = __ldg(reinterpret_cast<const float4*>(gamma + col));
            const float4 v4 = attn_ld4<T>(V + static_cast<int64_t>(row) * H + col);
            attn_st4<T>(out + static_cast<int64_t>(row) * H + col,
                        make_float4(fmaf(gm.x, o4.x, v4.x), fmaf(gm.y, o4.y, v4.y), fmaf(gm.z, o4.z, v4.z), fmaf(gm.w, o4.w, v4.w)));
        }
    }
}

// ------------------------------------------------------------------------------------------------ backward, query-major
// dOut: gradient of out [B*N, H];  P: saved probabilities [B, N, Np];  o: saved P h [B*N, H] (fp32)
// writes dS [B, N, Np] (fp32), df [B*N, D] (T), dgamma_part [gridDim.y * gridDim.x, H] (fp32 partial column sums)
template <typename T, int H, int D>
__global__ void __launch_bounds__(kAttnThreads)
attn_bwd_q_kernel(const T* __restrict__ dOut, const float* __restrict__ P, const float* __restrict__ o,
                  const T* __restrict__ g, const T* __restrict__ h, const float* __restrict__ gamma,
                  float* __restrict__ dS, T* __restrict__ df, float* __restrict__ dgamma_part, int N, int Np) {
    constexpr int PF = D + 4, PH = H + 4, PP = kAttnTile + 4, NC = H / 64, ND = D / 16;
    extern __shared__ __align__(16) float attn_smem[];
    float* dos = attn_smem;                      // [64][PH]  do = gamma * dOut of the query block
    float* hs = dos + kAttnTile * PH;            // [64][PH]  values of the key block (also the d(gamma) scratch)
    float* gs = hs + kAttnTile * PH;             // [64][PF]  keys
    float* dSs = gs + kAttnTile * PF;            // [64][PP]

    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int b = blockIdx.y, q0 = blockIdx.x * kAttnTile;
    const int64_t node0 = static_cast<int64_t>(b) * N;
    dOut += node0 * H; o += node0 * H; g += node0 * D; h += node0 * H; df += node0 * D;
    P += node0 * Np; dS += node0 * Np;

    // D_i = do_i . o_i  and the block's partial d(gamma)[c] = sum_i dOut[i][c] * o[i][c]
    float Drow[4];
    float dgp[NC * 4];
#pragma unroll
    for (int c = 0; c < NC * 4; ++c) dgp[c] = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = q0 + ty * 4 + i;
        float part = 0.f;
        if (row < N) {
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                const int col = c * 64 + tx * 4;
                const float4 d4 = attn_ld4<T>(dOut + static_cast<int64_t>(row) * H + col);
                const float4 o4 = __ldg(reinterpret_cast<const float4*>(o + static_cast<int64_t>(row) * H + col));
                const float4 gm = __ldg(reinterpret_cast<const float4*>(gamma + col));
                part += gm.x * d4.x * o4.x + gm.y * d4.y * o4.y + gm.z * d4.z * o4.z + gm.w * d4.w * o4.w;
                dgp[c * 4 + 0] += d4.x * o4.x; dgp[c * 4 + 1] += d4.y * o4.y;
                dgp[c * 4 + 2] += d4.z * o4.z; dgp[c * 4 + 3] += d4.w * o4.w;
            }
        }
        Drow[i] = attn_row_sum(part);
    }
    // reduce d(gamma) over the 16 row groups through shared memory (fixed order: deterministic)
#pragma unroll
    for (int c = 0; c < NC; ++c)
        *reinterpret_cast<float4*>(hs + ty * PH + c * 64 + tx * 4) = make_float4(dgp[c * 4], dgp[c * 4 + 1], dgp[c * 4 + 2], dgp[c * 4 + 3]);
    attn_load_rows<T, H, PH>(dos, dOut, H, q0, N, gamma, tid);
    __syncthreads();
    if (tid < H) {
        float s = 0.f;
#pragma unroll
        for (int r = 0; r < 16; ++r) s += hs[r * PH + tid];
        dgamma_part[(static_cast<int64_t>(blockIdx.y) * gridDim.x + blockIdx.x) * H + tid] = s;
    }

    float dfa[4][ND];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int e = 0; e < ND; ++e) dfa[i][e] = 0.f;

    const int n_kb = (N + kAttnTile - 1) / kAttnTile;
    for (int kb = 0; kb < n_kb; ++kb) {
        __syncthreads();                         // readers of hs / gs / dSs of the previous block are done
        attn_load_rows<T, H, PH>(hs, h, H, kb * kAttnTile, N, nullptr, tid);
        attn_load_rows<T, D, PF>(gs, g, D, kb * kAttnTile, N, nullptr, tid);
        __syncthreads();
        float dp[4][4];
        attn_tile_nt<H, PH, PH>(dos, hs, ty, tx, dp);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = q0 + ty * 4 + i;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int col = kb * kAttnTile + tx + 16 * j;
                float ds = 0.f;
                if (row < N) {                    // P is zero beyond column N (written so by the forward)
                    ds = __ldg(P + static_cast<int64_t>(row) * Np + col) * (dp[i][j] - Drow[i]);
                    dS[static_cast<int64_t>(row) * Np + col] = ds;
                }
                dSs[(ty * 4 + i) * PP + tx + 16 * j] = ds;
            }
        }
        __syncthreads();
        // df[i][tx + 16 e] += sum_k dS[i][k] * g[k][tx + 16 e]
#pragma unroll 4
        for (int k = 0; k < kAttnTile; k += 4) {
            float4 a[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(dSs + (ty * 4 + i) * PP + k);
#pragma unroll
            for (int e = 0; e < ND; ++e) {
                const float b0 = gs[(k + 0) * PF + tx + 16 * e], b1 = gs[(k + 1) * PF + tx + 16 * e];
                const float b2 = gs[(k + 2) * PF + tx + 16 * e], b3 = gs[(k + 3) * PF + tx + 16 * e];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    dfa[i][e] = fmaf(a[i].x, b0, dfa[i][e]);
                    dfa[i][e] = fmaf(a[i].y, b1, dfa[i][e]);
                    dfa[i][e] = fmaf(a[i].z, b2, dfa[i][e]);
                    dfa[i][e] = fmaf(a[i].w, b3, dfa[i][e]);
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = q0 + ty * 4 + i;
        if (row >= N) continue;
#pragma unroll
        for (int e = 0; e < ND; ++e) df[static_cast<int64_t>(row) * D + tx + 16 * e] = from_float<T>(dfa[i][e]);
    }
}

// ------------------------------------------------------------------------------------------------ backward, key-major
// dh[j] = sum_i P[i][j] * do[i],  dg[j] = sum_i dS[i][j] * f[i]   (CTA = 64 key rows j of one graph; i in order)
template <typename T, int H, int D>
__global__ void __launch_bounds__(kAttnThreads)
attn_bwd_k_kernel(const T* __restrict__ dOut, const float* __restrict__ P, const float* __restrict__ dS,
                  const T* __restrict__ f, const float* __restrict__ gamma, T* __restrict__ dh, T* __restrict__ dg,
                  int N, int Np) {
    constexpr int PF = D + 4, PH = H + 4, PP = kAttnTile + 4, NC = H / 64, ND = D / 16;
    extern __shared__ __align__(16) float attn_smem[];
    float* dos = attn_smem;                      // [64 queries][PH]
    float* Ps = dos + kAttnTile * PH;            // [64 queries][PP]  P[i][j0 + .]
    float* dSs = Ps + kAttnTile * PP;            // [64 queries][PP]
    float* fs = dSs + kAttnTile * PP;            // [64 queries][PF]

    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int b = blockIdx.y, j0 = blockIdx.x * kAttnTile;
    const int64_t node0 = static_cast<int64_t>(b) * N;
    dOut += node0 * H; f += node0 * D; dh += node0 * H; dg += node0 * D;
    P += node0 * Np; dS += node0 * Np;

    float acc[4][NC * 4];
    float dga[4][ND];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int c = 0; c < NC * 4; ++c) acc[i][c] = 0.f;
#pragma unroll
        for (int e = 0; e < ND; ++e) dga[i][e] = 0.f;
    }
    const int n_qb = (N + kAttnTile - 1) / kAttnTile;
    for (int qb = 0; qb < n_qb; ++qb) {
        const int i0 = qb * kAttnTile;
        __syncthreads();
        attn_load_rows<T, H, PH>(dos, dOut, H, i0, N, gamma, tid);
        attn_load_rows<T, D, PF>(fs, f, D, i0, N, nullptr, tid);
        attn_load_rows<float, kAttnTile, PP>(Ps, P + j0, Np, i0, N, nullptr, tid);      // columns j0 .. j0+63 < Np
        attn_load_rows<float, kAttnTile, PP>(dSs, dS + j0, Np, i0, N, nullptr, tid);
        __syncthreads();
        attn_mac_tn<NC, PP, PH>(Ps, dos, ty, tx, acc);
#pragma unroll 4
        for (int k = 0; k < kAttnTile; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(dSs + k * PP + ty * 4);
#pragma unroll
            for (int e = 0; e < ND; ++e) {
                const float bv = fs[k * PF + tx + 16 * e];
                dga[0][e] = fmaf(a.x, bv, dga[0][e]);
                dga[1][e] = fmaf(a.y, bv, dga[1][e]);
                dga[2][e] = fmaf(a.z, bv, dga[2][e]);
                dga[3][e] = fmaf(a.w, bv, dga[3][e]);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = j0 + ty * 4 + i;
        if (row >= N) continue;
#pragma unroll
        for (int c = 0; c < NC; ++c)
            attn_st4<T>(dh + static_cast<int64_t>(row) * H + c * 64 + tx * 4,
                        make_float4(acc[i][c * 4], acc[i][c * 4 + 1], acc[i][c * 4 + 2], acc[i][c * 4 + 3]));
#pragma unroll
        for (int e = 0; e < ND; ++e) dg[static_cast<int64_t>(row) * D + tx + 16 * e] = from_float<T>(dga[i][e]);
    }
}

// ------------------------------------------------------------------------------------------------ any (H, D)
// Plain kernels for widths the tiled kernels are not instantiated for (the model allows any net_size; the tests use
// net_size 32 / 64).  One warp per row, scalar loads, the same arithmetic; P_save is required (it doubles as the score
// row buffer).  Not tuned: the tiled kernels serve the benchmark shapes.
template <typename T>
__global__ void __launch_bounds__(256)
attn_generic_fwd_kernel(const T* __restrict__ f, const T* __restrict__ g, const T* __restrict__ h, const T* __restrict__ V,
                        const float* __restrict__ gamma, T* __restrict__ out, float* __restrict__ o_save,
                        float* __restrict__ P, int B, int N, int Np, int H, int D) {
    const int lane = threadIdx.x & 31;
    const int64_t row_g = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;     // b*N + i
    if (row_g >= static_cast<int64_t>(B) * N) return;
    const int64_t node0 = (row_g / N) * N;
    float* Prow = P + row_g * Np;
    float m = -INFINITY;
    for (int j = lane; j < N; j += 32) {
        float sc = 0.f;
        for (int k = 0; k < D; ++k) sc = fmaf(to_float(f[row_g * D + k]), to_float(g[(node0 + j) * D + k]), sc);
        Prow[j] = sc;
        m = fmaxf(m, sc);
    }
    for (int o = 16; o >= 1; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float l = 0.f;
    for (int j = lane; j < N; j += 32) {
        const float e = expf(Prow[j] - m);
        Prow[j] = e;
        l += e;
    }
    for (int o = 16; o >= 1; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
    const float inv = 1.0f / l;
    for (int j = lane; j < Np; j += 32) Prow[j] = j < N ? Prow[j] * inv : 0.f;
    __syncwarp();
    for (int c = lane; c < H; c += 32) {
        float acc = 0.f;
        for (int j = 0; j < N; ++j) acc = fmaf(Prow[j], to_float(h[(node0 + j) * H + c]), acc);
        if (o_save) o_save[row_g * H + c] = acc;
        out[row_g * H + c] = from_float<T>(fmaf(gamma[c], acc, to_float(V[row_g * H + c])));
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
attn_generic_bwd_q_kernel(const T* __restrict__ dOut, const float* __restrict__ P, const float* __restrict__ o,
                          const T* __restrict__ g, const T* __restrict__ h, const float* __restrict__ gamma,
                          float* __restrict__ dS, T* __restrict__ df, int B, int N, int Np, int H, int D) {
    const int lane = threadIdx.x & 31;
    const int64_t row_g = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    if (row_g >= static_cast<int64_t>(B) * N) return;
    const int64_t node0 = (row_g / N) * N;
    float Dr = 0.f;
    for (int c = lane; c < H; c += 32) Dr += gamma[c] * to_float(dOut[row_g * H + c]) * o[row_g * H + c];
    for (int of = 16; of >= 1; of >>= 1) Dr += __shfl_xor_sync(0xffffffffu, Dr, of);
    for (int j = lane; j < Np; j += 32) {
        float ds = 0.f;
        if (j < N) {
            float dp = 0.f;
            for (int c = 0; c < H; ++c) dp = fmaf(gamma[c] * to_float(dOut[row_g * H + c]), to_float(h[(node0 + j) * H + c]), dp);
            ds = P[row_g * Np + j] * (dp - Dr);
        }
        dS[row_g * Np + j] = ds;
    }
    __syncwarp();
    for (int e = lane; e < D; e += 32) {
        float acc = 0.f;
        for (int j = 0; j < N; ++j) acc = fmaf(dS[row_g * Np + j], to_float(g[(node0 + j) * D + e]), acc);
        df[row_g * D + e] = from_float<T>(acc);
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
attn_generic_bwd_k_kernel(const T* __restrict__ dOut, const float* __restrict__ P, const float* __restrict__ dS,
                          const T* __restrict__ f, const float* __restrict__ gamma, T* __restrict__ dh, T* __restrict__ dg,
                          int B, int N, int Np, int H, int D) {
    const int lane = threadIdx.x & 31;
    const int64_t row_g = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;     // b*N + j (key)
    if (row_g >= static_cast<int64_t>(B) * N) return;
    const int64_t node0 = (row_g / N) * N;
    const int j = static_cast<int>(row_g - node0);
    for (int c = lane; c < H; c += 32) {
        float acc = 0.f;
        for (int i = 0; i < N; ++i) acc = fmaf(P[(node0 + i) * Np + j], gamma[c] * to_float(dOut[(node0 + i) * H + c]), acc);
        dh[row_g * H + c] = from_float<T>(acc);
    }
    for (int e = lane; e < D; e += 32) {
        float acc = 0.f;
        for (int i = 0; i < N; ++i) acc = fmaf(dS[(node0 + i) * Np + j], to_float(f[(node0 + i) * D + e]), acc);
        dg[row_g * D + e] = from_float<T>(acc);
    }
}

// dgamma_part[b * n_qb + qb][c] = sum over the 64 rows of the block of dOut * o   (same layout as the tiled kernel)
template <typename T>
__global__ void __launch_bounds__(256)
attn_generic_dgamma_kernel(const T* __restrict__ dOut, const float* __restrict__ o, float* __restrict__ part, int N, int H) {
    const int b = blockIdx.y, qb = blockIdx.x;
    const int64_t node0 = static_cast<int64_t>(b) * N;
    for (int c = threadIdx.x; c < H; c += blockDim.x) {
        float s = 0.f;
        for (int r = qb * kAttnTile; r < min(N, (qb + 1) * kAttnTile); ++r)
            s += to_float(dOut[(node0 + r) * H + c]) * o[(node0 + r) * H + c];
        part[(static_cast<int64_t>(b) * gridDim.x + qb) * H + c] = s;
    }
}

template <typename T>
static int launch_attn_generic(int which, const void* a0, const void* a1, const void* a2, const void* a3, const void* a4,
                               const void* a5, void* o0, void* o1, void* o2, int B, int N, int Np, int H, int D,
                               cudaStream_t st) {
    const int64_t rows = static_cast<int64_t>(B) * N;
    const int blocks = static_cast<int>((rows + 7) / 8);
    if (which == 0) {
        if (!o2) {
            set_error("grl_selfatten_fwd: (H=%d, d=%d) runs on the generic kernel, which needs P_save", H, D);
            return GRL_E_BADARG;
        }
        attn_generic_fwd_kernel<T><<<blocks, 256, 0, st>>>(static_cast<const T*>(a0), static_cast<const T*>(a1),
                                                          static_cast<const T*>(a2), static_cast<const T*>(a3),
                                                          static_cast<const float*>(a4), static_cast<T*>(o0),
                                                          static_cast<float*>(o1), static_cast<float*>(o2), B, N, Np, H, D);
        GRL_LAUNCH_CHECK("attn_generic_fwd_kernel");
    } else if (which == 1) {
        attn_generic_bwd_q_kernel<T><<<blocks, 256, 0, st>>>(static_cast<const T*>(a0), static_cast<const float*>(a1),
                                                            static_cast<const float*>(a2), static_cast<const T*>(a3),
                                                            static_cast<const T*>(a4), static_cast<const float*>(a5),
                                                            static_cast<float*>(o0), static_cast<T*>(o1), B, N, Np, H, D);
        GRL_LAUNCH_CHECK("attn_generic_bwd_q_kernel");
        attn_generic_dgamma_kernel<T><<<dim3((N + kAttnTile - 1) / kAttnTile, B), 256, 0, st>>>(
            static_cast<const T*>(a0), static_cast<const float*>(a2), static_cast<float*>(o2), N, H);
        GRL_LAUNCH_CHECK("attn_generic_dgamma_kernel");
    } else {
        attn_generic_bwd_k_kernel<T><<<blocks, 256, 0, st>>>(static_cast<const T*>(a0), static_cast<const float*>(a1),
                                                            static_cast<const float*>(a2), static_cast<const T*>(a3),
                                                            static_cast<const float*>(a4), static_cast<T*>(o0),
                                                            static_cast<T*>(o1), B, N, Np, H, D);
        GRL_LAUNCH_CHECK("attn_generic_bwd_k_kernel");
    }
    return 0;
}

template <typename T, int H, int D>
static int launch_attn(int which, const void* a0, const void* a1, const void* a2, const void* a3, const void* a4,
                       const void* a5, void* o0, void* o1, void* o2, int B, int N, int Np, cudaStream_t st) {
    constexpr int PF = D + 4, PH = H + 4, PP = kAttnTile + 4;
    const dim3 grid((N + kAttnTile - 1) / kAttnTile, B);
    if (which == 0) {
        constexpr size_t smem = sizeof(float) * kAttnTile * (2 * PF + PP + PH);
        auto kern = attn_fwd_kernel<T, H, D>;
        static bool attr = false;
        if (!attr) {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
            if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(attn_fwd_kernel)");
            attr = true;
        }
        kern<<<grid, kAttnThreads, smem, st>>>(static_cast<const T*>(a0), static_cast<const T*>(a1), static_cast<const T*>(a2),
                                              static_cast<const T*>(a3), static_cast<const float*>(a4), static_cast<T*>(o0),
                                              static_cast<float*>(o1), static_cast<float*>(o2), N, Np);
        GRL_LAUNCH_CHECK("attn_fwd_kernel");
    } else if (which == 1) {
        constexpr size_t smem = sizeof(float) * kAttnTile * (2 * PH + PF + PP);
        auto kern = attn_bwd_q_kernel<T, H, D>;
        static bool attr = false;
        if (!attr) {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
            if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(attn_bwd_q_kernel)");
            attr = true;
        }
        kern<<<grid, kAttnThreads, smem, st>>>(static_cast<const T*>(a0), static_cast<const float*>(a1),
                                              static_cast<const float*>(a2), static_cast<const T*>(a3),
                                              static_cast<const T*>(a4), static_cast<const float*>(a5),
                                              static_cast<float*>(o0), static_cast<T*>(o1), static_cast<float*>(o2), N, Np);
        GRL_LAUNCH_CHECK("attn_bwd_q_kernel");
    } else {
        constexpr size_t smem = sizeof(float) * kAttnTile * (PH + 2 * PP + PF);
        auto kern = attn_bwd_k_kernel<T, H, D>;
        static bool attr = false;
        if (!attr) {
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
            if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(attn_bwd_k_kernel)");
            attr = true;
        }
        kern<<<grid, kAttnThreads, smem, st>>>(static_cast<const T*>(a0), static_cast<const float*>(a1),
                                              static_cast<const float*>(a2), static_cast<const T*>(a3),
                                              static_cast<const float*>(a4), static_cast<T*>(o0), static_cast<T*>(o1), N, Np);
        GRL_LAUNCH_CHECK("attn_bwd_k_kernel");
    }
    return 0;
}

// dispatch on (dtype, H, D): the model has D = H / 8 and H = net_size / 2 in {128, 256}; anything else -> generic kernels
static int dispatch_attn(int which, int dtype, int H, int D, const void* a0, const void* a1, const void* a2, const void* a3,
                         const void* a4, const void* a5, void* o0, void* o1, void* o2, int B, int N, int Np,
                         cudaStream_t st) {
#define GRL_ATTN_CASE(TT, HH, DD) \
    if (H == HH && D == DD) return launch_attn<TT, HH, DD>(which, a0, a1, a2, a3, a4, a5, o0, o1, o2, B, N, Np, st)
    if (dtype == GRL_F32) {
        GRL_ATTN_CASE(float, 64, 16); GRL_ATTN_CASE(float, 128, 16); GRL_ATTN_CASE(float, 256, 32);
    } else if (dtype == GRL_BF16) {
        GRL_ATTN_CASE(__nv_bfloat16, 64, 16); GRL_ATTN_CASE(__nv_bfloat16, 128, 16); GRL_ATTN_CASE(__nv_bfloat16, 256, 32);
    }
#undef GRL_ATTN_CASE
    if (dtype == GRL_F32) return launch_attn_generic<float>(which, a0, a1, a2, a3, a4, a5, o0, o1, o2, B, N, Np, H, D, st);
    if (dtype == GRL_BF16)
        return launch_attn_generic<__nv_bfloat16>(which, a0, a1, a2, a3, a4, a5, o0, o1, o2, B, N, Np, H, D, st);
    set_error("grl_selfatten: dtype %d unsupported", dtype);
    return GRL_E_UNSUP;
}

}  // namespace grl

using namespace grl;

extern "C" int grl_selfatten_fwd(const void* f, const void* g, const void* h, const void* V, const float* gamma, void* out,
                                 float* o_save, float* P_save, int dtype, int B, int N, int H, int D, int Np, void* stream) {
    GRL_CHECK_ARG(f && g && h && V && gamma && out, GRL_E_BADARG, "grl_selfatten_fwd: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0 && B <= 65535, GRL_E_BADARG, "grl_selfatten_fwd: bad shape (B=%d, N=%d)", B, N);
    GRL_CHECK_ARG(H > 0 && D > 0, GRL_E_BADARG, "grl_selfatten_fwd: H and D must be positive");
    GRL_CHECK_ARG(!P_save || Np == (N + kAttnTile - 1) / kAttnTile * kAttnTile, GRL_E_BADARG,
                  "grl_selfatten_fwd: Np must be N rounded up to %d", kAttnTile);
    return dispatch_attn(0, dtype, H, D, f, g, h, V, gamma, nullptr, out, o_save, P_save, B, N, Np,
                         static_cast<cudaStream_t>(stream));
}

extern "C" int grl_selfatten_bwd(const void* dOut, const float* P, const float* o, const void* f, const void* g,
                                 const void* h, const float* gamma, float* dS, void* df, void* dg, void* dh,
                                 float* dgamma_part, int dtype, int B, int N, int H, int D, int Np, void* stream) {
    GRL_CHECK_ARG(dOut && P && o && f && g && h && gamma && dS && df && dg && dh && dgamma_part, GRL_E_BADARG,
                  "grl_selfatten_bwd: null pointer");
    GRL_CHECK_ARG(B > 0 && N > 0 && B <= 65535 && Np == (N + kAttnTile - 1) / kAttnTile * kAttnTile, GRL_E_BADARG,
                  "grl_selfatten_bwd: bad shape (B=%d, N=%d, Np=%d)", B, N, Np);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc = dispatch_attn(1, dtype, H, D, dOut, P, o, g, h, gamma, dS, df, dgamma_part, B, N, Np, st);
    if (rc) return rc;
    return dispatch_attn(2, dtype, H, D, dOut, P, dS, f, gamma, nullptr, dh, dg, nullptr, B, N, Np, st);
}
