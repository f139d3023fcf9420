// HBM microbenchmark: what a pure-write / pure-read / copy kernel reaches on this box (the SpMM writes 5x what it reads).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o membw membw.cu ; run: ./membw
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>

enum { ST_DEFAULT = 0, ST_CS = 1, ST_WT = 2, ST_CG = 3 };

template <int MODE>
__device__ __forceinline__ void st16(uint4* p, uint4 v) {
    if (MODE == ST_CS) __stcs(p, v);
    else if (MODE == ST_WT) __stwt(p, v);
    else if (MODE == ST_CG) __stcg(p, v);
    else *p = v;
}

// every warp writes `seg` contiguous bytes per instruction group: lanes cover 512 B (seg=512) or two 256-B pieces
// `pitch` bytes apart (seg=256) -- the slab kernel's pattern
template <int MODE>
__global__ void write_kernel(uint4* out, size_t n16, int seg16, size_t pitch16) {
    const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const uint4 v = make_uint4(tid, 1, 2, 3);
    if (seg16 == 32) {
        for (size_t i = tid; i < n16; i += stride) st16<MODE>(out + i, v);
    } else {
        // row-major [rows, pitch16]; a group of seg16 lanes writes seg16*16 bytes of one row; groups of a warp take
        // consecutive rows (same column chunk) -- like the (graph, chunk) CTAs
        const size_t rows = n16 / pitch16;
        const int chunks = (int)(pitch16 / seg16);
        const int lane = threadIdx.x & 31, g = lane / seg16, gl = lane % seg16, G = 32 / seg16;
        const size_t warp = tid >> 5, nwarps = stride >> 5;
        for (size_t w = warp; w < (rows / G) * chunks; w += nwarps) {
            const size_t chunk = w % chunks, r = (w / chunks) * G + g;
            st16<MODE>(out + r * pitch16 + chunk * seg16 + gl, v);
        }
    }
}

__global__ void read_kernel(const uint4* in, size_t n16, uint32_t* sink) {
    const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    uint32_t acc = 0;
    for (size_t i = tid; i < n16; i += stride) {
        const uint4 v = __ldg(in + i);
        acc += v.x ^ v.y ^ v.z ^ v.w;
    }
    if (acc == 0x12345678u) *sink = acc;
}

__global__ void copy_kernel(const uint4* in, uint4* out, size_t n16) {
    const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = tid; i < n16; i += stride) __stcs(out + i, __ldg(in + i));
}

template <typename F>
static float time_ms(F f, int iters = 10) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) f();
    cudaEventRecord(e0);
    for (int i = 0; i < iters; ++i) f();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    return ms / iters;
}

int main() {
    const size_t bytes = 917504000;            // the gcn1 new_V of cfg2
    const size_t n16 = bytes / 16;
    uint4 *a, *b; uint32_t* sink;
    cudaMalloc(&a, bytes); cudaMalloc(&b, bytes); cudaMalloc(&sink, 4);
    cudaMemset(a, 1, bytes); cudaMemset(b, 2, bytes);
    const int grids[] = {148 * 4, 148 * 8, 148 * 16, 148 * 32};
    const char* names[] = {"default", "cs", "wt", "cg"};
    for (int g : grids) {
        for (int mode = 0; mode < 4; ++mode) {
            float ms = 0;
            auto run = [&](int seg16, size_t pitch16) {
                switch (mode) {
                    case 0: return time_ms([&] { write_kernel<0><<<g, 256>>>(a, n16, seg16, pitch16); });
                    case 1: return time_ms([&] { write_kernel<1><<<g, 256>>>(a, n16, seg16, pitch16); });
                    case 2: return time_ms([&] { write_kernel<2><<<g, 256>>>(a, n16, seg16, pitch16); });
                    default: return time_ms([&] { write_kernel<3><<<g, 256>>>(a, n16, seg16, pitch16); });
                }
            };
            ms = run(32, 0);
            printf("write  st.%-7s grid %5d linear      : %.3f ms  %.0f GB/s\n", names[mode], g, ms, bytes / ms / 1e6);
            ms = run(16, 64);
            printf("write  st.%-7s grid %5d 256B of 1KB : %.3f ms  %.0f GB/s\n", names[mode], g, ms, bytes / ms / 1e6);
        }
        float ms = time_ms([&] { read_kernel<<<g, 256>>>(a, n16, sink); });
        printf("read   grid %5d: %.3f ms  %.0f GB/s\n", g, ms, bytes / ms / 1e6);
        ms = time_ms([&] { copy_kernel<<<g, 256>>>(a, b, n16); });
        printf("copy   grid %5d: %.3f ms  %.0f GB/s (read+write)\n", g, ms, 2.0 * bytes / ms / 1e6);
    }
    float ms = time_ms([&] { cudaMemsetAsync(a, 0, bytes); });
    printf("cudaMemsetAsync: %.3f ms  %.0f GB/s\n", ms, bytes / ms / 1e6);
    ms = time_ms([&] { cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice); });
    printf("cudaMemcpyAsync D2D: %.3f ms  %.0f GB/s (read+write)\n", ms, 2.0 * bytes / ms / 1e6);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
