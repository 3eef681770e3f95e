// Probe: how fast can [64,3,640,640] f32 be WRITTEN with different CTA -> address mappings? (B200, sm_100a)
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o store_pattern store_pattern.cu
#include <cstdio>
#include <cuda_runtime.h>
constexpr int N = 64, H = 640, W = 640;
__device__ __forceinline__ void st4(float* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__global__ void linear_fill(float4* o, size_t n4, float v) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) o[i] = make_float4(v, v, v, v);
}
// tile: TWx x TH rows x 3 planes per CTA of 256 threads; each warp writes rows, lanes 16 B each over TW=128 px
template <int TH, bool HINT>
__global__ void tile128(float* o, float v) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tx0 = blockIdx.x * 128, ty0 = blockIdx.y * TH, z = blockIdx.z;
    for (int rr = 0; rr < TH / 8; rr++) {
        const int oy = ty0 + warp * (TH / 8) + rr;
        for (int c = 0; c < 3; c++) {
            float* d = o + (((size_t)z * 3 + c) * H + oy) * W + tx0 + lane * 4;
            if (HINT) st4(d, make_float4(v, v, v, v)); else *reinterpret_cast<float4*>(d) = make_float4(v, v, v, v);
        }
    }
}
// full-row tiles: CTA of 160 threads per row segment of 640 px; TH rows x 3 planes
template <int TH>
__global__ void rowtile(float* o, float v) {
    const int ty0 = blockIdx.x * TH, z = blockIdx.y;
    for (int i = threadIdx.x; i < TH * 3 * (W / 4); i += blockDim.x) {
        const int x4 = i % (W / 4), rc = i / (W / 4), c = rc / TH, r = rc % TH;   // plane-major inside the CTA: TH contiguous rows per plane
        st4(o + (((size_t)z * 3 + c) * H + ty0 + r) * W + x4 * 4, make_float4(v, v, v, v));
    }
}
template <typename F> float timeit(F f) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 3; i++) f();
    float best = 1e9;
    for (int i = 0; i < 20; i++) { cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms; }
    return best;
}
int main() {
    float *o, *o2; size_t bytes = (size_t)N * 3 * H * W * 4;
    cudaMalloc(&o, bytes); cudaMalloc(&o2, bytes);
    auto rep = [&](const char* n, float ms) { printf("%-44s %.4f ms  %.0f GB/s\n", n, ms, bytes / ms / 1e6); };
    rep("linear fill (grid-stride, 148*8 CTAs)", timeit([&] { linear_fill<<<148 * 8, 256>>>((float4*)o, bytes / 16, 1.f); }));
    rep("linear fill (one float4 per thread)", timeit([&] { linear_fill<<<(unsigned)(bytes / 16 / 256), 256>>>((float4*)o, bytes / 16, 1.f); }));
    rep("tile 128x32x3 (st no_allocate)", timeit([&] { tile128<32, true><<<dim3(5, 20, N), 256>>>(o, 1.f); }));
    rep("tile 128x32x3 (plain st)", timeit([&] { tile128<32, false><<<dim3(5, 20, N), 256>>>(o, 1.f); }));
    rep("tile 128x16x3", timeit([&] { tile128<16, true><<<dim3(5, 40, N), 256>>>(o, 1.f); }));
    rep("tile 128x8x3", timeit([&] { tile128<8, true><<<dim3(5, 80, N), 256>>>(o, 1.f); }));
    rep("row tile 640x8x3 (256 thr)", timeit([&] { rowtile<8><<<dim3(80, N), 256>>>(o, 1.f); }));
    rep("row tile 640x16x3 (256 thr)", timeit([&] { rowtile<16><<<dim3(40, N), 256>>>(o, 1.f); }));
    rep("row tile 640x32x3 (512 thr)", timeit([&] { rowtile<32><<<dim3(20, N), 512>>>(o, 1.f); }));
    rep("row tile 640x4x3 (256 thr)", timeit([&] { rowtile<4><<<dim3(160, N), 256>>>(o, 1.f); }));
    cudaError_t e = cudaDeviceSynchronize(); printf("%s\n", cudaGetErrorString(e));
    return 0;
}
