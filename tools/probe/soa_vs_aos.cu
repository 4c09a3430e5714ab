// Probe (not part of the product): what does the path-state layout cost?  One thread per path gathers NR words and
// scatters NW words, (a) as 4-byte loads from NR + NW structure-of-arrays planes (the layout of the wave state),
// (b) as 16-byte loads from (NR + NW) / 4 planes of float4 records.  Same bytes; prints GB/s of each.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/probe/soa_vs_aos tools/probe/soa_vs_aos.cu
#include <cstdio>
#include <cuda_runtime.h>
constexpr int NR = 40, NW = 16;
__global__ void __launch_bounds__(128) k_soa(const float* __restrict__ in, float* __restrict__ out, size_t cap, unsigned n, int dense) {
    for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x) {
        const unsigned i = dense ? q : (q * 2u) % n + (q * 2u >= n ? 1u : 0u);
        float s[NW];
#pragma unroll
        for (int w = 0; w < NW; w++) s[w] = 0.f;
#pragma unroll
        for (int r = 0; r < NR; r++) s[r % NW] += in[(size_t)r * cap + i];
#pragma unroll
        for (int w = 0; w < NW; w++) out[(size_t)w * cap + q] = s[w];
    }
}
__global__ void __launch_bounds__(128) k_aos(const float4* __restrict__ in, float4* __restrict__ out, size_t cap, unsigned n, int dense) {
    for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < n; q += gridDim.x * blockDim.x) {
        const unsigned i = dense ? q : (q * 2u) % n + (q * 2u >= n ? 1u : 0u);
        float4 s[NW / 4];
#pragma unroll
        for (int w = 0; w < NW / 4; w++) s[w] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < NR / 4; r++) {
            const float4 v = in[(size_t)r * cap + i];
            s[r % (NW / 4)].x += v.x; s[r % (NW / 4)].y += v.y; s[r % (NW / 4)].z += v.z; s[r % (NW / 4)].w += v.w;
        }
#pragma unroll
        for (int w = 0; w < NW / 4; w++) out[(size_t)w * cap + q] = s[w];
    }
}
int main() {
    const size_t cap = 1u << 25;
    const unsigned n = 30000000u;
    float *in, *out;
    cudaMalloc(&in, cap * NR * 4); cudaMalloc(&out, cap * NW * 4);
    cudaMemset(in, 0, cap * NR * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int blocks = 8; blocks <= 16; blocks += 8)
    for (int dense = 1; dense >= 0; dense--)
    for (int aos = 0; aos < 2; aos++) {
        float best = 1e9f;
        for (int rep = 0; rep < 5; rep++) {
            cudaEventRecord(e0);
            if (aos) k_aos<<<148 * blocks, 128>>>((const float4*)in, (float4*)out, cap, n, dense);
            else k_soa<<<148 * blocks, 128>>>(in, out, cap, n, dense);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        printf("%s %s blocks/SM %2d: %.3f ms  %.0f GB/s\n", aos ? "float4 records" : "4-byte planes ", dense ? "dense  " : "strided", blocks, best,
               (double)n * (NR + NW) * 4 / best / 1e6);
    }
    return 0;
}
