// texprobe.cu — diagnostic: what does tex2D<float4> return for the reference's texture
// descriptor (OptixRenderer.cpp:592-603: wrap, linear, normalised float, normalised coords)?
// Writes coordinates and results so that oracle/ref_harness.cpp's software filter can be
// pinned to the texture unit.  Not part of the product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o texprobe texprobe.cu && ./texprobe out.bin
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>

__global__ void k(cudaTextureObject_t t, const float2* xy, float4* out, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = tex2D<float4>(t, xy[i].x, xy[i].y);
}

int main(int argc, char** argv) {
    const int W = 8, H = 4, N = 1 << 20;
    std::vector<uint8_t> px(W * H * 4);
    uint32_t s = 12345;
    for (auto& p : px) { s = s * 1664525u + 1013904223u; p = (uint8_t)(s >> 24); }
    cudaArray_t arr;
    cudaChannelFormatDesc cd = cudaCreateChannelDesc<uchar4>();
    cudaMallocArray(&arr, &cd, W, H);
    cudaMemcpy2DToArray(arr, 0, 0, px.data(), W * 4, W * 4, H, cudaMemcpyHostToDevice);
    cudaResourceDesc rd = {};
    rd.resType = cudaResourceTypeArray;
    rd.res.array.array = arr;
    cudaTextureDesc td = {};
    td.addressMode[0] = td.addressMode[1] = cudaAddressModeWrap;
    td.filterMode = cudaFilterModeLinear;
    td.readMode = cudaReadModeNormalizedFloat;
    td.normalizedCoords = 1;
    td.maxAnisotropy = 1;
    td.maxMipmapLevelClamp = 99;
    td.minMipmapLevelClamp = 0;
    td.mipmapFilterMode = cudaFilterModePoint;
    td.sRGB = 0;
    cudaTextureObject_t tex;
    cudaCreateTextureObject(&tex, &rd, &td, nullptr);
    std::vector<float2> xy(N);
    for (int i = 0; i < N; i++) {
        if (i < N / 4) {   // fine sweep along x across one texel boundary pair, y at a texel centre
            xy[i].x = (1.5f + 2.0f * (float)i / (float)(N / 4)) / W;
            xy[i].y = 1.5f / H;
        } else if (i < N / 2) {   // fine sweep along y
            xy[i].x = 2.5f / W;
            xy[i].y = (0.5f + 2.0f * (float)(i - N / 4) / (float)(N / 4)) / H;
        } else {   // random, including wrap on both sides
            s = s * 1664525u + 1013904223u; xy[i].x = (float)(s >> 8) / 16777216.0f * 3.0f - 1.0f;
            s = s * 1664525u + 1013904223u; xy[i].y = (float)(s >> 8) / 16777216.0f * 3.0f - 1.0f;
        }
    }
    float2* dxy; float4* dout;
    cudaMalloc(&dxy, N * sizeof(float2));
    cudaMalloc(&dout, N * sizeof(float4));
    cudaMemcpy(dxy, xy.data(), N * sizeof(float2), cudaMemcpyHostToDevice);
    k<<<(N + 255) / 256, 256>>>(tex, dxy, dout, N);
    std::vector<float4> out(N);
    cudaError_t e = cudaMemcpy(out.data(), dout, N * sizeof(float4), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { fprintf(stderr, "cuda: %s\n", cudaGetErrorString(e)); return 1; }
    FILE* f = fopen(argc > 1 ? argv[1] : "texprobe.bin", "wb");
    int hdr[4] = {W, H, N, 0};
    fwrite(hdr, sizeof(int), 4, f);
    fwrite(px.data(), 1, px.size(), f);
    fwrite(xy.data(), sizeof(float2), N, f);
    fwrite(out.data(), sizeof(float4), N, f);
    fclose(f);
    printf("wrote %d samples\n", N);
    return 0;
}
