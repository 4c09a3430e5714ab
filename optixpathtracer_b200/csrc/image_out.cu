// image_out.cu — K10: the display transform and the image writers, headless.
//
// Replaces, for the step right after accumulation:
//   * the GL display pass  Renderer/OpenGL/Shader/Final.frag:12-36 — LinearToSRGB8 (the sRGB OETF
//     with glm::step / mix) applied to the running mean; the window's 8-bit target becomes a
//     device buffer that is downloaded or written to a file;
//   * Renderer/Images/WriteImage.cpp:9-33  WriteBMP — clamp to [0,1], (uint8_t)(c * 255.0f)
//     (truncation), rows from the top of the image down;
//   * Renderer/Images/WriteImage.cpp:36-99 WriteEXR — three FLOAT channels named B, G, R, rows from
//     the top down, a pixel with a NaN component written as (0, 0, 0).
// The file containers are written here without third-party code (the reference uses stb_image_write
// and tinyexr): 24-bit BMP, PNG with stored deflate blocks, scan-line OpenEXR without compression.
//
// Row convention of the API buffers (b2pt.h): row 0 = bottom of the image (the reference's
// colorBuffer / glm::vec3 h_pixels[]); every writer below emits the TOP row first like the
// reference's loops `for (y = size.y - 1; y >= 0; --y)`.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <vector>

#include "render.h"

namespace b2pt {

namespace {

// Final.frag:12-17.  step(edge, x) = x < edge ? 0 : 1;  mix(a, b, m) = a*(1-m) + b*m with m in {0,1}.
// pow through double: the shader's pow is a GPU approximation anyway; a rounded double pow is the
// correctly rounded value on all but ~0.1 % of inputs.
__device__ __forceinline__ float linear_to_srgb_1(float c) {
    const float m = c < 0.0031308f ? 0.0f : 1.0f;
    const float lo = 12.92f * c;
    const float hi = 1.055f * (float)pow((double)c, (double)(1.0f / 2.4f)) - 0.055f;
    return m == 0.0f ? lo : hi;   // mix with m in {0,1}; avoids NaN * 0 from pow of a negative value
}

// mode 0: display (OETF, clamp, truncate to 8 bits), 3 bytes per pixel RGB
// top_first: output row 0 = image top
__global__ void __launch_bounds__(256)
k_display(int W, int H, const float* __restrict__ sum, float inv_spp, int apply_oetf, int top_first,
          uint8_t* __restrict__ rgb8) {
    const size_t n = (size_t)W * H;
    for (size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x; p < n; p += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(p / W), x = (int)(p - (size_t)y * W);
        const int yo = top_first ? H - 1 - y : y;
        uint8_t* o = rgb8 + 3 * ((size_t)yo * W + x);
#pragma unroll
        for (int c = 0; c < 3; c++) {
            float v = sum[3 * p + c] * inv_spp;
            if (apply_oetf) v = linear_to_srgb_1(v);
            v = fminf(fmaxf(v, 0.0f), 1.0f);       // glm::clamp; NaN -> 0 through fmaxf
            o[c] = (uint8_t)(v * 255.0f);          // WriteImage.cpp:16-18 (truncation)
        }
    }
}

// B, G, R planes, top row first, NaN pixel -> 0 (WriteImage.cpp:50-63, 66-69)
__global__ void __launch_bounds__(256)
k_exr_planes(int W, int H, const float* __restrict__ sum, float inv_spp, float* __restrict__ planes) {
    const size_t n = (size_t)W * H;
    for (size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x; p < n; p += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(p / W), x = (int)(p - (size_t)y * W);
        const size_t q = (size_t)(H - 1 - y) * W + x;
        float r = sum[3 * p] * inv_spp, g = sum[3 * p + 1] * inv_spp, b = sum[3 * p + 2] * inv_spp;
        if (isnan(r) || isnan(g) || isnan(b)) r = g = b = 0.0f;
        planes[q] = b;
        planes[n + q] = g;
        planes[2 * n + q] = r;
    }
}

float inv_spp_of(const B2ptRenderer* r) { return r->spp_accumulated ? 1.0f / (float)r->spp_accumulated : 0.0f; }

void put32(std::vector<uint8_t>& v, uint32_t x) { for (int i = 0; i < 4; i++) v.push_back((uint8_t)(x >> (8 * i))); }
void put16(std::vector<uint8_t>& v, uint32_t x) { v.push_back((uint8_t)x); v.push_back((uint8_t)(x >> 8)); }
void put32be(std::vector<uint8_t>& v, uint32_t x) { for (int i = 3; i >= 0; i--) v.push_back((uint8_t)(x >> (8 * i))); }
void put64(std::vector<uint8_t>& v, uint64_t x) { for (int i = 0; i < 8; i++) v.push_back((uint8_t)(x >> (8 * i))); }
void puts0(std::vector<uint8_t>& v, const char* s) { while (*s) v.push_back((uint8_t)*s++); v.push_back(0); }

void write_file(const char* path, const std::vector<uint8_t>& bytes) {
    FILE* f = fopen(path, "wb");
    B2PT_REQUIRE(f != nullptr, B2PT_E_INVALID, "cannot open output file");
    const size_t w = fwrite(bytes.data(), 1, bytes.size(), f);
    fclose(f);
    B2PT_REQUIRE(w == bytes.size(), B2PT_E_INVALID, "short write");
}

uint32_t crc32_of(const uint8_t* p, size_t n, uint32_t crc = 0) {
    static uint32_t table[256];
    static bool init = false;
    if (!init) {
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            table[i] = c;
        }
        init = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xff] ^ (crc >> 8);
    return ~crc;
}

void png_chunk(std::vector<uint8_t>& out, const char* type, const std::vector<uint8_t>& data) {
    put32be(out, (uint32_t)data.size());
    const size_t at = out.size();
    out.insert(out.end(), type, type + 4);
    out.insert(out.end(), data.begin(), data.end());
    put32be(out, crc32_of(out.data() + at, 4 + data.size()));
}

}  // namespace

// 8-bit image on the host: rgb8[3 * (row * W + x)], row 0 = top when top_first
void display_to_host(B2ptRenderer* r, bool apply_oetf, bool top_first, uint8_t* h_rgb8) {
    B2PT_REQUIRE(r->width > 0 && r->height > 0, B2PT_E_STATE, "resize first");
    const size_t n = (size_t)r->width * r->height * 3;
    DeviceBuffer<uint8_t> d;
    d.alloc(n);
    k_display<<<r->sm_count * 4, 256, 0, r->stream>>>(r->width, r->height, r->d_accum.get(), inv_spp_of(r),
                                                     apply_oetf ? 1 : 0, top_first ? 1 : 0, d.get());
    r->kernel_launches++;
    B2PT_CUDA(cudaGetLastError());
    d.download(h_rgb8, n, r->stream);
    B2PT_CUDA(cudaStreamSynchronize(r->stream));
}

// stb-style 24-bit BMP: bottom-up rows of B,G,R, each padded to 4 bytes
void write_bmp(B2ptRenderer* r, const char* path, bool apply_oetf) {
    const int W = r->width, H = r->height;
    std::vector<uint8_t> px((size_t)W * H * 3);
    display_to_host(r, apply_oetf, /*top_first=*/false, px.data());   // BMP stores the bottom row first
    const int pad = (4 - (W * 3) % 4) % 4;
    const uint32_t data_bytes = (uint32_t)((W * 3 + pad) * H);
    std::vector<uint8_t> f;
    f.push_back('B'); f.push_back('M');
    put32(f, 14 + 40 + data_bytes); put16(f, 0); put16(f, 0); put32(f, 14 + 40);
    put32(f, 40); put32(f, (uint32_t)W); put32(f, (uint32_t)H); put16(f, 1); put16(f, 24);
    put32(f, 0); put32(f, 0); put32(f, 0); put32(f, 0); put32(f, 0); put32(f, 0);
    for (int y = 0; y < H; y++) {
        for (int x = 0; x < W; x++) {
            const uint8_t* p = &px[3 * ((size_t)y * W + x)];
            f.push_back(p[2]); f.push_back(p[1]); f.push_back(p[0]);
        }
        for (int k = 0; k < pad; k++) f.push_back(0);
    }
    write_file(path, f);
}

// PNG, 8-bit RGB, filter 0 on every row, zlib stream of stored (uncompressed) deflate blocks
void write_png(B2ptRenderer* r, const char* path, bool apply_oetf) {
    const int W = r->width, H = r->height;
    std::vector<uint8_t> px((size_t)W * H * 3);
    display_to_host(r, apply_oetf, /*top_first=*/true, px.data());
    std::vector<uint8_t> raw;
    raw.reserve((size_t)H * (W * 3 + 1));
    for (int y = 0; y < H; y++) {
        raw.push_back(0);
        raw.insert(raw.end(), px.begin() + (size_t)y * W * 3, px.begin() + (size_t)(y + 1) * W * 3);
    }
    std::vector<uint8_t> z;
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;   // adler32
    size_t pos = 0;
    while (pos < raw.size() || raw.empty()) {
        const size_t len = std::min<size_t>(65535, raw.size() - pos);
        z.push_back(pos + len >= raw.size() ? 1 : 0);
        put16(z, (uint32_t)len); put16(z, (uint32_t)(~len & 0xffff));
        for (size_t i = 0; i < len; i++) {
            const uint8_t c = raw[pos + i];
            z.push_back(c);
            a = (a + c) % 65521u;
            b = (b + a) % 65521u;
        }
        pos += len;
        if (raw.empty()) break;
    }
    put32be(z, (b << 16) | a);
    std::vector<uint8_t> f = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::vector<uint8_t> ihdr;
    put32be(ihdr, (uint32_t)W); put32be(ihdr, (uint32_t)H);
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    png_chunk(f, "IHDR", ihdr);
    png_chunk(f, "IDAT", z);
    png_chunk(f, "IEND", {});
    write_file(path, f);
}

// Scan-line OpenEXR 2.0, no compression, channels B G R (FLOAT), increasing-y line order with the
// image's top row as y = 0 — what tinyexr's SaveEXRImageToFile writes for WriteImage.cpp:36-99 minus
// the ZIP compression of the pixel data.
void write_exr(B2ptRenderer* r, const char* path) {
    B2PT_REQUIRE(r->width > 0 && r->height > 0, B2PT_E_STATE, "resize first");
    const int W = r->width, H = r->height;
    const size_t n = (size_t)W * H;
    DeviceBuffer<float> d;
    d.alloc(3 * n);
    k_exr_planes<<<r->sm_count * 4, 256, 0, r->stream>>>(W, H, r->d_accum.get(), inv_spp_of(r), d.get());
    r->kernel_launches++;
    B2PT_CUDA(cudaGetLastError());
    std::vector<float> planes(3 * n);
    d.download(planes.data(), 3 * n, r->stream);
    B2PT_CUDA(cudaStreamSynchronize(r->stream));

    std::vector<uint8_t> f;
    put32(f, 20000630u);   // magic
    put32(f, 2u);          // version 2, scan-line, single part
    auto attr = [&](const char* name, const char* type, const std::vector<uint8_t>& v) {
        puts0(f, name); puts0(f, type); put32(f, (uint32_t)v.size());
        f.insert(f.end(), v.begin(), v.end());
    };
    std::vector<uint8_t> ch;
    for (const char* c : {"B", "G", "R"}) {
        puts0(ch, c); put32(ch, 2u /*FLOAT*/); ch.push_back(0); ch.push_back(0); ch.push_back(0); ch.push_back(0);
        put32(ch, 1u); put32(ch, 1u);
    }
    ch.push_back(0);
    attr("channels", "chlist", ch);
    attr("compression", "compression", {0});
    std::vector<uint8_t> box;
    put32(box, 0); put32(box, 0); put32(box, (uint32_t)(W - 1)); put32(box, (uint32_t)(H - 1));
    attr("dataWindow", "box2i", box);
    attr("displayWindow", "box2i", box);
    attr("lineOrder", "lineOrder", {0});
    std::vector<uint8_t> one; { float v = 1.0f; uint32_t u; memcpy(&u, &v, 4); put32(one, u); }
    attr("pixelAspectRatio", "float", one);
    std::vector<uint8_t> v2; put32(v2, 0); put32(v2, 0);
    attr("screenWindowCenter", "v2f", v2);
    attr("screenWindowWidth", "float", one);
    f.push_back(0);   // end of header
    const size_t line_bytes = (size_t)W * 4 * 3;
    const uint64_t table_at = f.size();
    uint64_t off = table_at + (uint64_t)H * 8;
    for (int y = 0; y < H; y++) { put64(f, off); off += 8 + line_bytes; }
    for (int y = 0; y < H; y++) {
        put32(f, (uint32_t)y); put32(f, (uint32_t)line_bytes);
        for (int c = 0; c < 3; c++) {   // channels in alphabetical order: B, G, R = plane order
            const uint8_t* src = reinterpret_cast<const uint8_t*>(&planes[c * n + (size_t)y * W]);
            f.insert(f.end(), src, src + (size_t)W * 4);
        }
    }
    write_file(path, f);
}

}  // namespace b2pt
