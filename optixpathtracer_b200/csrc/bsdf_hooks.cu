// bsdf_hooks.cu — b2pt_bsdf_f / b2pt_bsdf_sample / b2pt_bsdf_pdf: the device
// BSDF functions of bsdf.cuh (the code the K7 shade kernels run) on caller
// rows, one thread per row, for parity against the reference's PBRT/*.h.
#include <vector>

#include "bsdf_rows.cuh"
#include "host_util.h"
#include "wave_common.cuh"

using namespace b2pt;

namespace {

template <class Rng>
__global__ void k_bsdf_f(int bsdf, int64_t n, const float* albedo, const float* roughness, const float* wo,
                         const float* wi, const uint32_t* seed, float* f_out, uint32_t* seed_out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) bsdf_row_f<Rng>(bsdf, i, albedo, roughness, wo, wi, seed, f_out, seed_out);
}
template <class Rng>
__global__ void k_bsdf_sample(int bsdf, int64_t n, const float* albedo, const float* roughness, const float* wo,
                              const uint32_t* seed, uint8_t* ok, float* f_out, float* wi_out, float* pdf_out,
                              uint8_t* flags_out, uint32_t* seed_out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) bsdf_row_sample<Rng>(bsdf, i, albedo, roughness, wo, seed, ok, f_out, wi_out, pdf_out, flags_out, seed_out);
}
__global__ void k_bsdf_pdf(int bsdf, int64_t n, const float* roughness, const float* wo, const float* wi,
                           int reflection, int transmission, float* pdf_out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) bsdf_row_pdf(bsdf, i, roughness, wo, wi, reflection != 0, transmission != 0, pdf_out);
}

thread_local std::string g_hook_error;

template <typename F>
int hook_guard(int device, F&& f) {
    try {
        B2PT_CUDA(cudaSetDevice(device));
        f();
        B2PT_CUDA(cudaGetLastError());
        B2PT_CUDA(cudaDeviceSynchronize());
        return B2PT_OK;
    } catch (const Error& e) {
        g_hook_error = e.msg;
        return e.code;
    }
}

bool bad_bsdf(int b) { return b < B2PT_BSDF_CONDUCTOR || b > B2PT_BSDF_LAYERED; }
// rng PHILOX = the production instantiation: for the layered BSDF the walk engine's per-lane logic
// (fast_layered.cuh), for the others the same bsdf.cuh functions drawing from PhiloxRng as K7 does
bool bad_lcg(int m, int /*bsdf*/) {
    return m != B2PT_RNG_PHILOX && m != B2PT_RNG_LCG_DEVICE && m != B2PT_RNG_LCG_HOST;
}
int blocks(int64_t n) { return (int)((n + 127) / 128); }

}  // namespace

extern "C" {

int b2pt_bsdf_f(int32_t device, int32_t bsdf, int32_t lcg_mode, int64_t n, const float* albedo,
                const float* roughness, const float* wo, const float* wi, const uint32_t* seed, float* f_out,
                uint32_t* seed_out) {
    if (n < 0 || bad_bsdf(bsdf) || bad_lcg(lcg_mode, bsdf) || (n > 0 && (!albedo || !roughness || !wo || !wi || !f_out)))
        return B2PT_E_INVALID;
    if (n == 0) return B2PT_OK;
    return hook_guard(device, [&] {
        DeviceBuffer<float> d_a, d_r, d_wo, d_wi, d_f;
        DeviceBuffer<uint32_t> d_s, d_so;
        std::vector<uint32_t> zeros;
        d_a.upload(albedo, (size_t)n * 3); d_r.upload(roughness, (size_t)n);
        d_wo.upload(wo, (size_t)n * 3); d_wi.upload(wi, (size_t)n * 3);
        if (!seed) { zeros.assign((size_t)n, 0u); seed = zeros.data(); }
        d_s.upload(seed, (size_t)n);
        d_f.alloc((size_t)n * 3); d_so.alloc((size_t)n);
        d_so.zero();
        if (lcg_mode == B2PT_RNG_PHILOX && bsdf == B2PT_BSDF_LAYERED)
            launch_fast_layered_rows_f(n, d_a.get(), d_r.get(), d_wo.get(), d_wi.get(), d_s.get(), d_f.get());
        else if (lcg_mode == B2PT_RNG_PHILOX)
            k_bsdf_f<PhiloxRng><<<blocks(n), 128>>>(bsdf, n, d_a.get(), d_r.get(), d_wo.get(), d_wi.get(), d_s.get(), d_f.get(), d_so.get());
        else if (lcg_mode == B2PT_RNG_LCG_HOST)
            k_bsdf_f<LcgRng<true>><<<blocks(n), 128>>>(bsdf, n, d_a.get(), d_r.get(), d_wo.get(), d_wi.get(), d_s.get(), d_f.get(), d_so.get());
        else
            k_bsdf_f<LcgRng<false>><<<blocks(n), 128>>>(bsdf, n, d_a.get(), d_r.get(), d_wo.get(), d_wi.get(), d_s.get(), d_f.get(), d_so.get());
        B2PT_CUDA(cudaDeviceSynchronize());
        d_f.download(f_out, (size_t)n * 3);
        if (seed_out) d_so.download(seed_out, (size_t)n);
    });
}

int b2pt_bsdf_sample(int32_t device, int32_t bsdf, int32_t lcg_mode, int64_t n, const float* albedo,
                     const float* roughness, const float* wo, const uint32_t* seed, uint8_t* ok, float* f_out,
                     float* wi_out, float* pdf_out, uint8_t* flags_out, uint32_t* seed_out) {
    if (n < 0 || bad_bsdf(bsdf) || bad_lcg(lcg_mode, bsdf) ||
        (n > 0 && (!albedo || !roughness || !wo || !seed || !ok || !f_out || !wi_out || !pdf_out || !flags_out)))
        return B2PT_E_INVALID;
    if (n == 0) return B2PT_OK;
    return hook_guard(device, [&] {
        DeviceBuffer<float> d_a, d_r, d_wo, d_f, d_wi, d_p;
        DeviceBuffer<uint32_t> d_s, d_so;
        DeviceBuffer<uint8_t> d_ok, d_fl;
        d_a.upload(albedo, (size_t)n * 3); d_r.upload(roughness, (size_t)n); d_wo.upload(wo, (size_t)n * 3);
        d_s.upload(seed, (size_t)n);
        d_f.alloc((size_t)n * 3); d_wi.alloc((size_t)n * 3); d_p.alloc((size_t)n);
        d_so.alloc((size_t)n); d_ok.alloc((size_t)n); d_fl.alloc((size_t)n);
        d_so.zero();
        if (lcg_mode == B2PT_RNG_PHILOX && bsdf == B2PT_BSDF_LAYERED)
            launch_fast_layered_rows_sample(n, d_a.get(), d_r.get(), d_wo.get(), d_s.get(), d_ok.get(), d_f.get(), d_wi.get(),
                                            d_p.get(), d_fl.get());
        else if (lcg_mode == B2PT_RNG_PHILOX)
            k_bsdf_sample<PhiloxRng><<<blocks(n), 128>>>(bsdf, n, d_a.get(), d_r.get(), d_wo.get(), d_s.get(), d_ok.get(), d_f.get(), d_wi.get(), d_p.get(), d_fl.get(), d_so.get());
        else if (lcg_mode == B2PT_RNG_LCG_HOST)
            k_bsdf_sample<LcgRng<true>><<<blocks(n), 128>>>(bsdf, n, d_a.get(), d_r.get(), d_wo.get(), d_s.get(), d_ok.get(), d_f.get(), d_wi.get(), d_p.get(), d_fl.get(), d_so.get());
        else
            k_bsdf_sample<LcgRng<false>><<<blocks(n), 128>>>(bsdf, n, d_a.get(), d_r.get(), d_wo.get(), d_s.get(), d_ok.get(), d_f.get(), d_wi.get(), d_p.get(), d_fl.get(), d_so.get());
        B2PT_CUDA(cudaDeviceSynchronize());
        d_ok.download(ok, (size_t)n); d_f.download(f_out, (size_t)n * 3); d_wi.download(wi_out, (size_t)n * 3);
        d_p.download(pdf_out, (size_t)n); d_fl.download(flags_out, (size_t)n);
        if (seed_out) d_so.download(seed_out, (size_t)n);
    });
}

int b2pt_bsdf_pdf(int32_t device, int32_t bsdf, int64_t n, const float* roughness, const float* wo, const float* wi,
                  int32_t reflection, int32_t transmission, float* pdf_out) {
    if (n < 0 || bad_bsdf(bsdf) || (n > 0 && (!roughness || !wo || !wi || !pdf_out))) return B2PT_E_INVALID;
    if (n == 0) return B2PT_OK;
    return hook_guard(device, [&] {
        DeviceBuffer<float> d_r, d_wo, d_wi, d_p;
        d_r.upload(roughness, (size_t)n); d_wo.upload(wo, (size_t)n * 3); d_wi.upload(wi, (size_t)n * 3);
        d_p.alloc((size_t)n);
        k_bsdf_pdf<<<blocks(n), 128>>>(bsdf, n, d_r.get(), d_wo.get(), d_wi.get(), reflection, transmission, d_p.get());
        B2PT_CUDA(cudaDeviceSynchronize());
        d_p.download(pdf_out, (size_t)n);
    });
}

}  // extern "C"
