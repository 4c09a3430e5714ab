// host_util.h — host-side helpers: error plumbing and an owning device buffer.
//
// The reference wraps cudaMalloc/cudaMemcpy in the non-owning `CUDABuffer`
// (3rdParty/OptixSample/CUDABuffer.h:26-82) and reports errors by throwing
// (CUDA_CHECK, optix7.h:26-36) or exit(2) (OPTIX_CHECK / CUDA_SYNC_CHECK,
// :43-62).  Behind a C-ABI neither is acceptable: every failure becomes a
// B2PT_E_* code plus a message the caller can fetch with b2pt_last_error.
#pragma once

#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <utility>

#include "../../include/b2pt.h"

namespace b2pt {

struct Error {
    int code;
    std::string msg;
};

// Thrown only inside the library and always caught at the ABI boundary.
#define B2PT_CUDA(call)                                                                     \
    do {                                                                                    \
        cudaError_t e_ = (call);                                                            \
        if (e_ != cudaSuccess) {                                                            \
            char buf_[512];                                                                 \
            snprintf(buf_, sizeof buf_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                     __FILE__, __LINE__);                                                   \
            throw ::b2pt::Error{e_ == cudaErrorMemoryAllocation ? B2PT_E_NOMEM : B2PT_E_CUDA, buf_}; \
        }                                                                                   \
    } while (0)

#define B2PT_REQUIRE(cond, code, message)                   \
    do {                                                    \
        if (!(cond)) throw ::b2pt::Error{(code), (message)}; \
    } while (0)

template <typename T>
class DeviceBuffer {
  public:
    DeviceBuffer() = default;
    DeviceBuffer(const DeviceBuffer&) = delete;
    DeviceBuffer& operator=(const DeviceBuffer&) = delete;
    DeviceBuffer(DeviceBuffer&& o) noexcept : p_(o.p_), n_(o.n_) { o.p_ = nullptr; o.n_ = 0; }
    DeviceBuffer& operator=(DeviceBuffer&& o) noexcept {
        if (this != &o) { release(); p_ = o.p_; n_ = o.n_; o.p_ = nullptr; o.n_ = 0; }
        return *this;
    }
    ~DeviceBuffer() { release(); }

    void alloc(size_t n) {
        if (n == n_ && p_) return;
        release();
        if (n == 0) return;
        B2PT_CUDA(cudaMalloc(&p_, n * sizeof(T)));
        n_ = n;
    }
    void release() {
        if (p_) cudaFree(p_);
        p_ = nullptr;
        n_ = 0;
    }
    void upload(const T* h, size_t n, cudaStream_t s = 0) {
        alloc(n);
        if (n) B2PT_CUDA(cudaMemcpyAsync(p_, h, n * sizeof(T), cudaMemcpyHostToDevice, s));
    }
    void download(T* h, size_t n, cudaStream_t s = 0) const {
        if (n) B2PT_CUDA(cudaMemcpyAsync(h, p_, n * sizeof(T), cudaMemcpyDeviceToHost, s));
    }
    void zero(cudaStream_t s = 0) {
        if (n_) B2PT_CUDA(cudaMemsetAsync(p_, 0, n_ * sizeof(T), s));
    }
    T* get() const { return static_cast<T*>(p_); }
    size_t size() const { return n_; }
    size_t bytes() const { return n_ * sizeof(T); }

  private:
    void* p_ = nullptr;
    size_t n_ = 0;
};

}  // namespace b2pt
