// Shared helpers for the stofnet_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

#include "../../include/stofnet_b200.h"

namespace stof {

// thread-local error text behind stof_last_error()
void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what, const char *file, int line);

extern std::atomic<unsigned long long> g_launches;
inline void count_launch(unsigned long long n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define STOF_CUDA(call)                                                         \
    do {                                                                        \
        cudaError_t e__ = (call);                                               \
        if (e__ != cudaSuccess) return ::stof::cuda_fail(e__, #call, __FILE__, __LINE__); \
    } while (0)

#define STOF_LAUNCH_CHECK()                                                     \
    do {                                                                        \
        ::stof::count_launch();                                                 \
        cudaError_t e__ = cudaGetLastError();                                   \
        if (e__ != cudaSuccess) return ::stof::cuda_fail(e__, "kernel launch", __FILE__, __LINE__); \
    } while (0)

#define STOF_REQUIRE(cond, code, ...)                                           \
    do {                                                                        \
        if (!(cond)) { ::stof::set_error(__VA_ARGS__); return (code); }         \
    } while (0)

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per device: done once per (kernel, device) - a process may hold one
// model handle per GPU
struct SmemAttrOnce {
    std::atomic<unsigned long long> done{0};
};
template <typename K>
inline int ensure_dyn_smem(K kernel, int bytes, SmemAttrOnce &once) {
    int dev = 0;
    STOF_CUDA(cudaGetDevice(&dev));
    const unsigned long long bit = 1ull << (dev & 63);
    if (!(once.done.load(std::memory_order_acquire) & bit)) {
        STOF_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        once.done.fetch_or(bit, std::memory_order_release);
    }
    return STOF_OK;
}

// RAII: make `dev` the calling thread's current device for the scope and restore the caller's device afterwards (a process
// may hold models on several GPUs; the library must not leak its device into the caller's thread)
struct DeviceGuard {
    int prev = -1;
    bool ok = false;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
        if (prev == dev) { ok = true; prev = -1; return; }
        ok = cudaSetDevice(dev) == cudaSuccess;
        if (!ok) cudaGetLastError();
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceGuard(const DeviceGuard &) = delete;
    DeviceGuard &operator=(const DeviceGuard &) = delete;
};

constexpr int kSMs = 148;           // B200: 2 dies x 74 SMs
constexpr float kLeakySlope = 0.01f;  // nn.LeakyReLU() / F.leaky_relu default (models/stofnet.py:56,89,95)

__device__ __forceinline__ float leaky(float v) { return v > 0.0f ? v : v * kLeakySlope; }

// Activation layout used between our kernels for F=64-channel tensors ("C4L4"):
//   act[n][c/4][l][c%4]  - one 16-byte vector holds 4 channels of one position; positions are
//   contiguous, so a one-sample tap shift is a 16-byte address shift (what the UMMA descriptors of
//   the tensor-core conv need) and warps that walk along l issue fully coalesced 128-bit accesses.
__host__ __device__ __forceinline__ size_t c4l4_index(int c, int l, int L) {
    return ((size_t)(c >> 2) * L + l) * 4 + (c & 3);
}

}  // namespace stof
