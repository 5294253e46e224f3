// Shared host/device helpers for the nerfw_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/nerfw_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "nerfw_b200 is written for sm_100a (B200) only"
#endif

namespace nwb {

constexpr int kNumSMs = 148;

// ---- error plumbing: thread-local text, never throws / exits ---------------------------------
char* err_buf();
void note_launch();             // api.cu: counts kernel launches (nwb_kernel_launches)
int set_error(int code, const char* fmt, ...);

#define NWB_REQUIRE(cond, ...)                                   \
    do {                                                         \
        if (!(cond)) return nwb::set_error(NWB_E_ARG, __VA_ARGS__); \
    } while (0)

#define NWB_CUDA(call)                                                                    \
    do {                                                                                  \
        cudaError_t e__ = (call);                                                         \
        if (e__ != cudaSuccess)                                                           \
            return nwb::set_error(NWB_E_CUDA, "%s failed: %s (%s:%d)", #call,             \
                                  cudaGetErrorString(e__), __FILE__, __LINE__);           \
    } while (0)

#define NWB_LAUNCH_CHECK()                                                                \
    do {                                                                                  \
        nwb::note_launch();                                                               \
        cudaError_t e__ = cudaGetLastError();                                             \
        if (e__ != cudaSuccess)                                                           \
            return nwb::set_error(NWB_E_CUDA, "kernel launch failed: %s (%s:%d)",         \
                                  cudaGetErrorString(e__), __FILE__, __LINE__);           \
    } while (0)

// ---- per-device one-time kernel configuration --------------------------------------------------
// cudaFuncSetAttribute is a per-device setting: every launcher keeps one flag per device ordinal (a process may
// drive several GPUs from several host threads; the flag write is idempotent) and checks the SM count the
// persistent grids were sized for.
constexpr int kMaxDevices = 64;
int current_device_checked(int* dev_out);       // cudaGetDevice + "148 SMs" check, cached per device (api.cu)

#define NWB_CONFIGURE_ONCE(flags, ...)                                          \
    do {                                                                        \
        int dev__ = 0;                                                          \
        int rc__ = nwb::current_device_checked(&dev__);                         \
        if (rc__ != NWB_OK) return rc__;                                        \
        if (!(flags)[dev__]) { __VA_ARGS__; (flags)[dev__] = true; }            \
    } while (0)

// ---- torch.linspace(0,1,n)[i] exactly as ATen computes it in fp32 ----------------------------
// step = (end-start)/(n-1); lower half counts up from start, upper half counts down from end
// with a fused multiply-add (verified bit-exact against torch.linspace for n = 2..1024).
__host__ __device__ __forceinline__ float linspace01(int i, int n) {
    if (n == 1) return 0.0f;
    const float step = 1.0f / static_cast<float>(n - 1);
#ifdef __CUDA_ARCH__
    return (i < n / 2) ? __fmul_rn(step, static_cast<float>(i))
                       : __fmaf_rn(-step, static_cast<float>(n - 1 - i), 1.0f);
#else
    return (i < n / 2) ? step * static_cast<float>(i) : fmaf(-step, static_cast<float>(n - 1 - i), 1.0f);
#endif
}

// ---- Philox4x32-10 counter RNG (timed runs; parity runs pass explicit tensors) ---------------
// `draw` names the draw site (0 = coarse stratified jitter, 1 = inverse-CDF uniforms): it occupies a counter word of
// its own, so two sites keyed by the same seed never share a counter block.
__device__ __forceinline__ uint4 philox4x32(uint64_t seed, uint64_t ctr, uint32_t draw) {
    uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
    uint32_t c0 = static_cast<uint32_t>(ctr), c1 = static_cast<uint32_t>(ctr >> 32), c2 = draw,
             c3 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        c0 = hi1 ^ c1 ^ k0;
        c1 = lo1;
        c2 = hi0 ^ c3 ^ k1;
        c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}
__device__ __forceinline__ float u01(uint32_t x) {   // [0,1) with 24 random bits
    return static_cast<float>(x >> 8) * (1.0f / 16777216.0f);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace nwb
