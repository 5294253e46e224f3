// Adam on the flat parameter / gradient buffers (reference: utils/__init__.py:24-33 builds
// torch.optim.Adam(lr, eps=1e-8, weight_decay) over every trainable tensor; train.py:139-142).
// One vectorised elementwise pass over the 1.3 M parameters: reads p, g, m, v, writes p, m, v and
// (optionally) clears g for the next step's atomic accumulation -- HBM-bound, 7 floats per element.
#include "common.cuh"
#include <algorithm>
#include <math.h>

namespace nwb {

__global__ void __launch_bounds__(256)
adam_kernel(float4* __restrict__ p, float4* __restrict__ g, float4* __restrict__ m, float4* __restrict__ v, int64_t n4,
            float* __restrict__ pt, float* __restrict__ gt, float* __restrict__ mt, float* __restrict__ vt, int tail,
            float step_size, float beta1, float beta2, float inv_sqrt_bc2, float eps, float wd, float gscale,
            int zero_grad, const float* __restrict__ hyper) {
    if (hyper != nullptr) {               // device-resident step state (CUDA-graph replays): {step, lr, lr/bc1, 1/sqrt(bc2)}
        step_size = hyper[2];
        inv_sqrt_bc2 = hyper[3];
    }
    auto upd = [&](float& pp, float gg, float& mm, float& vv) {
        gg = gg * gscale + wd * pp;
        mm = beta1 * mm + (1.0f - beta1) * gg;
        vv = beta2 * vv + (1.0f - beta2) * gg * gg;
        pp -= step_size * mm / (sqrtf(vv) * inv_sqrt_bc2 + eps);
    };
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n4;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        float4 pp = p[i], gg = g[i], mm = m[i], vv = v[i];
        upd(pp.x, gg.x, mm.x, vv.x); upd(pp.y, gg.y, mm.y, vv.y);
        upd(pp.z, gg.z, mm.z, vv.z); upd(pp.w, gg.w, mm.w, vv.w);
        p[i] = pp; m[i] = mm; v[i] = vv;
        if (zero_grad) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (blockIdx.x == 0 && static_cast<int>(threadIdx.x) < tail) {
        const int i = threadIdx.x;
        float pp = pt[i], mm = mt[i], vv = vt[i];
        upd(pp, gt[i], mm, vv);
        pt[i] = pp; mt[i] = mm; vt[i] = vv;
        if (zero_grad) gt[i] = 0.f;
    }
}

// advances the device-resident step state: state = {int step; float lr; float lr/bc1; float 1/sqrt(bc2)}
__global__ void adam_tick_kernel(float* state, float beta1, float beta2) {
    int* step = reinterpret_cast<int*>(state);
    const int t = *step + 1;
    *step = t;
    const double bc1 = 1.0 - pow(static_cast<double>(beta1), t), bc2 = 1.0 - pow(static_cast<double>(beta2), t);
    state[2] = static_cast<float>(static_cast<double>(state[1]) / bc1);
    state[3] = static_cast<float>(1.0 / sqrt(bc2));
}

}  // namespace nwb

using namespace nwb;

extern "C" int nwb_adam_step(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                             float eps, float weight_decay, int step, float grad_scale, int zero_grad, void* stream) {
    NWB_REQUIRE(n >= 0 && step >= 1, "nwb_adam_step: bad size / step");
    if (n == 0) return NWB_OK;
    NWB_REQUIRE(p && g && m && v, "nwb_adam_step: null pointer");
    NWB_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                  reinterpret_cast<uintptr_t>(v)) & 15u) == 0, "nwb_adam_step: buffers must be 16 B aligned");
    const double bc1 = 1.0 - pow(static_cast<double>(beta1), step), bc2 = 1.0 - pow(static_cast<double>(beta2), step);
    const int64_t n4 = n / 4;
    const int tail = static_cast<int>(n - n4 * 4);
    const int blocks = static_cast<int>(std::min<int64_t>((n4 + 255) / 256 + 1, kNumSMs * 8));
    adam_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<float4*>(p), reinterpret_cast<float4*>(g), reinterpret_cast<float4*>(m),
        reinterpret_cast<float4*>(v), n4, p + n4 * 4, g + n4 * 4, m + n4 * 4, v + n4 * 4, tail,
        static_cast<float>(lr / bc1), beta1, beta2, static_cast<float>(1.0 / sqrt(bc2)), eps, weight_decay, grad_scale,
        zero_grad, nullptr);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

extern "C" int nwb_adam_step_dev(float* p, float* g, float* m, float* v, int64_t n, float beta1, float beta2, float eps,
                                 float weight_decay, float* state, float grad_scale, int zero_grad, void* stream) {
    NWB_REQUIRE(n >= 0, "nwb_adam_step_dev: bad size");
    if (n == 0) return NWB_OK;
    NWB_REQUIRE(p && g && m && v && state, "nwb_adam_step_dev: null pointer");
    NWB_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                  reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(state)) & 15u) == 0,
                "nwb_adam_step_dev: buffers must be 16 B aligned");
    const int64_t n4 = n / 4;
    const int tail = static_cast<int>(n - n4 * 4);
    const int blocks = static_cast<int>(std::min<int64_t>((n4 + 255) / 256 + 1, kNumSMs * 8));
    adam_tick_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(state, beta1, beta2);
    adam_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<float4*>(p), reinterpret_cast<float4*>(g), reinterpret_cast<float4*>(m),
        reinterpret_cast<float4*>(v), n4, p + n4 * 4, g + n4 * 4, m + n4 * 4, v + n4 * 4, tail,
        0.0f, beta1, beta2, 1.0f, eps, weight_decay, grad_scale, zero_grad, state);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
