// (f1) NerfWLoss -- losses.py:19-50 (NeRF-W eq. 13) -- forward and backward as ONE kernel each.
// Written with torch ops the objective is ~40 launches of 2-3 us over (N,3) / (N,S) tensors per training
// iteration (10 % of a 1024-ray step); here the four terms come out of one grid-wide deterministic reduction
// and the four cotangents (rgb_coarse, rgb_fine, beta, transient_sigmas) out of one elementwise pass.
//   c_l = coef * 0.5 * mean((rgb_coarse - t)^2)                         losses.py:37
//   f_l = coef * mean((rgb_fine - t)^2 / (2 beta^2))                    losses.py:42-43
//   b_l = coef * (3 + mean(log beta))                                   losses.py:44
//   s_l = coef * lambda_u * mean(transient_sigmas)                      losses.py:45
// Without beta (no transient head) f_l = coef * 0.5 * mean((rgb_fine - t)^2) (:40) and b_l, s_l are absent;
// without rgb_fine only c_l exists.  terms[4] = their sum.
#include "objective.cuh"
#include <algorithm>

namespace nwb {

constexpr int kObjThreads = 256;
constexpr int kObjMaxBlocks = 148;
static_assert(kObjMaxBlocks <= kObjThreads, "the last block reduces one partial per thread");

struct ObjArgs {
    const float* rgb_c; const float* rgb_f; const float* beta; const float* ts; const float* tgt;
    int n; long long n_ts;
    float lambda_u, coef;
};

__device__ __forceinline__ float block_sum(float v, float* red) {
    v = warp_sum(v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float s = 0.0f;
    if (warp == 0) {
        s = (lane < kObjThreads / 32) ? red[lane] : 0.0f;
        s = warp_sum(s);
    }
    return s;                                   // valid in warp 0
}

// scratch: [kObjMaxBlocks][4] partial sums followed by one uint32 ticket; the last block to finish adds the
// partials in a fixed tree order (deterministic) and writes terms[0..4]
__global__ void __launch_bounds__(kObjThreads)
nerfw_loss_fwd_kernel(const ObjArgs a, float* __restrict__ scratch, float* __restrict__ terms) {
    __shared__ float red[kObjThreads / 32];
    __shared__ bool is_last;
    const long long tid = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
    const long long n3 = 3ll * a.n;
    float s_c = 0.0f, s_f = 0.0f, s_b = 0.0f, s_s = 0.0f;
    for (long long i = tid; i < n3; i += nth) {
        const float t = __ldg(a.tgt + i);
        const float dc = __ldg(a.rgb_c + i) - t;
        s_c += dc * dc;
        if (a.rgb_f != nullptr) {
            const float df = __ldg(a.rgb_f + i) - t;
            if (a.beta != nullptr) {
                const float b = __ldg(a.beta + i / 3);
                s_f += df * df / (2.0f * b * b);
            } else {
                s_f += df * df;
            }
        }
    }
    if (a.beta != nullptr) {
        for (long long i = tid; i < a.n; i += nth) s_b += logf(__ldg(a.beta + i));
        if (a.ts != nullptr) {
            // the one large input (n x S): float4 loads, two independent accumulators per thread
            long long done = 0;
            if ((reinterpret_cast<uintptr_t>(a.ts) & 15u) == 0) {
                const float4* t4 = reinterpret_cast<const float4*>(a.ts);
                const long long n4 = a.n_ts >> 2;
                float s1 = 0.0f;
                long long i = tid;
                for (; i + nth < n4; i += 2 * nth) {
                    const float4 u = __ldg(t4 + i), w = __ldg(t4 + i + nth);
                    s_s += (u.x + u.y) + (u.z + u.w);
                    s1 += (w.x + w.y) + (w.z + w.w);
                }
                if (i < n4) { const float4 u = __ldg(t4 + i); s_s += (u.x + u.y) + (u.z + u.w); }
                s_s += s1;
                done = n4 << 2;
            }
            for (long long i = done + tid; i < a.n_ts; i += nth) s_s += __ldg(a.ts + i);
        }
    }
    const float r0 = block_sum(s_c, red), r1 = block_sum(s_f, red), r2 = block_sum(s_b, red), r3 = block_sum(s_s, red);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + 4 * kObjMaxBlocks);   // fixed slot behind the partials
    if (threadIdx.x == 0) {
        float* p = scratch + 4 * blockIdx.x;
        p[0] = r0; p[1] = r1; p[2] = r2; p[3] = r3;
        __threadfence();
        is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {                               // whole block: partials of block b go to thread b, fixed tree order
        __threadfence();
        float q0 = 0, q1 = 0, q2 = 0, q3 = 0;
        if (threadIdx.x < gridDim.x) {
            const volatile float* p = scratch + 4 * threadIdx.x;
            q0 = p[0]; q1 = p[1]; q2 = p[2]; q3 = p[3];
        }
        const float t0 = block_sum(q0, red), t1 = block_sum(q1, red), t2 = block_sum(q2, red), t3 = block_sum(q3, red);
        if (threadIdx.x == 0) {
            const float inv3 = 1.0f / static_cast<float>(n3), invn = 1.0f / static_cast<float>(a.n);
            const float c_l = a.coef * 0.5f * t0 * inv3;
            float f_l = 0.0f, b_l = 0.0f, s_l = 0.0f;
            if (a.rgb_f != nullptr) f_l = a.coef * (a.beta != nullptr ? t1 * inv3 : 0.5f * t1 * inv3);
            if (a.beta != nullptr) {
                b_l = a.coef * (t2 * invn + 3.0f);
                if (a.ts != nullptr) s_l = a.coef * a.lambda_u * t3 / static_cast<float>(a.n_ts);
            }
            terms[0] = c_l; terms[1] = f_l; terms[2] = b_l; terms[3] = s_l; terms[4] = ((c_l + f_l) + b_l) + s_l;
            *ticket = 0u;                        // ready for the next call (graph replays reuse the scratch)
        }
    }
}

// g[0..4] = cotangents of (c_l, f_l, b_l, s_l, total), device scalars, any may be null (= 0)
__global__ void __launch_bounds__(kObjThreads)
nerfw_loss_bwd_kernel(const ObjArgs a, const float* g0, const float* g1, const float* g2, const float* g3, const float* g4,
                      float* __restrict__ d_rgb_c, float* __restrict__ d_rgb_f, float* __restrict__ d_beta,
                      float* __restrict__ d_ts) {
    const LossScalars sc = loss_scalars(g0, g1, g2, g3, g4, a.coef, a.n);
    const long long tid = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long r = tid; r < a.n; r += nth) {
        float acc = 0.0f;
        const float b = a.beta ? __ldg(a.beta + r) : 1.0f;
        const float ib2 = loss_inv_beta2(b);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float t = __ldg(a.tgt + 3 * r + c);
            d_rgb_c[3 * r + c] = loss_d_rgb_coarse(sc, __ldg(a.rgb_c + 3 * r + c), t);
            if (a.rgb_f != nullptr) {
                const float df = __fsub_rn(__ldg(a.rgb_f + 3 * r + c), t);
                d_rgb_f[3 * r + c] = loss_d_rgb_fine(sc, df, ib2, a.beta != nullptr);
                acc = __fmaf_rn(df, df, acc);
            }
        }
        if (a.beta != nullptr && d_beta != nullptr) d_beta[r] = loss_d_beta(sc, acc, ib2, b);
    }
    if (a.ts != nullptr && d_ts != nullptr) {
        const float v = loss_d_tsigma(sc, a.lambda_u, a.n_ts);
        for (long long i = tid; i < a.n_ts; i += nth) d_ts[i] = v;
    }
}


// metrics.py:4-13: mse = mean((pred - gt)^2), psnr = -10 log10(mse); same deterministic last-block reduction.
// clip01: eval.py:201 clips the prediction to [0,1] before it is compared (eval.py:208-210).
__global__ void __launch_bounds__(kObjThreads)
mse_psnr_kernel(const float* __restrict__ pred, const float* __restrict__ gt, long long n, int clip01,
                float* __restrict__ scratch, float* __restrict__ out) {
    __shared__ float red[kObjThreads / 32];
    __shared__ bool is_last;
    const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
    float s = 0.0f;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += nth) {
        float p = __ldg(pred + i);
        if (clip01) p = fminf(fmaxf(p, 0.0f), 1.0f);
        const float d = p - __ldg(gt + i);
        s += d * d;
    }
    const float r = block_sum(s, red);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + 4 * kObjMaxBlocks);
    if (threadIdx.x == 0) {
        scratch[4 * blockIdx.x] = r;
        __threadfence();
        is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        float q = 0.0f;
        if (threadIdx.x < gridDim.x) q = *reinterpret_cast<const volatile float*>(scratch + 4 * threadIdx.x);
        const float t = block_sum(q, red);
        if (threadIdx.x == 0) {
            const float mse = t / static_cast<float>(n);
            out[0] = mse;
            out[1] = -10.0f * log10f(mse);
            *ticket = 0u;
        }
    }
}

}  // namespace nwb

using namespace nwb;

static int obj_grid(const ObjArgs& a) {
    const long long work = std::max<long long>(3ll * a.n, a.ts ? a.n_ts : 0);
    return static_cast<int>(std::max<long long>(1, std::min<long long>(kObjMaxBlocks, (work + 4 * kObjThreads - 1) / (4 * kObjThreads))));
}

extern "C" size_t nwb_nerfw_loss_scratch_bytes(void) { return (4 * kObjMaxBlocks + 4) * sizeof(float); }

extern "C" int nwb_nerfw_loss_fwd(const float* rgb_coarse, const float* rgb_fine, const float* beta,
                                  const float* transient_sigmas, const float* targets, int n_rays, int64_t n_ts,
                                  float lambda_u, float coef, void* scratch, float* terms, void* stream) {
    NWB_REQUIRE(n_rays >= 1 && rgb_coarse && targets && scratch && terms, "nwb_nerfw_loss_fwd: null pointer / empty batch");
    NWB_REQUIRE(!(beta && !rgb_fine), "nwb_nerfw_loss_fwd: beta without rgb_fine");
    NWB_REQUIRE(!transient_sigmas || n_ts >= 1, "nwb_nerfw_loss_fwd: bad transient_sigmas size");
    ObjArgs a{rgb_coarse, rgb_fine, beta, transient_sigmas, targets, n_rays, n_ts, lambda_u, coef};
    nerfw_loss_fwd_kernel<<<obj_grid(a), kObjThreads, 0, static_cast<cudaStream_t>(stream)>>>(a, static_cast<float*>(scratch), terms);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

// ---- the same forward without caller scratch (no counterpart in the reference; added
// for callers that cannot allocate: graph capture, serving loops, harnesses that forbid cudaMalloc) ---------
// The partial sums and the ticket live in a file-scope device array, one per GPU that loads the library, held until exit.
// Every call on a device shares it: calls that overlap on the GPU (two streams) race on the partials and the ticket.
constexpr int kObjStaticMaxRays = 1 << 20;                   // NWB_LOSS_STATIC_MAX_RAYS in the header
constexpr int kObjScratchFloats = 4 * kObjMaxBlocks + 4;
static __device__ __align__(256) float g_obj_static_scratch[kObjScratchFloats];

extern "C" int nwb_nerfw_loss_fwd_static(const float* rgb_coarse, const float* rgb_fine, const float* beta,
                                         const float* transient_sigmas, const float* targets, int n_rays, int64_t n_ts,
                                         float lambda_u, float coef, float* terms, void* stream) {
    NWB_REQUIRE(n_rays >= 1 && rgb_coarse && targets && terms, "nwb_nerfw_loss_fwd_static: null pointer / empty batch");
    NWB_REQUIRE(n_rays <= kObjStaticMaxRays,
                "nwb_nerfw_loss_fwd_static: n_rays = %d is beyond the static scratch's maximum of %d rays "
                "(use nwb_nerfw_loss_fwd with caller-owned scratch)", n_rays, kObjStaticMaxRays);
    NWB_REQUIRE(!(beta && !rgb_fine), "nwb_nerfw_loss_fwd_static: beta without rgb_fine");
    NWB_REQUIRE(!transient_sigmas || n_ts >= 1, "nwb_nerfw_loss_fwd_static: bad transient_sigmas size");
    // In host code the array's name is a host-side stand-in, not a device address: the address is looked up once per
    // device, with that device current, and it is the looked-up pointer that goes to the kernel.
    static float* scratch_of[kMaxDevices] = {};
    static bool resolved[kMaxDevices] = {};
    int dev = 0;
    int rc = current_device_checked(&dev);
    if (rc != NWB_OK) return rc;
    NWB_CONFIGURE_ONCE(resolved, {
        float* ptr = nullptr;
        NWB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_obj_static_scratch));
        NWB_REQUIRE(ptr != nullptr, "nwb_nerfw_loss_fwd_static: no device address for the static scratch");
        scratch_of[dev] = ptr;
    });
    float* scratch = scratch_of[dev];
    const cudaStream_t st = static_cast<cudaStream_t>(stream);
    // the array keeps the last call's contents: the ticket must start at zero (a call that was cut short may have left it
    // counting); every partial slot the grid reads is written by the grid first
    NWB_CUDA(cudaMemsetAsync(scratch + 4 * kObjMaxBlocks, 0, 4 * sizeof(float), st));
    ObjArgs a{rgb_coarse, rgb_fine, beta, transient_sigmas, targets, n_rays, n_ts, lambda_u, coef};
    nerfw_loss_fwd_kernel<<<obj_grid(a), kObjThreads, 0, st>>>(a, scratch, terms);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

extern "C" int nwb_nerfw_loss_bwd(const float* rgb_coarse, const float* rgb_fine, const float* beta,
                                  const float* transient_sigmas, const float* targets, int n_rays, int64_t n_ts,
                                  float lambda_u, float coef, const float* const* g_terms, float* d_rgb_coarse,
                                  float* d_rgb_fine, float* d_beta, float* d_transient_sigmas, void* stream) {
    NWB_REQUIRE(n_rays >= 1 && rgb_coarse && targets && g_terms && d_rgb_coarse, "nwb_nerfw_loss_bwd: null pointer / empty batch");
    NWB_REQUIRE(!rgb_fine || d_rgb_fine, "nwb_nerfw_loss_bwd: d_rgb_fine is null");
    ObjArgs a{rgb_coarse, rgb_fine, beta, transient_sigmas, targets, n_rays, n_ts, lambda_u, coef};
    nerfw_loss_bwd_kernel<<<obj_grid(a), kObjThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        a, g_terms[0], g_terms[1], g_terms[2], g_terms[3], g_terms[4], d_rgb_coarse, d_rgb_fine, d_beta, d_transient_sigmas);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

extern "C" int nwb_mse_psnr(const float* pred, const float* gt, int64_t n, int clip01, void* scratch, float* out, void* stream) {
    NWB_REQUIRE(n >= 1 && pred && gt && scratch && out, "nwb_mse_psnr: null pointer / empty input");
    const int grid = static_cast<int>(std::max<long long>(1, std::min<long long>(kObjMaxBlocks, (n + 4 * kObjThreads - 1) / (4 * kObjThreads))));
    mse_psnr_kernel<<<grid, kObjThreads, 0, static_cast<cudaStream_t>(stream)>>>(pred, gt, n, clip01, static_cast<float*>(scratch), out);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
