// Adam on the flat parameter / gradient buffers (reference: utils/__init__.py:24-33 builds
// torch.optim.Adam(lr, eps=1e-8, weight_decay) over every trainable tensor; train.py:139-142).
// One vectorised elementwise pass over the 1.3 M parameters: reads p, g, m, v, writes p, m, v and
// (optionally) clears g for the next step's atomic accumulation -- HBM-bound, 7 floats per element.
#include "field_pack.cuh"
#include "field_dev.cuh"
#include <algorithm>
#include <math.h>

namespace nwb {

// torch.optim.Adam's update; every rounding spelled out so that the plain and the cache-refreshing kernels (and any
// future variant) produce the same bits
__device__ __forceinline__ void adam_update(float& pp, float gg, float& mm, float& vv, float beta1, float beta2, float eps,
                                            float wd, float gscale, float step_size, float inv_sqrt_bc2) {
#ifdef NWB_ADAM_NOMATH            // timing experiment only
    mm = mm + gg; vv = vv + gg; pp = pp + step_size * mm * inv_sqrt_bc2 + eps + beta1 + beta2 + wd + gscale; return;
#endif
    gg = __fmaf_rn(wd, pp, __fmul_rn(gg, gscale));
    mm = __fmaf_rn(beta1, mm, __fmul_rn(1.0f - beta1, gg));
    vv = __fmaf_rn(beta2, vv, __fmul_rn(__fmul_rn(1.0f - beta2, gg), gg));
    pp = __fsub_rn(pp, __fdiv_rn(__fmul_rn(step_size, mm), __fmaf_rn(__fsqrt_rn(vv), inv_sqrt_bc2, eps)));
}

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
        adam_update(pp, gg, mm, vv, beta1, beta2, eps, wd, gscale, step_size, inv_sqrt_bc2);
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


// ---- (f2) Adam that also refreshes the packed operand caches ------------------------------------------------------
// One launch walks (1) every element of every field's bf16 weight images and (2) constant block -- each maps to exactly
// one fp32 parameter (field_pack.cuh) which is updated in place and written straight into the cache -- and (3) `rest`,
// the flat offsets no image or constant refers to (per-ray tail columns of dir_encoding / transient_encoding.0, their
// biases, the embedding tables, alignment padding).  The step state advances inside the same launch: every block derives
// the new step's scalars from state[0], the last block to finish stores them (ticket in state[4]).
constexpr int kAdamMaxFields = 4;
#ifndef NWB_ADAM_BLOCKS_X
#define NWB_ADAM_BLOCKS_X 24
#endif
constexpr int kAdamPackBlocksX = NWB_ADAM_BLOCKS_X;      // blocks per grid slice: ~650 blocks of 256 threads = 1.5 waves at 3 blocks/SM

struct AdamPackArgs {
    PackArgs f[kAdamMaxFields];
    uint8_t* packed[kAdamMaxFields];
    int y0[kAdamMaxFields + 1];          // first blockIdx.y of each field's slices (its layers, then its constants)
    int n_fields;
    float *p, *g, *m, *v;
    const int32_t* rest;
    int n_rest;
    unsigned char* cover;                // mark mode (setup): cover[off] = 1 for every referenced parameter, no update
    float beta1, beta2, eps, wd, gscale;
    int zero_grad;
    int no_advance;                      // NWB_ADAM_NO_ADVANCE: this launch is one of several of the same step
    float* state;
    // multi-GPU (csrc/p2p.cu): the summed gradient of flat offset i sits in rank (i / slice)'s reduced buffer -- read it from
    // there (peer memory over NVLink) instead of from g; g is still cleared.  world == 0: plain g.
    const float* red[16];
    long long slice;
    int world;
};

__device__ __forceinline__ float4 ld_sys_f4(const float* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_sys_f(const float* p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
// gradient of flat offset off (grad4: off % 4 == 0; a slice is a multiple of 4 floats, so the four share a slice)
__device__ __forceinline__ float grad1(const AdamPackArgs& a, size_t off) {
    if (a.world == 0) return a.g[off];
    const long long q = static_cast<long long>(off) / a.slice;
    return ld_sys_f(a.red[q] + (off - q * a.slice));
}
__device__ __forceinline__ float4 grad4(const AdamPackArgs& a, size_t off) {
    if (a.world == 0) return *reinterpret_cast<const float4*>(a.g + off);
    const long long q = static_cast<long long>(off) / a.slice;
    return ld_sys_f4(a.red[q] + (off - q * a.slice));
}

#ifdef NWB_ADAM_TRACE               // tools/trace_adam.py only: per-block %globaltimer stamps (start, after the prologue, end)
__device__ unsigned long long g_adam_trace[3 * 4096];
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define NWB_ADAM_STAMP(k) do { if (threadIdx.x == 0) g_adam_trace[3 * (blockIdx.y * gridDim.x + blockIdx.x) + (k)] = gtime(); } while (0)
#else
#define NWB_ADAM_STAMP(k) do { } while (0)
#endif

__global__ void __launch_bounds__(256, 3)
adam_pack_kernel(const AdamPackArgs a) {
    NWB_ADAM_STAMP(0);
    __shared__ float s_hyp[2];
    __shared__ double s_pow[2];
    const bool mark = a.cover != nullptr;
    int t_new = 0;
    if (threadIdx.x == 0 && !mark) {
        // beta^t as running products in double (state words 6..9, valid for the step in word 5): every block of the grid
        // needs the step's scalars before its first update, and two double-precision pow() per block were the longest
        // thing this launch did (FP64 runs at 1/64 rate here).  A state written from outside (resume: word 0 set, word 5
        // not) falls back to pow() once.
#ifdef NWB_ADAM_NOPROLOGUE       // timing experiment only: last step's scalars, no double-precision math
        s_hyp[0] = a.state[2]; s_hyp[1] = a.state[3]; s_pow[0] = s_pow[1] = 0.5;
        if (false) {
#else
        {
#endif
        const int t_old = *reinterpret_cast<const volatile int*>(a.state);
        t_new = t_old + 1;
        const volatile double* pw = reinterpret_cast<const volatile double*>(a.state + 6);
        double b1t, b2t;
        if (*reinterpret_cast<const volatile int*>(a.state + 5) == t_old && t_old > 0) {
            b1t = pw[0] * static_cast<double>(a.beta1);
            b2t = pw[1] * static_cast<double>(a.beta2);
        } else {
            b1t = pow(static_cast<double>(a.beta1), t_new);
            b2t = pow(static_cast<double>(a.beta2), t_new);
        }
        s_pow[0] = b1t; s_pow[1] = b2t;
        const double bc1 = 1.0 - b1t, bc2 = 1.0 - b2t;
        s_hyp[0] = static_cast<float>(static_cast<double>(a.state[1]) / bc1);
        s_hyp[1] = static_cast<float>(1.0 / sqrt(bc2));
        }
    }
    __syncthreads();
    NWB_ADAM_STAMP(1);
    const float step_size = mark ? 0.0f : s_hyp[0], inv_sqrt_bc2 = mark ? 1.0f : s_hyp[1];
    // returns the updated parameter
    auto step_at = [&](const float* src) -> float {
        const size_t off = static_cast<size_t>(src - a.p);
        if (mark) { a.cover[off] = 1; return 0.0f; }
        float pp = a.p[off], mm = a.m[off], vv = a.v[off];
        adam_update(pp, grad1(a, off), mm, vv, a.beta1, a.beta2, a.eps, a.wd, a.gscale, step_size, inv_sqrt_bc2);
        a.p[off] = pp; a.m[off] = mm; a.v[off] = vv;
        if (a.zero_grad) a.g[off] = 0.0f;
        return pp;
    };
    const int y = blockIdx.y;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
#ifdef NWB_ADAM_LAYERS_ONLY       // timing experiment only: rest list and constant blocks skipped
    if (y >= a.y0[a.n_fields]) return;
    { int f2 = 0; while (f2 + 1 < a.n_fields && y >= a.y0[f2 + 1]) ++f2; if (y - a.y0[f2] >= a.f[f2].fs.n_packed_layers()) return; }
#endif
    if (y >= a.y0[a.n_fields]) {
        // the rest list takes as many grid slices as give every thread one element: walked by one slice each thread did
        // ~18 dependent (index -> p, g, m, v -> store) round trips to memory one after the other, the longest thing in
        // this launch
        const int rs = y - a.y0[a.n_fields], n_rs = static_cast<int>(gridDim.y) - a.y0[a.n_fields];
        for (int i = rs * nth + tid; i < a.n_rest; i += n_rs * nth) step_at(a.p + a.rest[i]);
    } else {
        int f = 0;
        while (f + 1 < a.n_fields && y >= a.y0[f + 1]) ++f;
        const PackArgs& pa = a.f[f];
        const FieldShape& fs = pa.fs;
        const int g = y - a.y0[f];
        if (g < fs.n_packed_layers()) {
            // one thread per 16-byte image chunk = 8 consecutive k of one row: when the 8 sources are contiguous and
            // 16 B aligned (every layer whose row length is a multiple of 4) p, g, m, v move as float4 pairs
            const int N = fs.packed_n(g), K8 = fs.layer_kb(g) * 8;
            uint8_t* dst = a.packed[f] + fs.w_off(g);
            // Tensors whose rows are not a multiple of 4 floats (trunk layer 0: in_xyz = 63 columns; the skip layer: 63+256;
            // dir_encoding.0: 256+27+48) cannot be walked by image chunk with 16-byte accesses -- 8 scalar loads and
            // stores per array and chunk, each touching 32 sectors per warp, made these slices the longest of the launch
            // ([B200, tools/trace_adam.py] 16 us against 5 us for an aligned layer).  They are walked FLAT instead: the
            // tensor's storage is contiguous and starts 16 B aligned, so p, g, m, v move as float4 and only the bf16 image
            // elements are scattered (2-byte stores into L2-resident images).  The whole tensor is covered, columns that
            // have no image element included (the per-ray tail of dir_encoding.0), so the rest list leaves them alone.
            const float* flat_base = nullptr;
            int flat_rows = 0, flat_ld = 0;
            if (g == 0) { flat_base = pa.p.trunk_w[0]; flat_rows = kW; flat_ld = fs.in_xyz; }
            else if (g == 4) { flat_base = pa.p.trunk_w[4]; flat_rows = kW; flat_ld = fs.in_xyz + kW; }
            else if (g == 9) { flat_base = pa.p.dir_w; flat_rows = kHalf; flat_ld = kW + fs.in_dir + fs.in_a; }
            const bool flat = flat_base != nullptr && (flat_ld & 3) != 0 && (reinterpret_cast<uintptr_t>(flat_base) & 15u) == 0;
            if (flat) {
                const int total = flat_rows * flat_ld;
                const size_t off0 = static_cast<size_t>(flat_base - a.p);
                for (int j = tid; 4 * j < total; j += nth) {
                    const int i0 = 4 * j, cnt = min(4, total - i0);
                    const size_t off = off0 + i0;
                    if (mark) {
                        for (int e = 0; e < cnt; ++e) a.cover[off + e] = 1;
                        continue;
                    }
                    float pp[4], gg[4], mm[4], vv[4];
                    if (cnt == 4) {
                        const float4 p4 = *reinterpret_cast<const float4*>(a.p + off), g4 = grad4(a, off);
                        const float4 m4 = *reinterpret_cast<const float4*>(a.m + off), v4 = *reinterpret_cast<const float4*>(a.v + off);
                        pp[0] = p4.x; pp[1] = p4.y; pp[2] = p4.z; pp[3] = p4.w;  gg[0] = g4.x; gg[1] = g4.y; gg[2] = g4.z; gg[3] = g4.w;
                        mm[0] = m4.x; mm[1] = m4.y; mm[2] = m4.z; mm[3] = m4.w;  vv[0] = v4.x; vv[1] = v4.y; vv[2] = v4.z; vv[3] = v4.w;
                    } else {
                        for (int e = 0; e < 4; ++e) {
                            const bool in = e < cnt;
                            pp[e] = in ? a.p[off + e] : 0.0f; gg[e] = in ? grad1(a, off + e) : 0.0f;
                            mm[e] = in ? a.m[off + e] : 0.0f; vv[e] = in ? a.v[off + e] : 0.0f;
                        }
                    }
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        adam_update(pp[e], gg[e], mm[e], vv[e], a.beta1, a.beta2, a.eps, a.wd, a.gscale, step_size, inv_sqrt_bc2);
                    if (cnt == 4) {
                        *reinterpret_cast<float4*>(a.p + off) = make_float4(pp[0], pp[1], pp[2], pp[3]);
                        *reinterpret_cast<float4*>(a.m + off) = make_float4(mm[0], mm[1], mm[2], mm[3]);
                        *reinterpret_cast<float4*>(a.v + off) = make_float4(vv[0], vv[1], vv[2], vv[3]);
                        if (a.zero_grad) *reinterpret_cast<float4*>(a.g + off) = make_float4(0.f, 0.f, 0.f, 0.f);
                    } else {
                        for (int e = 0; e < cnt; ++e) {
                            a.p[off + e] = pp[e]; a.m[off + e] = mm[e]; a.v[off + e] = vv[e];
                            if (a.zero_grad) a.g[off + e] = 0.0f;
                        }
                    }
                    int n = i0 / flat_ld, k = i0 - n * flat_ld;
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        if (e < cnt) {
                            // image column of tensor column k (field_pack.cuh pack_ptr, inverted)
                            int kc = k;
                            if (g == 4 && k >= fs.in_xyz) kc = 64 * fs.kx + (k - fs.in_xyz);
                            if (g == 9 && k >= kW) kc = -1;
                            if (kc >= 0)
                                *reinterpret_cast<__nv_bfloat16*>(dst + pack_dst_off(N, n, kc)) = __float2bfloat16_rn(pp[e]);
                        }
                        if (++k == flat_ld) { k = 0; ++n; }
                    }
                }
            }
            // the chunk walk: every layer with aligned rows, and the transient_encoding.0 rows of layer 9 next to a flat dir half
            const int n_lo = (flat && g == 9) ? kHalf : (flat ? N : 0);
            for (int c = tid + n_lo * K8; c < N * K8; c += nth) {
                const int n = c / K8, k0 = (c - n * K8) << 3;
                const float* s0 = pack_ptr(pa, g, n, k0);
                const float* s7 = pack_ptr(pa, g, n, k0 + 7);
                float w[8];
                if (!mark && s0 != nullptr && s7 == s0 + 7 && (reinterpret_cast<uintptr_t>(s0) & 15u) == 0) {
                    const size_t off = static_cast<size_t>(s0 - a.p);
                    float4 pp[2], gg[2], mm[2], vv[2];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        pp[h] = reinterpret_cast<const float4*>(a.p + off)[h];
                        gg[h] = grad4(a, off + 4 * h);
                        mm[h] = reinterpret_cast<const float4*>(a.m + off)[h];
                        vv[h] = reinterpret_cast<const float4*>(a.v + off)[h];
                    }
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        adam_update(pp[h].x, gg[h].x, mm[h].x, vv[h].x, a.beta1, a.beta2, a.eps, a.wd, a.gscale, step_size, inv_sqrt_bc2);
                        adam_update(pp[h].y, gg[h].y, mm[h].y, vv[h].y, a.beta1, a.beta2, a.eps, a.wd, a.gscale, step_size, inv_sqrt_bc2);
                        adam_update(pp[h].z, gg[h].z, mm[h].z, vv[h].z, a.beta1, a.beta2, a.eps, a.wd, a.gscale, step_size, inv_sqrt_bc2);
                        adam_update(pp[h].w, gg[h].w, mm[h].w, vv[h].w, a.beta1, a.beta2, a.eps, a.wd, a.gscale, step_size, inv_sqrt_bc2);
                        reinterpret_cast<float4*>(a.p + off)[h] = pp[h];
                        reinterpret_cast<float4*>(a.m + off)[h] = mm[h];
                        reinterpret_cast<float4*>(a.v + off)[h] = vv[h];
                        if (a.zero_grad) reinterpret_cast<float4*>(a.g + off)[h] = make_float4(0.f, 0.f, 0.f, 0.f);
                        w[4 * h] = pp[h].x; w[4 * h + 1] = pp[h].y; w[4 * h + 2] = pp[h].z; w[4 * h + 3] = pp[h].w;
                    }
                } else if (mark) {
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const float* src = pack_ptr(pa, g, n, k0 + q);
                        w[q] = src ? step_at(src) : 0.0f;
                    }
                } else {
                    // rows that are not 16 B aligned (in_xyz = 63: layers 0 and 4; the 331-float rows of dir_encoding.0):
                    // all eight elements' loads are issued before the first update -- element by element each store could
                    // alias the next load, so the eight memory round trips ran one after the other (the longest threads of
                    // this launch: [B200, ncu r02ze] 28 us with the SMs active half of the time)
                    int off8[8];                                      // flat offsets (the buffer is far below 2^31 floats)
                    float pp[8], gg[8], mm[8], vv[8];
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const float* src = pack_ptr(pa, g, n, k0 + q);
                        off8[q] = src ? static_cast<int>(src - a.p) : -1;
                    }
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        if (off8[q] >= 0) {
                            pp[q] = a.p[off8[q]]; gg[q] = grad1(a, off8[q]); mm[q] = a.m[off8[q]]; vv[q] = a.v[off8[q]];
                        }
                    }
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        w[q] = 0.0f;
                        if (off8[q] >= 0) {
                            adam_update(pp[q], gg[q], mm[q], vv[q], a.beta1, a.beta2, a.eps, a.wd, a.gscale, step_size, inv_sqrt_bc2);
                            a.p[off8[q]] = pp[q]; a.m[off8[q]] = mm[q]; a.v[off8[q]] = vv[q];
                            if (a.zero_grad) a.g[off8[q]] = 0.0f;
                            w[q] = pp[q];
                        }
                    }
                }
                if (!mark) {
                    uint4 o;
                    o.x = pack_bf16(w[0], w[1]); o.y = pack_bf16(w[2], w[3]);
                    o.z = pack_bf16(w[4], w[5]); o.w = pack_bf16(w[6], w[7]);
#ifdef NWB_ADAM_NOIMAGE           // timing experiment only
                    if (o.x == 0x12345678u && o.y == 0x9abcdef0u)
#endif
                    *reinterpret_cast<uint4*>(dst + pack_dst_off(N, n, k0)) = o;
                }
            }
        } else {
            float* c = reinterpret_cast<float*>(a.packed[f] + fs.w_bytes());
            for (int i = tid; i < ConstOff::count; i += nth) {
                const float* src = const_ptr(pa, i);
                const float w = src ? step_at(src) : 0.0f;
                if (!mark) c[i] = w;
            }
        }
    }
    if (mark) return;
#ifdef NWB_ADAM_TRACE
    __syncthreads();
    NWB_ADAM_STAMP(2);
#endif
    if (a.no_advance) return;         // a later launch of the same step (same state, same scalars) advances it
#ifdef NWB_ADAM_NOTICKET          // timing experiment only: the step state is not advanced
    return;
#endif
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int* ticket = reinterpret_cast<unsigned int*>(a.state + 4);
        __threadfence();
        if (atomicAdd(ticket, 1u) == gridDim.x * gridDim.y - 1) {
            *reinterpret_cast<int*>(a.state) = t_new;
            a.state[2] = s_hyp[0];
            a.state[3] = s_hyp[1];
            *reinterpret_cast<int*>(a.state + 5) = t_new;
            reinterpret_cast<double*>(a.state + 6)[0] = s_pow[0];
            reinterpret_cast<double*>(a.state + 6)[1] = s_pow[1];
            *ticket = 0u;
        }
    }
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

extern "C" int64_t nwb_p2p_slice_floats(int64_t n_floats, int world);
extern "C" size_t nwb_p2p_reduced_offset(int64_t n_floats);

static int adam_pack_launch(float* p, float* g, float* m, float* v, int64_t n, float beta1, float beta2, float eps,
                            float weight_decay, float* state, float grad_scale, int zero_grad, int n_fields,
                            const nwb_field_desc* const* descs, const nwb_field_params* const* params,
                            void* const* packed, const int32_t* rest, int64_t n_rest, unsigned char* cover, void* stream,
                            const char* who, void* const* regions = nullptr, int world = 0) {
    NWB_REQUIRE(n >= 1 && n < (1ll << 31) && p && n_fields >= 1 && n_fields <= kAdamMaxFields && descs && params && packed,
                "%s: bad size / null pointer (at most %d fields)", who, kAdamMaxFields);
    NWB_REQUIRE(n_rest >= 0 && n_rest < (1ll << 31) && (n_rest == 0 || rest), "%s: bad rest list", who);
    AdamPackArgs a{};
    a.n_fields = n_fields;
    int y = 0;
    for (int f = 0; f < n_fields; ++f) {
        int rc = make_shape(descs[f], &a.f[f].fs);
        if (rc != NWB_OK) return rc;
        rc = check_pack_params(a.f[f].fs, params[f], packed[f], who);
        if (rc != NWB_OK) return rc;
        a.f[f].p = *params[f];
        // every parameter the caches are built from must live inside the flat buffer
        const float* const* slots = reinterpret_cast<const float* const*>(params[f]);
        for (size_t i = 0; i < sizeof(nwb_field_params) / sizeof(float*); ++i)
            NWB_REQUIRE(slots[i] == nullptr || (slots[i] >= p && slots[i] < p + n),
                        "%s: parameter slot %d of field %d is outside the flat buffer", who, static_cast<int>(i), f);
        a.packed[f] = static_cast<uint8_t*>(packed[f]);
        a.y0[f] = y;
        y += a.f[f].fs.n_packed_layers() + 1;
    }
    a.y0[n_fields] = y;
    a.p = p; a.g = g; a.m = m; a.v = v;
    a.rest = rest; a.n_rest = static_cast<int>(n_rest);
    a.cover = cover;
    a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.wd = weight_decay; a.gscale = grad_scale;
    a.zero_grad = zero_grad & 1; a.no_advance = (zero_grad & NWB_ADAM_NO_ADVANCE) != 0;
    a.state = state;
    if (regions != nullptr && world >= 2) {
        NWB_REQUIRE(world <= 16, "%s: world > 16", who);
        a.world = world;
        a.slice = nwb_p2p_slice_floats(n, world);
        const size_t roff = nwb_p2p_reduced_offset(n);
        for (int q = 0; q < world; ++q) {
            NWB_REQUIRE(regions[q] != nullptr, "%s: region %d is null", who, q);
            a.red[q] = reinterpret_cast<const float*>(static_cast<const uint8_t*>(regions[q]) + roff);
        }
    }
    const int rest_slices = std::max<int>(1, static_cast<int>((n_rest + kAdamPackBlocksX * 256 - 1) / (kAdamPackBlocksX * 256)));
    dim3 grid(kAdamPackBlocksX, y + std::min(rest_slices, 64));
    adam_pack_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(a);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

#ifdef NWB_ADAM_TRACE
extern "C" int nwb_debug_adam_trace(unsigned long long* host_out, int n) {
    NWB_CUDA(cudaMemcpyFromSymbol(host_out, g_adam_trace, sizeof(unsigned long long) * n));
    return NWB_OK;
}
#endif

extern "C" int nwb_adam_pack_cover(float* p, int64_t n, int n_fields, const nwb_field_desc* const* descs,
                                   const nwb_field_params* const* params, void* const* packed, unsigned char* cover,
                                   void* stream) {
    NWB_REQUIRE(cover, "nwb_adam_pack_cover: cover is null");
    return adam_pack_launch(p, nullptr, nullptr, nullptr, n, 0.f, 0.f, 0.f, 0.f, nullptr, 0.f, 0, n_fields, descs, params,
                            packed, nullptr, 0, cover, stream, "nwb_adam_pack_cover");
}

// the same launch as the second half of the multi-GPU gradient exchange: gradients are read from the ranks' reduced slices
// (after nwb_p2p_reduce_scatter on the same stream), g is only cleared.  regions as for nwb_p2p_allreduce.
extern "C" int nwb_adam_pack_step_p2p(float* p, float* g, float* m, float* v, int64_t n, float beta1, float beta2, float eps,
                                      float weight_decay, float* state, float grad_scale, int zero_grad, int n_fields,
                                      const nwb_field_desc* const* descs, const nwb_field_params* const* params,
                                      void* const* packed, const int32_t* rest, int64_t n_rest, void* const* regions, int world,
                                      void* stream) {
    NWB_REQUIRE(g && m && v && state && regions, "nwb_adam_pack_step_p2p: null pointer");
    NWB_REQUIRE((reinterpret_cast<uintptr_t>(state) & 15u) == 0, "nwb_adam_pack_step_p2p: state must be 16 B aligned (48 B long)");
    NWB_REQUIRE(world >= 2, "nwb_adam_pack_step_p2p: world must be >= 2");
    return adam_pack_launch(p, g, m, v, n, beta1, beta2, eps, weight_decay, state, grad_scale, zero_grad, n_fields, descs,
                            params, packed, rest, n_rest, nullptr, stream, "nwb_adam_pack_step_p2p", regions, world);
}

extern "C" int nwb_adam_pack_step_dev(float* p, float* g, float* m, float* v, int64_t n, float beta1, float beta2, float eps,
                                      float weight_decay, float* state, float grad_scale, int zero_grad, int n_fields,
                                      const nwb_field_desc* const* descs, const nwb_field_params* const* params,
                                      void* const* packed, const int32_t* rest, int64_t n_rest, void* stream) {
    NWB_REQUIRE(g && m && v && state, "nwb_adam_pack_step_dev: null pointer");
    NWB_REQUIRE((reinterpret_cast<uintptr_t>(state) & 15u) == 0, "nwb_adam_pack_step_dev: state must be 16 B aligned (48 B long)");
    return adam_pack_launch(p, g, m, v, n, beta1, beta2, eps, weight_decay, state, grad_scale, zero_grad, n_fields, descs,
                            params, packed, rest, n_rest, nullptr, stream, "nwb_adam_pack_step_dev");
}
