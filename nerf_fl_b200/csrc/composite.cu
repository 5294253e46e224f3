// (a5) volume rendering, forward and backward -- models/rendering.py:141-226.
// One warp per ray.  Per 32-sample chunk the warp pulls the AoS raw rows (C = 1|4|9 floats per
// sample) through a coalesced, float4-vectorised copy into its shared-memory slab, each lane
// then owns one sample; transmittance is a warp product-scan carried across chunks.  Backward
// keeps T in shared memory and runs the division-free reverse suffix scan of SURVEY.md A.3.
#include "objective.cuh"
#include <algorithm>

namespace nwb {

constexpr int kCompWarps = 4;

// ---- cp.async staging ring ------------------------------------------------------------------------------
// These kernels are latency-bound: with one chunk prefetched in registers a warp keeps 2 x 1.3 KB in flight and the SM
// ~30 KB, which Little's law turns into ~50 % of HBM peak (ncu r01d).  Here each warp owns a ring of kRing stages in
// shared memory filled by cp.async (no registers held by loads in flight), kRing-1 chunks ahead of the one it consumes.
constexpr int kRing = 4;
__device__ __forceinline__ uint32_t comp_smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(comp_smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(void* dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(comp_smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int C>
struct RingStage {
    float raw[32 * C];
    float z[36];                   // z[base .. base+32] (33 values: lane i's z_next is z[i+1])
};
// one chunk (32 samples of `ray` from `base`) -> stage; every lane commits exactly one group per call
template <int C>
__device__ __forceinline__ void ring_issue(RingStage<C>* st, const float* __restrict__ raw, const float* __restrict__ z,
                                           int ray, int n_rays, int base, int S, int lane) {
    if (ray < n_rays) {
        const int64_t row = static_cast<int64_t>(ray) * S;
        const int cnt = min(32, S - base);
        const float* src = raw + (row + base) * C;
        const int nfl = cnt * C;
        if (cnt == 32 && (reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
            // full chunk: 8 C float4, a fixed number of copies per lane (no loop counters, no tail tests)
#pragma unroll
            for (int k = 0; k < (8 * C + 31) / 32; ++k) {
                const int i = lane + 32 * k;
                if (8 * C % 32 == 0 || k + 1 < (8 * C + 31) / 32 || i < 8 * C) cp_async16(st->raw + 4 * i, src + 4 * i);
            }
        } else if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0 && (nfl & 3) == 0) {
            for (int i = lane; i < (nfl >> 2); i += 32) cp_async16(st->raw + 4 * i, src + 4 * i);
        } else {
            for (int i = lane; i < nfl; i += 32) cp_async4(st->raw + i, src + i);
        }
        if (lane < cnt) cp_async4(st->z + lane, z + row + base + lane);
        if (lane == 0 && base + cnt < S) cp_async4(st->z + cnt, z + row + base + cnt);
    }
    cp_async_commit();
}

__device__ __forceinline__ float excl_prod_scan(float om, float& carry, int lane) {
    // returns carry * prod_{j<lane} om_j ; updates carry *= prod_{j<32} om_j
    float p = om;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const float t = __shfl_up_sync(0xffffffffu, p, o);
        if (lane >= o) p *= t;
    }
    float ex = __shfl_up_sync(0xffffffffu, p, 1);
    if (lane == 0) ex = 1.0f;
    const float T = carry * ex;
    carry *= __shfl_sync(0xffffffffu, p, 31);
    return T;
}

// kV: bit 0 = transient (C == 9 only), bit 1 = test_time extras (transient only), bit 2 = weights / opacity only -- compile-time,
// so that the per-chunk body carries no flag tests (the kernel issues on every scheduler slot it gets: ncu r02z)
template <int C, uint32_t kV>
__global__ void __launch_bounds__(kCompWarps * 32)
composite_fwd_kernel(const float* __restrict__ raw, const float* __restrict__ z,
                     const float* __restrict__ noise, float noise_std, int n_rays, int S,
                     uint32_t flags, float beta_min, nwb_composite_out o) {
    __shared__ __align__(16) RingStage<C> ring[kCompWarps][kRing];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (o.clear_word != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *o.clear_word = 0u;
    RingStage<C>* my = ring[warp];
    constexpr bool transient = (C == 9) && (kV & 1u);
    const bool white = flags & NWB_COMP_WHITE_BACK;
    constexpr bool test_time = (kV & 2u) != 0;
    constexpr bool weights_only = (C == 1) || (kV & 4u);
    // the warp's chunk sequence: (ray, 0), (ray, 32), ..., (ray + stride, 0), ...; `pf_*` is the next chunk to issue
    const int stride = gridDim.x * kCompWarps;
    int pf_ray = blockIdx.x * kCompWarps + warp, pf_base = 0;
    uint32_t it = 0;
#pragma unroll
    for (int k = 0; k < kRing - 1; ++k) {
        ring_issue<C>(&my[k], raw, z, pf_ray, n_rays, pf_base, S, lane);
        pf_base += 32;
        if (pf_base >= S) { pf_base = 0; pf_ray += stride; }
    }

    for (int ray = blockIdx.x * kCompWarps + warp; ray < n_rays; ray += stride) {
        const int64_t row = static_cast<int64_t>(ray) * S;
        float cT = 1.0f, cTs = 1.0f, cTt = 1.0f;
        float opac = 0, depth = 0, beta = 0, sr = 0, sg = 0, sb = 0, tr = 0, tg = 0, tb = 0;
        float xsr = 0, xsg = 0, xsb = 0, xsd = 0, xtr = 0, xtg = 0, xtb = 0, xtd = 0;
        for (int base = 0; base < S; base += 32, ++it) {
            const int cnt = min(32, S - base);
            // refill the stage consumed one iteration ago (every lane passed the __syncwarp below since), then wait
            // for this iteration's chunk: kRing groups are outstanding, the oldest is the one needed
            ring_issue<C>(&my[(it + kRing - 1) % kRing], raw, z, pf_ray, n_rays, pf_base, S, lane);
            pf_base += 32;
            if (pf_base >= S) { pf_base = 0; pf_ray += stride; }
            cp_async_wait<kRing - 1>();
            __syncwarp();
            const RingStage<C>* st = &my[it % kRing];
            const int i = base + lane;
            const bool live = lane < cnt;
            float zi = 0, delta = 0, sig_s = 0, sig_t = 0, cr = 0, cg = 0, cb = 0, qr = 0, qg = 0, qb = 0, bt = 0;
            if (live) {
                zi = st->z[lane];
                delta = (i + 1 < S) ? __fsub_rn(st->z[lane + 1], zi) : 1e2f;
                const float* v = st->raw + lane * C;
                if (C == 1) {
                    sig_s = v[0];
                } else {
                    cr = v[0]; cg = v[1]; cb = v[2]; sig_s = v[3];
                }
                if (C == 9) { qr = v[4]; qg = v[5]; qb = v[6]; sig_t = v[7]; bt = v[8]; }
            }
            __syncwarp();
            float alpha, a_s = 0, a_t = 0;
            if (transient) {
                a_s = __fsub_rn(1.0f, expf(-__fmul_rn(delta, sig_s)));
                a_t = __fsub_rn(1.0f, expf(-__fmul_rn(delta, sig_t)));
                alpha = __fsub_rn(1.0f, expf(-__fmul_rn(delta, __fadd_rn(sig_s, sig_t))));
            } else {
                float s = sig_s;
                if (noise != nullptr && live) s = __fadd_rn(s, __fmul_rn(__ldg(noise + row + i), noise_std));
                alpha = __fsub_rn(1.0f, expf(-__fmul_rn(delta, fmaxf(s, 0.0f))));
            }
            if (!live) { alpha = 0; a_s = 0; a_t = 0; }
            const float T = excl_prod_scan(__fsub_rn(1.0f, alpha), cT, lane);
            const float w = alpha * T;
            if (live) {
                if (o.weights) o.weights[row + i] = w;
                if (transient && o.transient_sigmas) o.transient_sigmas[row + i] = sig_t;
            }
            opac += w;
            if (!weights_only) {
                depth += w * zi;
                if (transient) {
                    const float ws = a_s * T, wt = a_t * T;
                    sr += ws * cr; sg += ws * cg; sb += ws * cb;
                    tr += wt * qr; tg += wt * qg; tb += wt * qb;
                    beta += wt * bt;
                    if (test_time) {
                        const float Ts = excl_prod_scan(__fsub_rn(1.0f, a_s), cTs, lane);
                        const float Tt = excl_prod_scan(__fsub_rn(1.0f, a_t), cTt, lane);
                        const float w1 = a_s * Ts, w2 = a_t * Tt;
                        xsr += w1 * cr; xsg += w1 * cg; xsb += w1 * cb; xsd += w1 * zi;
                        xtr += w2 * qr; xtg += w2 * qg; xtb += w2 * qb; xtd += w2 * zi;
                    }
                } else {
                    sr += w * cr; sg += w * cg; sb += w * cb;
                }
            }
        }
        opac = warp_sum(opac);
        if (lane == 0 && o.opacity) o.opacity[ray] = opac;
        if (weights_only) continue;
        depth = warp_sum(depth);
        sr = warp_sum(sr); sg = warp_sum(sg); sb = warp_sum(sb);
        const float bg = white ? (1.0f - opac) : 0.0f;
        if (transient) {
            tr = warp_sum(tr); tg = warp_sum(tg); tb = warp_sum(tb);
            beta = warp_sum(beta);
            if (test_time) {
                xsr = warp_sum(xsr); xsg = warp_sum(xsg); xsb = warp_sum(xsb); xsd = warp_sum(xsd);
                xtr = warp_sum(xtr); xtg = warp_sum(xtg); xtb = warp_sum(xtb); xtd = warp_sum(xtd);
            }
            if (lane == 0) {
                const float mr = sr + bg, mg = sg + bg, mb = sb + bg;
                if (o.rgb_static) { o.rgb_static[ray * 3 + 0] = mr; o.rgb_static[ray * 3 + 1] = mg; o.rgb_static[ray * 3 + 2] = mb; }
                if (o.rgb_transient) { o.rgb_transient[ray * 3 + 0] = tr; o.rgb_transient[ray * 3 + 1] = tg; o.rgb_transient[ray * 3 + 2] = tb; }
                if (o.rgb) { o.rgb[ray * 3 + 0] = mr + tr; o.rgb[ray * 3 + 1] = mg + tg; o.rgb[ray * 3 + 2] = mb + tb; }
                if (o.beta) o.beta[ray] = beta + beta_min;
                if (o.depth) o.depth[ray] = depth;
                if (test_time) {
                    // static-only map keeps the COMBINED opacity in its white-back term (:203-204)
                    if (o.tt_rgb_static) { o.tt_rgb_static[ray * 3 + 0] = xsr + bg; o.tt_rgb_static[ray * 3 + 1] = xsg + bg; o.tt_rgb_static[ray * 3 + 2] = xsb + bg; }
                    if (o.tt_depth_static) o.tt_depth_static[ray] = xsd;
                    if (o.tt_rgb_transient) { o.tt_rgb_transient[ray * 3 + 0] = xtr; o.tt_rgb_transient[ray * 3 + 1] = xtg; o.tt_rgb_transient[ray * 3 + 2] = xtb; }
                    if (o.tt_depth_transient) o.tt_depth_transient[ray] = xtd;
                }
            }
        } else if (lane == 0) {
            if (o.rgb) { o.rgb[ray * 3 + 0] = sr + bg; o.rgb[ray * 3 + 1] = sg + bg; o.rgb[ray * 3 + 2] = sb + bg; }
            if (o.depth) o.depth[ray] = depth;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// backward.  With G_j the cotangent reaching each weight kind and q_j = G w + Gs ws + Gt wt,
//   dL/dsigma^x_j = delta_j * [ T_j (G_j e_j + G^x_j e^x_j) - sum_{k>j} q_k ]     
// pass 1 (forward order): recompute T, stash T_j in smem; pass 2 (reverse order): suffix scan.
// dynamic smem per warp: kRing chunk stages (cp.async ring) + S floats of T.
// ------------------------------------------------------------------------------------------------
template <int C, bool kTr>
__global__ void __launch_bounds__(kCompWarps * 32)
composite_bwd_kernel(const float* __restrict__ raw, const float* __restrict__ z,
                     const float* __restrict__ noise, float noise_std, int n_rays, int S,
                     uint32_t flags, nwb_composite_grads g, float* __restrict__ d_raw, const LossSeedDev seed,
                     unsigned int* __restrict__ amax_bits) {
    extern __shared__ __align__(16) float dsm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int Sp = (S + 3) & ~3;
    // seeded: the cotangents of rgb / beta / transient_sigmas under NerfWLoss are formed here from the maps themselves
    // (objective.cuh) and ADDED to whatever g carries for them
    const bool seeded = seed.tgt != nullptr;
    LossScalars sc{};
    float seed_ts = 0.0f;
    if (seeded) {
        sc = loss_scalars(seed.g[0], seed.g[1], seed.g[2], seed.g[3], seed.g[4], seed.coef, seed.n);
        if (seed.fine && seed.beta != nullptr && seed.n_ts > 0) seed_ts = loss_d_tsigma(sc, seed.lambda_u, seed.n_ts);
    }
    float amax = 0.0f;                                       // max |d_raw| this lane wrote (amax_bits != nullptr)
    constexpr int kStageFloats = sizeof(RingStage<C>) / sizeof(float);
    float* wbase = dsm + warp * (kRing * kStageFloats + Sp);
    RingStage<C>* my = reinterpret_cast<RingStage<C>*>(wbase);
    float* Tbuf = wbase + kRing * kStageFloats;
    constexpr bool transient = (C == 9) && kTr;
    const bool white = flags & NWB_COMP_WHITE_BACK;
    // The warp's chunk sequence per ray: chunks 0..n-1 (pass 1), then n-1..0 (pass 2), then the next ray: one cp.async
    // ring runs through all of it, kRing-1 chunks ahead of the consumer (the re-read of a chunk hits L1/L2).
    const int n_chunks = (S + 31) >> 5;
    const int stride = gridDim.x * kCompWarps;
    int pf_ray = blockIdx.x * kCompWarps + warp, pf_k = 0;
    uint32_t iss = 0, it = 0;
    auto issue_next = [&]() {
        const int c = pf_k < n_chunks ? pf_k : 2 * n_chunks - 1 - pf_k;
        ring_issue<C>(&my[iss % kRing], raw, z, pf_ray, n_rays, c * 32, S, lane);
        ++iss;
        if (++pf_k == 2 * n_chunks) { pf_k = 0; pf_ray += stride; }
    };
    // consumer side: refill the stage freed one iteration ago, then wait for the oldest outstanding chunk
    auto acquire = [&]() -> RingStage<C>* {
        issue_next();
        cp_async_wait<kRing - 1>();
        __syncwarp();
        return &my[(it++) % kRing];
    };
#pragma unroll
    for (int k = 0; k < kRing - 1; ++k) issue_next();

    for (int ray = blockIdx.x * kCompWarps + warp; ray < n_rays; ray += stride) {
        const int64_t row = static_cast<int64_t>(ray) * S;
        // per-ray cotangents
        const float g_op = g.d_opacity ? __ldg(g.d_opacity + ray) : 0.0f;
        const float g_dp = g.d_depth ? __ldg(g.d_depth + ray) : 0.0f;
        float g_be = (transient && g.d_beta) ? __ldg(g.d_beta + ray) : 0.0f;
        float seed_rgb[3] = {0.0f, 0.0f, 0.0f};
        if (seeded) {
            const bool has_beta = seed.fine && seed.beta != nullptr;
            const float b = has_beta ? __ldg(seed.beta + ray) : 1.0f;
            const float ib2 = loss_inv_beta2(b);
            float err = 0.0f;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float t = __ldg(seed.tgt + ray * 3 + c), x = __ldg(seed.rgb + ray * 3 + c);
                if (seed.fine) {
                    const float df = __fsub_rn(x, t);
                    seed_rgb[c] = loss_d_rgb_fine(sc, df, ib2, has_beta);
                    err = __fmaf_rn(df, df, err);
                } else {
                    seed_rgb[c] = loss_d_rgb_coarse(sc, x, t);
                }
            }
            if (transient && has_beta) g_be += loss_d_beta(sc, err, ib2, b);
        }
        float Gs[3], Gt[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float gr = (g.d_rgb ? __ldg(g.d_rgb + ray * 3 + c) : 0.0f) + seed_rgb[c];
            Gs[c] = gr + ((transient && g.d_rgb_static) ? __ldg(g.d_rgb_static + ray * 3 + c) : 0.0f);
            Gt[c] = gr + ((transient && g.d_rgb_transient) ? __ldg(g.d_rgb_transient + ray * 3 + c) : 0.0f);
        }
        const float g_bg = white ? -(Gs[0] + Gs[1] + Gs[2]) : 0.0f;   // d(1 - sum w)/dw

        // pass 1: transmittance
        const int last_base = ((S - 1) / 32) * 32;
        float cT = 1.0f;
        for (int base = 0; base < S; base += 32) {
            const RingStage<C>* st = acquire();
            const int i = base + lane;
            const bool live = i < S;
            float alpha = 0.0f;
            if (live) {
                const float zi = st->z[lane];
                const float delta = (i + 1 < S) ? __fsub_rn(st->z[lane + 1], zi) : 1e2f;
                float s = (C == 1) ? st->raw[lane] : st->raw[lane * C + 3];
                if (transient) s = __fadd_rn(s, st->raw[lane * C + 7]);
                else {
                    if (noise != nullptr) s = __fadd_rn(s, __fmul_rn(__ldg(noise + row + i), noise_std));
                    s = fmaxf(s, 0.0f);
                }
                alpha = __fsub_rn(1.0f, expf(-__fmul_rn(delta, s)));
            }
            __syncwarp();                                   // every lane is done with the stage before it is refilled
            const float T = excl_prod_scan(__fsub_rn(1.0f, alpha), cT, lane);
            if (live) Tbuf[i] = T;
        }
        __syncwarp();

        // pass 2: reverse chunks, suffix sums of q
        float suffix = 0.0f;   // sum of q over all samples after the current chunk
        for (int base = last_base; base >= 0; base -= 32) {
            const int cnt = min(32, S - base);
            RingStage<C>* st = acquire();
            float* slab = st->raw;                          // overwritten in place with the chunk's d_raw rows below
            const float zi_r = st->z[lane], zn_r = st->z[lane + 1];
            const int i = base + lane;
            const bool live = lane < cnt;
            float q = 0.0f, dsig_s = 0.0f, dsig_t = 0.0f, dcs[3] = {0, 0, 0}, dct[3] = {0, 0, 0}, dbeta = 0.0f;
            float delta = 0, T = 0, head_s = 0, head_t = 0;
            bool pass_s = true;
            if (live) {
                const float* v = slab + lane * C;
                const float zi = zi_r;
                delta = (i + 1 < S) ? __fsub_rn(zn_r, zi) : 1e2f;
                T = Tbuf[i];
                float Gw = g_op + g_dp * zi + g_bg + (g.d_weights ? __ldg(g.d_weights + row + i) : 0.0f);
                if (transient) {
                    const float ss = v[3], st = v[7];
                    const float es = expf(-__fmul_rn(delta, ss)), et = expf(-__fmul_rn(delta, st));
                    const float e = expf(-__fmul_rn(delta, __fadd_rn(ss, st)));
                    const float ws = (1.0f - es) * T, wt = (1.0f - et) * T, w = (1.0f - e) * T;
                    const float Gws = Gs[0] * v[0] + Gs[1] * v[1] + Gs[2] * v[2];
                    const float Gwt = Gt[0] * v[4] + Gt[1] * v[5] + Gt[2] * v[6] + g_be * v[8];
                    q = Gw * w + Gws * ws + Gwt * wt;
                    head_s = T * (Gw * e + Gws * es);
                    head_t = T * (Gw * e + Gwt * et);
#pragma unroll
                    for (int c = 0; c < 3; ++c) { dcs[c] = ws * Gs[c]; dct[c] = wt * Gt[c]; }
                    dbeta = wt * g_be;
                } else {
                    float s = (C == 1) ? v[0] : v[3];
                    if (noise != nullptr) s = __fadd_rn(s, __fmul_rn(__ldg(noise + row + i), noise_std));
                    pass_s = s > 0.0f;
                    const float e = expf(-__fmul_rn(delta, fmaxf(s, 0.0f)));
                    const float w = (1.0f - e) * T;
                    if (C != 1) {
                        Gw += Gs[0] * v[0] + Gs[1] * v[1] + Gs[2] * v[2];
#pragma unroll
                        for (int c = 0; c < 3; ++c) dcs[c] = w * Gs[c];
                    }
                    q = Gw * w;
                    head_s = T * Gw * e;
                }
            }
            __syncwarp();
            // exclusive suffix sum within the chunk (lanes above me) + carried suffix
            float incl = q;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const float t = __shfl_down_sync(0xffffffffu, incl, o);
                if (lane + o < 32) incl += t;
            }
            const float R = suffix + (incl - q);
            suffix += __shfl_sync(0xffffffffu, incl, 0);
            if (live) {
                dsig_s = pass_s ? delta * (head_s - R) : 0.0f;
                float* v = slab + lane * C;
                if (C == 1) {
                    v[0] = dsig_s;
                } else {
                    v[0] = dcs[0]; v[1] = dcs[1]; v[2] = dcs[2]; v[3] = dsig_s;
                }
                if (C == 9) {
                    if (transient) {
                        dsig_t = delta * (head_t - R) + ((g.d_transient_sigmas ? __ldg(g.d_transient_sigmas + row + i) : 0.0f) + seed_ts);
                        v[4] = dct[0]; v[5] = dct[1]; v[6] = dct[2]; v[7] = dsig_t; v[8] = dbeta;
                    } else {
                        v[4] = v[5] = v[6] = v[7] = v[8] = 0.0f;
                    }
                }
                if (amax_bits != nullptr) {
#pragma unroll
                    for (int c = 0; c < C; ++c) amax = fmaxf(amax, fabsf(v[c]));       // fmaxf drops NaN operands
                }
            }
            __syncwarp();
            // coalesced write-back of the chunk's d_raw rows
            float* dst = d_raw + (row + base) * C;
            const int nfl = cnt * C;
            if ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0 && (nfl & 3) == 0) {
                for (int k = lane; k < (nfl >> 2); k += 32)
                    reinterpret_cast<float4*>(dst)[k] = reinterpret_cast<const float4*>(slab)[k];
            } else {
                for (int k = lane; k < nfl; k += 32) dst[k] = slab[k];
            }
            __syncwarp();
        }
    }
    if (amax_bits != nullptr) {
        // max |d_raw| of the pass for the field backward's 8-bit scale (what dz_amax_kernel finds in a launch of its own):
        // non-negative floats order like their bit patterns, one atomicMax per block
        __shared__ float amax_red[kCompWarps];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        if (lane == 0) amax_red[warp] = amax;
        __syncthreads();
        if (threadIdx.x == 0) {
            float m = amax_red[0];
#pragma unroll
            for (int w = 1; w < kCompWarps; ++w) m = fmaxf(m, amax_red[w]);
            if (m > 0.0f && m < INFINITY) atomicMax(amax_bits, __float_as_uint(m));
        }
    }
}


// ------------------------------------------------------------------------------------------------
// The same backward with the WHOLE ray resident in shared memory (S <= 32*NCH samples, NCH chunks of 32): one round of
// cp.async per ray, no second read for the reverse pass, and the chunks' scans run as NCH independent chains instead of
// one after the other -- the carries (transmittance in, suffix sum out) only enter a final multiply / add, exactly where
// the kernel above applies them, so every value comes out bit for bit the same.  At the training batch (one ray per warp,
// 7 warps per SM) the kernel above is a latency chain of 2*S/32 chunk steps: [B200, tools/graph_timeline.py] 16 us for
// 1024 x 128 x 9 at the head of the backward's critical path.
// ------------------------------------------------------------------------------------------------
template <int C, bool kTr, int NCH>
__global__ void __launch_bounds__(kCompWarps * 32)
composite_bwd_resident_kernel(const float* __restrict__ raw, const float* __restrict__ z,
                              const float* __restrict__ noise, float noise_std, int n_rays, int S,
                              uint32_t flags, nwb_composite_grads g, float* __restrict__ d_raw, const LossSeedDev seed,
                              unsigned int* __restrict__ amax_bits) {
    static_assert(C == 4 || C == 9, "sigma-only passes have no backward");
    extern __shared__ __align__(16) float dsm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    RingStage<C>* my = reinterpret_cast<RingStage<C>*>(dsm) + warp * NCH;
    constexpr bool transient = (C == 9) && kTr;
    const bool white = flags & NWB_COMP_WHITE_BACK;
    const bool seeded = seed.tgt != nullptr;
    LossScalars sc{};
    float seed_ts = 0.0f;
    if (seeded) {
        sc = loss_scalars(seed.g[0], seed.g[1], seed.g[2], seed.g[3], seed.g[4], seed.coef, seed.n);
        if (seed.fine && seed.beta != nullptr && seed.n_ts > 0) seed_ts = loss_d_tsigma(sc, seed.lambda_u, seed.n_ts);
    }
    float amax = 0.0f;
    const int n_chunks = (S + 31) >> 5;                      // <= NCH (host)
    const int stride = gridDim.x * kCompWarps;
    for (int ray = blockIdx.x * kCompWarps + warp; ray < n_rays; ray += stride) {
        const int64_t row = static_cast<int64_t>(ray) * S;
#pragma unroll
        for (int c = 0; c < NCH; ++c)
            if (c < n_chunks) ring_issue<C>(&my[c], raw, z, ray, n_rays, c * 32, S, lane);
        // per-ray cotangents (their loads overlap the copies)
        const float g_op = g.d_opacity ? __ldg(g.d_opacity + ray) : 0.0f;
        const float g_dp = g.d_depth ? __ldg(g.d_depth + ray) : 0.0f;
        float g_be = (transient && g.d_beta) ? __ldg(g.d_beta + ray) : 0.0f;
        float seed_rgb[3] = {0.0f, 0.0f, 0.0f};
        if (seeded) {
            const bool has_beta = seed.fine && seed.beta != nullptr;
            const float b = has_beta ? __ldg(seed.beta + ray) : 1.0f;
            const float ib2 = loss_inv_beta2(b);
            float err = 0.0f;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float t = __ldg(seed.tgt + ray * 3 + c), x = __ldg(seed.rgb + ray * 3 + c);
                if (seed.fine) {
                    const float df = __fsub_rn(x, t);
                    seed_rgb[c] = loss_d_rgb_fine(sc, df, ib2, has_beta);
                    err = __fmaf_rn(df, df, err);
                } else {
                    seed_rgb[c] = loss_d_rgb_coarse(sc, x, t);
                }
            }
            if (transient && has_beta) g_be += loss_d_beta(sc, err, ib2, b);
        }
        float Gs[3], Gt[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float gr = (g.d_rgb ? __ldg(g.d_rgb + ray * 3 + c) : 0.0f) + seed_rgb[c];
            Gs[c] = gr + ((transient && g.d_rgb_static) ? __ldg(g.d_rgb_static + ray * 3 + c) : 0.0f);
            Gt[c] = gr + ((transient && g.d_rgb_transient) ? __ldg(g.d_rgb_transient + ray * 3 + c) : 0.0f);
        }
        const float g_bg = white ? -(Gs[0] + Gs[1] + Gs[2]) : 0.0f;   // d(1 - sum w)/dw
        cp_async_wait<0>();
        __syncwarp();

        // pass 1: transmittance -- NCH local scans side by side, then the carries in chunk order
        float T[NCH];
        {
            float p[NCH];
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                float alpha = 0.0f;
                const int i = c * 32 + lane;
                if (c < n_chunks && i < S) {
                    const RingStage<C>* st = &my[c];
                    const float zi = st->z[lane];
                    const float delta = (i + 1 < S) ? __fsub_rn(st->z[lane + 1], zi) : 1e2f;
                    float sg = st->raw[lane * C + 3];
                    if (transient) sg = __fadd_rn(sg, st->raw[lane * C + 7]);
                    else {
                        if (noise != nullptr) sg = __fadd_rn(sg, __fmul_rn(__ldg(noise + row + i), noise_std));
                        sg = fmaxf(sg, 0.0f);
                    }
                    alpha = __fsub_rn(1.0f, expf(-__fmul_rn(delta, sg)));
                }
                p[c] = __fsub_rn(1.0f, alpha);
            }
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    const float t = __shfl_up_sync(0xffffffffu, p[c], o);
                    if (lane >= o) p[c] *= t;
                }
            }
            float carry = 1.0f;
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                float ex = __shfl_up_sync(0xffffffffu, p[c], 1);
                if (lane == 0) ex = 1.0f;
                T[c] = carry * ex;
                carry *= __shfl_sync(0xffffffffu, p[c], 31);
            }
        }
        // pass 2a: per chunk (independent) everything that does not need the suffix sum: the finished columns go into the
        // slab in place of the raw values, head_s / head_t wait in the sigma columns
        float q[NCH];
        uint32_t pass_mask = 0;                                // bit c: sigma_c > 0 (non-transient clamp)
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            q[c] = 0.0f;
            const int i = c * 32 + lane;
            if (c < n_chunks && i < S) {
                RingStage<C>* st = &my[c];
                float* v = st->raw + lane * C;
                const float zi = st->z[lane];
                const float delta = (i + 1 < S) ? __fsub_rn(st->z[lane + 1], zi) : 1e2f;
                const float Tc = T[c];
                float Gw = g_op + g_dp * zi + g_bg + (g.d_weights ? __ldg(g.d_weights + row + i) : 0.0f);
                if (transient) {
                    const float ss = v[3], stt = v[7];
                    const float es = expf(-__fmul_rn(delta, ss)), et = expf(-__fmul_rn(delta, stt));
                    const float e = expf(-__fmul_rn(delta, __fadd_rn(ss, stt)));
                    const float ws = (1.0f - es) * Tc, wt = (1.0f - et) * Tc, w = (1.0f - e) * Tc;
                    const float Gws = Gs[0] * v[0] + Gs[1] * v[1] + Gs[2] * v[2];
                    const float Gwt = Gt[0] * v[4] + Gt[1] * v[5] + Gt[2] * v[6] + g_be * v[8];
                    q[c] = Gw * w + Gws * ws + Gwt * wt;
                    const float head_s = Tc * (Gw * e + Gws * es);
                    const float head_t = Tc * (Gw * e + Gwt * et);
                    v[0] = ws * Gs[0]; v[1] = ws * Gs[1]; v[2] = ws * Gs[2]; v[3] = head_s;
                    v[4] = wt * Gt[0]; v[5] = wt * Gt[1]; v[6] = wt * Gt[2]; v[7] = head_t; v[8] = wt * g_be;
                    pass_mask |= 1u << c;
                } else {
                    float sg = v[3];
                    if (noise != nullptr) sg = __fadd_rn(sg, __fmul_rn(__ldg(noise + row + i), noise_std));
                    if (sg > 0.0f) pass_mask |= 1u << c;
                    const float e = expf(-__fmul_rn(delta, fmaxf(sg, 0.0f)));
                    const float w = (1.0f - e) * Tc;
                    Gw += Gs[0] * v[0] + Gs[1] * v[1] + Gs[2] * v[2];
                    q[c] = Gw * w;
                    v[0] = w * Gs[0]; v[1] = w * Gs[1]; v[2] = w * Gs[2]; v[3] = Tc * Gw * e;
                    if (C == 9) v[4] = v[5] = v[6] = v[7] = v[8] = 0.0f;
                }
            }
        }
        // NCH local inclusive suffix scans side by side, then the carried suffix in reverse chunk order
        float R[NCH];
        {
            float incl[NCH];
#pragma unroll
            for (int c = 0; c < NCH; ++c) incl[c] = q[c];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    const float t = __shfl_down_sync(0xffffffffu, incl[c], o);
                    if (lane + o < 32) incl[c] += t;
                }
            }
            float suffix = 0.0f;
#pragma unroll
            for (int c = NCH - 1; c >= 0; --c) {
                if (c < n_chunks) {
                    R[c] = suffix + (incl[c] - q[c]);
                    suffix += __shfl_sync(0xffffffffu, incl[c], 0);
                } else {
                    R[c] = 0.0f;
                }
            }
        }
        // pass 2b: the two columns that needed it, then the coalesced write-back of each chunk's rows
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            if (c < n_chunks) {
                const int base = c * 32, i = base + lane, cnt = min(32, S - base);
                RingStage<C>* st = &my[c];
                if (lane < cnt) {
                    float* v = st->raw + lane * C;
                    const float delta = (i + 1 < S) ? __fsub_rn(st->z[lane + 1], st->z[lane]) : 1e2f;
                    v[3] = ((pass_mask >> c) & 1u) ? delta * (v[3] - R[c]) : 0.0f;
                    if (transient)
                        v[7] = delta * (v[7] - R[c]) + ((g.d_transient_sigmas ? __ldg(g.d_transient_sigmas + row + i) : 0.0f) + seed_ts);
                    if (amax_bits != nullptr) {
#pragma unroll
                        for (int k = 0; k < C; ++k) amax = fmaxf(amax, fabsf(v[k]));
                    }
                }
                __syncwarp();
                float* dst = d_raw + (row + base) * C;
                const float* slab = st->raw;
                const int nfl = cnt * C;
                if ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0 && (nfl & 3) == 0) {
                    for (int k = lane; k < (nfl >> 2); k += 32)
                        reinterpret_cast<float4*>(dst)[k] = reinterpret_cast<const float4*>(slab)[k];
                } else {
                    for (int k = lane; k < nfl; k += 32) dst[k] = slab[k];
                }
            }
        }
        __syncwarp();                                          // the stages are refilled for the warp's next ray
    }
    if (amax_bits != nullptr) {
        __shared__ float amax_red[kCompWarps];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        if (lane == 0) amax_red[warp] = amax;
        __syncthreads();
        if (threadIdx.x == 0) {
            float m = amax_red[0];
#pragma unroll
            for (int w = 1; w < kCompWarps; ++w) m = fmaxf(m, amax_red[w]);
            if (m > 0.0f && m < INFINITY) atomicMax(amax_bits, __float_as_uint(m));
        }
    }
}

}  // namespace nwb

using namespace nwb;

extern "C" int nwb_composite_fwd(const float* raw, int C, const float* z, const float* noise,
                                 float noise_std, int n_rays, int S, uint32_t flags,
                                 float beta_min, const nwb_composite_out* out, void* stream) {
    NWB_REQUIRE(n_rays == 0 || (raw && z && out), "nwb_composite_fwd: null pointer");
    NWB_REQUIRE(C == 1 || C == 4 || C == 9, "nwb_composite_fwd: C must be 1, 4 or 9 (got %d)", C);
    NWB_REQUIRE(S >= 1 && n_rays >= 0, "nwb_composite_fwd: bad shape");
    NWB_REQUIRE(!(flags & NWB_COMP_TRANSIENT) || C == 9, "nwb_composite_fwd: transient needs C==9");
    if (n_rays == 0) return NWB_OK;
    const int blocks = std::min((n_rays + kCompWarps - 1) / kCompWarps, kNumSMs * 16);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool tr = (flags & NWB_COMP_TRANSIENT) != 0, tt = (flags & NWB_COMP_TEST_TIME) != 0;
    const bool wo = tt && (flags & NWB_COMP_COARSE);
#define NWB_FWD(CC, VV) composite_fwd_kernel<CC, VV><<<blocks, kCompWarps * 32, 0, st>>>(raw, z, noise, noise_std, n_rays, S, flags, beta_min, *out)
    if (C == 1) NWB_FWD(1, 0u);
    else if (C == 4) { if (wo) NWB_FWD(4, 4u); else NWB_FWD(4, 0u); }
    else if (wo) NWB_FWD(9, 4u);
    else if (!tr) NWB_FWD(9, 0u);
    else if (tt) NWB_FWD(9, 3u);
    else NWB_FWD(9, 1u);
#undef NWB_FWD
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

// which compositing backward runs: 0 (default) = the whole-ray-resident kernel for S <= 128, 1 = always the ring kernel,
// 2 = the resident kernel wherever it fits (S <= 256).  Process-wide; tests walk all three.
static int g_comp_bwd_gen = 0;
extern "C" int nwb_composite_bwd_select(int gen) {
    NWB_REQUIRE(gen >= 0 && gen <= 2, "nwb_composite_bwd_select: gen must be 0, 1 or 2");
    g_comp_bwd_gen = gen;
    return NWB_OK;
}

static int composite_bwd_launch(const float* raw, int C, const float* z, const float* noise, float noise_std, int n_rays, int S,
                                uint32_t flags, const nwb_composite_grads* g, const nwb_loss_seed* seed,
                                unsigned int* amax_bits, float* d_raw, void* stream, const char* who) {
    NWB_REQUIRE(n_rays == 0 || (raw && z && d_raw), "%s: null pointer", who);
    NWB_REQUIRE(C == 1 || C == 4 || C == 9, "%s: C must be 1, 4 or 9 (got %d)", who, C);
    NWB_REQUIRE(S >= 1 && n_rays >= 0, "%s: bad shape", who);
    NWB_REQUIRE(!(flags & NWB_COMP_TEST_TIME), "%s: test_time outputs carry no gradient", who);
    NWB_REQUIRE(!(flags & NWB_COMP_TRANSIENT) || C == 9, "%s: transient needs C==9", who);
    LossSeedDev sd{};
    if (seed != nullptr) {
        NWB_REQUIRE(seed->rgb && seed->targets && seed->n_rays >= 1, "%s: seed needs rgb, targets and the batch size", who);
        NWB_REQUIRE(C != 1, "%s: a sigma-only pass has no rgb to seed", who);
        NWB_REQUIRE(!seed->beta || (seed->fine && (flags & NWB_COMP_TRANSIENT)), "%s: seed beta needs the fine transient pass", who);
        sd.rgb = seed->rgb; sd.beta = seed->beta; sd.tgt = seed->targets;
        for (int i = 0; i < 5; ++i) sd.g[i] = seed->g_terms[i];
        sd.coef = seed->coef; sd.lambda_u = seed->lambda_u; sd.n = seed->n_rays; sd.n_ts = seed->n_ts; sd.fine = seed->fine;
    }
    if (n_rays == 0) return NWB_OK;
    const nwb_composite_grads none{};
    if (g == nullptr) g = &none;
    const int blocks = std::min((n_rays + kCompWarps - 1) / kCompWarps, kNumSMs * 16);
    const size_t smem = static_cast<size_t>(kCompWarps) * (kRing * (32 * C + 36) + ((S + 3) & ~3)) * sizeof(float);
    NWB_REQUIRE(smem <= 200 * 1024, "%s: S too large for shared memory", who);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (amax_bits != nullptr && !(flags & NWB_COMP_AMAX_CLEARED)) NWB_CUDA(cudaMemsetAsync(amax_bits, 0, sizeof(unsigned int), st));
    // the kernel that keeps the whole ray in shared memory (2, 4 or 8 chunks of 32 samples).  [B200, tools/perf_hbm.py, 16384
    // rays] C = 4, S = 128: 26.7 us against the ring kernel's 32.8; C = 9, S = 256 (8 chunks, 41 KB per block): 92 against 84 --
    // so by default up to 4 chunks only; gen 2 forces it wherever it fits (tests)
    if (C != 1 && S <= (g_comp_bwd_gen == 2 ? 256 : 128) && g_comp_bwd_gen != 1) {
        const int nch = S <= 64 ? 2 : (S <= 128 ? 4 : 8);
        const size_t rsmem = static_cast<size_t>(kCompWarps) * nch * (32 * C + 36) * sizeof(float);
        const bool tr = (flags & NWB_COMP_TRANSIENT) != 0;
#define NWB_RES(CC, TR, NN) composite_bwd_resident_kernel<CC, TR, NN><<<blocks, kCompWarps * 32, rsmem, st>>>( \
            raw, z, noise, noise_std, n_rays, S, flags, *g, d_raw, sd, amax_bits)
#define NWB_RES_N(CC, TR) do { if (nch == 2) NWB_RES(CC, TR, 2); else if (nch == 4) NWB_RES(CC, TR, 4); else NWB_RES(CC, TR, 8); } while (0)
        if (C == 4) NWB_RES_N(4, false);
        else if (tr) NWB_RES_N(9, true);
        else NWB_RES_N(9, false);
#undef NWB_RES_N
#undef NWB_RES
        NWB_LAUNCH_CHECK();
        return NWB_OK;
    }
#define NWB_BWD(CC, TR)                                                                             \
    do {                                                                                            \
        if (smem > 48 * 1024)                                                                       \
            NWB_CUDA(cudaFuncSetAttribute(composite_bwd_kernel<CC, TR>,                             \
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        composite_bwd_kernel<CC, TR><<<blocks, kCompWarps * 32, smem, st>>>(raw, z, noise, noise_std, \
                                                                            n_rays, S, flags, *g, d_raw, sd, amax_bits); \
    } while (0)
    if (C == 1) NWB_BWD(1, false);
    else if (C == 4) NWB_BWD(4, false);
    else if (flags & NWB_COMP_TRANSIENT) NWB_BWD(9, true);
    else NWB_BWD(9, false);
#undef NWB_BWD
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

extern "C" int nwb_composite_bwd(const float* raw, int C, const float* z, const float* noise,
                                 float noise_std, int n_rays, int S, uint32_t flags,
                                 const nwb_composite_grads* g, float* d_raw, void* stream) {
    NWB_REQUIRE(n_rays == 0 || g, "nwb_composite_bwd: null pointer");
    return composite_bwd_launch(raw, C, z, noise, noise_std, n_rays, S, flags, g, nullptr, nullptr, d_raw, stream,
                                "nwb_composite_bwd");
}

extern "C" int nwb_composite_bwd_seeded(const float* raw, int C, const float* z, const float* noise,
                                        float noise_std, int n_rays, int S, uint32_t flags,
                                        const nwb_composite_grads* g, const nwb_loss_seed* seed,
                                        unsigned int* amax_bits, float* d_raw, void* stream) {
    return composite_bwd_launch(raw, C, z, noise, noise_std, n_rays, S, flags, g, seed, amax_bits, d_raw, stream,
                                "nwb_composite_bwd_seeded");
}
