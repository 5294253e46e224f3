// Cotangents of NerfWLoss (losses.py:35-50) with respect to the maps the compositing pass produced -- shared by
// nerfw_loss_bwd_kernel (objective.cu) and by the compositing backward when it seeds its own cotangents
// (composite.cu, nwb_composite_bwd_seeded): one definition, every rounding spelled out, so both give the same bits.
#pragma once
#include "common.cuh"

namespace nwb {

// device view of nwb_loss_seed (include/nerfw_b200.h); tgt == nullptr: not seeded
struct LossSeedDev {
    const float* rgb;
    const float* beta;
    const float* tgt;
    const float* g[5];
    float coef, lambda_u;
    int n;
    long long n_ts;
    int fine;
};

struct LossScalars {
    float gc, gf, gb, gs;      // (cotangent of the term + cotangent of the total) * coef
    float inv3, invn;          // 1 / (3 n), 1 / n
};

__device__ __forceinline__ LossScalars loss_scalars(const float* g0, const float* g1, const float* g2, const float* g3,
                                                    const float* g4, float coef, int n) {
    const float gt = g4 ? __ldg(g4) : 0.0f;
    LossScalars s;
    s.gc = __fmul_rn(__fadd_rn(g0 ? __ldg(g0) : 0.0f, gt), coef);
    s.gf = __fmul_rn(__fadd_rn(g1 ? __ldg(g1) : 0.0f, gt), coef);
    s.gb = __fmul_rn(__fadd_rn(g2 ? __ldg(g2) : 0.0f, gt), coef);
    s.gs = __fmul_rn(__fadd_rn(g3 ? __ldg(g3) : 0.0f, gt), coef);
    s.inv3 = __fdiv_rn(1.0f, static_cast<float>(3ll * n));
    s.invn = __fdiv_rn(1.0f, static_cast<float>(n));
    return s;
}

// d c_l / d rgb_coarse: d/dx [0.5 mean((x - t)^2)]
__device__ __forceinline__ float loss_d_rgb_coarse(const LossScalars& s, float x, float t) {
    return __fmul_rn(__fmul_rn(s.gc, __fsub_rn(x, t)), s.inv3);
}
// 1 / beta^2 (1 without a transient head)
__device__ __forceinline__ float loss_inv_beta2(float b) { return __fdiv_rn(1.0f, __fmul_rn(b, b)); }
// d f_l / d rgb_fine; df = rgb_fine - t
__device__ __forceinline__ float loss_d_rgb_fine(const LossScalars& s, float df, float ib2, bool has_beta) {
    const float v = __fmul_rn(s.gf, df);
    return has_beta ? __fmul_rn(__fmul_rn(v, ib2), s.inv3) : __fmul_rn(v, s.inv3);
}
// d (f_l + b_l) / d beta; err = sum_c df_c^2 accumulated with __fmaf_rn in channel order
//   d/dbeta [ mean(err / (2 beta^2)) ] = -err / beta^3 / (3N);  d/dbeta [ mean(log beta) ] = 1 / (beta N)
__device__ __forceinline__ float loss_d_beta(const LossScalars& s, float err, float ib2, float b) {
    const float a = __fdiv_rn(__fmul_rn(__fmul_rn(-s.gf, err), ib2), b);
    return __fmaf_rn(a, s.inv3, __fdiv_rn(__fmul_rn(s.gb, s.invn), b));
}
// d s_l / d transient_sigmas (the same for every sample)
__device__ __forceinline__ float loss_d_tsigma(const LossScalars& s, float lambda_u, long long n_ts) {
    return __fdiv_rn(__fmul_rn(s.gs, lambda_u), static_cast<float>(n_ts));
}

}  // namespace nwb
