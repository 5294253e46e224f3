// Where every element of a field's packed operand cache comes from: shared by the repack kernel
// (field_pack.cu) and by the optimizer step that refreshes the cache in the same pass (optim.cu).
#pragma once
#include "field.cuh"
#include "ptx.cuh"

namespace nwb {

struct PackArgs {
    FieldShape fs;
    nwb_field_params p;
};

// fp32 source of element (n, k) of GEMM layer g's weight image; nullptr = zero padding
__device__ __forceinline__ const float* pack_ptr(const PackArgs& a, int g, int n, int k) {
    const FieldShape& fs = a.fs;
    if (g == 0) return (k < fs.in_xyz) ? a.p.trunk_w[0] + n * fs.in_xyz + k : nullptr;
    if (g == 4) {
        const int ld = fs.in_xyz + kW;
        if (k < 64 * fs.kx) return (k < fs.in_xyz) ? a.p.trunk_w[4] + n * ld + k : nullptr;
        return a.p.trunk_w[4] + n * ld + fs.in_xyz + (k - 64 * fs.kx);
    }
    if (g <= 7) return a.p.trunk_w[g] + n * kW + k;
    if (g == 8) return a.p.final_w + n * kW + k;
    if (g == 9) {
        if (n < kHalf) return a.p.dir_w + n * (kW + fs.in_dir + fs.in_a) + k;
        return a.p.tenc_w[0] + (n - kHalf) * (kW + fs.in_t) + k;
    }
    return a.p.tenc_w[g - 9] + n * kHalf + k;
}

// byte offset of element (n, k) inside layer g's image (N rows, SWIZZLE_128B k-blocks of 64 columns)
__device__ __forceinline__ size_t pack_dst_off(int N, int n, int k) {
    const int kb = k >> 6, chunk = (k & 63) >> 3;
    return static_cast<size_t>(kb) * N * 128 + ptx::image_chunk_off(n, chunk) + (k & 7) * 2;
}

// fp32 source of slot i of the constant block (ConstOff); nullptr = unused slot (0)
__device__ __forceinline__ const float* const_ptr(const PackArgs& a, int i) {
    const FieldShape& fs = a.fs;
    const bool t = fs.has_t();
    if (i < ConstOff::final_b) return a.p.trunk_b[i >> 8] + (i & 255);
    if (i < ConstOff::tenc_b) return a.p.final_b + (i - ConstOff::final_b);
    if (i < ConstOff::sigma_w) { const int j = i - ConstOff::tenc_b; return t ? a.p.tenc_b[1 + (j >> 7)] + (j & 127) : nullptr; }
    if (i < ConstOff::rgb_w) return a.p.sigma_w + (i - ConstOff::sigma_w);
    if (i < ConstOff::tsigma_w) return a.p.rgb_w + (i - ConstOff::rgb_w);
    if (i < ConstOff::trgb_w) return t ? a.p.tsigma_w + (i - ConstOff::tsigma_w) : nullptr;
    if (i < ConstOff::tbeta_w) return t ? a.p.trgb_w + (i - ConstOff::trgb_w) : nullptr;
    if (i < ConstOff::head_b) return t ? a.p.tbeta_w + (i - ConstOff::tbeta_w) : nullptr;
    const int j = i - ConstOff::head_b;
    if (j == 0) return a.p.sigma_b;
    if (j < 4) return a.p.rgb_b + (j - 1);
    if (!t) return nullptr;
    if (j == 4) return a.p.tsigma_b;
    if (j < 8) return a.p.trgb_b + (j - 5);
    if (j == 8) return a.p.tbeta_b;
    return nullptr;
}

// host-side validation shared by nwb_field_pack and nwb_adam_pack_step_dev
inline int check_pack_params(const FieldShape& fs, const nwb_field_params* p, const void* packed, const char* who) {
    NWB_REQUIRE(p && packed, "%s: null pointer", who);
    NWB_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 1023u) == 0, "%s: packed buffer must be 1 KB aligned", who);
    for (int i = 0; i < 8; ++i) NWB_REQUIRE(p->trunk_w[i] && p->trunk_b[i], "%s: trunk parameter %d is null", who, i);
    NWB_REQUIRE(p->final_w && p->final_b && p->dir_w && p->dir_b && p->sigma_w && p->sigma_b && p->rgb_w && p->rgb_b,
                "%s: static head parameter is null", who);
    if (fs.has_t()) {
        for (int i = 0; i < 4; ++i) NWB_REQUIRE(p->tenc_w[i] && p->tenc_b[i], "%s: transient parameter %d is null", who, i);
        NWB_REQUIRE(p->tsigma_w && p->tsigma_b && p->trgb_w && p->trgb_b && p->tbeta_w && p->tbeta_b,
                    "%s: transient head parameter is null", who);
    }
    return NWB_OK;
}

}  // namespace nwb
