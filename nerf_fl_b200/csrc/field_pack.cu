// Packed operand cache of one NeRF field: bf16 SWIZZLE_128B weight slab images (one per GEMM
// layer, see field.cuh) followed by an fp32 block of biases and head weights.  Rebuilt from the
// fp32 nn.Linear parameters (state_dict layout) after every optimizer step.
#include "field.cuh"
#include "ptx.cuh"

namespace nwb {

struct PackArgs {
    FieldShape fs;
    nwb_field_params p;
};

__device__ __forceinline__ float pack_src(const PackArgs& a, int g, int n, int k) {
    const FieldShape& fs = a.fs;
    if (g == 0) return (k < fs.in_xyz) ? a.p.trunk_w[0][n * fs.in_xyz + k] : 0.0f;
    if (g == 4) {
        const int ld = fs.in_xyz + kW;
        if (k < 64 * fs.kx) return (k < fs.in_xyz) ? a.p.trunk_w[4][n * ld + k] : 0.0f;
        return a.p.trunk_w[4][n * ld + fs.in_xyz + (k - 64 * fs.kx)];
    }
    if (g <= 7) return a.p.trunk_w[g][n * kW + k];
    if (g == 8) return a.p.final_w[n * kW + k];
    if (g == 9) {
        if (n < kHalf) return a.p.dir_w[n * (kW + fs.in_dir + fs.in_a) + k];
        return a.p.tenc_w[0][(n - kHalf) * (kW + fs.in_t) + k];
    }
    return a.p.tenc_w[g - 9][n * kHalf + k];
}

__global__ void __launch_bounds__(256)
field_pack_kernel(PackArgs a, uint8_t* __restrict__ packed) {
    const FieldShape& fs = a.fs;
    const int g = blockIdx.y;
    const int N = fs.packed_n(g), K = fs.layer_kb(g) * 64;
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(packed + fs.w_off(g));
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < N * K; idx += gridDim.x * blockDim.x) {
        const int n = idx / K, k = idx - n * K;
        const int kb = k >> 6, chunk = (k & 63) >> 3;
        const size_t off = static_cast<size_t>(kb) * N * 128 + ptx::image_chunk_off(n, chunk) + (k & 7) * 2;
        dst[off >> 1] = __float2bfloat16_rn(pack_src(a, g, n, k));
    }
    // fp32 constants (block (0,0) only)
    if (g == 0 && blockIdx.x == 0) {
        float* c = reinterpret_cast<float*>(packed + fs.w_bytes());
        for (int i = threadIdx.x; i < ConstOff::count; i += blockDim.x) {
            float v = 0.0f;
            if (i < ConstOff::final_b) v = a.p.trunk_b[i >> 8][i & 255];
            else if (i < ConstOff::tenc_b) v = a.p.final_b[i - ConstOff::final_b];
            else if (i < ConstOff::sigma_w) { if (fs.has_t()) { const int j = i - ConstOff::tenc_b; v = a.p.tenc_b[1 + (j >> 7)][j & 127]; } }
            else if (i < ConstOff::rgb_w) v = a.p.sigma_w[i - ConstOff::sigma_w];
            else if (i < ConstOff::tsigma_w) v = a.p.rgb_w[i - ConstOff::rgb_w];
            else if (i < ConstOff::trgb_w) { if (fs.has_t()) v = a.p.tsigma_w[i - ConstOff::tsigma_w]; }
            else if (i < ConstOff::tbeta_w) { if (fs.has_t()) v = a.p.trgb_w[i - ConstOff::trgb_w]; }
            else if (i < ConstOff::head_b) { if (fs.has_t()) v = a.p.tbeta_w[i - ConstOff::tbeta_w]; }
            else {
                const int j = i - ConstOff::head_b;
                if (j == 0) v = a.p.sigma_b[0];
                else if (j < 4) v = a.p.rgb_b[j - 1];
                else if (fs.has_t()) {
                    if (j == 4) v = a.p.tsigma_b[0];
                    else if (j < 8) v = a.p.trgb_b[j - 5];
                    else if (j == 8) v = a.p.tbeta_b[0];
                }
            }
            c[i] = v;
        }
    }
}

}  // namespace nwb

using namespace nwb;

extern "C" size_t nwb_field_pack_bytes(const nwb_field_desc* d) {
    FieldShape fs;
    if (make_shape(d, &fs) != NWB_OK) return 0;
    return fs.w_bytes() + ConstOff::count * sizeof(float);
}

extern "C" int nwb_field_pack(const nwb_field_desc* d, const nwb_field_params* p, void* packed, void* stream) {
    FieldShape fs;
    int rc = make_shape(d, &fs);
    if (rc != NWB_OK) return rc;
    NWB_REQUIRE(p && packed, "nwb_field_pack: null pointer");
    NWB_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 1023u) == 0, "nwb_field_pack: packed buffer must be 1 KB aligned");
    for (int i = 0; i < 8; ++i) NWB_REQUIRE(p->trunk_w[i] && p->trunk_b[i], "nwb_field_pack: trunk parameter %d is null", i);
    NWB_REQUIRE(p->final_w && p->final_b && p->dir_w && p->dir_b && p->sigma_w && p->sigma_b && p->rgb_w && p->rgb_b,
                "nwb_field_pack: static head parameter is null");
    if (fs.has_t()) {
        for (int i = 0; i < 4; ++i) NWB_REQUIRE(p->tenc_w[i] && p->tenc_b[i], "nwb_field_pack: transient parameter %d is null", i);
        NWB_REQUIRE(p->tsigma_w && p->tsigma_b && p->trgb_w && p->trgb_b && p->tbeta_w && p->tbeta_b,
                    "nwb_field_pack: transient head parameter is null");
    }
    PackArgs a{fs, *p};
    dim3 grid(64, fs.n_packed_layers());
    field_pack_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(a, static_cast<uint8_t*>(packed));
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
