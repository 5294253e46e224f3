// Packed operand cache of one NeRF field: bf16 SWIZZLE_128B weight slab images (one per GEMM
// layer, see field.cuh) followed by an fp32 block of biases and head weights.  Rebuilt from the
// fp32 nn.Linear parameters (state_dict layout) after every optimizer step.
#include "field_pack.cuh"

namespace nwb {

__global__ void __launch_bounds__(256)
field_pack_kernel(PackArgs a, uint8_t* __restrict__ packed) {
    const FieldShape& fs = a.fs;
    const int g = blockIdx.y;
    const int N = fs.packed_n(g), K = fs.layer_kb(g) * 64;
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(packed + fs.w_off(g));
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < N * K; idx += gridDim.x * blockDim.x) {
        const int n = idx / K, k = idx - n * K;
        const float* src = pack_ptr(a, g, n, k);
        dst[pack_dst_off(N, n, k) >> 1] = __float2bfloat16_rn(src ? *src : 0.0f);
    }
    // fp32 constants (block (0,0) only)
    if (g == 0 && blockIdx.x == 0) {
        float* c = reinterpret_cast<float*>(packed + fs.w_bytes());
        for (int i = threadIdx.x; i < ConstOff::count; i += blockDim.x) {
            const float* src = const_ptr(a, i);
            c[i] = src ? *src : 0.0f;
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
    rc = check_pack_params(fs, p, packed, "nwb_field_pack");
    if (rc != NWB_OK) return rc;
    PackArgs a{fs, *p};
    dim3 grid(64, fs.n_packed_layers());
    field_pack_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(a, static_cast<uint8_t*>(packed));
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
