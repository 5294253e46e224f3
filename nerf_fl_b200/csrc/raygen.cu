// (f3) on-device ray generation and (f4) the eval output pack -- the data formats either side of render_rays.
//   nwb_gen_rays     datasets/ray_utils.py:5-55 (get_ray_directions + get_rays) and the (N,8) row blender.py:97-100
//                    assembles: [rays_o(3), rays_d(3), near, far].  A pixel range of one view becomes rays inside
//                    the kernel, so eval.py:190's 32 B/ray host->device copy disappears (48 B of pose instead).
//   nwb_pack_rgb8    eval.py:201-203: clip(rgb, 0, 1) * 255 -> uint8 (numpy's astype truncates), done on the
//                    device so 3 B/ray cross PCIe instead of 12.
// Both are streaming kernels: 32 B written per ray / 15 B moved per ray; HBM-bound, no reuse.
#include "common.cuh"
#include <algorithm>

namespace nwb {

struct RayGenArgs {
    int H, W;
    float fx, fy, cx, cy;
    float c2w[12];                 // (3,4) row-major camera-to-world
    float near_, far_;
    long long first, n;
};

__global__ void __launch_bounds__(256)
gen_rays_kernel(const RayGenArgs a, float4* __restrict__ rays) {
    const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
    for (long long r = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; r < a.n; r += nth) {
        const long long pix = a.first + r;
        const int j = static_cast<int>(pix / a.W), i = static_cast<int>(pix - static_cast<long long>(j) * a.W);
        // ray_utils.py:23-24, no +0.5 pixel centring
        const float dx = __fdiv_rn(__fsub_rn(static_cast<float>(i), a.cx), a.fx);
        const float dy = __fdiv_rn(-__fsub_rn(static_cast<float>(j), a.cy), a.fy);
        const float dz = -1.0f;
        // ray_utils.py:46: directions @ c2w[:, :3]^T
        float d[3];
#pragma unroll
        for (int k = 0; k < 3; ++k)
            d[k] = __fmaf_rn(dz, a.c2w[4 * k + 2], __fmaf_rn(dy, a.c2w[4 * k + 1], __fmul_rn(dx, a.c2w[4 * k + 0])));
        // :48 rays_d / norm(rays_d)
        const float nrm = __fsqrt_rn(__fmaf_rn(d[2], d[2], __fmaf_rn(d[1], d[1], __fmul_rn(d[0], d[0]))));
        d[0] = __fdiv_rn(d[0], nrm); d[1] = __fdiv_rn(d[1], nrm); d[2] = __fdiv_rn(d[2], nrm);
        rays[2 * r] = make_float4(a.c2w[3], a.c2w[7], a.c2w[11], d[0]);
        rays[2 * r + 1] = make_float4(d[1], d[2], a.near_, a.far_);
    }
}

__device__ __forceinline__ unsigned int to_u8(float x) {
    // np.clip(x, 0, 1) * 255 in fp32, then numpy's float -> uint8 conversion (truncation); NaN -> 0
    const float c = fminf(fmaxf(x, 0.0f), 1.0f);
    return static_cast<unsigned int>(__float2int_rz(__fmul_rn(c, 255.0f)));
}

__global__ void __launch_bounds__(256)
pack_rgb8_kernel(const float* __restrict__ x, long long n, unsigned char* __restrict__ out) {
    const long long n4 = n >> 2;
    const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
    const long long tid = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    for (long long q = tid; q < n4; q += nth) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(x) + q);
        reinterpret_cast<unsigned int*>(out)[q] = to_u8(v.x) | (to_u8(v.y) << 8) | (to_u8(v.z) << 16) | (to_u8(v.w) << 24);
    }
    for (long long i = 4 * n4 + tid; i < n; i += nth) out[i] = static_cast<unsigned char>(to_u8(__ldg(x + i)));
}

}  // namespace nwb

using namespace nwb;

static int stream_grid(long long work_items) {
    return static_cast<int>(std::max<long long>(1, std::min<long long>(kNumSMs * 8, (work_items + 255) / 256)));
}

extern "C" int nwb_gen_rays(int H, int W, float fx, float fy, float cx, float cy, const float* c2w_host, float near_,
                            float far_, int64_t first_pixel, int64_t n_rays, float* rays, void* stream) {
    NWB_REQUIRE(H >= 1 && W >= 1 && fx != 0.0f && fy != 0.0f, "nwb_gen_rays: bad image size / focal length");
    NWB_REQUIRE(c2w_host != nullptr, "nwb_gen_rays: c2w is null");
    NWB_REQUIRE(first_pixel >= 0 && n_rays >= 0 && first_pixel + n_rays <= static_cast<int64_t>(H) * W,
                "nwb_gen_rays: pixel range [%lld, %lld) outside the %d x %d image", static_cast<long long>(first_pixel),
                static_cast<long long>(first_pixel + n_rays), H, W);
    if (n_rays == 0) return NWB_OK;
    NWB_REQUIRE(rays != nullptr && (reinterpret_cast<uintptr_t>(rays) & 15u) == 0, "nwb_gen_rays: rays must be 16 B aligned");
    RayGenArgs a{};
    a.H = H; a.W = W; a.fx = fx; a.fy = fy; a.cx = cx; a.cy = cy; a.near_ = near_; a.far_ = far_;
    a.first = first_pixel; a.n = n_rays;
    for (int i = 0; i < 12; ++i) a.c2w[i] = c2w_host[i];
    gen_rays_kernel<<<stream_grid(n_rays), 256, 0, static_cast<cudaStream_t>(stream)>>>(a, reinterpret_cast<float4*>(rays));
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

extern "C" int nwb_pack_rgb8(const float* x, int64_t n, unsigned char* out, void* stream) {
    NWB_REQUIRE(n >= 0, "nwb_pack_rgb8: bad size");
    if (n == 0) return NWB_OK;
    NWB_REQUIRE(x && out, "nwb_pack_rgb8: null pointer");
    NWB_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15u) == 0 && (reinterpret_cast<uintptr_t>(out) & 3u) == 0,
                "nwb_pack_rgb8: x must be 16 B and out 4 B aligned");
    pack_rgb8_kernel<<<stream_grid(n / 4 + 1), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, n, out);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
