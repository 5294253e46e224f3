// Self-test of the tcgen05 tile engine: D(128,N) = A * B^T out of bf16 tile images, exercising the
// exact primitives the field kernels use (bulk copy -> mbarrier, SW128 descriptors in both
// major-nesses, tcgen05.mma / commit, TMEM alloc + 32x32b loads).  tests/ drives it; descriptor
// parameters can be overridden from the host so an encoding question costs one GPU run, not a build.
#include "common.cuh"
#include "ptx.cuh"

namespace nwb {

struct SelftestParams {
    uint32_t idesc;
    uint32_t a_lbo, a_sbo, a_kstep, a_every, a_kblock;
    uint32_t b_lbo, b_sbo, b_kstep, b_every, b_kblock;
    uint32_t n_ksteps, N, a_bytes, b_bytes, f8, noswz;
};

__global__ void __launch_bounds__(128, 1)
umma_selftest_kernel(const uint8_t* __restrict__ a_img, const uint8_t* __restrict__ b_img,
                     SelftestParams p, float* __restrict__ d_out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar_full, bar_mma;
    __shared__ uint32_t tmem_base_slot;
    uint8_t* sA = smem + ((1024u - (ptx::smem_u32(smem) & 1023u)) & 1023u);   // swizzle atoms need 1 KB alignment
    uint8_t* sB = sA + ((p.a_bytes + 1023u) & ~1023u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        ptx::mbar_init(&bar_full, 1);
        ptx::mbar_init(&bar_mma, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 0) ptx::tmem_alloc<256>(&tmem_base_slot);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = tmem_base_slot;

    if (threadIdx.x == 0) {
        ptx::mbar_arrive_expect_tx(&bar_full, p.a_bytes + p.b_bytes);
        ptx::bulk_g2s(sA, a_img, p.a_bytes, &bar_full);
        ptx::bulk_g2s(sB, b_img, p.b_bytes, &bar_full);
        ptx::mbar_wait(&bar_full, 0);
        ptx::tc_fence_after_sync();
        const uint32_t a0 = ptx::smem_u32(sA), b0 = ptx::smem_u32(sB);
        for (uint32_t k = 0; k < p.n_ksteps; ++k) {
            const uint32_t ao = (k / p.a_every) * p.a_kblock + (k % p.a_every) * p.a_kstep;
            const uint32_t bo = (k / p.b_every) * p.b_kblock + (k % p.b_every) * p.b_kstep;
            if (p.noswz)           // noswz = 1 + descriptor layout code (6 = 32 B swizzle)
                ptx::umma_f8(tmem, ptx::umma_desc_any(a0 + ao, p.a_lbo, p.a_sbo, p.noswz - 1),
                             ptx::umma_desc_any(b0 + bo, p.b_lbo, p.b_sbo, p.noswz - 1), p.idesc, k > 0);
            else if (p.f8)
                ptx::umma_f8(tmem, ptx::umma_desc_sw128(a0 + ao, p.a_lbo, p.a_sbo),
                             ptx::umma_desc_sw128(b0 + bo, p.b_lbo, p.b_sbo), p.idesc, k > 0);
            else
                ptx::umma_bf16(tmem, ptx::umma_desc_sw128(a0 + ao, p.a_lbo, p.a_sbo),
                               ptx::umma_desc_sw128(b0 + bo, p.b_lbo, p.b_sbo), p.idesc, k > 0);
        }
        ptx::umma_commit(&bar_mma);
    }
    __syncwarp();
    ptx::mbar_wait(&bar_mma, 0);
    ptx::tc_fence_after_sync();

    const int row = warp * 32 + lane;
    for (uint32_t c0 = 0; c0 < p.N; c0 += 32) {
        uint32_t v[32];
        ptx::tmem_ld32(tmem + (static_cast<uint32_t>(warp * 32) << 16) + c0, v);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (c0 + j < p.N) d_out[static_cast<size_t>(row) * p.N + c0 + j] = __uint_as_float(v[j]);
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<256>(tmem);
}

}  // namespace nwb

using namespace nwb;

// variant 0: A image of (128 x K), B image of (N x K), both K-major  (forward / dgrad shape)
// variant 1: A image of (K x 128), B image of (K x N), both MN-major (wgrad shape: D = A^T B)
// variant 4: the 8-bit wgrad shape as the field kernels store it (SWIZZLE_32B images, field.cuh)
// variant 3: the same contraction out of SWIZZLE_128B 8-bit images: A image of (K x 128) e5m2, B image of (K x N) e4m3, both MN-major, images cut into
//            128-column blocks of (K rows x 128 B); tcgen05.mma kind::f8f6f4, K = 32 per instruction
// overrides (host pointer, may be NULL): 11 uint32 replacing idesc, a_{lbo,sbo,kstep,every,kblock},
// b_{...}; a zero entry keeps the default.
extern "C" int nwb_selftest_umma(const void* a_img, const void* b_img, int N, int K, int variant,
                                 const uint32_t* overrides, float* d_out, void* stream) {
    NWB_REQUIRE(a_img && b_img && d_out, "nwb_selftest_umma: null pointer");
    NWB_REQUIRE(N >= 16 && N <= 256 && N % 16 == 0, "nwb_selftest_umma: N must be a multiple of 16 in [16,256]");
    NWB_REQUIRE(variant >= 0 && variant <= 4, "nwb_selftest_umma: variant must be 0..4");
    SelftestParams p{};
    p.N = N;
    if (variant == 0) {
        NWB_REQUIRE(K % 64 == 0 && K >= 64 && K <= 256, "nwb_selftest_umma: variant 0 needs K in {64,128,192,256}");
        p.idesc = ptx::umma_idesc_bf16(128, N, false, false);
        p.a_bytes = 128 * K * 2;
        p.b_bytes = N * K * 2;
        p.n_ksteps = K / 16;
        p.a_lbo = 16; p.a_sbo = 1024; p.a_kstep = 32; p.a_every = 4; p.a_kblock = 128 * 128;
        p.b_lbo = 16; p.b_sbo = 1024; p.b_kstep = 32; p.b_every = 4; p.b_kblock = N * 128;
    } else if (variant == 2) {
        // dgrad shape: A image of (128 x K) K-major; B image of (K x N) read MN-major (D = A * B)
        NWB_REQUIRE(K % 64 == 0 && K >= 64 && K <= 256, "nwb_selftest_umma: variant 2 needs K in {64,128,192,256}");
        NWB_REQUIRE(N % 64 == 0, "nwb_selftest_umma: variant 2 needs N multiple of 64");
        p.idesc = ptx::umma_idesc_bf16(128, N, false, true);
        p.a_bytes = 128 * K * 2;
        p.b_bytes = K * N * 2;
        p.n_ksteps = K / 16;
        p.a_lbo = 16; p.a_sbo = 1024; p.a_kstep = 32; p.a_every = 4; p.a_kblock = 128 * 128;
        p.b_lbo = K * 128; p.b_sbo = 1024; p.b_kstep = 2048; p.b_every = 1u << 30; p.b_kblock = 0;
    } else if (variant == 4) {
        // variant 4: the layout the field kernels use for their 8-bit images: both MN-major, SWIZZLE_32B: 32-column groups of (K rows x 32 B), 8-row groups 256 B apart, the two
        // 16-byte chunks of a row swapped when bit 2 of the row is set
        NWB_REQUIRE(K % 32 == 0 && K >= 32 && K <= 128, "nwb_selftest_umma: variant 4 needs K multiple of 32, <=128");
        p.f8 = 1; p.noswz = 1 + 6;
        p.idesc = ptx::umma_idesc_f8(128, N, 1, 0, true, true);
        p.a_bytes = K * 128;
        p.b_bytes = K * N;
        p.n_ksteps = K / 32;
        p.a_lbo = K * 32; p.a_sbo = 256; p.a_kstep = 1024; p.a_every = 1u << 30; p.a_kblock = 0;
        p.b_lbo = K * 32; p.b_sbo = 256; p.b_kstep = 1024; p.b_every = 1u << 30; p.b_kblock = 0;
    } else if (variant == 3) {
        NWB_REQUIRE(K % 32 == 0 && K >= 32 && K <= 128, "nwb_selftest_umma: variant 3 needs K multiple of 32, <=128");
        NWB_REQUIRE(N % 128 == 0, "nwb_selftest_umma: variant 3 needs N multiple of 128");
        p.f8 = 1;
        p.idesc = ptx::umma_idesc_f8(128, N, 1, 0, true, true);
        p.a_bytes = K * 128;              // one 128-column block of (K rows x 128 B)
        p.b_bytes = K * N;
        p.n_ksteps = K / 32;
        p.a_lbo = K * 128; p.a_sbo = 1024; p.a_kstep = 4096; p.a_every = 1u << 30; p.a_kblock = 0;
        p.b_lbo = K * 128; p.b_sbo = 1024; p.b_kstep = 4096; p.b_every = 1u << 30; p.b_kblock = 0;
    } else {
        NWB_REQUIRE(K % 16 == 0 && K >= 16 && K <= 128, "nwb_selftest_umma: variant 1 needs K multiple of 16, <=128");
        NWB_REQUIRE(N % 64 == 0, "nwb_selftest_umma: variant 1 needs N multiple of 64");
        p.idesc = ptx::umma_idesc_bf16(128, N, true, true);
        p.a_bytes = K * 128 * 2;          // 2 column blocks of (K rows x 128 B)
        p.b_bytes = K * N * 2;
        p.n_ksteps = K / 16;
        p.a_lbo = K * 128; p.a_sbo = 1024; p.a_kstep = 2048; p.a_every = 1u << 30; p.a_kblock = 0;
        p.b_lbo = K * 128; p.b_sbo = 1024; p.b_kstep = 2048; p.b_every = 1u << 30; p.b_kblock = 0;
    }
    if (overrides) {
        uint32_t* f = &p.idesc;
        for (int i = 0; i < 11; ++i)
            if (overrides[i]) f[i] = overrides[i];
    }
    const size_t smem = ((p.a_bytes + 1023u) & ~1023u) + p.b_bytes + 1024;
    NWB_REQUIRE(smem <= 227 * 1024, "nwb_selftest_umma: tile does not fit shared memory");
    NWB_CUDA(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  static_cast<int>(smem)));
    umma_selftest_kernel<<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const uint8_t*>(a_img), static_cast<const uint8_t*>(b_img), p, d_out);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
