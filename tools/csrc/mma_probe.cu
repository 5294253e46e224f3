// Standalone probe: drain rate of back-to-back tcgen05.mma (M=128, N=256, K=16, bf16, cta_group::1) with the B operand
// K-major (forward) vs MN-major (the dgrad chain reads the forward's weight slabs transposed) vs both operands MN-major.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I nerf_fl_b200/csrc -I include -o tools/mma_probe tools/csrc/mma_probe.cu
#include <cstdio>
#include "ptx.cuh"
using namespace nwb;

__global__ void __launch_bounds__(128, 1) probe(int rounds, int variant, int n_cols, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
    for (int i = threadIdx.x; i < 98304 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    ptx::fence_proxy_async_smem();
    if (warp == 0) ptx::tmem_alloc<512>(&tmem_slot);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (threadIdx.x == 0) {
        const uint32_t a0 = ptx::smem_u32(smem), b0 = a0 + 32768;
        const bool a_mn = variant == 2, b_mn = variant >= 1;
        const uint32_t idesc = ptx::umma_idesc_bf16(128, n_cols, a_mn, b_mn);
        const long long t0 = clock64();
        for (int r = 0; r < rounds; ++r) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                // K-major: step K by 32 B inside the 128 B row; MN-major: 16 contraction rows = 2 groups of 8 rows, 2 KB apart
                const uint64_t ad = a_mn ? ptx::umma_desc_sw128(a0 + j * 2048, 16384, 1024) : ptx::umma_desc_sw128(a0 + j * 32, 16, 1024);
                const uint64_t bd = b_mn ? ptx::umma_desc_sw128(b0 + j * 2048, 8192, 1024) : ptx::umma_desc_sw128(b0 + j * 32, 16, 1024);
                ptx::umma_bf16(tmem, ad, bd, idesc, 1);
            }
        }
        ptx::umma_commit(&bar);
        ptx::mbar_wait(&bar, 0);
        out[0] = clock64() - t0;
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tmem);
}

int main() {
    long long* out; cudaMalloc(&out, 8);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 98304 + 1024);
    const char* names[3] = {"A K-major, B K-major (forward)", "A K-major, B MN-major (dgrad)", "A MN-major, B MN-major (wgrad, bf16)"};
    for (int n : {256, 128})
        for (int v = 0; v < 3; ++v) {
            const int rounds = 2000;
            probe<<<1, 128, 98304 + 1024>>>(rounds, v, n, out);
            if (cudaDeviceSynchronize() != cudaSuccess) { printf("failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
            long long h; cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
            printf("M=128 N=%d K=16 %-40s %7.1f cycles per MMA\n", n, names[v], double(h) / (rounds * 4));
        }
    return 0;
}
