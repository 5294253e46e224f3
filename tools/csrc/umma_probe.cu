// Micro-probe of the tcgen05 issue path (tools/probe_umma.py only): how many cycles does ONE thread need to issue
// tcgen05.mma / tcgen05.commit, as a function of the MMA's N, and how long do the MMAs take to drain?
#include "common.cuh"
#include "ptx.cuh"

namespace nwb {

// out[0] = cycles to issue `rounds` x (4 MMAs [+ commit]); out[1] = cycles until the final commit completes
__global__ void __launch_bounds__(128, 1)
umma_probe_kernel(int N, int rounds, int commit_every, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar[2];
    __shared__ uint32_t tmem_slot;
    if (threadIdx.x == 0) {
        ptx::mbar_init(&bar[0], 1);
        ptx::mbar_init(&bar[1], 1);
        ptx::fence_mbar_init();
    }
    for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    ptx::fence_proxy_async_smem();
    if (threadIdx.x < 32) ptx::tmem_alloc<256>(&tmem_slot);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = ptx::umma_idesc_bf16(128, N, false, false);
        const uint32_t a_lo = ptx::umma_desc_lo(ptx::smem_u32(smem)), b_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + 16384));
        const uint32_t junk = ptx::smem_u32(&bar[1]);
        const long long t0 = clock64();
        for (int r = 0; r < rounds; ++r) {
            ptx::umma_bf16_lo(tmem, a_lo, b_lo, idesc, 1);
            ptx::umma_bf16_lo(tmem, a_lo + 2, b_lo + 2, idesc, 1);
            ptx::umma_bf16_lo(tmem, a_lo + 4, b_lo + 4, idesc, 1);
            ptx::umma_bf16_lo(tmem, a_lo + 6, b_lo + 6, idesc, 1);
            if (commit_every && (r % commit_every) == commit_every - 1) ptx::umma_commit_a(junk);
        }
        const long long t1 = clock64();
        ptx::umma_commit(&bar[0]);
        ptx::mbar_wait(&bar[0], 0);
        const long long t2 = clock64();
        out[0] = t1 - t0;
        out[1] = t2 - t0;
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (threadIdx.x < 32) ptx::tmem_dealloc<256>(tmem);
}

}  // namespace nwb

extern "C" int nwb_debug_umma_probe(int N, int rounds, int commit_every, long long* out, void* stream) {
    using namespace nwb;
    const int smem = 16384 + 32768;
    NWB_CUDA(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    umma_probe_kernel<<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(N, rounds, commit_every, out);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
