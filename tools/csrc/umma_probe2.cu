// Interference probe (tools/probe_umma2.py only): how fast does the tensor pipe drain back-to-back
// tcgen05.mma M=128 N=256 K=16 while other warps of the CTA do what the field kernels' other roles do --
// st.shared traffic, fence.proxy.async, tcgen05.ld, bulk (TMA-engine) copies from L2 into shared memory?
#include "common.cuh"
#include "ptx.cuh"

namespace nwb {

constexpr int kP2Threads = 128 + 512;
constexpr int kP2OffA = 0;                 // 16 KB A tile
constexpr int kP2OffB = 16384;             // 32 KB B tile
constexpr int kP2OffS = 49152;             // 64 KB scratch the interfering warps write
constexpr int kP2OffR = 49152 + 65536;     // 3 x 32 KB ring for bulk copies
constexpr int kP2Smem = kP2OffR + 3 * 32768;

// mode bits: 1 = st.shared.v4 (16 B/thread/iter), 2 = + fence.proxy.async per iter, 4 = bulk g2s stream (warp 1),
//            8 = tcgen05.ld x16 per iter, 16 = pace the interfering threads (~epilogue rate: one iter per ~40 cycles),
//            32 / 64 = issue every 32 KB ring stage as 8 x 4 KB / 2 x 16 KB bulk copies
__global__ void __launch_bounds__(kP2Threads, 1)
umma_probe2_kernel(int rounds, int mode, const uint8_t* gsrc, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar[8];
    __shared__ uint32_t tmem_slot;
    __shared__ volatile int done;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int i = 0; i < 8; ++i) ptx::mbar_init(&bar[i], 1);
        ptx::fence_mbar_init();
        done = 0;
    }
    for (int i = threadIdx.x; i < kP2OffS / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    ptx::fence_proxy_async_smem();
    if (warp == 0) ptx::tmem_alloc<512>(&tmem_slot);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = tmem_slot;
    if (threadIdx.x == 0) {
        constexpr uint32_t idesc = ptx::umma_idesc_bf16(128, 256, false, false);
        const uint32_t a_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + kP2OffA)), b_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + kP2OffB));
        const long long t0 = clock64();
        for (int r = 0; r < rounds; ++r) {
            ptx::umma_bf16_lo(tmem, a_lo, b_lo, idesc, 1);
            ptx::umma_bf16_lo(tmem, a_lo + 2, b_lo + 2, idesc, 1);
            ptx::umma_bf16_lo(tmem, a_lo + 4, b_lo + 4, idesc, 1);
            ptx::umma_bf16_lo(tmem, a_lo + 6, b_lo + 6, idesc, 1);
        }
        ptx::umma_commit(&bar[0]);
        ptx::mbar_wait(&bar[0], 0);
        const long long t1 = clock64();
        done = 1;
        out[blockIdx.x] = t1 - t0;
    } else if (warp == 1) {
        if ((mode & 4) && lane == 0) {                 // stream 32 KB bulk copies from L2 into a 3-slot ring
            uint32_t it = 0;
            unsigned long long bytes = 0;
            while (!done) {
                const uint32_t s = it % 3;
                if (it >= 3) ptx::mbar_wait(&bar[1 + s], ((it / 3) - 1) & 1u);
                ptx::mbar_arrive_expect_tx(&bar[1 + s], 32768);
                const int pieces = (mode & 32) ? 8 : ((mode & 64) ? 2 : 1);      // 32 KB as 1 x 32 KB, 2 x 16 KB or 8 x 4 KB copies
                const uint32_t pb = 32768u / pieces;
                for (int pc = 0; pc < pieces; ++pc)
                    ptx::bulk_g2s(smem + kP2OffR + s * 32768 + pc * pb, gsrc + (static_cast<size_t>(it) % 40) * 32768 + pc * pb, pb, &bar[1 + s]);
                ++it;
                bytes += 32768;
            }
            for (uint32_t k = (it >= 3 ? it - 3 : 0); k < it; ++k) ptx::mbar_wait(&bar[1 + k % 3], (k / 3) & 1u);
            out[gridDim.x + blockIdx.x] = static_cast<long long>(bytes);
        }
    } else if (warp >= 4) {
        const int t = threadIdx.x - 128;
        const int row = t & 127, sub = t >> 7;
        uint8_t* dst = smem + kP2OffS;
        uint32_t v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = t + j;
        const uint32_t tl = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16) + 256 + sub * 16;
        int it = 0;
        while (!done) {
            if (mode & 8) { ptx::tmem_ld16(tl + (it & 3) * 64, v); ptx::tmem_ld_wait(); }
            if (mode & 1) {
                uint8_t* blk = dst + (it & 3) * 16384;
                *reinterpret_cast<uint4*>(blk + ptx::image_chunk_off(row, 2 * sub)) = make_uint4(v[0], v[1], v[2], v[3] + it);
                *reinterpret_cast<uint4*>(blk + ptx::image_chunk_off(row, 2 * sub + 1)) = make_uint4(v[4], v[5], v[6], v[7]);
            }
            if (mode & 2) ptx::fence_proxy_async_smem();
            if (mode & 16) __nanosleep(100);
            ++it;
        }
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tmem);
}

}  // namespace nwb

extern "C" int nwb_debug_umma_probe2(int grid, int rounds, int mode, const void* gsrc, long long* out, void* stream) {
    using namespace nwb;
    NWB_CUDA(cudaFuncSetAttribute(umma_probe2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kP2Smem));
    umma_probe2_kernel<<<grid, kP2Threads, kP2Smem, static_cast<cudaStream_t>(stream)>>>(rounds, mode, static_cast<const uint8_t*>(gsrc), out);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
