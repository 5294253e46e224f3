// Standalone probe (no torch): what do the epilogue's building blocks cost on one SM?
//   * tcgen05.ld 32x32b.x32 / .x16 throughput and latency with W warps reading TMEM (is TMEM read bandwidth per SM or
//     per scheduler?), optionally while one thread keeps the tensor pipe busy with back-to-back MMAs
//   * F2FP (fp32 pair -> bf16x2 / e4m3x2 / e5m2x2) issue rate
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I nerf_fl_b200/csrc -o tools/tmem_probe tools/csrc/tmem_probe.cu
// run:   tools/tmem_probe
#include <cstdio>
#include <cstdlib>
#include "ptx.cuh"

using namespace nwb;

// mode 0: x32 loads back to back, one wait at the end of every `batch` loads; mode 1: x16 likewise
// mma = 1: thread 0 of warp W issues MMAs (M=128 N=256 K=16) continuously meanwhile
__global__ void __launch_bounds__(1024, 1)
tmem_ld_probe(int warps, int iters, int batch, int x16, int mma, long long* out, float* sink) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tmem_slot;
    __shared__ __align__(8) uint64_t bar;
    __shared__ volatile int done;
    const int warp = threadIdx.x >> 5;
    __shared__ int finished;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); done = 0; finished = 0; }
    for (int i = threadIdx.x; i < 49152 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    ptx::fence_proxy_async_smem();
    if (warp == 0) ptx::tmem_alloc<512>(&tmem_slot);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = tmem_slot;
    float acc = 0.0f;
    long long dt = 0;
    if (warp < warps) {
        const uint32_t tl = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            for (int b = 0; b < batch; ++b) {
                const uint32_t col = ((it * batch + b) * 32 + (warp >> 2) * 64) & 255;      // stay inside columns 0..255+32
                if (x16) {
                    uint32_t v[16];
                    ptx::tmem_ld16(tl + col, v);
                    if (b == batch - 1) ptx::tmem_ld_wait();
                    acc += __uint_as_float(v[0]) + __uint_as_float(v[15]);
                } else {
                    uint32_t v[32];
                    ptx::tmem_ld32(tl + col, v);
                    if (b == batch - 1) ptx::tmem_ld_wait();
                    acc += __uint_as_float(v[0]) + __uint_as_float(v[31]);
                }
            }
        }
        ptx::tmem_ld_wait();
        dt = clock64() - t0;
    } else if (mma && warp == warps && (threadIdx.x & 31) == 0) {
        constexpr uint32_t idesc = ptx::umma_idesc_bf16(128, 256, false, false);
        const uint32_t a_lo = ptx::umma_desc_lo(ptx::smem_u32(smem)), b_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + 16384));
        uint32_t ph = 0;
        while (!done) {
            for (int r = 0; r < 16; ++r) ptx::umma_bf16_lo(tmem + 256, a_lo + 2 * (r & 3), b_lo + 2 * (r & 3), idesc, 1);
            ptx::umma_commit(&bar);
            ptx::mbar_wait(&bar, ph);     // 16 x 128 tensor cycles per batch, a few hundred cycles of bubble between batches
            ph ^= 1u;
        }
    }
    if (warp < warps) {
        __syncwarp();
        if ((threadIdx.x & 31) == 0) out[warp] = dt;
        sink[threadIdx.x] = acc;
    }
    // all readers done -> stop the MMA thread
    if (warp < warps && (threadIdx.x & 31) == 0) atomicAdd(&finished, 1);
    if (warp == 0 && (threadIdx.x & 31) == 0) { while (atomicAdd(&finished, 0) < warps) { } done = 1; }
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tmem);
}

// F2FP issue rate: each thread converts `iters` x 16 pairs; kind 0 = bf16x2, 1 = e4m3x2, 2 = e5m2x2, 3 = FADD2 (reference)
__global__ void __launch_bounds__(1024, 1)
f2fp_probe(int iters, int kind, long long* out, uint32_t* sink) {
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
    uint32_t acc = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
            uint32_t r;
            if (kind == 0) { __nv_bfloat162 t = __floats2bfloat162_rn(a[i], a[i + 1]); r = *reinterpret_cast<uint32_t*>(&t); }
            else if (kind == 1) r = ptx::cvt_e4m3x2(a[i], a[i + 1]);
            else if (kind == 2) r = ptx::cvt_e5m2x2(a[i], a[i + 1]);
            else { float x, y; ptx::fadd2(x, y, a[i], a[i + 1], 1.0f, 2.0f); r = __float_as_uint(x) ^ __float_as_uint(y); }
            acc ^= r;
            a[i] = __uint_as_float((__float_as_uint(a[i]) & 0x3FFFFFFFu) | (r & 1u));       // data dependence, 1 LOP3
        }
    }
    const long long dt = clock64() - t0;
    if ((threadIdx.x & 31) == 0) out[threadIdx.x >> 5] = dt;
    sink[threadIdx.x] = acc;
}

int main() {
    long long* out; float* sink; uint32_t* sink2;
    cudaMalloc(&out, 64 * sizeof(long long)); cudaMalloc(&sink, 1024 * 4); cudaMalloc(&sink2, 1024 * 4);
    cudaFuncSetAttribute(tmem_ld_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    long long h[64];
    const int iters = 2000;
    printf("tcgen05.ld throughput on one SM (bytes per clock over all reading warps; batch = loads per wait::ld)\n");
    for (int mma = 0; mma < 2; ++mma)
        for (int x16 = 0; x16 < 2; ++x16)
            for (int batch : {1, 2, 4})
                for (int warps : {1, 4, 8, 16}) {
                    tmem_ld_probe<<<1, 32 * (warps + 1), 65536>>>(warps, iters, batch, x16, mma, out, sink);
                    if (cudaDeviceSynchronize() != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
                    cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
                    long long mx = 0;
                    for (int w = 0; w < warps; ++w) mx = h[w] > mx ? h[w] : mx;
                    const double bytes = double(warps) * iters * batch * 32 * (x16 ? 16 : 32) * 4;
                    printf("mma=%d %s batch=%d warps=%2d: %8lld clk  %7.1f B/clk  %6.1f clk per load per warp\n", mma, x16 ? "x16" : "x32", batch, warps, mx,
                           bytes / mx, double(mx) / (iters * batch));
                }
    printf("conversion / add issue rate, 32 warps on one SM (8 ops + 8 LOP3 + 8 LOP3 per thread-iteration)\n");
    for (int kind = 0; kind < 4; ++kind) {
        f2fp_probe<<<1, 1024>>>(4000, kind, out, sink2);
        cudaDeviceSynchronize();
        cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
        long long mx = 0;
        for (int w = 0; w < 32; ++w) mx = h[w] > mx ? h[w] : mx;
        printf("kind %d (%s): %lld clk, %.2f clk per warp-op-triple per scheduler\n", kind, kind == 0 ? "bf16x2" : kind == 1 ? "e4m3x2" : kind == 2 ? "e5m2x2" : "fadd2",
               mx, double(mx) / (4000.0 * 8 * 8));
    }
    return 0;
}
