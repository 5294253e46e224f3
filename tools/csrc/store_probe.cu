// Standalone probe: sustained global-store throughput per SM, all SMs busy -- registers (st.global.v8.b32, 32 B per thread,
// a warp = 1 KB contiguous) vs the TMA engine (cp.async.bulk.global.shared::cta of 16 / 32 KB from shared memory).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I nerf_fl_b200/csrc -I include -o tools/store_probe tools/csrc/store_probe.cu
#include <cstdio>
#include "ptx.cuh"
using namespace nwb;

__global__ void __launch_bounds__(512, 1) reg_store(uint8_t* out, int iters, size_t region, long long* cyc) {
    uint8_t* base = out + static_cast<size_t>(blockIdx.x) * region;
    const uint4 a = make_uint4(threadIdx.x, 1, 2, 3), b = make_uint4(4, 5, 6, threadIdx.x);
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        uint8_t* p = base + (static_cast<size_t>(it) * 16384) % region + threadIdx.x * 32;
        ptx::st_global_256(p, a, b);
    }
    __syncthreads();
    if (threadIdx.x == 0) cyc[blockIdx.x] = clock64() - t0;
}

__global__ void __launch_bounds__(128, 1) tma_store(uint8_t* out, int iters, size_t region, int bytes, long long* cyc) {
    extern __shared__ __align__(1024) uint8_t smem[];
    for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = i;
    ptx::fence_proxy_async_smem();
    __syncthreads();
    uint8_t* base = out + static_cast<size_t>(blockIdx.x) * region;
    if (threadIdx.x == 0) {
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            ptx::bulk_s2g(base + (static_cast<size_t>(it) * bytes) % region, smem + (it & 1) * 32768, bytes);
            ptx::bulk_commit();
            ptx::bulk_wait_read<1>();          // two copies in flight
        }
        ptx::bulk_wait_all<0>();
        cyc[blockIdx.x] = clock64() - t0;
    }
}

int main() {
    const size_t region = 8u << 20;
    uint8_t* out; long long* cyc;
    cudaMalloc(&out, region * 148); cudaMalloc(&cyc, 148 * 8);
    long long h[148];
    auto report = [&](const char* name, double bytes_per_cta) {
        cudaDeviceSynchronize();
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        long long mx = 0; for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
        printf("%-44s %6.1f B/clk/SM  (%.2f TB/s at 1.9 GHz over 148 SMs)\n", name, bytes_per_cta / mx, bytes_per_cta / mx * 148 * 1.9e9 / 1e12);
    };
    const int iters = 4000;
    reg_store<<<148, 512>>>(out, iters, region, cyc);
    report("st.global.v8 from registers, 512 threads", double(iters) * 16384);
    cudaFuncSetAttribute(tma_store, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    for (int bytes : {4096, 16384, 32768}) {
        const int n = iters * 16384 / bytes;
        tma_store<<<148, 128, 65536>>>(out, n, region, bytes, cyc);
        char nm[64]; snprintf(nm, sizeof nm, "cp.async.bulk s2g, %d KB copies, 2 in flight", bytes >> 10);
        report(nm, double(n) * bytes);
    }
    // one SM alone
    reg_store<<<1, 512>>>(out, iters, region, cyc);
    cudaDeviceSynchronize(); cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s %6.1f B/clk\n", "registers, ONE SM alone", double(iters) * 16384 / h[0]);
    tma_store<<<1, 128, 65536>>>(out, iters / 2, region, 32768, cyc);
    cudaDeviceSynchronize(); cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s %6.1f B/clk\n", "bulk 32 KB, ONE SM alone", double(iters / 2) * 32768 / h[0]);
    return 0;
}
