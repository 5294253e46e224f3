// Standalone probe (no torch): reciprocal issue throughput of the epilogue's ALU / conversion instructions on one SM.
// Every thread runs 16 independent dependency chains of ONE instruction kind (the result feeds the same chain's next
// input), 32 warps per SM = 8 per scheduler -> 128 chains per scheduler: throughput-bound, no other instructions in the loop.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I nerf_fl_b200/csrc -I include -o tools/op_probe tools/csrc/op_probe.cu
#include <cstdio>
#include "ptx.cuh"
using namespace nwb;

template <int kKind>
__device__ __forceinline__ void op(uint32_t& a, uint32_t& b) {
    if (kKind == 0) { asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(a) : "f"(__uint_as_float(a)), "f"(__uint_as_float(b))); }
    else if (kKind == 1) { asm volatile("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(a) : "f"(__uint_as_float(a)), "f"(__uint_as_float(b))); }
    else if (kKind == 2) { uint16_t r; asm volatile("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(__uint_as_float(a)), "f"(__uint_as_float(b))); a = r | 0x3f000000u; }
    else if (kKind == 3) { uint16_t r; asm volatile("cvt.rn.satfinite.relu.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(__uint_as_float(a)), "f"(__uint_as_float(b))); a = r | 0x3f000000u; }
    else if (kKind == 4) { uint16_t r; asm volatile("cvt.rn.satfinite.e5m2x2.f32 %0, %1, %2;" : "=h"(r) : "f"(__uint_as_float(a)), "f"(__uint_as_float(b))); a = r | 0x3f000000u; }
    else if (kKind == 5) { asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(a) : "f"(__uint_as_float(a)), "f"(__uint_as_float(b))); }
    else if (kKind == 6) { float x, y; ptx::fadd2(x, y, __uint_as_float(a), __uint_as_float(b), 1.0f, 2.0f); a = __float_as_uint(x); b = __float_as_uint(y); }
    else if (kKind == 7) { asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(*reinterpret_cast<float*>(&a)) : "f"(__uint_as_float(a)), "f"(__uint_as_float(b))); }
    else if (kKind == 8) { asm volatile("max.bf16x2 %0, %1, %2;" : "=r"(a) : "r"(a), "r"(b)); }
    else if (kKind == 9) { asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(a) : "r"(a), "r"(b), "r"(0x12345u)); }
    else if (kKind == 10) { asm volatile("prmt.b32 %0, %1, %2, 0x5410;" : "=r"(a) : "r"(a), "r"(b)); }
    else if (kKind == 11) { asm volatile("{ .reg .pred p; setp.lt.u32 p, %1, 77; selp.b32 %0, %1, %2, p; }" : "=r"(a) : "r"(a), "r"(b)); }
    else if (kKind == 12) { asm volatile("fma.rn.f32 %0, %1, %2, %1;" : "=f"(*reinterpret_cast<float*>(&a)) : "f"(__uint_as_float(a)), "f"(__uint_as_float(b))); }
    else if (kKind == 13) { asm volatile("shf.r.clamp.b32 %0, %1, %2, 3;" : "=r"(a) : "r"(a), "r"(b)); }
    else if (kKind == 14) { uint16_t r; asm volatile("cvt.rn.satfinite.e5m2x2.f16x2 %0, %1;" : "=h"(r) : "r"(a)); a = r | 0x3c000000u; }
}
// kinds 2-4, 14 carry one extra LOP3 (the or) per op; it is subtracted from the report using kind 9's rate

template <int kKind>
__global__ void __launch_bounds__(1024, 1) probe(int iters, long long* out, uint32_t* sink) {
    uint32_t a[16], b[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) { a[i] = 0x3f800000u + threadIdx.x * 977u + i * 131u; b[i] = 0x3f000000u + i * 17u + threadIdx.x; }
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) op<kKind>(a[i], b[i]);
    }
    const long long dt = clock64() - t0;
    uint32_t acc = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) acc ^= a[i] ^ b[i];
    if ((threadIdx.x & 31) == 0) out[threadIdx.x >> 5] = dt;
    sink[threadIdx.x] = acc;
}

template <int kKind>
double run(const char* name, long long* out, uint32_t* sink, double lop = 0.0) {
    const int iters = 2000;
    probe<kKind><<<1, 1024>>>(iters, out, sink);
    cudaDeviceSynchronize();
    long long h[32];
    cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int w = 0; w < 32; ++w) mx = h[w] > mx ? h[w] : mx;
    const double per = double(mx) / (double(iters) * 16 * 8) - lop;     // 8 warps per scheduler
    printf("%-34s %6.2f clk per warp-instruction per scheduler  (%5.1f lanes/clk/SM)\n", name, per, 128.0 / per);
    return per;
}

int main() {
    long long* out; uint32_t* sink;
    cudaMalloc(&out, 32 * sizeof(long long)); cudaMalloc(&sink, 4096);
    const double lop = run<9>("lop3", out, sink);
    run<7>("add.f32", out, sink);
    run<12>("fma.f32", out, sink);
    run<6>("add.f32x2 (FADD2)", out, sink);
    run<8>("max.bf16x2 (HMNMX2)", out, sink);
    run<10>("prmt", out, sink);
    run<13>("shf", out, sink);
    run<11>("setp+selp (2 instr)", out, sink);
    run<0>("cvt.rn.bf16x2.f32", out, sink);
    run<1>("cvt.rn.relu.bf16x2.f32", out, sink);
    run<5>("cvt.rn.f16x2.f32", out, sink);
    run<2>("cvt.rn.satfinite.e4m3x2.f32", out, sink, lop);
    run<3>("cvt.rn.satfinite.relu.e4m3x2.f32", out, sink, lop);
    run<4>("cvt.rn.satfinite.e5m2x2.f32", out, sink, lop);
    run<14>("cvt.rn.satfinite.e5m2x2.f16x2", out, sink, lop);
    return 0;
}
