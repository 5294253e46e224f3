// Device helpers shared by the field forward / backward kernels.
#pragma once
#include "field.cuh"
#include "ptx.cuh"

namespace nwb {

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float softplus_f(float x) { return x > 20.0f ? x : log1pf(expf(x)); }   // nn.Softplus default
__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + expf(-x)); }
// named barrier of the 128 epilogue threads (barrier 0 is __syncthreads)
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// 32 fp32 values of one row -> bf16 -> 4 x 16 B chunks of a SWIZZLE_128B tile image at column c_out
__device__ __forceinline__ void store_slice_bf16(uint8_t* img, int row, int c_out, const float (&x)[32]) {
    uint8_t* blk = img + (c_out >> 6) * 16384;
    const int chunk0 = (c_out & 63) >> 3;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        uint4 pk;
        pk.x = pack_bf16(x[8 * q], x[8 * q + 1]);
        pk.y = pack_bf16(x[8 * q + 2], x[8 * q + 3]);
        pk.z = pack_bf16(x[8 * q + 4], x[8 * q + 5]);
        pk.w = pack_bf16(x[8 * q + 6], x[8 * q + 7]);
        *reinterpret_cast<uint4*>(blk + ptx::image_chunk_off(row, chunk0 + q)) = pk;
    }
}

}  // namespace nwb
