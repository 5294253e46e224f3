// (a2,a3,a4) fused NeRF-W field forward -- models/nerf.py:153-212 + the per-sample input assembly of
// models/rendering.py:96-139, as one persistent warp-specialised tcgen05 kernel.
//
// One CTA owns one 128-point tile at a time (two CTAs are co-resident per SM so one CTA's epilogue
// overlaps the other's MMAs).  Roles:
//   warp 0 / lane 0   weight producer: streams the packed bf16 slab images (field_pack.cu) from L2
//                     into a 2-stage ring with linear bulk copies (TMA engine) + mbarrier tx counts
//   warp 1 / lane 0   MMA issuer: tcgen05.mma M=128 N=128 K=16, A = activation tile image in smem,
//                     B = ring stage, D = fp32 accumulators in TMEM (256 columns)
//   warps 2..5        epilogue, thread == point: build the encoded-xyz operand tile (o + d*z,
//                     sin/cos octaves, bf16, written straight into the swizzled A image), then per
//                     layer tcgen05.ld -> +bias (+per-ray bias) -> ReLU -> bf16 -> next layer's A
//                     image in place; sigma / rgb / transient heads on CUDA cores in fp32.
// Activations never leave the SM unless NWB_MLP_SAVE asks for them (bf16 tile images for backward).
#include "field_dev.cuh"
#include <algorithm>

namespace nwb {

constexpr int kFwdThreads = 192;
constexpr int kStageBytes = 16384;            // 128 weight rows x 64 k (one SW128 K-block)
constexpr int kNumStages = 2;
constexpr int kOffX = 0;                      // 16 KB  encoded xyz, 128 x 64
constexpr int kOffH = 16384;                  // 64 KB  activations, 128 x 256 (4 K-blocks)
constexpr int kOffW = kOffH + 65536;          // ring
constexpr int kOffBar = kOffW + kNumStages * kStageBytes;
constexpr int kFwdSmem = kOffBar + 256;

struct FwdArgs {
    FieldShape fs;
    const uint8_t* packed;
    const float* rays;
    int ray_stride;
    const float* z;
    const float* x_enc;
    int n_freq_xyz;
    const float* ray_bias;
    int S;
    long long n_points;
    int n_tiles;
    uint32_t mode;
    float* raw_out;
    uint8_t* saved;
    size_t save_tile_bytes;
};

// One 32-column slice of a layer epilogue for this thread's row.
//   x = acc + bias[c] (+ rbias[c]);  optional ReLU;  optional head dot products (fp32, unrounded x);
//   bf16 pack -> 4 x 16 B chunks of the A image at column c_out (same row).
template <bool kRelu, int kHeads>
__device__ __forceinline__ void epi_slice(uint32_t taddr, const float* __restrict__ bias,
                                          const float* __restrict__ rbias, const float* __restrict__ hw, int hw_ld,
                                          float* hacc, uint8_t* sH, int row, int c_out, uint32_t* mask_dst) {
    uint32_t v[32];
    ptx::tmem_ld32(taddr, v);
    float b[32];
    if (bias != nullptr) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(bias) + q);
            b[4 * q] = t.x; b[4 * q + 1] = t.y; b[4 * q + 2] = t.z; b[4 * q + 3] = t.w;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) b[j] = 0.0f;
    }
    if (rbias != nullptr) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(rbias) + q);
            b[4 * q] += t.x; b[4 * q + 1] += t.y; b[4 * q + 2] += t.z; b[4 * q + 3] += t.w;
        }
    }
    ptx::tmem_ld_wait();
    float x[32];
    uint32_t neg = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        x[j] = __uint_as_float(v[j]) + b[j];
        if (kRelu) {
            neg = __funnelshift_l(__float_as_uint(x[j]), neg, 1);   // bit (31-j) = sign of the pre-activation
            x[j] = fmaxf(x[j], 0.0f);
        }
    }
    if (kRelu && mask_dst != nullptr) *mask_dst = neg;
    if (kHeads > 0) {
#pragma unroll
        for (int h = 0; h < kHeads; ++h) {
            float a = hacc[h];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(hw + h * hw_ld) + q);
                a = fmaf(x[4 * q], t.x, a); a = fmaf(x[4 * q + 1], t.y, a);
                a = fmaf(x[4 * q + 2], t.z, a); a = fmaf(x[4 * q + 3], t.w, a);
            }
            hacc[h] = a;
        }
    }
    store_slice_bf16(sH, row, c_out, x);
}

__global__ void __launch_bounds__(kFwdThreads, 2)
field_fwd_kernel(const FwdArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sX = smem + kOffX;
    uint8_t* sH = smem + kOffH;
    uint8_t* sW = smem + kOffW;
    uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem + kOffBar);   // [kNumStages]
    uint64_t* bar_empty = bar_full + kNumStages;                        // [kNumStages]
    uint64_t* bar_acc = bar_empty + kNumStages;                         // MMA -> epilogue
    uint64_t* bar_act = bar_acc + 1;                                    // epilogue -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_act + 1);
    uint32_t* s_woff = tmem_slot + 2;                                   // [kMaxG] slab offsets

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const FieldShape& fs = a.fs;
    const int nl = fs.n_layers(a.mode);
    const bool transient = (a.mode & NWB_MLP_TRANSIENT) != 0;
    const bool sigma_only = (a.mode & NWB_MLP_SIGMA_ONLY) != 0;
    const bool save = (a.mode & NWB_MLP_SAVE) != 0;

    if (threadIdx.x == 0) {
        if ((ptx::smem_u32(smem) & 1023u) != 0) {
            printf("nwb: dynamic shared memory is not 1 KB aligned\n");
            __trap();
        }
        for (int s = 0; s < kNumStages; ++s) { ptx::mbar_init(&bar_full[s], 1); ptx::mbar_init(&bar_empty[s], 1); }
        ptx::mbar_init(bar_acc, 1);
        ptx::mbar_init(bar_act, 128);
        ptx::fence_mbar_init();
        for (int g = 0; g < kMaxG; ++g) s_woff[g] = static_cast<uint32_t>(fs.w_off(g));
    }
    if (warp == 1) ptx::tmem_alloc<256>(tmem_slot);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ============================ weight producer ============================================
        if (lane == 0) {
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
                for (int g = 0; g < nl; ++g) {
                    const uint8_t* wsrc = a.packed + s_woff[g];
                    const int KB = fs.layer_kb(g), NH = fs.layer_n(g, a.mode) >> 7, PN = fs.packed_n(g);
                    for (int kb = 0; kb < KB; ++kb) {
                        for (int nh = 0; nh < NH; ++nh, ++it) {
                            const uint32_t s = it % kNumStages;
                            ptx::mbar_wait(&bar_empty[s], ((it / kNumStages) & 1u) ^ 1u);
                            ptx::mbar_arrive_expect_tx(&bar_full[s], kStageBytes);
                            ptx::bulk_g2s(sW + s * kStageBytes, wsrc + static_cast<size_t>(kb) * PN * 128 + nh * kStageBytes,
                                          kStageBytes, &bar_full[s]);
                        }
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ============================ MMA issuer =================================================
        if (lane == 0) {
            constexpr uint32_t idesc = ptx::umma_idesc_bf16(128, 128, false, false);
            const uint32_t xa = ptx::smem_u32(sX), ha = ptx::smem_u32(sH), wa = ptx::smem_u32(sW);
            uint32_t it = 0, ev = 0;
            for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
                for (int g = 0; g < nl; ++g) {
                    ptx::mbar_wait(bar_act, ev & 1u);
                    ++ev;
                    ptx::tc_fence_after_sync();
                    const int KB = fs.layer_kb(g), NH = fs.layer_n(g, a.mode) >> 7;
                    for (int kb = 0; kb < KB; ++kb) {
                        uint32_t abase;
                        if (g == 0) abase = xa;
                        else if (g == 4) abase = (kb == 0) ? xa : ha + (kb - 1) * 16384;
                        else if (g >= 10) abase = ha + (kb + 2) * 16384;
                        else abase = ha + kb * 16384;
                        for (int nh = 0; nh < NH; ++nh, ++it) {
                            const uint32_t s = it % kNumStages;
                            ptx::mbar_wait(&bar_full[s], (it / kNumStages) & 1u);
                            ptx::tc_fence_after_sync();
                            const uint32_t wb = wa + s * kStageBytes;
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                ptx::umma_bf16(tmem + nh * 128, ptx::umma_desc_sw128(abase + k * 32, 16, 1024),
                                               ptx::umma_desc_sw128(wb + k * 32, 16, 1024), idesc, (kb | k) != 0);
                            ptx::umma_commit(&bar_empty[s]);
                        }
                    }
                    ptx::umma_commit(bar_acc);
                }
            }
        }
        __syncwarp();
    } else {
        // ============================ epilogue warps (thread == point) ===========================
        const int quad = warp & 3;                         // TMEM lane quadrant this warp may read
        const int row = quad * 32 + lane;
        const uint32_t tlane = tmem + (static_cast<uint32_t>(quad * 32) << 16);
        const bool leader = (threadIdx.x == 64);
        const float* cst = reinterpret_cast<const float*>(a.packed + fs.w_bytes());
        uint32_t acc_ev = 0;
        for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
            const long long p = static_cast<long long>(tile) * kTileM + row;
            const bool valid = p < a.n_points;
            const long long ray = valid ? p / a.S : 0;
            uint8_t* save_tile = save ? a.saved + static_cast<size_t>(tile) * a.save_tile_bytes : nullptr;
            const SaveLayout sl{fs.kx, transient};

            // ---- encoded-xyz operand tile ------------------------------------------------------
            {
                float e[64];
#pragma unroll
                for (int c = 0; c < 64; ++c) e[c] = 0.0f;
                if (valid) {
                    if (a.x_enc != nullptr) {
                        const float* xr = a.x_enc + p * fs.in_xyz;
#pragma unroll
                        for (int c = 0; c < 64; ++c)
                            if (c < fs.in_xyz) e[c] = __ldg(xr + c);
                    } else {
                        const float* rr = a.rays + ray * a.ray_stride;
                        const float zz = __ldg(a.z + p);
                        float sn[3], cs[3];
#pragma unroll
                        for (int i = 0; i < 3; ++i) {
                            e[i] = __fadd_rn(__ldg(rr + i), __fmul_rn(__ldg(rr + 3 + i), zz));   // rendering.py:261
                            sincosf(e[i], &sn[i], &cs[i]);
                        }
                        // octave k+1 from octave k by the double-angle identities: the error grows to
                        // ~2^k ulp (<= 6e-5 at k = 9), far below the bf16 rounding of the operand
#pragma unroll
                        for (int k = 0; k < 10; ++k) {
                            if (k < a.n_freq_xyz) {
#pragma unroll
                                for (int i = 0; i < 3; ++i) {
                                    e[3 + 6 * k + i] = sn[i];
                                    e[6 + 6 * k + i] = cs[i];
                                    const float s2 = 2.0f * sn[i] * cs[i];
                                    const float c2 = fmaf(-2.0f * sn[i], sn[i], 1.0f);
                                    sn[i] = s2; cs[i] = c2;
                                }
                            }
                        }
                    }
                }
                if (save) {                                 // previous tile's last bulk store must have read sH/sX
                    if (leader) ptx::bulk_wait_read<0>();
                    epi_bar();
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    uint4 pk;
                    pk.x = pack_bf16(e[8 * q], e[8 * q + 1]);
                    pk.y = pack_bf16(e[8 * q + 2], e[8 * q + 3]);
                    pk.z = pack_bf16(e[8 * q + 4], e[8 * q + 5]);
                    pk.w = pack_bf16(e[8 * q + 6], e[8 * q + 7]);
                    *reinterpret_cast<uint4*>(sX + ptx::image_chunk_off(row, q)) = pk;
                }
                ptx::fence_proxy_async_smem();
                if (save) {
                    epi_bar();
                    if (leader) { ptx::bulk_s2g(save_tile + sl.x_off(), sX, 16384); ptx::bulk_commit(); }
                }
                ptx::tc_fence_before_sync();
                ptx::mbar_arrive(bar_act);
            }

            float sigma = 0.0f, rgb[3] = {0, 0, 0};
            const float* rb = (a.ray_bias != nullptr) ? a.ray_bias + ray * 256 : nullptr;
            for (int g = 0; g < nl; ++g) {
                ptx::mbar_wait(bar_acc, acc_ev & 1u);
                ++acc_ev;
                ptx::tc_fence_after_sync();
                if (save) {                                 // the bulk store of layer g-1 must have read sH
                    if (leader) ptx::bulk_wait_read<0>();
                    epi_bar();
                }
                float hacc[5] = {0, 0, 0, 0, 0};
                uint32_t* mrow = save ? reinterpret_cast<uint32_t*>(save_tile + sl.mask_off(g, nl)) + row * 8 : nullptr;
                if (g < 7) {
#pragma unroll 1
                    for (int c0 = 0; c0 < 256; c0 += 32)
                        epi_slice<true, 0>(tlane + c0, cst + ConstOff::trunk_b + g * 256 + c0, nullptr, nullptr, 0, hacc, sH, row, c0, mrow ? mrow + (c0 >> 5) : nullptr);
                } else if (g == 7) {
#pragma unroll 1
                    for (int c0 = 0; c0 < 256; c0 += 32)
                        epi_slice<true, 1>(tlane + c0, cst + ConstOff::trunk_b + 7 * 256 + c0, nullptr, cst + ConstOff::sigma_w + c0, 0,
                                           hacc, sH, row, c0, mrow ? mrow + (c0 >> 5) : nullptr);
                    sigma = softplus_f(hacc[0] + __ldg(cst + ConstOff::head_b));                // nerf.py:190
                } else if (g == 8) {
#pragma unroll 1
                    for (int c0 = 0; c0 < 256; c0 += 32)
                        epi_slice<false, 0>(tlane + c0, cst + ConstOff::final_b + c0, nullptr, nullptr, 0, hacc, sH, row, c0, mrow ? mrow + (c0 >> 5) : nullptr);
                } else if (g == 9) {
                    // per-ray bias already contains dir_encoding.0.bias / transient_encoding.0.bias
#pragma unroll 1
                    for (int c0 = 0; c0 < 128; c0 += 32)
                        epi_slice<true, 3>(tlane + c0, nullptr, rb + c0, cst + ConstOff::rgb_w + c0, 128, hacc, sH, row, c0, mrow ? mrow + (c0 >> 5) : nullptr);
#pragma unroll
                    for (int i = 0; i < 3; ++i) rgb[i] = sigmoid_f(hacc[i] + __ldg(cst + ConstOff::head_b + 1 + i));   // nerf.py:197
                    if (transient) {
#pragma unroll 1
                        for (int c0 = 128; c0 < 256; c0 += 32)
                            epi_slice<true, 0>(tlane + c0, nullptr, rb + c0, nullptr, 0, hacc, sH, row, c0, mrow ? mrow + (c0 >> 5) : nullptr);
                    }
                } else if (g < 12) {
#pragma unroll 1
                    for (int c0 = 0; c0 < 128; c0 += 32)
                        epi_slice<true, 0>(tlane + c0, cst + ConstOff::tenc_b + (g - 10) * 128 + c0, nullptr, nullptr, 0, hacc, sH, row, 128 + c0, mrow ? mrow + (c0 >> 5) : nullptr);
                } else {
                    // transient heads: hw rows = [tsigma_w | trgb_w(3) | tbeta_w], contiguous, ld 128
#pragma unroll 1
                    for (int c0 = 0; c0 < 128; c0 += 32)
                        epi_slice<true, 5>(tlane + c0, cst + ConstOff::tenc_b + 2 * 128 + c0, nullptr, cst + ConstOff::tsigma_w + c0, 128,
                                           hacc, sH, row, 128 + c0, mrow ? mrow + (c0 >> 5) : nullptr);
                }
                ptx::fence_proxy_async_smem();
                if (save) {
                    epi_bar();
                    if (leader) {
                        const uint32_t bytes = (g <= 8) ? 65536u : (g == 9 ? (transient ? 65536u : 32768u) : 32768u);
                        const uint8_t* src = (g >= 10) ? sH + 32768 : sH;
                        ptx::bulk_s2g(save_tile + sl.act_off(g), src, bytes);
                        ptx::bulk_commit();
                    }
                }
                // ---- tile outputs ---------------------------------------------------------------
                if (g == nl - 1 && valid) {
                    if (sigma_only) {
                        a.raw_out[p] = sigma;
                    } else {
                        float* o = a.raw_out + p * (transient ? 9 : 4);
                        o[0] = rgb[0]; o[1] = rgb[1]; o[2] = rgb[2]; o[3] = sigma;               // nerf.py:198
                        if (transient) {
                            const float* hb = cst + ConstOff::head_b;
                            o[4] = sigmoid_f(hacc[1] + __ldg(hb + 5));                           // nerf.py:206
                            o[5] = sigmoid_f(hacc[2] + __ldg(hb + 6));
                            o[6] = sigmoid_f(hacc[3] + __ldg(hb + 7));
                            o[7] = softplus_f(hacc[0] + __ldg(hb + 4));                          // nerf.py:205
                            o[8] = softplus_f(hacc[4] + __ldg(hb + 8));                          // nerf.py:207
                        }
                    }
                }
                if (g < nl - 1) {
                    ptx::tc_fence_before_sync();
                    ptx::mbar_arrive(bar_act);
                }
            }
        }
        if (save && leader) ptx::bulk_wait_all<0>();
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<256>(tmem);
}

}  // namespace nwb

using namespace nwb;

extern "C" size_t nwb_mlp_save_bytes(const nwb_field_desc* d, int64_t n_points, uint32_t mode) {
    FieldShape fs;
    if (make_shape(d, &fs) != NWB_OK || n_points <= 0) return 0;
    const SaveLayout sl{fs.kx, (mode & NWB_MLP_TRANSIENT) != 0};
    const int64_t n_tiles = (n_points + kTileM - 1) / kTileM;
    return static_cast<size_t>(n_tiles) * sl.tile_bytes(fs.n_layers(mode));
}

extern "C" int nwb_mlp_fwd(const nwb_field_desc* d, const void* packed, const float* rays, int ray_stride,
                           const float* z, const float* x_enc, int n_freq_xyz, const float* ray_bias, int n_rays,
                           int S, uint32_t mode, float* raw_out, void* saved, void* stream) {
    FieldShape fs;
    int rc = make_shape(d, &fs);
    if (rc != NWB_OK) return rc;
    NWB_REQUIRE(n_rays >= 0 && S >= 1, "nwb_mlp_fwd: bad shape");
    if (n_rays == 0) return NWB_OK;
    NWB_REQUIRE(packed && raw_out, "nwb_mlp_fwd: null pointer");
    NWB_REQUIRE((z != nullptr) != (x_enc != nullptr), "nwb_mlp_fwd: exactly one of z / x_enc must be given");
    NWB_REQUIRE(z == nullptr || (rays != nullptr && ray_stride >= 8), "nwb_mlp_fwd: z needs rays (stride >= 8)");
    NWB_REQUIRE(z == nullptr || (fs.in_xyz == 3 + 6 * n_freq_xyz && n_freq_xyz <= 10),
                "nwb_mlp_fwd: in_xyz must equal 3+6*n_freq_xyz with n_freq_xyz <= 10");
    const bool sigma_only = (mode & NWB_MLP_SIGMA_ONLY) != 0, transient = (mode & NWB_MLP_TRANSIENT) != 0;
    NWB_REQUIRE(!(sigma_only && transient), "nwb_mlp_fwd: sigma_only excludes transient");
    NWB_REQUIRE(!transient || fs.has_t(), "nwb_mlp_fwd: transient requested but the field has no transient head");
    NWB_REQUIRE(sigma_only || ray_bias != nullptr, "nwb_mlp_fwd: ray_bias is required unless sigma_only");
    NWB_REQUIRE(!(mode & NWB_MLP_SAVE) || saved != nullptr, "nwb_mlp_fwd: NWB_MLP_SAVE needs a saved buffer");
    NWB_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 1023u) == 0, "nwb_mlp_fwd: packed buffer must be 1 KB aligned");
    NWB_REQUIRE(!(mode & NWB_MLP_SAVE) || (reinterpret_cast<uintptr_t>(saved) & 15u) == 0, "nwb_mlp_fwd: saved buffer must be 16 B aligned");

    FwdArgs a{};
    a.fs = fs; a.packed = static_cast<const uint8_t*>(packed); a.rays = rays; a.ray_stride = ray_stride;
    a.z = z; a.x_enc = x_enc; a.n_freq_xyz = n_freq_xyz; a.ray_bias = ray_bias; a.S = S;
    a.n_points = static_cast<long long>(n_rays) * S;
    a.n_tiles = static_cast<int>((a.n_points + kTileM - 1) / kTileM);
    a.mode = mode; a.raw_out = raw_out; a.saved = static_cast<uint8_t*>(saved);
    const SaveLayout sl{fs.kx, transient};
    a.save_tile_bytes = sl.tile_bytes(fs.n_layers(mode));

    static bool configured = false;
    if (!configured) {
        NWB_CUDA(cudaFuncSetAttribute(field_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem));
        NWB_CUDA(cudaFuncSetAttribute(field_fwd_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        configured = true;
    }
    const int grid = std::min(a.n_tiles, 2 * kNumSMs);
    field_fwd_kernel<<<grid, kFwdThreads, kFwdSmem, static_cast<cudaStream_t>(stream)>>>(a);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
