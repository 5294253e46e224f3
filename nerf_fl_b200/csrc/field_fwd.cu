// (a2,a3,a4) fused NeRF-W field forward -- models/nerf.py:153-212 + the per-sample input assembly of
// models/rendering.py:96-139, as one persistent warp-specialised tcgen05 kernel.
//
// One CTA per SM owns one 128-point tile at a time and pipelines the layer chain INSIDE the tile:
//   * activations ping-pong between two 64 KB tile images (layer g reads H[(g-1)&1], writes H[g&1]) and the
//     fp32 accumulators ping-pong between the two halves of TMEM (layer g accumulates into columns
//     [(g&1)*256, +256)), so the epilogue of layer g and the MMAs of layer g+1 never touch the same bytes;
//   * the epilogue walks a layer's output in four 64-column quarters -- exactly the K-blocks of the next
//     layer's A operand -- and signals one mbarrier per quarter, the MMA issuer runs K-outer and starts
//     K-block kb of layer g+1 as soon as quarter kb of layer g is in shared memory.  The tensor pipe idles
//     only for the first quarter's latency per layer.
// Roles:
//   warp 0 / lane 0   weight producer: streams the packed bf16 slab images (field_pack.cu) from L2 into a
//                     5-stage ring with linear bulk copies (TMA engine, SASS UBLKCP) + mbarrier tx counts
//   warp 1 / lane 0   MMA issuer: tcgen05.mma M=128 N=128 K=16, A = activation tile image in smem,
//                     B = ring stage, D = fp32 accumulators in TMEM
//   warps 2..17       epilogue, four threads per point (column halves of each quarter): build the encoded-
//                     xyz operand tile (o + d*z, sin/cos octaves, bf16, written straight into the swizzled
//                     A image), then per layer tcgen05.ld -> +bias (+per-ray bias) -> ReLU -> bf16 -> next
//                     layer's A image; sigma / rgb / transient heads on CUDA cores in fp32.
// Activations never leave the SM unless NWB_MLP_SAVE asks for them: then every epilogue thread also writes its 16
// columns of each quarter as e4m3 (one 16-byte store, straight from registers, behind the quarter's handoff) into the
// layer's 8-bit image for the weight-gradient kernel, plus the ReLU sign masks for the dgrad chain.
#include "field_dev.cuh"
#include <algorithm>

namespace nwb {

constexpr int kMmaWarps = 1;                  // MMA-issuing warps, each owns 1/kMmaWarps of a layer's output columns
constexpr int kEpiWarp0 = 4;                  // first epilogue warp: warps 0..3 are one warpgroup (producer, MMA issuer, two spare
                                              // warps) so that setmaxnreg can move its registers to the epilogue warpgroups
constexpr int kEpiWarps = 16;
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kFwdThreads = (kEpiWarp0 + kEpiWarps) * 32;
static_assert(1 + kMmaWarps <= kEpiWarp0, "producer + MMA issuers live in warpgroup 0");
// Registers: setmaxnreg moves registers inside the CTA's own launch allocation (640 threads x 96 = 61 440; 104 at
// launch would not fit the 65 536 of an SM).  Warpgroup 0 drops to 64 and the four epilogue warpgroups rise to 104:
// 128 x 64 + 512 x 104 = 61 440.  ptxas: 28 bytes of spills in the saving pair kernel, none in the other three.
constexpr int kFwdMaxRegs = 96;
constexpr int kFwdRegsWg0 = 64;
constexpr int kFwdRegsEpi = 104;
static_assert(kFwdThreads * kFwdMaxRegs >= 128 * kFwdRegsWg0 + kEpiThreads * kFwdRegsEpi, "setmaxnreg pool is the launch allocation");
constexpr int kStageBytes = 32768;            // 256 weight rows x 64 k (one SW128 K-block of a whole layer)
constexpr int kNumStages = 3;
constexpr int kHSlots = 6;                    // activation K-block slots (128 rows x 64 columns, 16 KB each)
constexpr int kSlotX = 6;                     // barrier index of the encoded-xyz tile
constexpr int kOffX = 0;                      // 16 KB  encoded xyz, 128 x 64
constexpr int kOffH = 16384;                  // 6 x 16 KB activation K-block slots, used round-robin
constexpr int kOffW = kOffH + kHSlots * 16384;               // weight ring
constexpr int kOffC = kOffW + kNumStages * kStageBytes;      // 16 KB  fp32 biases + head weights (ConstOff block)
constexpr int kOffBar = kOffC + 16384;
constexpr int kFwdSmem = kOffBar + 320;
static_assert(ConstOff::count * 4 <= 16384, "constant block does not fit its shared-memory slot");
static_assert(kFwdSmem <= 232448, "forward kernel exceeds the 227 KB shared-memory limit");

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
    int first_tile;        // one-tile kernel: tiles [first_tile, n_tiles); pair kernel: tiles [0, n_tiles)
    uint32_t mode;
    float* raw_out;
    uint8_t* saved;
    size_t save_tile_bytes;
    long long* trace;      // debug (-DNWB_FWD_TRACE): block 0 timestamps for tools/trace_fwd.py, else unused
};

// named barrier of the epilogue threads (barrier 0 is __syncthreads)
__device__ __forceinline__ void epi_bar_all() { asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory"); }

// One 16-column slice of a layer epilogue for this thread's row.
//   x = acc + bias[c] (+ rbias[c]);  optional head dot products (fp32, unrounded x, after ReLU);
//   bf16 pack, ReLU on the packed pairs (rounding is monotonic and keeps the sign, so max(rn(x), 0) ==
//   rn(max(x, 0))) -> chunks chunk0, chunk0+1 (16 B each) of this row in the K-block image blk.
//   kSave: the sign bits of the 16 pre-activations are returned in neg_out, bit (15-j) = column j, and the 16
//   activations as e4m3 (ReLU applied by the conversion) in p8 -- one 16-byte chunk of the saved 8-bit image.
//   l8 (kSave, layers that feed a head, may be null): the e4m3 chunk of what p8 left over (the residual image)
template <bool kRelu, int kHeads, bool kSave>
__device__ __forceinline__ void epi_slice16(uint32_t taddr, const float* __restrict__ bias,
                                            const float* __restrict__ rbias, const float* __restrict__ hw, int hw_ld,
                                            float* hacc, uint8_t* blk, int row, int chunk0, uint32_t& neg_out, uint4& p8,
                                            uint4* l8 = nullptr) {
    uint32_t v[16];
    ptx::tmem_ld16(taddr, v);
    float b[16];
    if (bias != nullptr) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float4 t = reinterpret_cast<const float4*>(bias)[q];                 // shared memory
            b[4 * q] = t.x; b[4 * q + 1] = t.y; b[4 * q + 2] = t.z; b[4 * q + 3] = t.w;
        }
    } else {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(rbias) + q);
            b[4 * q] = t.x; b[4 * q + 1] = t.y; b[4 * q + 2] = t.z; b[4 * q + 3] = t.w;
        }
    }
    ptx::tmem_ld_wait();
    float x[16];
#pragma unroll
    for (int j = 0; j < 8; ++j)                  // packed fp32 adds (FADD2): the epilogue is issue-bound
        ptx::fadd2(x[2 * j], x[2 * j + 1], __uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]), b[2 * j], b[2 * j + 1]);
    uint32_t pk[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) pk[j] = pack_bf16(x[2 * j], x[2 * j + 1]);
    if (kSave) {
        p8.x = ptx::pack4_e4m3<kRelu>(x[0], x[1], x[2], x[3]);    p8.y = ptx::pack4_e4m3<kRelu>(x[4], x[5], x[6], x[7]);
        p8.z = ptx::pack4_e4m3<kRelu>(x[8], x[9], x[10], x[11]);  p8.w = ptx::pack4_e4m3<kRelu>(x[12], x[13], x[14], x[15]);
        if (kHeads > 0 && l8 != nullptr) {
            l8->x = ptx::residual4_e4m3<kRelu>(p8.x, x[0], x[1], x[2], x[3], kLoScale);    l8->y = ptx::residual4_e4m3<kRelu>(p8.y, x[4], x[5], x[6], x[7], kLoScale);
            l8->z = ptx::residual4_e4m3<kRelu>(p8.z, x[8], x[9], x[10], x[11], kLoScale);  l8->w = ptx::residual4_e4m3<kRelu>(p8.w, x[12], x[13], x[14], x[15], kLoScale);
        }
    }
    if (kRelu && kSave) {
        // packed word j: bit 15 = sign of column 2j, bit 31 = sign of column 2j+1.  One shift + one LOP3 per word
        // drops them at bits 15-j and 31-j; the 16-bit entry is then [even columns, MSB first | odd columns, MSB first]
        // (relu_mask_bit below is the one place that knows the order)
        uint32_t acc = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc |= (pk[j] >> j) & (0x80008000u >> j);
        const uint32_t neg = (acc & 0xFF00u) | (acc >> 24);
        neg_out = neg & 0xFFFFu;
    }
    if (kRelu) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            __nv_bfloat162 t = *reinterpret_cast<__nv_bfloat162*>(&pk[j]);
            t = __hmax2(t, __floats2bfloat162_rn(0.0f, 0.0f));
            pk[j] = *reinterpret_cast<uint32_t*>(&t);
        }
    }
    if (kHeads > 0) {
#pragma unroll
        for (int h = 0; h < kHeads; ++h) {
            float a = hacc[h];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 t = reinterpret_cast<const float4*>(hw + h * hw_ld)[q];     // shared memory
                a = fmaf(kRelu ? fmaxf(x[4 * q], 0.0f) : x[4 * q], t.x, a);
                a = fmaf(kRelu ? fmaxf(x[4 * q + 1], 0.0f) : x[4 * q + 1], t.y, a);
                a = fmaf(kRelu ? fmaxf(x[4 * q + 2], 0.0f) : x[4 * q + 2], t.z, a);
                a = fmaf(kRelu ? fmaxf(x[4 * q + 3], 0.0f) : x[4 * q + 3], t.w, a);
            }
            hacc[h] = a;
        }
    }
    *reinterpret_cast<uint4*>(blk + ptx::image_chunk_off(row, chunk0)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
    *reinterpret_cast<uint4*>(blk + ptx::image_chunk_off(row, chunk0 + 1)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
}

// The same residual chunk, recomputed behind the handoff from what is still at hand: the bf16 activations of this
// thread's 16 columns in the shared-memory image (chunks chunk0, chunk0+1 of its row) and the e4m3 chunk it kept.
// (The pair kernel hands a whole layer over at once; holding four more chunks until then would spill.)
__device__ __forceinline__ uint4 residual_chunk(const uint8_t* blk, int row, int chunk0, const uint4& p8) {
    const uint4 b0 = *reinterpret_cast<const uint4*>(blk + ptx::image_chunk_off(row, chunk0));
    const uint4 b1 = *reinterpret_cast<const uint4*>(blk + ptx::image_chunk_off(row, chunk0 + 1));
    auto lo = [](uint32_t w) { return __uint_as_float(w << 16); };
    auto hi = [](uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); };
    uint4 l8;
    l8.x = ptx::residual4_e4m3<false>(p8.x, lo(b0.x), hi(b0.x), lo(b0.y), hi(b0.y), kLoScale);
    l8.y = ptx::residual4_e4m3<false>(p8.y, lo(b0.z), hi(b0.z), lo(b0.w), hi(b0.w), kLoScale);
    l8.z = ptx::residual4_e4m3<false>(p8.z, lo(b1.x), hi(b1.x), lo(b1.y), hi(b1.y), kLoScale);
    l8.w = ptx::residual4_e4m3<false>(p8.w, lo(b1.z), hi(b1.z), lo(b1.w), hi(b1.w), kLoScale);
    return l8;
}

#ifdef NWB_FWD_TRACE      // tools/trace_fwd.py: build with -DNWB_FWD_TRACE to get block 0's timeline
__device__ __forceinline__ void trace_ev(long long* tr, int slot) {
    if (tr != nullptr && blockIdx.x == 0) tr[slot] = clock64();
}
#define NWB_TRACE(cond, slot) do { if (cond) trace_ev(a.trace, slot); } while (0)
#else
#define NWB_TRACE(cond, slot) do { } while (0)
#endif

// ---- activation slot schedule --------------------------------------------------------------------------
// Output quarters (64 columns = one K-block of the next layer's A operand) are written to the six 16 KB slots
// round-robin, counted from the start of the tile: quarter q of layer g lives in slot (first_quarter(g)+q) % 6.
// A layer's epilogue starts only when ALL of its MMAs are complete, so its input slots are dead by then, and
// its outputs (<= 4 slots) never collide with the next layer's inputs (they ARE its inputs).
__device__ __forceinline__ int layer_quarters(int g, bool transient) {
    if (g <= 8) return 4;
    if (g == 9) return transient ? 4 : 2;
    return 2;
}
__device__ __forceinline__ int first_quarter(int g) { return g <= 10 ? 4 * g : 42 + 2 * (g - 11); }   // transient chain: 40, 42, 44
__device__ __forceinline__ int out_slot(int g, int q) { return (first_quarter(g) + q) % kHSlots; }

// ---- one layer's epilogue of the one-tile kernel, straight-line (see dual_layer_epi in field_fwd_dual.cuh for the why) ------
// Differences from the pair kernel's: every 64-column quarter is handed to the MMA warp on its own barrier as soon as it is
// in shared memory (intra-tile pipelining), quarters land in the six rotating K-block slots, and the 8-bit chunks (image
// proper + residual image for head-feeding layers) are stored right behind each quarter's handoff.
//   slot0        slot of quarter 0 (quarter q -> (slot0 + q) mod 6);  act_s: shared address of bar_act[0]
//   feed_from    first quarter that feeds an MMA (kNQ: none -- last layer; 2: the dir half of layer 9 in transient mode)
//   save_act / save_lo   this tile's image of the layer / residual image (kSave)
template <bool kRelu, int kHeads, int kHeadQ, int kNQ, bool kRayBias, bool kSave>
__device__ __forceinline__ void single_layer_epi(uint32_t tacc, uint32_t bias_s, const float* __restrict__ rbias, uint32_t hw_s,
                                                 uint32_t sH_s, int slot0, uint32_t act_s, int feed_from, uint32_t off0, uint32_t off1,
                                                 int lane, float (&hacc)[5], uint8_t* save_act, uint8_t* save_lo, uint32_t so0, uint32_t so1,
                                                 uint32_t& mlo, uint32_t& mhi) {
#pragma unroll
    for (int q = 0; q < kNQ; ++q) {
        int slot = slot0 + q;
        if (slot >= kHSlots) slot -= kHSlots;
        const uint32_t blk = sH_s + slot * 16384;
        uint32_t v[16];
        ptx::tmem_ld16(tacc + q * 64, v);
        float4 b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (kRayBias) b[i] = __ldg(reinterpret_cast<const float4*>(rbias + q * 64) + i);
            else b[i] = ptx::lds_f4(bias_s + q * 256 + i * 16);
        }
        ptx::tmem_ld_wait();
        float x[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            ptx::fadd2(x[4 * i], x[4 * i + 1], __uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]), b[i].x, b[i].y);
            ptx::fadd2(x[4 * i + 2], x[4 * i + 3], __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]), b[i].z, b[i].w);
        }
        uint32_t pk[8];
        if (kRelu && !kSave) {
#pragma unroll
            for (int j = 0; j < 8; ++j) pk[j] = ptx::pack_bf16_relu(x[2 * j], x[2 * j + 1]);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) pk[j] = pack_bf16(x[2 * j], x[2 * j + 1]);
        }
        uint4 p8 = make_uint4(0, 0, 0, 0), l8 = make_uint4(0, 0, 0, 0);
        if (kSave) {
            p8.x = ptx::pack4_e4m3<kRelu>(x[0], x[1], x[2], x[3]);    p8.y = ptx::pack4_e4m3<kRelu>(x[4], x[5], x[6], x[7]);
            p8.z = ptx::pack4_e4m3<kRelu>(x[8], x[9], x[10], x[11]);  p8.w = ptx::pack4_e4m3<kRelu>(x[12], x[13], x[14], x[15]);
            if (kHeads > 0 && q < kHeadQ) {
                l8.x = ptx::residual4_e4m3<kRelu>(p8.x, x[0], x[1], x[2], x[3], kLoScale);    l8.y = ptx::residual4_e4m3<kRelu>(p8.y, x[4], x[5], x[6], x[7], kLoScale);
                l8.z = ptx::residual4_e4m3<kRelu>(p8.z, x[8], x[9], x[10], x[11], kLoScale);  l8.w = ptx::residual4_e4m3<kRelu>(p8.w, x[12], x[13], x[14], x[15], kLoScale);
            }
            if (kRelu) {
                uint32_t acc = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j) acc |= (pk[j] >> j) & (0x80008000u >> j);
                const uint32_t neg = ((acc & 0xFF00u) | (acc >> 24)) & 0xFFFFu;
                if (q < 2) mlo |= neg << (16 * q); else mhi |= neg << (16 * (q - 2));
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    __nv_bfloat162 t = *reinterpret_cast<__nv_bfloat162*>(&pk[j]);
                    t = __hmax2(t, __floats2bfloat162_rn(0.0f, 0.0f));
                    pk[j] = *reinterpret_cast<uint32_t*>(&t);
                }
            }
        }
        if (kHeads > 0 && q < kHeadQ) {
#pragma unroll
            for (int h = 0; h < kHeads; ++h) {
                float acc = hacc[h];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 t = ptx::lds_f4(hw_s + h * 512 + q * 256 + i * 16);
                    acc = fmaf(kRelu ? fmaxf(x[4 * i], 0.0f) : x[4 * i], t.x, acc);
                    acc = fmaf(kRelu ? fmaxf(x[4 * i + 1], 0.0f) : x[4 * i + 1], t.y, acc);
                    acc = fmaf(kRelu ? fmaxf(x[4 * i + 2], 0.0f) : x[4 * i + 2], t.z, acc);
                    acc = fmaf(kRelu ? fmaxf(x[4 * i + 3], 0.0f) : x[4 * i + 3], t.w, acc);
                }
                hacc[h] = acc;
            }
        }
        ptx::sts_u4(blk + off0, pk[0], pk[1], pk[2], pk[3]);
        ptx::sts_u4(blk + off1, pk[4], pk[5], pk[6], pk[7]);
        ptx::fence_proxy_async_smem();
        ptx::tc_fence_before_sync();
        __syncwarp();
        if (lane == 0 && q >= feed_from) {
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(act_s + 8 * slot) : "memory");
        }
        if (kSave) {
            const uint32_t at = (q >> 1) * 16384 + ((q & 1) ? so1 : so0);
            *reinterpret_cast<uint4*>(save_act + at) = p8;
            if (kHeads > 0 && q < kHeadQ) *reinterpret_cast<uint4*>(save_lo + at) = l8;
        }
    }
}

// kKx = 64-column blocks of the encoded-xyz operand.  kKx == 2 (in_channels_xyz up to 128: N_emb_xyz = 15 -> 93 columns,
// models/nerf.py:83,121-128) keeps the second block where the third ring stage would be and runs a two-stage ring.
template <bool kSave, int kKx>
__global__ void __maxnreg__(kFwdMaxRegs)
field_fwd_kernel(const FwdArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    constexpr int kStages = (kKx == 2) ? kNumStages - 1 : kNumStages;
    uint8_t* sX = smem + kOffX;
    uint8_t* sX1 = smem + kOffW + (kNumStages - 1) * kStageBytes;       // kKx == 2 only: columns 64..127 of the encoded xyz
    uint8_t* sH = smem + kOffH;                                          // six 16 KB K-block slots
    uint8_t* sW = smem + kOffW;
    float* s_part = reinterpret_cast<float*>(smem + kOffX);             // [3][5][128] head partial sums: sX is dead
                                                                        // once layer 4's MMAs are done
    uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem + kOffBar);   // [kNumStages] producer -> MMA
    uint64_t* bar_empty = bar_full + kNumStages;                        // [kNumStages] MMA -> producer
    uint64_t* bar_acc = bar_empty + kNumStages;                         // [1] MMA -> epilogue (layer accumulator complete)
    uint64_t* bar_act = bar_acc + 1;                                    // [7] epilogue -> MMA, per slot (+X)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_act + 7);
    uint32_t* s_woff = tmem_slot + 2;                                   // [kMaxG] slab offsets

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const FieldShape& fs = a.fs;
    const int nl = fs.n_layers(a.mode);
    const bool transient = (a.mode & NWB_MLP_TRANSIENT) != 0;
    const bool sigma_only = (a.mode & NWB_MLP_SIGMA_ONLY) != 0;
    constexpr bool save = kSave;

    if (threadIdx.x == 0) {
        if ((ptx::smem_u32(smem) & 1023u) != 0) {
            printf("nwb: dynamic shared memory is not 1 KB aligned\n");
            __trap();
        }
        for (int s = 0; s < kStages; ++s) { ptx::mbar_init(&bar_full[s], 1); ptx::mbar_init(&bar_empty[s], kMmaWarps); }
        ptx::mbar_init(bar_acc, kMmaWarps);
        for (int q = 0; q < 7; ++q) ptx::mbar_init(&bar_act[q], kEpiWarps);   // one arrival per epilogue warp
        ptx::fence_mbar_init();
        for (int g = 0; g < kMaxG; ++g) s_woff[g] = static_cast<uint32_t>(fs.w_off(g));
    }
    {   // biases and head weights: L1 is carved down to nothing by the 227 KB of shared memory, so keep them here
        const float4* src = reinterpret_cast<const float4*>(a.packed + fs.w_bytes());
        float4* dst = reinterpret_cast<float4*>(smem + kOffC);
        for (int i = threadIdx.x; i < ConstOff::count / 4; i += kFwdThreads) dst[i] = __ldg(src + i);
    }
    if (warp == 1) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp < kEpiWarp0) {
    ptx::setmaxnreg_dec<kFwdRegsWg0>();
    if (warp == 0) {
        // ============================ weight producer ============================================
        // one stage = one 64-wide K-block of a whole layer: N x 64 bf16 = 32 KB (N = 256) or 16 KB (N = 128),
        // contiguous in the packed image (field_pack.cu), one bulk copy
        if (lane == 0) {
            uint32_t s = 0, ph = 0;
            for (int tile = a.first_tile + blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
                for (int g = 0; g < nl; ++g) {
                    const uint8_t* wsrc = a.packed + s_woff[g];
                    const int KB = fs.layer_kb(g), PN = fs.packed_n(g);
                    const uint32_t bytes = static_cast<uint32_t>(fs.layer_n(g, a.mode)) * 128u;
                    for (int kb = 0; kb < KB; ++kb) {
                        ptx::mbar_wait_sleepy(&bar_empty[s], ph ^ 1u);
                        ptx::mbar_arrive_expect_tx(&bar_full[s], bytes);
                        ptx::bulk_g2s(sW + s * kStageBytes, wsrc + static_cast<size_t>(kb) * PN * 128, bytes, &bar_full[s]);
                        if (++s == kStages) { s = 0; ph ^= 1u; }
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp <= kMmaWarps) {
        // ============================ MMA issuers ================================================
        // kMmaWarps warps, each owning a column range of the layer's output (its own TMEM columns and its own
        // rows of every weight stage), so they need no ordering between them.  A single issuing thread spends
        // ~10 cycles per instruction on waits, descriptor setup and commits -- several hundred cycles per
        // K-block -- during which only the MMAs it has already queued keep the tensor pipe busy; two issuers
        // fill each other's gaps.  The whole warp walks the schedule (uniform control flow), one elected lane
        // issues.  Per stage and warp: 4 x tcgen05.mma M=128 N=(256|128)/kMmaWarps K=16 + the commits.
        {
            const int mw = warp - 1;
            constexpr int n256 = 256 / kMmaWarps, n128 = 128 / kMmaWarps;
            constexpr uint32_t idesc256 = ptx::umma_idesc_bf16(128, n256, false, false);
            constexpr uint32_t idesc128 = ptx::umma_idesc_bf16(128, n128, false, false);
            const uint32_t x_lo = ptx::umma_desc_lo(ptx::smem_u32(sX)), h_lo = ptx::umma_desc_lo(ptx::smem_u32(sH));
            const uint32_t x1_lo = ptx::umma_desc_lo(ptx::smem_u32(sX1));
            const uint32_t w_lo = ptx::umma_desc_lo(ptx::smem_u32(sW));
            const uint32_t full_a = ptx::smem_u32(bar_full), empty_a = ptx::smem_u32(bar_empty);
            const uint32_t acc_a = ptx::smem_u32(bar_acc), act_a = ptx::smem_u32(bar_act);
            uint32_t s = 0, ph = 0;                          // ring stage and its phase
            uint32_t act_par = 0;                            // bit i = parity of the next phase of bar_act[i]
            // tcgen05.mma issue is asynchronous but the queue behind it is shallow, and the waits + fence that open
            // a K-block cost a few hundred cycles on one thread (tools/probe_umma.py).  So the waits for K-block
            // kb+1 of the SAME layer are taken between the 2nd and 3rd MMA of K-block kb: two queued MMAs keep the
            // tensor pipe busy meanwhile.  (Never across a layer boundary: the next layer's first K-block is
            // produced by an epilogue that itself waits for this layer's last MMAs.)
            for (int tile = a.first_tile + blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
                int isl = 0;                                 // slot of the next activation K-block (round-robin, per tile)
                for (int g = 0; g < nl; ++g) {
                    const int KB = fs.layer_kb(g);
                    const bool wide = fs.layer_n(g, a.mode) == 256;
                    const uint32_t idesc = wide ? idesc256 : idesc128;
                    const int ncols = wide ? n256 : n128;                                // this warp's output columns
                    const uint32_t dcol = tmem + (g & 1) * 256 + mw * ncols;
                    const uint32_t wrow_lo = (mw * ncols * 128) >> 4;                    // this warp's rows of a weight stage
                    if (g == 10) { isl += 2; if (isl >= kHSlots) isl -= kHSlots; }       // layer 9's dir quarters feed no MMA
                    uint32_t a_lo = 0, b_lo = 0, empty_cur = 0;
                    // opens K-block kb: waits for its activation quarter and its weight stage, returns the descriptors
                    auto open_kb = [&](int kb, uint32_t& a_out, uint32_t& b_out, uint32_t& empty_out) {
                        const bool is_x = (g == 0) || (g == 4 && kb < kKx);
                        const int sl = is_x ? kSlotX : isl;
                        a_out = is_x ? ((kKx == 2 && kb == 1) ? x1_lo : x_lo) : h_lo + isl * 1024;   // 16 KB >> 4
                        if (!is_x && ++isl == kHSlots) isl = 0;
                        if (!is_x || (g == 0 && kb == 0)) {                               // X (all its blocks): awaited once, by layer 0
                            ptx::mbar_wait_a(act_a + 8 * sl, (act_par >> sl) & 1u);
                            act_par ^= 1u << sl;
                        }
                        ptx::mbar_wait_a(full_a + 8 * s, ph);
                        ptx::tc_fence_after_sync();
                        b_out = w_lo + s * (kStageBytes >> 4) + wrow_lo;
                        empty_out = empty_a + 8 * s;
                        if (++s == kStages) { s = 0; ph ^= 1u; }
                    };
                    open_kb(0, a_lo, b_lo, empty_cur);
                    NWB_TRACE(tile == 0 && lane == 0 && mw == 0, g * 16);                 // first operand ready
                    for (int kb = 0; kb < KB; ++kb) {
                        const bool issuer = ptx::elect_one();
                        if (issuer) {
                            ptx::umma_bf16_lo(dcol, a_lo, b_lo, idesc, kb != 0);
                            ptx::umma_bf16_lo(dcol, a_lo + 2, b_lo + 2, idesc, 1);
                        }
                        __syncwarp();
                        uint32_t a_nx = 0, b_nx = 0, empty_nx = 0;
                        if (kb + 1 < KB) open_kb(kb + 1, a_nx, b_nx, empty_nx);
                        if (issuer) {
                            ptx::umma_bf16_lo(dcol, a_lo + 4, b_lo + 4, idesc, 1);
                            ptx::umma_bf16_lo(dcol, a_lo + 6, b_lo + 6, idesc, 1);
                            ptx::umma_commit_a(empty_cur);
                            if (kb == KB - 1) ptx::umma_commit_a(acc_a);                 // this warp's columns of the layer are complete
                        }
                        __syncwarp();
                        NWB_TRACE(tile == 0 && lane == 0 && mw == 0, g * 16 + 5 + kb);    // issued
                        a_lo = a_nx; b_lo = b_nx; empty_cur = empty_nx;
                    }
                }
            }
        }
    }
    } else {
        // ============================ epilogue warps (four threads per point) ====================
        ptx::setmaxnreg_inc<kFwdRegsEpi>();
        const int ew = warp - kEpiWarp0;                   // 0..15
        const int quad = warp & 3;                         // TMEM lane quadrant this warp may read
        const int sub = ew >> 2;                           // 16-column slice of every 64-column quarter
        const int row = quad * 32 + lane;
        const uint32_t tlane = tmem + (static_cast<uint32_t>(quad * 32) << 16);
        const float* cst = reinterpret_cast<const float*>(smem + kOffC);
        const SaveLayout sl{fs.kx, transient};
        uint32_t acc_ev = 0;                               // layers consumed (parity of bar_acc)
        // this thread's 16-byte chunk in an 8-bit image: quarter q = columns [64 q + 16 sub, +16) = chunk 4 q + sub
        const uint32_t so0 = image8_chunk_off(row, sub), so1 = image8_chunk_off(row, 4 + sub);
        const uint32_t smem_s = ptx::smem_u32(smem), cst_s = smem_s + kOffC + sub * 64, sH_s = smem_s + kOffH, act_s = ptx::smem_u32(bar_act);
        const uint32_t off0 = ptx::image_chunk_off(row, 2 * sub), off1 = ptx::image_chunk_off(row, 2 * sub + 1);
        for (int tile = a.first_tile + blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
            const long long p = static_cast<long long>(tile) * kTileM + row;
            const bool valid = p < a.n_points;
            const long long ray = valid ? p / a.S : 0;
            uint8_t* save_tile = save ? a.saved + static_cast<size_t>(tile) * a.save_tile_bytes : nullptr;

            // ---- encoded-xyz operand tile: this thread packs columns [16*sub, 16*sub+16) of its row ----
            if constexpr (kKx == 1) {
                float e[64];
#pragma unroll
                for (int c = 0; c < 64; ++c) e[c] = 0.0f;
                if (valid) {
                    if (a.x_enc != nullptr) {
                        const float* xr = a.x_enc + p * fs.in_xyz;
#pragma unroll
                        for (int c = 0; c < 64; ++c)
                            if (c < fs.in_xyz && (c >> 4) == sub) e[c] = __ldg(xr + c);
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
                uint32_t pk[8];
#pragma unroll
                for (int s4 = 0; s4 < 4; ++s4) {
                    if (s4 == sub) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) pk[j] = pack_bf16(e[16 * s4 + 2 * j], e[16 * s4 + 2 * j + 1]);
                    }
                }
                uint4 x8 = make_uint4(0, 0, 0, 0);
                if (save) {
#pragma unroll
                    for (int s4 = 0; s4 < 4; ++s4) {
                        if (s4 == sub) {
                            x8.x = ptx::pack4_e4m3<false>(e[16 * s4], e[16 * s4 + 1], e[16 * s4 + 2], e[16 * s4 + 3]);
                            x8.y = ptx::pack4_e4m3<false>(e[16 * s4 + 4], e[16 * s4 + 5], e[16 * s4 + 6], e[16 * s4 + 7]);
                            x8.z = ptx::pack4_e4m3<false>(e[16 * s4 + 8], e[16 * s4 + 9], e[16 * s4 + 10], e[16 * s4 + 11]);
                            x8.w = ptx::pack4_e4m3<false>(e[16 * s4 + 12], e[16 * s4 + 13], e[16 * s4 + 14], e[16 * s4 + 15]);
                        }
                    }
                }
                // sX doubles as the head scratch of the previous tile: its last combine ended with a barrier
                *reinterpret_cast<uint4*>(sX + ptx::image_chunk_off(row, 2 * sub)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                *reinterpret_cast<uint4*>(sX + ptx::image_chunk_off(row, 2 * sub + 1)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                ptx::fence_proxy_async_smem();
                ptx::tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&bar_act[kSlotX]);
                // saved 8-bit image of the encoded xyz: columns [16 sub, 16 sub + 16) = chunk `sub` of this row's line
                if (save) *reinterpret_cast<uint4*>(save_tile + sl.x_off() + so0) = x8;
            } else {
                // two 64-column blocks (in_channels_xyz in (64, 128]): this thread owns columns [16 sub, 16 sub + 16) of
                // EACH block; values go straight to their place (static register index, predicated on the owner) so that
                // no 128-entry array is ever live
                float v0[16], v1[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) { v0[j] = 0.0f; v1[j] = 0.0f; }
                auto put = [&](int c, float val) {           // c is a compile-time constant after unrolling
                    if (((c >> 4) & 3) == sub) { if (c < 64) v0[c & 15] = val; else v1[c & 15] = val; }
                };
                if (valid) {
                    if (a.x_enc != nullptr) {
                        const float* xr = a.x_enc + p * fs.in_xyz;
#pragma unroll
                        for (int c = 0; c < 128; ++c)
                            if (c < fs.in_xyz && ((c >> 4) & 3) == sub) put(c, __ldg(xr + c));
                    } else {
                        const float* rr = a.rays + ray * a.ray_stride;
                        const float zz = __ldg(a.z + p);
                        float x3[3], sn[3], cs[3];
#pragma unroll
                        for (int i = 0; i < 3; ++i) {
                            x3[i] = __fadd_rn(__ldg(rr + i), __fmul_rn(__ldg(rr + 3 + i), zz));   // rendering.py:261
                            put(i, x3[i]);
                            sincosf(x3[i], &sn[i], &cs[i]);
                        }
                        // double-angle recurrences, re-anchored with an accurate sincosf at octave 8 (2^8 x is exact in fp32
                        // and far below sincosf's slow-path threshold): the recurrence error stays <= ~2^7 ulp
#pragma unroll
                        for (int k = 0; k < 15; ++k) {
                            if (k < a.n_freq_xyz) {
                                if (k == 8) {
#pragma unroll
                                    for (int i = 0; i < 3; ++i) sincosf(256.0f * x3[i], &sn[i], &cs[i]);
                                }
#pragma unroll
                                for (int i = 0; i < 3; ++i) {
                                    put(3 + 6 * k + i, sn[i]);
                                    put(6 + 6 * k + i, cs[i]);
                                    const float s2 = 2.0f * sn[i] * cs[i];
                                    const float c2 = fmaf(-2.0f * sn[i], sn[i], 1.0f);
                                    sn[i] = s2; cs[i] = c2;
                                }
                            }
                        }
                    }
                }
                uint32_t pk0[8], pk1[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) { pk0[j] = pack_bf16(v0[2 * j], v0[2 * j + 1]); pk1[j] = pack_bf16(v1[2 * j], v1[2 * j + 1]); }
                *reinterpret_cast<uint4*>(sX + ptx::image_chunk_off(row, 2 * sub)) = make_uint4(pk0[0], pk0[1], pk0[2], pk0[3]);
                *reinterpret_cast<uint4*>(sX + ptx::image_chunk_off(row, 2 * sub + 1)) = make_uint4(pk0[4], pk0[5], pk0[6], pk0[7]);
                *reinterpret_cast<uint4*>(sX1 + ptx::image_chunk_off(row, 2 * sub)) = make_uint4(pk1[0], pk1[1], pk1[2], pk1[3]);
                *reinterpret_cast<uint4*>(sX1 + ptx::image_chunk_off(row, 2 * sub + 1)) = make_uint4(pk1[4], pk1[5], pk1[6], pk1[7]);
                ptx::fence_proxy_async_smem();
                ptx::tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&bar_act[kSlotX]);
                if (save) {
                    // saved 8-bit image: ONE 128-column block; columns [16 sub, +16) = chunk sub, columns 64 + [16 sub, +16) = chunk 4 + sub
                    uint4 x8, y8;
                    x8.x = ptx::pack4_e4m3<false>(v0[0], v0[1], v0[2], v0[3]);     x8.y = ptx::pack4_e4m3<false>(v0[4], v0[5], v0[6], v0[7]);
                    x8.z = ptx::pack4_e4m3<false>(v0[8], v0[9], v0[10], v0[11]);   x8.w = ptx::pack4_e4m3<false>(v0[12], v0[13], v0[14], v0[15]);
                    y8.x = ptx::pack4_e4m3<false>(v1[0], v1[1], v1[2], v1[3]);     y8.y = ptx::pack4_e4m3<false>(v1[4], v1[5], v1[6], v1[7]);
                    y8.z = ptx::pack4_e4m3<false>(v1[8], v1[9], v1[10], v1[11]);   y8.w = ptx::pack4_e4m3<false>(v1[12], v1[13], v1[14], v1[15]);
                    *reinterpret_cast<uint4*>(save_tile + sl.x_off() + so0) = x8;
                    *reinterpret_cast<uint4*>(save_tile + sl.x_off() + so1) = y8;
                }
            }

            float sigma = 0.0f, rgb[3] = {0, 0, 0};
            const float* rb = (a.ray_bias != nullptr) ? a.ray_bias + ray * 256 : nullptr;
            for (int g = 0; g < nl; ++g) {
                const uint32_t par = acc_ev & 1u;
                ++acc_ev;
                const uint32_t tacc = tlane + (g & 1) * 256 + sub * 16;
                const bool last = (g == nl - 1);
                ptx::mbar_wait_sleepy(bar_acc, par);
                ptx::tc_fence_after_sync();
                NWB_TRACE(tile == 0 && warp == kEpiWarp0 && lane == 0, g * 16 + 10);          // accumulator seen
                float hacc[5] = {0, 0, 0, 0, 0};
                // sign masks: per layer [sub][row] x 4 uint16 (one per quarter: a warp's store is 256 contiguous bytes); a thread
                // owns the four entries of its sub and writes them as ONE 8-byte store after the layer
                uint32_t mlo = 0, mhi = 0;
                const int slot0 = first_quarter(g) % kHSlots;
                uint8_t* sv_act = save ? save_tile + sl.act_off(g) : nullptr;
                uint8_t* sv_lo = save ? save_tile + sl.lo_off(g == 7 ? 0 : (g == 9 ? 1 : 2), nl) : nullptr;
                const int none = last ? 4 : 0;               // the last layer feeds no MMA
                if (g < 7) {
                    single_layer_epi<true, 0, 0, 4, false, kSave>(tacc, cst_s + (ConstOff::trunk_b + g * 256) * 4, nullptr, 0, sH_s, slot0, act_s, none, off0, off1,
                                                                  lane, hacc, sv_act, sv_lo, so0, so1, mlo, mhi);
                } else if (g == 7) {
                    single_layer_epi<true, 1, 4, 4, false, kSave>(tacc, cst_s + (ConstOff::trunk_b + 7 * 256) * 4, nullptr, cst_s + ConstOff::sigma_w * 4, sH_s, slot0,
                                                                  act_s, none, off0, off1, lane, hacc, sv_act, sv_lo, so0, so1, mlo, mhi);
                } else if (g == 8) {
                    single_layer_epi<false, 0, 0, 4, false, kSave>(tacc, cst_s + ConstOff::final_b * 4, nullptr, 0, sH_s, slot0, act_s, none, off0, off1, lane, hacc,
                                                                   sv_act, sv_lo, so0, so1, mlo, mhi);
                } else if (g == 9) {
                    // per-ray bias already contains dir_encoding.0.bias / transient_encoding.0.bias; the dir half feeds no MMA
                    if (transient) single_layer_epi<true, 3, 2, 4, true, kSave>(tacc, 0, rb + sub * 16, cst_s + ConstOff::rgb_w * 4, sH_s, slot0, act_s, last ? 4 : 2,
                                                                                 off0, off1, lane, hacc, sv_act, sv_lo, so0, so1, mlo, mhi);
                    else single_layer_epi<true, 3, 2, 2, true, kSave>(tacc, 0, rb + sub * 16, cst_s + ConstOff::rgb_w * 4, sH_s, slot0, act_s, 2, off0, off1, lane,
                                                                      hacc, sv_act, sv_lo, so0, so1, mlo, mhi);
                } else if (g < 12) {
                    single_layer_epi<true, 0, 0, 2, false, kSave>(tacc, cst_s + (ConstOff::tenc_b + (g - 10) * 128) * 4, nullptr, 0, sH_s, slot0, act_s, none, off0, off1,
                                                                  lane, hacc, sv_act, sv_lo, so0, so1, mlo, mhi);
                } else {
                    // transient heads: hw rows = [tsigma_w | trgb_w(3) | tbeta_w], contiguous, ld 128
                    single_layer_epi<true, 5, 2, 2, false, kSave>(tacc, cst_s + (ConstOff::tenc_b + 2 * 128) * 4, nullptr, cst_s + ConstOff::tsigma_w * 4, sH_s, slot0,
                                                                  act_s, 2, off0, off1, lane, hacc, sv_act, sv_lo, so0, so1, mlo, mhi);
                }
                if (save && g != 8) *reinterpret_cast<uint2*>(save_tile + sl.mask_off(g, nl) + sub * 1024 + row * 8) = make_uint2(mlo, mhi);
                // ---- heads: combine the four column slices of a row through shared memory (fixed order) ----
                const int nh_heads = (g == 7) ? 1 : (g == 9 ? 3 : (g == 12 ? 5 : 0));
                if (nh_heads > 0) {
                    if (sub > 0) {
#pragma unroll
                        for (int h = 0; h < 5; ++h)          // static indices: hacc stays in registers
                            if (h < nh_heads) s_part[((sub - 1) * 5 + h) * 128 + row] = hacc[h];
                    }
                    epi_bar_all();
                    if (sub == 0) {
#pragma unroll
                        for (int h = 0; h < 5; ++h)
                            if (h < nh_heads)
                                hacc[h] = ((hacc[h] + s_part[h * 128 + row]) + s_part[(5 + h) * 128 + row]) + s_part[(10 + h) * 128 + row];
                        if (g == 7) sigma = softplus_f(hacc[0] + cst[ConstOff::head_b]);                // nerf.py:190
                        if (g == 9) {
#pragma unroll
                            for (int i = 0; i < 3; ++i) rgb[i] = sigmoid_f(hacc[i] + cst[ConstOff::head_b + 1 + i]);   // nerf.py:197
                        }
                    }
                    if (last) epi_bar_all();                // scratch is read before the next tile's X overwrites it
                }
                // ---- tile outputs ---------------------------------------------------------------
                if (last && valid && sub == 0) {
                    if (sigma_only) {
                        a.raw_out[p] = sigma;
                    } else {
                        float* o = a.raw_out + p * (transient ? 9 : 4);
                        o[0] = rgb[0]; o[1] = rgb[1]; o[2] = rgb[2]; o[3] = sigma;               // nerf.py:198
                        if (transient) {
                            const float* hb = cst + ConstOff::head_b;
                            o[4] = sigmoid_f(hacc[1] + hb[5]);                           // nerf.py:206
                            o[5] = sigmoid_f(hacc[2] + hb[6]);
                            o[6] = sigmoid_f(hacc[3] + hb[7]);
                            o[7] = softplus_f(hacc[0] + hb[4]);                          // nerf.py:205
                            o[8] = softplus_f(hacc[4] + hb[8]);                          // nerf.py:207
                        }
                    }
                }
            }
        }
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem);
}

}  // namespace nwb

#include "field_fwd_dual.cuh"

using namespace nwb;

#ifdef NWB_FWD_TRACE
static long long* g_fwd_trace = nullptr;
// tools/trace_fwd.py (a -DNWB_FWD_TRACE build only): device buffer of >= 16*13 int64 that block 0 fills with clock64() stamps
extern "C" void nwb_debug_set_fwd_trace(void* buf) { g_fwd_trace = static_cast<long long*>(buf); }
#endif

// which forward kernel nwb_mlp_fwd launches: 0 = per launch, whichever is faster (default); 1 = always the
// one-tile kernel; 2 = always the CTA-pair kernel.  Process-wide, read at launch time.
static int g_fwd_gen = 0;
extern "C" int nwb_mlp_fwd_select(int gen) {
    NWB_REQUIRE(gen >= 0 && gen <= 2, "nwb_mlp_fwd_select: gen must be 0, 1 or 2");
    g_fwd_gen = gen;
    return NWB_OK;
}

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
    NWB_REQUIRE(z == nullptr || (fs.in_xyz == 3 + 6 * n_freq_xyz && n_freq_xyz <= 15),
                "nwb_mlp_fwd: in_xyz must equal 3+6*n_freq_xyz with n_freq_xyz <= 15");
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
#ifdef NWB_FWD_TRACE
    a.trace = g_fwd_trace;
#endif

    static bool configured[kMaxDevices] = {};
    NWB_CONFIGURE_ONCE(configured, {
        NWB_CUDA(cudaFuncSetAttribute(field_fwd_kernel<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem));
        NWB_CUDA(cudaFuncSetAttribute(field_fwd_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem));
        NWB_CUDA(cudaFuncSetAttribute(field_fwd_kernel<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem));
        NWB_CUDA(cudaFuncSetAttribute(field_fwd_kernel<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem));
        NWB_CUDA(cudaFuncSetAttribute(field_fwd_dual_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDSmem));
        NWB_CUDA(cudaFuncSetAttribute(field_fwd_dual_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDSmem));
    });
    // second-generation kernel (CTA pairs, two tiles in flight per CTA); nwb_mlp_fwd_select(1) forces the first one
    // Two kernels, each used where it is faster [B200, tools/perf_fwd.py]: a round of the pair kernel (74 clusters x
    // 4 tiles) costs ~1.55x (inference) / ~1.75x (saving activations) a round of the one-tile kernel (148 CTAs x 1
    // tile).  So: whole rounds of 296 tiles go to the pair kernel, and a remainder of <= 148 tiles -- which would
    // cost it a full round -- goes to the one-tile kernel, launched behind it on the same stream.
    const int gen = g_fwd_gen;                                      // nwb_mlp_fwd_select: 1 / 2 force a kernel
    constexpr int kRound = 4 * (kNumSMs / 2);
    int pair_tiles = a.n_tiles;                                     // tiles [0, pair_tiles) -> pair kernel
    if (gen == 1 || fs.kx == 2) pair_tiles = 0;                     // two xyz blocks do not fit the pair kernel's shared memory
    else if (gen != 2 && a.n_tiles % kRound != 0 && a.n_tiles % kRound <= kNumSMs) pair_tiles = a.n_tiles - a.n_tiles % kRound;
    if (pair_tiles > 0) {
        FwdArgs a2 = a;
        a2.n_tiles = pair_tiles;
        const int grid2 = 2 * std::min((pair_tiles + 3) / 4, kNumSMs / 2);
        if (mode & NWB_MLP_SAVE) field_fwd_dual_kernel<true><<<grid2, kFwdThreads, kDSmem, static_cast<cudaStream_t>(stream)>>>(a2);
        else field_fwd_dual_kernel<false><<<grid2, kFwdThreads, kDSmem, static_cast<cudaStream_t>(stream)>>>(a2);
        NWB_LAUNCH_CHECK();
        if (pair_tiles == a.n_tiles) return NWB_OK;
    }
    a.first_tile = pair_tiles;
    const int grid = std::min(a.n_tiles - a.first_tile, kNumSMs);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (fs.kx == 2) {
        if (mode & NWB_MLP_SAVE) field_fwd_kernel<true, 2><<<grid, kFwdThreads, kFwdSmem, st>>>(a);
        else field_fwd_kernel<false, 2><<<grid, kFwdThreads, kFwdSmem, st>>>(a);
    } else {
        if (mode & NWB_MLP_SAVE) field_fwd_kernel<true, 1><<<grid, kFwdThreads, kFwdSmem, st>>>(a);
        else field_fwd_kernel<false, 1><<<grid, kFwdThreads, kFwdSmem, st>>>(a);
    }
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
