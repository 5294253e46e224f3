// Static description of the NeRF-W field network (models/nerf.py:121-151) as a schedule of
// tensor-core GEMM layers, shared by the pack / forward / backward / wgrad kernels.
//
// GEMM layers g (N = output features, K-blocks of 64 input features):
//   g=0      xyz_encoding_1            N=256  A = [X]                (X = encoded xyz, 64-col blocks)
//   g=1..3   xyz_encoding_2..4         N=256  A = [H0..H3]
//   g=4      xyz_encoding_5 (skip)     N=256  A = [X, H0..H3]        (input first, nerf.py:187)
//   g=5..7   xyz_encoding_6..8         N=256  A = [H0..H3]           (g=7 epilogue: sigma head)
//   g=8      xyz_encoding_final        N=256  A = [H0..H3]           (no activation)
//   g=9      dir_encoding | tenc.0     N=128|256  A = [H0..H3]       (feature part; dir/a/t parts
//                                                                     arrive as a per-ray bias)
//   g=10..12 transient_encoding.2,4,6  N=128  A = [H2,H3]            (g=12 epilogue: transient heads)
// Heads (N = 1|3|5) run on CUDA cores inside the producing layer's epilogue.
#pragma once
#include "common.cuh"

namespace nwb {

constexpr int kW = 256;          // trunk width
constexpr int kHalf = 128;       // W/2
constexpr int kTileM = 128;      // points per tile
constexpr int kMaxG = 13;

struct FieldShape {
    int in_xyz, in_dir, in_a, in_t;
    int kx;                      // 64-column blocks of the encoded-xyz operand (1 or 2)
    __host__ __device__ bool has_t() const { return in_t > 0; }
    __host__ __device__ int n_layers(uint32_t mode) const {
        if (mode & NWB_MLP_SIGMA_ONLY) return 8;
        return (mode & NWB_MLP_TRANSIENT) ? 13 : 10;
    }
    __host__ __device__ int layer_n(int g, uint32_t mode) const {      // MMA N of layer g
        if (g <= 8) return kW;
        if (g == 9) return (mode & NWB_MLP_TRANSIENT) ? kW : kHalf;
        return kHalf;
    }
    __host__ __device__ int packed_n(int g) const {                    // rows stored in the pack
        if (g <= 8) return kW;
        if (g == 9) return has_t() ? kW : kHalf;
        return kHalf;
    }
    __host__ __device__ int layer_kb(int g) const {                    // number of 64-wide K-blocks
        if (g == 0) return kx;
        if (g == 4) return kx + 4;
        if (g >= 10) return 2;
        return 4;
    }
    // byte offset of layer g's slab image inside the packed buffer; slab kb = rows x 128 B
    __host__ __device__ size_t w_off(int g) const {
        size_t o = 0;
        for (int i = 0; i < g; ++i) o += static_cast<size_t>(packed_n(i)) * 128 * layer_kb(i);
        return o;
    }
    __host__ __device__ int n_packed_layers() const { return has_t() ? 13 : 10; }
    __host__ __device__ size_t w_bytes() const { return w_off(n_packed_layers()); }
};

// fp32 constant block that follows the weight slabs in the packed buffer (float offsets)
// Bit of a 16-column sign-mask entry that holds column c (0..15) of its group: the forward epilogue builds the entry
// from packed bf16 pairs, so even columns land in the high byte and odd columns in the low byte, each MSB first.
__host__ __device__ constexpr int relu_mask_bit(int c) { return (c & 1) ? 7 - (c >> 1) : 15 - (c >> 1); }

struct ConstOff {
    static constexpr int trunk_b = 0;              // 8 x 256
    static constexpr int final_b = 8 * 256;        // 256
    static constexpr int tenc_b = 9 * 256;         // 3 x 128 (transient_encoding.2,4,6)
    static constexpr int sigma_w = tenc_b + 3 * 128;   // 256
    static constexpr int rgb_w = sigma_w + 256;        // 3 x 128
    static constexpr int tsigma_w = rgb_w + 384;       // 128
    static constexpr int trgb_w = tsigma_w + 128;      // 3 x 128
    static constexpr int tbeta_w = trgb_w + 384;       // 128
    static constexpr int head_b = tbeta_w + 128;       // sigma_b, rgb_b[3], tsigma_b, trgb_b[3], tbeta_b
    static constexpr int count = head_b + 16;
};

inline int make_shape(const nwb_field_desc* d, FieldShape* s) {
    NWB_REQUIRE(d != nullptr, "field desc is null");
    NWB_REQUIRE(d->in_xyz >= 3 && d->in_xyz <= 128,
                "in_channels_xyz = %d unsupported (at most two 64-column xyz blocks: N_emb_xyz <= 20 explicit, <= 15 in-kernel)",
                d->in_xyz);
    NWB_REQUIRE(d->in_dir >= 3 && d->in_dir <= 256 && d->in_a >= 0 && d->in_a <= 256 && d->in_t >= 0 && d->in_t <= 256,
                "unsupported dir/appearance/transient widths");
    s->in_xyz = d->in_xyz; s->in_dir = d->in_dir; s->in_a = d->in_a; s->in_t = d->in_t;
    s->kx = (d->in_xyz + 63) / 64;
    return NWB_OK;
}

// ---- 8-bit images (what the forward saves and dgrad writes for the weight-gradient kernel) ------------------------
// An 8-bit tile image of a 128 x 32k matrix: 32-column group m (32 B per row) at m * 4 KB, row r at r * 32, its two
// 16-byte chunks (16 columns each) swapped when bit 2 of r is set -- the SWIZZLE_32B canonical layout, read MN-major
// (rows = contraction index = points) by tcgen05.mma kind::f8f6f4.  Why 32 B and not the 128 B swizzle of the bf16
// images: the writers hold ROWS in lanes (a warp = 32 rows of one TMEM lane quadrant), so with 128-byte lines every
// warp store touched 32 different lines (LSU: one line per cycle -- 2560 cycles of store issue per forward layer and
// slot, profiles/r02h_trace_fwd_dual_save1.log); here a warp's 32 rows are 1 KB of contiguous bytes (8 lines).
//   saved activations: e4m3, unscaled (post-ReLU values; the encoded-xyz block and `feat` are signed)
//   dZ / dpre:         e5m2, times ONE power-of-two scale per backward pass (DzHeader), so that amax|d_raw| -> 2^kDzTop
// tools/fp8_wgrad_study.py: against the fp32 oracle the model gradient moves from 1.29e-3 (bf16 images) to 1.55e-3.
constexpr float kLoScale = 16.0f;
constexpr int kDzTop = 10;       // e5m2 tops out at 57344 ~ 2^15.8: 2^5.8 of headroom for growth along the chain

// saved-activation layout per tile (forward NWB_MLP_SAVE): e4m3 images followed by ReLU sign masks
//   [X: 16K (64 of 128 columns written)][h1..h8: 8 x 32K][feat: 32K][dir|t0: 16K or 32K][t2,t4,t6: 3 x 16K][masks: 13 x 4K]
//   [lo h8: 32K][lo dir: 16K][lo t6: 16K]
// masks: per layer g four 1 KB sub-blocks [sub][row] of 4 x uint16: bit relu_mask_bit(j) of entry q of (sub, row) = 1 iff the
// pre-activation of column 64q + 16sub + j of that row was negative (ReLU gate closed).  A forward epilogue thread owns
// one (sub, row) entry group: its warp writes 256 contiguous bytes.
struct SaveLayout {
    int kx; bool transient;
    __host__ __device__ size_t x_off() const { return 0; }
    __host__ __device__ size_t act_off(int g) const {   // image written by layer g's epilogue
        size_t o = 16384;                                // encoded xyz: one 128-column block whatever kx is
        if (g <= 8) return o + static_cast<size_t>(g) * 32768;
        o += 9 * 32768;
        if (g == 9) return o;
        o += transient ? 32768 : 16384;
        return o + static_cast<size_t>(g - 10) * 16384;
    }
    __host__ __device__ size_t acts_bytes(int n_layers) const {
        if (n_layers <= 8) return act_off(7) + 32768;
        if (n_layers <= 10) return act_off(9) + (transient ? 32768 : 16384);
        return act_off(12) + 16384;
    }
    __host__ __device__ size_t mask_off(int g, int n_layers) const { return acts_bytes(n_layers) + static_cast<size_t>(g) * 4096; }
    // residual images of the three activations that feed the fp32 heads (h8 -> sigma, dir hidden -> rgb, t6 -> transient
    // heads): lo = e4m3(kLoScale (a - e4m3(a))), so that hi + lo / kLoScale carries ~7 significant bits into the
    // head-weight gradients (a single 8-bit contraction there costs the coarse network 4x its gradient error,
    // tools/fp8_wgrad_study.py heads).  The residual is at most 2^-4 of |a|; times 16 it stays below |a| <= 448 and
    // clear of e4m3's subnormals (2^-9 steps below 2^-6), where a unit-scale residual of a small activation would vanish.
    //   k = 0: h8 (32K, layer 7), k = 1: dir hidden (16K, layer 9 quarters 0-1), k = 2: t6 (16K, layer 12)
    __host__ __device__ size_t lo_off(int k, int n_layers) const {
        return acts_bytes(n_layers) + kMaxG * 4096 + (k == 0 ? 0 : (k == 1 ? 32768 : 49152));
    }
    __host__ __device__ size_t tile_bytes(int n_layers) const {
        return acts_bytes(n_layers) + kMaxG * 4096 + (n_layers <= 8 ? 32768 : (n_layers <= 10 ? 49152 : 65536));
    }
};

// backward workspace (written by the dgrad kernel, read by the wgrad kernel): a 256-byte header, then per tile
//   [dZ g0..g8: 9 x 32K][dZ9: 32K (dir|t0) or 16K (dir)][dZ10..12: 3 x 16K][dpre: 16K]
// dZ_g = cotangent of layer g's pre-activation (e5m2 image, scaled); dpre = cotangents of the head pre-activations
// [sigma, rgb(3), t_sigma, t_rgb(3), t_beta] as a THREE-term e5m2 expansion (term k = e5m2 of what terms < k left
// over): one 32-column group (the first 4 KB of its 16 KB slot), column 9k + n = term k of head n, columns 27..31 zero.
// byte offset of element (row, 16-byte chunk index c16 = column / 16) inside an 8-bit image (see above)
__host__ __device__ __forceinline__ uint32_t image8_chunk_off(int row, int c16) {
    return static_cast<uint32_t>((c16 >> 1) * 4096 + row * 32 + (((c16 & 1) ^ ((row >> 2) & 1)) << 4));
}
constexpr uint32_t kImg8GroupBytes = 4096;       // one 32-column group of 128 rows
constexpr uint32_t kImg8DescLbo = 4096, kImg8DescSbo = 256, kImg8DescLayout = 6, kImg8KStep = 1024;   // UMMA descriptor terms (MN-major)

struct DzHeader {
    unsigned int amax_bits;      // max |d_raw| of the pass as fp32 bits (non-negative floats order like their bits)
    unsigned int pad[63];
};
constexpr size_t kDzHeaderBytes = sizeof(DzHeader);
static_assert(kDzHeaderBytes == 256, "header keeps the tile records 16 B aligned");
// scale = 2^(kDzTop - ceil(log2(amax))), from the bits alone (every CTA of dgrad and wgrad derives the same number)
__host__ __device__ __forceinline__ float dz_scale_from_amax_bits(unsigned int bits) {
    if (bits == 0u || bits >= 0x7F800000u) return 1.0f;
    int e = static_cast<int>(bits >> 23) - 127 + ((bits & 0x7FFFFFu) ? 1 : 0);     // ceil(log2(amax)) (normal numbers)
    int p = kDzTop - e;
    p = p < -100 ? -100 : (p > 100 ? 100 : p);
    union { unsigned int u; float f; } c;
    c.u = static_cast<unsigned int>(p + 127) << 23;
    return c.f;
}
// Weight-gradient partials (behind the tile records of the workspace): one slot per CTA of the wgrad launch, written
// with plain stores and summed in CTA order by wgrad_reduce_kernel -- no atomics, bit-reproducible gradients.
//   [main: 128 rows x 256 columns, row-major like the sinks][bias: 128][heads: 9 x 128]   (floats)
constexpr int kWgSlotMain = 128 * 256, kWgSlotBias = kWgSlotMain, kWgSlotHead = kWgSlotBias + 128;
constexpr int kWgSlotFloats = kWgSlotHead + 9 * 128;
constexpr size_t kWgPartialBytes = static_cast<size_t>(kNumSMs) * kWgSlotFloats * sizeof(float);

struct DzLayout {
    bool transient;
    __host__ __device__ size_t dz_off(int g) const {
        if (g <= 9) return static_cast<size_t>(g) * 32768;
        return 9 * 32768 + 32768 + static_cast<size_t>(g - 10) * 16384;
    }
    __host__ __device__ size_t dpre_off() const { return transient ? dz_off(12) + 16384 : 9 * 32768 + 16384; }
    __host__ __device__ size_t tile_bytes() const { return dpre_off() + 16384; }
};

}  // namespace nwb
