// Backward of the fused field network, part 2: weight / bias / head-weight gradients.
//
// Every parameter gradient of the field is a contraction over points of two 8-bit tile images that
// already sit in HBM: dW_g = dZ_g^T . A_g (dZ_g, e5m2 times the pass's power-of-two scale, from the dgrad
// kernel's workspace; A_g, e4m3, the layer's input activation saved by the forward).  Both images have
// rows == points, so both are fed to tcgen05.mma kind::f8f6f4 MN-major with no transpose anywhere (K = 32
// points per instruction, fp32 accumulate -- the products of two 8-bit floats are exact in fp32).  The work
// is cut into "units" -- one 128-column block of one layer's dZ (the M side, 16 KB per tile) against one
// activation image (the N side, <= 256 columns, <= 32 KB per tile) -- and each unit is split over points
// across CTAs in proportion to its byte traffic (the kernel is HBM-bound: 48 KB of operands per 512 tensor
// cycles).  A CTA keeps its (128 x <=256) fp32 partial in TMEM over its whole tile range, and flushes it once
// (times 1/scale) with fp32 atomics.  Bias gradients ride along as one more N=16 MMA per k-step against a
// constant block of ones; head-weight gradients are units whose M side is the head's input activation -- its
// e4m3 image AND the e4m3 image of what that left over, two MMAs into one accumulator -- and whose N side is the
// dpre block (three e5m2 terms per head, summed at the flush): ~8 significant bits on either side, because the heads
// are fp32 in the forward and a single 8-bit contraction there is the largest error of the whole 8-bit scheme
// (tools/fp8_wgrad_study.py heads).
#include "field_dev.cuh"
#include <algorithm>

namespace nwb {

constexpr int kWgThreads = 192;
constexpr int kWgStages = 4;                     // 4 x 48 KB in flight per SM: the kernel lives on HBM latency x bandwidth
constexpr int kWgMBytes = 16384;                 // M side: 128 points x 128 columns, 8-bit
constexpr int kWgNBytes = 32768;                 // N side: 128 points x <= 256 columns, 8-bit
constexpr int kWgStageBytes = kWgMBytes + kWgNBytes;
constexpr int kWgOffOnes = kWgStages * kWgStageBytes;
constexpr int kWgOffBar = kWgOffOnes + 4096;
constexpr int kWgSmem = kWgOffBar + 128;
constexpr int kWgMaxUnits = 40;

struct WUnit {
    uint32_t m_off, n_off;        // byte offsets inside the tile record
    uint8_t m_src, n_src;         // 0 = dgrad workspace, 1 = forward save buffer
    uint8_t n_blocks;             // N side 128-column blocks (1..2)
    uint16_t n_mma;               // MMA N (a multiple of 16 up to 256; 32 for head units)
    uint8_t m_e5m2;               // 1: M side is a dZ image (e5m2), N side e4m3; 0: head unit, M side e4m3, N side e5m2
    uint8_t ones;                 // also accumulate sum_p dZ[p, m] (bias gradient)
    uint8_t head_lo, head_hi;     // head unit: D[m][n] + D[m][9+n] + D[m][18+n], n in [head_lo, head_hi) -> head_out[n][m_base + m]
    uint32_t lo_off;              // head unit: the residual image of the M side (forward save buffer), staged behind the dpre block
    uint16_t n_valid;             // valid N columns (e.g. 63 of 64 for the encoded-xyz block)
    uint16_t m_base;              // first output row (normal) / first head column (head unit)
    uint16_t n_base;              // first output column
    uint16_t ld;                  // output row stride (floats)
    float* out;                   // normal unit: out[(m_base+m)*ld + n_base + n]
    float* out_b;                 // ones: out_b[m_base + m]
    uint16_t cta0, splits;        // CTAs [cta0, cta0+splits) work on this unit
};

struct WgArgs {
    const DzHeader* hdr;
    const uint8_t* dz;
    const uint8_t* saved;
    size_t dz_tile_bytes, save_tile_bytes;
    int n_tiles, n_units;
    float* head_out[9];
    float* partial;               // kNumSMs slots of kWgSlotFloats (field.cuh), behind the tile records of the workspace
    WUnit units[kWgMaxUnits];
};

__global__ void __launch_bounds__(kWgThreads, 1)
field_wgrad_kernel(const __grid_constant__ WgArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sOnes = smem + kWgOffOnes;
    uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem + kWgOffBar);
    uint64_t* bar_empty = bar_full + kWgStages;
    uint64_t* bar_done = bar_empty + kWgStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_done + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // which unit / which slice of the tiles is mine
    int ui = 0;
    while (ui + 1 < a.n_units && static_cast<int>(blockIdx.x) >= a.units[ui + 1].cta0) ++ui;
    const WUnit& u = a.units[ui];
    const int split = blockIdx.x - u.cta0;
    const int t0 = static_cast<int>(static_cast<long long>(a.n_tiles) * split / u.splits);
    const int t1 = static_cast<int>(static_cast<long long>(a.n_tiles) * (split + 1) / u.splits);
    const uint32_t m_stride = static_cast<uint32_t>(u.m_src ? a.save_tile_bytes : a.dz_tile_bytes);
    const uint32_t n_stride = static_cast<uint32_t>(u.n_src ? a.save_tile_bytes : a.dz_tile_bytes);
    const uint8_t* m_base = (u.m_src ? a.saved : a.dz) + u.m_off;
    const uint8_t* n_base = (u.n_src ? a.saved : a.dz) + u.n_off;
    // whole 32-column groups travel (the dpre block of a head unit is one group), the MMA reads n_mma columns
    const uint32_t n_bytes = ((static_cast<uint32_t>(u.n_mma) + 31u) >> 5) * kImg8GroupBytes;
    const bool head = u.head_hi > u.head_lo;

    if (threadIdx.x == 0) {
        if ((ptx::smem_u32(smem) & 1023u) != 0) {
            printf("nwb: dynamic shared memory is not 1 KB aligned\n");
            __trap();
        }
        for (int s = 0; s < kWgStages; ++s) { ptx::mbar_init(&bar_full[s], 1); ptx::mbar_init(&bar_empty[s], 1); }
        ptx::mbar_init(bar_done, 1);
        ptx::fence_mbar_init();
    }
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) reinterpret_cast<uint32_t*>(sOnes)[i] = 0x38383838u;  // e4m3 1.0
    ptx::fence_proxy_async_smem();
    if (warp == 1) ptx::tmem_alloc<512>(tmem_slot);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;
            for (int t = t0; t < t1; ++t, ++it) {
                const uint32_t s = it % kWgStages;
                ptx::mbar_wait(&bar_empty[s], ((it / kWgStages) & 1u) ^ 1u);
                ptx::mbar_arrive_expect_tx(&bar_full[s], kWgMBytes + n_bytes + (head ? 16384u : 0u));
                uint8_t* st = smem + s * kWgStageBytes;
                ptx::bulk_g2s(st, m_base + static_cast<size_t>(t) * m_stride, kWgMBytes, &bar_full[s]);
                ptx::bulk_g2s(st + kWgMBytes, n_base + static_cast<size_t>(t) * n_stride, n_bytes, &bar_full[s]);
                if (head)      // residual image of the M side, staged in the second half of the N slot
                    ptx::bulk_g2s(st + kWgMBytes + 16384, a.saved + u.lo_off + static_cast<size_t>(t) * a.save_tile_bytes, 16384, &bar_full[s]);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            // D(128 m x N) += M_img^T (32 points x 128 m, MN-major) . N_img (32 points x N, MN-major)
            const uint32_t idesc = ptx::umma_idesc_f8(128, u.n_mma, u.m_e5m2 ? 1 : 0, u.m_e5m2 ? 0 : 1, true, true);
            constexpr uint32_t idesc1 = ptx::umma_idesc_f8(128, 16, 1, 0, true, true);
            const uint32_t ones = ptx::smem_u32(sOnes);
            uint32_t it = 0;
            for (int t = t0; t < t1; ++t, ++it) {
                const uint32_t s = it % kWgStages;
                ptx::mbar_wait(&bar_full[s], (it / kWgStages) & 1u);
                ptx::tc_fence_after_sync();
                const uint32_t ma = ptx::smem_u32(smem + s * kWgStageBytes), na = ma + kWgMBytes;
#pragma unroll
                for (int j = 0; j < 4; ++j) {            // K = 32 points per instruction = 1 KB of every 32-column group
                    const uint64_t ad = ptx::umma_desc_any(ma + j * kImg8KStep, kImg8DescLbo, kImg8DescSbo, kImg8DescLayout);
                    const uint64_t nd = ptx::umma_desc_any(na + j * kImg8KStep, kImg8DescLbo, kImg8DescSbo, kImg8DescLayout);
                    ptx::umma_f8(tmem, ad, nd, idesc, (it | j) != 0);
                    if (u.ones) ptx::umma_f8(tmem + 256, ad, ptx::umma_desc_any(ones, kImg8DescLbo, kImg8DescSbo, kImg8DescLayout), idesc1, (it | j) != 0);
                    if (head)      // residual image (second half of the N slot) against the same dpre block, its own accumulator
                        ptx::umma_f8(tmem + 32, ptx::umma_desc_any(na + 16384 + j * kImg8KStep, kImg8DescLbo, kImg8DescSbo, kImg8DescLayout), nd, idesc, (it | j) != 0);
                }
                ptx::umma_commit(&bar_empty[s]);
            }
            ptx::umma_commit(bar_done);
        }
        __syncwarp();
    } else if (t1 > t0) {
        // ---- flush: thread == D row m; raw partial sums to this CTA's slot with plain stores (row-major, so that the reduce
        // kernel reads and writes along rows), scaled and summed in CTA order by wgrad_reduce_kernel ------------
        const int quad = warp & 3;
        const int m = quad * 32 + lane;
        const uint32_t tlane = tmem + (static_cast<uint32_t>(quad * 32) << 16);
        float* slot = a.partial + static_cast<size_t>(blockIdx.x) * kWgSlotFloats;
        ptx::mbar_wait(bar_done, 0);
        ptx::tc_fence_after_sync();
        if (u.head_hi > u.head_lo) {
            uint32_t v[32], w[32];              // image proper / residual image (times kLoScale)
            ptx::tmem_ld32(tlane, v);
            ptx::tmem_ld32(tlane + 32, w);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int n = 0; n < 9; ++n)
                if (n >= u.head_lo && n < u.head_hi) {
                    const float hi = (__uint_as_float(v[n]) + __uint_as_float(v[9 + n])) + __uint_as_float(v[18 + n]);
                    const float lo = (__uint_as_float(w[n]) + __uint_as_float(w[9 + n])) + __uint_as_float(w[18 + n]);
                    slot[kWgSlotHead + n * 128 + m] = fmaf(lo, 1.0f / kLoScale, hi);
                }
        } else {
            for (int c0 = 0; c0 < u.n_valid; c0 += 32) {
                uint32_t v[32];
                ptx::tmem_ld32(tlane + c0, v);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<uint4*>(slot + m * 256 + c0 + 4 * j) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            }
            if (u.ones && u.out_b != nullptr) {
                uint32_t v[32];
                ptx::tmem_ld32(tlane + 256, v);
                ptx::tmem_ld_wait();
                slot[kWgSlotBias + m] = __uint_as_float(v[0]);
            }
        }
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc<512>(tmem);
}

// Second phase: every output element belongs to exactly one unit; its value is the sum of that unit's CTAs' partials,
// taken in CTA order (bit-reproducible), times the pass's inverse scale, ADDED to the sink (caller-zeroed accumulate
// semantics; no other kernel touches these elements while the backward runs).  grid = (units, kWgReduceY).
constexpr int kWgReduceY = 16;
__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const __grid_constant__ WgArgs a) {
    const WUnit& u = a.units[blockIdx.x];
    const float inv = 1.0f / dz_scale_from_amax_bits(__ldg(&a.hdr->amax_bits));     // exact: a power of two
    const int ns = u.splits;                    // every split has tiles (the host never gives a unit more CTAs than tiles)
    const float* base = a.partial + static_cast<size_t>(u.cta0) * kWgSlotFloats;
    const int tid = blockIdx.y * blockDim.x + threadIdx.x, nth = gridDim.y * blockDim.x;
    if (u.head_hi > u.head_lo) {
        const int nh = u.head_hi - u.head_lo;
        for (int i = tid; i < nh * 128; i += nth) {
            const int n = u.head_lo + i / 128, m = i % 128;
            float acc = 0.0f;
            for (int sp = 0; sp < ns; ++sp) acc += base[static_cast<size_t>(sp) * kWgSlotFloats + kWgSlotHead + n * 128 + m];
            a.head_out[n][u.m_base + m] += acc * inv;
        }
        return;
    }
    // thread = (row m, column quad q), q fastest: partial rows and sink rows are both walked along their columns
    const int nq = (u.n_valid + 3) >> 2;
    for (int i = tid; i < nq * 128; i += nth) {
        const int m = i / nq, q = i - m * nq;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int sp = 0; sp < ns; ++sp) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(base + static_cast<size_t>(sp) * kWgSlotFloats + m * 256 + 4 * q));
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        float* o = u.out + static_cast<size_t>(u.m_base + m) * u.ld + u.n_base + 4 * q;
        const int left = u.n_valid - 4 * q;
        if (left >= 4 && (reinterpret_cast<uintptr_t>(o) & 15u) == 0) {
            float4 w = *reinterpret_cast<float4*>(o);
            w.x += acc.x * inv; w.y += acc.y * inv; w.z += acc.z * inv; w.w += acc.w * inv;
            *reinterpret_cast<float4*>(o) = w;
        } else {
            o[0] += acc.x * inv;
            if (left > 1) o[1] += acc.y * inv;
            if (left > 2) o[2] += acc.z * inv;
            if (left > 3) o[3] += acc.w * inv;
        }
    }
    if (u.ones && u.out_b != nullptr) {
        for (int m = tid; m < 128; m += nth) {
            float acc = 0.0f;
            for (int sp = 0; sp < ns; ++sp) acc += base[static_cast<size_t>(sp) * kWgSlotFloats + kWgSlotBias + m];
            u.out_b[u.m_base + m] += acc * inv;
        }
    }
}

}  // namespace nwb

using namespace nwb;

int nwb_wgrad_launch(const FieldShape& fs, const nwb_field_grads* g, const uint8_t* saved, size_t save_tile_bytes,
                     const uint8_t* ws, size_t dz_tile_bytes, int n_tiles, uint32_t mode, cudaStream_t st) {
    const bool transient = (mode & NWB_MLP_TRANSIENT) != 0;
    const SaveLayout sl{fs.kx, transient};
    const DzLayout dl{transient};
    WgArgs a{};                            // passed by value at launch
    a.hdr = reinterpret_cast<const DzHeader*>(ws);
    a.dz = ws + kDzHeaderBytes; a.saved = saved; a.dz_tile_bytes = dz_tile_bytes; a.save_tile_bytes = save_tile_bytes;
    a.n_tiles = n_tiles;
    a.partial = reinterpret_cast<float*>(const_cast<uint8_t*>(ws) + kDzHeaderBytes + static_cast<size_t>(n_tiles) * dz_tile_bytes);
    int nu = 0;
    // n_cols: columns of the N-side image the MMA reads (64 for the encoded-xyz block, else 128 per block)
    auto add = [&](uint32_t m_off, int m_src, uint32_t n_off, int n_src, int n_cols, int n_valid, bool ones, float* out,
                   int ld, int m_base_, int n_base_, float* out_b) -> WUnit& {
        WUnit& u = a.units[nu++];
        u.m_off = m_off; u.m_src = static_cast<uint8_t>(m_src); u.n_off = n_off; u.n_src = static_cast<uint8_t>(n_src);
        u.n_blocks = static_cast<uint8_t>((n_cols + 127) / 128); u.n_mma = static_cast<uint16_t>(n_cols);
        u.n_valid = static_cast<uint16_t>(n_valid); u.ones = ones ? 1 : 0; u.m_e5m2 = 1;
        u.out = out; u.ld = static_cast<uint16_t>(ld); u.m_base = static_cast<uint16_t>(m_base_);
        u.n_base = static_cast<uint16_t>(n_base_); u.out_b = out_b; u.head_lo = u.head_hi = 0;
        return u;
    };
    const int xcols = (fs.in_xyz + 15) & ~15;        // MMA N: a multiple of 16 covering the valid encoded-xyz columns
    for (int gq = 0; gq <= 8; ++gq) {
        float* W = (gq < 8) ? g->trunk_w[gq] : g->final_w;
        float* B = (gq < 8) ? g->trunk_b[gq] : g->final_b;
        for (int mh = 0; mh < 2; ++mh) {
            const uint32_t m_off = static_cast<uint32_t>(dl.dz_off(gq)) + mh * 16384u;
            if (gq == 0) {
                add(m_off, 0, static_cast<uint32_t>(sl.x_off()), 1, xcols, fs.in_xyz, true, W, fs.in_xyz, mh * 128, 0, B);
            } else if (gq == 4) {
                add(m_off, 0, static_cast<uint32_t>(sl.x_off()), 1, xcols, fs.in_xyz, false, W, fs.in_xyz + kW, mh * 128, 0, nullptr);
                add(m_off, 0, static_cast<uint32_t>(sl.act_off(3)), 1, 256, 256, true, W, fs.in_xyz + kW, mh * 128, fs.in_xyz, B);
            } else {
                add(m_off, 0, static_cast<uint32_t>(sl.act_off(gq - 1)), 1, 256, 256, true, W, kW, mh * 128, 0, B);
            }
        }
    }
    // g = 9: feature columns of dir_encoding.0 / transient_encoding.0 (their other columns and biases
    // come from the per-ray operands, ray_consts.cu)
    add(static_cast<uint32_t>(dl.dz_off(9)), 0, static_cast<uint32_t>(sl.act_off(8)), 1, 256, 256, false, g->dir_w,
        kW + fs.in_dir + fs.in_a, 0, 0, nullptr);
    if (transient) {
        add(static_cast<uint32_t>(dl.dz_off(9)) + 16384u, 0, static_cast<uint32_t>(sl.act_off(8)), 1, 256, 256, false, g->tenc_w[0],
            kW + fs.in_t, 0, 0, nullptr);
        for (int gq = 10; gq <= 12; ++gq) {
            const uint32_t n_off = (gq == 10) ? static_cast<uint32_t>(sl.act_off(9)) + 16384u : static_cast<uint32_t>(sl.act_off(gq - 1));
            add(static_cast<uint32_t>(dl.dz_off(gq)), 0, n_off, 1, 128, 128, true, g->tenc_w[gq - 9], kHalf, 0, 0, g->tenc_b[gq - 9]);
        }
    }
    // heads: M side = the head's input activation (e4m3 image + e4m3 residual image), N side = dpre block (e5m2, 32 columns read)
    const uint32_t dpre = static_cast<uint32_t>(dl.dpre_off());
    const int nl = fs.n_layers(mode);
    auto add_head = [&](uint32_t m_off, uint32_t lo_off, int m_base_, int lo, int hi) {
        WUnit& u = add(m_off, 1, dpre, 0, 32, 32, false, nullptr, 0, m_base_, 0, nullptr);
        u.m_e5m2 = 0; u.head_lo = static_cast<uint8_t>(lo); u.head_hi = static_cast<uint8_t>(hi); u.lo_off = lo_off;
    };
    for (int mh = 0; mh < 2; ++mh)
        add_head(static_cast<uint32_t>(sl.act_off(7)) + mh * 16384u, static_cast<uint32_t>(sl.lo_off(0, nl)) + mh * 16384u, mh * 128, 0, 1);
    add_head(static_cast<uint32_t>(sl.act_off(9)), static_cast<uint32_t>(sl.lo_off(1, nl)), 0, 1, 4);
    if (transient) add_head(static_cast<uint32_t>(sl.act_off(12)), static_cast<uint32_t>(sl.lo_off(2, nl)), 0, 4, 9);
    a.head_out[0] = g->sigma_w;
    for (int i = 0; i < 3; ++i) a.head_out[1 + i] = g->rgb_w + i * kHalf;
    if (transient) {
        a.head_out[4] = g->tsigma_w;
        for (int i = 0; i < 3; ++i) a.head_out[5 + i] = g->trgb_w + i * kHalf;
        a.head_out[8] = g->tbeta_w;
    }
    a.n_units = nu;
    // split every unit over points in proportion to its operand bytes per tile
    double cost[kWgMaxUnits], total = 0;
    int sp[kWgMaxUnits], cta = 0;
    for (int i = 0; i < nu; ++i) {
        // empirical weights [B200, profiles/r02i_wgrad_split.log]: a unit's time per tile follows its 128-column blocks rather
        // than its bytes (a head unit -- 36 KB in three copies, eight small MMAs -- takes as long as a 48 KB trunk unit);
        // splitting by bytes left the head units' CTAs running 20 % longer than the rest
        cost[i] = kWgMBytes + 16384.0 * a.units[i].n_blocks + (a.units[i].head_hi > a.units[i].head_lo ? 16384.0 : 0.0);
        total += cost[i];
    }
    const int budget = std::max(nu, std::min(kNumSMs, nu * n_tiles));      // one CTA per SM, never a second wave
    for (int i = 0; i < nu; ++i) { sp[i] = std::max(1, std::min(static_cast<int>(cost[i] / total * budget), n_tiles)); cta += sp[i]; }
    while (cta > budget) {                       // trim the unit with the least work per CTA
        int k = -1;
        for (int i = 0; i < nu; ++i) if (sp[i] > 1 && (k < 0 || cost[i] / sp[i] < cost[k] / sp[k])) k = i;
        if (k < 0) break;
        --sp[k]; --cta;
    }
    while (cta < budget) {                       // grow the unit with the most work per CTA
        int k = -1;
        for (int i = 0; i < nu; ++i) if (sp[i] < n_tiles && (k < 0 || cost[i] / sp[i] > cost[k] / sp[k])) k = i;
        if (k < 0) break;
        ++sp[k]; ++cta;
    }
    cta = 0;
    for (int i = 0; i < nu; ++i) {
        a.units[i].cta0 = static_cast<uint16_t>(cta);
        a.units[i].splits = static_cast<uint16_t>(sp[i]);
        cta += sp[i];
    }
    static bool configured[kMaxDevices] = {};
    NWB_CONFIGURE_ONCE(configured, {
        NWB_CUDA(cudaFuncSetAttribute(field_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgSmem));
    });
    field_wgrad_kernel<<<cta, kWgThreads, kWgSmem, st>>>(a);
    NWB_LAUNCH_CHECK();
    wgrad_reduce_kernel<<<dim3(nu, kWgReduceY), 256, 0, st>>>(a);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
