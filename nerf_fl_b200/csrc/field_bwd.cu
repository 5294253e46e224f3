// Backward of the fused field network, part 1: the dgrad chain.
//
// Mirror image of field_fwd.cu: one persistent warp-specialised CTA per 128-point tile walks the
// layers in reverse.  For every layer g it forms dZ_g = dA_g (.) relu'(z_g) -- dA_g comes out of
// TMEM (the previous step's tcgen05 product dZ_{g+1} . W_{g+1}) and/or from the head cotangents on
// CUDA cores, the ReLU gate from the sign masks the forward saved -- writes it as a bf16 tile image
// into shared memory (the A operand of the next product, in place), writes the same values as an e5m2
// image straight from registers to the workspace for the wgrad kernel (field_wgrad.cu; 32-byte stores,
// one full sector per 32 columns of a row), and lets the MMA warp issue
// dA_{g-1} = dZ_g . W_g with W_g read MN-major from the very slab images the forward uses.
// Also produced here: the head-bias gradients, the per-ray cotangent d_ray_bias (sum of dZ_9 over
// the samples of a ray) and the head pre-activation cotangents (dpre block) for the wgrad kernel.
// The whole chain runs on cotangents multiplied by ONE power of two per pass (DzHeader, field.cuh: amax|d_raw| ->
// 2^kDzTop, found by dz_amax_kernel in front of this kernel) so that the 8-bit images use e5m2's range; bf16 and fp32
// have the exponent range to spare, and everything that leaves in fp32 is multiplied by the inverse.
#include "field_dev.cuh"
#include <algorithm>

namespace nwb {

// Eight epilogue warps per CTA (two column halves x four TMEM lane quadrants): with two CTAs per SM that is 16 working
// warps, four per scheduler, as in the forward.  The four-warp version (one thread per row and layer, 256 columns
// each) left two warps per scheduler to hide tcgen05.ld and shared-memory latencies, and its epilogue phase was twice
// the MMA phase (profiles/r01g_stalls_dgrad_fine.md: a third of the epilogue warps' time was the accumulator wait of
// the OTHER phase).  320 threads x <= 96 registers x 2 CTAs fit the register file.
constexpr int kBwdEpiWarps = 8;
constexpr int kBwdEpiThreads = kBwdEpiWarps * 32;
constexpr int kBwdThreads = 64 + kBwdEpiThreads;
__device__ __forceinline__ void epi_bar8() { asm volatile("bar.sync 1, %0;" ::"n"(kBwdEpiThreads) : "memory"); }
constexpr int kBStageBytes = 8192;                 // this CTA's half of a stage: 2 of the 4 k-blocks x 32 n-rows x 128 B
constexpr int kBChunkBytes = 4096;                 // one k-block's 32 rows
constexpr int kBNumStages = 6;                  // the kernel is bound by the latency of the weight stream: depth matters
constexpr int kBOffD = 0;                       // 64 KB  dZ image (A operand)
constexpr int kBOffW = kBOffD + 65536;
constexpr int kBOffBar = kBOffW + kBNumStages * kBStageBytes;
constexpr int kBwdSmem = kBOffBar + 256;

// Waits of the single-thread roles.  ncu (profiles/r02e): 55 % of all warp-instructions this kernel issued were spin-wait
// loops of the producer / MMA threads and of the epilogue warps waiting for an accumulator -- on schedulers they share
// with the working epilogue warps of BOTH resident CTAs.  try_wait with a suspend-time hint parks the warp instead.
#ifdef NWB_BWD_SPIN
#define NWB_BWD_WAIT_A(addr, par) ptx::mbar_wait_a(addr, par)
#else
#define NWB_BWD_WAIT_A(addr, par) ptx::mbar_wait_sleepy_a(addr, par)
#endif
// the two waits on the layer's critical path (MMA thread: operand image complete; polling epilogue warp: accumulator complete)
#ifdef NWB_BWD_SPIN_ACT
#define NWB_BWD_WAIT_ACT(addr, par) ptx::mbar_wait_a(addr, par)
#else
#define NWB_BWD_WAIT_ACT(addr, par) NWB_BWD_WAIT_A(addr, par)
#endif
#ifdef NWB_BWD_SPIN_ACC
#define NWB_BWD_WAIT_ACC(addr, par) ptx::mbar_wait_a(addr, par)
#else
#define NWB_BWD_WAIT_ACC(addr, par) ptx::mbar_wait_sleepy_a(addr, par)
#endif

struct BwdArgs {
    FieldShape fs;
    const uint8_t* packed;
    const float* raw;
    const float* d_raw;
    const uint8_t* saved;
    size_t save_tile_bytes;
    const DzHeader* hdr;
    uint8_t* dz;              // first tile record (behind the header)
    size_t dz_tile_bytes;
    float* d_ray_bias;
    float* g_head_b[5];       // sigma_b, rgb_b, tsigma_b, trgb_b, tbeta_b gradient sinks
    int S;
    long long n_points;
    int n_tiles;
    uint32_t mode;
    long long* trace;        // debug (-DNWB_BWD_TRACE): block 0 / first tile clock64 stamps, else unused
};

#ifdef NWB_BWD_TRACE
#define NWB_BTRACE(cond, slot) do { if ((cond) && a.trace != nullptr && blockIdx.x == 0) a.trace[slot] = clock64(); } while (0)
#else
#define NWB_BTRACE(cond, slot) do { } while (0)
#endif

// one 32-column slice of dZ for this thread's row, straight-line (every address term that does not depend on the thread
// is a compile-time constant at the call site)
//   kAcc: take dA from TMEM;  kHeads > 0: add sum_h dp[h] * hw[h][c];  negmask: two 16-column entries (columns 0-15 in
//   the high half), bit relu_mask_bit(c) of an entry set = gate closed
//   Head weights arrive lane-distributed (lane l holds hw[h][c_out + l], one coalesced load per warp issued ahead of
//   the accumulator wait) and are broadcast with shuffles: every thread of the warp needs the same 32 values, and a
//   warp-uniform global load per value costs an L2 round trip each (shared memory is full, L1 is carved away).
//   blk_s: shared address of the 16 KB bf16 image block that holds the slice; cb: its first 16-byte chunk there (0 | 4);
//   rowx = this row's line offset | (row % 8) << 4, so that chunk c of the row sits at blk_s + (rowx ^ (c << 4))
//   g8: this row's 32 bytes in 32-column group 0 of the 128-column block of the layer's e5m2 image that holds the slice
//   (nullptr: ghost tile); ps: the slice's group inside that block (0..3, 4 KB apart).  The row's two 16-byte chunks
//   are swapped when bit 2 of the row is set (SWIZZLE_32B); a warp's 32 rows are 1 KB of contiguous bytes
template <bool kAcc, int kHeads, bool kMask>
__device__ __forceinline__ void bwd_slice(uint32_t taddr, const float* hwl, const float* dp, uint32_t negmask,
                                          uint32_t blk_s, uint32_t rowx, int cb, uint8_t* g8, int ps, int row, uint32_t s8 = 0) {
    float x[32];
    if (kAcc) {
        uint32_t v[32];
        ptx::tmem_ld32(taddr, v);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(v[j]);
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) x[j] = 0.0f;
    }
    if (kHeads > 0) {
#pragma unroll
        for (int h = 0; h < kHeads; ++h) {
            const float d = dp[h], wl = hwl[h];
#pragma unroll
            for (int j = 0; j < 32; ++j) x[j] = fmaf(d, __shfl_sync(0xffffffffu, wl, j), x[j]);
        }
    }
    if (kMask) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if ((negmask >> ((j < 16 ? 16 : 0) + relu_mask_bit(j & 15))) & 1u) x[j] = 0.0f;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q)
        ptx::sts_u4(blk_s + (rowx ^ static_cast<uint32_t>((cb + q) << 4)), pack_bf16(x[8 * q], x[8 * q + 1]), pack_bf16(x[8 * q + 2], x[8 * q + 3]),
                    pack_bf16(x[8 * q + 4], x[8 * q + 5]), pack_bf16(x[8 * q + 6], x[8 * q + 7]));
    // s8 != 0: the same 32 bytes go to a shared-memory copy of the image block instead (staged for one bulk store per layer)
    if (g8 != nullptr || s8 != 0) {
        uint4 lo, hi;
        lo.x = ptx::pack4_e5m2(x[0], x[1], x[2], x[3]);     lo.y = ptx::pack4_e5m2(x[4], x[5], x[6], x[7]);
        lo.z = ptx::pack4_e5m2(x[8], x[9], x[10], x[11]);   lo.w = ptx::pack4_e5m2(x[12], x[13], x[14], x[15]);
        hi.x = ptx::pack4_e5m2(x[16], x[17], x[18], x[19]); hi.y = ptx::pack4_e5m2(x[20], x[21], x[22], x[23]);
        hi.z = ptx::pack4_e5m2(x[24], x[25], x[26], x[27]); hi.w = ptx::pack4_e5m2(x[28], x[29], x[30], x[31]);
        if (s8 != 0) {
            const uint32_t d = s8 + ps * kImg8GroupBytes;
            if (row & 4) { ptx::sts_u4(d, hi.x, hi.y, hi.z, hi.w); ptx::sts_u4(d + 16, lo.x, lo.y, lo.z, lo.w); }
            else { ptx::sts_u4(d, lo.x, lo.y, lo.z, lo.w); ptx::sts_u4(d + 16, hi.x, hi.y, hi.z, hi.w); }
        } else {
            uint8_t* dst = g8 + ps * kImg8GroupBytes;
            if (row & 4) ptx::st_global_256(dst, hi, lo); else ptx::st_global_256(dst, lo, hi);
        }
    }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kBwdThreads, 2)
field_dgrad_kernel(const BwdArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sD = smem + kBOffD;
    uint8_t* sW = smem + kBOffW;
    uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem + kBOffBar);
    uint64_t* bar_empty = bar_full + kBNumStages;
    uint64_t* bar_acc = bar_empty + kBNumStages;
    uint64_t* bar_act = bar_acc + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_act + 1);
    uint32_t* s_woff = tmem_slot + 2;              // [kMaxG]
    float* s_hb = reinterpret_cast<float*>(s_woff + kMaxG + 1);   // [9] head-bias gradient partials

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // CTA pairs: one tcgen05.mma.cta_group::2 per K-step covers this CTA's tile and its partner's (M = 256), and
    // each CTA stages only HALF of every weight stage (its half of the output k-blocks) -- the kernel waits for
    // the weight stream 30 % of the time, and two CTAs per SM leave each only 48 KB of ring.
    const uint32_t rank = ptx::cluster_ctarank();
    const int pair0 = static_cast<int>(ptx::cluster_id_x()), pair_step = static_cast<int>(ptx::cluster_count_x());
    const int n_pairs = (a.n_tiles + 1) >> 1;      // an odd tail tile is paired with a ghost (protocol only)
    const FieldShape& fs = a.fs;
    const int nl = fs.n_layers(a.mode);            // 10 or 13
    const bool transient = (a.mode & NWB_MLP_TRANSIENT) != 0;

    if (threadIdx.x == 0) {
        if ((ptx::smem_u32(smem) & 1023u) != 0) {
            printf("nwb: dynamic shared memory is not 1 KB aligned\n");
            __trap();
        }
        // leader's full barrier: own producer (+ bytes) and the partner's relay; completions are multicast
        for (int s = 0; s < kBNumStages; ++s) { ptx::mbar_init(&bar_full[s], rank == 0 ? 2 : 1); ptx::mbar_init(&bar_empty[s], 1); }
        ptx::mbar_init(bar_acc, 1);
        ptx::mbar_init(bar_act, 2);                 // one arrival per CTA (its epilogue leader), on the leader CTA's copy
        ptx::fence_mbar_init();
        for (int g = 0; g < kMaxG; ++g) s_woff[g] = static_cast<uint32_t>(fs.w_off(g));
        for (int i = 0; i < 9; ++i) s_hb[i] = 0.0f;
    }
    if (warp == 1) ptx::tmem_alloc_pair<256>(tmem_slot);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::cluster_sync_all();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ============================ weight producer ============================================
        if (lane == 0) {
            // one stage = a 32-row chunk of the contraction index n across ALL output k-blocks of the layer:
            // KBO pieces of 32 rows x 128 B (4 KB, contiguous in the slab image), laid [kbo][32 rows] in the stage,
            // so one tcgen05.mma covers the layer's whole output width (N = 64*KBO) and the dZ operand is read once
            // per K-step instead of once per 64-column output block
            uint32_t it = 0;
            const uint32_t empty_s = ptx::smem_u32(bar_empty);
            for (int pair = pair0; pair < n_pairs; pair += pair_step) {
                for (int g = nl - 1; g >= 1; --g) {
                    const uint8_t* wsrc = a.packed + s_woff[g];
                    const int KH = (g >= 10) ? 1 : 2, NC = fs.layer_n(g, a.mode) >> 5, PN = fs.packed_n(g);   // my k-blocks
                    const int kshift = ((g == 4) ? fs.kx : 0) + static_cast<int>(rank) * KH;   // layer 4: skip the xyz slab
                    for (int c = 0; c < NC; ++c, ++it) {
                        const uint32_t s = it % kBNumStages;
                        ptx::mbar_wait_sleepy_a(empty_s + 8 * s, ((it / kBNumStages) & 1u) ^ 1u);
                        ptx::mbar_arrive_expect_tx(&bar_full[s], static_cast<uint32_t>(KH) * kBChunkBytes);
                        for (int kl = 0; kl < KH; ++kl)
                            ptx::bulk_g2s(sW + s * kBStageBytes + kl * kBChunkBytes,
                                          wsrc + static_cast<size_t>(kl + kshift) * PN * 128 + c * kBChunkBytes, kBChunkBytes,
                                          &bar_full[s]);
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ============================ MMA issuer =================================================
        if (lane == 0 && rank == 0) {
            // D(256 x 64*KBO) += dZ(128 x 16 n per CTA, K-major) . W(16 n-rows x 64*KBO k-cols, MN-major; each CTA holds
            // its half of the k-blocks: 8-row groups 1 KB apart, k-blocks kBChunkBytes apart inside the stage)
            const uint32_t da = ptx::smem_u32(sD), wa = ptx::smem_u32(sW);
            const uint32_t full_a = ptx::smem_u32(bar_full), empty_a = ptx::smem_u32(bar_empty);
            const uint32_t acc_a = ptx::smem_u32(bar_acc), act_a = ptx::smem_u32(bar_act);
            uint32_t it = 0, ev = 0;
            for (int pair = pair0; pair < n_pairs; pair += pair_step) {
                for (int g = nl - 1; g >= 1; --g) {
                    NWB_BWD_WAIT_ACT(act_a, ev & 1u);
                    ++ev;
                    ptx::tc_fence_after_sync();
                    NWB_BTRACE(pair == pair0 + pair_step, g * 8 + 0);      // dZ_g seen by the MMA thread
                    const int KBO = (g >= 10) ? 2 : 4, NC = fs.layer_n(g, a.mode) >> 5;
                    const uint32_t idesc = ptx::umma_idesc_bf16(256, 64 * KBO, false, true);
                    const uint32_t dcol = (g == 10) ? 128u : 0u;
                    for (int c = 0; c < NC; ++c, ++it) {
                        const uint32_t s = it % kBNumStages;
                        NWB_BWD_WAIT_A(full_a + 8 * s, (it / kBNumStages) & 1u);
                        ptx::tc_fence_after_sync();
                        const uint32_t wb = wa + s * kBStageBytes;
                        const uint32_t ab = da + (c >> 1) * 16384 + (c & 1) * 64;
#pragma unroll
                        for (int j = 0; j < 2; ++j)
                            ptx::umma2_bf16(tmem + dcol, ptx::umma_desc_sw128(ab + j * 32, 16, 1024),
                                            ptx::umma_desc_sw128(wb + j * 2048, kBChunkBytes, 1024), idesc, (c | j) != 0);
                        ptx::umma2_commit_both(empty_a + 8 * s);
                    }
                    ptx::umma2_commit_both(acc_a);
                    NWB_BTRACE(pair == pair0 + pair_step, g * 8 + 1);      // all MMAs of layer g issued
                }
            }
        } else if (lane == 0) {
            // partner: tell the leader when MY half of a weight stage has landed
            const uint32_t full_a = ptx::smem_u32(bar_full);
            const uint32_t full_leader = ptx::mapa(full_a, 0);
            uint32_t it = 0;
            for (int pair = pair0; pair < n_pairs; pair += pair_step) {
                for (int g = nl - 1; g >= 1; --g) {
                    const int NC = fs.layer_n(g, a.mode) >> 5;
                    for (int c = 0; c < NC; ++c, ++it) {
                        const uint32_t s = it % kBNumStages;
                        NWB_BWD_WAIT_A(full_a + 8 * s, (it / kBNumStages) & 1u);
                        ptx::mbar_arrive_cluster(full_leader + 8 * s);
                    }
                }
            }
        }
        __syncwarp();
    } else {
        // ============================ epilogue warps (two threads per point: column halves) =======
        const int quad = warp & 3;                          // TMEM lane quadrant this warp may read
        const int ch = (warp - 2) >> 2;                     // column half of every layer this thread converts
        const int row = quad * 32 + lane;
        const int et = threadIdx.x - 64;                    // 0..255 among the epilogue threads
        const uint32_t tlane = tmem + (static_cast<uint32_t>(quad * 32) << 16);
        const bool leader = (et == 0);
        const float* cst = reinterpret_cast<const float*>(a.packed + fs.w_bytes());
        const SaveLayout sl{fs.kx, transient};
        const DzLayout dl{transient};
        const int C = transient ? 9 : 4;
        const float scale = dz_scale_from_amax_bits(__ldg(&a.hdr->amax_bits));
        const float inv_scale = 1.0f / scale;               // both exact powers of two
        const uint32_t line_off = static_cast<uint32_t>((row >> 3) * 1024 + (row & 7) * 128);   // this row inside an image block
        const uint32_t rowx = line_off | static_cast<uint32_t>((row & 7) << 4);
        const uint32_t sD_s = ptx::smem_u32(sD), acc_s = ptx::smem_u32(bar_acc);
        const uint32_t sh16 = 16u * ch;
        uint32_t acc_ev = 0;
        const uint32_t act_leader = ptx::mapa(ptx::smem_u32(bar_act), 0);
        for (int pair = pair0; pair < n_pairs; pair += pair_step) {
            const int tile = 2 * pair + static_cast<int>(rank);
            const bool ghost = tile >= a.n_tiles;           // partner of an odd tail tile: all rows invalid, nothing stored
            const long long p0 = static_cast<long long>(tile) * kTileM;
            const long long p = p0 + row;
            const bool valid = p < a.n_points;
            const uint8_t* save_tile = a.saved + static_cast<size_t>(tile) * a.save_tile_bytes;
            uint8_t* dz_tile = a.dz + static_cast<size_t>(tile) * a.dz_tile_bytes;

            // ---- head pre-activation cotangents (nerf.py:190,197,205-207 backward): both threads of a row ----------
            float dp[9];                                    // [sigma, rgb3, tsigma, trgb3, tbeta]
#pragma unroll
            for (int i = 0; i < 9; ++i) dp[i] = 0.0f;
            if (valid) {
                const float* y = a.raw + p * C;
                const float* dy = a.d_raw + p * C;
#pragma unroll
                for (int i = 0; i < 3; ++i) { const float v = __ldg(y + i); dp[1 + i] = __ldg(dy + i) * v * (1.0f - v); }
                dp[0] = __ldg(dy + 3) * (1.0f - expf(-__ldg(y + 3)));          // softplus' = sigmoid(pre) = 1 - exp(-softplus)
                if (transient) {
#pragma unroll
                    for (int i = 0; i < 3; ++i) { const float v = __ldg(y + 4 + i); dp[5 + i] = __ldg(dy + 4 + i) * v * (1.0f - v); }
                    dp[4] = __ldg(dy + 7) * (1.0f - expf(-__ldg(y + 7)));
                    dp[8] = __ldg(dy + 8) * (1.0f - expf(-__ldg(y + 8)));
                }
            }
            if (ch == 0) {                                  // head-bias gradients: one thread per row
#pragma unroll
                for (int i = 0; i < 9; ++i) {
                    const float s = warp_sum(dp[i]);
                    if (lane == 0 && s != 0.0f) atomicAdd(&s_hb[i], s);
                }
            }
#pragma unroll
            for (int i = 0; i < 9; ++i) dp[i] *= scale;     // from here on the chain carries scaled cotangents
            if (!ghost && ch == 0) {
                // dpre block (e5m2, three-term expansion: column 9k + n = e5m2 of what terms < k left of head n's scaled
                // cotangent): one 32-column group, which is all the wgrad kernel's head units read
                uint32_t by[32];
#pragma unroll
                for (int i = 27; i < 32; ++i) by[i] = 0u;
#pragma unroll
                for (int i = 0; i < 9; ++i) {
                    float r = dp[i];
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        const uint32_t b = ptx::cvt_e5m2x2(r, 0.0f) & 0xFFu;
                        by[9 * k + i] = b;
                        r -= ptx::dec_e5m2(b);
                    }
                }
                uint32_t w[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) w[j] = by[4 * j] | (by[4 * j + 1] << 8) | (by[4 * j + 2] << 16) | (by[4 * j + 3] << 24);
                const uint4 lo = make_uint4(w[0], w[1], w[2], w[3]), hi = make_uint4(w[4], w[5], w[6], w[7]);
                uint8_t* dst = dz_tile + dl.dpre_off() + row * 32;
                if (row & 4) ptx::st_global_256(dst, hi, lo); else ptx::st_global_256(dst, lo, hi);
            }

            for (int g = nl - 1; g >= 0; --g) {
                // ReLU gates of layer g, forward layout [sub][row] x 4 uint16 (entry q of sub-block `sub`: bit
                // relu_mask_bit(j) = column 64q + 16sub + j).  The loads are issued here; their first use comes after
                // the accumulator wait, so their L2/DRAM round trip overlaps the wait
                uint2 a0 = make_uint2(0u, 0u), a1 = a0, a2 = a0, a3 = a0;
                const bool gated = (g != 8 && !ghost);
                if (gated) {
                    const uint2* mp = reinterpret_cast<const uint2*>(save_tile + sl.mask_off(g, nl)) + row;
                    a0 = __ldg(mp); a1 = __ldg(mp + 128); a2 = __ldg(mp + 256); a3 = __ldg(mp + 384);
                }
                // head weights of this layer's head term, lane-distributed over this thread's slices
                float hwl[10];
                if (g == 12) {          // columns [64 ch, +64): slices i = 0, 1, heads [tsigma | trgb(3) | tbeta] contiguous, ld 128
#pragma unroll
                    for (int i = 0; i < 10; ++i) hwl[i] = __ldg(cst + ConstOff::tsigma_w + (i % 5) * 128 + 64 * ch + (i / 5) * 32 + lane);
                } else if (g == 9) {    // dir columns [64 ch, +64)
#pragma unroll
                    for (int i = 0; i < 6; ++i) hwl[i] = __ldg(cst + ConstOff::rgb_w + (i % 3) * 128 + 64 * ch + (i / 3) * 32 + lane);
                } else if (g == 7) {    // columns [128 ch, +128)
#pragma unroll
                    for (int i = 0; i < 4; ++i) hwl[i] = __ldg(cst + ConstOff::sigma_w + 128 * ch + i * 32 + lane);
                }
                // Accumulator wait: ONE warp polls the mbarrier, the other seven park at the named barrier that follows
                // anyway (a blocked bar.sync issues nothing; eight polling warps were 29 % of all instructions the kernel
                // issued, profiles/r02g -- a parked try_wait wakes at every mbarrier event of the CTA, and the weight ring
                // produces those all the time)
                if (g < nl - 1) {
                    if (warp == 2) NWB_BWD_WAIT_ACC(acc_s, acc_ev & 1u);
                    ++acc_ev;
                }
                // mask word of a 32-column slice: entries (s0, q) | (s0 + 1, q), q chosen by word (.x: q 0-1, .y: q 2-3) and shift
                auto mword = [](uint32_t w_hi, uint32_t w_lo, uint32_t sh) { return (((w_hi >> sh) & 0xFFFFu) << 16) | ((w_lo >> sh) & 0xFFFFu); };
                NWB_BTRACE(leader && pair == pair0 + pair_step, g * 8 + 2);    // accumulator seen
                epi_bar8();                                 // accumulator complete; everyone is done reading sD (the d_ray_bias sums of layer 9)
                ptx::tc_fence_after_sync();
                NWB_BTRACE(leader && pair == pair0 + pair_step, g * 8 + 3);    // sD free
                // this row in group 0 of block 0 / block 1 of the layer's e5m2 image
                uint8_t* g8a = ghost ? nullptr : dz_tile + dl.dz_off(g) + row * 32;
                uint8_t* g8b = ghost ? nullptr : g8a + 16384;
                if (g >= 10) {
                    // 128 columns: mine are [64 ch, +64) = bf16 image block ch, e5m2 groups 2 ch + i, gate entries q = ch
                    const uint32_t blk = sD_s + ch * 16384;
                    if (g == 12) {
#pragma unroll
                        for (int i = 0; i < 2; ++i)
                            bwd_slice<false, 5, true>(0, hwl + 5 * i, dp + 4, mword(i ? a2.x : a0.x, i ? a3.x : a1.x, sh16), blk, rowx, 4 * i, g8a, 2 * ch + i, row);
                    } else {
#pragma unroll
                        for (int i = 0; i < 2; ++i)
                            bwd_slice<true, 0, true>(tlane + 64 * ch + 32 * i, nullptr, nullptr, mword(i ? a2.x : a0.x, i ? a3.x : a1.x, sh16), blk, rowx, 4 * i,
                                                     g8a, 2 * ch + i, row);
                    }
                } else if (g == 9) {
                    // dir half: columns [64 ch, +64), head term only; transient half: columns 128 + [64 ch, +64), accumulator only
#pragma unroll
                    for (int i = 0; i < 2; ++i)
                        bwd_slice<false, 3, true>(0, hwl + 3 * i, dp + 1, mword(i ? a2.x : a0.x, i ? a3.x : a1.x, sh16), sD_s + ch * 16384, rowx, 4 * i, g8a,
                                                  2 * ch + i, row);
                    if (transient) {
#pragma unroll
                        for (int i = 0; i < 2; ++i)
                            bwd_slice<true, 0, true>(tlane + 128 + 64 * ch + 32 * i, nullptr, nullptr, mword(i ? a2.y : a0.y, i ? a3.y : a1.y, sh16),
                                                     sD_s + (2 + ch) * 16384, rowx, 4 * i, g8b, 2 * ch + i, row);
                    }
                } else {
                    // 256 columns: mine are [128 ch, +128) = bf16 image blocks 2 ch, 2 ch + 1, e5m2 block ch, gate entries q = 2 ch + i / 2
                    const uint32_t w0 = ch ? a0.y : a0.x, w1 = ch ? a1.y : a1.x, w2 = ch ? a2.y : a2.x, w3 = ch ? a3.y : a3.x;
                    uint8_t* g8 = ch ? g8b : g8a;
                    const uint32_t blk = sD_s + 2 * ch * 16384;
                    const uint32_t tcol = tlane + 128 * ch;
                    if (g == 8) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            bwd_slice<true, 0, false>(tcol + 32 * i, nullptr, nullptr, 0u, blk + (i >> 1) * 16384, rowx, 4 * (i & 1), g8, i, row);
                    } else if (g == 7) {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            bwd_slice<true, 1, true>(tcol + 32 * i, hwl + i, dp, mword((i & 1) ? w2 : w0, (i & 1) ? w3 : w1, 16 * (i >> 1)), blk + (i >> 1) * 16384,
                                                     rowx, 4 * (i & 1), g8, i, row);
                    } else {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            bwd_slice<true, 0, true>(tcol + 32 * i, nullptr, nullptr, mword((i & 1) ? w2 : w0, (i & 1) ? w3 : w1, 16 * (i >> 1)), blk + (i >> 1) * 16384,
                                                     rowx, 4 * (i & 1), g8, i, row);
                    }
                }
                NWB_BTRACE(leader && pair == pair0 + pair_step, g * 8 + 4);    // leader's slices written
                ptx::fence_proxy_async_smem();
                ptx::tc_fence_before_sync();                // this thread's tcgen05.ld of the accumulator are done
                epi_bar8();
                NWB_BTRACE(leader && pair == pair0 + pair_step, g * 8 + 5);    // all slices written
                if (leader && g >= 1) { if (rank == 0) ptx::mbar_arrive(bar_act); else ptx::mbar_arrive_cluster(act_leader); }
                if (g == 9) {
                    // per-ray cotangent of the dir/appearance/transient bias: sum of dZ_9 over the ray's samples.  Runs
                    // AFTER the handoff, next to layer 9's MMAs (both only read the image);
                    // the next layer's epilogue rewrites it only behind its barrier.  One column per thread.
                    const int ncols = transient ? 256 : 128;
                    const int c = et;
                    if (c < ncols) {
                        const uint32_t colp = sD_s + (c >> 6) * 16384 + (c & 7) * 2;
                        const int chunk = (c & 63) >> 3;
                        long long ray = p0 / a.S;
                        const int nrow = static_cast<int>(min(static_cast<long long>(kTileM), a.n_points - p0));
                        int r = 0;
                        int seg_end = static_cast<int>(min(static_cast<long long>(nrow), (ray + 1) * a.S - p0));
                        while (r < nrow) {                      // one segment = the rows of one ray inside this tile
                            float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
                            for (; r + 8 <= seg_end; r += 8) {  // r is a multiple of 8 here: rows r..r+7 = one 1 KB group
                                uint32_t w[8];
#pragma unroll
                                for (int k = 0; k < 8; ++k) w[k] = ptx::lds_u16(colp + ptx::image_chunk_off(r + k, chunk));
                                s0 += __uint_as_float(w[0] << 16); s1 += __uint_as_float(w[1] << 16);
                                s2 += __uint_as_float(w[2] << 16); s3 += __uint_as_float(w[3] << 16);
                                s0 += __uint_as_float(w[4] << 16); s1 += __uint_as_float(w[5] << 16);
                                s2 += __uint_as_float(w[6] << 16); s3 += __uint_as_float(w[7] << 16);
                            }
                            for (; r < seg_end; ++r) s0 += __uint_as_float(ptx::lds_u16(colp + ptx::image_chunk_off(r, chunk)) << 16);
                            atomicAdd(a.d_ray_bias + ray * 256 + c, ((s0 + s1) + (s2 + s3)) * inv_scale);
                            ++ray;
                            seg_end = min(nrow, seg_end + a.S);
                        }
                    }
                }
            }
        }
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::cluster_sync_all();                        // the partner has finished with my shared memory / TMEM too
    if (warp == 1) ptx::tmem_dealloc_pair<256>(tmem);
    if (threadIdx.x < 9) {
        const int i = threadIdx.x;
        float* dst = (i == 0) ? a.g_head_b[0] : (i < 4) ? a.g_head_b[1] + (i - 1) : (i == 4) ? a.g_head_b[2]
                   : (i < 8) ? a.g_head_b[3] + (i - 5) : a.g_head_b[4];
        if (dst != nullptr && (i < 4 || transient)) atomicAdd(dst, s_hb[i]);
    }
}

// max |d_raw| of the pass -> DzHeader.amax_bits (zeroed by a memset node in front): non-negative floats order like
// their bit patterns, so one atomicMax per block does it; NaNs are skipped (they would poison the scale)
__global__ void __launch_bounds__(256)
dz_amax_kernel(const float* __restrict__ d_raw, long long n, DzHeader* __restrict__ hdr) {
    float m = 0.0f;
    const long long tid = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
    const long long n4 = ((reinterpret_cast<uintptr_t>(d_raw) & 15u) == 0) ? (n >> 2) : 0;
    const float4* p4 = reinterpret_cast<const float4*>(d_raw);
    for (long long i = tid; i < n4; i += nth) {
        const float4 v = __ldg(p4 + i);
        m = fmaxf(m, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));    // fmaxf drops NaN operands
    }
    for (long long i = (n4 << 2) + tid; i < n; i += nth) m = fmaxf(m, fabsf(__ldg(d_raw + i)));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    __shared__ float red[8];
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) m = fmaxf(m, red[w]);
        if (m > 0.0f && m < INFINITY) atomicMax(&hdr->amax_bits, __float_as_uint(m));
    }
}

}  // namespace nwb

#include "field_bwd_dual.cuh"

using namespace nwb;

// which dgrad kernel nwb_mlp_bwd / nwb_mlp_dgrad launch: 0 (default) and 2 = the two-slot pair kernel; 1 = the first-
// generation kernel (two CTAs per SM, one tile each).  Process-wide, read at launch time; results are identical up to the
// fp32 summation order of the per-ray bias cotangent (tests walk both).
// [B200, profiles/r02t_*, r02u_dgrad_ab.log] 1024 rays, kernels timed alone: static-only pass (10 layers, all 256 wide)
// 0.121 ms two-slot vs 0.129 ms; a+t pass (13 layers, three of them 128 wide: short MMAs behind a fixed handoff chain) 0.283
// vs 0.264 ms -- but inside the training step's graph, where the coarse and fine backward chains share the GPU, two-slot
// everywhere is the fastest: 1.114 ms per step vs 1.118 (mixed) vs 1.127 (first generation everywhere).
static int g_dgrad_gen = 0;
extern "C" int nwb_mlp_dgrad_select(int gen) {
    NWB_REQUIRE(gen >= 0 && gen <= 2, "nwb_mlp_dgrad_select: gen must be 0, 1 or 2");
    g_dgrad_gen = gen;
    return NWB_OK;
}

#ifdef NWB_BWD_TRACE
static long long* g_bwd_trace = nullptr;
// tools/trace_bwd.py (a -DNWB_BWD_TRACE build only): device buffer of >= 8*13 int64 that block 0 fills with clock64() stamps
extern "C" void nwb_debug_set_bwd_trace(void* buf) { g_bwd_trace = static_cast<long long*>(buf); }
#endif

int nwb_wgrad_launch(const FieldShape& fs, const nwb_field_grads* grads, const uint8_t* saved, size_t save_tile_bytes,
                     const uint8_t* workspace, size_t dz_tile_bytes, int n_tiles, uint32_t mode, cudaStream_t st);

extern "C" size_t nwb_mlp_bwd_workspace_bytes(const nwb_field_desc* d, int64_t n_points, uint32_t mode) {
    FieldShape fs;
    if (make_shape(d, &fs) != NWB_OK || n_points <= 0) return 0;
    const DzLayout dl{(mode & NWB_MLP_TRANSIENT) != 0};
    const int64_t n_tiles = (n_points + kTileM - 1) / kTileM;
    return kDzHeaderBytes + static_cast<size_t>(n_tiles) * dl.tile_bytes() + kWgPartialBytes;
}

static int mlp_bwd_phases(const nwb_field_desc* d, const void* packed, const float* raw, const float* d_raw,
                          const void* saved, int n_rays, int S, uint32_t mode, const nwb_field_grads* grads,
                          float* d_ray_bias, void* workspace, void* stream, bool run_dgrad, bool run_wgrad) {
    FieldShape fs;
    int rc = make_shape(d, &fs);
    if (rc != NWB_OK) return rc;
    NWB_REQUIRE(n_rays >= 0 && S >= 1, "nwb_mlp_bwd: bad shape");
    if (n_rays == 0) return NWB_OK;
    NWB_REQUIRE(packed && raw && d_raw && saved && grads && d_ray_bias && workspace, "nwb_mlp_bwd: null pointer");
    const bool transient = (mode & NWB_MLP_TRANSIENT) != 0;
    NWB_REQUIRE(!(mode & NWB_MLP_SIGMA_ONLY), "nwb_mlp_bwd: sigma_only passes carry no gradient (test_time only)");
    NWB_REQUIRE(!transient || fs.has_t(), "nwb_mlp_bwd: transient requested but the field has no transient head");
    NWB_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 1023u) == 0, "nwb_mlp_bwd: packed buffer must be 1 KB aligned");
    NWB_REQUIRE((reinterpret_cast<uintptr_t>(saved) & 15u) == 0 && (reinterpret_cast<uintptr_t>(workspace) & 15u) == 0,
                "nwb_mlp_bwd: saved / workspace buffers must be 16 B aligned");
    for (int i = 0; i < 8; ++i) NWB_REQUIRE(grads->trunk_w[i] && grads->trunk_b[i], "nwb_mlp_bwd: trunk gradient sink %d is null", i);
    NWB_REQUIRE(grads->final_w && grads->final_b && grads->dir_w && grads->sigma_w && grads->sigma_b && grads->rgb_w && grads->rgb_b,
                "nwb_mlp_bwd: static gradient sink is null");
    if (transient) {
        for (int i = 0; i < 4; ++i) NWB_REQUIRE(grads->tenc_w[i] && (i == 0 || grads->tenc_b[i]), "nwb_mlp_bwd: transient gradient sink %d is null", i);
        NWB_REQUIRE(grads->tsigma_w && grads->tsigma_b && grads->trgb_w && grads->trgb_b && grads->tbeta_w && grads->tbeta_b,
                    "nwb_mlp_bwd: transient head gradient sink is null");
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool amax_given = (mode & NWB_MLP_DZ_AMAX_GIVEN) != 0;      // nwb_composite_bwd_seeded left it in the header
    mode &= ~NWB_MLP_DZ_AMAX_GIVEN;

    BwdArgs a{};
    a.fs = fs; a.packed = static_cast<const uint8_t*>(packed); a.raw = raw; a.d_raw = d_raw;
    a.saved = static_cast<const uint8_t*>(saved);
    a.n_points = static_cast<long long>(n_rays) * S;
    a.n_tiles = static_cast<int>((a.n_points + kTileM - 1) / kTileM);
    const SaveLayout sl{fs.kx, transient};
    const DzLayout dl{transient};
    a.save_tile_bytes = sl.tile_bytes(fs.n_layers(mode));
    a.hdr = static_cast<const DzHeader*>(workspace);
    a.dz = static_cast<uint8_t*>(workspace) + kDzHeaderBytes;
    a.dz_tile_bytes = dl.tile_bytes();
    a.d_ray_bias = d_ray_bias;
    a.g_head_b[0] = grads->sigma_b; a.g_head_b[1] = grads->rgb_b; a.g_head_b[2] = grads->tsigma_b;
    a.g_head_b[3] = grads->trgb_b; a.g_head_b[4] = grads->tbeta_b;
    a.S = S; a.mode = mode;
#ifdef NWB_BWD_TRACE
    a.trace = g_bwd_trace;
#endif

    if (run_dgrad) {
        // the pass's scale: amax |d_raw| into the workspace header (a memset node and one small reduction)
        if (!amax_given) {
            NWB_CUDA(cudaMemsetAsync(workspace, 0, kDzHeaderBytes, st));
            const long long n_raw = a.n_points * (transient ? 9 : 4);
            const int ablocks = static_cast<int>(std::min<long long>((n_raw / 4 + 255) / 256 + 1, kNumSMs * 2));
            dz_amax_kernel<<<ablocks, 256, 0, st>>>(d_raw, n_raw, static_cast<DzHeader*>(workspace));
            NWB_LAUNCH_CHECK();
        }
        static bool configured[kMaxDevices] = {};
        NWB_CONFIGURE_ONCE(configured, {
            NWB_CUDA(cudaFuncSetAttribute(field_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem));
            NWB_CUDA(cudaFuncSetAttribute(field_dgrad_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
            NWB_CUDA(cudaFuncSetAttribute(field_dgrad_dual_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kB2Smem));
        });
        if (g_dgrad_gen == 1) {
            const int grid = 2 * std::min((a.n_tiles + 1) / 2, kNumSMs);      // CTA pairs, two CTAs per SM
            field_dgrad_kernel<<<grid, kBwdThreads, kBwdSmem, st>>>(a);
        } else {
            const int grid = 2 * std::min((a.n_tiles + 3) / 4, kNumSMs / 2);  // CTA pairs, one CTA per SM, two tiles in flight each
            field_dgrad_dual_kernel<<<grid, kB2Threads, kB2Smem, st>>>(a);
        }
        NWB_LAUNCH_CHECK();
    }
    if (run_wgrad)
        return nwb_wgrad_launch(fs, grads, a.saved, a.save_tile_bytes, static_cast<const uint8_t*>(workspace), a.dz_tile_bytes,
                                a.n_tiles, mode, st);
    return NWB_OK;
}

extern "C" int nwb_mlp_bwd(const nwb_field_desc* d, const void* packed, const float* raw, const float* d_raw,
                           const void* saved, int n_rays, int S, uint32_t mode, const nwb_field_grads* grads,
                           float* d_ray_bias, void* workspace, void* stream) {
    return mlp_bwd_phases(d, packed, raw, d_raw, saved, n_rays, S, mode, grads, d_ray_bias, workspace, stream, true, true);
}
extern "C" int nwb_mlp_dgrad(const nwb_field_desc* d, const void* packed, const float* raw, const float* d_raw,
                             const void* saved, int n_rays, int S, uint32_t mode, const nwb_field_grads* grads,
                             float* d_ray_bias, void* workspace, void* stream) {
    return mlp_bwd_phases(d, packed, raw, d_raw, saved, n_rays, S, mode, grads, d_ray_bias, workspace, stream, true, false);
}
extern "C" int nwb_mlp_wgrad(const nwb_field_desc* d, const void* packed, const float* raw, const float* d_raw,
                             const void* saved, int n_rays, int S, uint32_t mode, const nwb_field_grads* grads,
                             float* d_ray_bias, void* workspace, void* stream) {
    return mlp_bwd_phases(d, packed, raw, d_raw, saved, n_rays, S, mode, grads, d_ray_bias, workspace, stream, false, true);
}
