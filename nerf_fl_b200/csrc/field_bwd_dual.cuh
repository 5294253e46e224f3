// dgrad chain, second generation: CTA PAIRS with TWO TILES IN FLIGHT per CTA -- the forward pair kernel's structure
// (field_fwd_dual.cuh) applied to the backward chain.  Included by field_bwd.cu (shares BwdArgs and bwd_slice with the
// first-generation kernel, which stays selectable: nwb_mlp_dgrad_select).
//
// Why (DESIGN.md section 4b, round 2 timelines): with two CTAs per SM of eight epilogue warps each, a layer of a tile is
// 4 100 cycles of slices + 900 of handoff + 4 300 of MMA phase, strictly in series inside a CTA, and the two CTAs only
// overlap by luck: the tensor pipe is busy 2 x 2 048 of every 9 700 cycles.  Here ONE CTA per SM owns two tiles
// (slots): all 16 epilogue warps convert slot A's layer while the tensor pipe runs slot B's products, and the whole
// shared memory that two CTAs needed for rings is one ring of large stages (16 KB bulk copies), so the MMA thread stops
// waiting for the weight stream.
//   TMEM   512 columns = 2 slots x 256 fp32 accumulator columns
//   smem   2 x 64 KB dZ images (A operand, rewritten in place per layer), 3 x 32 KB weight ring (this CTA's half of
//          every stage: 128 contraction rows x its 2 of the 4 output k-blocks, 16 KB per copy)
//   epilogue thread = (row, column quarter): two 32-column slices per 256-wide layer, one per 128-wide layer
// Handoffs as in the forward: bar_act[t] (epilogue warps of BOTH CTAs -> rank 0's MMA thread: slot t's dZ image is
// complete) and bar_acc[t] (MMA -> epilogues: slot t's products are complete).
#pragma once

namespace nwb {

constexpr int kB2EpiWarp0 = 4, kB2EpiWarps = 16;
constexpr int kB2Threads = (kB2EpiWarp0 + kB2EpiWarps) * 32;
constexpr int kB2MaxRegs = 96, kB2RegsWg0 = 64, kB2RegsEpi = 104;
static_assert(kB2Threads * kB2MaxRegs >= 128 * kB2RegsWg0 + kB2EpiWarps * 32 * kB2RegsEpi, "setmaxnreg pool is the launch allocation");
// weight ring: a stage = 128 contraction rows x my half of the output k-blocks = 2 (or 1) copies of 16 KB -- a k-block's
// 128 rows are contiguous in the slab image.  (32-row stages of 4 KB copies, as in the first-generation kernel, deliver
// ~29 B/clk/SM from L2 against the 32 B/clk the tensor pipe consumes: the MMA phase of a layer took 4 700 cycles for
// 2 048 of tensor work, profiles/r02r_trace_bwd_dual.log; 16 KB copies sustain ~55-64 B/clk.)
#ifndef NWB_DGRAD_STAGE       // default: 3 x 32 KB ring of 16 KB copies, e5m2 images stored from registers
constexpr int kB2Stages = 3;
constexpr int kB2RowsLog2 = 7;                               // contraction rows per stage = 128
constexpr bool kB2Stage8 = false;
#else                         // experiment (-DNWB_DGRAD_STAGE): e5m2 image staged in shared memory + one bulk store per layer, paid
                              // for with a 4 x 16 KB ring of 8 KB copies.  [B200, profiles/r02w_*] SLOWER: 0.309 vs 0.279 ms (a+t
                              // pass), 0.133 vs 0.119 (static) -- the slices got longer, not shorter (shared memory now carries the
                              // bf16 image, the staged image, the bulk store's read and both MMA operands), and the MMA phase too
                              // (8 KB copies)
constexpr int kB2Stages = 4;
constexpr int kB2RowsLog2 = 6;                               // 64
constexpr bool kB2Stage8 = true;
#endif
constexpr int kB2Rows = 1 << kB2RowsLog2;
constexpr int kB2CopyBytes = kB2Rows * 128;
constexpr int kB2StageBytes = 2 * kB2CopyBytes;
constexpr int kB2OffD = 0;                                   // 2 x 64 KB
constexpr int kB2OffW = 2 * 65536;
// (experiment) staged e5m2 image of the layer being converted (32 KB): register -> global stores move ~32 B/clk/SM and cost
// a step of the chain ~1 000 of its ~4 000 cycles (profiles/r02t NOSTORE experiment); the idea was shared-memory stores plus
// one bulk copy per layer issued by an otherwise idle warp
constexpr int kB2OffS = kB2OffW + kB2Stages * kB2StageBytes;
constexpr int kB2OffBar = kB2OffS + (kB2Stage8 ? 32768 : 0);
constexpr int kB2Smem = kB2OffBar + 512;
static_assert(kB2Smem <= 232448, "dual-slot dgrad exceeds the 227 KB shared-memory limit");

// -DNWB_BWD_TRACE: clock64 stamps of cluster 0 / rank 0 during its SECOND quad, slot ((g * 2 + t) * 8 + k): k = 0 MMA thread sees
// the dZ image, 1 its MMAs are issued, 2 epilogue (first epilogue thread) sees the accumulator, 3 its slices are written, 4 handoff out
#ifdef NWB_BWD_TRACE
#define NWB_B2TRACE(cond, g, t, k) do { if ((cond) && a.trace != nullptr && blockIdx.x == 0 && quad == quad0 + quad_step) a.trace[((g) * 2 + (t)) * 8 + (k)] = clock64(); } while (0)
#else
#define NWB_B2TRACE(cond, g, t, k) do { } while (0)
#endif

__device__ __forceinline__ void epi_bar16() { asm volatile("bar.sync 1, %0;" ::"n"(kB2EpiWarps * 32) : "memory"); }

// head pre-activation cotangents of one row (nerf.py:190,197,205-207 backward), scaled
__device__ __forceinline__ float dp_sigma(const float* y, const float* dy) { return __ldg(dy + 3) * (1.0f - expf(-__ldg(y + 3))); }   // softplus' = 1 - exp(-softplus)
__device__ __forceinline__ float dp_sig(const float* y, const float* dy, int i) { const float v = __ldg(y + i); return __ldg(dy + i) * v * (1.0f - v); }
__device__ __forceinline__ float dp_sp(const float* y, const float* dy, int i) { return __ldg(dy + i) * (1.0f - expf(-__ldg(y + i))); }

__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(kB2MaxRegs)
field_dgrad_dual_kernel(const BwdArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sW = smem + kB2OffW;
    uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem + kB2OffBar);
    uint64_t* bar_empty = bar_full + kB2Stages;
    uint64_t* bar_acc = bar_empty + kB2Stages;                  // [2]
    uint64_t* bar_act = bar_acc + 2;                            // [2]
    uint64_t* bar_staged = bar_act + 2;                         // epilogue warps -> store thread: the layer's e5m2 image is in sS
    uint64_t* bar_free = bar_staged + 1;                        // store thread -> epilogue warps: sS has been read out
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_free + 1);
    uint32_t* s_woff = tmem_slot + 2;                           // [kMaxG]
    float* s_hb = reinterpret_cast<float*>(s_woff + kMaxG + 1); // [9] head-bias gradient partials

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = ptx::cluster_ctarank();
    const int quad0 = static_cast<int>(ptx::cluster_id_x()), quad_step = static_cast<int>(ptx::cluster_count_x());
    const int n_quads = (a.n_tiles + 3) >> 2;                   // a cluster step covers 4 tiles: tile = 4 quad + 2 slot + rank
    const FieldShape& fs = a.fs;
    const int nl = fs.n_layers(a.mode);                         // 10 or 13
    const bool transient = (a.mode & NWB_MLP_TRANSIENT) != 0;

    if (threadIdx.x == 0) {
        if ((ptx::smem_u32(smem) & 1023u) != 0) {
            printf("nwb: dynamic shared memory is not 1 KB aligned\n");
            __trap();
        }
        for (int s = 0; s < kB2Stages; ++s) { ptx::mbar_init(&bar_full[s], rank == 0 ? 2 : 1); ptx::mbar_init(&bar_empty[s], 1); }
        for (int t = 0; t < 2; ++t) { ptx::mbar_init(&bar_acc[t], 1); ptx::mbar_init(&bar_act[t], 2 * kB2EpiWarps); }
        ptx::mbar_init(bar_staged, kB2EpiWarps);
        ptx::mbar_init(bar_free, 1);
        ptx::fence_mbar_init();
        for (int g = 0; g < kMaxG; ++g) s_woff[g] = static_cast<uint32_t>(fs.w_off(g));
        for (int i = 0; i < 9; ++i) s_hb[i] = 0.0f;
    }
    if (warp == 1) ptx::tmem_alloc_pair<512>(tmem_slot);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::cluster_sync_all();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp < kB2EpiWarp0) {
    ptx::setmaxnreg_dec<kB2RegsWg0>();
    if (warp == 0) {
        // ============================ weight producer ============================================
        // one stage = a 32-row chunk of the contraction index n across my half of the layer's output k-blocks (see the
        // first-generation kernel); order (layer, slot, chunk) = the MMA thread's
        if (lane == 0) {
            uint32_t it = 0;
            const uint32_t empty_s = ptx::smem_u32(bar_empty);
            for (int quad = quad0; quad < n_quads; quad += quad_step) {
                for (int g = nl - 1; g >= 1; --g) {
                    const uint8_t* wsrc = a.packed + s_woff[g];
                    const int KH = (g >= 10) ? 1 : 2, NC = fs.layer_n(g, a.mode) >> kB2RowsLog2, PN = fs.packed_n(g);
                    const int kshift = ((g == 4) ? fs.kx : 0) + static_cast<int>(rank) * KH;   // layer 4: skip the xyz slab
                    for (int t = 0; t < 2; ++t) {
                        for (int c = 0; c < NC; ++c, ++it) {
#ifdef NWB_DGRAD_NOWEIGHTS      // timing experiment only (results are wrong): no weight stream at all
                            continue;
#endif
                            const uint32_t s = it % kB2Stages;
                            ptx::mbar_wait_sleepy_a(empty_s + 8 * s, ((it / kB2Stages) & 1u) ^ 1u);
                            ptx::mbar_arrive_expect_tx(&bar_full[s], static_cast<uint32_t>(KH) * kB2CopyBytes);
                            for (int kl = 0; kl < KH; ++kl)     // rows [kB2Rows c, +kB2Rows) of k-block kl + kshift: contiguous in the slab image
                                ptx::bulk_g2s(sW + s * kB2StageBytes + kl * kB2CopyBytes,
                                              wsrc + static_cast<size_t>(kl + kshift) * PN * 128 + c * kB2CopyBytes, kB2CopyBytes, &bar_full[s]);
                        }
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ============================ MMA issuer (rank 0) / weight relay (rank 1) ==================
        if (lane == 0 && rank == 0) {
            const uint32_t wa = ptx::smem_u32(sW);
            const uint32_t full_a = ptx::smem_u32(bar_full), empty_a = ptx::smem_u32(bar_empty);
            const uint32_t acc_a = ptx::smem_u32(bar_acc), act_a = ptx::smem_u32(bar_act);
            uint32_t it = 0, act_par = 0;
            for (int quad = quad0; quad < n_quads; quad += quad_step) {
                for (int g = nl - 1; g >= 1; --g) {
                    const int KBO = (g >= 10) ? 2 : 4, NC = fs.layer_n(g, a.mode) >> kB2RowsLog2;
                    const uint32_t idesc = ptx::umma_idesc_bf16(256, 64 * KBO, false, true);
                    for (int t = 0; t < 2; ++t) {
                        const uint32_t da = ptx::smem_u32(smem + kB2OffD + t * 65536);
                        const uint32_t dcol = tmem + t * 256 + ((g == 10) ? 128u : 0u);
                        ptx::mbar_wait_sleepy_a(act_a + 8 * t, (act_par >> t) & 1u);
                        act_par ^= 1u << t;
                        ptx::tc_fence_after_sync();
                        NWB_B2TRACE(true, g, t, 0);
                        for (int c = 0; c < NC; ++c, ++it) {
                            const uint32_t s = it % kB2Stages;
#ifndef NWB_DGRAD_NOWEIGHTS
                            ptx::mbar_wait_sleepy_a(full_a + 8 * s, (it / kB2Stages) & 1u);
#endif
                            ptx::tc_fence_after_sync();
                            const uint32_t wb = wa + s * kB2StageBytes;
                            // D(256 x 64 KBO) += dZ(128 rows per CTA x 16 contraction columns, K-major) . W(16 n-rows x 64 KBO k-columns,
                            // MN-major: 8-row groups 1 KB apart, my k-blocks kB2CopyBytes apart inside the stage)
#pragma unroll
                            for (int j = 0; j < kB2Rows / 16; ++j) {
                                const int col = c * kB2Rows + 16 * j;            // first contraction column of this MMA
                                ptx::umma2_bf16(dcol, ptx::umma_desc_sw128(da + (col >> 6) * 16384 + ((col & 63) >> 4) * 32, 16, 1024),
                                                ptx::umma_desc_sw128(wb + j * 2048, kB2CopyBytes, 1024), idesc, (c | j) != 0);
                            }
                            ptx::umma2_commit_both(empty_a + 8 * s);
                        }
                        ptx::umma2_commit_both(acc_a + 8 * t);
                        NWB_B2TRACE(true, g, t, 1);
                    }
                }
            }
        } else if (lane == 0) {
            const uint32_t full_a = ptx::smem_u32(bar_full);
            const uint32_t full_leader = ptx::mapa(full_a, 0);
            uint32_t it = 0;
            for (int quad = quad0; quad < n_quads; quad += quad_step) {
                for (int g = nl - 1; g >= 1; --g) {
                    const int n = 2 * (fs.layer_n(g, a.mode) >> kB2RowsLog2);
                    for (int c = 0; c < n; ++c, ++it) {
#ifdef NWB_DGRAD_NOWEIGHTS
                        continue;
#endif
                        const uint32_t s = it % kB2Stages;
                        ptx::mbar_wait_sleepy_a(full_a + 8 * s, (it / kB2Stages) & 1u);
                        ptx::mbar_arrive_cluster(full_leader + 8 * s);
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 2 && kB2Stage8) {
        // ============================ image store: one bulk copy per (layer, slot) =================
        if (lane == 0) {
            const DzLayout dl{transient};
            uint32_t k = 0;
            for (int quad = quad0; quad < n_quads; quad += quad_step) {
                for (int g = nl - 1; g >= 0; --g) {
                    const uint32_t bytes = (g <= 8 || (g == 9 && transient)) ? 32768u : 16384u;
                    for (int t = 0; t < 2; ++t, ++k) {
                        const int tile = 4 * quad + 2 * t + static_cast<int>(rank);
                        ptx::mbar_wait_sleepy(bar_staged, k & 1u);
                        if (tile < a.n_tiles) {
                            ptx::bulk_s2g(a.dz + static_cast<size_t>(tile) * a.dz_tile_bytes + dl.dz_off(g), smem + kB2OffS, bytes);
                            ptx::bulk_commit();
                            ptx::bulk_wait_read<0>();
                        }
                        ptx::mbar_arrive(bar_free);
                    }
                }
            }
            ptx::bulk_wait_all<0>();
        }
        __syncwarp();
    }
    } else {
        // ============================ epilogue warps: thread = (row, column quarter) ================
        ptx::setmaxnreg_inc<kB2RegsEpi>();
        const int quad_l = warp & 3;                        // TMEM lane quadrant this warp may read
        const int cq = (warp - kB2EpiWarp0) >> 2;           // column quarter of every layer this thread converts
        const int row = quad_l * 32 + lane;
        const int et = threadIdx.x - kB2EpiWarp0 * 32;      // 0..511
        const uint32_t tlane = tmem + (static_cast<uint32_t>(quad_l * 32) << 16);
        const float* cst = reinterpret_cast<const float*>(a.packed + fs.w_bytes());
        const SaveLayout sl{fs.kx, transient};
        const DzLayout dl{transient};
        const int C = transient ? 9 : 4;
        const float scale = dz_scale_from_amax_bits(__ldg(&a.hdr->amax_bits));
        const float inv_scale = 1.0f / scale;               // both exact powers of two
        const uint32_t rowx = static_cast<uint32_t>((row >> 3) * 1024 + (row & 7) * 128) | static_cast<uint32_t>((row & 7) << 4);
        const uint32_t smem_s = ptx::smem_u32(smem), acc_s = ptx::smem_u32(bar_acc);
        const uint32_t act_leader = ptx::mapa(ptx::smem_u32(bar_act), 0);
        uint32_t acc_par = 0, step_k = 0;
        const uint32_t s8row = kB2Stage8 ? smem_s + kB2OffS + row * 32 : 0u;     // this row in group 0 of block 0 of the staged image
        auto hand_over = [&](int t, bool feeds_mma) {
            ptx::fence_proxy_async_smem();
            ptx::tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) {
                if (feeds_mma) { if (rank == 0) ptx::mbar_arrive(&bar_act[t]); else ptx::mbar_arrive_cluster(act_leader + 8 * t); }
                if (kB2Stage8) ptx::mbar_arrive(bar_staged);
            }
        };
        auto mword = [](uint32_t e_hi, uint32_t e_lo) { return (e_hi << 16) | e_lo; };
        // ReLU gate entries of this thread's columns for step (g, t) of a quad (forward layout [sub][row] x 4 uint16, entry q of
        // sub-block `sub` = columns 64 q + 16 sub ..+15).  256-wide layers: my columns [64 cq, +64) = forward quarter cq, subs
        // 0..3; 128-wide layers (and each half of layer 9): my columns [32 cq, +32) = quarter cq / 2, subs 2 (cq & 1), +1.
        // Loaded ONE STEP AHEAD (the masks come from DRAM: loaded at the top of their own step, the first use of the
        // values -- hoisted in front of the accumulator wait by the compiler -- exposed ~2 500 cycles per step).
        // (32-bit loads of the word that holds the entry, extracted only behind the accumulator wait: a 16-bit load is
        // followed at once by its zero-extension mask, i.e. waited for on the spot -- 11 % of all stall samples, profiles/r02s)
        auto load_gates = [&](int quad, int g, int t, uint32_t& w0, uint32_t& w1, uint32_t& w2, uint32_t& w3) {
            w0 = w1 = w2 = w3 = 0u;
            const int tile = 4 * quad + 2 * t + static_cast<int>(rank);
            if (g == 8 || tile >= a.n_tiles) return;
            const uint32_t* mp = reinterpret_cast<const uint32_t*>(a.saved + static_cast<size_t>(tile) * a.save_tile_bytes + sl.mask_off(g, nl)) + row * 2;
            if (g <= 7) {
                const int h = cq >> 1;
                w0 = __ldg(mp + h); w1 = __ldg(mp + 256 + h); w2 = __ldg(mp + 512 + h); w3 = __ldg(mp + 768 + h);
            } else {
                const int sb = 2 * (cq & 1);
                w0 = __ldg(mp + sb * 256); w1 = __ldg(mp + (sb + 1) * 256);
                if (g == 9 && transient) { w2 = __ldg(mp + sb * 256 + 1); w3 = __ldg(mp + (sb + 1) * 256 + 1); }
            }
        };
        for (int quad = quad0; quad < n_quads; quad += quad_step) {
            uint32_t n0, n1, n2, n3;                            // gates of the coming step
            load_gates(quad, nl - 1, 0, n0, n1, n2, n3);
            for (int g = nl - 1; g >= 0; --g) {
#pragma unroll 1
                for (int t = 0; t < 2; ++t) {
                    NWB_B2TRACE(et == 0, g, t, 6);
                    uint32_t e0 = n0, e1 = n1, e2 = n2, e3 = n3;       // raw words until the accumulator wait is over
                    if (t == 0) load_gates(quad, g, 1, n0, n1, n2, n3);
                    else if (g > 0) load_gates(quad, g - 1, 0, n0, n1, n2, n3);
                    const int tile = 4 * quad + 2 * t + static_cast<int>(rank);
                    const bool ghost = tile >= a.n_tiles;       // protocol only: rows invalid, nothing stored
                    const long long p0 = static_cast<long long>(tile) * kTileM;
                    const long long p = p0 + row;
                    const bool valid = p < a.n_points;
                    const float* y = a.raw + (valid ? p : 0) * C;
                    const float* dy = a.d_raw + (valid ? p : 0) * C;
                    const uint8_t* save_tile = a.saved + static_cast<size_t>(ghost ? 0 : tile) * a.save_tile_bytes;
                    uint8_t* dz_tile = a.dz + static_cast<size_t>(ghost ? 0 : tile) * a.dz_tile_bytes;
                    const uint32_t sD_s = smem_s + kB2OffD + t * 65536;
                    const uint32_t tacc = tlane + t * 256;

                    if (g == nl - 1 && cq == 0) {
                        // ---- once per tile: head-bias gradients and the dpre block (e5m2, three-term expansion, one
                        // 32-column group: all the wgrad kernel's head units read) ----
                        float dp[9];
#pragma unroll
                        for (int i = 0; i < 9; ++i) dp[i] = 0.0f;
                        if (valid) {
                            dp[0] = dp_sigma(y, dy);
#pragma unroll
                            for (int i = 0; i < 3; ++i) dp[1 + i] = dp_sig(y, dy, i);
                            if (transient) {
                                dp[4] = dp_sp(y, dy, 7);
#pragma unroll
                                for (int i = 0; i < 3; ++i) dp[5 + i] = dp_sig(y, dy, 4 + i);
                                dp[8] = dp_sp(y, dy, 8);
                            }
                        }
#pragma unroll
                        for (int i = 0; i < 9; ++i) {
                            const float s = warp_sum(dp[i]);
                            if (lane == 0 && s != 0.0f) atomicAdd(&s_hb[i], s);
                            dp[i] *= scale;
                        }
                        if (!ghost) {
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
                    }

                    // head weights of this layer's head term, lane-distributed over this thread's slice(s); head cotangents
                    float hwl[5], dph[5];
                    if (g == 12) {
#pragma unroll
                        for (int h = 0; h < 5; ++h) hwl[h] = __ldg(cst + ConstOff::tsigma_w + h * 128 + 32 * cq + lane);
                        dph[0] = valid ? dp_sp(y, dy, 7) * scale : 0.0f;
#pragma unroll
                        for (int i = 0; i < 3; ++i) dph[1 + i] = valid ? dp_sig(y, dy, 4 + i) * scale : 0.0f;
                        dph[4] = valid ? dp_sp(y, dy, 8) * scale : 0.0f;
                    } else if (g == 9) {
#pragma unroll
                        for (int h = 0; h < 3; ++h) { hwl[h] = __ldg(cst + ConstOff::rgb_w + h * 128 + 32 * cq + lane); dph[h] = valid ? dp_sig(y, dy, h) * scale : 0.0f; }
                    } else if (g == 7) {
                        hwl[0] = __ldg(cst + ConstOff::sigma_w + 64 * cq + lane);
                        hwl[1] = __ldg(cst + ConstOff::sigma_w + 64 * cq + 32 + lane);
                        dph[0] = valid ? dp_sigma(y, dy) * scale : 0.0f;
                    }
                    NWB_B2TRACE(et == 0, g, t, 5);
                    if (g < nl - 1) {
                        ptx::mbar_wait_sleepy_a(acc_s + 8 * t, (acc_par >> t) & 1u);
                        acc_par ^= 1u << t;
                        ptx::tc_fence_after_sync();
                    }
                    NWB_B2TRACE(et == 0, g, t, 2);
                    {   // first use of the gate words, pinned behind the wait: entry = 16 bits at (g <= 7 ? cq & 1 : cq >> 1)
                        asm volatile("" : "+r"(e0), "+r"(e1), "+r"(e2), "+r"(e3));
                        const uint32_t sh = 16u * static_cast<uint32_t>(g <= 7 ? (cq & 1) : (cq >> 1));
                        e0 = (e0 >> sh) & 0xFFFFu; e1 = (e1 >> sh) & 0xFFFFu; e2 = (e2 >> sh) & 0xFFFFu; e3 = (e3 >> sh) & 0xFFFFu;
                    }
                    if (kB2Stage8) {                            // the previous step's image has left the staging buffer
                        if (step_k > 0) ptx::mbar_wait_sleepy(bar_free, (step_k - 1) & 1u);
                        ++step_k;
                    }
                    if (g == 8) epi_bar16();                    // every warp is done with layer 9's d_ray_bias sums over this image
                    // this row in group 0 of block 0 / block 1 of the layer's e5m2 image
#ifdef NWB_DGRAD_NOSTORE        // timing experiment only: no dZ images are written
                    uint8_t* g8a = nullptr;
                    uint8_t* g8b = nullptr;
#else
                    uint8_t* g8a = (ghost || kB2Stage8) ? nullptr : dz_tile + dl.dz_off(g) + row * 32;
                    uint8_t* g8b = (ghost || kB2Stage8) ? nullptr : g8a + 16384;
#endif
                    const uint32_t s8a = s8row, s8b = s8row ? s8row + 16384 : 0u;
                    if (g >= 10) {
                        const uint32_t blk = sD_s + (cq >> 1) * 16384;
                        if (g == 12) bwd_slice<false, 5, true>(0, hwl, dph, mword(e0, e1), blk, rowx, 4 * (cq & 1), g8a, cq, row, s8a);
                        else bwd_slice<true, 0, true>(tacc + 32 * cq, nullptr, nullptr, mword(e0, e1), blk, rowx, 4 * (cq & 1), g8a, cq, row, s8a);
                    } else if (g == 9) {
                        bwd_slice<false, 3, true>(0, hwl, dph, mword(e0, e1), sD_s + (cq >> 1) * 16384, rowx, 4 * (cq & 1), g8a, cq, row, s8a);
                        if (transient)
                            bwd_slice<true, 0, true>(tacc + 128 + 32 * cq, nullptr, nullptr, mword(e2, e3), sD_s + (2 + (cq >> 1)) * 16384, rowx, 4 * (cq & 1),
                                                     g8b, cq, row, s8b);
                    } else {
                        uint8_t* g8 = (cq >> 1) ? g8b : g8a;
                        const uint32_t s8 = (cq >> 1) ? s8b : s8a;
                        const uint32_t blk = sD_s + cq * 16384;
                        const uint32_t tcol = tacc + 64 * cq;
                        const int ps0 = 2 * (cq & 1);
                        if (g == 8) {
                            bwd_slice<true, 0, false>(tcol, nullptr, nullptr, 0u, blk, rowx, 0, g8, ps0, row, s8);
                            bwd_slice<true, 0, false>(tcol + 32, nullptr, nullptr, 0u, blk, rowx, 4, g8, ps0 + 1, row, s8);
                        } else if (g == 7) {
                            bwd_slice<true, 1, true>(tcol, hwl, dph, mword(e0, e1), blk, rowx, 0, g8, ps0, row, s8);
                            bwd_slice<true, 1, true>(tcol + 32, hwl + 1, dph, mword(e2, e3), blk, rowx, 4, g8, ps0 + 1, row, s8);
                        } else {
                            bwd_slice<true, 0, true>(tcol, nullptr, nullptr, mword(e0, e1), blk, rowx, 0, g8, ps0, row, s8);
                            bwd_slice<true, 0, true>(tcol + 32, nullptr, nullptr, mword(e2, e3), blk, rowx, 4, g8, ps0 + 1, row, s8);
                        }
                    }
                    NWB_B2TRACE(et == 0, g, t, 3);
                    hand_over(t, g >= 1);
                    NWB_B2TRACE(et == 0, g, t, 4);
                    if (g == 9) {
                        // per-ray cotangent of the dir/appearance/transient bias: sum of dZ_9 over the ray's samples, after the
                        // handoff, next to layer 9's MMAs (both only read the image).  All 16 warps must have written their
                        // slices first: the sums read every row.  One column per thread.
                        epi_bar16();
                        const int ncols = transient ? 256 : 128;
                        const int c = et;
                        if (c < ncols && !ghost) {
                            const uint32_t colp = sD_s + (c >> 6) * 16384 + (c & 7) * 2;
                            const int chunk = (c & 63) >> 3;
                            long long ray = p0 / a.S;
                            const int nrow = static_cast<int>(min(static_cast<long long>(kTileM), a.n_points - p0));
                            int r = 0;
                            int seg_end = static_cast<int>(min(static_cast<long long>(nrow), (ray + 1) * a.S - p0));
                            while (r < nrow) {                  // one segment = the rows of one ray inside this tile
                                float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
                                for (; r + 8 <= seg_end; r += 8) {
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
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::cluster_sync_all();                        // the partner has finished with my shared memory / TMEM too
    if (warp == 1) ptx::tmem_dealloc_pair<512>(tmem);
    if (threadIdx.x < 9) {
        const int i = threadIdx.x;
        float* dst = (i == 0) ? a.g_head_b[0] : (i < 4) ? a.g_head_b[1] + (i - 1) : (i == 4) ? a.g_head_b[2]
                   : (i < 8) ? a.g_head_b[3] + (i - 5) : a.g_head_b[4];
        if (dst != nullptr && (i < 4 || transient)) atomicAdd(dst, s_hb[i]);
    }
}

}  // namespace nwb
