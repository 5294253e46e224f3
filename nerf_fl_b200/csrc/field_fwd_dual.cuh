// Fused field forward, second generation: CTA PAIRS with TWO TILES IN FLIGHT per CTA.  Included by field_fwd.cu
// (shares FwdArgs, epi_slice16, layer_quarters with the first-generation kernel, which stays as the reference).
//
// Why: in the one-tile kernel the epilogue is the critical resource -- the MMA warp spends
// 50-60 % of its time waiting for activation quarters, because a layer's MMAs cannot start before the previous
// layer's epilogue has produced their operand, and that epilogue cannot start before the previous MMAs are done.
// Here every CTA works on two independent 128-point tiles (slots 0 and 1).  The 16 epilogue warps alternate between
// them -- epilogue(slot 0, layer g), epilogue(slot 1, layer g), epilogue(slot 0, layer g+1), ... -- and while they
// are busy with one slot the tensor pipe runs the other slot's next layer, whose operand is already complete.
//   TMEM     512 columns = 2 slots x 256 fp32 accumulator columns (no ping-pong inside a tile any more)
//   smem     2 x 16 KB encoded xyz, 2 x 64 KB activations (a layer's output overwrites its input IN PLACE: the
//            input is dead once the layer's MMAs have completed), 3 x 16 KB weight ring, 16 KB constants
//   weights  48 KB of ring only feeds two tiles because the kernel runs as CTA pairs: tcgen05.mma.cta_group::2,
//            M = 256 = slot t of both CTAs, each CTA staging HALF of every weight K-block (its N/2 rows).
//            Rank 0 issues; rank 1's MMA warp relays "my half has landed"; completions are multicast.
// One handoff per (slot, layer) in each direction: bar_act[t] (epilogue -> MMA: the slot's operand image is
// complete, arrivals from the epilogue warps of BOTH CTAs on rank 0's barrier) and bar_acc[t] (MMA -> epilogue).
// NWB_MLP_SAVE: behind each handoff every epilogue thread writes its 16 columns of each quarter as e4m3 (16-byte
// stores from registers into the layer's 8-bit image for the weight-gradient kernel) and its ReLU sign-mask words.
#pragma once

namespace nwb {

// waits of the MMA-issuing / relay threads: parked with a suspend-time hint (default) or spinning (-DNWB_FWD_SPIN).
// [B200, profiles/r02f_perf_*.log] 32 768-ray inference chunk: 4.71 ms parked vs 4.91 ms spinning (the spinning thread
// shares a scheduler with four epilogue warps); no difference at 1024 rays.
#ifdef NWB_FWD_SPIN
#define NWB_FWD_WAIT_A(addr, par) ptx::mbar_wait_a(addr, par)
#else
#define NWB_FWD_WAIT_A(addr, par) ptx::mbar_wait_sleepy_a(addr, par)
#endif

// -DNWB_FWD_TRACE: clock64 stamps of cluster 0 / rank 0 during its SECOND quad of tiles (steady state), slot (g*2+t)*8 + k:
//   k = 0 MMA thread sees the operand image, 1 its MMAs are issued, 2 epilogue (warp kEpiWarp0) sees the accumulator,
//   3 its slices are written, 4 handoff issued, 5 save stores / head combine done
#ifdef NWB_FWD_TRACE
#define NWB_DTRACE(cond, g, t, k) do { if ((cond) && a.trace != nullptr && blockIdx.x == 0 && quad == quad0 + quad_step) a.trace[((g) * 2 + (t)) * 8 + (k)] = clock64(); } while (0)
#else
#define NWB_DTRACE(cond, g, t, k) do { } while (0)
#endif

constexpr int kDOffX = 0;                         // 2 x 16 KB
constexpr int kDOffH = 32768;                     // 2 x 64 KB
constexpr int kDStageBytes = 16384;               // this CTA's half of a weight K-block
constexpr int kDNumStages = 3;
constexpr int kDOffW = kDOffH + 2 * 65536;
constexpr int kDOffC = kDOffW + kDNumStages * kDStageBytes;
constexpr int kDOffBar = kDOffC + 16384;
constexpr int kDSmem = kDOffBar + 256;
static_assert(kDSmem <= 232448, "dual-tile forward exceeds the 227 KB shared-memory limit");

// first activation K-block (16 KB slot of the slot's H image) layer g reads, after the optional X block
__device__ __forceinline__ int dual_in_slot0(int g) { return g == 10 ? 2 : 0; }


// ---- one layer's epilogue for one slot, straight-line -------------------------------------------------------------
// The quarter loop is fully unrolled and the layer kinds are template instances (dispatched once per (layer, slot) by
// a switch), shared memory is addressed by 32-bit shared-space addresses: ncu counted 293 warp-instructions per
// (warp, quarter) in the runtime-dispatched version (profiles/r02e_*), a third of them register moves, zeroing and
// generic-address tests -- the kernel issues on 60 % of its scheduler slots, so instruction count is time.
//   tacc    TMEM address of this thread's lane quadrant, slot and 16-column slice (quarter q adds 64 columns)
//   bias_s  shared address of bias[16 sub] (quarter q adds 256 B); kRayBias: rbias = global per-ray bias + 16 sub
//   hw_s    shared address of head 0's weights [16 sub] (head h adds 512 B -- every multi-head block has ld 128 --, quarter
//           q adds 256 B); heads read quarters q < kHeadQ only
//   img_s   shared address of the slot's activation image (quarter q = K-block q, 16 KB apart); off0 / off1 = this
//           thread's two 16-byte chunks inside a K-block
//   kSave:  p8[q] = the 16 activations as e4m3 (ReLU applied by the conversion); mlo / mhi = sign-mask entries of
//           quarters 0-1 / 2-3 (16 bits each, relu_mask_bit order)
template <bool kRelu, int kHeads, int kHeadQ, int kNQ, bool kRayBias, bool kSave>
__device__ __forceinline__ void dual_layer_epi(uint32_t tacc, uint32_t bias_s, const float* __restrict__ rbias, uint32_t hw_s,
                                               uint32_t img_s, uint32_t off0, uint32_t off1, float (&hacc)[5], uint4 (&p8)[4],
                                               uint32_t& mlo, uint32_t& mhi) {
#ifdef NWB_FWD_LD_PAIRS
    // inference: the accumulator slices of two quarters are requested together (tcgen05.ld costs ~230 cycles from issue to
    // wait::ld alone, ~160 each in pairs: tools/csrc/tmem_probe.cu); the saving build has no registers to spare
    uint32_t vn[16];
#endif
#pragma unroll
    for (int q = 0; q < kNQ; ++q) {
        uint32_t v[16];
#ifdef NWB_FWD_LD_PAIRS
        if (!kSave) {
            if ((q & 1) == 0) {
                ptx::tmem_ld16(tacc + q * 64, v);
                if (q + 1 < kNQ) ptx::tmem_ld16(tacc + (q + 1) * 64, vn);
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = vn[i];
            }
        } else
#endif
        ptx::tmem_ld16(tacc + q * 64, v);
        float4 b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (kRayBias) b[i] = __ldg(reinterpret_cast<const float4*>(rbias + q * 64) + i);
            else b[i] = ptx::lds_f4(bias_s + q * 256 + i * 16);
        }
#ifdef NWB_FWD_LD_PAIRS
        if (kSave || (q & 1) == 0)
#endif
        ptx::tmem_ld_wait();
        float x[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            ptx::fadd2(x[4 * i], x[4 * i + 1], __uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]), b[i].x, b[i].y);
            ptx::fadd2(x[4 * i + 2], x[4 * i + 3], __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]), b[i].z, b[i].w);
        }
        uint32_t pk[8];
        if (kRelu && !kSave) {
            // inference: ReLU folded into the bf16 conversion (max(rn(x), 0) == rn(max(x, 0)))
#pragma unroll
            for (int j = 0; j < 8; ++j) pk[j] = ptx::pack_bf16_relu(x[2 * j], x[2 * j + 1]);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) pk[j] = pack_bf16(x[2 * j], x[2 * j + 1]);
        }
        if (kSave) {
            p8[q].x = ptx::pack4_e4m3<kRelu>(x[0], x[1], x[2], x[3]);    p8[q].y = ptx::pack4_e4m3<kRelu>(x[4], x[5], x[6], x[7]);
            p8[q].z = ptx::pack4_e4m3<kRelu>(x[8], x[9], x[10], x[11]);  p8[q].w = ptx::pack4_e4m3<kRelu>(x[12], x[13], x[14], x[15]);
            if (kRelu) {
                // sign bits of the packed pre-activations (bit 15 = column 2j, bit 31 = column 2j+1) -> one 16-bit entry,
                // see epi_slice16 / relu_mask_bit
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
                float a = hacc[h];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 t = ptx::lds_f4(hw_s + h * 512 + q * 256 + i * 16);
                    a = fmaf(kRelu ? fmaxf(x[4 * i], 0.0f) : x[4 * i], t.x, a);
                    a = fmaf(kRelu ? fmaxf(x[4 * i + 1], 0.0f) : x[4 * i + 1], t.y, a);
                    a = fmaf(kRelu ? fmaxf(x[4 * i + 2], 0.0f) : x[4 * i + 2], t.z, a);
                    a = fmaf(kRelu ? fmaxf(x[4 * i + 3], 0.0f) : x[4 * i + 3], t.w, a);
                }
                hacc[h] = a;
            }
        }
        ptx::sts_u4(img_s + q * 16384 + off0, pk[0], pk[1], pk[2], pk[3]);
        ptx::sts_u4(img_s + q * 16384 + off1, pk[4], pk[5], pk[6], pk[7]);
    }
}

// residual chunk of a head-feeding layer, recomputed behind the handoff from the bf16 image (see residual_chunk)
__device__ __forceinline__ uint4 residual_chunk_s(uint32_t blk_s, uint32_t off0, uint32_t off1, const uint4& p8) {
    const uint4 b0 = ptx::lds_u4(blk_s + off0), b1 = ptx::lds_u4(blk_s + off1);
    auto lo = [](uint32_t w) { return __uint_as_float(w << 16); };
    auto hi = [](uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); };
    uint4 l8;
    l8.x = ptx::residual4_e4m3<false>(p8.x, lo(b0.x), hi(b0.x), lo(b0.y), hi(b0.y), kLoScale);
    l8.y = ptx::residual4_e4m3<false>(p8.y, lo(b0.z), hi(b0.z), lo(b0.w), hi(b0.w), kLoScale);
    l8.z = ptx::residual4_e4m3<false>(p8.z, lo(b1.x), hi(b1.x), lo(b1.y), hi(b1.y), kLoScale);
    l8.w = ptx::residual4_e4m3<false>(p8.w, lo(b1.z), hi(b1.z), lo(b1.w), hi(b1.w), kLoScale);
    return l8;
}

template <bool kSave>
__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(kFwdMaxRegs)
field_fwd_dual_kernel(const FwdArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sW = smem + kDOffW;
    uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem + kDOffBar);  // [3] producer (+ partner relay) -> MMA
    uint64_t* bar_empty = bar_full + kDNumStages;                       // [3] MMA -> producers of both CTAs
    uint64_t* bar_acc = bar_empty + kDNumStages;                        // [2] MMA -> epilogue, per slot
    uint64_t* bar_act = bar_acc + 2;                                    // [2] epilogues of both CTAs -> MMA, per slot
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_act + 2);
    uint32_t* s_woff = tmem_slot + 2;                                   // [kMaxG]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = ptx::cluster_ctarank();
    const int quad0 = static_cast<int>(ptx::cluster_id_x()), quad_step = static_cast<int>(ptx::cluster_count_x());
    const int n_quads = (a.n_tiles + 3) >> 2;          // a cluster step covers 4 tiles: tile = 4*quad + 2*slot + rank
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
        for (int s = 0; s < kDNumStages; ++s) { ptx::mbar_init(&bar_full[s], rank == 0 ? 2 : 1); ptx::mbar_init(&bar_empty[s], 1); }
        for (int t = 0; t < 2; ++t) {
            ptx::mbar_init(&bar_acc[t], 1);
            ptx::mbar_init(&bar_act[t], 2 * kEpiWarps);
        }
        ptx::fence_mbar_init();
        for (int g = 0; g < kMaxG; ++g) s_woff[g] = static_cast<uint32_t>(fs.w_off(g));
    }
    {
        const float4* src = reinterpret_cast<const float4*>(a.packed + fs.w_bytes());
        float4* dst = reinterpret_cast<float4*>(smem + kDOffC);
        for (int i = threadIdx.x; i < ConstOff::count / 4; i += kFwdThreads) dst[i] = __ldg(src + i);
    }
    if (warp == 1) ptx::tmem_alloc_pair<512>(tmem_slot);
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::cluster_sync_all();
    ptx::tc_fence_after_sync();
    const uint32_t tmem = *tmem_slot;

    if (warp < kEpiWarp0) {
    ptx::setmaxnreg_dec<kFwdRegsWg0>();
    if (warp == 0) {
        // ============================ weight producer ============================================
        // per (layer, slot, K-block): my N/2 rows x 64 k of the packed slab image, one bulk copy
        if (lane == 0) {
            uint32_t s = 0, ph = 0;
            for (int quad = quad0; quad < n_quads; quad += quad_step) {
                for (int g = 0; g < nl; ++g) {
                    const int KB = fs.layer_kb(g), PN = fs.packed_n(g);
                    const uint32_t bytes = static_cast<uint32_t>(fs.layer_n(g, a.mode)) * 64u;
                    const uint8_t* wsrc = a.packed + s_woff[g] + rank * bytes;
                    for (int t = 0; t < 2; ++t) {
                        for (int kb = 0; kb < KB; ++kb) {
                            ptx::mbar_wait_sleepy(&bar_empty[s], ph ^ 1u);
                            ptx::mbar_arrive_expect_tx(&bar_full[s], bytes);
                            ptx::bulk_g2s(sW + s * kDStageBytes, wsrc + static_cast<size_t>(kb) * PN * 128, bytes, &bar_full[s]);
                            if (++s == kDNumStages) { s = 0; ph ^= 1u; }
                        }
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ============================ MMA issuer (rank 0) / weight relay (rank 1) ==================
        if (rank == 0) {
            constexpr uint32_t idesc256 = ptx::umma_idesc_bf16(256, 256, false, false);
            constexpr uint32_t idesc128 = ptx::umma_idesc_bf16(256, 128, false, false);
            const uint32_t w_lo = ptx::umma_desc_lo(ptx::smem_u32(sW));
            const uint32_t full_a = ptx::smem_u32(bar_full), empty_a = ptx::smem_u32(bar_empty);
            const uint32_t acc_a = ptx::smem_u32(bar_acc), act_a = ptx::smem_u32(bar_act);
            uint32_t s = 0, ph = 0;
            uint32_t act_par = 0;                            // bit t = parity of the next phase of bar_act[t]
            for (int quad = quad0; quad < n_quads; quad += quad_step) {
                for (int g = 0; g < nl; ++g) {
                    const int KB = fs.layer_kb(g);
                    const uint32_t idesc = (fs.layer_n(g, a.mode) == 256) ? idesc256 : idesc128;
                    for (int t = 0; t < 2; ++t) {
                        const uint32_t x_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + kDOffX + t * 16384));
                        const uint32_t h_lo = ptx::umma_desc_lo(ptx::smem_u32(smem + kDOffH + t * 65536));
                        const uint32_t dcol = tmem + t * 256;
                        // the slot's operand image (X for layer 0, the previous layer's output otherwise) is complete in
                        // both CTAs, and its previous accumulator has been read out
                        NWB_FWD_WAIT_A(act_a + 8 * t, (act_par >> t) & 1u);
                        act_par ^= 1u << t;
                        ptx::tc_fence_after_sync();
                        NWB_DTRACE(lane == 0, g, t, 0);
                        int hs = dual_in_slot0(g);
                        for (int kb = 0; kb < KB; ++kb) {
                            const bool is_x = (g == 0) || (g == 4 && kb == 0);
                            const uint32_t a_lo = is_x ? x_lo : h_lo + (hs++) * 1024;
                            NWB_FWD_WAIT_A(full_a + 8 * s, ph);
                            ptx::tc_fence_after_sync();
                            const uint32_t b_lo = w_lo + s * (kDStageBytes >> 4);
                            if (ptx::elect_one()) {
                                ptx::umma2_bf16_lo(dcol, a_lo, b_lo, idesc, kb != 0);
                                ptx::umma2_bf16_lo(dcol, a_lo + 2, b_lo + 2, idesc, 1);
                                ptx::umma2_bf16_lo(dcol, a_lo + 4, b_lo + 4, idesc, 1);
                                ptx::umma2_bf16_lo(dcol, a_lo + 6, b_lo + 6, idesc, 1);
                                ptx::umma2_commit_both(empty_a + 8 * s);
                                if (kb == KB - 1) ptx::umma2_commit_both(acc_a + 8 * t);
                            }
                            __syncwarp();
                            if (++s == kDNumStages) { s = 0; ph ^= 1u; }
                        }
                        NWB_DTRACE(lane == 0, g, t, 1);
                    }
                }
            }
        } else if (lane == 0) {
            const uint32_t full_a = ptx::smem_u32(bar_full);
            const uint32_t full_leader = ptx::mapa(full_a, 0);
            uint32_t s = 0, ph = 0;
            for (int quad = quad0; quad < n_quads; quad += quad_step) {
                for (int g = 0; g < nl; ++g) {
                    const int n = 2 * fs.layer_kb(g);
                    for (int i = 0; i < n; ++i) {
                        NWB_FWD_WAIT_A(full_a + 8 * s, ph);
                        ptx::mbar_arrive_cluster(full_leader + 8 * s);
                        if (++s == kDNumStages) { s = 0; ph ^= 1u; }
                    }
                }
            }
        }
        __syncwarp();
    }
    } else {
        // ============================ epilogue warps (four threads per point) ====================
        ptx::setmaxnreg_inc<kFwdRegsEpi>();
        const int ew = warp - kEpiWarp0;                   // 0..15
        const int quad_l = warp & 3;                       // TMEM lane quadrant this warp may read
        const int sub = ew >> 2;                           // 16-column slice of every 64-column quarter
        const int row = quad_l * 32 + lane;
        const uint32_t tlane = tmem + (static_cast<uint32_t>(quad_l * 32) << 16);
        const float* cst = reinterpret_cast<const float*>(smem + kDOffC);
        const SaveLayout sl{fs.kx, transient};
        const uint32_t act_leader = ptx::mapa(ptx::smem_u32(bar_act), 0);
        uint32_t acc_par = 0;                              // bit t = parity of the next phase of bar_acc[t]
        const uint32_t smem_s = ptx::smem_u32(smem), cst_s = smem_s + kDOffC + sub * 64, acc_s = ptx::smem_u32(bar_acc);
        const uint32_t off0 = ptx::image_chunk_off(row, 2 * sub), off1 = ptx::image_chunk_off(row, 2 * sub + 1);   // bf16 image chunks
        // this thread's 16-byte chunk in the 8-bit image of a layer: quarter q = columns [64 q + 16 sub, +16) = chunk 4 q + sub
        const uint32_t so0 = image8_chunk_off(row, sub), so1 = image8_chunk_off(row, 4 + sub);                        // quarters 0, 1 (2, 3: + 16 KB)
        auto hand_over = [&](int t, bool feeds_mma) {      // after this thread's shared-memory writes of the image
            ptx::fence_proxy_async_smem();
            ptx::tc_fence_before_sync();
            __syncwarp();
            if (lane == 0 && feeds_mma) { if (rank == 0) ptx::mbar_arrive(&bar_act[t]); else ptx::mbar_arrive_cluster(act_leader + 8 * t); }
        };
        for (int quad = quad0; quad < n_quads; quad += quad_step) {
            // per-slot state lives in scalar registers selected by t (arrays indexed by the runtime slot number would sit
            // in local memory: their reloads queue behind the epilogue's global stores -- a third of all stall samples of
            // the first 8-bit build, profiles/r02b); what is cheap to recompute from the tile number is recomputed
            int ray0 = 0, ray1 = 0;
            // ---- encoded-xyz operand tiles of both slots: this thread packs columns [16*sub, 16*sub+16) of its row ----
#pragma unroll 1
            for (int t = 0; t < 2; ++t) {
                const int tile = 4 * quad + 2 * t + static_cast<int>(rank);
                const bool ghost_t = tile >= a.n_tiles;
                const long long p_t = static_cast<long long>(tile) * kTileM + row;
                const bool valid_t = p_t < a.n_points;
                const int ray_t = valid_t ? static_cast<int>(p_t / a.S) : 0;
                if (t == 0) ray0 = ray_t; else ray1 = ray_t;
                uint8_t* save_tile_t = save ? a.saved + static_cast<size_t>(tile) * a.save_tile_bytes : nullptr;
                uint8_t* sX = smem + kDOffX + t * 16384;
                float e[64];
#pragma unroll
                for (int c = 0; c < 64; ++c) e[c] = 0.0f;
                if (valid_t) {
                    if (a.x_enc != nullptr) {
                        const float* xr = a.x_enc + p_t * fs.in_xyz;
#pragma unroll
                        for (int c = 0; c < 64; ++c)
                            if (c < fs.in_xyz && (c >> 4) == sub) e[c] = __ldg(xr + c);
                    } else {
                        const float* rr = a.rays + static_cast<long long>(ray_t) * a.ray_stride;
                        const float zz = __ldg(a.z + p_t);
                        float sn[3], cs[3];
#pragma unroll
                        for (int i = 0; i < 3; ++i) {
                            e[i] = __fadd_rn(__ldg(rr + i), __fmul_rn(__ldg(rr + 3 + i), zz));   // rendering.py:261
                            sincosf(e[i], &sn[i], &cs[i]);
                        }
                        // octave k+1 from octave k by the double-angle identities (see the first-generation kernel)
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
                // X_t doubles as the head scratch of the previous quad (its last combine ended with a barrier)
                *reinterpret_cast<uint4*>(sX + ptx::image_chunk_off(row, 2 * sub)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                *reinterpret_cast<uint4*>(sX + ptx::image_chunk_off(row, 2 * sub + 1)) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                hand_over(t, true);
                // saved 8-bit image of the encoded xyz: columns [16 sub, 16 sub + 16) = chunk `sub` of this row's line
                if (save && !ghost_t) *reinterpret_cast<uint4*>(save_tile_t + sl.x_off() + so0) = x8;
            }

            float sg0 = 0.0f, sg1 = 0.0f, c00 = 0.0f, c01 = 0.0f, c02 = 0.0f, c10 = 0.0f, c11 = 0.0f, c12 = 0.0f;   // sigma / rgb of slot 0 / 1
            for (int g = 0; g < nl; ++g) {
                const bool last = (g == nl - 1);
#pragma unroll 1
                for (int t = 0; t < 2; ++t) {
                    const int tile = 4 * quad + 2 * t + static_cast<int>(rank);
                    const uint32_t img_s = smem_s + kDOffH + t * 65536;
                    const uint32_t part_s = smem_s + kDOffX + t * 16384 + row * 4;     // [3][5][128] head partial sums: X_t is
                                                                                         // dead once layer 4 is done
                    const uint32_t tacc = tlane + t * 256 + sub * 16;
                    ptx::mbar_wait_sleepy_a(acc_s + 8 * t, (acc_par >> t) & 1u);
                    acc_par ^= 1u << t;
                    ptx::tc_fence_after_sync();
                    NWB_DTRACE(threadIdx.x == kEpiWarp0 * 32, g, t, 2);
                    float hacc[5] = {0, 0, 0, 0, 0};
                    uint4 p8[4];                           // e4m3 chunks of this thread's quarters, kept until the handoff is out
                    uint32_t mlo = 0, mhi = 0;             // sign-mask entries (one 8-byte store per thread after the handoff)
                    if (g < 7) {
                        dual_layer_epi<true, 0, 0, 4, false, kSave>(tacc, cst_s + (ConstOff::trunk_b + g * 256) * 4, nullptr, 0, img_s, off0, off1, hacc, p8, mlo, mhi);
                    } else if (g == 7) {
                        dual_layer_epi<true, 1, 4, 4, false, kSave>(tacc, cst_s + (ConstOff::trunk_b + 7 * 256) * 4, nullptr, cst_s + ConstOff::sigma_w * 4,
                                                                    img_s, off0, off1, hacc, p8, mlo, mhi);
                    } else if (g == 8) {
                        dual_layer_epi<false, 0, 0, 4, false, kSave>(tacc, cst_s + ConstOff::final_b * 4, nullptr, 0, img_s, off0, off1, hacc, p8, mlo, mhi);
                    } else if (g == 9) {
                        // per-ray bias already contains dir_encoding.0.bias / transient_encoding.0.bias; rgb heads read the dir half
                        const float* rb = a.ray_bias + static_cast<long long>(t ? ray1 : ray0) * 256 + sub * 16;
                        if (transient) dual_layer_epi<true, 3, 2, 4, true, kSave>(tacc, 0, rb, cst_s + ConstOff::rgb_w * 4, img_s, off0, off1, hacc, p8, mlo, mhi);
                        else dual_layer_epi<true, 3, 2, 2, true, kSave>(tacc, 0, rb, cst_s + ConstOff::rgb_w * 4, img_s, off0, off1, hacc, p8, mlo, mhi);
                    } else if (g < 12) {
                        dual_layer_epi<true, 0, 0, 2, false, kSave>(tacc, cst_s + (ConstOff::tenc_b + (g - 10) * 128) * 4, nullptr, 0, img_s, off0, off1, hacc, p8, mlo, mhi);
                    } else {
                        // transient heads: hw rows = [tsigma_w | trgb_w(3) | tbeta_w], contiguous, ld 128
                        dual_layer_epi<true, 5, 2, 2, false, kSave>(tacc, cst_s + (ConstOff::tenc_b + 2 * 128) * 4, nullptr, cst_s + ConstOff::tsigma_w * 4,
                                                                    img_s, off0, off1, hacc, p8, mlo, mhi);
                    }
                    NWB_DTRACE(threadIdx.x == kEpiWarp0 * 32, g, t, 3);
                    hand_over(t, !last);
                    NWB_DTRACE(threadIdx.x == kEpiWarp0 * 32, g, t, 4);
#ifdef NWB_FWD_NOSTORE          // timing experiment only: conversions done, nothing written
                    if (save && tile < a.n_tiles && p8[0].x == 0x12345678u && mlo == 0x9abcdefu) {
#else
                    if (save && tile < a.n_tiles) {
#endif
                        // saved 8-bit image: quarter q = columns [64q, 64q+64) -> block q/2, chunks 4(q%2) .. +3; this thread's
                        // 16 columns are chunk 4(q%2) + sub of its row's line
                        const bool four = (g <= 8) || (g == 9 && transient);
                        uint8_t* save_tile_t = a.saved + static_cast<size_t>(tile) * a.save_tile_bytes;
                        uint8_t* ln = save_tile_t + sl.act_off(g);
                        *reinterpret_cast<uint4*>(ln + so0) = p8[0];
                        *reinterpret_cast<uint4*>(ln + so1) = p8[1];
                        if (four) {
                            *reinterpret_cast<uint4*>(ln + 16384 + so0) = p8[2];
                            *reinterpret_cast<uint4*>(ln + 16384 + so1) = p8[3];
                        }
                        if (g != 8) *reinterpret_cast<uint2*>(save_tile_t + sl.mask_off(g, nl) + sub * 1024 + row * 8) = make_uint2(mlo, mhi);
                        if (g == 7 || g == 9 || g == 12) {
                            // residual image of a layer that feeds a head: same place as in the image proper (of layer 9
                            // only the dir half, quarters 0-1)
                            uint8_t* lo = save_tile_t + sl.lo_off(g == 7 ? 0 : (g == 9 ? 1 : 2), nl);
                            *reinterpret_cast<uint4*>(lo + so0) = residual_chunk_s(img_s, off0, off1, p8[0]);
                            *reinterpret_cast<uint4*>(lo + so1) = residual_chunk_s(img_s + 16384, off0, off1, p8[1]);
                            if (g == 7) {
                                *reinterpret_cast<uint4*>(lo + 16384 + so0) = residual_chunk_s(img_s + 2 * 16384, off0, off1, p8[2]);
                                *reinterpret_cast<uint4*>(lo + 16384 + so1) = residual_chunk_s(img_s + 3 * 16384, off0, off1, p8[3]);
                            }
                        }
                    }
                    // ---- heads: combine the four column slices of a row through shared memory (fixed order) ----
                    const int nh_heads = (g == 7) ? 1 : (g == 9 ? 3 : (g == 12 ? 5 : 0));
                    if (nh_heads > 0) {
                        if (sub > 0) {
#pragma unroll
                            for (int h = 0; h < 5; ++h)      // static indices: hacc stays in registers
                                if (h < nh_heads) ptx::sts_f(part_s + ((sub - 1) * 5 + h) * 512, hacc[h]);
                        }
                        epi_bar_all();
                        if (sub == 0) {
#pragma unroll
                            for (int h = 0; h < 5; ++h)
                                if (h < nh_heads)
                                    hacc[h] = ((hacc[h] + ptx::lds_f(part_s + h * 512)) + ptx::lds_f(part_s + (5 + h) * 512)) + ptx::lds_f(part_s + (10 + h) * 512);
                            if (g == 7) {                                                                      // nerf.py:190
                                const float v = softplus_f(hacc[0] + cst[ConstOff::head_b]);
                                if (t == 0) sg0 = v; else sg1 = v;
                            }
                            if (g == 9) {                                                                      // nerf.py:197
                                const float v0 = sigmoid_f(hacc[0] + cst[ConstOff::head_b + 1]);
                                const float v1 = sigmoid_f(hacc[1] + cst[ConstOff::head_b + 2]);
                                const float v2 = sigmoid_f(hacc[2] + cst[ConstOff::head_b + 3]);
                                if (t == 0) { c00 = v0; c01 = v1; c02 = v2; } else { c10 = v0; c11 = v1; c12 = v2; }
                            }
                        }
                        epi_bar_all();                       // scratch is read before anyone rewrites it
                    }
                    NWB_DTRACE(threadIdx.x == kEpiWarp0 * 32, g, t, 5);
                    // ---- tile outputs ---------------------------------------------------------------
                    const long long p_t = static_cast<long long>(tile) * kTileM + row;
                    if (last && p_t < a.n_points && sub == 0) {
                        if (sigma_only) {
                            a.raw_out[p_t] = t ? sg1 : sg0;
                        } else {
                            float* o = a.raw_out + p_t * (transient ? 9 : 4);
                            o[0] = t ? c10 : c00; o[1] = t ? c11 : c01; o[2] = t ? c12 : c02; o[3] = t ? sg1 : sg0;   // nerf.py:198
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
    }
    ptx::tc_fence_before_sync();
    __syncthreads();
    ptx::cluster_sync_all();                               // the partner has finished with my shared memory / TMEM too
    if (warp == 1) ptx::tmem_dealloc_pair<512>(tmem);
}

}  // namespace nwb
