// Per-ray operands of dir_encoding / transient_encoding.0 (models/rendering.py:113-130,
// models/nerf.py:195,203).  The reference repeats dir/appearance/transient rows to every sample
// and concatenates them into the layer input; here the ray-constant part of the product,
//     W[:, 256:] . [PosEmbedding(dir) | a | t]  + b,
// is evaluated once per ray in fp32 and enters the fused MLP as a per-ray bias.  Forward: one
// warp per ray.  Backward: per-ray cotangents -> weight-column / bias / embedding-row gradients.
#include "field.cuh"
#include <algorithm>

namespace nwb {

constexpr int kRcRays = 16;            // rays per block pass
constexpr int kRcThreads = 256;        // thread j = output feature j (0..127 dir_encoding, 128..255 transient_encoding.0)
constexpr int kMaxFeat = 768;

// Forward.  Phase 1: the block builds the feature rows [PosEmbedding(dir) | a | t] of kRcRays rays in shared memory
// (and writes them out for the backward).  Phase 2: thread j keeps 32 weights of its row W[j, 256 + k0 .. +32) in
// REGISTERS and sweeps the block's rays against them (features read as warp-uniform float4 broadcasts), so the
// fp32 weight columns are read from L2 once per block instead of once per ray, and never uncoalesced per FMA.
__global__ void __launch_bounds__(kRcThreads)
ray_consts_fwd_kernel(FieldShape fs, nwb_field_params p, const float* __restrict__ dirs, int dir_stride,
                      int dir_is_encoded, int n_freq_dir, const int64_t* __restrict__ ts,
                      const float* __restrict__ emb_a, const float* __restrict__ emb_t, int n_vocab,
                      const float* __restrict__ a_rows, const float* __restrict__ t_rows, int n_rays,
                      uint32_t mode, float* __restrict__ ray_feat, float* __restrict__ ray_bias,
                      int* __restrict__ err_flag) {
    extern __shared__ __align__(16) float sm[];
    const bool use_t = (mode & NWB_MLP_TRANSIENT) != 0;
    const int nd = fs.in_dir, na = fs.in_a, nt = use_t ? fs.in_t : 0;
    const int nf = nd + na + nt;
    const int ka = nd + na;                                  // contraction length of the dir_encoding half
    const int ka_p = (ka + 3) & ~3, kt_p = (nt + 3) & ~3;   // padded row lengths (float4 reads)
    const int feat_ld = fs.in_dir + fs.in_a + fs.in_t;
    float* sfa = sm;                                         // [kRcRays][ka_p]  dir | a
    float* sft = sfa + kRcRays * ka_p;                       // [kRcRays][kt_p]  t
    const int j = threadIdx.x;
    const bool is_t = j >= kHalf;
    const int jj = is_t ? j - kHalf : j;
    const int K = is_t ? nt : ka;
    const float* wrow = is_t ? (use_t ? p.tenc_w[0] + static_cast<int64_t>(jj) * (kW + fs.in_t) + kW : nullptr)
                             : p.dir_w + static_cast<int64_t>(jj) * (kW + fs.in_dir + fs.in_a) + kW;
    const float bias = is_t ? (use_t ? p.tenc_b[0][jj] : 0.0f) : p.dir_b[jj];
    const float* sf = is_t ? sft : sfa;
    const int sf_ld = is_t ? kt_p : ka_p;

    for (int r0 = blockIdx.x * kRcRays; r0 < n_rays; r0 += gridDim.x * kRcRays) {
        const int nr = min(kRcRays, n_rays - r0);
        // ---- phase 1: feature rows ----
        for (int idx = threadIdx.x; idx < nr * nf; idx += kRcThreads) {
            const int r = idx / nf, e = idx - r * nf;
            const int ray = r0 + r;
            int64_t t_idx = 0;
            if (e >= nd && ((e < ka && a_rows == nullptr) || (e >= ka && t_rows == nullptr))) {
                t_idx = ts[ray];
                if (t_idx < 0 || t_idx >= n_vocab) {
                    atomicExch(err_flag, 1);
                    t_idx = 0;
                }
            }
            float v;
            if (e < nd) {
                if (dir_is_encoded || e < 3) {
                    v = dirs[static_cast<int64_t>(ray) * dir_stride + e];
                } else {
                    const int k = (e - 3) / 6, rr = (e - 3) % 6;
                    const float arg = __fmul_rn(dirs[static_cast<int64_t>(ray) * dir_stride + (rr % 3)],
                                                exp2f(static_cast<float>(k)));
                    v = (rr < 3) ? sinf(arg) : cosf(arg);
                }
                sfa[r * ka_p + e] = v;
            } else if (e < ka) {
                const int q = e - nd;
                v = a_rows ? a_rows[static_cast<int64_t>(ray) * na + q] : emb_a[t_idx * na + q];
                sfa[r * ka_p + e] = v;
            } else {
                const int q = e - ka;
                v = t_rows ? t_rows[static_cast<int64_t>(ray) * nt + q] : emb_t[t_idx * nt + q];
                sft[r * kt_p + q] = v;
            }
            if (ray_feat) ray_feat[static_cast<int64_t>(ray) * feat_ld + e] = v;
        }
        for (int idx = threadIdx.x; idx < nr * (ka_p - ka); idx += kRcThreads) sfa[(idx / (ka_p - ka)) * ka_p + ka + idx % (ka_p - ka)] = 0.0f;
        if (kt_p > nt)
            for (int idx = threadIdx.x; idx < nr * (kt_p - nt); idx += kRcThreads) sft[(idx / (kt_p - nt)) * kt_p + nt + idx % (kt_p - nt)] = 0.0f;
        __syncthreads();
        // ---- phase 2: ray_bias[ray, j] = b[j] + sum_k W[j, 256 + k] feat[ray, k] ----
        float acc[kRcRays];
#pragma unroll
        for (int r = 0; r < kRcRays; ++r) acc[r] = bias;
        for (int k0 = 0; k0 < K; k0 += 32) {
            float w[32];
#pragma unroll
            for (int q = 0; q < 32; ++q) w[q] = (k0 + q < K) ? __ldg(wrow + k0 + q) : 0.0f;
            const int nq4 = min(8, (sf_ld - k0) >> 2);                      // float4 groups left in the padded row
#pragma unroll
            for (int r = 0; r < kRcRays; ++r) {
                if (r < nr) {
                    const float4* fr = reinterpret_cast<const float4*>(sf + r * sf_ld + k0);
                    float a = acc[r];
#pragma unroll
                    for (int q4 = 0; q4 < 8; ++q4) {
                        if (q4 < nq4) {
                            const float4 f = fr[q4];
                            a = fmaf(w[4 * q4], f.x, a); a = fmaf(w[4 * q4 + 1], f.y, a);
                            a = fmaf(w[4 * q4 + 2], f.z, a); a = fmaf(w[4 * q4 + 3], f.w, a);
                        }
                    }
                    acc[r] = a;
                }
            }
        }
#pragma unroll
        for (int r = 0; r < kRcRays; ++r)
            if (r < nr) ray_bias[static_cast<int64_t>(r0 + r) * 256 + j] = (is_t && !use_t) ? 0.0f : acc[r];
        __syncthreads();
    }
}

// backward: one 256-thread block per kRcbRays rays; d_ray_bias (n,256) -> dW[:, 256:], db, d_emb rows.
// Thread j owns output feature j (0..127 dir_encoding, 128..255 transient_encoding.0): it walks the
// block's rays once per 32-column chunk of the per-ray feature row, keeping 32 partial sums of
// dW[j, 256+k] in registers, so each block issues ONE atomic per weight element (n_rays/kRcbRays-way
// contention instead of one atomic per ray group and element).  Embedding-row cotangents
// W[:, 256+nd:]^T g are taken against a shared-memory copy of the weight columns.
constexpr int kRcbRays = 16;
constexpr int kRcbThreads = 256;

__global__ void __launch_bounds__(kRcbThreads)
ray_consts_bwd_kernel(FieldShape fs, nwb_field_params p, nwb_field_grads g, const float* __restrict__ ray_feat,
                      const float* __restrict__ d_ray_bias, const int64_t* __restrict__ ts, int n_vocab,
                      int n_rays, uint32_t mode, float* __restrict__ d_emb_a, float* __restrict__ d_emb_t,
                      float* __restrict__ d_a_rows, float* __restrict__ d_t_rows) {
    extern __shared__ float sm[];
    const bool use_t = (mode & NWB_MLP_TRANSIENT) != 0;
    const int nd = fs.in_dir, na = fs.in_a, nt = use_t ? fs.in_t : 0;
    const int feat_ld = fs.in_dir + fs.in_a + fs.in_t;
    const int ld_dir = kW + fs.in_dir + fs.in_a, ld_t = kW + fs.in_t;
    const int r0 = blockIdx.x * kRcbRays, nr = min(n_rays - r0, kRcbRays);
    if (nr <= 0) return;
    float* sg = sm;                                   // [kRcbRays][256]   d_ray_bias rows
    float* sf = sg + kRcbRays * 256;                  // [kRcbRays][feat_ld] per-ray features
    float* sw = sf + kRcbRays * feat_ld;              // [128][na + nt]    weight columns of the embedding inputs
    for (int i = threadIdx.x; i < nr * 256; i += kRcbThreads) sg[i] = d_ray_bias[static_cast<int64_t>(r0) * 256 + i];
    for (int i = threadIdx.x; i < nr * feat_ld; i += kRcbThreads) sf[i] = ray_feat[static_cast<int64_t>(r0) * feat_ld + i];
    const int ne = na + nt;
    for (int i = threadIdx.x; i < kHalf * ne; i += kRcbThreads) {
        const int j = i / ne, e = i - j * ne;
        sw[i] = (e < na) ? p.dir_w[static_cast<int64_t>(j) * ld_dir + kW + nd + e]
                         : p.tenc_w[0][static_cast<int64_t>(j) * ld_t + kW + (e - na)];
    }
    __syncthreads();
    // (1) weight-column and bias gradients
    {
        const int j = threadIdx.x;                    // 0..255
        const bool is_t = j >= kHalf;
        if (!is_t || use_t) {
            const int jj = is_t ? j - kHalf : j;
            const int k_lo = is_t ? nd + na : 0, k_n = is_t ? nt : nd + na;
            float* wrow = is_t ? g.tenc_w[0] + static_cast<int64_t>(jj) * ld_t + kW
                               : g.dir_w + static_cast<int64_t>(jj) * ld_dir + kW;
            float bsum = 0.0f;
            for (int r = 0; r < nr; ++r) bsum += sg[r * 256 + j];
            atomicAdd((is_t ? g.tenc_b[0] : g.dir_b) + jj, bsum);
            for (int k0 = 0; k0 < k_n; k0 += 32) {
                float acc[32];
#pragma unroll
                for (int q = 0; q < 32; ++q) acc[q] = 0.0f;
                for (int r = 0; r < nr; ++r) {
                    const float gv = sg[r * 256 + j];
                    const float* fr = sf + r * feat_ld + k_lo + k0;
#pragma unroll
                    for (int q = 0; q < 32; ++q)
                        if (k0 + q < k_n) acc[q] = fmaf(gv, fr[q], acc[q]);
                }
#pragma unroll
                for (int q = 0; q < 32; ++q)
                    if (k0 + q < k_n) atomicAdd(wrow + k0 + q, acc[q]);
            }
        }
    }
    // (2) embedding-row gradients: d a = W_dir[:, 256+nd:]^T . g_dir ; d t = W_t0[:, 256:]^T . g_t
    for (int idx = threadIdx.x; idx < nr * ne; idx += kRcbThreads) {
        const int r = idx / ne, e = idx - r * ne;
        const float* gr = sg + r * 256 + (e < na ? 0 : kHalf);
        float acc = 0.0f;
#pragma unroll 8
        for (int j = 0; j < kHalf; ++j) acc = fmaf(sw[j * ne + e], gr[j], acc);
        // an index outside [0, n_vocab) was reported by the forward (err_flag) and gathered nothing there: it scatters
        // nothing here either (the reference device-asserts, README.md:124)
        if (e < na) {
            if (d_a_rows) d_a_rows[static_cast<int64_t>(r0 + r) * na + e] = acc;
            else if (d_emb_a) {
                const int64_t t = ts[r0 + r];
                if (t >= 0 && t < n_vocab) atomicAdd(d_emb_a + t * na + e, acc);
            }
        } else {
            const int k = e - na;
            if (d_t_rows) d_t_rows[static_cast<int64_t>(r0 + r) * nt + k] = acc;
            else if (d_emb_t) {
                const int64_t t = ts[r0 + r];
                if (t >= 0 && t < n_vocab) atomicAdd(d_emb_t + t * nt + k, acc);
            }
        }
    }
}

}  // namespace nwb

using namespace nwb;

extern "C" int nwb_ray_consts_fwd(const nwb_field_desc* d, const nwb_field_params* p, const float* dirs,
                                  int dir_stride, int dir_is_encoded, int n_freq_dir, const int64_t* ts,
                                  const float* emb_a, const float* emb_t, int n_vocab, const float* a_rows,
                                  const float* t_rows, int n_rays, uint32_t mode, float* ray_feat,
                                  float* ray_bias, int* err_flag, void* stream) {
    FieldShape fs;
    int rc = make_shape(d, &fs);
    if (rc != NWB_OK) return rc;
    if (n_rays == 0) return NWB_OK;
    NWB_REQUIRE(p && dirs && ray_bias && err_flag, "nwb_ray_consts_fwd: null pointer");
    NWB_REQUIRE(p->dir_w && p->dir_b, "nwb_ray_consts_fwd: dir_encoding parameters are null");
    NWB_REQUIRE(dir_is_encoded || fs.in_dir == 3 + 6 * n_freq_dir, "nwb_ray_consts_fwd: in_dir != 3+6*n_freq_dir");
    NWB_REQUIRE(dir_stride >= (dir_is_encoded ? fs.in_dir : 3), "nwb_ray_consts_fwd: dir_stride too small");
    const bool use_t = (mode & NWB_MLP_TRANSIENT) != 0;
    NWB_REQUIRE(!use_t || fs.in_t > 0, "nwb_ray_consts_fwd: transient requested but in_t == 0");
    NWB_REQUIRE(!use_t || (p->tenc_w[0] && p->tenc_b[0]), "nwb_ray_consts_fwd: transient parameters are null");
    NWB_REQUIRE(fs.in_a == 0 || a_rows || (emb_a && ts), "nwb_ray_consts_fwd: appearance needs a_rows or (emb_a, ts)");
    NWB_REQUIRE(!use_t || t_rows || (emb_t && ts), "nwb_ray_consts_fwd: transient needs t_rows or (emb_t, ts)");
    NWB_REQUIRE(fs.in_dir + fs.in_a + fs.in_t <= kMaxFeat, "nwb_ray_consts_fwd: feature row too wide");
    const int blocks = std::min((n_rays + kRcRays - 1) / kRcRays, kNumSMs * 4);
    const size_t smem = static_cast<size_t>(kRcRays) * (((fs.in_dir + fs.in_a + 3) & ~3) + ((fs.in_t + 3) & ~3)) * sizeof(float);
    ray_consts_fwd_kernel<<<blocks, kRcThreads, smem, static_cast<cudaStream_t>(stream)>>>(
        fs, *p, dirs, dir_stride, dir_is_encoded, n_freq_dir, ts, emb_a, emb_t, n_vocab, a_rows, t_rows, n_rays,
        mode, ray_feat, ray_bias, err_flag);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

extern "C" int nwb_ray_consts_bwd(const nwb_field_desc* d, const nwb_field_params* p, const float* ray_feat,
                                  const float* d_ray_bias, const int64_t* ts, int n_vocab, int n_rays,
                                  uint32_t mode, const nwb_field_grads* grads, float* d_emb_a, float* d_emb_t,
                                  float* d_a_rows, float* d_t_rows, void* stream) {
    FieldShape fs;
    int rc = make_shape(d, &fs);
    if (rc != NWB_OK) return rc;
    if (n_rays == 0) return NWB_OK;
    NWB_REQUIRE(p && ray_feat && d_ray_bias && grads, "nwb_ray_consts_bwd: null pointer");
    NWB_REQUIRE(grads->dir_w && grads->dir_b, "nwb_ray_consts_bwd: dir_encoding gradient sinks are null");
    const bool use_t = (mode & NWB_MLP_TRANSIENT) != 0;
    NWB_REQUIRE(!use_t || (grads->tenc_w[0] && grads->tenc_b[0]), "nwb_ray_consts_bwd: transient gradient sinks are null");
    NWB_REQUIRE(fs.in_a == 0 || d_a_rows || !d_emb_a || ts, "nwb_ray_consts_bwd: ts needed for the embedding scatter");
    const int feat_ld = fs.in_dir + fs.in_a + fs.in_t;
    const int ne = fs.in_a + (use_t ? fs.in_t : 0);
    const size_t smem = (static_cast<size_t>(kRcbRays) * (256 + feat_ld) + static_cast<size_t>(kHalf) * ne) * sizeof(float);
    NWB_REQUIRE(smem <= 200 * 1024, "nwb_ray_consts_bwd: feature rows too wide for the staging buffer");
    static size_t configured[kMaxDevices] = {};
    {
        int dev = 0;
        rc = current_device_checked(&dev);
        if (rc != NWB_OK) return rc;
        if (smem > configured[dev]) {
            NWB_CUDA(cudaFuncSetAttribute(ray_consts_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
            configured[dev] = smem;
        }
    }
    const int blocks = (n_rays + kRcbRays - 1) / kRcbRays;
    ray_consts_bwd_kernel<<<blocks, kRcbThreads, smem, static_cast<cudaStream_t>(stream)>>>(
        fs, *p, *grads, ray_feat, d_ray_bias, ts, n_vocab, n_rays, mode, d_emb_a, d_emb_t, d_a_rows, d_t_rows);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
