// (a1) coarse depth sampling, (a2) standalone PosEmbedding, (a6) sample_pdf + merge.
// HBM-bound integer/float work: coalesced loads, one warp per ray, everything else in smem/regs.
#include "common.cuh"
#include <math.h>
#include <algorithm>

namespace nwb {

// ------------------------------------------------------------------------------------------------
// (a1) models/rendering.py:244-259.  One thread per (ray, sample); z written coalesced.
// Explicit round-to-nearest intrinsics keep the op order of the reference's separate ATen ops
// (no FMA contraction), so z is bit-identical to the oracle's.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float coarse_z(float near, float far, int i, int n, bool disp) {
    const float s = linspace01(i, n);
    const float oms = __fsub_rn(1.0f, s);
    if (!disp) return __fadd_rn(__fmul_rn(near, oms), __fmul_rn(far, s));
    const float a = __fmul_rn(__fdiv_rn(1.0f, near), oms);
    const float b = __fmul_rn(__fdiv_rn(1.0f, far), s);
    return __fdiv_rn(1.0f, __fadd_rn(a, b));
}

__global__ void __launch_bounds__(256)
sample_coarse_kernel(const float* __restrict__ rays, int ray_stride, int n_rays, int n_samples,
                     int use_disp, float perturb, const float* __restrict__ perturb_rand,
                     uint64_t seed, float* __restrict__ z_out) {
    const int64_t total = static_cast<int64_t>(n_rays) * n_samples;
    for (int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; idx < total;
         idx += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int r = static_cast<int>(idx / n_samples);
        const int i = static_cast<int>(idx - static_cast<int64_t>(r) * n_samples);
        const float near = __ldg(rays + static_cast<int64_t>(r) * ray_stride + 6);
        const float far = __ldg(rays + static_cast<int64_t>(r) * ray_stride + 7);
        float z = coarse_z(near, far, i, n_samples, use_disp != 0);
        if (perturb > 0.0f) {
            // lower/upper interval bounds around sample i (rendering.py:253-256)
            const float zp = (i > 0) ? coarse_z(near, far, i - 1, n_samples, use_disp != 0) : z;
            const float zn = (i + 1 < n_samples) ? coarse_z(near, far, i + 1, n_samples, use_disp != 0) : z;
            const float lower = (i > 0) ? __fmul_rn(0.5f, __fadd_rn(zp, z)) : z;
            const float upper = (i + 1 < n_samples) ? __fmul_rn(0.5f, __fadd_rn(z, zn)) : z;
            float u;
            if (perturb_rand != nullptr) {
                u = __ldg(perturb_rand + idx);
            } else {
                const uint4 rnd = philox4x32(seed, static_cast<uint64_t>(idx) >> 2, 0u);
                const uint32_t w = (idx & 3) == 0 ? rnd.x : (idx & 3) == 1 ? rnd.y : (idx & 3) == 2 ? rnd.z : rnd.w;
                u = u01(w);
            }
            z = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), __fmul_rn(perturb, u)));
        }
        z_out[idx] = z;
    }
}

// Four consecutive samples of a ray per thread (n_samples % 4 == 0): one Philox block serves its four words (the scalar
// kernel computes the block four times), the six z values the four intervals need are evaluated once, one 16-byte store.
// Same counters, same words, same op order per value: bit-identical to the scalar kernel.
__global__ void __launch_bounds__(256)
sample_coarse4_kernel(const float* __restrict__ rays, int ray_stride, int n_rays, int n_samples,
                      int use_disp, float perturb, const float* __restrict__ perturb_rand,
                      uint64_t seed, float* __restrict__ z_out) {
    const int q_per_ray = n_samples >> 2;
    const int64_t total = static_cast<int64_t>(n_rays) * q_per_ray;
    const bool disp = use_disp != 0;
    for (int64_t qd = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; qd < total;
         qd += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int r = static_cast<int>(qd / q_per_ray);
        const int i0 = 4 * static_cast<int>(qd - static_cast<int64_t>(r) * q_per_ray);
        const float2 nf = __ldg(reinterpret_cast<const float2*>(rays + static_cast<int64_t>(r) * ray_stride + 6));
        const float near = nf.x, far = nf.y;
        float zz[6];                                     // z of samples i0-1 .. i0+4 (ends clamped, unused there)
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const int i = min(max(i0 - 1 + k, 0), n_samples - 1);
            zz[k] = coarse_z(near, far, i, n_samples, disp);
        }
        float out[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) out[k] = zz[k + 1];
        if (perturb > 0.0f) {
            float u[4];
            if (perturb_rand != nullptr) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(perturb_rand) + qd);
                u[0] = t.x; u[1] = t.y; u[2] = t.z; u[3] = t.w;
            } else {
                const uint4 rnd = philox4x32(seed, static_cast<uint64_t>(qd), 0u);    // idx >> 2 == qd
                u[0] = u01(rnd.x); u[1] = u01(rnd.y); u[2] = u01(rnd.z); u[3] = u01(rnd.w);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int i = i0 + k;
                const float z = zz[k + 1];
                const float lower = (i > 0) ? __fmul_rn(0.5f, __fadd_rn(zz[k], z)) : z;
                const float upper = (i + 1 < n_samples) ? __fmul_rn(0.5f, __fadd_rn(z, zz[k + 2])) : z;
                out[k] = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), __fmul_rn(perturb, u[k])));
            }
        }
        reinterpret_cast<float4*>(z_out)[qd] = make_float4(out[0], out[1], out[2], out[3]);
    }
}

// ------------------------------------------------------------------------------------------------
// (a2) models/nerf.py:19-32.  One thread per output element -> fully coalesced stores; sinf/cosf
// are the accurate libdevice versions (arguments reach 2^14*|x| rad; no fast-math).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
posenc_kernel(const float* __restrict__ x, int64_t n, int n_freqs, float* __restrict__ out) {
    const int width = 3 + 6 * n_freqs;
    const int64_t total = n * width;
    for (int64_t idx = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; idx < total;
         idx += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int64_t b = idx / width;
        const int c = static_cast<int>(idx - b * width);
        float v;
        if (c < 3) {
            v = __ldg(x + b * 3 + c);
        } else {
            const int k = (c - 3) / 6, r = (c - 3) % 6;
            const float arg = __fmul_rn(__ldg(x + b * 3 + (r % 3)), exp2f(static_cast<float>(k)));
            v = (r < 3) ? sinf(arg) : cosf(arg);
        }
        out[idx] = v;
    }
}

// Ascending bitonic sort of 32*R keys held in shared memory `a` (padded with +inf), done in registers: element
// e = lane + 32*r lives in register r of `lane`, so partners at distance j < 32 are another lane's register r
// (__shfl_xor) and partners at distance j >= 32 are this lane's register r ^ (j/32).
template <int R>
__device__ __forceinline__ void warp_bitonic_regs(float* a, int lane) {
    float v[R];
#pragma unroll
    for (int r = 0; r < R; ++r) v[r] = a[lane + 32 * r];
#pragma unroll
    for (int k = 2; k <= 32 * R; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j >= 32) {
                const int jr = j >> 5;
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if ((r & jr) == 0) {
                        const int e = lane + 32 * r;
                        const bool up = (e & k) == 0;
                        const float lo = fminf(v[r], v[r | jr]), hi = fmaxf(v[r], v[r | jr]);
                        v[r] = up ? lo : hi;
                        v[r | jr] = up ? hi : lo;
                    }
                }
            } else {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const int e = lane + 32 * r;
                    const float o = __shfl_xor_sync(0xffffffffu, v[r], j);
                    const bool up = (e & k) == 0, lower = (lane & j) == 0;
                    v[r] = (up == lower) ? fminf(v[r], o) : fmaxf(v[r], o);
                }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) a[lane + 32 * r] = v[r];
    __syncwarp();
}

// ------------------------------------------------------------------------------------------------
// (a6) models/rendering.py:7-46 + :267-272.  One warp per ray:
//   stage z / w rows in smem (coalesced), warp-scan the CDF, per-lane binary search
//   (searchsorted right=True), inverse-CDF lerp.  sort(cat[z, z_new]) is a MERGE: z_coarse is sorted, z_new is
//   sorted whenever u is (det, or a sorted u_in) -- checked with a warp vote, otherwise z_new alone goes through a
//   bitonic network -- and every element then finds its output slot by one binary search in the other list
//   (rank of a coarse z = i + #{z_new < z}, rank of a new z = j + #{z_coarse <= z}: a stable merge, no collisions).
// smem per warp: cdf[Ns-1], bins[Ns-1], zc[Ns], zn[P] with P = max(32, next_pow2(Ni)), out[Ns+Ni].
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
sample_pdf_merge_kernel(const float* __restrict__ z_coarse, const float* __restrict__ w_coarse,
                        int n_rays, int Ns, int Ni, int det, const float* __restrict__ u_in,
                        uint64_t seed, const float* __restrict__ cdf_in, float* __restrict__ z_out,
                        float* __restrict__ z_new_out, int64_t* __restrict__ inds_out, int P) {
    extern __shared__ float smem[];
    const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int M = Ns - 2;        // number of pdf bins (weights[:,1:-1])
    const int nb = Ns - 1;       // number of bin edges == cdf entries
    const int per_warp = 2 * nb + Ns + P + Ns + Ni;
    float* cdf = smem + warp * per_warp;
    float* bins = cdf + nb;
    float* keys = bins + nb;          // zc[Ns] followed by zn[P]
    float* zn = keys + Ns;
    float* outb = zn + P;
    const float eps = 1e-5f;

    for (int ray = blockIdx.x * warps + warp; ray < n_rays; ray += gridDim.x * warps) {
        const float* zr = z_coarse + static_cast<int64_t>(ray) * Ns;
        const float* wr = w_coarse + static_cast<int64_t>(ray) * Ns;
        // z row -> keys[0..Ns), midpoints -> bins
        for (int i = lane; i < Ns; i += 32) keys[i] = __ldg(zr + i);
        __syncwarp();
        for (int i = lane; i < nb; i += 32) bins[i] = __fmul_rn(0.5f, __fadd_rn(keys[i], keys[i + 1]));
        // pdf normaliser
        float part = 0.0f;
        for (int i = lane; i < M; i += 32) part += __ldg(wr + 1 + i) + eps;
        const float total = warp_sum(part);
        // cdf = [0, cumsum(pdf)]
        if (cdf_in != nullptr) {
            for (int i = lane; i < nb; i += 32) cdf[i] = __ldg(cdf_in + static_cast<int64_t>(ray) * nb + i);
        } else {
            float carry = 0.0f;
            if (lane == 0) cdf[0] = 0.0f;
            for (int base = 0; base < M; base += 32) {
                const int i = base + lane;
                float v = (i < M) ? __fdiv_rn(__ldg(wr + 1 + i) + eps, total) : 0.0f;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const float t = __shfl_up_sync(0xffffffffu, v, o);
                    if (lane >= o) v += t;
                }
                v += carry;
                if (i < M) cdf[i + 1] = v;
                carry = __shfl_sync(0xffffffffu, v, 31);
            }
        }
        __syncwarp();
        // inverse-CDF samples
        for (int j = lane; j < Ni; j += 32) {
            float u;
            if (det) {
                u = linspace01(j, Ni);
            } else if (u_in != nullptr) {
                u = __ldg(u_in + static_cast<int64_t>(ray) * Ni + j);
            } else {
                const uint64_t ctr = static_cast<uint64_t>(ray) * Ni + j;
                const uint4 rnd = philox4x32(seed, ctr >> 2, 1u);
                const uint32_t w = (ctr & 3) == 0 ? rnd.x : (ctr & 3) == 1 ? rnd.y : (ctr & 3) == 2 ? rnd.z : rnd.w;
                u = u01(w);
            }
            // first index with cdf[idx] > u  (searchsorted right=True), idx in [0, nb]
            int lo = 0, hi = nb;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (cdf[mid] > u) hi = mid; else lo = mid + 1;
            }
            const int below = max(lo - 1, 0);
            const int above = min(lo, M);
            const float c0 = cdf[below], c1 = cdf[above];
            const float b0 = bins[below], b1 = bins[above];
            float denom = __fsub_rn(c1, c0);
            if (denom < eps) denom = 1.0f;
            const float t = __fdiv_rn(__fsub_rn(u, c0), denom);
            const float s = __fadd_rn(b0, __fmul_rn(t, __fsub_rn(b1, b0)));
            keys[Ns + j] = s;
            if (z_new_out != nullptr) z_new_out[static_cast<int64_t>(ray) * Ni + j] = s;
            if (inds_out != nullptr) inds_out[static_cast<int64_t>(ray) * Ni + j] = lo;
        }
        for (int i = Ni + lane; i < P; i += 32) zn[i] = INFINITY;
        __syncwarp();
        // is z_new already ascending?
        bool ok = true;
        for (int j = lane; j + 1 < Ni; j += 32) ok = ok && (zn[j] <= zn[j + 1]);
        const bool unsorted = !__all_sync(0xffffffffu, ok);
        if (unsorted && P <= 256) {
            // stochastic u: bitonic sort in REGISTERS (P/32 keys per lane, exchanges across lanes by shuffle) -- the
            // shared-memory network below costs a load/compare/store/__syncwarp round trip per stage (28 stages at P = 128)
            if (P <= 32) warp_bitonic_regs<1>(zn, lane);
            else if (P == 64) warp_bitonic_regs<2>(zn, lane);
            else if (P == 128) warp_bitonic_regs<4>(zn, lane);
            else warp_bitonic_regs<8>(zn, lane);
        } else if (unsorted) {
            // warp bitonic sort of the P new keys in smem
            for (int k = 2; k <= P; k <<= 1) {
                for (int j = k >> 1; j > 0; j >>= 1) {
                    for (int t = lane; t < (P >> 1); t += 32) {
                        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));   // index with bit j clear
                        const int l = i | j;
                        const bool up = ((i & k) == 0);
                        const float a = zn[i], b = zn[l];
                        if ((a > b) == up) { zn[i] = b; zn[l] = a; }
                    }
                    __syncwarp();
                }
            }
        }
        // merge by rank
        for (int i = lane; i < Ns; i += 32) {
            const float a = keys[i];
            int lo = 0, hi = Ni;                       // #{z_new < a}
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (zn[mid] < a) lo = mid + 1; else hi = mid;
            }
            outb[i + lo] = a;
        }
        for (int j = lane; j < Ni; j += 32) {
            const float b = zn[j];
            int lo = 0, hi = Ns;                       // #{z_coarse <= b}
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (keys[mid] <= b) lo = mid + 1; else hi = mid;
            }
            outb[j + lo] = b;
        }
        __syncwarp();
        float* zo = z_out + static_cast<int64_t>(ray) * (Ns + Ni);
        for (int i = lane; i < Ns + Ni; i += 32) zo[i] = outb[i];
        __syncwarp();
    }
}


// Ascending bitonic sort of 32*R keys held BLOCKED in registers: element e = lane*R + r is register r of `lane`.  Partners
// at distance j < R are this lane's register r ^ j (two FMNMX per pair), partners at distance j >= R are register r of lane
// ^ (j/R) (one shuffle and one FMNMX per key).  For R = 4 and 128 keys that is 13 register steps and 15 shuffle steps
// (the strided layout above: 3 and 25).
template <int R>
__device__ __forceinline__ void warp_bitonic_blocked(float (&v)[R], int lane) {
#pragma unroll
    for (int k = 2; k <= 32 * R; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j < R) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if ((r & j) == 0) {
                        const bool up = ((lane * R + r) & k) == 0;
                        const float a = v[r], b = v[r | j];
                        v[r] = up ? fminf(a, b) : fmaxf(a, b);
                        v[r | j] = up ? fmaxf(a, b) : fminf(a, b);
                    }
                }
            } else {
                const int jl = j / R;
                const bool keep_min = (((lane * R) & k) == 0) == ((lane & jl) == 0);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const float o = __shfl_xor_sync(0xffffffffu, v[r], jl);
                    v[r] = keep_min ? fminf(v[r], o) : fmaxf(v[r], o);
                }
            }
        }
    }
}

// #{k < n : !(a[k] > x)} for ascending a[0..n), n a power of two, without a data-dependent branch: log2(n) probes of the
// first n-1 entries and one of the entry the walk stops at.  kStrict counts a[k] < x instead.  The same result as the
// bisection loops of the general kernel on ascending input.
template <int n, bool kStrict>
__device__ __forceinline__ int count_below(const float* a, float x) {
    int pos = 0;
#pragma unroll
    for (int s = n >> 1; s > 0; s >>= 1) {
        const float c = a[pos + s - 1];
        if (kStrict ? (c < x) : !(c > x)) pos += s;
    }
    const float c = a[pos];
    return pos + ((kStrict ? (c < x) : !(c > x)) ? 1 : 0);
}


// (a6) for Ns = 32*RS, Ni = 32*RQ (RS, RQ in {2, 4}): the same arithmetic per value as sample_pdf_merge_kernel -- bit-
// identical z_out / z_new / inds -- with every loop unrolled, the coarse row and the pdf kept in registers, the bisections
// replaced by fixed-trip-count walks (RQ independent ones in flight per lane), the new samples BLOCKED over the lanes
// (lane owns RQ consecutive j: one Philox block or one vector load of u per lane, the ascending check and the sort on
// registers) and the row written back with 8-byte stores.
// [B200, profiles/r02z9_*] the general kernel executes ~2000 warp instructions per ray at 128+128 and is issue-bound
// (70 % of the issue slots, 0.10 of the copy bandwidth).
template <int RS, int RQ>
__global__ void __launch_bounds__(128)
sample_pdf_merge_fast_kernel(const float* __restrict__ z_coarse, const float* __restrict__ w_coarse, int n_rays, int det,
                             const float* __restrict__ u_in, uint64_t seed, const float* __restrict__ cdf_in,
                             float* __restrict__ z_out, float* __restrict__ z_new_out, int64_t* __restrict__ inds_out) {
    constexpr int Ns = 32 * RS, Ni = 32 * RQ;
    constexpr int M = Ns - 2, nb = Ns - 1;
    constexpr int per_warp = 3 * Ns + Ni + (Ns + Ni);        // cdf[Ns] (Ns-1 used), bins[Ns], zc[Ns], zn[Ni], out[Ns+Ni]
    extern __shared__ __align__(16) float smem[];
    const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* cdf = smem + warp * per_warp;
    float* bins = cdf + Ns;
    float* zc = bins + Ns;
    float* zn = zc + Ns;
    float* outb = zn + Ni;
    const float eps = 1e-5f;

    for (int ray = blockIdx.x * warps + warp; ray < n_rays; ray += gridDim.x * warps) {
        const float* zr = z_coarse + static_cast<int64_t>(ray) * Ns;
        const float* wr = w_coarse + static_cast<int64_t>(ray) * Ns;
        float zv[RS], pw[RS];                                  // element i = lane + 32 r
#pragma unroll
        for (int r = 0; r < RS; ++r) {
            const int i = lane + 32 * r;
            zv[r] = __ldg(zr + i);
            pw[r] = (i < M) ? __ldg(wr + 1 + i) + eps : 0.0f;
            zc[i] = zv[r];
        }
        __syncwarp();
#pragma unroll
        for (int r = 0; r < RS; ++r) {
            const int i = lane + 32 * r;
            if (i < nb) bins[i] = __fmul_rn(0.5f, __fadd_rn(zv[r], zc[i + 1]));
        }
        // pdf normaliser: the general kernel's order (lane-strided partial sums, then the warp tree)
        float part = 0.0f;
#pragma unroll
        for (int r = 0; r < RS; ++r)
            if (lane + 32 * r < M) part += pw[r];
        const float total = warp_sum(part);
        if (cdf_in != nullptr) {
#pragma unroll
            for (int r = 0; r < RS; ++r) {
                const int i = lane + 32 * r;
                if (i < nb) cdf[i] = __ldg(cdf_in + static_cast<int64_t>(ray) * nb + i);
            }
        } else {
            float carry = 0.0f;
            if (lane == 0) cdf[0] = 0.0f;
#pragma unroll
            for (int r = 0; r < RS; ++r) {
                const int i = lane + 32 * r;
                float v = (i < M) ? __fdiv_rn(pw[r], total) : 0.0f;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const float t = __shfl_up_sync(0xffffffffu, v, o);
                    if (lane >= o) v += t;
                }
                v += carry;
                if (i < M) cdf[i + 1] = v;
                carry = __shfl_sync(0xffffffffu, v, 31);
            }
        }
        __syncwarp();
        // inverse-CDF samples j = RQ*lane + q
        float u[RQ], sv[RQ];
        const int j0 = RQ * lane;
        if (det) {
#pragma unroll
            for (int q = 0; q < RQ; ++q) u[q] = linspace01(j0 + q, Ni);
        } else if (u_in != nullptr) {
            const float* up = u_in + static_cast<int64_t>(ray) * Ni + j0;
            if constexpr (RQ == 4) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(up));
                u[0] = t.x; u[1] = t.y; u[2] = t.z; u[3] = t.w;
            } else {
                const float2 t = __ldg(reinterpret_cast<const float2*>(up));
                u[0] = t.x; u[1] = t.y;
            }
        } else {
            // counters ray*Ni + j: the RQ consecutive j of a lane sit in one Philox block (Ni and j0 are multiples of RQ)
            const uint64_t ctr = static_cast<uint64_t>(ray) * Ni + j0;
            const uint4 rnd = philox4x32(seed, ctr >> 2, 1u);
            const uint32_t w4[4] = {rnd.x, rnd.y, rnd.z, rnd.w};
#pragma unroll
            for (int q = 0; q < RQ; ++q) u[q] = u01(w4[(j0 + q) & 3]);
        }
        int lo[RQ];                                            // first index with cdf[idx] > u, idx in [0, nb]
#pragma unroll
        for (int q = 0; q < RQ; ++q) lo[q] = 0;
#pragma unroll
        for (int s2 = Ns >> 1; s2 > 0; s2 >>= 1) {             // the walk covers exactly the nb = Ns-1 entries
#pragma unroll
            for (int q = 0; q < RQ; ++q)
                if (!(cdf[lo[q] + s2 - 1] > u[q])) lo[q] += s2;
        }
#pragma unroll
        for (int q = 0; q < RQ; ++q) {
            const int below = max(lo[q] - 1, 0);
            const int above = min(lo[q], M);
            const float c0 = cdf[below], c1 = cdf[above];
            const float b0 = bins[below], b1 = bins[above];
            float denom = __fsub_rn(c1, c0);
            if (denom < eps) denom = 1.0f;
            const float t = __fdiv_rn(__fsub_rn(u[q], c0), denom);
            sv[q] = __fadd_rn(b0, __fmul_rn(t, __fsub_rn(b1, b0)));
            if (z_new_out != nullptr) z_new_out[static_cast<int64_t>(ray) * Ni + j0 + q] = sv[q];
            if (inds_out != nullptr) inds_out[static_cast<int64_t>(ray) * Ni + j0 + q] = lo[q];
        }
        // ascending already (det, or a sorted u)?  otherwise the register sort
        bool ok = true;
#pragma unroll
        for (int q = 0; q + 1 < RQ; ++q) ok = ok && (sv[q] <= sv[q + 1]);
        const float nxt = __shfl_down_sync(0xffffffffu, sv[0], 1);
        ok = ok && (lane == 31 || sv[RQ - 1] <= nxt);
        if (!__all_sync(0xffffffffu, ok)) warp_bitonic_blocked<RQ>(sv, lane);
#pragma unroll
        for (int q = 0; q < RQ; ++q) zn[j0 + q] = sv[q];
        __syncwarp();
        // merge by rank (stable: a coarse z goes before an equal new z)
#pragma unroll
        for (int r = 0; r < RS; ++r) outb[lane + 32 * r + count_below<Ni, true>(zn, zv[r])] = zv[r];
#pragma unroll
        for (int q = 0; q < RQ; ++q) outb[j0 + q + count_below<Ns, false>(zc, sv[q])] = sv[q];
        __syncwarp();
        float2* zo = reinterpret_cast<float2*>(z_out + static_cast<int64_t>(ray) * (Ns + Ni));
#pragma unroll
        for (int i = 0; i < (RS + RQ) / 2; ++i) zo[lane + 32 * i] = reinterpret_cast<const float2*>(outb)[lane + 32 * i];
        __syncwarp();
    }
}

}  // namespace nwb

using namespace nwb;

extern "C" int nwb_sample_coarse(const float* rays, int ray_stride, int n_rays, int n_samples,
                                 int use_disp, float perturb, const float* perturb_rand,
                                 uint64_t seed, float* z_out, void* stream) {
    NWB_REQUIRE(n_rays == 0 || (rays && z_out), "nwb_sample_coarse: null pointer");
    NWB_REQUIRE(ray_stride >= 8 && n_rays >= 0 && n_samples >= 1, "nwb_sample_coarse: bad shape");
    if (n_rays == 0) return NWB_OK;
    const int64_t total = static_cast<int64_t>(n_rays) * n_samples;
    const bool vec4 = (n_samples & 3) == 0 && (ray_stride & 1) == 0 && (reinterpret_cast<uintptr_t>(rays) & 7u) == 0 &&
                      (reinterpret_cast<uintptr_t>(z_out) & 15u) == 0 && (perturb_rand == nullptr || (reinterpret_cast<uintptr_t>(perturb_rand) & 15u) == 0);
    if (vec4) {
        const int blocks = static_cast<int>(std::min<int64_t>((total / 4 + 255) / 256, kNumSMs * 8));
        sample_coarse4_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
            rays, ray_stride, n_rays, n_samples, use_disp, perturb, perturb_rand, seed, z_out);
        NWB_LAUNCH_CHECK();
        return NWB_OK;
    }
    const int blocks = static_cast<int>(std::min<int64_t>((total + 255) / 256, kNumSMs * 8));
    sample_coarse_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        rays, ray_stride, n_rays, n_samples, use_disp, perturb, perturb_rand, seed, z_out);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

extern "C" int nwb_posenc(const float* x, int64_t n, int n_freqs, float* out, void* stream) {
    NWB_REQUIRE(n == 0 || (x && out), "nwb_posenc: null pointer");
    NWB_REQUIRE(n >= 0 && n_freqs >= 0 && n_freqs <= 24, "nwb_posenc: bad shape");
    if (n == 0) return NWB_OK;
    const int64_t total = n * (3 + 6 * n_freqs);
    const int blocks = static_cast<int>(std::min<int64_t>((total + 255) / 256, kNumSMs * 8));
    posenc_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, n, n_freqs, out);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}

extern "C" int nwb_sample_pdf_merge(const float* z_coarse, const float* w_coarse, int n_rays,
                                    int n_samples, int n_importance, int det, const float* u,
                                    uint64_t seed, const float* cdf_in, float* z_out, float* z_new,
                                    int64_t* inds, void* stream) {
    NWB_REQUIRE(n_rays == 0 || (z_coarse && w_coarse && z_out), "nwb_sample_pdf_merge: null pointer");
    NWB_REQUIRE(n_samples >= 3 && n_importance >= 1 && n_rays >= 0,
                "nwb_sample_pdf_merge: need n_samples>=3, n_importance>=1");
    NWB_REQUIRE(n_samples + n_importance <= 4096, "nwb_sample_pdf_merge: too many samples per ray");
    if (n_rays == 0) return NWB_OK;
    const cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool vec_ok = (reinterpret_cast<uintptr_t>(z_out) & 7u) == 0 && (reinterpret_cast<uintptr_t>(u) & 15u) == 0;
    if (vec_ok && (n_samples == 64 || n_samples == 128) && (n_importance == 64 || n_importance == 128)) {
        const int warps = 4;
        const int blocks = std::min((n_rays + warps - 1) / warps, kNumSMs * 8);
        const size_t smem = static_cast<size_t>(warps) * (4 * n_samples + 2 * n_importance) * sizeof(float);
#define NWB_PDF_FAST(RS, RQ) sample_pdf_merge_fast_kernel<RS, RQ><<<blocks, warps * 32, smem, st>>>( \
            z_coarse, w_coarse, n_rays, det, u, seed, cdf_in, z_out, z_new, inds)
        if (n_samples == 64 && n_importance == 64) NWB_PDF_FAST(2, 2);
        else if (n_samples == 64) NWB_PDF_FAST(2, 4);
        else if (n_importance == 64) NWB_PDF_FAST(4, 2);
        else NWB_PDF_FAST(4, 4);
#undef NWB_PDF_FAST
        NWB_LAUNCH_CHECK();
        return NWB_OK;
    }
    int P = 32;                                  // sort width: >= one key per lane (register sort), +inf padded
    while (P < n_importance) P <<= 1;
    const int warps = 4;
    const size_t smem = static_cast<size_t>(warps) * (2 * (n_samples - 1) + P + 2 * n_samples + n_importance) * sizeof(float);
    if (smem > 48 * 1024)        // per device and per size: set on every such call (a host-side attribute write)
        NWB_CUDA(cudaFuncSetAttribute(sample_pdf_merge_kernel,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    const int blocks = std::min((n_rays + warps - 1) / warps, kNumSMs * 8);
    sample_pdf_merge_kernel<<<blocks, warps * 32, smem, static_cast<cudaStream_t>(stream)>>>(
        z_coarse, w_coarse, n_rays, n_samples, n_importance, det, u, seed, cdf_in, z_out, z_new,
        inds, P);
    NWB_LAUNCH_CHECK();
    return NWB_OK;
}
