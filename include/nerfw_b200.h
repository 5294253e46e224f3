/*
 * nerfw_b200.h -- C ABI of the B200-native NeRF-W per-ray hot path.
 *
 * The reference (nmerty/nerf-fl) is pure Python on top of ATen; it has no FFI of its own, so
 * every entry point below cites the reference Python lines whose work it replaces (paths relative
 * to /root/reference).  The binding a maintainer adds is a ctypes stub -- see INTEGRATION.md.
 *
 * Conventions
 *   - plain pointers + sizes, no torch types; all pointers are DEVICE pointers unless noted
 *   - every call is stream-ordered on `stream` (a cudaStream_t passed as void*), never syncs the
 *     device, never allocates: the caller owns every buffer (workspaces included)
 *   - return 0 on success, <0 on error (NWB_E_*); nwb_last_error() gives the thread-local text
 *   - fp32 row-major tensors unless stated; "image" buffers are opaque bf16 tile images whose
 *     size is given by the nwb_*_bytes() queries
 */
#ifndef NERFW_B200_H
#define NERFW_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NWB_OK 0
#define NWB_E_ARG (-1)      /* bad argument / unsupported shape                     */
#define NWB_E_CUDA (-2)     /* a CUDA runtime call or launch failed                 */
#define NWB_E_RANGE (-3)    /* embedding index out of range (README.md:124 hazard)  */

int nwb_abi_version(void);
const char* nwb_last_error(void);
/* number of kernels this library has launched in this process (diagnostics; bench.py gpu_launches) */
long long nwb_kernel_launches(void);

/* ------------------------------------------------------------------------------------------
 * (a1) coarse depth sampling -- models/rendering.py:244-259
 * rays: (n_rays, ray_stride>=8) [o3 d3 near far]; z_out: (n_rays, n_samples).
 * perturb>0: stratified jitter with `perturb_rand` (n_rays,n_samples) U[0,1) if non-NULL, else an
 * in-kernel Philox4x32-10 stream keyed by `seed`.
 * ---------------------------------------------------------------------------------------- */
int nwb_sample_coarse(const float* rays, int ray_stride, int n_rays, int n_samples, int use_disp,
                      float perturb, const float* perturb_rand, uint64_t seed, float* z_out,
                      void* stream);

/* ------------------------------------------------------------------------------------------
 * (a2) PosEmbedding.forward -- models/nerf.py:19-32 (logscale freqs 2^k, nerf.py:15)
 * x: (n, 3) -> out: (n, 3 + 6*n_freqs)
 * ---------------------------------------------------------------------------------------- */
int nwb_posenc(const float* x, int64_t n, int n_freqs, float* out, void* stream);

/* ------------------------------------------------------------------------------------------
 * (a6) sample_pdf + merge -- models/rendering.py:7-46 and :267-272
 * z_coarse, w_coarse: (n_rays, n_samples).  bins = midpoints, weights = w_coarse[:,1:-1].
 * det!=0: u = linspace(0,1,n_importance); else u (n_rays,n_importance) if non-NULL, else Philox.
 * z_out: (n_rays, n_samples + n_importance) sorted ascending.
 * Optional debug outputs (may be NULL): z_new (n_rays,n_importance) unsorted samples,
 * inds (n_rays,n_importance) int64 searchsorted(right) indices.
 * cdf_in (may be NULL): (n_rays, n_samples-1) externally supplied CDF used instead of the one
 * built in-kernel (the "bit-exact given identical CDFs" contract).
 * ---------------------------------------------------------------------------------------- */
int nwb_sample_pdf_merge(const float* z_coarse, const float* w_coarse, int n_rays, int n_samples,
                         int n_importance, int det, const float* u, uint64_t seed,
                         const float* cdf_in, float* z_out, float* z_new, int64_t* inds,
                         void* stream);

/* ------------------------------------------------------------------------------------------
 * (a5) volume rendering -- models/rendering.py:141-226
 * raw: (n_rays, S, C) with C = 1 (sigma only), 4 (rgb,sigma) or 9 (+rgb_t,sigma_t,beta).
 * ---------------------------------------------------------------------------------------- */
#define NWB_COMP_TRANSIENT 1u   /* C==9 and output_transient                     */
#define NWB_COMP_WHITE_BACK 2u
#define NWB_COMP_TEST_TIME 4u
#define NWB_COMP_COARSE 8u      /* typ == 'coarse' (test_time: weights/opacity only) */
#define NWB_COMP_AMAX_CLEARED 16u /* nwb_composite_bwd_seeded: *amax_bits is already 0 (nwb_composite_out.clear_word) */

typedef struct nwb_composite_out {
    float* weights;          /* (n,S)  weights_{typ}                    rendering.py:162,165 */
    float* opacity;          /* (n)    opacity_{typ}                    :163,166             */
    float* rgb;              /* (n,3)  rgb_{typ}; transient: rgb_fine   :190,219-223         */
    float* depth;            /* (n)    depth_{typ}                      :225                 */
    float* transient_sigmas; /* (n,S)                                   :168                 */
    float* beta;             /* (n)    incl. +beta_min                  :182-185             */
    float* rgb_static;       /* (n,3)  _rgb_fine_static                 :174-177,188         */
    float* rgb_transient;    /* (n,3)  _rgb_fine_transient              :179-181,189         */
    float* tt_rgb_static;    /* (n,3)  rgb_fine_static    (test_time)   :196-205             */
    float* tt_depth_static;  /* (n)    depth_fine_static                :206-207             */
    float* tt_rgb_transient; /* (n,3)  rgb_fine_transient               :209-215             */
    float* tt_depth_transient; /* (n)  depth_fine_transient             :216-217             */
    unsigned int* clear_word;  /* may be NULL: the launch stores 0 there -- the amax word of the backward's workspace,
                                * cleared a pass ahead so that nwb_composite_bwd_seeded needs no memset node
                                * (NWB_COMP_AMAX_CLEARED)                                                             */
} nwb_composite_out;

int nwb_composite_fwd(const float* raw, int C, const float* z, const float* noise,
                      float noise_std, int n_rays, int S, uint32_t flags, float beta_min,
                      const nwb_composite_out* out, void* stream);

/* cotangents of the training-mode outputs (NULL = zero).  d_raw: (n_rays,S,C) is overwritten. */
typedef struct nwb_composite_grads {
    const float* d_weights;          /* (n,S) */
    const float* d_opacity;          /* (n)   */
    const float* d_rgb;              /* (n,3) rgb_{typ} / rgb_fine */
    const float* d_depth;            /* (n)   */
    const float* d_transient_sigmas; /* (n,S) */
    const float* d_beta;             /* (n)   */
    const float* d_rgb_static;       /* (n,3) _rgb_fine_static    */
    const float* d_rgb_transient;    /* (n,3) _rgb_fine_transient */
} nwb_composite_grads;

int nwb_composite_bwd(const float* raw, int C, const float* z, const float* noise,
                      float noise_std, int n_rays, int S, uint32_t flags,
                      const nwb_composite_grads* g, float* d_raw, void* stream);
/* which kernel the compositing backward launches: 0 (default) = the one that keeps the whole ray in shared memory for S <= 128,
 * 1 = always the chunk-ring kernel, 2 = the resident one wherever it fits (S <= 256).  Process-wide; same results (tests walk all). */
int nwb_composite_bwd_select(int gen);

/* The same backward seeding its own cotangents under NerfWLoss (losses.py:35-50; SURVEY.md 8f-1 "fold the loss into the
 * compositing epilogue"): with `seed` the cotangents of this pass's rgb (and, fine pass with a transient head, of beta and
 * transient_sigmas) are formed in the kernel from the maps and the targets -- the arithmetic of nwb_nerfw_loss_bwd, bit for
 * bit -- and ADDED to whatever `g` carries (g may be NULL = all zero); no loss-backward launch, no (n,3)/(n)/(n,S)
 * cotangent tensors.  g_terms: DEVICE scalars, cotangents of (c_l, f_l, b_l, s_l, total), NULL = 0.
 * amax_bits (may be NULL): receives max |d_raw| as fp32 bits (cleared by a 4-byte memset node first) -- point it at the
 * first word of the nwb_mlp_bwd workspace and pass NWB_MLP_DZ_AMAX_GIVEN there to skip that call's own reduction launch. */
typedef struct nwb_loss_seed {
    const float* rgb;          /* (n,3) rgb_coarse (fine == 0) or rgb_fine (fine != 0), as the forward produced it */
    const float* beta;         /* (n)   fine pass with a transient head, else NULL                  losses.py:42-44 */
    const float* targets;      /* (n,3)                                                                            */
    const float* g_terms[5];
    float coef, lambda_u;      /* NerfWLoss(coef, lambda_u)                                         losses.py:27-33 */
    int n_rays;                /* batch size the means run over                                                     */
    int64_t n_ts;              /* numel(transient_sigmas); 0: no s_l term                           losses.py:45   */
    int fine;
} nwb_loss_seed;

int nwb_composite_bwd_seeded(const float* raw, int C, const float* z, const float* noise,
                             float noise_std, int n_rays, int S, uint32_t flags,
                             const nwb_composite_grads* g, const nwb_loss_seed* seed,
                             unsigned int* amax_bits, float* d_raw, void* stream);

/* ------------------------------------------------------------------------------------------
 * (a3,a4) field network -- models/nerf.py:81-212 + the per-sample input assembly of
 * models/rendering.py:96-139.  W=256, D=8, skips=[4] (the only shape the reference ever builds).
 * ---------------------------------------------------------------------------------------- */
typedef struct nwb_field_desc {
    int in_xyz;        /* 3+6*N_emb_xyz, <= 128 (> 64: one-tile kernel) nerf.py:83  */
    int in_dir;        /* 3+6*N_emb_dir                              nerf.py:83  */
    int in_a;          /* appearance width, 0 = no encode_appearance nerf.py:84  */
    int in_t;          /* transient width,  0 = no encode_transient  nerf.py:85  */
} nwb_field_desc;

/* fp32 parameter pointers with the reference's state_dict names */
typedef struct nwb_field_params {
    const float* trunk_w[8];   /* xyz_encoding_{1..8}.0.weight  (256, in) */
    const float* trunk_b[8];   /* xyz_encoding_{1..8}.0.bias              */
    const float* final_w;      /* xyz_encoding_final.weight (256,256)     */
    const float* final_b;
    const float* dir_w;        /* dir_encoding.0.weight (128, 256+in_dir+in_a) */
    const float* dir_b;
    const float* sigma_w;      /* static_sigma.0 (1,256) */
    const float* sigma_b;
    const float* rgb_w;        /* static_rgb.0 (3,128) */
    const float* rgb_b;
    const float* tenc_w[4];    /* transient_encoding.{0,2,4,6}.weight */
    const float* tenc_b[4];
    const float* tsigma_w;     /* transient_sigma.0 (1,128) */
    const float* tsigma_b;
    const float* trgb_w;       /* transient_rgb.0 (3,128) */
    const float* trgb_b;
    const float* tbeta_w;      /* transient_beta.0 (1,128) */
    const float* tbeta_b;
} nwb_field_params;

/* same layout, mutable: gradient sinks (accumulated into, caller zeroes) */
typedef struct nwb_field_grads {
    float* trunk_w[8];
    float* trunk_b[8];
    float* final_w;
    float* final_b;
    float* dir_w;
    float* dir_b;
    float* sigma_w;
    float* sigma_b;
    float* rgb_w;
    float* rgb_b;
    float* tenc_w[4];
    float* tenc_b[4];
    float* tsigma_w;
    float* tsigma_b;
    float* trgb_w;
    float* trgb_b;
    float* tbeta_w;
    float* tbeta_b;
} nwb_field_grads;

/* bytes of the packed operand cache (bf16 swizzled weight slabs + fp32 head/bias block) */
size_t nwb_field_pack_bytes(const nwb_field_desc* d);
/* (re)build the packed cache from fp32 parameters; call after every optimizer step */
int nwb_field_pack(const nwb_field_desc* d, const nwb_field_params* p, void* packed, void* stream);

/* modes of one field evaluation */
#define NWB_MLP_SIGMA_ONLY 1u   /* NeRF.forward(sigma_only=True)            nerf.py:191-192 */
#define NWB_MLP_TRANSIENT 2u    /* output_transient=True                    nerf.py:200-212 */
#define NWB_MLP_SAVE 4u         /* keep activations for the backward pass                   */
#define NWB_MLP_DZ_AMAX_GIVEN 8u /* nwb_mlp_bwd / nwb_mlp_dgrad: workspace word 0 already holds max|d_raw| (fp32 bits) */

/* per-ray operands of dir_encoding / transient_encoding.0: gathers embedding rows, encodes the
 * view direction and folds them through the matching weight columns into per-ray bias vectors
 * (rendering.py:113-130 repeat+cat, never materialised per sample).
 *   dirs      (n_rays, dir_stride) raw directions (PosEmbedding applied here, n_freq_dir) or, if
 *             dir_is_encoded, already-encoded (n_rays, in_dir) rows
 *   a_rows/t_rows: (n_rays, in_a)/(n_rays, in_t) explicit rows (kwargs a_embedded/t_embedded,
 *             rendering.py:277-285) or NULL to gather emb_a/emb_t[ts]
 *   ray_feat  (n_rays, in_dir+in_a+in_t) fp32 out: [dir_enc | a | t], kept for the backward
 *   ray_bias  (n_rays, 256) fp32 out: [dir bias (128) | transient bias (128)]
 * returns NWB_E_RANGE if any ts is outside [0, n_vocab) (checked on device, reported via
 * *err_flag, a device int the caller reads when it next syncs).
 */
int nwb_ray_consts_fwd(const nwb_field_desc* d, const nwb_field_params* p, const float* dirs,
                       int dir_stride, int dir_is_encoded, int n_freq_dir, const int64_t* ts,
                       const float* emb_a, const float* emb_t, int n_vocab, const float* a_rows,
                       const float* t_rows, int n_rays, uint32_t mode, float* ray_feat,
                       float* ray_bias, int* err_flag, void* stream);

/* bytes of the saved-activation buffer of one nwb_mlp_fwd(NWB_MLP_SAVE) over n_points */
size_t nwb_mlp_save_bytes(const nwb_field_desc* d, int64_t n_points, uint32_t mode);

/* fused field forward over n_rays*S points.
 *   xyz source: either (rays, z) -- points o + d*z are generated and encoded in-kernel
 *   (rendering.py:261,273 + nerf.py:27-32, n_freq_xyz) -- or x_enc (n_points, in_xyz) explicit
 *   encoded rows (the NeRF.forward surface); exactly one of z / x_enc is non-NULL.
 *   raw_out: (n_points, C) C = 1 | 4 | 9 by mode (column order nerf.py:198,209-212)
 */
int nwb_mlp_fwd(const nwb_field_desc* d, const void* packed, const float* rays, int ray_stride,
                const float* z, const float* x_enc, int n_freq_xyz, const float* ray_bias,
                int n_rays, int S, uint32_t mode, float* raw_out, void* saved, void* stream);

/* which kernel nwb_mlp_fwd launches: 0 (default) = chosen per launch -- whole rounds of 296 tiles on the CTA-pair
 * kernel, a remainder of <= 148 tiles on the one-tile kernel; 1 = always the one-tile kernel; 2 = always the pair
 * kernel.  Process-wide; results are identical, only speed differs (tests walk every branch with it). */
int nwb_mlp_fwd_select(int gen);

/* which dgrad kernel the backward launches: 0 (default) or 2 = CTA pairs with two tiles in flight per CTA, 1 = the
 * first-generation kernel (two CTAs per SM).  Process-wide; only speed differs. */
int nwb_mlp_dgrad_select(int gen);

/* backward of nwb_mlp_fwd: raw (the forward's output) and its cotangent d_raw (n_points, C) ->
 * parameter gradients (ACCUMULATED into the sinks with fp32 atomics; the caller zeroes them) and the
 * per-ray cotangent d_ray_bias (n_rays,256), also accumulated (caller zeroes).  `saved` is the buffer
 * the forward filled under NWB_MLP_SAVE; workspace: nwb_mlp_bwd_workspace_bytes() bytes, 16 B aligned.
 * Two kernels: the dgrad chain (tcgen05, W read MN-major from the packed cache) and the weight-gradient
 * contraction over points (tcgen05, both operands MN-major).  grads->dir_b / tenc_b[0] and the
 * [256:] columns of dir_w / tenc_w[0] are produced by nwb_ray_consts_bwd from d_ray_bias. */
size_t nwb_mlp_bwd_workspace_bytes(const nwb_field_desc* d, int64_t n_points, uint32_t mode);
int nwb_mlp_bwd(const nwb_field_desc* d, const void* packed, const float* raw, const float* d_raw,
                const void* saved, int n_rays, int S, uint32_t mode, const nwb_field_grads* grads,
                float* d_ray_bias, void* workspace, void* stream);
/* the two kernels of nwb_mlp_bwd as separate calls (same arguments; dgrad first, then wgrad on the
 * workspace it filled) -- lets a caller time them apart or overlap wgrad with other work */
int nwb_mlp_dgrad(const nwb_field_desc* d, const void* packed, const float* raw, const float* d_raw,
                  const void* saved, int n_rays, int S, uint32_t mode, const nwb_field_grads* grads,
                  float* d_ray_bias, void* workspace, void* stream);
int nwb_mlp_wgrad(const nwb_field_desc* d, const void* packed, const float* raw, const float* d_raw,
                  const void* saved, int n_rays, int S, uint32_t mode, const nwb_field_grads* grads,
                  float* d_ray_bias, void* workspace, void* stream);

/* backward of nwb_ray_consts_fwd: d_ray_bias -> dir_encoding / transient_encoding.0 weight
 * columns [256:] (+ their biases) and, scattered by ts, the embedding-table gradients. */
int nwb_ray_consts_bwd(const nwb_field_desc* d, const nwb_field_params* p, const float* ray_feat,
                       const float* d_ray_bias, const int64_t* ts, int n_vocab, int n_rays,
                       uint32_t mode, const nwb_field_grads* grads, float* d_emb_a,
                       float* d_emb_t, float* d_a_rows, float* d_t_rows, void* stream);

/* ------------------------------------------------------------------------------------------
 * optimizer step on flat buffers -- utils/__init__.py:24-33 (torch.optim.Adam(lr, eps=1e-8,
 * weight_decay)), train.py:139-142.  p, g, m, v: n fp32 each (16 B aligned); `step` counts from 1;
 * g is multiplied by grad_scale first (1/world_size turns an allreduce-sum into DDP's mean);
 * zero_grad != 0 clears g afterwards (the kernels of the next step accumulate into it).
 * ---------------------------------------------------------------------------------------- */
int nwb_adam_step(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                  float eps, float weight_decay, int step, float grad_scale, int zero_grad, void* stream);

/* same update with the step state resident on the device, so that a captured CUDA graph of the training step
 * can be replayed: state = 16 B {int32 step; float lr; float lr/(1-beta1^step); float 1/sqrt(1-beta2^step)};
 * the call first advances step (and refreshes the two derived scalars), then applies the update.  The host
 * writes lr (schedulers, utils/__init__.py:44-61) and may zero step. */
int nwb_adam_step_dev(float* p, float* g, float* m, float* v, int64_t n, float beta1, float beta2, float eps,
                      float weight_decay, float* state, float grad_scale, int zero_grad, void* stream);

/* (f2) the same update in ONE launch that also refreshes the packed operand caches of
 * up to 4 fields whose parameters live inside the flat buffer [p, p+n): every weight-image element and constant
 * slot maps to one parameter, which is updated in place and written to the cache in the same pass (ends the
 * per-step nwb_field_pack launches); `rest` lists the flat offsets no cache refers to (int32, from
 * nwb_adam_pack_cover).  state is 48 B, zeroed by the caller once: the 16 B of nwb_adam_step_dev, a uint32 ticket, the
 * int32 step for which the two doubles that follow (beta1^step, beta2^step, kept as running products) are valid;
 * the step advances inside the launch.  Results are identical to
 * nwb_adam_step_dev followed by nwb_field_pack (tests/test_optim_gpu.py).
 * zero_grad | NWB_ADAM_NO_ADVANCE: the launch updates its fields with the scalars of step state+1 but leaves the state where
 * it is -- one optimizer step split over several launches (a model whose backward is done is updated while the other's
 * still runs); the LAST launch of the step comes without the flag and after the others have finished. */
#define NWB_ADAM_NO_ADVANCE 2
int nwb_adam_pack_step_dev(float* p, float* g, float* m, float* v, int64_t n, float beta1, float beta2, float eps,
                           float weight_decay, float* state, float grad_scale, int zero_grad, int n_fields,
                           const nwb_field_desc* const* descs, const nwb_field_params* const* params,
                           void* const* packed, const int32_t* rest, int64_t n_rest, void* stream);
/* setup: cover[off] = 1 for every flat offset the caches of these fields are built from (cover: n bytes, zeroed
 * by the caller); the complement is the `rest` list above */
int nwb_adam_pack_cover(float* p, int64_t n, int n_fields, const nwb_field_desc* const* descs,
                        const nwb_field_params* const* params, void* const* packed, unsigned char* cover,
                        void* stream);

/* n <= 8 device-to-device copies in ONE launch: a training step's batch (rays, ts, rgbs; train.py:158-161 receives them from
 * the DataLoader) into the static input buffers of the step's CUDA graph.  Buffers must not overlap. */
int nwb_copy_segments(int n, const void* const* src, void* const* dst, const int64_t* bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * (f1) NerfWLoss -- losses.py:19-50 -- the objective over the maps render_rays returns, forward and
 * backward as one launch each.  terms[5] = {c_l, f_l, b_l, s_l, their sum} (device); rgb_fine / beta /
 * transient_sigmas may be NULL (coarse-only and no-transient variants, losses.py:37-45).  scratch:
 * nwb_nerfw_loss_scratch_bytes() bytes, zeroed ONCE by the caller, reused by every call (the kernel resets
 * its ticket; grid-wide reduction in block order, deterministic).  g_terms: host array of 5 DEVICE pointers to
 * the cotangents of terms[0..4] (any may be NULL = 0).
 * ---------------------------------------------------------------------------------------- */
size_t nwb_nerfw_loss_scratch_bytes(void);
int nwb_nerfw_loss_fwd(const float* rgb_coarse, const float* rgb_fine, const float* beta,
                       const float* transient_sigmas, const float* targets, int n_rays, int64_t n_ts,
                       float lambda_u, float coef, void* scratch, float* terms, void* stream);
/* nwb_nerfw_loss_fwd without the scratch argument, for callers that cannot allocate (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  It has no counterpart in the reference.
 * The partial sums and the ticket live in a file-scope device array of the library
 * (2.4 KB, one per GPU that loads the library, held until exit); the call allocates nothing and calls no allocating
 * routine: one 16-byte memset node (the ticket must start at zero) and the same kernel, so terms[0..4] are bit-identical
 * to nwb_nerfw_loss_fwd's.  n_rays <= NWB_LOSS_STATIC_MAX_RAYS, else NWB_E_ARG with a message.  Every call on a device
 * shares the array: calls that overlap on the GPU (different streams) race.  The first call on a device looks the
 * array's address up (cudaGetSymbolAddress): make it outside a capture. */
#define NWB_LOSS_STATIC_MAX_RAYS (1 << 20)
int nwb_nerfw_loss_fwd_static(const float* rgb_coarse, const float* rgb_fine, const float* beta,
                              const float* transient_sigmas, const float* targets, int n_rays, int64_t n_ts,
                              float lambda_u, float coef, float* terms, void* stream);
int nwb_nerfw_loss_bwd(const float* rgb_coarse, const float* rgb_fine, const float* beta,
                       const float* transient_sigmas, const float* targets, int n_rays, int64_t n_ts,
                       float lambda_u, float coef, const float* const* g_terms, float* d_rgb_coarse,
                       float* d_rgb_fine, float* d_beta, float* d_transient_sigmas, void* stream);

/* metrics.py:4-13 (mse, psnr) over n floats: out[2] = {mean((pred-gt)^2), -10 log10 of it} (device), one launch,
 * deterministic.  clip01 != 0 clamps pred to [0,1] first (eval.py:201).  scratch: nwb_nerfw_loss_scratch_bytes()
 * bytes, zeroed once by the caller, not shared with a concurrently running loss call. */
int nwb_mse_psnr(const float* pred, const float* gt, int64_t n, int clip01, void* scratch, float* out, void* stream);

/* ------------------------------------------------------------------------------------------
 * (f3) rays of a pixel range of one view, generated on the device -- datasets/ray_utils.py:5-55
 * (get_ray_directions: no +0.5 centring; get_rays: directions @ c2w[:, :3]^T, normalised; origin = c2w[:, 3])
 * and the (N,8) row [rays_o, rays_d, near, far] of datasets/blender.py:97-100.  Pixels are row-major,
 * pixel = j * W + i; c2w is a HOST pointer to 12 floats ((3,4) row-major), passed by value to the kernel.
 * rays: device (n_rays, 8) fp32, 16 B aligned.
 * ---------------------------------------------------------------------------------------- */
int nwb_gen_rays(int H, int W, float fx, float fy, float cx, float cy, const float* c2w_host, float near_,
                 float far_, int64_t first_pixel, int64_t n_rays, float* rays, void* stream);

/* (f4) eval.py:201-203: out[i] = uint8(clip(x[i], 0, 1) * 255) (truncation, as numpy's astype) over n floats */
int nwb_pack_rgb8(const float* x, int64_t n, unsigned char* out, void* stream);

/* ------------------------------------------------------------------------------------------
 * (e) multi-GPU gradient exchange over NVLink / NVSwitch peer memory -- replaces the all-reduce Lightning's DDP runs for
 * the reference (train.py:235-236; one process per GPU).  Every rank owns one exchange region
 *     [signal block 1 KB][flat gradient buffer: n floats][reduced slice: ceil(n / world) floats]
 * allocated by nwb_p2p_alloc (plain cudaMalloc), exported as a 64-byte CUDA IPC handle, opened by every other rank.
 * nwb_p2p_allreduce launches ONE kernel on `stream`: two-shot all-reduce (reduce-scatter by P2P loads, cross-GPU flag
 * barrier, all-gather by P2P loads), sums in rank order -> bit-identical result on all ranks; in place on the gradient
 * buffer at nwb_p2p_grad_offset() inside the own region.  No host synchronisation: capturable in a CUDA graph.  All
 * ranks must launch it the same number of times; a peer that never arrives makes the kernel trap after ~2 s.
 * ---------------------------------------------------------------------------------------- */
size_t nwb_p2p_region_bytes(int64_t n_floats, int world);
size_t nwb_p2p_grad_offset(void);
int nwb_p2p_alloc(size_t bytes, void** ptr);                       /* zero-filled */
int nwb_p2p_free(void* ptr);
int nwb_p2p_export(const void* ptr, unsigned char handle[64]);
int nwb_p2p_open(const unsigned char handle[64], void** ptr);       /* peer access enabled lazily */
int nwb_p2p_close(void* ptr);
/* regions[q]: base of rank q's region as mapped in this process (q == rank: the own pointer); counter: 16 zeroed bytes of
 * local device memory owned by the communicator */
int nwb_p2p_allreduce(void* const* regions, int rank, int world, int64_t n_floats, void* counter, void* stream);
/* first half only (reduce-scatter): afterwards slice q of the summed gradient sits in rank q's reduced buffer, at
 * nwb_p2p_reduced_offset(n) inside its region, nwb_p2p_slice_floats(n, world) floats per slice; the consumer is
 * nwb_adam_pack_step_p2p, which reads the slices from the peers itself (the all-gather half never runs) */
int nwb_p2p_reduce_scatter(void* const* regions, int rank, int world, int64_t n_floats, void* counter, void* stream);
int64_t nwb_p2p_slice_floats(int64_t n_floats, int world);
size_t nwb_p2p_reduced_offset(int64_t n_floats);
/* nwb_adam_pack_step_dev with the gradient taken from the ranks' reduced slices (g itself is only cleared) */
int nwb_adam_pack_step_p2p(float* p, float* g, float* m, float* v, int64_t n, float beta1, float beta2, float eps,
                           float weight_decay, float* state, float grad_scale, int zero_grad, int n_fields,
                           const nwb_field_desc* const* descs, const nwb_field_params* const* params,
                           void* const* packed, const int32_t* rest, int64_t n_rest, void* const* regions, int world,
                           void* stream);

/* ------------------------------------------------------------------------------------------
 * self-test of the tcgen05 tile engine (used by tests/ only): D(128,N) = A(128,K) * B(N,K)^T
 * from bf16 tile images, through the same smem layouts / descriptors the field kernels use.
 * ---------------------------------------------------------------------------------------- */
int nwb_selftest_umma(const void* a_img, const void* b_img, int N, int K, int variant,
                      const uint32_t* overrides /* host, may be NULL */, float* d_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NERFW_B200_H */
