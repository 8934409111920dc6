/*
 * b200_sde.h -- C ABI of libb200sde.so: the sm_100a (B200) kernels behind the reverse-SDE
 * predictor-corrector sampling hot path of XDzzzzzZyq/b-pinn-kalman-filter.
 *
 * Conventions
 *   - Plain pointers and sizes only; all pointers are DEVICE pointers unless a name ends in `_host`.
 *   - The caller owns every buffer; no entry point allocates device memory or synchronises.
 *   - Every launcher takes the CUDA stream to launch on and returns 0 (B200_OK) or a negative code:
 *       -1 bad argument, -2 unsupported shape, -3 driver entry point missing, -(1000+cudaError_t) launch failure.
 *   - dtype: 0 = fp32, 1 = bf16 (activations inside the score engine).  Accumulation is always fp32.
 *   - Channels-last ("nhwc") tensors are [B, H, W, C] with C contiguous; "planes" tensors are [N*C, H, W].
 *
 * Each entry point cites the reference interface it replaces (paths relative to the reference repo).
 */
#ifndef B200_SDE_H_
#define B200_SDE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st* cudaStream_t;
#endif

/* library / build info */
int b200_abi_version(void);
const char* b200_build_info(void);

/* ---- upfirdn2d ------------------------------------------------------------------------------
 * replaces: upfirdn2d_op.upfirdn2d pybind (op/upfirdn2d.cpp:12-23), kernels op/upfirdn2d_kernel.cu:49-207,
 *           python dispatcher op/upfirdn2d.py:145-156 (CPU restatement there: upfirdn2d_native :159-200). */
int b200_upfirdn2d_out_size(int in_size, int k, int up, int down, int pad0, int pad1);
int b200_upfirdn2d_f32(const float* in, const float* k, float* out, int n_planes, int in_h, int in_w, int kh, int kw,
                       int up_x, int up_y, int down_x, int down_y, int pad_x0, int pad_x1, int pad_y0, int pad_y1,
                       cudaStream_t stream);
int b200_upfirdn2d_nhwc(const void* in, const float* k, void* out, int dtype, int batch, int in_h, int in_w,
                        int channels, int kh, int kw, int up, int down, int pad0, int pad1, cudaStream_t stream);

/* ---- fused bias + activation ----------------------------------------------------------------
 * replaces: fused.fused_bias_act pybind (op/fused_bias_act.cpp:11-20), kernel op/fused_bias_act_kernel.cu:18-49.
 * act: 1 linear, 3 leaky-relu; grad: 0 fwd, 1 (uses sign of `ref`), 2 (zeros). bias/ref may be NULL. */
int b200_fused_bias_act_f32(float* out, const float* x, const float* bias, const float* ref, long long n,
                            long long step_b, int size_b, int act, int grad, float alpha, float scale,
                            cudaStream_t stream);

/* ---- Navier-Stokes simulator step -----------------------------------------------------------
 * replaces: ns_step_forward.update_{density,velocity,pressure} pybind (op/ns_step.cpp:45-107) and the kernels in
 * op/ns_step_kernel.cu:96-234.  Fields are [B, C, H, W] fp32.  `scratch` (B*2*H*W floats) is only needed when
 * bug_compat = 1 (reproduces the reference's cross-sample plane mix-up for B > 1, op/ns_step.cpp:70-72). */
int b200_ns_update_density(float* dens_out, const float* dens, const float* vel, int B, int H, int W, float dt, float dx,
                           cudaStream_t stream);
int b200_ns_update_velocity(float* vel_out, const float* vel, const float* pres, float* scratch, int B, int H, int W,
                            float dt, float dx, int bug_compat, cudaStream_t stream);
int b200_ns_update_pressure(float* pres_out, const float* pres, const float* vel, int B, int H, int W, float dt,
                            float dx, cudaStream_t stream);
/* one simulator step (pinn_kalman/simulator.py:55-57, ukf_utils.py:109-111): velocity -> pressure -> density into the
 * caller's output buffers (must not alias the inputs); three launches on `stream`, no allocation */
int b200_ns_step_fused(float* dens_out, float* vel_out, float* pres_out, const float* dens, const float* vel,
                       const float* pres, int B, int H, int W, float dt, float dx, cudaStream_t stream);
/* building blocks: v* = v - dt grad(p) (op/ns_step_kernel.cu:180-202) and CIP advection of an arbitrary plane set
 * (op/ns_step_kernel.cu:114-158); strides in elements. */
int b200_ns_velocity_nonadvect(float* vstar, const float* vel, const float* pres, int B, int H, int W, float dt,
                               float dx, cudaStream_t stream);
int b200_ns_cip_advect(float* out, const float* field, const float* vel, long long field_bstride,
                       long long out_bstride, int B, int H, int W, float dt, float dx, cudaStream_t stream);

/* ---- reverse-SDE state updates --------------------------------------------------------------
 * replaces the elementwise/reduction torch calls of sampling.py:181-187, 195-200, 213-233, 262-282, 299-319.
 * x, score, z, x_mean: [B, n_per] fp32.  z == NULL -> Philox4x32-10 noise generated in registers from
 * (seed, offset, elem_base + element index); elem_base (multiple of 4) is the shard's global element offset.
 * `step` != NULL -> coefficient index is *step for every sample (per-step tables) and the Philox stream id becomes
 * offset + *step * 16, so a captured CUDA graph draws fresh noise on every replay. */
int b200_sample_sumsq(float* ss_score, float* ss_noise, const float* score, const float* z, int B, long long n_per,
                      uint64_t seed, uint64_t offset, long long elem_base, const int* step, cudaStream_t stream);
int b200_langevin_norm_sums(float* sums2, const float* ss_score, const float* ss_noise, int B, cudaStream_t stream);
/* Batch-sharded runs (one process per GPU): the reference averages the Langevin norms over the WHOLE batch
 * (sampling.py:276-278), so the ranks' partial sums must meet.  b200_langevin_norm_sums_xchg is b200_langevin_norm_sums
 * fused with that exchange over NVLink peer memory: the reducing block stores its two (exact, fixed-point int64)
 * partials into every peer's CUDA-IPC-mapped slot table with a release store, waits for the peers' slots of the same
 * epoch and adds them in rank order -- no NCCL call, no host involvement, graph-capturable, and bit-identical to the
 * unsharded batch's sums.  b200_xchg_alloc / _open / _close / _free set the slot tables up once per sampler
 * (world <= b200_xchg_max_ranks() = 8 GPUs of one node); every rank must make the same sequence of calls. */
int b200_xchg_max_ranks(void);
int b200_xchg_alloc(void** table, unsigned char* handle64);
int b200_xchg_open(const unsigned char* handle64, void** table);
int b200_xchg_close(void* peer_table);
int b200_xchg_free(void* table);
int b200_langevin_norm_sums_xchg(float* sums2, const float* ss_score, const float* ss_noise, int B, void* const* tables,
                                 int world, int rank, cudaStream_t stream);
/* NO COUNTERPART IN THE REFERENCE.  Allocation-free form of the two calls
 * above for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it):
 * sums2[0] = sum_b ||score_b||, sums2[1] = sum_b ||z_b|| with the per-sample [2][B] scratch in a file-scope static
 * device array (128 KB, held by every GPU that loads the library until exit).  B <= 16384
 * (b200_langevin_norms_static_max_batch()); larger B returns -4 ("size beyond the static scratch capacity").  Calls no
 * cudaMalloc / cudaFree.  Every call shares the array, which keeps the last call's contents (nothing needs clearing:
 * entries [0, B) are written before they are read); calls that overlap on the GPU (different streams) race. */
int b200_langevin_norms_static_max_batch(void);
int b200_langevin_norms_static(float* sums2, const float* score, const float* z, int B, long long n_per, uint64_t seed,
                               uint64_t offset, long long elem_base, const int* step, cudaStream_t stream);
/* controllable generation (controllable_generation.py:44-54, 132-141): after a predictor / corrector update the known
 * part of the state is replaced by freshly noised data, y = T^-1(T(x)(1-m) + (a T(data) + s z) m), y_mean likewise with
 * the noise-free data term; T = identity (inpainting, M = Minv = NULL) or the 3x3 colour basis (colourisation, C == 3;
 * M, Minv are HOST pointers to 9 floats).  [B,C,H,W] fp32, C <= 4; a (NULL = 1), s: [B] or per-step tables with `step`. */
int b200_mask_mix(float* x, float* x_mean, const float* data, const float* mask, const float* z, const float* a,
                  const float* s, const int* step, const float* M, const float* Minv, int B, int C, long long HW,
                  uint64_t seed, uint64_t offset, long long elem_base, cudaStream_t stream);
int b200_sde_update(float* x, float* x_mean, const float* score, const float* z, const float* a, const float* b,
                    const float* c, const int* step, int langevin, const float* norm_sums, float inv_batch_total,
                    float snr, int B, long long n_per, uint64_t seed, uint64_t offset, long long elem_base,
                    cudaStream_t stream);
int b200_fill_normal(float* out, long long total, float scale, uint64_t seed, uint64_t offset, long long elem_base,
                     cudaStream_t stream);
int b200_step_advance(int* step, int delta, cudaStream_t stream);

/* ---- score-network building blocks (channels-last activations) -------------------------------
 * GroupNorm(+SiLU): replaces nn.GroupNorm + act pairs (models/layerspp.py:219-264, models/layers.py:618-642).
 * Two sources = the skip-connection concat of models/ncsnpp.py:318 without materialising it. */
int b200_gn_nsplit(int B, int HW);
int b200_gn_partial(const void* x1, int C1, const void* x2, int C2, int dtype, int B, int HW, int G, int nsplit,
                    float* partial, cudaStream_t stream);
int b200_gn_apply(const void* x1, int C1, const void* x2, int C2, int dtype, int B, int HW, int G, int nsplit,
                  const float* partial, const float* gamma, const float* beta, float eps, int act, void* out,
                  cudaStream_t stream);
int b200_gn_apply_stats(const void* x1, int C1, const float* st1, int T1, const void* x2, int C2, const float* st2,
                        int T2, int dtype, int B, int HW, int G, int nsplit, const float* gamma, const float* beta,
                        float eps, int act, void* out, cudaStream_t stream);
/* b200_gn_apply_stats uses a thread-per-group kernel (no shared memory, no barrier; bit-identical output) for images of
 * at most 64 pixels with at most 4 statistics slots.  on = 0 switches it off (A/B measurements, bit-identity tests);
 * returns the previous setting.  Not a reference interface. */
int b200_gn_small_enable(int on);
/* layout changes at the API boundary (NCHW fp32 <-> channels-last), incl. `2x-1` (ncsnpp.py:261) and `h/sigma` (:379) */
int b200_nchw_to_nhwc(const float* x, void* out, int dtype, int B, int C, int H, int W, int Cp, float scale,
                      float shift, cudaStream_t stream);
int b200_nhwc_to_nchw(const void* x, float* out, int dtype, int B, int C, int H, int W, int Cs, const float* mul,
                      cudaStream_t stream);
/* time embedding: kind 0 Gaussian Fourier (layerspp.py:32-41), 1 sinusoidal (layers.py:500-514) */
int b200_temb_embed(const float* labels, const int* step, int kind, const float* fourier_w, int nf, const float* sigmas,
                    int n_sigmas, int scale_by_sigma, float* emb, float* out_mul, int B, cudaStream_t stream);
/* out[b,n] = act_out(sum_k act_in(x[b,k]) W[n,k] + bias[n])  -- temb MLP and all Dense_0 layers at once */
int b200_linear_f32(const float* x, const float* W, const float* bias, float* out, int B, int K, int N, int act_in,
                    int act_out, cudaStream_t stream);
int b200_add_scale(const void* a, const void* b, void* out, int dtype, long long n, float s, cudaStream_t stream);
int b200_nearest_up2_nhwc(const void* x, void* out, int dtype, int B, int H, int W, int C, cudaStream_t stream);
/* Strict-mode convolution on the tensor cores: fp32 NHWC activations, fp32 weights [Cout][kh][kw][Cin] pre-split into
 * (w_hi, w_lo) = (tf32(w), w - tf32(w)) by b200_split_tf32; tcgen05.mma kind::tf32, three-term split x w ~= xh wh +
 * xl wh + xh wl with fp32 accumulation (terms = 3: fp32-faithful, for `precision="fp32"`), or xh wh only (terms = 1: plain
 * TF32, what cudnn.allow_tf32 gives the reference).  out = ((conv + bias[n] + temb[b,n]) + res) * out_scale, fp32 NHWC.
 * replaces nn.Conv2d 3x3 / 1x1 and NIN in fp32 (models/layers.py:96-118, 537-546).  C1 % 32 == C2 % 32 == 0,
 * Cout % 128 == 0, ksize 1 or 3 (stride 1, pad ksize / 2). */
int b200_split_tf32(const float* w, float* hi, float* lo, long long n, cudaStream_t stream);
int b200_conv2d_tc32_supported(int C1, int C2, int B, int H, int W, int Cout, int ksize);
int b200_conv2d_tc32(const float* x1, int C1, const float* x2, int C2, const float* w_hi, const float* w_lo,
                     const float* bias, const float* temb, int temb_stride, const float* res, float out_scale,
                     float* out, int B, int H, int W, int Cout, int ksize, int terms, cudaStream_t stream);
/* the same kernel as a batched GEMM, out[b][m][n] = scale * sum_k a[b][m][k] bm[b][n][k] (a's rows lda floats apart, bm
 * pre-split), and a row softmax: the strict-mode attention q k^T -> softmax -> p v (models/layerspp.py:82-86) */
int b200_gemm_tc32_batched(const float* a, long long lda, const float* b_hi, const float* b_lo, float* out, int B, int M,
                           int N, int K, float scale, int terms, cudaStream_t stream);
int b200_softmax_rows_f32(float* x, long long rows, int n, cudaStream_t stream);
/* qkv [B][L][3C] -> k as (hi, lo) [B][L][C] and v transposed as (hi, lo) [B][C][L]: the B operands of the two products */
int b200_attn32_prep(const float* qkv, float* k_hi, float* k_lo, float* vt_hi, float* vt_lo, int B, int L, int C,
                     cudaStream_t stream);
/* implicit-GEMM convolution on CUDA cores (strict fp32 path + odd shapes); epilogue
 * out = ((conv + bias[n] + temb[b,n]) + res) * out_scale.  replaces nn.Conv2d / NIN (models/layers.py:96-118,537-546) */
int b200_conv2d_simt(const void* x1, int C1, const void* x2, int C2, const void* w, const float* bias,
                     const float* temb, int temb_stride, const void* res, float out_scale, void* out, int out_mode,
                     const float* sample_mul, int dtype, int B, int H, int W, int Ho, int Wo, int Cout, int KH, int KW,
                     int stride, int pad_t, int pad_l, cudaStream_t stream);
/* the same convolution on the tcgen05 tensor cores (bf16 channels-last, stride 1, "same" padding, ksize 1 or 3,
 * C1 % 64 == C2 % 64 == 0, Cout % 128 == 0): TMA-fed implicit GEMM with TMEM accumulators, same fused epilogue.
 * w: packed [Cout][ksize*ksize*(C1+C2)] bf16.  b200_conv2d_tc_supported returns 1 when the shape is taken.
 * out_mode 0: channels-last bf16 [B,H,W,Cout].  out_mode 1 (network head, Cout <= 16): NCHW fp32 [B,Cout,H,W] scaled
 * by sample_mul[b] (`h / used_sigmas`, models/ncsnpp.py:377-379); w then has Cout rows.  out_mode 2 (Cout == 8 stored
 * channels, the 3-channel image pyramids of progressive='output_skip', models/ncsnpp.py:343-365): channels-last bf16
 * [B,H,W,8], out = (conv + bias + res) * out_scale with res of the same layout; w has 8 rows (unused ones zero). */
int b200_conv2d_tc_supported(int C1, int C2, int B, int H, int W, int Cout, int ksize, int stride, int out_mode);
int b200_conv2d_tc(const void* x1, int C1, const void* x2, int C2, const void* w, const float* bias, const float* temb,
                   int temb_stride, const void* res, float out_scale, void* out, int out_mode, const float* sample_mul,
                   float* stats, int B, int H, int W, int Cout, int ksize, cudaStream_t stream);
/* `stats` (may be NULL): by-product for the GroupNorm that consumes `out` -- per (image, slot, channel) sum and sum of
 * squares of the bf16 output, [B][slots][Cout][2] fp32 with slots = b200_conv2d_tc_stat_slots(B, H, W); consumed by
 * b200_gn_apply_stats, so the separate statistics pass over the activation disappears. */
int b200_conv2d_tc_stat_slots(int B, int H, int W);
/* GroupNorm + SiLU of the INPUT fused into the convolution's operand pipeline: `h = act(GroupNorm(x)); h = Conv(h)` of
 * the residual blocks (models/layerspp.py:242-246, 264-266; models/layers.py:640-647) without ever writing h.
 * b200_gn_coef folds the producers' `stats` by-products into coef = [B][C1+C2][2] fp32 (scale/2, shift/2) per (image,
 * channel) of GroupNorm(G groups, eps) over cat(x1, x2); b200_conv2d_tc_gn then convolves the RAW x1 / x2: extra warps
 * rewrite each TMA-landed activation stage in shared memory (h = fma(x, scale/2, shift/2); h + h tanh(h)) before the
 * MMA reads it; zero padding stays zero.  Bit-identical to b200_gn_apply_stats followed by b200_conv2d_tc.  Taken for
 * the 3x3 shapes b200_conv2d_tc_gn_supported accepts (the halo + CTA-pair tile variants: H % 16 == 0, W % 8 == 0). */
int b200_gn_coef(const float* st1, int T1, int C1, const float* st2, int T2, int C2, int B, int HW, int G,
                 const float* gamma, const float* beta, float eps, float* coef, cudaStream_t stream);
/* [B][T][C][2] -> [B][1][C][2]: the per-tile statistics slots of a producing convolution summed once, in a fixed order,
 * for the GroupNorm consumers of large images (T = 256 / 1024 at 128x128 / 256x256) */
int b200_gn_fold_stats(const float* stats, float* out, int B, int T, int C, cudaStream_t stream);
int b200_conv2d_tc_gn_supported(int C1, int C2, int B, int H, int W, int Cout, int ksize);
/* Conv_1 of a residual block with its 1x1 shortcut `Conv_2(x)` folded in (models/layerspp.py:266-274):
 * out = (conv3x3(h; W1) + conv1x1(cat(xs1, xs2); W2) + bias) * out_scale with bias = b1 + b2 and
 * w = [Cout][9 (C1 + C2) + Cs1 + Cs2] = [W1 | W2]; the shortcut's K blocks ride the same K loop at the centre tap, so
 * Conv_2's launch, its output tensor and Conv_1's residual read disappear.  gn_coef (may be NULL) as b200_conv2d_tc_gn.
 * Shapes: whatever b200_conv2d_tc_supported(C1, C2, B, H, W, Cout, 3, 1, 0) takes, Cs1 % 64 == Cs2 % 64 == 0. */
int b200_conv2d_tc_sc(const void* x1, int C1, const void* x2, int C2, const float* gn_coef, const void* xs1, int Cs1,
                      const void* xs2, int Cs2, const void* w, const float* bias, const float* temb, int temb_stride,
                      float out_scale, void* out, float* stats, int B, int H, int W, int Cout, cudaStream_t stream);
/* 1 where the fused form is also the FASTER one (256-wide tiles; the 128-wide ones are shared-memory-bandwidth bound
 * and lose to conv + standalone GroupNorm): what the engine's planner asks */
int b200_conv2d_tc_gn_preferred(int C1, int C2, int B, int H, int W, int Cout, int ksize);
int b200_conv2d_tc_gn(const void* x1, int C1, const void* x2, int C2, const float* gn_coef, const void* w,
                      const float* bias, const float* temb, int temb_stride, const void* res, float out_scale, void* out,
                      float* stats, int B, int H, int W, int Cout, int ksize, cudaStream_t stream);
/* 3x3 stride-2 VALID conv (zero beyond the input extent) on the same tensor-core kernel (TMA element strides):
 * F.conv2d(stride=2) of conv_downsample_2d (models/up_or_down_sampling.py:174-178) on the FIR-padded input, and
 * DDPM's pad(0,1,0,1)+conv3x3/2 (models/layers.py:596-608).  x [B,Hin,Win,C] bf16, out [B,Ho,Wo,Cout] bf16. */
int b200_conv2d_tc_s2_supported(int C, int B, int Ho, int Wo, int Cout);
int b200_conv2d_tc_s2(const void* x, int C, const void* w, const float* bias, const void* res, float out_scale,
                      void* out, float* stats, int B, int Hin, int Win, int Ho, int Wo, int Cout, cudaStream_t stream);
/* softmax(q k^T scale) v per image on CUDA cores (models/layerspp.py:82-86); ld* = row strides in elements */
int b200_attention_simt(const void* q, const void* k, const void* v, void* out, int dtype, int B, int L, int C,
                        int ldq, int ldk, int ldv, int ldo, float scale, cudaStream_t stream);

/* the same attention on the tcgen05 tensor cores for L = 256 tokens, C = 256 channels (every shipped network's
 * attention shape): qkv = fused projection [B, L, 3C] bf16 (q | k | v), out [B, L, C] bf16; scores stay in TMEM. */
int b200_attention_tc_supported(int L, int C);
int b200_attention_tc(const void* qkv, void* out, int B, int L, int C, float scale, cudaStream_t stream);

/* ---- grid_sample (op/grid_sample.py:15-22; models/flownet.py:8-25 is the caller) --------------------------------------
 * dim = 2: input [N,C,H,W], grid [N,Ho,Wo,2] (x,y in [-1,1]), out [N,C,Ho,Wo]; dim = 3: input [N,C,D,H,W],
 * grid [N,Do,Ho,Wo,3] (x,y,z), out [N,C,Do,Ho,Wo] (pass D = Do = 1 for dim 2).  fp32, contiguous.  Bilinear/trilinear,
 * padding_mode 0 = zeros, 1 = border (the two op/grid_sample.py:16 accepts), align_corners as in F.grid_sample.
 * fwd   replaces F.grid_sample(mode='bilinear')                       (op/grid_sample.py:38-39, 107-108)
 * bwd   replaces aten::grid_sampler_{2d,3d}_backward                  (op/grid_sample.py:54-60, 124-130);
 *       g_in is ACCUMULATED into (caller zero-fills it); g_in or g_grid may be NULL (output_mask).
 * grad2 replaces gridsample_grad2.grad2_{2d,3d}(g2_gin, g2_ggrid, g_out, input, grid, padding_mode, align_corners)
 *       -> [gg_out, g_in, g_grid] (op/grid_sample.cpp:26-56, op/grid_sample_kernel.cu:27-533); g_in accumulated. */
int b200_grid_sample_fwd(int dim, float* out, const float* in, const float* grid, int N, int C, int D, int H, int W,
                         int Do, int Ho, int Wo, int padding_mode, int align_corners, cudaStream_t stream);
int b200_grid_sample_bwd(int dim, float* g_in, float* g_grid, const float* g_out, const float* in, const float* grid,
                         int N, int C, int D, int H, int W, int Do, int Ho, int Wo, int padding_mode, int align_corners,
                         cudaStream_t stream);
int b200_grid_sample_grad2(int dim, float* gg_out, float* g_in, float* g_grid, const float* g2_gin, const float* g2_ggrid,
                           const float* g_out, const float* in, const float* grid, int N, int C, int D, int H, int W,
                           int Do, int Ho, int Wo, int padding_mode, int align_corners, cudaStream_t stream);

/* ---- input gradient of the score network (DPS: inverse/conditional_sampling.py:116-136 takes
 * torch.autograd.grad(logp, x_t) through x0_hat = x_t/mean + std^2 score(x_t), i.e. J_model^T v w.r.t. the image) ------
 * Convolution backward-data = the forward convolution entry points on flipped / transposed weights (packed by the host
 * layer); the remaining backward pieces are below.  No weight gradients. */
/* GroupNorm(+SiLU) backward over one or two concatenated sources; part_fwd = b200_gn_partial of the same sources */
int b200_gn_bwd_partial(const void* x1, int C1, const void* x2, int C2, int dtype, int B, int HW, int G, int nsplit,
                        const float* part_fwd, const float* gamma, const float* beta, float eps, int act, const void* dy,
                        float* part_bwd, cudaStream_t stream);
/* the same pass, also leaving the per-channel sums chan_sums[B][nsplit][C][2] = (sum dz xhat, sum dz): summed over B and
 * nsplit they are d gamma / d beta (loss.backward() on nn.GroupNorm's affine parameters, losses.py:196) */
int b200_gn_bwd_partial2(const void* x1, int C1, const void* x2, int C2, int dtype, int B, int HW, int G, int nsplit,
                         const float* part_fwd, const float* gamma, const float* beta, float eps, int act,
                         const void* dy, float* part_bwd, float* chan_sums, cudaStream_t stream);
int b200_gn_bwd_apply(const void* x1, int C1, const void* x2, int C2, int dtype, int B, int HW, int G, int nsplit,
                      const float* part_fwd, const float* part_bwd, const float* gamma, const float* beta, float eps,
                      int act, const void* dy, void* dx1, int acc1, void* dx2, int acc2, cudaStream_t stream);
/* out = sa * a + sb * b (b may be NULL): residual branches and gradient accumulation */
int b200_axpby(const void* a, float sa, const void* b, float sb, void* out, int dtype, long long n, cudaStream_t stream);
/* adjoint of b200_nearest_up2_nhwc: dx [B,H,W,C] (+)= 2x2 block sums of dy [B,2H,2W,C] */
int b200_nearest_up2_bwd_nhwc(const void* dy, void* dx, int dtype, int B, int H, int W, int C, int acc, cudaStream_t stream);
/* out [B,Ho,Wo,C] = 0 except out[:, 2i+1, 2j+1] = dy[:, i, j]: makes the backward-data of a 3x3 stride-2 VALID
 * convolution a 3x3 stride-1 "same" convolution with the flipped kernel (models/layers.py:596-608) */
int b200_zero_stuff2_nhwc(const void* dy, void* out, int dtype, int B, int Hi, int Wi, int Ho, int Wo, int C,
                          cudaStream_t stream);
/* backward of softmax(q k^T scale) v (models/layers.py:566-570): qkv, dqkv [B,L,3C]; dout [B,L,C];
 * scratch 2*B*L*L floats */
int b200_attention_bwd(const void* qkv, const void* dout, void* dqkv, float* scratch, int dtype, int B, int L, int C,
                       float scale, cudaStream_t stream);

/* ---- adaptive Runge-Kutta on the device (replaces the scipy.integrate.solve_ivp round trips of
 * sampling.py:459-473, likelihood.py:86-99, inverse/conditional_sampling.py:10-29: the state stays in HBM, only the
 * 8-byte error norm crosses to the host per step).  y, y_out fp64 [n]; K fp32 stage derivatives, stage j at
 * K + j*kstride; a_host / e_host: ns coefficients in HOST memory (passed to the kernel by value). */
#define B200_RK_MAX_STAGES 7
#define B200_RK_PARTIALS 1184
/* x32_out (may be NULL) = fp32(y + h sum_j a[j] K[j]); y_out (may be NULL) = the same in fp64 */
int b200_rk_stage(float* x32_out, double* y_out, const double* y, const float* K, long long kstride, int ns,
                  const double* a_host, double h, long long n, cudaStream_t stream);
/* partial[0..B200_RK_PARTIALS) = block sums of (h sum_j e[j] K[j] / (atol + rtol max(|y|,|y_new|)))^2 (unused slots 0);
 * ns = 7 (RK45) or 4 (RK23).  error_norm = sqrt(sum(partial) / n) as scipy/integrate/_ivp/rk.py:108-109 */
int b200_rk_error(double* partial, const double* y, const double* y_new, const float* K, long long kstride, int ns,
                  const double* e_host, double h, double atol, double rtol, long long n, cudaStream_t stream);

/* attention backward on tensor cores (bf16, L = C = 256; warp-level MMA): what b200_attention_bwd dispatches to */
int b200_attention_bwd_mma_supported(int L, int C);
int b200_attention_bwd_mma(const void* qkv, const void* dout, void* dqkv, void* scratch, int B, int L, int C, float scale,
                           cudaStream_t stream);
/* ---- training step (losses.py:165-222 step_fn: loss.backward(); optimize_fn = warm-up + clip_grad_norm_ + Adam,
 * losses.py:48-66; models/ema.py:36-56) -------------------------------------------------------------------------------
 * Weight gradients are contracted by b200_conv2d_tc / b200_conv2d_simt over K = pixels: b200_transpose2d gives dY^T
 * [Cout][P], b200_im2col_t the transposed im2col [(tap, ci)][P]; the product lands in the packed weight layout.
 * b200_colsum: bias / time-embedding gradients; b200_gn_param_grad: GroupNorm affine gradients (per-sample partials);
 * b200_gemm_f32 / b200_silu_bwd_f32: backward of the time-embedding MLP (models/ncsnpp.py:251-255);
 * b200_score_loss: SDE / SMLD / DDPM score-matching losses (losses.py:69-163) and their cotangent;
 * b200_dropout: nn.Dropout in train mode (layerspp.py:265), mask regenerated from (seed, offset, step) for the backward;
 * b200_grad_sumsq + b200_adam_ema: gradient norm, clipping, Adam and the EMA shadow in one pass over flat buffers. */
int b200_transpose2d(const void* in, void* out, int dtype, long long R, int C, cudaStream_t stream);
int b200_im2col_t(const void* x1, int C1, const void* x2, int C2, void* out, int dtype, int B, int H, int W, int Ho, int Wo,
                  int ksize, int stride, int pad, cudaStream_t stream);
int b200_colsum(const void* in, float* out, int dtype, int Bn, long long R, int C, long long out_stride, float scale,
                int accumulate, cudaStream_t stream);
int b200_gn_param_grad(const void* x1, int C1, const void* x2, int C2, const void* dy, int dtype, int B, int HW, int G,
                       int nsplit, const float* stats, const float* gamma, const float* beta, float eps, int act,
                       float* partial, cudaStream_t stream);
int b200_gemm_f32(float* C, const float* A, const float* B, int M, int N, int K, int ta, int tb, float alpha, float beta,
                  cudaStream_t stream);
int b200_silu_f32(float* y, const float* x, long long n, cudaStream_t stream);
int b200_silu_bwd_f32(float* dx, const float* dy, const float* x, long long n, cudaStream_t stream);
int b200_score_loss(float* losses, float* loss, float* dout, const float* out, const float* z, const float* alpha,
                    const float* beta, const float* weight, const float* mul, float r, int B, long long n_per,
                    cudaStream_t stream);
int b200_dropout(void* y, const void* x, int dtype, long long n, float p, uint64_t seed, uint64_t offset, const int* step,
                 cudaStream_t stream);
int b200_grad_sumsq(float* sumsq, const float* g, long long n, cudaStream_t stream);
int b200_adam_ema(float* p, float* m, float* v, float* ema, const float* g, const float* hyper, const float* sumsq,
                  float grad_scale, long long n, cudaStream_t stream);
/* split-K batched GEMM on the tensor-core 1x1 path: out[s][m][n] = sum over the s-th of S K chunks of a[m][k] bm[n][k]
 * (bf16, row strides lda / ldb elements); the weight-gradient GEMMs over K = pixels (autograd's conv weight backward,
 * losses.py:196).  M % 128 == N % 128 == 0, (K / S) % 64 == 0.  b200_sum_partials: out[i] = sum_s in[s][i] (fp32), i < n, n % 8 == 0. */
/* weight gradient of a stride-1 "same" convolution, split over S chunks of pixels, straight from the channels-last
 * tensors (no transpose, no im2col): out[s][co][(tap, ci)] = sum_{p in chunk s} gy[p][co] x[p + tap][ci]; gy = dY [P][M],
 * x1 / x2 = the conv's source(s) [P][C] (P = B H W); MN-major MMA operands, the tap shift and the zero padding are TMA box
 * coordinates.  M % 128 == 0, C % 64 == 0, (P / S) % 64 == 0, rows of <= 64 pixels tile 64-pixel K blocks. */
int b200_wgrad_tc_supported(int M, int C1, int C2, int B, int H, int W, int ksize, int S);
int b200_wgrad_tc(const void* gy, const void* x1, int C1, const void* x2, int C2, void* out, int M, int B, int H, int W,
                  int ksize, int S, cudaStream_t stream);
int b200_sum_partials(const void* in, float* out, int dtype, int S, long long n, cudaStream_t stream);
int b200_gemm_tc_splitk(const void* a, long long lda, const void* bm, long long ldb, void* out, int M, int N, long long K,
                        int S, cudaStream_t stream);
/* One launch for a table of <= 4-D strided copies with scale and fp32 <-> bf16 conversion (dst[i] = scale * (src[j(i)] +
 * src2[j(i)])): what keeps the plans' packed / flipped weight copies and the flat gradient buffer in step with the flat
 * parameters after optimizer.step() (the reference re-reads its nn.Parameters in place; no counterpart).  Descriptor
 * layout: csrc/train.cu CopyDesc; built by models/copylist.py. */
int b200_copy_desc_bytes(void);
int b200_copy_chunk(void);
int b200_copy_batch(const void* descs, const void* blocks, int nblocks, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_SDE_H_ */
