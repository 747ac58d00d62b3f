/*
 * gance_b200.h - C ABI of libgance_b200.so: the B200 (sm_100a) kernels behind GANce's
 * audio -> latents -> StyleGAN2 frames path.
 *
 * What this replaces. The reference (esologic/GANce) is pure Python and has no FFI of its own;
 * its device code is reached through TF1/dnnlib calls made in
 *   gance/network_interface/network_functions.py:151-158  Gs.run(z, None, truncation_psi=1.2, ...)
 *   gance/network_interface/network_functions.py:167-169  Gs.components.synthesis.run(w+, ...)
 * whose CUDA side is the (missing) submodule gance/stylegan2 = NVlabs/stylegan2:
 *   dnnlib/tflib/ops/fused_bias_act.cu (FusedBiasAct), dnnlib/tflib/ops/upfirdn_2d.cu (UpFirDn2D),
 *   cuDNN grouped conv inside training/networks_stylegan2.py:modulated_conv2d_layer,
 * and the numpy/scipy audio code in
 *   gance/apply_spectrogram.py:20-118, gance/vector_sources/vector_sources_common.py:136-428,
 *   gance/data_into_network_visualization/visualization_inputs.py:53-270.
 * Each entry below cites the reference code it stands in for. The Python host
 * (gance_b200/network_interface, gance_b200/vector_sources) binds these with ctypes and keeps the
 * reference's call signatures; INTEGRATION.md shows the stub a GANce maintainer would add.
 *
 * Conventions (every entry):
 *   - returns 0 on success, a positive GANCE_E_* code for a rejected argument, or
 *     -(cudaError_t) if a launch failed; gance_last_error_string() has the message;
 *   - never throws, never synchronises, never allocates device memory (callers pass scratch);
 *   - every pointer is a DEVICE pointer unless named host_*, owned by the caller (PyTorch
 *     tensor.data_ptr()), alive until the work on `stream` has completed;
 *   - `stream` is a cudaStream_t passed as void*; calls on distinct streams are thread-safe;
 *   - activations are NHWC; `dtype` selects the activation/weight element type:
 *       GANCE_F32  = 0  fp32 validation mode (CUDA-core FMA, fp32 everywhere)
 *       GANCE_BF16 = 1  bf16 operands, fp32 accumulate in TMEM (tcgen05), fp32 epilogue
 *       GANCE_F16  = 2  as GANCE_BF16 with IEEE half operands (same kernels, same speed,
 *                       11-bit instead of 8-bit significand; |x| must stay below 65504)
 *   - there is NO CPU fallback: without a CUDA device every compute entry fails.
 */
#ifndef GANCE_B200_H
#define GANCE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GANCE_ABI_VERSION 1

#define GANCE_F32 0
#define GANCE_BF16 1
#define GANCE_F16 2

#define GANCE_OK 0
#define GANCE_E_NULL 1        /* a required pointer was NULL */
#define GANCE_E_SHAPE 2       /* a dimension is unsupported by this kernel */
#define GANCE_E_DTYPE 3       /* unknown dtype code */
#define GANCE_E_ALIGN 4       /* a pointer is not aligned as the kernel requires (16 B) */
#define GANCE_E_NOT_INIT 5    /* gance_init() has not succeeded on this device */
#define GANCE_E_DRIVER 6      /* CUDA driver entry point (TMA descriptor encode) unavailable */
#define GANCE_E_NONFINITE 7   /* a rendered frame held inf/NaN pixels: a 16-bit activation overflowed (see
                                 gance_set_nonfinite_flag); raised by the host binding after a batch, never by a launch */

/* ---- lifecycle ------------------------------------------------------------------------------
 * Replaces dnnlib.tflib.init_tf() + custom_ops.get_plugin() (nvcc build of the TF plugin) called
 * from network_functions.py:101-103: selects the device, checks it is sm_100, resolves the
 * driver's cuTensorMapEncodeTiled and raises the kernels' dynamic shared-memory limits. */
int gance_abi_version(void);
int gance_init(int device);
const char* gance_last_error_string(void);
/* Number of kernel launches issued through this library since load (bench.py's gpu_launches). */
unsigned long long gance_launch_count(void);
/* Range guard of the 16-bit modes (no counterpart in the reference, whose arithmetic is fp32 throughout). Activations
 * are stored as half / bfloat16 pre-multiplied by the next layer's style; a network with a large style overflows half's
 * 65504 to inf, which every later layer carries into the RGB sum. Registers a device int (zeroed by the caller) into
 * which every entry that writes uint8 frames (gance_modconv3x3_torgb, gance_torgb_skip, gance_to_uint8_nhwc) ORs 1 when
 * a pixel of the fp32 image it converts is inf or NaN - the flag is sticky until the caller clears it. NULL unregisters.
 * The registration is per device (the calling thread's current device), process-wide, and is read at launch time (a
 * CUDA graph keeps the pointer it was captured with).
 * The host reads the flag back with the frames and reports GANCE_E_NONFINITE. */
int gance_set_nonfinite_flag(int* device_flag);
/* Tuning switch (no counterpart in the reference): how the generic 3x3 implicit GEMM (Cout % 128 == 0, plain or fused
 * ToRGB) is scheduled. 0 = one CTA per tile (tcgen05 cta_group::1); 1 (default) = CTA pairs on the two SMs of a TPC
 * sharing one M = 256 MMA (cta_group::2: each SM reads its own pixels and HALF of the weight tile) for layers with at
 * least one work item per co-resident pair; 2 = CTA pairs whenever the shape is eligible (tests: small shapes, odd tile
 * counts). Results are the same in every mode. Process-wide, read at launch time; returns the previous mode. */
int gance_set_conv_pairing(int mode);
/* Tuning switch (no counterpart in the reference): 1 (default) = the 32 -> 32-channel SAME convolution (1024^2 Conv1 of
 * config-f) runs with TWO image rows per accumulator row (N = 64 MMAs, tiles of 32 x 8 pixels) when H % 32 == 0; 0 = one row
 * per accumulator row (N = 32 MMAs, tiles of 16 x 8). Results are bit-identical. Process-wide, read at launch time; returns
 * the previous setting. */
int gance_set_conv_two_rows(int enabled);

/* ---- G_mapping + truncation ------------------------------------------------------------------
 * networks_stylegan2.G_mapping + the 'Truncation' block of G_main, as run by Gs.run
 * (network_functions.py:151-158). z [B,512] f32 -> wplus [B,num_ws,512] f32.
 * weights [8][512][512] (in,out) and biases [8][512] are pre-multiplied by their runtime
 * coefficients (lrmul/sqrt(fan_in), lrmul). psi == 1 skips the lerp. One launch per dense layer (the batch
 * spreads each layer over every SM); the layers ping-pong between rows 0 and 1 of w+, so num_ws >= 2. */
int gance_mapping(const float* z, int batch, const float* weights, const float* biases,
                  const float* dlatent_avg, float truncation_psi, int num_ws, float* wplus,
                  void* stream);

/* ---- style affine + demodulation coefficients -------------------------------------------------
 * The `s = dense(w) + bias + 1` and `d = rsqrt(sum (w*s)^2 + 1e-8)` lines of
 * modulated_conv2d_layer. styles[b,ci] = wplus[b,w_index,:] @ mod_weight[:,ci] + mod_bias[ci] + 1;
 * demod[b,co] = rsqrt(sum_ci styles[b,ci]^2 * wsq[ci,co] + 1e-8) with wsq = sum over taps of the
 * squared runtime-scaled conv weight. wsq/demod may be NULL (ToRGB: demodulate=False).
 * mod_weight [512,Cin] is pre-multiplied by 1/sqrt(512). */
int gance_style_demod(const float* wplus, int batch, int num_ws, int w_index,
                      const float* mod_weight, const float* mod_bias, const float* wsq, int c_in,
                      int c_out, float* styles, float* demod, void* stream);

/* ---- styles + demodulation coefficients of every layer at once ---------------------------------
 * The same arithmetic as gance_style_demod for ALL modulated layers of a generator in two launches
 * (the per-layer entry costs one launch per layer: 26 at 1024^2). mod_weight_cat [512,total_cols] is
 * the column-wise concatenation of the layers' (pre-scaled) affine weights, mod_bias_cat
 * [total_cols]. style_chunks [n][4] int32, one row per 32 columns: {w_index, first column of the
 * layer, Cin of the layer, first column of the chunk inside the layer}. demod_chunks [n][6] int32,
 * one row per 32 output channels of a demodulated layer: {float offset of the layer's wsq [Cin,Cout]
 * in wsq_cat, Cin, Cout, first style column of the layer, first output channel of the chunk, first
 * demod column of the layer}. Outputs are layer-major: layer L's [B,C] block starts at B*col0(L). */
int gance_styles_all(const float* wplus, int batch, int num_ws, const float* mod_weight_cat,
                     int total_cols, const float* mod_bias_cat, const int* style_chunks,
                     int n_style_chunks, float* styles_out, const float* wsq_cat,
                     const int* demod_chunks, int n_demod_chunks, float* demod_out, void* stream);

/* ---- learned constant input ------------------------------------------------------------------
 * tile(const, B) * styles: out[b,h,w,c] = konst[h,w,c] * styles[b,c]; konst is [4,4,C] f32. */
int gance_const_input(const float* konst, const float* styles, int batch, int channels, int dtype,
                      void* out, void* stream);

/* ---- per-sample channel scaling -----------------------------------------------------------------
 * out[b,h,w,c] = x[b,h,w,c] * styles[b,c]: the `ww = w * s` modulation of modulated_conv2d_layer moved
 * onto the activations, for inputs whose producer could not pre-scale them in its epilogue. */
int gance_scale_channels(const void* x, int dtype, int batch, int h, int w, int channels,
                         const float* styles, void* out, void* stream);

/* ---- 3x3 modulated convolution with fused epilogue ---------------------------------------------
 * modulated_conv2d_layer(kernel=3, demodulate=True) + noise + fused_bias_act('lrelu'), computed
 * as conv(x*s, W) * d (SURVEY.md Appendix A.4 identity): `x` is ALREADY multiplied by this
 * layer's styles. out = lrelu(acc*demod[b,co] + noise*noise_strength + bias[co], 0.2)*sqrt(2),
 * then * next_styles[b,co] when next_styles != NULL (pre-scaling for the next conv).
 *   x      [B, H+2-2*pad, W+2-2*pad, Cin]   pad = 1: SAME (zero border), pad = 0: VALID
 *   weight GANCE_BF16 / GANCE_F16: [Cout][3][3][Cin] bf16 / half (K-major GEMM operand, fed by TMA)
 *          GANCE_F32 : [3][3][Cin][Cout] f32 (the TF HWIO layout)
 *          both pre-multiplied by 1/sqrt(9*Cin)
 *   noise  [H,W] f32 (noise_batch_stride = 0) or [B,H,W] (stride = H*W); may be NULL
 *   out    [B,H,W,Cout]
 * GANCE_BF16 / GANCE_F16 run the tcgen05/TMEM implicit GEMM and need Cin % 32 == 0, Cout % 32 == 0. */
int gance_modconv3x3(const void* x, int dtype, int batch, int h_out, int w_out, int c_in,
                     int c_out, int pad, const void* weight, const float* demod,
                     const float* noise, long long noise_batch_stride, float noise_strength,
                     const float* bias, const float* next_styles, void* out, void* stream);

/* gance_modconv3x3 with TWO outputs from one epilogue: `out` = the activated layer output as G_synthesis_stylegan2's
 * block() hands it to torgb() (no next_styles), `out_scaled` = the same values * next_styles[b,co], i.e. the
 * modulated input `x * s` of the next block's Conv0_up (modulated_conv2d_layer's `x *= s`), rounded once from the
 * fp32 accumulator instead of being re-read and re-rounded by gance_scale_channels. GANCE_BF16 / GANCE_F16,
 * Cin % 64 == 0 and Cin >= 128 (the 4^2..64^2 Conv layers of config-e/f); arguments otherwise as gance_modconv3x3. */
int gance_modconv3x3_dual(const void* x, int dtype, int batch, int h_out, int w_out, int c_in,
                          int c_out, int pad, const void* weight, const float* demod,
                          const float* noise, long long noise_batch_stride, float noise_strength,
                          const float* bias, const float* next_styles, void* out_scaled, void* out,
                          void* stream);

/* ---- fp32-grade validation mode on the tensor core ("3xTF32") -----------------------------------
 * north_star: "tcgen05/TMEM implicit-GEMM ... bf16 with fp32 accumulate, plus an fp32/TF32 validation mode". The same
 * modulated 3x3 convolution + noise + fused_bias_act as gance_modconv3x3(GANCE_F32) - the reference's fp32 arithmetic in
 * networks_stylegan2.modulated_conv2d_layer - computed by tcgen05.mma.kind::tf32 with error compensation: every operand
 * is split x = hi + lo (hi = x with its low 13 significand bits cleared, exactly a tf32 value; lo = x - hi, exact) and
 * D += lo*hi + hi*lo + hi*hi with fp32 accumulation in TMEM (the dropped lo*lo term is ~2^-22 relative).
 * gance_split_tf32: x [B,H,W,C] f32 (times styles [B,C] when styles != NULL) -> hi, lo [B,H,W,C] f32. C % 4 == 0.
 * gance_modconv3x3_tf32x3: x_hi/x_lo [B,H+2-2pad,W+2-2pad,Cin] f32 (already modulated), w_hi/w_lo [Cout][3][3][Cin] f32
 * (runtime-scaled, split the same way), epilogue and remaining arguments as gance_modconv3x3; out [B,H,W,Cout] f32.
 * Cin % 32 == 0, Cout % 32 == 0. The up-convolutions of this mode run FIR-first: gance_upfir2x_pad, split, pad = 0. */
int gance_split_tf32(const float* x, const float* styles, int batch, int h, int w, int channels, float* hi,
                     float* lo, void* stream);
int gance_modconv3x3_tf32x3(const float* x_hi, const float* x_lo, int batch, int h_out, int w_out, int c_in,
                            int c_out, int pad, const float* w_hi, const float* w_lo, const float* demod,
                            const float* noise, long long noise_batch_stride, float noise_strength,
                            const float* bias, const float* next_styles, float* out, void* stream);

/* ---- 2x FIR upsampling feeding the up-convolution ----------------------------------------------
 * First half of upsample_conv_2d in FIR-first form (tests/test_oracle_stylegan2.py proves
 * conv_transpose(stride 2) -> upfirdn([1,3,3,1]*4, pad 1,1) == VALID 3x3 conv of this tensor):
 * u[b,j,i,c] over j,i in [0,2H+2): per axis u[2a] = .75 x[a-1] + .25 x[a], u[2a+1] = .25 x[a-1]
 * + .75 x[a], x[-1] = x[H] = 0; multiplied by styles[b,c] when styles != NULL.
 * Replaces UpFirDn2D (upfirdn_2d.cu) on this path. x [B,H,W,C] -> u [B,2H+2,2W+2,C]. */
int gance_upfir2x_pad(const void* x, int dtype, int batch, int h, int w, int channels,
                      const float* styles, void* u, void* stream);

/* ---- up-convolution = upfir2x_pad + VALID modconv3x3 ---------------------------------------------
 * modulated_conv2d_layer(up=True) + noise + bias/lrelu. `x` is the UNSCALED input [B,H,W,Cin];
 * styles [B,Cin] are applied inside; scratch holds [B,2H+2,2W+2,Cin] elements of `dtype`. */
int gance_modconv3x3_up2_fir(const void* x, int dtype, int batch, int h_in, int w_in, int c_in,
                             int c_out, const float* styles, const void* weight,
                             const float* demod, const float* noise,
                             long long noise_batch_stride, float noise_strength, const float* bias,
                             const float* next_styles, void* scratch, void* out, void* stream);

/* ---- up-convolution as ONE implicit GEMM at the input resolution ("folded") -----------------------
 * modulated_conv2d_layer(up=True) = conv2d_transpose(stride 2) -> upfirdn_2d([1,3,3,1]) is linear in x,
 * so each of the 4 output phases (py,px) is a SAME 3x3 convolution of x with its own 3x3 weight
 *   Wf[py,px][dy][dx] = sum_{ky,kx} A_py[dy][ky] A_px[dx][kx] w[ky][kx],
 *   A_0 = [[.75,.25,0],[.25,.75,.75],[0,0,.25]], A_1 = [[.25,0,0],[.75,.75,.25],[0,.25,.75]]
 * (tests/test_oracle_stylegan2.py proves the identity against the transposed-conv + upfirdn form).
 * One GEMM with N = 4*Cout computes all phases from the same A tiles; the epilogue scatters column
 * n = (py*2+px)*Cout + co to pixel (2h+py, 2w+px). No (2H+2)^2 intermediate ever reaches HBM.
 *   x [B,H,W,Cin] ALREADY multiplied by this layer's styles; weight_folded [4*Cout][3][3][Cin]
 *   (GANCE_BF16 / GANCE_F16 only; the fp32 validation mode uses gance_modconv3x3_up2_fir)
 *   noise [2H,2W] or [B,2H,2W]; out [B,2H,2W,Cout]. */
int gance_modconv3x3_up2_folded(const void* x, int dtype, int batch, int h_in, int w_in, int c_in,
                                int c_out, const void* weight_folded, const float* demod,
                                const float* noise, long long noise_batch_stride,
                                float noise_strength, const float* bias, const float* next_styles,
                                void* out, void* stream);

/* ---- up-convolution at its algorithmic cost: 4-phase transposed-convolution GEMM + FIR pass -------------------------
 * The folded form above executes 36 Cin*Cout MACs per input pixel; upsample_conv_2d needs 9. With
 * t = conv2d_transpose(x, flip(w), stride 2) (upfirdn_2d.py:upsample_conv_2d), phase (a,b) of t at cell (m,n) is
 *   t[2m+a, 2n+b] = sum_{dy,dx in {0,1}} x[m-dy, n-dx] * w[iy(dy,a)][ix(dx,b)],  iy(0,0)=2, iy(0,1)=1, iy(1,0)=0
 * so the input shifted by (dy,dx) feeds 4/2/2/1 phases: one tcgen05 GEMM row per cell with columns
 * [P01|P00|P10|P11] x 64 channels and MMAs of N = 256/128/128/64. Launch 1 writes t*demod as 16-bit NHWC into
 * scratch [B,2H+2,2W+2,Cout]; launch 2 applies the [1,3,3,1] FIR (pad 1,1) + noise + bias + lrelu*sqrt(2) + next_styles.
 *   x [B,H,W,Cin] ALREADY multiplied by this layer's styles; H >= 8, W >= 16, Cin % 64 == 0, Cout % 64 == 0
 *   weight_phase [9*Cout][Cin]: rows ordered shift (0,0),(1,0),(0,1),(1,1) -> 64-channel chunk -> phase block ->
 *   channel (gance_b200/synthesis.py:phase_up_weight), pre-multiplied by 1/sqrt(9*Cin); GANCE_BF16 / GANCE_F16 only. */
int gance_modconv3x3_up2_phase(const void* x, int dtype, int batch, int h_in, int w_in, int c_in,
                               int c_out, const void* weight_phase, const float* demod,
                               const float* noise, long long noise_batch_stride, float noise_strength,
                               const float* bias, const float* next_styles, void* scratch, void* out,
                               void* stream);

/* ---- gance_modconv3x3_up2_phase for callers that cannot allocate device memory ---------------------------------------
 * No counterpart in the reference: added for callers that may not call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Same two launches, same
 * arguments and same results as gance_modconv3x3_up2_phase without `scratch`: t is kept in a file-scope
 * `static __device__` array of GANCE_PHASE_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library holds
 * it until exit); its device address is resolved with cudaGetSymbolAddress on the first call on each device. The entry
 * calls no cudaMalloc / cudaFree / allocating routine.
 *   limit: 2 * B * (2H+2) * (2W+2) * Cout bytes <= GANCE_PHASE_STATIC_SCRATCH_BYTES, else GANCE_E_SHAPE (use the entry above);
 *   every call on a device shares the array: pass 1 overwrites all of the range pass 2 reads, so nothing is cleared and
 *   calls in stream order are independent, but calls that OVERLAP on the GPU (other streams / threads) race on it. */
#define GANCE_PHASE_STATIC_SCRATCH_BYTES (4u << 20)
int gance_modconv3x3_up2_phase_static(const void* x, int dtype, int batch, int h_in, int w_in, int c_in,
                                      int c_out, const void* weight_phase, const float* demod,
                                      const float* noise, long long noise_batch_stride,
                                      float noise_strength, const float* bias, const float* next_styles,
                                      void* out, void* stream);

/* ---- the same up-convolution in ONE launch: 4-phase GEMM whose epilogue applies the FIR on chip -------------------------
 * upconv_fused.cu: tiles of 16 x 8 cells (a 32 x 16-pixel patch of t, kept as 16-bit in shared memory) produce 28 x 12 output
 * pixels each; the 2-cell halo is recomputed by the tensor core (1.52x the algorithmic MACs) and t never reaches HBM.
 * Same arguments as gance_modconv3x3_up2_phase without `scratch`; Cout % 64 == 0 (64-channel accumulator chunks) or
 * Cout == 32 (one 32-channel chunk); weight_phase rows use that chunk size (synthesis.py:phase_up_weight(w, chunk)). */
int gance_modconv3x3_up2_fused(const void* x, int dtype, int batch, int h_in, int w_in, int c_in,
                               int c_out, const void* weight_phase, const float* demod,
                               const float* noise, long long noise_batch_stride, float noise_strength,
                               const float* bias, const float* next_styles, void* out, void* stream);

/* ---- Conv1 + ToRGB + skip (+ uint8) in one kernel ------------------------------------------------
 * gance_modconv3x3 (SAME) whose epilogue also evaluates torgb() on the fp32 activations it has in
 * registers: rgb = sum_co v[co] * rgb_styles[b,co] * rgb_weight[co,:] + rgb_bias, y = upsample_2d(y_prev)
 * + rgb -> y_out [B,H,W,4] f32 and/or convert_images_to_uint8 -> frames [B,H,W,3]. The activation
 * tensor is never re-read (gance_torgb_skip would) and `out` may be NULL on the last block, where nothing
 * consumes it. next_styles (the NEXT up-convolution's styles) is applied to `out` only.
 * GANCE_BF16 / GANCE_F16; Cout in {32,64,128,256} (all channels of a pixel in one accumulator tile). */
int gance_modconv3x3_torgb(const void* x, int dtype, int batch, int h, int w, int c_in, int c_out,
                           const void* weight, const float* demod, const float* noise,
                           long long noise_batch_stride, float noise_strength, const float* bias,
                           const float* next_styles, void* out, const float* rgb_styles,
                           const float* rgb_weight, const float* rgb_bias, const float* y_prev,
                           float* y_out, uint8_t* frames, void* stream);

/* ---- ToRGB + skip accumulation (+ uint8 conversion) ---------------------------------------------
 * torgb(): modulated 1x1 conv without demodulation + bias, y = upsample_2d(y_prev) + t, and on
 * the last block tflib.convert_images_to_uint8(nchw_to_nhwc=True): trunc(clamp(y*127.5+128)).
 *   x [B,H,W,C] unscaled; styles [B,C]; weight [C,3] f32 pre-multiplied by 1/sqrt(C); bias [3]
 *   y_prev [B,H/2,W/2,4] f32 or NULL; y_out [B,H,W,4] f32 or NULL; frames [B,H,W,3] u8 or NULL */
int gance_torgb_skip(const void* x, int dtype, int batch, int h, int w, int channels,
                     const float* styles, const float* weight, const float* bias,
                     const float* y_prev, float* y_out, uint8_t* frames, void* stream);

/* Stand-alone convert_images_to_uint8 on a [B,H,W,4] f32 skip tensor -> [B,H,W,3] u8. */
int gance_to_uint8_nhwc(const float* y, int batch, int h, int w, uint8_t* frames, void* stream);

/* ---- fresh noise inputs ---------------------------------------------------------------------------
 * Gs.run leaves randomize_noise=True (network_functions.py:151-158), so every call draws
 * tf.random_normal([B,1,H,W]) per noise input. out[4g..4g+3] = BoxMuller(Philox4x32-10(counter = g + offset,
 * key = seed)): N(0,1) float32, reproducible from (seed, offset), independent of the launch shape. count % 4 == 0.
 * One launch fills the noise of every layer of a batch (the layers are slices of one buffer). */
int gance_normal_noise(float* out, long long count, unsigned long long seed, unsigned long long offset,
                       void* stream);
/* The same stream with its position kept in device memory: offset = *counter, then *counter += count/4 (a second,
 * one-thread launch). Two calls never share noise even when they are replays of one captured CUDA graph. */
int gance_normal_noise_next(float* out, long long count, unsigned long long seed,
                            unsigned long long* counter, void* stream);

/* ---- audio: WAV samples -> float32 -> resampled so that one frame = vector_length samples --------
 * gance_pcm_to_float: music.read_wav_file (gance/vector_sources/music.py:190-205) +
 * remap_values_into_range (vector_sources_common.py:44-61): PCM of `bits` (16, 32, 8) mapped from
 * its integer range onto [-1,1] in float64, stored float32. Bit-exact with the reference.
 * gance_sinc_resample: resampy 0.2.2 resample(filter='kaiser_best') as called by
 * _scale_wav_to_sample_rate (music.py:212-230). interp_win / interp_delta are the float64 half
 * window (already multiplied by the ratio when down-sampling) and its first difference, n_win
 * long; scale = min(1, ratio); times[t] is the float64 running sum of 1/ratio (host-built by
 * sequential addition). y [n_out] f32. Same operation order and rounding as the numba loop. */
int gance_pcm_to_float(const void* pcm, int bits, long long num_samples, float* out, void* stream);
int gance_sinc_resample(const float* x, int n_in, const double* times, int n_out,
                        const double* interp_win, const double* interp_delta, int n_win,
                        double scale, int num_table, float* y, void* stream);

/* ---- audio: windowed real DFT -> magnitude ------------------------------------------------------
 * apply_spectrogram.compute_spectrogram (apply_spectrogram.py:49-82) up to |X|: frames of
 * window = vector_length-2 samples at hop vector_length, periodic Hann np.hanning(window+1)[:-1],
 * DFT bins [0, window/2), all in float64. audio f32 [T*vector_length] -> mag f64 [T, window/2].
 * global_max is an 8-byte device slot that receives max|X| (needed by the dB conversion). */
int gance_stft_mag(const float* audio, int num_frames, int vector_length, double* mag,
                   double* global_max, void* stream);

/* 20*log10(mag/max) (apply_spectrogram.py:80-81) followed by the Fourier resample bins ->
 * vector_length of reshape_spectrogram_to_vectors (apply_spectrogram.py:20-46,
 * vector_sources_common.py:211-230): scipy.signal.resample is linear, so it is applied as the
 * fixed matrix resample_matrix_t [bins, vector_length] f64 (host-built, transposed for coalescing).
 * out [T, vector_length] f64. minmax is a 16-byte device slot that receives the global min and max
 * of `out` in an order-preserving integer encoding; pass it on to gance_minmax_savgol_time. */
int gance_db_resample(const double* mag, const double* global_max, int num_frames, int bins,
                      int vector_length, const double* resample_matrix_t, double* out,
                      double* minmax, void* stream);

/* Min-max scale to [range_lo, range_hi] fused with the Savitzky-Golay filter along TIME
 * (smooth_across_vectors, window 7 poly 3; vector_sources_common.py:136-166). `coeffs` is the
 * [window, window] f64 matrix whose row r gives output sample r of an edge fit (rows
 * 0..half-1 and half+1..window-1) and whose middle row is the interior FIR. */
int gance_minmax_savgol_time(const double* in, const double* minmax, double range_lo,
                             double range_hi, int num_frames, int vector_length,
                             const double* coeffs, int window, double* out, void* stream);

/* Savitzky-Golay within each vector (smooth_each_vector, vector_sources_common.py:169-188),
 * optionally preceded by the circular left shift of rotate_vectors_over_time
 * (vector_sources_common.py:408-428): in row t is read at (i + shift[t]) mod L. shift may be
 * NULL. coeffs as above. in/out [T, L] f64. */
int gance_roll_savgol_vector(const double* in, const int* shift, int num_frames,
                             int vector_length, const double* coeffs, int window, double* out,
                             void* stream);

/* Alpha blends (visualization_inputs.py:144 and 233-243).
 * noise-blend:      z[t,i]     = noise[t,i]*(1-alpha) + spec[t,i]*alpha           (noise f32)
 * projection-blend: w[t,r,i]   = r < blend_depth ? proj[t/dup,i]*(1-alpha) + spec[t,i]*alpha
 *                                                : proj[t/dup,i]                  (proj f32 [F,L])
 * `a*(1-alpha)` is a float32 product as in the reference (float32 array times a Python float); the
 * sum is float64 without FMA contraction, so results are bit-identical to numpy's.
 * Outputs (either may be NULL): latents_f64 in the reference's layout - (T*L,) resp. (depth, T*L) -
 * and latents_f32 in the generator's layout - [T,L] resp. [T,depth,L]. */
int gance_alpha_blend_noise(const double* spec, const float* noise, double alpha, int num_frames,
                            int vector_length, double* latents_f64, float* latents_f32,
                            void* stream);
int gance_alpha_blend_wplus(const double* spec, const float* projected, int duplication,
                            double alpha, int blend_depth, int depth, int num_frames,
                            int vector_length, double* latents_f64, float* latents_f32,
                            void* stream);

/* ---- a9: Gaussian noise source, wrap-around filter along time ------------------------------------
 * scipy.ndimage.gaussian_filter(latents, sigma=(sigma,0,0), mode="wrap") of gance/vector_sources/primatives.py:69, as
 * scipy's NI_Correlate1D evaluates it (double accumulation, symmetric pairs, outermost tap first, no FMA): bit-exact.
 *   in / out [T, L] f32 (distinct buffers); weights [radius+1] f64 = the normalised kernel from its centre outwards
 *   (gance_b200/audio_device.py:gaussian_kernel1d builds it the way scipy does). The RNG stays on the host (legacy
 *   MT19937 stream of RandomState(1234)), as do the two global reductions that follow (RMS, min-max). */
int gance_gaussian_filter_time_wrap(const float* in, int num_vectors, int vector_length,
                                    const double* weights, int radius, float* out, void* stream);

/* ---- output-side frame scaling: bicubic resize of uint8 NHWC frames ------------------------------------
 * Replaces cv2.resize(image, (side, side), interpolation=cv2.INTER_CUBIC) in
 * gance/image_sources/video_common.py:399-429 (resize_source; used by scale_square_source_duplicate, 432-455, from
 * gance/projection_file_blend.py:184-199). Bit-exact with OpenCV's own 8-bit code path (resize.cpp: 11-bit fixed-point
 * coefficients, integer horizontal pass, float vertical pass for the first floor(n/8)*8 elements of a destination row,
 * integer tail; oracle/resize_ref.py).
 *   src [B,src_h,src_w,3] u8 -> dst [B,dst_h,dst_w,3] u8 (dst 4-byte aligned)
 *   xofs [dst_w] i32 / xcoef [dst_w,4] i16, yofs [dst_h] / ycoef [dst_h,4]: per-axis tap-1 source index and
 *   round(c*2048) coefficients (gance_b200/image_sources/video_common.py:cubic_tables); 8-byte aligned tables.
 *   win_w / win_h: the largest source extent, taps included, any 128-column / 16-row destination tile needs. */
int gance_resize_cubic_u8(const uint8_t* src, int batch, int src_h, int src_w, uint8_t* dst, int dst_h,
                          int dst_w, const int* xofs, const short* xcoef, const int* yofs,
                          const short* ycoef, int win_w, int win_h, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GANCE_B200_H */
