/* ddsp_b200.h -- C ABI of the B200-native DDSP synthesis hot path (libddsp_b200.so).
 *
 * The reference (kureta/ddsp-pytorch) is pure Python and has no FFI layer; its boundary for this
 * path is the torch module API of model/ddsp/{harmonic_oscillator,filtered_noise,reverb}.py and
 * loss/mss_loss.py.  Each entry point below is what a binding for one of those module methods
 * calls (the ctypes binding is ddsp_pytorch_b200/_lib.py), and cites the
 * reference lines it replaces.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to contiguous row-major fp32 unless stated otherwise;
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream); calls only enqueue work,
 *     they never synchronise and never allocate;
 *   - scratch/state buffers are caller-owned; sizes come from the *_bytes() queries;
 *   - return value 0 = success, non-zero = error (ddsp_last_error() returns a message for the
 *     calling thread).  Arguments are validated before anything is launched.
 *   - shapes: B voices, T frames, hop samples per frame, N = T*hop samples, K harmonics,
 *     NB noise bands (FIR length M = 2(NB-1) <= hop), L reverb taps.
 *   - hop: any value the reference accepts, within the kernels' reach -- oscillator: hop % 4 == 0 up to 4096,
 *     every other hop (odd ones included) up to 1024; filtered noise: 2..4096 (fast path when hop % 16 == 0).
 */
#ifndef DDSP_B200_H
#define DDSP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDSP_B200_ABI_VERSION 1

int         ddsp_abi_version(void);
const char* ddsp_last_error(void);
/* Number of kernels this library has launched (or recorded into a capturing stream) in this process: the bench's
 * `gpu_launches` claim is counted from it, not assumed.  No reference counterpart. */
unsigned long long ddsp_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * Harmonic oscillator bank  --  OscillatorBank.forward, harmonic_oscillator.py:57-62
 *   (prepare_harmonics :24-37, rescale :52-55, generate_phases :39-43, generate_signal :45-50)
 * ------------------------------------------------------------------------------------------ */

/* Bytes of the frame-rate state ddsp_osc_forward fills and ddsp_osc_backward reads:
 * normalised masked amplitudes [B,T,K], their pre-normalisation sums [B,T], active-harmonic
 * counts [B,T] and the fp64 phase prefix (cycles) at each segment start [B,T+1]. */
size_t ddsp_osc_state_bytes(int B, int T, int K);

/* y[B,N] (+)= sum_k a_up * c'_up * sin(phi_k).
 *   f0 [B,T] Hz, c [B,T,K], a [B,T];  z_save [B,N] or NULL: harmonic sum before the loudness
 *   multiply, needed by ddsp_osc_backward;  accumulate != 0 adds into y instead of overwriting. */
int ddsp_osc_forward(const float* f0, const float* c, const float* a, float* y, float* z_save,
                     void* state, int accumulate, int B, int T, int K, int hop, int sample_rate,
                     void* stream);

/* The two launches of ddsp_osc_forward on their own (frame-rate preparation; sample-rate synthesis). */
int ddsp_osc_prepare(const float* f0, const float* c, void* state, int B, int T, int K, int hop,
                     int sample_rate, void* stream);
int ddsp_osc_synthesize(const float* f0, const float* a, const void* state, float* y, float* z_save,
                        int accumulate, int B, int T, int K, int hop, int sample_rate, void* stream);

/* Scratch for ddsp_osc_backward: per-segment partial gradients [B,T+1,K,2] + [B,T+1,2]. */
size_t ddsp_osc_backward_scratch_bytes(int B, int T, int K);

/* Gradients of sum(g*y) w.r.t. c -> gc [B,T,K] and a -> ga [B,T]  (autograd of :24-50; f0 carries
 * no gradient on the training path).  state/z_save are those of the forward call. */
int ddsp_osc_backward(const float* g, const float* f0, const float* a, const float* z_save,
                      const void* state, void* scratch, float* gc, float* ga, int B, int T, int K,
                      int hop, int sample_rate, void* stream);

/* Gradient of sum(g*y) w.r.t. f0 -> gf0 [B,T] (autograd of :26-43 in the reference: k*f0 -> rad/sample -> upsample ->
 * cumsum -> sin; the Nyquist comparison carries no gradient there either).  Separate from ddsp_osc_backward because the
 * training path treats f0 as data; costs about one more synthesis pass.  state: that of the forward call;
 * scratch: ddsp_osc_backward_f0_scratch_bytes(B, T, hop) bytes (one float per sample). */
size_t ddsp_osc_backward_f0_scratch_bytes(int B, int T, int hop);
int ddsp_osc_backward_f0(const float* g, const float* f0, const float* a, const void* state, void* scratch,
                         float* gf0, int B, int T, int K, int hop, int sample_rate, void* stream);

/* The two launches of ddsp_osc_backward on their own (per-segment sums over samples; frame-rate gather
 * + backward of mask/normalise). */
int ddsp_osc_grad_segments(const float* g, const float* f0, const float* a, const float* z_save,
                           const void* state, void* scratch, int B, int T, int K, int hop,
                           int sample_rate, void* stream);
int ddsp_osc_grad_finalize(const float* f0, const void* state, const void* scratch, float* gc,
                           float* ga, int B, int T, int K, int hop, int sample_rate, void* stream);

/* Allocation-free forward.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * ddsp_osc_forward(..., z_save = NULL, ...), bit for bit, but the frame-rate state lives in a static device array inside
 * the library (DDSP_OSC_NOALLOC_STATE_BYTES = 8 MiB, held by every GPU that loads the library until exit) instead of a
 * caller-owned buffer; the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 *   limit: ddsp_osc_state_bytes(B, T, K) <= DDSP_OSC_NOALLOC_STATE_BYTES (e.g. B=16, T=125, K=100, or B=1 real-time
 *     blocks of any shipped size); larger shapes are rejected with an error message, nothing is launched;
 *   every call on a device shares the one array: calls that may OVERLAP on the GPU (different streams, or graph replays
 *     launched concurrently) race -- serialise them; the array keeps the last call's contents and nothing in it has to
 *     start at zero (the preparation kernel rewrites all of it).  Forward only (no z_save: no backward from this call). */
#define DDSP_OSC_NOALLOC_STATE_BYTES (8u << 20)
int ddsp_osc_forward_noalloc(const float* f0, const float* c, const float* a, float* y, int accumulate,
                             int B, int T, int K, int hop, int sample_rate, void* stream);

/* No reference counterpart.  The harmonic loops of the two dominant kernels (synthesis with 8 samples per thread, gradient
 * with 16: every hop that is a multiple of 32, the shipped 512 included) also exist in a hand-scheduled form: the same
 * instructions in an order that feeds the FMA pipe from the operand reuse cache (ddsp_pytorch_b200/sass_sched.py; results are
 * bit-identical).  The image is loaded through the driver API on the first oscillator call of a device -- make that call
 * outside stream capture -- and used from then on.  mode 1 / 0 switches it on (default) / off for the calling process,
 * any other value only queries; returns bit 0 = synthesis, bit 1 = gradient, bit 2 = d/d f0 synthesis-pass kernel tuned and in
 * use on the current device. */
int ddsp_osc_tuned_sass(int mode);

/* OscillatorBank.live, harmonic_oscillator.py:64-75: as forward, but voice 0 starts harmonic k from the
 * carried phase last_in[k] (radians; arbitrary per-harmonic values are honoured) and last_out[k] receives
 * the wrapped phase of voice 0's last sample (:72).  last_in and last_out must not alias. */
int ddsp_osc_live(const float* f0, const float* c, const float* a, const float* last_in,
                  float* last_out, float* y, void* state, int B, int T, int K, int hop,
                  int sample_rate, void* stream);

/* OscillatorBank.rescale, harmonic_oscillator.py:52-55 (F.interpolate(..., scale_factor=hop, mode='linear'), half-pixel
 * centres): x [B,T,C] -> y [B,T*hop,C].  The synthesis entry points above never build this tensor; it backs the
 * reference's public helper methods (prepare_harmonics :24-37, generate_signal :45-50), which the drop-in keeps.
 * backward: gx [B,T,C] = adjoint applied to gy [B,T*hop,C] (fixed summation order). */
int ddsp_rescale_forward(const float* x, float* y, int B, int T, int C, int hop, void* stream);
int ddsp_rescale_backward(const float* gy, float* gx, int B, int T, int C, int hop, void* stream);

/* ------------------------------------------------------------------------------------------
 * Filtered noise  --  FilteredNoise.forward, filtered_noise.py:40-53
 *   (amp_to_impulse_response :7-22, fft_convolve :25-32; per frame, tails dropped, no overlap-add)
 * ------------------------------------------------------------------------------------------ */

/* y[B,N] (+)= per-frame causal convolution of uniform noise with the windowed zero-phase FIR of H.
 *   H [B,T,NB].  noise [B,T,hop] in [-1,1) or NULL: NULL draws the noise in-kernel with
 *   Philox4x32-10, key = seed, counter = (j>>2, t, voice_offset + b, offset), word j&3,
 *   u = (word>>8)*2^-24, noise = 2u-1  (oracle/philox.py is the CPU statement of this layout).
 *   offset_dev: NULL, or a DEVICE uint32 added to `offset` when the kernel runs -- a call counter that lives
 *   on the device lets a captured CUDA graph draw fresh noise on every replay. */
int ddsp_noise_forward(const float* H, const float* noise, float* y, int accumulate, uint64_t seed,
                       uint32_t offset, const uint32_t* offset_dev, uint32_t voice_offset, int B, int T,
                       int NB, int hop, void* stream);

/* gH [B,T,NB] = d sum(g*y) / dH with the same noise (regenerated from the same seed/offset). */
int ddsp_noise_backward(const float* g, const float* noise, float* gH, uint64_t seed,
                        uint32_t offset, const uint32_t* offset_dev, uint32_t voice_offset, int B, int T,
                        int NB, int hop, void* stream);

/* The noise block itself [B,T,hop] (for tests of the Philox layout and for injecting it elsewhere). */
int ddsp_noise_uniform(float* out, uint64_t seed, uint32_t offset, uint32_t voice_offset, int B,
                       int T, int hop, void* stream);

/* amp_to_impulse_response, filtered_noise.py:7-22: H [R,NB] -> ir [R,target]. */
int ddsp_amp_to_impulse_response(const float* H, float* ir, int R, int NB, int target, void* stream);

/* ------------------------------------------------------------------------------------------
 * Reverb  --  reverb.py
 * ------------------------------------------------------------------------------------------ */

/* Reverb.build_impulse, reverb.py:24-29: ir[L] = noise*exp(-softplus(-decay)*t*500)*sigmoid(wet), ir[0]=1.
 * noise [L], t [L], decay/wet: device scalars. */
int ddsp_reverb_impulse(const float* noise, const float* t, const float* decay, const float* wet,
                        float* ir, int L, void* stream);

/* Backward of build_impulse: gir [L] -> gnoise [L], gdecay [1], gwet [1]. */
int ddsp_reverb_impulse_backward(const float* gir, const float* noise, const float* t,
                                 const float* decay, const float* wet, float* gnoise, float* gdecay,
                                 float* gwet, int L, void* stream);

/* fft_convolve with a shared kernel, filtered_noise.py:25-32 as used by Reverb.forward reverb.py:31-38:
 * y[b,n] = sum_{m<=n, m<Lk} ir[m] * x[b,n-m], n < N   (ir is cropped to N when Lk > N).
 * scratch: ddsp_conv_scratch_bytes(B, N, Lk). */
size_t ddsp_conv_scratch_bytes(int B, int N, int Lk);
int ddsp_conv_forward(const float* x, const float* ir, float* y, void* scratch, int B, int N, int Lk,
                      void* stream);
/* gx[b,m] = sum_{n>=m} g[b,n]*ir[n-m];  gir[q] = sum_b sum_{n>=q} g[b,n]*x[b,n-q]. Either output may be NULL.
 * scratch_from_forward != 0: `scratch` is the buffer the matching ddsp_conv_forward call (same x, ir, B, N, Lk)
 * used and nothing has written to it since -- the block spectra of x and ir in it are reused instead of being
 * recomputed.  The call overwrites them, so a second backward on the same buffer must pass 0. */
int ddsp_conv_backward(const float* g, const float* x, const float* ir, float* gx, float* gir,
                       void* scratch, int B, int N, int Lk, int scratch_from_forward, void* stream);

/* fft_convolve with one kernel per row, filtered_noise.py:25-32 in its general call shape (signal [...,n],
 * kernel [...,n], flattened to `rows` rows): y[r,m] = sum_{q<=m} kernel[r,q] * signal[r,m-q].  n <= 4096.
 * backward: gsignal[r,j] = sum_{m>=j} g[r,m]*kernel[r,m-j], gkernel[r,q] = sum_{m>=q} g[r,m]*signal[r,m-q];
 * either output may be NULL. */
int ddsp_rowconv_forward(const float* signal, const float* kernel, float* y, long long rows, int n,
                         void* stream);
int ddsp_rowconv_backward(const float* g, const float* signal, const float* kernel, float* gsignal,
                          float* gkernel, long long rows, int n, void* stream);

/* Reverb.live_forward, reverb.py:40-49: buffer[L] <- shift left by n, append x[n]; y[n] = last n
 * samples of causal_conv(buffer, ir). */
int ddsp_reverb_live(float* buffer, const float* x, const float* ir, float* y, void* scratch, int L,
                     int n, void* stream);   /* scratch: >= L floats */

/* ------------------------------------------------------------------------------------------
 * Multi-scale spectral loss  --  loss/mss_loss.py:11-68 (torchaudio Spectrogram defaults: power
 * spectrogram, periodic Hann, center reflect pad, onesided; hop = int(n_fft*(1-overlap))).
 * ------------------------------------------------------------------------------------------ */

/* Elements of one scale's spectrogram: B * (n_fft/2+1) * (1 + N/hop)  (the `mean` denominators of :29-30). */
int64_t ddsp_mss_count(int B, int N, int n_fft, int hop);

/* Bytes of the caller-owned scratch ddsp_mss_forward fills and ddsp_mss_backward reads: per scale the windowed
 * gradient frames already overlap-added by their owners (B*N floats plus margins), the spills that crossed a work-item
 * boundary, and the per-CTA partial loss sums.  Depends on the arguments and the current device only. */
size_t ddsp_mss_scratch_bytes(const int* n_ffts_host, const int* hops_host, int n_scales, int B, int N);

/* MSSLoss.forward, mss_loss.py:57-68.  x_pred, x_true [B,N].
 *   loss_terms (device, double[1 + 2*n_scales]): [0] the loss
 *     sum_s ( sum|Sp-St| + alpha*sum|log2(St+eps)-log2(Sp+eps)| ) / count_s, then per scale the two raw sums;
 *   loss_f32 (device float[1]) or NULL: the loss rounded to fp32;
 *   scratch: ddsp_mss_scratch_bytes(...) bytes; need_grad != 0 also leaves d loss / d x_pred in it (per scale, owner-
 *     computes overlap-add -- no floating-point atomics: loss and gradient are bit-reproducible and a voice's gradient
 *     does not depend on which batch it is evaluated in, up to the exact ratio of the two means);
 *   n_ffts_host / hops_host: HOST int arrays; n_fft a power of two in [4,4096] (the reference accepts any n_fft:
 *     INTEGRATION.md, behavioural differences), n_fft/2 < N, at most 16 scales; hop = n_fft/4 (the reference's
 *     default overlap 0.75) runs the register-tiled kernels, any other hop a generic shared-memory kernel;
 *   windows_host: NULL, or a HOST array of n_scales device pointers (entries may be NULL) to the fp32 analysis
 *     windows [n_fft] -- the reference's `stft.window` buffers (mss_loss.py:23); NULL = periodic Hann rounded
 *     from fp64, which differs from torch.hann_window's fp32 table by an ulp in places: pass the table when the
 *     gradient has to match the reference bin for bin on near-silent input. */
int ddsp_mss_forward(const float* x_pred, const float* x_true, double* loss_terms, float* loss_f32,
                     void* scratch, int need_grad, const int* n_ffts_host, const int* hops_host,
                     const float* const* windows_host, int n_scales, float alpha, float eps, int B, int N,
                     void* stream);

/* Backward of the scalar loss: gx_out [B,N] = gloss[0] * d loss / d x_pred, gathered from the forward's scratch
 * (scales in order, then the reflect-padded margins folded back; gloss: device float scalar). */
int ddsp_mss_backward(const float* gloss, const void* scratch, float* gx_out, const int* n_ffts_host,
                      const int* hops_host, int n_scales, int B, int N, void* stream);

/* ------------------------------------------------------------------------------------------
 * DFT-as-GEMM variant of ONE scale of the spectral loss on the tcgen05 tensor cores (north_star item 3: "only for the
 * small FFT sizes where ncu shows it beats a shared-memory radix FFT").  Forward of SSSLoss, mss_loss.py:25-33, for
 * n_fft in {64, 128}, hop = n_fft/4: frames x windowed real-DFT matrix as a bf16 x 3 split GEMM (TMA-fed operands,
 * accumulators in TMEM), power spectra and both L1 sums in the epilogue.  Measured against the register-tiled FFT path of
 * ddsp_mss_forward in profiles/ (DESIGN.md 4.3); the product path (ddsp_mss_forward) does not call it.
 *   prepare: reflect pad + split of both signals into the workspace's block-row planes, DFT matrix from `window`
 *            (device fp32 [n_fft] or NULL = periodic Hann);
 *   forward: terms2 (device double[2]) = sum|Sp-St|, sum|log2(St+eps)-log2(Sp+eps)|  (the two raw sums of loss_terms);
 *            S_out: NULL, or device float [2][B][F][n_fft/2+1] (true, pred power spectrograms, frame-major) for accuracy
 *            checks; splits: 3 (six bf16 products, fp32-class accuracy) or 1 (one product, bf16 accuracy).
 * ------------------------------------------------------------------------------------------ */
size_t ddsp_dftgemm_workspace_bytes(int B, int N, int n_fft);
int ddsp_dftgemm_prepare(const float* x_pred, const float* x_true, const float* window, void* workspace,
                         int n_fft, int B, int N, void* stream);
int ddsp_dftgemm_sss_forward(void* workspace, double* terms2, float* S_out, int n_fft, int splits, float eps,
                             int B, int N, void* stream);

/* ------------------------------------------------------------------------------------------
 * Loudness feature  --  LoudnessEncoder.forward, model/autoencoder/encoder.py:141-156
 *   out[B,F] = mean_k( (20 log10(|STFT(x)[k]| + 1e-20) + a_weight[k]) / 90 + 1 ),  rectangular window,
 *   center=False, F = ddsp_loudness_frames(N, n_fft, hop) = 1 + (N - n_fft)/hop, a_weight [n_fft/2+1] in dB.
 *   n_fft a power of two in [8,4096].  Forward only (the encoder runs under no_grad, encoder.py:51).
 * ------------------------------------------------------------------------------------------ */
int ddsp_loudness_frames(int N, int n_fft, int hop);
int ddsp_loudness_forward(const float* x, const float* a_weight, float* out, int B, int N, int n_fft,
                          int hop, void* stream);

/* ------------------------------------------------------------------------------------------
 * Controller tail  --  Controller.forward, model/autoencoder/decoder.py:96-100 with the heads of
 * :74-76 and modified_sigmoid :110-116:
 *   c[R,K] = msig(x W_c^T + b_c), a[R,1] = msig(x W_a^T + b_a), H[R,NB] = msig(x W_h^T + b_h),
 *   msig(z) = 2 sigmoid(z)^2.3026 + 1e-7, x [R,U] (R = B*T rows of the last MLP's output), nn.Linear
 *   weight layout [out,in].  One launch; U % 4 == 0; matrices 16-byte aligned.
 *   z (optional) [R, ddsp_ctrl_heads_ldz(K,NB)] receives the pre-activations for the backward call.
 * backward: from g_c, g_a, g_h (any may be NULL = zero) to gx [R,U] (NULL to skip) and the six
 *   parameter gradients (overwritten, not accumulated); deterministic (fixed summation order).
 *   workspace: ddsp_ctrl_heads_workspace_bytes(R,U,K,NB) bytes, 16-byte aligned.
 * ------------------------------------------------------------------------------------------ */
int ddsp_ctrl_heads_ldz(int K, int NB);
size_t ddsp_ctrl_heads_workspace_bytes(int R, int U, int K, int NB);
int ddsp_ctrl_heads_forward(const float* x, const float* w_c, const float* b_c, const float* w_a,
                            const float* b_a, const float* w_h, const float* b_h, float* c, float* a,
                            float* H, float* z, int R, int U, int K, int NB, void* stream);
int ddsp_ctrl_heads_backward(const float* x, const float* w_c, const float* w_a, const float* w_h,
                             const float* z, const float* g_c, const float* g_a, const float* g_h,
                             float* gx, float* gw_c, float* gb_c, float* gw_a, float* gb_a, float* gw_h,
                             float* gb_h, void* workspace, int R, int U, int K, int NB, void* stream);
/* The same two calls on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM, operands by TMA): both
 * operands as bf16 x 3 splits, six products in fp32 -- an fp32-accurate product, deterministic.  For R of a few hundred
 * rows and more (below that the launches of the split pre-passes cost more than the product); U % 8 == 0.
 *   workspace: ddsp_ctrl_heads_tc_workspace_bytes(R,U,K,NB, backward != 0) bytes, 16-byte aligned (0: bad arguments).
 *   z as above; the tensor-core backward takes the z either forward wrote. */
size_t ddsp_ctrl_heads_tc_workspace_bytes(int R, int U, int K, int NB, int backward);
int ddsp_ctrl_heads_forward_tc(const float* x, const float* w_c, const float* b_c, const float* w_a,
                               const float* b_a, const float* w_h, const float* b_h, float* c, float* a,
                               float* H, float* z, void* workspace, int R, int U, int K, int NB,
                               void* stream);
int ddsp_ctrl_heads_backward_tc(const float* x, const float* w_c, const float* w_a, const float* w_h,
                                const float* z, const float* g_c, const float* g_a, const float* g_h,
                                float* gx, float* gw_c, float* gb_c, float* gw_a, float* gb_a,
                                float* gw_h, float* gb_h, void* workspace, int R, int U, int K, int NB,
                                void* stream);

/* ------------------------------------------------------------------------------------------
 * The transform underneath the reverb and the spectral loss, on its own: `batch` complex FFTs of
 * size n (power of two in [8,4096]), interleaved (re,im) fp32 (fp64 = 0) or fp64 (fp64 != 0), in
 * natural order; inverse != 0 is the unnormalised inverse.  Replaces the cuFFT calls the reference
 * reaches through torch.fft.rfft/irfft (filtered_noise.py:10,29) and torch.stft (mss_loss.py:23-27).
 * ------------------------------------------------------------------------------------------ */
int ddsp_fft_c2c(const void* in, void* out, int n, long long batch, int inverse, int fp64, void* stream);

/* ------------------------------------------------------------------------------------------
 * Measurement helper: dependent-free FFMA chains, for the measured FP32 roofline denominator.
 * Each thread executes iters*64 FFMAs; out[grid*block] keeps the compiler honest.
 * ------------------------------------------------------------------------------------------ */
int ddsp_probe_fp32_fma(float* out, int blocks, int threads, int iters, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DDSP_B200_H */
