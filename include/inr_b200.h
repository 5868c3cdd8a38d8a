/* inr_b200.h -- C-ABI of the B200-native INR-fitting hot path (libinr_b200.so).
 *
 * The reference (senyuanfan/inr-for-audio) has no FFI: its seam is Python classes.
 * Each entry point below replaces the reference code cited beside it; the Python host layer in
 * inr_for_audio_b200/ binds them with ctypes (INTEGRATION.md shows the stub a reference maintainer adds).
 *
 * Conventions
 *   - plain C: borrowed raw DEVICE pointers + sizes + a cudaStream_t passed as void*; no torch types.
 *   - every function returns 0 on success, <0 on error (INR_E_*); inr_last_error() gives the message.
 *   - the caller owns all buffers and keeps them alive until the stream work completes; all work is
 *     enqueued on `stream` and is asynchronous; functions are re-entrant per stream.
 *   - tensors are fp32, row-major, unless stated.  "flat params" is the concatenation of the model's
 *     parameters in state_dict order (net.0.linear.weight, net.0.linear.bias, ..., models.py:248-269).
 *   - there is NO CPU fallback: without a CUDA device every compute call returns INR_E_CUDA.
 */
#ifndef INR_B200_H
#define INR_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define INR_OK 0
#define INR_E_BADARG (-1)
#define INR_E_UNSUPPORTED (-2)
#define INR_E_CUDA (-3)
#define INR_E_WORKSPACE (-4)

/* arithmetic of the hidden HxH contractions */
#define INR_PREC_FP32 0 /* CUDA-core fp32 FMA everywhere, like the reference: matches it to ~1e-6 per step  */
#define INR_PREC_BF16 1 /* tcgen05 kind::f16, bf16 operands, fp32 TMEM accumulation: the KAN tensor-core mode.
                           SIREN descriptors are refused: 8-bit significands floor a SIREN fit at ~24 dB      */
#define INR_PREC_FP16 2 /* tcgen05 kind::f16, IEEE fp16 operands (11-bit significands) with a power-of-two loss
                           scale on dZ, fp32 TMEM accumulation; first layer, sin/cos arguments, loss, reductions
                           and Adam master weights in fp32: the fast SIREN tensor-core mode (fit floor ~31 dB
                           on gt_bach: profiles/r02_c1_noise_*.jsonl)                                          */
#define INR_PREC_FP16S 3 /* INR_PREC_FP16 with the FORWARD operands S_l and W_l as fp16 hi + lo pairs (22-bit
                           significands) and three MMAs per K step (hi*hi + lo*hi + hi*lo): fp32-grade forward,
                           fp16 backward -- the parity-grade tensor-core mode                                  */

/* time-domain loss, run.py:124-125,161-169 */
#define INR_LOSS_MSE 0
#define INR_LOSS_MAE 1

/* Shape of the MLP: models.py:244-269 Siren(in,H,L,out,outermost_linear,first_omega_0,hidden_omega_0)
 * == models.py:310-394 SirenWithSnakeTanh(num_sine=L, num_snake=0, num_tanh=0), and the general
 * SirenWithSnakeTanh stack (models.py:337-385), in this order:
 *   layer 0            SineLayer(in->H, is_first)           or, first_linear: nn.Linear(in->H) + Snake(H)
 *   hidden_layers  x   SineLayer(H->H)                      y = sin(omega z)
 *   num_snake      x   nn.Linear(H->H) + Snake(H)           y = z + sin^2(a z)/a, `a` (H) TRAINABLE (models.py:185-241)
 *   num_tanh       x   nn.Linear(H->H) + nn.Tanh()
 *   last               nn.Linear(H->1) (outermost_linear)   or SineLayer(H->1)
 * flat params = state_dict order: W, b per layer, with a Snake layer's `a` (H floats) right after its bias.
 * INR_PREC_FP32, INR_PREC_FP16 and INR_PREC_FP16S take every stack; the two tcgen05 modes need H in {128,256,512}, at
 * least one HxH layer and in_features <= H.  INR_PREC_BF16 is refused for a SIREN descriptor. */
typedef struct {
  int32_t in_features;     /* 1 (waveform), 2 (MDCT image), or the width of an encoded
                              coordinate (2*num_freq, run.py:81-82; fp16 mode: <= hidden_features) */
  int32_t hidden_features; /* H                                                   */
  int32_t hidden_layers;   /* number of HxH SineLayers (num_sine)                 */
  int32_t out_features;    /* must be 1                                           */
  int32_t outermost_linear;/* 1: final nn.Linear; 0: final SineLayer              */
  float first_omega_0;
  float hidden_omega_0;
  int32_t precision;       /* INR_PREC_*                                          */
  int32_t num_snake;       /* number of HxH Linear+Snake layers after the sine layers */
  int32_t num_tanh;        /* number of HxH Linear+Tanh layers after the snake layers  */
  int32_t first_linear;    /* 1: layer 0 is nn.Linear + Snake instead of a SineLayer   */
} inr_siren_desc;

/* Optimiser + scheduler state, resident in DEVICE memory (no per-step host sync; the reference syncs
 * 3-4x per step, run.py:171-187).  Initialise with inr_opt_init. */
typedef struct {
  float lr;            /* current learning rate                                    */
  float min_lr;        /* ReduceLROnPlateau min_lr                                 */
  float factor;        /* 0.8                                                      */
  float threshold;     /* 1e-4 (rel)                                               */
  float sched_eps;     /* 1e-8                                                     */
  float beta1, beta2, adam_eps;
  int32_t patience;    /* 200                                                      */
  int32_t num_bad;     /* plateau counter                                          */
  int32_t step;        /* number of optimiser steps taken (Adam's t)               */
  int32_t hist_cap;    /* capacity of the loss-history ring (floats)               */
  float best;          /* best loss seen by the scheduler                          */
  float last_loss;     /* loss of the most recent step                             */
  float step_size;     /* lr/(1-beta1^t) for the NEXT step, computed in float64    */
  float bc2_sqrt;      /* sqrt(1-beta2^t) for the NEXT step, computed in float64   */
  float gmax_acc;      /* max |d loss/d out| seen this step (float bits, atomicMax) */
  float gmax_prev;     /* ... of the previous step: sets the fp16 loss scale         */
  float omb1, omb2;    /* (float)(1.0 - beta1), (float)(1.0 - beta2) formed in float64, as torch passes them */
  int32_t hist_base;   /* Adam step at which loss_hist[0] was written (resume)      */
  int32_t reserved0;   /* ticket counter of k_adam (the last block out runs the scheduler); zero between launches */
  double loss_acc[4];  /* [0] time-loss numerator sum, [1] spectral loss, [2..3] scratch; zeroed per step */
} inr_opt_state;

const char* inr_last_error(void);
/* ABI version of the library that was loaded; a binding compares it with the INR_ABI_VERSION it was written against and
 * refuses a mismatch (struct layouts -- inr_opt_state grew in 101 -- and entry points change between versions).
 *   100  round 1      101  inr_opt_state + gmax/omb/hist_base fields, INR_PREC_FP16S, SIREN bf16 refused
 *   102  inr_siren_forward_static */
#define INR_ABI_VERSION 102
int inr_version(void);

/* Launch accounting: number of kernels this library has launched in this process (bench.py gpu_launches),
 * and an event-based per-kernel timer: between inr_profile_begin and inr_profile_end every launch is bracketed
 * by CUDA events on its stream; inr_profile_end synchronises the device and writes "tag,launches,total_ms\n"
 * lines into buf (returns the full length).  Not usable during CUDA-graph capture. */
int64_t inr_launch_count(void);

/* Grid-size limit for the persistent kernels: launches enqueued after this call occupy at most n_sms SMs (rounded down
 * to an even number, at least 8; 0 = all SMs).  Used by the clip-parallel batch fitter (BASELINE.json config 5: many independent
 * clips, engine.ClipBatchFitter) so that the per-clip steps it enqueues on several streams run concurrently on disjoint
 * SMs.  Scoped to the calling host thread (thread-local): it applies to the launches that thread enqueues until it calls
 * inr_set_sm_limit(0); returns the limit in effect (0 = none). */
int inr_set_sm_limit(int n_sms);
int inr_profile_begin(void);
int64_t inr_profile_end(char* buf, int64_t cap);

/* number of floats in the flat parameter vector of `d` */
int64_t inr_siren_num_params(const inr_siren_desc* d);

/* bytes of scratch inr_siren_forward / inr_siren_loss_backward need for n rows (activations live here;
 * rows are processed in chunks so the result is bounded) */
int64_t inr_siren_workspace_bytes(const inr_siren_desc* d, int64_t n_rows, int training);

/* utils.py:99-109 get_coord(sidelen, dim=1, scale): out[j] = linspace(-scale, scale, T_total)[i0 + j],
 * j < n, BIT-EXACT with torch.linspace on CPU. */
int inr_coords_linspace(float* out, int64_t i0, int64_t n, int64_t T_total, float scale, void* stream);

/* utils.py:382-403 MDCTFitting 2-D grid: out[(r*W + c), :] = (lin(H)[r], lin(W)[c]) for flat index
 * i0 <= r*W+c < i0+n. */
int inr_coords_grid2d(float* out, int64_t i0, int64_t n, int64_t height, int64_t width, void* stream);

/* Input encodings applied to the coordinates once, in front of the first layer (the network then has
 * in_features = the encoded width).  coords == NULL (in_features == 1 only): row j encodes
 * linspace(-scale,scale,T_total)[i0 + j] generated in the kernel; otherwise coords is (n, in_features).
 *
 * run.py:141-144 `rff.layers.GaussianEncoding(10.0, 1, num_freq)` (third-party `rff`, un-vendored and absent here:
 * PARITY UNPINNED, restated from its published definition): vp = (2*pi*v) @ b^T with b (encoded_size, in_features)
 * drawn once as sigma * randn by the caller; out (n, 2*encoded_size) = [cos(vp) | sin(vp)]. */
int inr_encode_gaussian(const float* coords, int in_features, int64_t n, int64_t i0, int64_t T_total, float scale,
                        const float* b, int encoded_size, float* out, void* stream);
/* models.py:14-59 PosEncodingNeRF.forward: out (n, in + 2*in*num_frequencies) =
 * [x, sin(s^0 pi x_0), cos(s^0 pi x_0), sin(s^0 pi x_1), ..., sin(s^1 pi x_0), ...], s = freq_scale,
 * num_frequencies <= 32. */
int inr_encode_nerf(const float* coords, int in_features, int64_t n, int64_t i0, int64_t T_total, float scale,
                    int num_frequencies, float freq_scale, float* out, void* stream);

/* models.py:271-275 / 388-394  Siren.forward:  out[n, 1] = net(coords).
 * coords == NULL (in_features == 1 only): row j uses linspace(-scale,scale,T_total)[i0 + j] generated in
 * the kernel (0 HBM bytes).  Otherwise coords is (n, in_features). */
int inr_siren_forward(const inr_siren_desc* d, const float* params, const float* coords, int64_t n,
                      int64_t i0, int64_t T_total, float coord_scale, float* out, void* workspace,
                      int64_t workspace_bytes, void* stream);

/* No counterpart in the reference (its tensors come from PyTorch's allocator): inr_siren_forward for callers that
 * cannot allocate device memory -- a real-time audio render callback, where cudaMalloc/cudaFree (driver-wide locks,
 * possible device synchronisation) are as forbidden as malloc; a process whose HBM is already owned by another pool so
 * a late cudaMalloc fails; a plain-C or scripting host with no CUDA runtime binding of its own.  Same arguments and
 * result as inr_siren_forward minus the workspace: the activations go through ONE array of INR_STATIC_SCRATCH_BYTES
 * built into the library (every GPU that loads the library holds it until the process exits).
 *   - the whole call must fit: inr_siren_workspace_bytes(d, n, 0) <= INR_STATIC_SCRATCH_BYTES, else INR_E_WORKSPACE
 *     (about 16 000 rows at H = 256, 8 000 at H = 512 in INR_PREC_FP32); split the rows yourself beyond that;
 *   - every call on a device shares the array, which keeps the previous call's bytes (the forward overwrites all it
 *     reads): calls whose GPU work may overlap -- two streams, two threads -- race; keep them in one stream order;
 *   - the first call on each device looks the array's device address up and loads the module there: make it a warm-up
 *     call, outside the latency-critical part and before any stream capture is begun;
 *   - calls no cudaMalloc/cudaFree or allocating library routine (unless the inr_profile_* event profiler is on). */
#define INR_STATIC_SCRATCH_BYTES (32 * 1024 * 1024)
int inr_siren_forward_static(const inr_siren_desc* d, const float* params, const float* coords, int64_t n, int64_t i0,
                             int64_t T_total, float coord_scale, float* out, void* stream);

/* run.py:158-169 + 185: forward, loss = weight * reduce(out, target), and backward to the flat gradient.
 *   grads (P floats)  += d loss / d params      (ACCUMULATES: zero it, or let inr_adam_step re-zero it)
 *   opt->loss_acc[0]  += sum_j (out-target)^2 or |out-target|   (caller divides by count)
 * grad_scale = weight / count  ( (1-alpha)/T_total for the mean losses of run.py:163-169 ).
 * out_or_null receives the forward output when non-NULL.
 * extra_gout_or_null: optional (n) upstream gradient added to d loss/d out (spectral losses). */
int inr_siren_loss_backward(const inr_siren_desc* d, const float* params, const float* coords,
                            const float* target, int64_t n, int64_t i0, int64_t T_total, float coord_scale,
                            int loss_kind, float grad_scale, const float* extra_gout_or_null,
                            float* grads, inr_opt_state* opt, float* out_or_null, void* workspace,
                            int64_t workspace_bytes, void* stream);

/* The same step split in two, so that a loss that needs the WHOLE model output (the windowed-transform losses
 * below; auraloss' SNR) can be evaluated between the halves without a second forward pass:
 *   inr_siren_train_forward   runs the network keeping every layer's activations in `workspace`, writes out[n];
 *   inr_siren_train_backward  resumes from the SAME, untouched workspace: time-domain loss against `target`
 *                             (may be NULL) plus the upstream gradient extra_gout (n floats, may be NULL).
 * Both need the n rows to fit ONE workspace chunk (INR_E_WORKSPACE otherwise). */
int inr_siren_train_forward(const inr_siren_desc* d, const float* params, const float* coords, int64_t n, int64_t i0,
                            int64_t T_total, float coord_scale, float* out, void* workspace, int64_t workspace_bytes,
                            void* stream);
int inr_siren_train_backward(const inr_siren_desc* d, const float* params, const float* coords, const float* target,
                             int64_t n, int64_t i0, int64_t T_total, float coord_scale, int loss_kind, float grad_scale,
                             const float* extra_gout_or_null, float* grads, inr_opt_state* opt, void* workspace,
                             int64_t workspace_bytes, void* stream);

/* ---- windowed-transform losses on the model output (spectral.cu) ----------------------------------------------
 * All of them ACCUMULATE d loss / d pred into gout (T floats: the caller zeroes it once per step) and add
 * weight * loss to opt->loss_acc[1]; `weight` is alpha of the mixing rule run.py:163-169 (times 1/count for means).
 *
 * mdct.py:51-72 STMDCT(data, N) on the device: frames of N samples at hop N/2 over the tail-zero-padded signal
 * (num_frames = T / (N/2), mdct.py:54-58), window[n] (N floats: window.py:42-61 KBDWindow, computed on the host in
 * float64), mdct.py:16-42 MDCT with its 2/N factor.  coef is FRAME-major: coef[f*(N/2) + k] (the reference returns
 * the transpose, (N/2, F), float64).  N in {256, 512, 1024, 2048, 4096}. */
int64_t inr_stmdct_num_frames(int64_t T, int N);
int inr_stmdct_forward(const float* x, int64_t T, int N, const float* window, float* coef, void* stream);
/* mdct.py:74-113 ISTMDCT(coef, N) on the device: per frame IMDCT (mdct.py:29-35,46-48, factor 2), the same window,
 * overlap-add at hop N/2, last half frame dropped: out has (N/2) * num_frames samples.  coef is FRAME-major as above
 * (the reference takes the transpose (N/2, F)).  Deterministic (one writer per output sample). */
int inr_istmdct(const float* coef, int64_t num_frames, int N, const float* window, float* out, void* stream);
/* MDCT-domain loss (BASELINE.json config 3): weight * sum |STMDCT(pred) - target_coef| (INR_LOSS_MAE) or squared
 * (INR_LOSS_MSE), and its exact adjoint.  Deterministic (one writer per output sample). */
int inr_mdct_loss(const float* pred, const float* target_coef, int64_t T, int N, const float* window, int kind,
                  float weight, float* gout, inr_opt_state* opt, void* stream);

/* auraloss.freq.STFTLoss() (run.py:128,160; third-party, un-vendored: PARITY UNPINNED): magnitudes
 * sqrt(clamp(re^2+im^2, eps)) of torch.stft(x, n_fft, hop, n_fft, window, center=True, reflect), frame-major
 * mag[f*(n_fft/2+1) + k], F = 1 + T/hop; loss = ||Y-X||_F / ||Y||_F + mean |log X - log Y|.
 * inr_stft_mag precomputes the target side (and adds sum Y^2 to *sumsq_dev, a device double the caller zeroed). */
int64_t inr_stft_num_frames(int64_t T, int hop);
int inr_stft_mag(const float* x, int64_t T, int n_fft, int hop, const float* window, float eps, float* mag,
                 double* sumsq_dev, void* stream);
int inr_stft_loss(const float* pred, const float* target_mag, double target_sumsq, int64_t T, int n_fft, int hop,
                  const float* window, float eps, float weight, float* gout, inr_opt_state* opt, void* stream);

/* auraloss.time.SNRLoss(zero_mean=True, eps) (run.py:126,165 loss_mode='snr'; PARITY UNPINNED):
 * -10 log10( sum gt_c^2 / (sum (pred_c - gt_c)^2 + eps) + eps ).  scratch2: two device doubles, zero on entry/exit. */
int inr_snr_loss(const float* pred, const float* target, int64_t n, float eps, float weight, float* gout,
                 inr_opt_state* opt, double* scratch2, void* stream);

/* torch.optim.Adam(lr) + ReduceLROnPlateau('min', 0.8, 200, min_lr), run.py:116-117,186-189.
 * inr_opt_init writes a fresh state to device memory. */
int inr_opt_init(inr_opt_state* opt_dev, float lr, float min_lr, int32_t hist_cap, void* stream);

/* One optimiser step on P params: Adam update with the state's lr, then the plateau scheduler fed with
 *   loss = loss_acc[0] * loss_scale + loss_acc[1]
 * then step += 1, loss_hist[step % hist_cap] = loss, loss_acc zeroed, grads zeroed. */
int inr_adam_step(float* params, float* grads, float* m, float* v, int64_t P, inr_opt_state* opt,
                  float loss_scale, float* loss_hist, void* stream);

/* The exchange step of a coordinate-sharded fit as ONE collective: inr_opt_loss_to_f32 writes the two
 * float64 loss accumulators as fp32 hi, hi, lo, lo into tail4 (the 4 floats right behind the P gradient floats), the
 * caller sum-allreduces P + 4 floats, inr_opt_loss_from_f32 rebuilds the accumulators (spectral term * spectral_scale:
 * 1/world when every rank evaluated the whole-output loss) and zeroes tail4. */
int inr_opt_loss_to_f32(const inr_opt_state* opt, float* tail4, void* stream);
int inr_opt_loss_from_f32(inr_opt_state* opt, float* tail4, float spectral_scale, void* stream);

/* Fused single-GPU step = inr_siren_loss_backward + inr_adam_step (run.py:158-189), graph-capturable. */
int inr_siren_fit_step(const inr_siren_desc* d, float* params, float* grads, float* m, float* v,
                       const float* coords, const float* target, int64_t n, int64_t i0, int64_t T_total,
                       float coord_scale, int loss_kind, float alpha, inr_opt_state* opt, float* loss_hist,
                       void* workspace, int64_t workspace_bytes, void* stream);

/* ---- efficient-kan KAN (kan.py:6-285) -----------------------------------------------------------------------
 * KAN(layers_hidden, grid_size=5, spline_order=3, ...) with update_grid=False, the only mode run.py:158 uses.
 * flat params, state_dict order per KANLinear (kan.py:36-45): base_weight (out,in), spline_weight (out,in,G+K),
 * spline_scaler (out,in).  `grids` = the layers' `grid` buffers concatenated, (in_l, G+2K+1) floats each (kan.py:26-35),
 * so grids refined by the reference's host-side update_grid are honoured.  spline_order must be 3. */
#define INR_KAN_MAX_LAYERS 8
typedef struct {
  int32_t num_layers;                       /* len(layers_hidden) - 1                    */
  int32_t widths[INR_KAN_MAX_LAYERS + 1];   /* layers_hidden                             */
  int32_t grid_size;                        /* G (5)                                     */
  int32_t spline_order;                     /* K (3)                                     */
  int32_t precision;                        /* INR_PREC_FP32 or INR_PREC_BF16 */
} inr_kan_desc;

int64_t inr_kan_num_params(const inr_kan_desc* d);
int64_t inr_kan_num_grid_floats(const inr_kan_desc* d);
int64_t inr_kan_workspace_bytes(const inr_kan_desc* d, int64_t n_rows, int training);

/* kan.py:78-111 b_splines: x (rows, in) -> bases (rows, in, G+K), Cox-de Boor on grid (in, G+2K+1) with the
 * reference's half-open order-0 test and operation order. */
int inr_kan_b_splines(const float* x, const float* grid, int64_t rows, int in_features, int grid_size,
                      int spline_order, float* bases, void* stream);

/* kan.py:274-279 KAN.forward(x): out (n, widths[L]).  coords NULL (widths[0] == 1): generated as in inr_siren_forward. */
int inr_kan_forward(const inr_kan_desc* d, const float* params, const float* grids, const float* coords, int64_t n,
                    int64_t i0, int64_t T_total, float coord_scale, float* out, void* workspace, int64_t workspace_bytes,
                    void* stream);

/* run.py:158-169 + 185 for arch='kan' (run.py:110): forward, loss, backward to the flat gradient (ACCUMULATES);
 * same conventions as inr_siren_loss_backward; widths[L] must be 1. */
int inr_kan_loss_backward(const inr_kan_desc* d, const float* params, const float* grids, const float* coords,
                          const float* target, int64_t n, int64_t i0, int64_t T_total, float coord_scale, int loss_kind,
                          float grad_scale, const float* extra_gout_or_null, float* grads, inr_opt_state* opt,
                          float* out_or_null, void* workspace, int64_t workspace_bytes, void* stream);

/* Test hooks and diagnostics (inr_dbg_*) are declared in include/inr_b200_debug.h: they are not part of the product ABI. */

#ifdef __cplusplus
}
#endif
#endif /* INR_B200_H */
