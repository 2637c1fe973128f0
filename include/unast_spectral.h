/* libunast_spectral -- C ABI of the B200 (sm_100a) spectral signal path of Lucaskabela/UNAST.
 *
 * The reference has no FFI layer for this path: the boundary is two Python functions
 * (src/utils.py:235 get_spectrograms, src/utils.py:281 spectrogram2wav), one script loop
 * (src/gl_vocoder.py:10 vocode) and the .npy layout written at src/prepare_data.py:33-34.
 * The entry points below are what a ctypes binding for those call sites needs; INTEGRATION.md
 * shows the reference-side stub.  Every entry point cites the reference lines it replaces.
 *
 * Conventions
 *   - all data pointers are DEVICE pointers owned by the caller (e.g. torch tensors' data_ptr());
 *     `*_host` arguments are host arrays read before the call returns;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), allocates nothing,
 *     and returns 0 on success or a negative UNAST_E_* code (message: unast_last_error());
 *   - one context per (device, host thread); audio constants are those of
 *     src/audio_parameters.py:1-20 (sr 22050, n_fft 2048, hop 275, win 1102, 80 mels,
 *     preemphasis 0.97, ref_db 20, max_db 100) and are compiled in;
 *   - there is no CPU fallback: without a CUDA device unast_create fails.
 */
#ifndef UNAST_SPECTRAL_H_
#define UNAST_SPECTRAL_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UNAST_ABI_VERSION 2

#define UNAST_OK 0
#define UNAST_E_CUDA (-1)      /* a CUDA runtime call failed */
#define UNAST_E_ARG (-2)       /* invalid argument */
#define UNAST_E_WORKSPACE (-3) /* workspace too small */
#define UNAST_E_DATA (-4)      /* data-dependent failure reported by a kernel (unast_check) */

/* sample formats of `wave` (librosa.load gives float32; LJSpeech files are PCM16 = value / 32768) */
#define UNAST_F32 0
#define UNAST_I16 1

/* arithmetic of the front-end FFT.  F64 mirrors numpy.fft (float64) as used by librosa.stft and is the
 * mode that meets the 1e-4 parity gate on every bin; F32 is ~2x faster, max error ~2.5e-4 on bins that
 * sit 90 dB below the frame peak ("precision"). */
#define UNAST_PREC_F32 0
#define UNAST_PREC_F64 1

/* mel output layouts */
#define UNAST_MEL_MAJOR_PER_UTT 0 /* utterance b: float[80][T_b] at 80*frame_off[b] == payload of the
                                     fortran_order (T_b, 80) .pt.npy written at prepare_data.py:33 */
#define UNAST_MEL_PADDED_BTM 1    /* float[B][T_max][80], rows t >= T_b untouched (caller zero-fills):
                                     the batch tensor of preprocess.py:82-121 collate_fn_transformer */

typedef struct unast_ctx unast_ctx;

int unast_abi_version(void);
const char* unast_last_error(void);

/* Builds the constant tables (periodic Hann, twiddles, Slaney filterbank = librosa.filters.mel at
 * utils.py:263, window-sum envelope) in float64 on the host and uploads them to `device`. */
int unast_create(int device, unast_ctx** out);
int unast_destroy(unast_ctx* ctx);
int unast_num_sms(const unast_ctx* ctx);

/* Blocks until `stream` is idle and returns UNAST_E_DATA if any kernel flagged a data error since the
 * last check (utterance too short for reflect padding, empty signal after trimming). */
int unast_check(unast_ctx* ctx, void* stream);

/* The 80 x 1025 filterbank the context uses (row-major float32), for inspection / tests. */
int unast_get_mel_basis(const unast_ctx* ctx, float* out_host);

/* ---- forward path: utils.get_spectrograms (utils.py:235-278) -------------------------------- */

/* librosa.effects.trim(y) of utils.py:248 / :304 / gl_vocoder.py:32 (top_db 60, frame 2048, hop 512).
 * wave: B utterances back to back; wave_off[B+1] (device, int64) their first samples.
 * bounds_out[B][2] (device, int32) receives start / end relative to each utterance.
 * max_len = longest utterance in samples (host-known).  workspace >= unast_trim_workspace_bytes. */
size_t unast_trim_workspace_bytes(int B, int max_len);
int unast_trim_bounds(unast_ctx* ctx, const void* wave, int dtype, const int64_t* wave_off, int B, int max_len,
                      int32_t* bounds_out, void* workspace, size_t workspace_bytes, void* stream);

/* frame_off[B+1] (device, int32) = prefix sum of T_b = 1 + (end_b - start_b) / 275 (librosa.stft, center=True). */
int unast_frame_offsets(unast_ctx* ctx, const int32_t* bounds, int B, int32_t* frame_off_out, void* stream);

/* utils.py:251-276: pre-emphasis, STFT(2048/275/1102 Hann), |.|, mel projection, dB, normalisation.
 * mag_out: float[sum T][1025] (utterance b's rows at frame_off[b]; == payload of the C-ordered
 * (T_b, 1025) .mag.npy of prepare_data.py:34) or NULL to skip it (loader path).
 * workspace >= unast_frontend_workspace_bytes(B). */
size_t unast_frontend_workspace_bytes(int B);
int unast_frontend(unast_ctx* ctx, const void* wave, int dtype, const int64_t* wave_off, const int32_t* bounds,
                   const int32_t* frame_off, int B, int precision, float* mel_out, int mel_layout, int T_max,
                   float* mag_out, void* workspace, size_t workspace_bytes, void* stream);

/* ---- inverse path: utils.spectrogram2wav (utils.py:281-306) / gl_vocoder.vocode (gl_vocoder.py:14-32) */

/* mag_norm: float[sum T][1025] normalised magnitudes; frames_host[B] = T_b (host; every T_b >= 5).
 * n_iter / momentum / angles0 select the variant:
 *   utils.griffin_lim (utils.py:309-320): momentum = 0, angles0 = NULL (zero phase), n_iter = ap.n_iter
 *   librosa.griffinlim (gl_vocoder.py:26): momentum = 0.99, n_iter = 32, angles0 = exp(2 pi i U)
 * angles0: complex64[sum T][1025] (interleaved re,im; frame-major) or NULL.
 * power = ap.power (1.2) de-normalises mag_norm as utils.py:289-298 does; power = 0 means mag_norm already is
 * the linear target magnitude S (utils.griffin_lim / invert_spectrogram called directly).
 * deemphasis != 0 applies lfilter([1],[1,-0.97]) in float64 (utils.py:301).
 * wav_out: float[sum 275 (T_b - 1)], utterance b at wav_off = prefix sum.
 * bounds_out (device int32 [B][2]) receives librosa.effects.trim bounds of each output (utils.py:304) or NULL. */
size_t unast_griffinlim_workspace_bytes(const int32_t* frames_host, int B, float momentum);
int unast_griffinlim(unast_ctx* ctx, const float* mag_norm, const int32_t* frames_host, int B, int n_iter,
                     float momentum, const float* angles0, float power, int deemphasis, float* wav_out,
                     int32_t* bounds_out, void* workspace, size_t workspace_bytes, void* stream);

/* ---- one long waveform sharded over GPUs by frame range (the loop of utils.py:309-320) ------
 * A *fragment* is the contiguous range [t0, t0 + n) of the frames of an utterance of tglob frames (n >= 9 unless the
 * fragment is the whole utterance).  Where two fragments of one utterance meet, the 827 samples their frames share (a
 * *linked seam*) exist in both fragments' overlap-add buffers; every iteration both sides add their partial sums into
 * both copies (plain stores / atomics over NVLink peer mappings) and publish their progress in the partner's mailbox.
 * No collective, no host round trip; the kernel is one persistent cooperative launch per call.
 *   side 0 = the seam at the fragment's first frame, side 1 = the seam after its last frame;
 *   peer_ola[k] = address, in the PARTNER's copy of rotating buffer k, of the seam's first sample: the partner's
 *                 workspace base + layout[k] + 4 * (partner's pad_off[fragment'] + 473 (+ 275 n' for its side 1)),
 *                 pad_off = prefix sum of 2048 + 275 (n - 1) over the partner's fragments;
 *   peer_mail   = address of the partner's mailbox slot for this link (its workspace base + layout[3] + 4 * slot);
 *   my_mail     = this side's mailbox slot (< 16) -- the partner's peer_mail must point at it.
 * epoch >= 1 must be the same on both sides and grow by at least n_iter + 2 from call to call (mailboxes are never
 * reset; a workspace must be zero-filled before its first use -- unast_peer_alloc does that).
 * raw_out receives, fragment after fragment, the samples each fragment owns (unast_griffinlim_fragment_samples) of
 * istft(S * angles): normalised by the window-sum envelope, NOT de-emphasised or trimmed (both need the whole
 * waveform: gather, then unast_deemph_trim). */
typedef struct unast_gl_link {
    int32_t fragment;
    int32_t side;
    void* peer_ola[3];
    void* peer_mail;
    int32_t my_mail;
} unast_gl_link;

/* Canonical summation blocks (optional; default 0 = off).  With frames = G >= 4 the overlap-add of every later
 * unast_griffinlim / unast_griffinlim_fragments call of this context restarts its running sums at every multiple of G
 * frames of an utterance, whatever the cut into runs, fragments and ranks: a sample is then always (sum over the frames
 * of one block, in frame order) + (the same of the next block), so the result is BIT-IDENTICAL for every sharding whose
 * linked seams sit on multiples of G (a one-GPU run of the whole waveform included).  Costs granularity (runs are whole
 * blocks): ~15 % on one GPU.  Off, the cut is chosen for speed and results of different shardings agree to ~1 ulp per
 * iteration at the seams (which Griffin-Lim then amplifies like any other rounding). */
int unast_griffinlim_set_block(unast_ctx* ctx, int frames);

/* layout_out[5] = byte offsets of the three rotating overlap-add buffers and of the mailbox array inside a workspace
 * for these fragment sizes, and the workspace size (== unast_griffinlim_workspace_bytes).  Host-only, deterministic. */
int unast_griffinlim_layout(const int32_t* frames_host, int B, float momentum, size_t* layout_out);
int unast_griffinlim_fragment_samples(int32_t t0, int32_t n, int32_t tglob, int64_t* first_out, int64_t* end_out);
int unast_griffinlim_fragments(unast_ctx* ctx, const float* mag_norm, const int32_t* frames_host, const int32_t* t0_host,
                               const int32_t* tglob_host, int B, const unast_gl_link* links_host, int n_links, int epoch,
                               int n_iter, float momentum, const float* angles0, float power, float* raw_out,
                               void* workspace, size_t workspace_bytes, void* stream);

/* scipy.signal.lfilter([1], [1, -0.97]) in float64 (utils.py:301) and librosa.effects.trim bounds (utils.py:304) of B
 * plain waveforms x (float32, off[B+1] int64 device offsets): the epilogue of the sharded run after the gather. */
size_t unast_deemph_trim_workspace_bytes(int B, int max_len);
int unast_deemph_trim(unast_ctx* ctx, const float* x, const int64_t* off, int B, int max_len, int deemphasis,
                      float* wav_out, int32_t* bounds_out, void* workspace, size_t workspace_bytes, void* stream);

/* Device memory other processes on the node can map (CUDA IPC; peer access over NVLink is enabled on open).
 * handle64: 64 bytes to hand to the peer process (e.g. torch.distributed.all_gather_object). */
int unast_peer_alloc(unast_ctx* ctx, size_t bytes, void** dptr_out, unsigned char* handle64_out);
int unast_peer_open(unast_ctx* ctx, const unsigned char* handle64, void** dptr_out);
int unast_peer_close(unast_ctx* ctx, void* dptr);
int unast_peer_free(unast_ctx* ctx, void* dptr);

/* ---- next row: the masking ops applied to the padded mel batch ----------- */

/* utils.specaugment (utils.py:51-75, called at train.py:236) and utils.noise_fn (utils.py:40-49, called at
 * network.py:205,307) on a float[B][T][D] batch, in one pass:
 *     out[b][t][:] = (t in [f_zero, f_zero+f) or t in [t_zero, t_zero+t) ? fill_b : x[b][t][:]) * keep[b][t]
 * spans: device int32 [B][4] = {f_zero, f, t_zero, t} as the reference draws them (both spans run along the TIME
 *        axis: `res[i][:][a:b]` slices rows), or NULL for no spans; fill_b = mean of the whole padded slab x[b]
 *        (utils.py:59), or 0 if replace_with_zero.
 * keep:  device float [B][T] = the reference's `zero_mask` (torch.bernoulli(1 - mask_p)), or NULL.
 * out may alias x only when spans is NULL or replace_with_zero (the mean is read from x).
 * workspace >= unast_mel_augment_workspace_bytes(B) (partial sums of the mean). */
size_t unast_mel_augment_workspace_bytes(int B);
int unast_mel_augment(unast_ctx* ctx, const float* x, int B, int T, int D, const int32_t* spans, const float* keep,
                      int replace_with_zero, float* out, void* workspace, size_t workspace_bytes, void* stream);

/* unast_mel_augment for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
 * No counterpart in the reference.  Same arguments and results as
 * unast_mel_augment without the workspace: the partial sums of the mean live in a file-scope __device__ array of the library
 * (UNAST_AUG_STATIC_MAX_B x 32 doubles = 256 KB, held by every GPU that loads the library until the process exits), and the
 * call performs no cudaMalloc / cudaFree / allocating library call.  B > UNAST_AUG_STATIC_MAX_B returns UNAST_E_ARG.
 * Every call on a device shares that array (it keeps the last call's contents; each call overwrites what it reads, nothing
 * needs clearing): calls that overlap on the GPU -- different streams, or two contexts of one device -- race; serialise them. */
#define UNAST_AUG_STATIC_MAX_B 1024
int unast_mel_augment_static(unast_ctx* ctx, const float* x, int B, int T, int D, const int32_t* spans, const float* keep,
                             int replace_with_zero, float* out, void* stream);

/* The masking ops above fused with the first layer of SpeechPrenet (module.py:93-95: fc1 = Linear(num_mels -> s_pre_hid),
 * ReLU), on the tensor cores (tcgen05.mma, A operand in tensor memory, float32 fidelity by operand splitting):
 *     y[b][t][:] = relu(W ((t in a span ? fill_b : x[b][t][:]) * keep[b][t]) + bias)          x: [B][T][80], y: [B][T][n_out]
 * unast_prenet_pack turns nn.Linear's weight [n_out][80] (+ bias [n_out] or NULL) into the image the kernel keeps in shared
 * memory (unast_prenet_packed_bytes(n_out) bytes, device, 16-byte aligned; n_out a multiple of 32, <= 256); repack after
 * every weight update.  spans / keep / replace_with_zero as in unast_mel_augment; relu = 0 returns the pre-activation.
 * Dropout (module.py:96) is left to the caller.  workspace >= unast_mel_augment_workspace_bytes(B). */
size_t unast_prenet_packed_bytes(int n_out);
int unast_prenet_pack(unast_ctx* ctx, const float* weight, const float* bias, int n_out, void* packed_out, void* stream);
int unast_mel_prenet_fc1(unast_ctx* ctx, const float* x, int B, int T, const int32_t* spans, const float* keep,
                         int replace_with_zero, const void* packed, int n_out, int relu, float* y, void* workspace,
                         size_t workspace_bytes, void* stream);

/* ---- diagnostics (used by the test-suite to pin the FFT core on its own) --------------------- */

/* frames: float[F][2048] whose samples [1102, 2048) are zero; out: complex64[F][1025] = rfft(frame). */
int unast_debug_rfft(unast_ctx* ctx, const float* frames, int F, int precision, float* out, void* stream);
/* spec: complex64[F][1025]; out: float[F][1102] = first 1102 samples of irfft(spec, 2048). */
int unast_debug_irfft(unast_ctx* ctx, const float* spec, int F, float* out, void* stream);

/* Sustained FMA issue-rate probe used for the roofline denominators that MEASURED_PEAKS.json lacks:
 * kind 0 = FFMA (float32), 1 = DFMA (float64), 2 = FFMA2 (packed float32x2).  Launches num_sms*8 CTAs of 256
 * threads, each thread `iters` x 8 independent FMAs; *flop_out (host) = flop the launch performs; out: float[num_sms*8*256]. */
int unast_debug_pipe_peak(unast_ctx* ctx, int kind, int iters, float* out, double* flop_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* UNAST_SPECTRAL_H_ */
