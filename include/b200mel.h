/*
 * b200mel.h - C ABI of libb200mel.so: the B200 (sm_100a) audio -> log-mel / MFCC / spectrogram
 * (+ SpecAugment masking, + delta / delta-delta) feature front-end.
 *
 * This is the drop-in boundary for ONE hot path of cosmoquester/speech-recognition:
 * speech_recognition/data.py:122-328.  The reference has no FFI layer of its own - its "operator
 * API" is a set of Python factory functions that tf.data maps over single utterances - so every
 * entry point below names the reference function it replaces (file:line in the reference tree).
 * INTEGRATION.md shows the ctypes binding a maintainer of the reference would add.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / TF types cross this boundary;
 *   - every function returns an int status (B200MEL_OK == 0); b200mel_last_error() returns the
 *     text for the calling thread's last failure; nothing aborts or throws across the ABI;
 *   - "device" pointers are CUDA device pointers on the handle's device; the caller owns every
 *     input and output buffer, the handle owns only its constant tables;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); launches are
 *     asynchronous with respect to the host;
 *   - a handle is immutable after creation and may be used from several host threads.
 *   - there is NO CPU fallback: without a CUDA device b200mel_create fails.
 */
#ifndef B200MEL_H_
#define B200MEL_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200MEL_VERSION 100

/* status codes */
#define B200MEL_OK 0
#define B200MEL_ERR_INVALID_ARGUMENT 1 /* maps to Python ValueError (TF: InvalidArgumentError / ValueError) */
#define B200MEL_ERR_UNSUPPORTED 2      /* size outside what the kernels support */
#define B200MEL_ERR_CUDA 3             /* a CUDA runtime call failed */
#define B200MEL_ERR_NO_DEVICE 4        /* no usable CUDA device: there is no CPU path */

/* DataConfig.audio_feature_type (configs/data_config.py:30) */
#define B200MEL_FEATURE_SPECTROGRAM 0 /* make_spectrogram          data.py:122-142 */
#define B200MEL_FEATURE_LOG_MEL 1     /* make_log_mel_spectrogram  data.py:145-189 */
#define B200MEL_FEATURE_MFCC 2        /* make_mfcc                 data.py:192-241 */

/* PCM element type of the input.  I16 is the decoder's native type (data.py:98,104: int16/32768);
 * the kernel applies the exact power-of-two scale 2^-15 itself. */
#define B200MEL_PCM_I16 0
#define B200MEL_PCM_F32 1

/* kernel selection for b200mel_forward */
#define B200MEL_KERNEL_AUTO 0    /* fastest kernel that supports the configuration */
#define B200MEL_KERNEL_GENERIC 1 /* shared-memory Stockham FFT kernel, any parameters */
#define B200MEL_KERNEL_TC 2      /* tcgen05 tensor-core DFT kernel (fails if unsupported) */

/* Every numeric key of DataConfig (configs/data_config.py:23-54) that parameterises the path.
 * Field names follow the YAML keys of resources/configs/{libri,kspon,clovacall}_config.yml. */
typedef struct b200mel_config {
  int32_t struct_size;          /* = sizeof(b200mel_config), for ABI checking */
  int32_t audio_feature_type;   /* B200MEL_FEATURE_* */
  int32_t sample_rate;
  int32_t frame_length;
  int32_t frame_step;
  int32_t fft_length;           /* <= 0: smallest power of two >= frame_length (data.py:128) */
  int32_t num_mel_bins;         /* log-mel / mfcc */
  int32_t num_mfcc;             /* mfcc */
  float lower_edge_hertz;
  float upper_edge_hertz;
  float epsilon;                /* data.py:153 default 1e-12 */
  int32_t use_delta_accelerate; /* delta_accelerate, data.py:310-328: C = 3 instead of 1 */
  /* SpecAugmentConfig (configs/data_config.py:10-20); W (time warp) is a separate stage: b200mel_time_warp */
  int32_t spec_augment_enable;
  int32_t sa_F, sa_m_F;         /* frequency masking, data.py:282-288 */
  int32_t sa_T, sa_m_T;         /* time masking, data.py:290-301 */
  float sa_p;
} b200mel_config;

typedef struct b200mel_handle b200mel_handle;

int b200mel_version(void);
const char* b200mel_last_error(void);

/* Build the plan (window, FFT twiddles, mel filterbank in float32 in TF's op order, DCT matrix)
 * and upload it to `device`.  Validates like the reference does at trace time: mel edge checks of
 * tf.signal.linear_to_mel_weight_matrix (ValueError), DataConfig.__post_init__ requirements
 * (configs/data_config.py:56-62). */
int b200mel_create(const b200mel_config* config, int device, b200mel_handle** out_handle);
int b200mel_destroy(b200mel_handle* handle);

/* T = max(0, 1 + (n_samples - frame_length) / frame_step): tf.signal.stft(pad_end=False),
 * asserted by the reference at tests/test_data.py:76-78. */
int b200mel_num_frames(const b200mel_handle* handle, int64_t n_samples, int64_t* n_frames);

/* frequency_dim and feature_dim of DataConfig (configs/data_config.py:64-75), and the resolved
 * fft_length. */
int b200mel_feature_shape(const b200mel_handle* handle, int32_t* frequency_dim,
                          int32_t* feature_dim, int32_t* fft_length);

/* Host copy of the [fft_length/2+1, num_mel_bins] float32 filterbank the kernels use
 * (tf.signal.linear_to_mel_weight_matrix, data.py:175-179), row-major. */
int b200mel_mel_weights(const b200mel_handle* handle, float* out_host);

/* 1 if the tensor-core kernel supports this handle's configuration. */
int b200mel_tc_supported(const b200mel_handle* handle);

/* Number of kernel launches this handle has issued so far. */
int64_t b200mel_launch_count(const b200mel_handle* handle);

/*
 * The batched hot path: for each utterance b, feature (data.py:122-241) -> SpecAugment masks
 * (data.py:282-301, if enabled) -> delta/delta-delta (data.py:310-328, if enabled), zero padded to
 * t_max frames exactly like padded_batch (run/train.py:192-197).
 *
 *   pcm        device; int16 or float32 samples.  Utterance b starts at element
 *              offsets[b] (if offsets != NULL; ragged) else b * row_stride (dense [B, row_stride]).
 *   lengths    device int32[B]: valid samples of each utterance.
 *   out        device float32 [B, t_max, frequency_dim, feature_dim].  Frames >= n_frames[b] are
 *              +0.0; frames >= t_max are dropped (slice_example, data.py:345-354).
 *   n_frames   device int32[B] or NULL: receives the unclipped frame count of each utterance.
 *   mask_table device int32 [B, 2*(m_F+m_T)] (start, width) pairs, frequency masks first, or NULL
 *              to draw them from Philox4x32-10 keyed by (seed, first_utterance_id + b)
 *              (same stream as b200mel_specaug_draw).  Ignored unless spec_augment_enable.
 *   status     device int32[1] or NULL: OR-ed with 1 if a draw hit TF's `maxval <= minval` error
 *              case (the mask is then empty; tf.random.uniform raises there, data.py:284-285, 297),
 *              with 0x100 if a pipeline wait inside the tensor-core kernel timed out (a bug, never
 *              a hang).  The call is asynchronous: a caller that wants the reference's error
 *              behaviour reads the word after synchronising; the host-buffer calls below do it.
 */
int b200mel_forward(b200mel_handle* handle, const void* pcm, int32_t pcm_dtype,
                    const int64_t* offsets, const int32_t* lengths, int32_t batch,
                    int64_t row_stride, float* out, int64_t t_max, int32_t* n_frames,
                    const int32_t* mask_table, uint64_t seed, uint64_t first_utterance_id,
                    int32_t* status, int32_t kernel, void* stream);

/* Same computation with HOST buffers: pinned staging, H2D copy, kernel, D2H copy, synchronised on
 * return.  This is the call the end-to-end (e2e) throughput is measured through.  offsets may be
 * NULL (dense).  n_frames may be NULL.  The step's status word comes back with the results: the call
 * (or _wait) returns B200MEL_ERR_INVALID_ARGUMENT when a SpecAugment draw was invalid (the reference
 * raises), B200MEL_ERR_CUDA on a kernel protocol timeout; `out_host` is complete in both cases. */
int b200mel_forward_host(b200mel_handle* handle, const void* pcm_host, int32_t pcm_dtype,
                         const int64_t* offsets_host, const int32_t* lengths_host, int32_t batch,
                         int64_t row_stride, int64_t total_samples, float* out_host, int64_t t_max,
                         int32_t* n_frames_host, uint64_t seed, uint64_t first_utterance_id,
                         int32_t kernel);

/* The same call split in two, for pipelines that keep the link busy in both directions: _submit enqueues
 * the step (H2D chunks, kernels, D2H chunks) on staging slot 0 or 1 and returns; _wait blocks until that
 * step's `out_host` / `n_frames_host` are complete.  Submitting step i+1 on the other slot before waiting
 * for step i lets its uploads overlap step i's downloads - what tf.data's prefetch does for the reference
 * (run/train.py:194).  Host buffers must stay valid (and should be pinned) until the wait returns; a
 * submit on a slot whose previous step has not been waited for waits for it first.
 * b200mel_forward_host == _submit(slot 0) + _wait(0). */
int b200mel_forward_host_submit(b200mel_handle* handle, const void* pcm_host, int32_t pcm_dtype,
                                const int64_t* offsets_host, const int32_t* lengths_host, int32_t batch,
                                int64_t row_stride, int64_t total_samples, float* out_host,
                                int64_t t_max, int32_t* n_frames_host, uint64_t seed,
                                uint64_t first_utterance_id, int32_t kernel, int32_t slot);
int b200mel_forward_host_wait(b200mel_handle* handle, int32_t slot);

/*
 * No counterpart in the reference: added for callers that
 * cannot allocate device memory inside the call (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  The MFCC path of b200mel_forward (int16 PCM, tensor-core log-mel stage ->
 * DCT / SpecAugment / delta stage) with the intermediate log-mel tensor in a file-scope device array of
 * B200MEL_NOALLOC_SCRATCH_BYTES instead of pool memory: the call allocates and frees nothing.
 *   - `handle` must be an MFCC configuration for which b200mel_tc_supported() is 1 (the array's
 *     address is resolved once per device in b200mel_create).
 *   - batch * t_max * num_mel_bins * 4 bytes must fit the array (16 MiB: 52428 frames of 80 mel bins,
 *     about 524 s of audio per call); larger calls return B200MEL_ERR_UNSUPPORTED.
 *   - `out` must be 32-byte aligned.  Other arguments as in b200mel_forward.
 *   - every call on a device shares the one array: calls that overlap on the GPU (different streams,
 *     or different handles / threads without ordering between them) race.  Nothing in the array has to
 *     start at zero: the first stage rewrites every row it later reads.
 */
#define B200MEL_NOALLOC_SCRATCH_BYTES (16u << 20)
int b200mel_forward_mfcc_noalloc(b200mel_handle* handle, const int16_t* pcm, const int64_t* offsets,
                                 const int32_t* lengths, int32_t batch, int64_t row_stride, float* out,
                                 int64_t t_max, int32_t* n_frames, const int32_t* mask_table,
                                 uint64_t seed, uint64_t first_utterance_id, int32_t* status,
                                 void* stream);

/*
 * Stand-alone stages, for callers that compose the reference's per-example maps one by one
 * (run/train.py:88-116) instead of using the fused call.
 */

/* spec_augment draw order (data.py:282-301): (f, f_0) x m_F then (t, t_0) x m_T with
 * tf.random.uniform's integer mapping lo + u32 % (hi - lo).  num_time: device int32[B].
 * table: device int32 [B, 2*(m_F+m_T)].  status as above. */
int b200mel_specaug_draw(const int32_t* num_time, int32_t batch, int32_t v, int32_t F,
                         int32_t m_F, int32_t T, float p, int32_t m_T, uint64_t seed,
                         uint64_t first_utterance_id, int32_t* table, int32_t* status,
                         void* stream);

/* audio *= mask (data.py:286-288, 299-301) on [B, t_max, v, C] features; in == out allowed. */
int b200mel_specaug_apply(const float* in, float* out, int32_t batch, int64_t t_max, int32_t v,
                          int32_t channels, const int32_t* table, int32_t m_F, int32_t m_T,
                          void* stream);

/* SpecAugment time warp (data.py:275-280: tfa.image.sparse_image_warp with one control point (src_t, v/2) ->
 * (dst_t, v/2) and num_boundary_points = 3) on [B, t_max, v, channels] features, per utterance over its own
 * n_frames[b] rows (clipped to t_max); rows behind them are written as +0.0.  src_t / dst_t are drawn like the
 * reference draws them (uniform(W, T - W) and src + uniform(-W, W); Philox stream 1 of (seed, utterance id)), or
 * taken from warp_table (device int32 [B, 2], e.g. recorded TF draws).  An utterance the reference's draw would
 * raise on (T - W <= W) is copied unchanged and bit 0 of *status is set.  Not in place.  tensorflow-addons is an
 * un-vendored dependency of the reference: parity unpinned (oracle/time_warp_oracle.py). */
int b200mel_time_warp(const float* in, float* out, const int32_t* n_frames, int32_t batch, int64_t t_max,
                      int32_t v, int32_t channels, int32_t W, const int32_t* warp_table, uint64_t seed,
                      uint64_t first_utterance_id, int32_t* status, void* stream);

/* delta_accelerate (data.py:310-328): [B, t_max, v, 1] -> [B, t_max, v, 3]. n_frames (device
 * int32[B] or NULL) keeps the padding rows at zero. */
int b200mel_delta(const float* in, float* out, const int32_t* n_frames, int32_t batch,
                  int64_t t_max, int32_t v, void* stream);

/*
 * Rows either side of the path: PCM / WAV ingest (N1), the feature cache's
 * record framing (N2), and the batch contract of the models (N3).
 */

/* Host only.  Walk the RIFF chunks of a WAV file image and return the position of its sample data
 * (what tfio.audio.AudioIOTensor(path, tf.int16) decodes, data.py:96-99).  format is the WAVE format
 * tag (1 = integer PCM).  Fails with B200MEL_ERR_INVALID_ARGUMENT if the bytes are not a RIFF/WAVE image
 * with a fmt and a data chunk (the reference raises at decode time). */
int b200mel_wav_parse(const void* file_bytes, int64_t n_bytes, int32_t* format, int32_t* channels,
                      int32_t* sample_rate, int32_t* bits_per_sample, int64_t* data_offset,
                      int64_t* data_bytes);

/* Device.  load_audio_file for 16-bit little-endian PCM payloads (data.py:94-117: the "wav" branch after
 * b200mel_wav_parse, and the raw "pcm" branch incl. its zero byte appended to odd-length files): utterance
 * b is the byte range [byte_offsets[b], byte_offsets[b] + byte_lengths[b]) of `bytes`, any alignment.
 *   channels   device int32[B], or NULL = every payload is mono.  Multi-channel payloads are reduced with
 *              reduce_mean over channels (data.py:116) and therefore need pcm_dtype F32.
 *   pcm_out    device int16 (samples as stored: the front-end's native input) or float32 (= int16 / 32768),
 *              dense [B, row_stride] (out_offsets NULL; longer payloads are cut at row_stride, the rest of
 *              each row is zero) or ragged at element offsets out_offsets[b] (device int64[B]).
 *   lengths_out device int32[B]: samples written per utterance - the `lengths` of b200mel_forward. */
int b200mel_ingest_pcm16(const void* bytes, const int64_t* byte_offsets, const int64_t* byte_lengths,
                         const int32_t* channels, int32_t batch, void* pcm_out, int32_t pcm_dtype,
                         const int64_t* out_offsets, int64_t row_stride, int32_t* lengths_out,
                         void* stream);

/* Host only.  The masked CRC-32C of TFRecord framing (u64 length, u32 crc(length), payload,
 * u32 crc(payload)) used by the reference's feature cache (run/make_tfrecord.py:55-56, data.py:64-79). */
uint32_t b200mel_tfrecord_masked_crc(const void* data, int64_t n_bytes);

/* Length of the models' conv-reduced padding mask for t_max frames: per layer
 * `t = (t - (kernel - stride)) / stride` (models/las.py:209-213, models/deepspeech2.py:71-73).
 * group_out receives the product of the strides.  Host only. */
int b200mel_audio_mask_length(int64_t t_max, const int32_t* time_kernel_sizes, const int32_t* time_strides,
                              int32_t n_layers, int64_t* reduced_out, int32_t* group_out);

/* Device.  _audio_mask of the models (models/las.py:205-217, models/deepspeech2.py:68-78) on a padded batch
 * feat [B, t_max, row_elems] float32 (row_elems = frequency_dim * feature_dim): frame t is real when any of
 * its values != 0 (AUDIO_PAD_VALUE), the first reduced * group frames are OR-ed in groups of `group`.
 * mask_out: device uint8 [B, reduced] (1 = real). */
int b200mel_audio_mask(const float* feat, int32_t batch, int64_t t_max, int64_t row_elems, int32_t group,
                       int64_t reduced, uint8_t* mask_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200MEL_H_ */
