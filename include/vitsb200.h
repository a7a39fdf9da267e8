/*
 * vitsb200 — C ABI of the B200-native flow-reverse + Generator engine.
 *
 * Drop-in boundary for emotional-vits' latent->waveform stage.  The reference
 * has no FFI/plugin registry; its boundary is Python method signatures on
 * nn.Modules.  Each entry point below names the reference
 * interface it replaces; paths are relative to emotional-vits/.
 *
 * Conventions
 *   - every function returns a vitsb_status (0 = ok); nothing throws across
 *     the ABI; vitsb_last_error(h) gives the message of the last failure.
 *   - the caller owns every input, output and workspace buffer.  The handle
 *     owns the packed weights, its options, a small cache of instantiated CUDA
 *     graphs (replay of the workspace-only launches of a repeated shape), the
 *     private stream those are captured on and two side streams for the decoder's
 *     concurrent MRF branches.  Calls are asynchronous on `stream`.
 *   - threads: every entry point takes the handle's mutex, so a handle may be
 *     shared between threads; calls on ONE handle are serialised on the host
 *     (their device work is ordered by the streams passed in) and concurrent
 *     calls must use distinct workspaces.  Use one handle per thread to issue
 *     from several threads at once.
 *   - tensors at the seam use the reference's layout: channel-major
 *     [B, C, T] fp32, contiguous.  `len` is a HOST array of B valid frame
 *     counts (the reference has them on the host too: infer.py:171,
 *     models.py:547); rows are zero beyond len[b].
 *   - there is no CPU fallback: without a CUDA device every compute call
 *     fails with VITSB_ERR_CUDA.
 */
#ifndef VITSB200_H
#define VITSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vitsb_handle vitsb_handle;
typedef void* vitsb_stream;            /* cudaStream_t */

typedef enum {
  VITSB_OK = 0,
  VITSB_ERR_BAD_ARG = 1,      /* null pointer, bad shape, B/T out of range */
  VITSB_ERR_UNSUPPORTED = 2,  /* config the kernels do not cover (e.g. resblock "1", modules.py:185-220) */
  VITSB_ERR_WEIGHTS = 3,      /* missing / mis-shaped state_dict entry */
  VITSB_ERR_CUDA = 4,         /* CUDA runtime / driver failure, or no device */
  VITSB_ERR_WORKSPACE = 5     /* workspace too small */
} vitsb_status;

typedef enum {
  VITSB_PREC_FP16 = 0,   /* fp16 operands on tcgen05 (kind::f16), fp32 accumulate + fp32 residual streams.
                            Default 16-bit path: the reference's production dtype (infer.py:66 model.half()) */
  VITSB_PREC_BF16 = 1,   /* bf16 operands, same kernels */
  VITSB_PREC_FP32 = 2    /* fp32-verification build: CUDA-core FFMA implicit GEMM, accurate tanh/exp */
} vitsb_precision;

#define VITSB_MAX_LIST 8

/* Mirrors the hot-path subset of configs/base.json "model" (models.py:416-448). */
typedef struct {
  int32_t inter_channels;                 /* 192 */
  int32_t hidden_channels;                /* 256 */
  int32_t gin_channels;                   /* 1024 */
  int32_t n_flows;                        /* 4   (models.py:441) */
  int32_t flow_kernel_size;               /* 5   (models.py:462) */
  int32_t flow_n_layers;                  /* 4   (models.py:462) */
  int32_t flow_dilation_rate[VITSB_MAX_LIST];
  int32_t n_resblock_kernels;             /* 3 */
  int32_t resblock_kernel_sizes[VITSB_MAX_LIST];                    /* 3,7,11 */
  int32_t n_resblock_dilations;           /* 3 */
  int32_t resblock_dilation_sizes[VITSB_MAX_LIST][VITSB_MAX_LIST];  /* [kernel][iter] 1,3,5 */
  int32_t n_upsamples;                    /* 4 */
  int32_t upsample_rates[VITSB_MAX_LIST];        /* 8,6,2,2 */
  int32_t upsample_kernel_sizes[VITSB_MAX_LIST]; /* 16,12,4,4 */
  int32_t upsample_initial_channel;       /* 512 */
  int32_t precision;                      /* vitsb_precision */
  int32_t enc_n_heads;                    /* 2   text-encoder attention heads (infer_p1 side; 0 = 2).  Every other
                                             encoder / duration-predictor size is read off the checkpoint tensors */
} vitsb_config;

typedef enum { VITSB_F32 = 0, VITSB_F16 = 1, VITSB_BF16 = 2 } vitsb_dtype;

/* One state_dict entry, HOST memory, contiguous, named by its reference key
 * (e.g. "dec.ups.0.weight_v"). */
typedef struct {
  const char* name;
  const void* data;
  int32_t dtype;        /* vitsb_dtype */
  int32_t ndim;
  int64_t shape[4];
} vitsb_tensor_desc;

/* Replaces: constructing models.Generator (models.py:282-304) and
 * models.ResidualCouplingBlock (models.py:192-217) for inference. */
int vitsb_create(const vitsb_config* cfg, int device, vitsb_handle** out);
void vitsb_destroy(vitsb_handle* h);
const char* vitsb_last_error(const vitsb_handle* h);   /* h may be NULL: last create() error */

/* Replaces: load_state_dict (export.py:59) + SynthesizerTrn.remove_weight_norm
 * (models.py:467-474) for the dec.* / flow.* keys.  Accepts folded `weight` or
 * `weight_g`/`weight_v` pairs; other keys are ignored.  Folds, repacks
 * (tap-major, Flip folded into pre/post, gate halves tiled, polyphase ups) and
 * uploads.  May be called again to swap checkpoints. */
int vitsb_load_weights(vitsb_handle* h, const vitsb_tensor_desc* tensors, int n);

/* Options: per handle, never read from the environment.  Keys (int values):
 *   "cuda_graph" 0/1 (1)   replay the workspace-only launches of a repeated shape as a CUDA graph
 *   "pdl" 0/1 (1)          programmatic dependent launch between the tensor-core kernels
 *   "tc_pair" 0/1 (1)      CTA-pair (cta_group::2) mode of the generic conv kernel and of the coupling kernel
 *   "branch_streams" 0/1 (1)  the three MRF branches of a stage run concurrently (two extra streams owned by the handle)
 *   "rb_nslab" 0..16 (0)   fused-ResBlock slab count override, 0 = automatic
 *   "ups_split_mask" -1..255 bit i: upsampler i runs 3-term split precision (16-bit paths; -1 = automatic: fp16 the
 *       first two, bf16 all); next vitsb_load_weights
 *   "ws_cap_mb" (24576)    workspace cap; batches whose layout exceeds it run in row chunks
 *   "rb_fused", "flow_fused", "flow_in_split" -1/0/1 (-1 = automatic): launch-plan / packing choices; they take
 *       effect at the next vitsb_load_weights.
 * vitsb_get_option also answers "plan.rb_fused", "plan.flow_fused", "plan.flow_in_split" (the plan in force) and
 * "plan.p1" (1 when the infer_p1 weights were loaded). */
int vitsb_set_option(vitsb_handle* h, const char* key, int value);
int vitsb_get_option(const vitsb_handle* h, const char* key, int* value);

/* Bytes of device workspace for a call with B rows of at most T_max frames (call after vitsb_load_weights).  Rows are
 * independent, so any workspace of at least vitsb_workspace_bytes(h, 1, T_max) is accepted: a batch whose full layout
 * does not fit the workspace handed in (or exceeds "ws_cap_mb" / 512 rows) is processed in row chunks. */
size_t vitsb_workspace_bytes(const vitsb_handle* h, int B, int T_max);

/* Replaces: ResidualCouplingBlock.infer(x, g, reverse=True) (models.py:228-235)
 * and the masked .forward(x, x_mask, g, reverse=True) (models.py:219-226):
 * rows beyond len[b] are treated as masked (zero).
 *   z_p, z : device fp32 [B, inter, T_max];  g : device fp32 [B, gin]. */
int vitsb_flow_reverse(vitsb_handle* h, const float* z_p, const float* g, const int32_t* len,
                       int B, int T_max, float* z, void* ws, size_t ws_bytes, vitsb_stream stream);

/* Replaces: Generator.forward(x, g) (models.py:306-318).
 *   ragged = 1: row b equals the B=1 call on its own len[b] frames, zero beyond
 *               (SynthesizerTrn.infer_p2 semantics, models.py:573-574)
 *   ragged = 0: every row is decoded over all T_max frames, unmasked
 *               (SynthesizerTrn.inference semantics, models.py:534)
 *   z : device fp32 [B, inter, T_max];  wav : device fp32 [B, 1, T_max*prod(upsample_rates)]. */
int vitsb_decode(vitsb_handle* h, const float* z, const float* g, const int32_t* len,
                 int B, int T_max, int ragged, float* wav, void* ws, size_t ws_bytes, vitsb_stream stream);

/* Replaces: the tail of SynthesizerTrn.infer_p2 (models.py:573-574) /
 * inference (models.py:533-534): flow-reverse then decode without leaving the
 * time-major layout.  z_out may be NULL. */
int vitsb_synthesize(vitsb_handle* h, const float* z_p, const float* g, const int32_t* len,
                     int B, int T_max, int ragged, float* z_out, float* wav,
                     void* ws, size_t ws_bytes, vitsb_stream stream);

/* vitsb_synthesize with the serving path's output post-processing fused into the last kernel:
 * pcm = int16(clip(wav * volume * 32767, -32768, 32767)), float32 products, truncation toward zero, exactly
 * VITSWrap.speaking's np.clip(...).astype(np.int16) (vits_wrap.py:195) on infer.py:182's float waveform.
 *   pcm : int16 [B, 1, T_max*prod(upsample_rates)], zero beyond len[b] when ragged.  Device memory, or page-locked host
 *         memory (device-accessible under unified addressing): the last kernel then writes the samples over PCIe and the
 *         caller needs no device buffer and no copy.  The same holds for `wav` of vitsb_synthesize / vitsb_decode. */
int vitsb_synthesize_pcm16(vitsb_handle* h, const float* z_p, const float* g, const int32_t* len,
                           int B, int T_max, int ragged, float volume, int16_t* pcm,
                           void* ws, size_t ws_bytes, vitsb_stream stream);

/* Replaces: the two librosa.resample calls of VITSWrap.speaking (vits_wrap.py:186-194: `pitch` resamples from
 * int(sr / pitch) to sr, `sampling_rate` from sr to the requested rate) and, with `pcm`, the int16 conversion after them
 * (vits_wrap.py:195), on the device.  The reference's resampler is whatever librosa defaults to (soxr_hq
 * today, unpinned, not in this image): this entry point is the polyphase method of scipy.signal.resample_poly -
 * y[n] = sum_k taps[(n + n_pre_remove) * down - k * up] * x[k], float64 accumulation - and is tested against scipy.
 *   x     : device fp32, row b at x + b * x_stride, n_in samples per row
 *   taps  : device float64 [n_taps]: the zero-padded low-pass filter scaled by `up` (vits_b200/resample.py computes it
 *           and n_pre_remove / n_out per (orig_sr, target_sr); up / down are the rates divided by their gcd)
 *   y/pcm : device (or page-locked host) fp32 / int16, row b at + b * y_stride, n_out samples per row; one may be NULL.
 *           pcm = int16(clip(y * volume * 32767, -32768, 32767)) as in vitsb_synthesize_pcm16. */
int vitsb_resample_poly(vitsb_handle* h, const float* x, int B, long long n_in, long long x_stride, int up, int down,
                        const double* taps, int n_taps, long long n_pre_remove, long long n_out, float* y, int16_t* pcm,
                        float volume, long long y_stride, vitsb_stream stream);

/* Replaces: the head of SynthesizerTrn.infer_p2 (models.py:570-572) - the two length-regulation matmuls and the
 * noise FMA - as one gather/FMA kernel.  attn is the duration path of commons.generate_path
 * (commons.py:143-155); zero weights are skipped, any other weights are summed like the matmul.
 *   attn : device fp32 [B, T, Tx];  m_p, s_p : device fp32 [B, inter, Tx];  noise, z_p : device fp32 [B, inter, T]. */
int vitsb_length_regulate(vitsb_handle* h, const float* attn, const float* m_p, const float* s_p, const float* noise,
                          int B, int T, int Tx, float* z_p, vitsb_stream stream);

/* Replaces: the host glue between infer_p1 and infer_p2 in EmoVITS.infer (infer.py:169-176: w = exp(logw) * rate,
 * w_ceil = ceil(w), y_length = clamp_min(sum, 1).item(), attn = infer_path(...); commons.py:143-155), the head of infer_p2
 * (models.py:570-572) and its tail (flow + decoder) in ONE stream-ordered call with no host synchronisation: durations,
 * their cumulative sum and the output lengths are computed on the device, the gather m_p + noise * s_p is written
 * straight into the time-major flow state (no dense attn, no channel-major z_p), and every kernel reads the lengths
 * from device memory (grids use upper bounds).  SURVEY 8(f) N1.
 *   m_p, s_p : device fp32 [B, inter, Tx_max];  logw : device fp32 [B, 1, Tx_max];  g : device fp32 [B, gin]
 *   x_len    : HOST int32 [B] phonemes per row, or NULL (all Tx_max);  Tx_max <= 1024
 *   noise    : device fp32, already scaled by noise_scale (infer.py:69); noise_stride >= T_cap: [B, inter, noise_stride];
 *              noise_stride = 0: the reference's compact form - row b is a contiguous [inter, y_len[b]] block placed
 *              after the blocks of the rows before it (infer.py:172-174 views a flat pool as [1, inter, y_length])
 *   T_cap    : frame capacity of the outputs; rows longer than T_cap are truncated to it
 *   wav / pcm: device fp32 / int16 [B, 1, T_cap * prod(upsample_rates)], one of them may be NULL; zero beyond y_len[b]
 *   z_out    : optional device fp32 [B, inter, T_cap];  y_len : optional device int32 [B] (frames per row)
 *   ws       : vitsb_workspace_bytes(h, B, T_cap) bytes (this entry point does not chunk rows). */
int vitsb_synthesize_durations(vitsb_handle* h, const float* m_p, const float* s_p, const float* logw, const float* g,
                               const int32_t* x_len, const float* noise, long long noise_stride, int B, int Tx_max, int T_cap,
                               float length_scale, int ragged, float* z_out, float* wav, int16_t* pcm, float volume,
                               int32_t* y_len, void* ws, size_t ws_bytes, vitsb_stream stream);

/* Replaces: SynthesizerTrn.infer_p1(x, emo, sid) (models.py:558-566) - speaker embedding, TextEncoder.infer
 * (models.py:180-189; attentions.py:48-54, 78-102, 160-166), exp(logs_p), DurationPredictor.infer (models.py:59-67) -
 * in fp32 (the durations feed ceil(exp(logw)), infer.py:169-171).  Available when vitsb_load_weights was given the
 * checkpoint's enc_p.* / dp.* / emb_g.* entries, else VITSB_ERR_WEIGHTS.  Rows are ragged: row b equals the B = 1 call on
 * its own x_len[b] phonemes (x_len = NULL: every row has Tx_max), outputs are zero beyond.
 *   text : device fp32 [B, Tx_max, text_channels];  emo : device fp32 [B, 1024];  sid, x_len : HOST int32 [B]
 *   m_p, s_p : device fp32 [B, inter, Tx_max];  logw : device fp32 [B, 1, Tx_max];  g : device fp32 [B, gin]
 *   Tx_max <= 1024. */
size_t vitsb_p1_workspace_bytes(const vitsb_handle* h, int B, int Tx_max);
int vitsb_infer_p1(vitsb_handle* h, const float* text, const float* emo, const int32_t* sid, const int32_t* x_len,
                   int B, int Tx_max, float* m_p, float* s_p, float* logw, float* g,
                   void* ws, size_t ws_bytes, vitsb_stream stream);

/* No counterpart in the reference: added for callers that cannot allocate
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc) and so cannot hand
 * vitsb_flow_reverse a workspace.  Same computation and arguments as vitsb_flow_reverse (fused coupling-layer plan
 * only: fp16 / bf16 precision), but the scratch - time-major flow state, its operand copy, the flow's conditioning
 * vectors, the device copy of len - lives in ONE static device array of VITSB_NOALLOC_SCRATCH_BYTES that every GPU
 * which loads the library holds until exit.  The call allocates nothing (no cudaMalloc / cudaFree, no stream or graph
 * creation).  Shapes whose scratch exceeds the array (more than about 5 000 padded frames B * T_max, plus 32 KB per batch row) are rejected with
 * VITSB_ERR_WORKSPACE and a message; other plans with VITSB_ERR_UNSUPPORTED.  Every call shares the array, which keeps
 * the previous call's contents (nothing in it needs to start at zero: each region is written before it is read within a
 * call).  Calls that overlap on the GPU - different streams, or different handles on one device - RACE on it:
 * serialise them. */
#define VITSB_NOALLOC_SCRATCH_BYTES (8u << 20)
int vitsb_flow_reverse_noalloc(vitsb_handle* h, const float* z_p, const float* g, const int32_t* len,
                               int B, int T_max, float* z, vitsb_stream stream);
size_t vitsb_noalloc_scratch_bytes(void);   /* = VITSB_NOALLOC_SCRATCH_BYTES, as built */

/* Introspection for tests and bench.py. */
int vitsb_total_upsample(const vitsb_handle* h);
/* Number of kernels one vitsb_synthesize call launches for (B, T_max). */
int vitsb_launches_per_call(const vitsb_handle* h, int B, int T_max);
/* Algorithmic FLOP per latent frame (2*Cin*Cout*k per output position). */
double vitsb_flop_per_frame(const vitsb_handle* h);
/* Per-launch device timing for bench.py / profiling: while enabled every kernel
 * launch of a call is bracketed by CUDA events on the call's stream and the call
 * ends with a stream synchronisation.  vitsb_profile_read returns one row per
 * launch family (tag: 0 pack, 1 cond, 2 unpack, 3 conv_post, 4 mrf_sum, 10-14 flow
 * pre/in/res/skip/post (layer-at-a-time flow), 15 fused coupling layer, 20 conv_pre, 30+s ups, 40+s ResBlock c1, 50+s ResBlock c2
 * of stage s) with summed milliseconds, launch count and algorithmic FLOP. */
int vitsb_profile(vitsb_handle* h, int enable);
int vitsb_profile_read(const vitsb_handle* h, int32_t* tags, double* ms, int64_t* launches, double* flops, int cap);
/* Library build info: "sm_100a;tcgen05;..." */
const char* vitsb_build_info(void);

/* Bring-up self-test of one implicit-GEMM conv tile configuration on the
 * tcgen05 path against the CUDA-core fp32 path (tests/test_conv_unit.py).
 * Returns max-abs difference in *max_err.  gate: 0 = plain conv with residual and fp32 + 16-bit outputs, 1 = gated,
 * 2 = 16-bit output only (the staged TMA-store epilogue; *max_err compares 16-bit values, so one ulp of the operand
 * type at the output scale is expected where the fp32 sums straddle a rounding boundary). */
int vitsb_selftest_conv(int device, int precision, int Cin, int N, int k, int dil, int L, int B,
                        int gate, uint32_t seed, double* max_err, double* ref_scale);

#ifdef __cplusplus
}
#endif
#endif /* VITSB200_H */
