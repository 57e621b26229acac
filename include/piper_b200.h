/*
 * piper_b200.h - C ABI of the B200-native Piper-VITS inference runtime.
 *
 * This is the drop-in boundary for the reference's hot path: everything under
 * PiperMetalRuntime.synthesize (Sources/PiperMetal/PiperMetalRuntime.swift:62-80) ->
 * GraphExecutor.executeOutput (Sources/PiperMetal/Execution/GraphExecutor.swift:156-327)
 * -> 2755 x executeNode -> MetalBackend -> the .metal kernels.  A Swift (or C++/Python)
 * host binds these symbols instead of constructing a GraphExecutor; see INTEGRATION.md
 * for the SwiftPM module and the PiperCUDARuntime wrapper that keeps the
 * synthesize(phonemeIDs:noiseScale:lengthScale:noiseW:) signature.
 *
 * Plain C: pointers and sizes only.  No exceptions cross this boundary; every call
 * returns a status code and piper_b200_last_error() gives the message (the reference
 * throws ExecutionError / NSError, CPUBackend.swift:3-17, PiperMetalRuntime.swift:74-78).
 *
 * Threading: one context per GPU; calls on one context must be serialised by the caller
 * (the reference's runtime is likewise single-caller, GraphExecutor.swift:27,33).
 */
#ifndef PIPER_B200_H
#define PIPER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PIPER_B200_ABI_VERSION 2

/* status codes */
#define PIPER_B200_OK 0
#define PIPER_B200_ERR_INVALID_ARGUMENT 1  /* ~ ExecutionError.shapeMismatch / typeMismatch */
#define PIPER_B200_ERR_IO 2                /* file unreadable */
#define PIPER_B200_ERR_DECODE 3            /* ~ ONNXDecodingError (ONNXLoader.swift:3-21) */
#define PIPER_B200_ERR_UNSUPPORTED 4       /* ~ ExecutionError.unsupportedOp: graph is not a Piper VITS export */
#define PIPER_B200_ERR_CUDA 5              /* ~ ExecutionError.metalUnavailable / MetalBackend NSError */
#define PIPER_B200_ERR_STATE 6             /* call order (not finalized, ...) */
#define PIPER_B200_ERR_CAPACITY 7          /* durations exploded past the frame capacity */

/* ONNX TensorProto.data_type values accepted by piper_b200_set_weight
 * (the three dtypes of TensorValue.swift:4-8 plus int32, which widens to int64, :84-100) */
#define PIPER_B200_DTYPE_F32 1
#define PIPER_B200_DTYPE_I32 6
#define PIPER_B200_DTYPE_I64 7
#define PIPER_B200_DTYPE_BOOL 9

/* arithmetic of the flow + HiFi-GAN convolutions (text encoder, duration predictor and
 * length regulator always run in fp32 so that ceil() decisions match the fp32 oracle) */
#define PIPER_B200_PRECISION_FP32 0        /* CUDA-core fp32 everywhere (parity mode) */
#define PIPER_B200_PRECISION_BF16 1        /* tcgen05 kind::f16, bf16 operands, fp32 accumulate in TMEM */
#define PIPER_B200_PRECISION_F16 2         /* tcgen05 kind::f16, fp16 operands, fp32 accumulate in TMEM */

/* request flags */
#define PIPER_B200_FLAG_IDS_ON_DEVICE 1u   /* phoneme_ids is a device pointer (lengths, speaker_ids stay host) */
#define PIPER_B200_FLAG_KEEP_ON_DEVICE 2u  /* leave audio in HBM: result.audio is a device pointer */
#define PIPER_B200_FLAG_KEEP_STAGES 4u     /* retain stage tensors for piper_b200_get_stage */
#define PIPER_B200_FLAG_PCM16 8u           /* also produce int16 PCM (WavFileWriter.swift:20-30 rule) */
#define PIPER_B200_FLAG_STREAM 32u         /* batch == 1: run text encoder, durations and flow, keep the latent in HBM and return
                                            * without audio; the waveform is then decoded chunk by chunk with
                                            * piper_b200_stream_next (the incremental decode synthesizeStream intends,
                                            * PiperMetalRuntime.swift:82-121) */
#define PIPER_B200_FLAG_PCM16_ONLY 64u      /* int16 PCM is the ONLY host output (WavFileWriter.swift:20-30 rule): result.audio is NULL,
                                            * half the device->host bytes; a caller-owned output buffer then receives int16 */
#define PIPER_B200_FLAG_GENERIC 128u        /* run the graph node by node through the generic op executor (the reference's
                                            * GraphExecutor.executeNode, GraphExecutor.swift:591-2663) instead of the fused pipeline;
                                            * implied for a graph the fused pipeline does not recognise.  Route A only, float host
                                            * audio only; durations are not reported (read them with piper_b200_run_graph) */
#define PIPER_B200_FLAG_PROFILE 16u        /* CUDA events around every GEMM-class kernel launch (PIPER_PROFILE, GraphExecutor.swift:163-175) */

typedef struct piper_b200_ctx piper_b200_ctx;

/* Replaces PiperConfig (PiperConfig.swift:3-46) + what GraphExecutor learns from the graph. */
typedef struct piper_b200_voice_info {
  int32_t sample_rate;      /* audio.sample_rate */
  int32_t num_symbols;      /* rows of enc_p.emb.weight */
  int32_t num_speakers;     /* 1 for single-speaker voices */
  int32_t hidden;           /* 192 */
  int32_t hop;              /* samples per frame (256) */
  int32_t dec_channels0;    /* upsample_initial_channel (256 medium / 512 high) */
  int32_t resblock;         /* 1 or 2 */
  int32_t n_ups;
  float noise_scale;        /* inference.noise_scale  (PiperConfig.swift:16-20) */
  float length_scale;       /* inference.length_scale */
  float noise_w;            /* inference.noise_w */
  int64_t n_params;         /* scalar count over all bound initializers */
  double dec_flops_per_frame;   /* algorithmic FLOPs of the HiFi-GAN generator per frame */
  double flow_flops_per_frame;
  double enc_flops_per_id;      /* text encoder + duration predictor dense FLOPs per phoneme id */
} piper_b200_voice_info;

/* Replaces ExecutionInputs (GraphExecutor.swift:5-15), widened to a ragged batch.
 * Utterance b owns ids[id_off_b .. id_off_b+lengths[b]) with id_off the running sum. */
typedef struct piper_b200_request {
  const int64_t* phoneme_ids;   /* concatenated over the batch */
  const int64_t* lengths;       /* [batch]  (graph feed input_lengths) */
  const int64_t* speaker_ids;   /* [batch] or NULL (graph feed sid) */
  int32_t batch;
  uint32_t flags;
  float noise_scale;            /* scales[0]; NaN -> voice default (PiperMetalRuntime.swift:63-65) */
  float length_scale;           /* scales[1] */
  float noise_w;                /* scales[2] */
  /* Determinism hook = the reference's `overrides` of the two RandomNormalLike outputs
   * (GraphExecutor.swift:190-193, 2647-2651).  NULL -> generated on the device with the
   * reference's own generator (elementwise.metal:132-163) seeded with `seed`. */
  const float* noise_dp;        /* host, per utterance [2, lengths[b]] concatenated */
  const float* noise_main;      /* host, per utterance [hidden, noise_main_stride] */
  int64_t noise_main_stride;    /* frames available per utterance in noise_main */
  uint64_t seed;                /* Options.seed (PiperMetalRuntime.swift:12-15); only low 32 bits used */
} piper_b200_request;

/* Replaces the [Float] return value (PiperMetalRuntime.swift:73-79).  Pointers stay valid
 * until the next synthesize/destroy on the same context. */
typedef struct piper_b200_result {
  const float* audio;           /* concatenated utterances; pinned host memory (or device, see flags) */
  const int16_t* pcm16;         /* only with PIPER_B200_FLAG_PCM16, same offsets */
  const int64_t* sample_offsets;/* [batch+1] */
  const int64_t* frame_lengths; /* [batch] y_lengths of the length regulator */
  const int32_t* durations;     /* [sum lengths] ceil(w) per phoneme id, host */
  int64_t total_samples;
  int64_t total_frames;
} piper_b200_result;

/* Replaces GraphExecutor.RunTimings (GraphExecutor.swift:29-38). */
typedef struct piper_b200_timings {
  double wall_ms;               /* host wall clock of the call */
  double gpu_ms;                /* CUDA events around all device work of the call */
  double front_ms;              /* text encoder + duration predictor + length regulator */
  double flow_ms;
  double decoder_ms;
  double h2d_ms, d2h_ms;
  int64_t kernel_launches;      /* kernels launched by the call */
  int64_t h2d_bytes, d2h_bytes;
  double dec_flops, flow_flops; /* algorithmic FLOPs executed by the call */
  /* per kernel class, only with PIPER_B200_FLAG_PROFILE: 0 = conv_tc_kernel (tcgen05 implicit-GEMM conv),
   * 1 = mrf_tc_kernel (fused generator stage), 2 = conv_f32_kernel (CUDA-core conv) */
  double class_ms[3];           /* summed device time of the class's launches (CUDA events on the stream) */
  double class_flops[3];        /* algorithmic FLOPs of those launches */
  int64_t class_launches[3];
} piper_b200_timings;

int piper_b200_abi_version(void);
const char* piper_b200_last_error(const piper_b200_ctx* ctx);   /* ctx may be NULL */

/* MetalContext.init (MetalContext.swift:11-32): device + stream + kernel module */
int piper_b200_device_count(int* count);   /* the reference has one MTLDevice (MetalContext.swift:12); a B200 box has up to 8 */
int piper_b200_create(int device, piper_b200_ctx** out);
void piper_b200_destroy(piper_b200_ctx* ctx);

/* Route A - C++ twin of PiperMetalRuntime.init(modelURL:configURL:) (PiperMetalRuntime.swift:37-41):
 * parses the .onnx (ONNXLoader.swift:25-61) and .onnx.json (PiperConfig.swift:41-45), binds, finalizes. */
int piper_b200_load_voice(piper_b200_ctx* ctx, const char* model_path, const char* config_path);

/* Route B - a host that already parsed the model (Swift PiperONNX) pushes initializers
 * (ONNXIR.swift:23-93 tensors) and the Conv/ConvTranspose attributes, then finalizes. */
int piper_b200_set_weight(piper_b200_ctx* ctx, const char* name, int32_t dtype, const int64_t* dims,
                          int32_t rank, const void* data, size_t bytes);
int piper_b200_set_conv_attr(piper_b200_ctx* ctx, const char* weight_name, int32_t stride,
                             int32_t dilation, int32_t group, int32_t pad_begin, int32_t pad_end);
/* Route B, graph side: one call per node of ONNXGraph.nodes (ONNXIR.swift:68-84; only Conv, ConvTranspose, Gather, Sub,
 * Mul and Exp nodes are kept, the others are ignored, so a host may simply pass them all).  Real piper_train exports do not
 * name every weight after its module (weight-normed flow convolutions become `onnx::Conv_NNNN`, the embedding table of a
 * single-speaker voice is the initializer called `sid`, exp(-logs) of the duration predictor is folded): the binder
 * follows these nodes exactly like route A does, which is what the reference's node-by-node interpreter gets for free
 * (GraphExecutor.swift:227-264).  stride / dilation / group / pads = the Conv attributes (first spatial axis; <= 0 or
 * pad_begin < 0: ONNX default).  The float value of a Constant node is passed with piper_b200_set_weight under the
 * node's OUTPUT name. */
int piper_b200_add_node(piper_b200_ctx* ctx, const char* op_type, const char* name, const char* const* inputs,
                        int32_t n_inputs, const char* const* outputs, int32_t n_outputs, int32_t stride, int32_t dilation,
                        int32_t group, int32_t pad_begin, int32_t pad_end);
int piper_b200_set_config(piper_b200_ctx* ctx, int32_t sample_rate, int32_t num_speakers,
                          float noise_scale, float length_scale, float noise_w);
int piper_b200_finalize(piper_b200_ctx* ctx);   /* recognise the VITS graph, repack to kernel layouts */

int piper_b200_get_voice_info(const piper_b200_ctx* ctx, piper_b200_voice_info* out);
int piper_b200_set_precision(piper_b200_ctx* ctx, int32_t precision);
int piper_b200_set_stream(piper_b200_ctx* ctx, void* cuda_stream);   /* NULL -> the context's own stream */

/* Caller-owned host output (the reference returns a fresh [Float] the caller owns, PiperMetalRuntime.swift:73-79;
 * SURVEY 8b "written into a caller-provided buffer", 8e "shared pinned host arena"): the next synthesize calls write
 * the concatenated float32 audio (int16 with PIPER_B200_FLAG_PCM16_ONLY) to `host` instead of the library's own pinned
 * buffer and result.audio / result.pcm16 point into it.  The memory must be page-locked for the copies to be
 * asynchronous: allocate it with piper_b200_host_alloc or pin an existing mapping (e.g. a /dev/shm arena shared by the
 * ranks of one box) with piper_b200_host_register.  A result larger than `capacity_bytes` fails with
 * PIPER_B200_ERR_CAPACITY (the message states the bytes needed) before any audio is written.  host == NULL restores
 * the library-owned buffer.  Several contexts (one per GPU) may write disjoint ranges of one arena concurrently. */
int piper_b200_set_output_buffer(piper_b200_ctx* ctx, void* host, size_t capacity_bytes);
int piper_b200_host_alloc(size_t bytes, void** out);       /* page-locked, portable across the GPUs of the process */
void piper_b200_host_free(void* p);
int piper_b200_host_register(void* p, size_t bytes);       /* page-lock caller memory (cudaHostRegisterPortable) */
int piper_b200_host_unregister(void* p);

/* GraphExecutor.executeOutput (GraphExecutor.swift:156) for a ragged batch. */
int piper_b200_synthesize(piper_b200_ctx* ctx, const piper_b200_request* req, piper_b200_result* res);
int piper_b200_get_timings(const piper_b200_ctx* ctx, piper_b200_timings* out);

/* Incremental decoding after a synthesize with PIPER_B200_FLAG_STREAM: runs the HiFi-GAN generator over the
 * next `max_frames` frames (plus the generator's receptive-field halo, recomputed) and returns their samples:
 * res->audio (pinned host, total_samples values; frame n covers samples [n*hop, (n+1)*hop)), res->total_frames
 * = frames decoded by this call.  *is_final = 1 with the last chunk.  The concatenated chunks are bit-identical
 * to the one-shot synthesize of the same request.  Replaces the chunk loop of PiperMetalRuntime.synthesizeStream
 * (PiperMetalRuntime.swift:84-121, PiperAudioChunk PiperAudio.swift:13-25). */
int piper_b200_stream_next(piper_b200_ctx* ctx, int64_t max_frames, piper_b200_result* res, int32_t* is_final);

/* Parity hooks = GraphExecutor.execute(inputs:maxNodeIndex:overrides:) (GraphExecutor.swift:75-152),
 * which returns intermediates for node-by-node diffing.  After a synthesize with
 * PIPER_B200_FLAG_KEEP_STAGES, copies stage `name` of utterance `utt` to dst in the graph's
 * [channels, length] layout.  Names: "x","m_p","logs_p","logw","z_p","z".
 * dst==NULL queries the shape only. */
int piper_b200_get_stage(piper_b200_ctx* ctx, const char* name, int32_t utt, float* dst,
                         int64_t capacity, int64_t* channels, int64_t* length);

/* The integer-exact part on its own: Exp, Mul, Ceil, ReduceSum, Clip, Cast,
 * CumSum, Range/Less path -> frame-to-phoneme index.  logw is host memory, concatenated
 * over the batch.  frame_index is written per utterance at frame_capacity stride, -1 marks
 * frames no phoneme covers. */
int piper_b200_length_regulate(piper_b200_ctx* ctx, const float* logw, const int64_t* lengths,
                               int32_t batch, float length_scale, int32_t* durations,
                               int64_t* frame_lengths, int32_t* frame_index, int64_t frame_capacity);

/* The same durations / frame counts for callers that must not allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO counterpart in the reference.
 * Calls no cudaMalloc / cudaFree or allocating routine: every device buffer lives in one file-scope
 * `static __device__` array of PIPER_B200_NOALLOC_MAX_IDS / _MAX_BATCH capacity (about 0.8 MB, held by every GPU that
 * loads the library until the process exits); larger inputs return PIPER_B200_ERR_CAPACITY.  All calls on a GPU share
 * that array (it keeps the last call's contents; the call clears what it needs): calls that overlap on one GPU race,
 * the caller serialises them.  logw, lengths, durations, frame_lengths are host memory as in
 * piper_b200_length_regulate; no frame_index output. */
#define PIPER_B200_NOALLOC_MAX_IDS 65536    /* phoneme ids over the whole batch */
#define PIPER_B200_NOALLOC_MAX_BATCH 1024   /* utterances per call */
int piper_b200_length_regulate_noalloc(piper_b200_ctx* ctx, const float* logw, const int64_t* lengths,
                                       int32_t batch, float length_scale, int32_t* durations,
                                       int64_t* frame_lengths);

/* IPA text -> phoneme ids with the voice's phoneme_id_map: BOS "^", id + blank "_" per Unicode scalar,
 * EOS "$" (ESpeakPhonemizer.phonemeIDs, Sources/PiperCLI/ESpeakPhonemizer.swift:73-100; idsFromIPA,
 * PiperCLI.swift:115-136).  Host-only table lookup, needs a voice loaded from files (route A).
 * Writes at most `capacity` ids and always the full count to *n_ids; a count above capacity returns
 * PIPER_B200_ERR_CAPACITY, an unmapped symbol PIPER_B200_ERR_INVALID_ARGUMENT naming it. */
int piper_b200_ipa_to_ids(piper_b200_ctx* ctx, const char* ipa_utf8, int64_t* ids, int64_t capacity, int64_t* n_ids);
/* the same lookup straight from a voice's `*.onnx.json` (PiperConfig.load, PiperConfig.swift:41-45): no context, no GPU */
int piper_b200_config_ipa_to_ids(const char* config_path, const char* ipa_utf8, int64_t* ids, int64_t capacity,
                                 int64_t* n_ids);

/* RandomNormalLike with the reference's generator (elementwise.metal:132-163). */
int piper_b200_random_normal(piper_b200_ctx* ctx, uint64_t seed, int64_t count, float* dst_host);

/* Single-kernel self-test, the counterpart of PiperMetalMicrobench
 * (Sources/PiperMetal/Microbench/PiperMetalMicrobench.swift:24-73): one convolution of the given
 * shape on random data through the tcgen05 tensor-core kernel and through the fp32 CUDA-core
 * kernel on identical operands.  mode 0 = Conv (+bias, +residual), 1 = WN gate (cout = 2*hidden),
 * 2 = res|skip two-group GEMM, 3 = ConvTranspose (kernel k, stride).  precision = BF16 or F16.
 * out4 = {max |tc - fp32|, max |fp32|, tc kernel ms, fp32 kernel ms}. */
int piper_b200_selftest_conv(piper_b200_ctx* ctx, int32_t mode, int32_t cin, int32_t cout, int32_t k,
                             int32_t dilation, int32_t stride, int32_t rows, int32_t precision, float* out4);

/* The generic op executor, one utterance: runs every node of the loaded graph (route A) as plain CUDA kernels and copies
 * the value called `value_name` (NULL = the graph output) to `out` - the reference's
 * GraphExecutor.execute(inputs:maxNodeIndex:overrides:) / executeOutput (GraphExecutor.swift:75-152, 156-327) over
 * executeNode (:591-2663), 50 op types (csrc/graph_exec.cu).  scales3 = {noise_scale, length_scale, noise_w} (the
 * `scales` feed), speaker_id feeds `sid` when the graph has that input.  dtype = ONNX data_type of the value (1 float32,
 * 7 int64, 9 bool = one byte), dims8 / rank its shape; out may be NULL to query the shape only.  An op outside the set
 * returns PIPER_B200_ERR_UNSUPPORTED (ExecutionError.unsupportedOp). */
int piper_b200_run_graph(piper_b200_ctx* ctx, const int64_t* phoneme_ids, int64_t n_ids, int64_t speaker_id,
                         const float* scales3, uint64_t seed, const char* value_name, void* out, size_t out_capacity_bytes,
                         int32_t* dtype, int64_t* dims8, int32_t* rank);

/* Op-level known-answer hook: ONE Conv (transposed = 0; odd k, stride 1, "same" padding dilation*(k-1)/2) or
 * ConvTranspose (transposed = 1; k a multiple of stride, padding (k-stride)/2, rows*stride output rows) node on
 * caller-supplied host operands - the reference's conv1d_f32 / convtranspose1d_f32 kernels
 * (Sources/PiperMetal/Kernels/conv1d.metal:28-71, 97-142; dispatch GraphExecutor.swift:1739-1860).  x is
 * [rows][cin] (time-major), w the ONNX initializer layout (Conv [cout][cin][k], ConvTranspose [cin][cout][k]), bias
 * [cout] or NULL, out [rows*stride][cout].  precision FP32 runs the CUDA-core kernel, F16 / BF16 the tcgen05 kernel on
 * the 16-bit rounding of x and w with fp32 accumulation (cin % 32 == 0, cout % 16 == 0).  The parity tests compare it
 * with torch.nn.functional.conv1d / conv_transpose1d. */
int piper_b200_conv1d(piper_b200_ctx* ctx, int32_t transposed, const float* x, int64_t rows, const float* w,
                      const float* bias, int32_t cin, int32_t cout, int32_t k, int32_t dilation, int32_t stride,
                      int32_t precision, float* out);

#ifdef __cplusplus
}
#endif
#endif /* PIPER_B200_H */
