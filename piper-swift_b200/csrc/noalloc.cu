// piper_b200_length_regulate_noalloc: the integer-exact length regulator for callers that must not
// allocate - code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  It has NO counterpart
// in the reference.
//
// Same arithmetic as piper_b200_length_regulate (durations_kernel<true> in kernels_f32.cu: w = exp(logw) * length_scale,
// ceil, per-utterance inclusive scan, y_len = max(sum, 1)), but every device buffer it needs - the staged logw, the
// utterance layout, dur, cum, y_len - lives in ONE file-scope `static __device__` array of a fixed, modest size
// instead of the context's cudaMalloc'ed arena.  Every GPU that loads the library holds that array until the process
// exits.  The array is shared by every call on a device and keeps the last call's contents: the kernel writes every
// element it later reads back (nothing has to start at zero; y_len is cleared anyway), and two calls that overlap on
// the same GPU race - the caller serialises them.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/piper_b200.h"
#include "pipeline.h"

struct piper_b200_ctx {   // same definition as api.cpp (one member; kept in sync by tests/test_zz_no_alloc.py running both)
  piper::Context c;
};

namespace {

constexpr int NA_MAX_IDS = PIPER_B200_NOALLOC_MAX_IDS;     // phoneme ids over the whole batch
constexpr int NA_MAX_BATCH = PIPER_B200_NOALLOC_MAX_BATCH;
constexpr int NA_MAX_DEVICES = 64;

// layout of the scratch array (byte offsets, each a multiple of 256)
constexpr size_t NA_OFF_LOGW = 0;                                        // float [NA_MAX_IDS]
constexpr size_t NA_OFF_DUR = NA_OFF_LOGW + (size_t)NA_MAX_IDS * 4;      // int   [NA_MAX_IDS]
constexpr size_t NA_OFF_CUM = NA_OFF_DUR + (size_t)NA_MAX_IDS * 4;       // int   [NA_MAX_IDS]
constexpr size_t NA_OFF_START = NA_OFF_CUM + (size_t)NA_MAX_IDS * 4;     // int   [NA_MAX_BATCH]
constexpr size_t NA_OFF_LEN = NA_OFF_START + (size_t)NA_MAX_BATCH * 4;   // int   [NA_MAX_BATCH]
constexpr size_t NA_OFF_YLEN = NA_OFF_LEN + (size_t)NA_MAX_BATCH * 4;    // int64 [NA_MAX_BATCH]
constexpr size_t NA_BYTES = NA_OFF_YLEN + (size_t)NA_MAX_BATCH * 8;      // 802 816 bytes
static_assert(NA_OFF_DUR % 256 == 0 && NA_OFF_CUM % 256 == 0 && NA_OFF_START % 256 == 0 && NA_OFF_LEN % 256 == 0 &&
              NA_OFF_YLEN % 256 == 0, "scratch sections must stay 256-byte aligned");

// In HOST code this name is only a stand-in for the symbol, never a device address: the address comes from
// cudaGetSymbolAddress (scratch_for_device below), once per device, with that device current.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

std::mutex g_mu;
void* g_scratch_ptr[NA_MAX_DEVICES] = {};
thread_local std::string g_na_error;

// returns cudaSuccess and the device address of the array on `device` (which must be the current device)
cudaError_t scratch_for_device(int device, unsigned char** out) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_scratch_ptr[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_scratch_ptr[device] = p;
  }
  *out = static_cast<unsigned char*>(g_scratch_ptr[device]);
  return cudaSuccess;
}

__device__ __forceinline__ int na_dur_from_w(float w) {   // Ceil + Cast (elementwise.metal:224-232, cast.metal:8-29)
  const float c = ceilf(w);
  if (isnan(c)) return 0;
  if (c >= 1073741824.f) return 1 << 30;
  if (c <= 0.f) return 0;
  return (int)c;
}

// one CTA per utterance: durations, their inclusive scan and the frame count (variant of durations_kernel<true>)
__global__ void __launch_bounds__(256) noalloc_durations_kernel(const float* __restrict__ logw, float length_scale,
                                                                 const int* __restrict__ utt_start,
                                                                 const int* __restrict__ utt_len, int* __restrict__ dur,
                                                                 int* __restrict__ cum, long long* __restrict__ y_len) {
  __shared__ int warp_tot[8];
  __shared__ int carry_s;
  const int b = blockIdx.x;
  const int T = utt_len[b], r0 = utt_start[b];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int t0 = 0; t0 < T; t0 += 256) {
    const int t = t0 + threadIdx.x;
    int d = 0;
    if (t < T) {
      d = na_dur_from_w(expf(logw[r0 + t]) * length_scale);
      dur[r0 + t] = d;
    }
    int v = d;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += n;
    }
    if (lane == 31) warp_tot[warp] = v;
    __syncthreads();
    int pre = carry_s;
    for (int w = 0; w < warp; ++w) pre += warp_tot[w];
    if (t < T) cum[r0 + t] = min(pre + v, 1 << 30);
    __syncthreads();
    if (threadIdx.x == 255) carry_s = min(pre + v, 1 << 30);
    __syncthreads();
  }
  if (threadIdx.x == 0) y_len[b] = carry_s < 1 ? 1 : carry_s;
}

int fail(piper_b200_ctx* ctx, int code, const std::string& msg) {
  g_na_error = msg;
  if (ctx) ctx->c.last_error = msg;
  return code;
}

}  // namespace

extern "C" int piper_b200_length_regulate_noalloc(piper_b200_ctx* ctx, const float* logw, const int64_t* lengths,
                                                  int32_t batch, float length_scale, int32_t* durations,
                                                  int64_t* frame_lengths) {
  if (!ctx || !logw || !lengths || !durations || !frame_lengths || batch <= 0)
    return fail(ctx, PIPER_B200_ERR_INVALID_ARGUMENT, "invalid argument: null pointer or empty batch");
  if (batch > NA_MAX_BATCH)
    return fail(ctx, PIPER_B200_ERR_CAPACITY, "capacity: piper_b200_length_regulate_noalloc takes at most " +
                                                  std::to_string(NA_MAX_BATCH) + " utterances per call (got " +
                                                  std::to_string(batch) + "); use piper_b200_length_regulate");
  int us[NA_MAX_BATCH], ul[NA_MAX_BATCH];   // host stack: no heap allocation either
  int64_t n = 0;
  for (int b = 0; b < batch; ++b) {
    if (lengths[b] <= 0) return fail(ctx, PIPER_B200_ERR_INVALID_ARGUMENT, "invalid argument: non-positive length");
    if (lengths[b] > NA_MAX_IDS || n + lengths[b] > NA_MAX_IDS)
      return fail(ctx, PIPER_B200_ERR_CAPACITY, "capacity: piper_b200_length_regulate_noalloc takes at most " +
                                                    std::to_string(NA_MAX_IDS) +
                                                    " phoneme ids per call; use piper_b200_length_regulate");
    us[b] = (int)n;
    ul[b] = (int)lengths[b];
    n += lengths[b];
  }
  piper::Context& c = ctx->c;
  if (c.device < 0 || c.device >= NA_MAX_DEVICES)
    return fail(ctx, PIPER_B200_ERR_INVALID_ARGUMENT, "invalid argument: device index out of range");
  cudaError_t e = cudaSetDevice(c.device);
  unsigned char* s = nullptr;
  if (e == cudaSuccess) e = scratch_for_device(c.device, &s);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(ctx, PIPER_B200_ERR_CUDA, std::string("cuda: ") + cudaGetErrorString(e) + " (no-alloc scratch set-up)");
  }
  float* d_logw = reinterpret_cast<float*>(s + NA_OFF_LOGW);
  int* d_dur = reinterpret_cast<int*>(s + NA_OFF_DUR);
  int* d_cum = reinterpret_cast<int*>(s + NA_OFF_CUM);
  int* d_us = reinterpret_cast<int*>(s + NA_OFF_START);
  int* d_ul = reinterpret_cast<int*>(s + NA_OFF_LEN);
  long long* d_yl = reinterpret_cast<long long*>(s + NA_OFF_YLEN);
  cudaStream_t st = c.stream;
  // the caller's host memory may be pageable: these copies return once the source has been staged, so the stack
  // arrays above are safe to reuse; none of them allocates
  e = cudaMemcpyAsync(d_us, us, (size_t)batch * 4, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_ul, ul, (size_t)batch * 4, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_logw, logw, (size_t)n * 4, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_yl, 0, (size_t)batch * 8, st);   // the array keeps the last call's contents
  if (e == cudaSuccess) {
    noalloc_durations_kernel<<<batch, 256, 0, st>>>(d_logw, length_scale, d_us, d_ul, d_dur, d_cum, d_yl);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(durations, d_dur, (size_t)n * 4, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(frame_lengths, d_yl, (size_t)batch * 8, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(ctx, PIPER_B200_ERR_CUDA, std::string("cuda: ") + cudaGetErrorString(e));
  }
  return PIPER_B200_OK;
}
