#include "pipeline.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <stdexcept>

namespace piper {

#define CUDA_OK(x)                                                                         \
  do {                                                                                     \
    cudaError_t e_ = (x);                                                                  \
    if (e_ != cudaSuccess)                                                                 \
      throw std::runtime_error(std::string("cuda: ") + cudaGetErrorString(e_) + " at " #x); \
  } while (0)

void Arena::reserve(size_t bytes) {
  if (bytes <= cap) return;
  if (base) CUDA_OK(cudaFree(base));
  base = nullptr;
  cap = 0;
  const size_t want = bytes + bytes / 8 + (1 << 20);
  CUDA_OK(cudaMalloc(reinterpret_cast<void**>(&base), want));
  cap = want;
}
void Arena::release() {
  if (base) cudaFree(base);
  base = nullptr;
  cap = off = 0;
}
void PinnedBuf::reserve(size_t bytes) {
  if (bytes <= cap) return;
  if (p) CUDA_OK(cudaFreeHost(p));
  p = nullptr;
  cap = 0;
  const size_t want = bytes + bytes / 8 + 4096;
  CUDA_OK(cudaMallocHost(&p, want));
  cap = want;
}
void PinnedBuf::release() {
  if (p) cudaFreeHost(p);
  p = nullptr;
  cap = 0;
}

void Context::init(int dev) {
  device = dev;
  int n = 0;
  CUDA_OK(cudaGetDeviceCount(&n));
  if (dev < 0 || dev >= n) throw std::runtime_error("cuda: device " + std::to_string(dev) + " not present");
  CUDA_OK(cudaSetDevice(dev));
  CUDA_OK(cudaStreamCreateWithFlags(&own_stream, cudaStreamNonBlocking));
  stream = own_stream;
  for (auto& e : ev) CUDA_OK(cudaEventCreate(&e));
  int lo = 0, hi = 0;
  CUDA_OK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  CUDA_OK(cudaStreamCreateWithPriority(&copy_stream, cudaStreamNonBlocking, hi));   // its small gather blocks go ahead of the next range's CTAs
  for (auto& e : ev_chunk) CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  CUDA_OK(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
}

void Context::destroy() {
  cudaSetDevice(device);
  if (stream) cudaStreamSynchronize(stream);
  tc.release();
  prof.release();
  voice.release();
  front.release();
  back.release();
  h_in.release();
  h_meta.release();
  h_audio.release();
  h_pcm.release();
  for (auto& e : ev)
    if (e) cudaEventDestroy(e);
  for (auto& e : ev_chunk)
    if (e) { cudaEventDestroy(e); e = nullptr; }
  if (ev_join) cudaEventDestroy(ev_join);
  ev_join = nullptr;
  if (copy_stream) cudaStreamDestroy(copy_stream);
  if (own_stream) cudaStreamDestroy(own_stream);
  own_stream = stream = copy_stream = nullptr;
}

void Context::finalize_voice() {
  CUDA_OK(cudaSetDevice(device));
  voice.finalize();
  // gap rows between utterances must cover the widest receptive-field halo at each level
  int gp = (voice.hp.ffn_kernel - 1);
  for (auto& l : voice.dp_dds) gp = std::max(gp, l.dil * (l.k - 1) / 2);
  for (auto& f : voice.dp_flows)
    for (auto& l : f.dds) gp = std::max(gp, l.dil * (l.k - 1) / 2);
  gap_p = std::max(gp, 1);
  int gf = std::max((voice.dec_pre.k - 1) / 2, (voice.hp.flow_kernel - 1) / 2);
  int U = 1;
  for (auto& s : voice.stages) {
    gf = std::max(gf, (s.up.taps + U - 1) / U + 1);
    U *= s.stride;
    for (auto& rb : s.rbs) {
      for (auto& c : rb.c1) gf = std::max(gf, (c.dil * (c.k - 1) / 2 + U - 1) / U);
      for (auto& c : rb.c2) gf = std::max(gf, (c.dil * (c.k - 1) / 2 + U - 1) / U);
    }
  }
  gf = std::max(gf, ((voice.post_k - 1) / 2 + U - 1) / U);
  gap_f = std::max(gf, 1);
}

// The integer-exact part on its own (parity entry point).
int Context::length_regulate(const float* logw_h, const int64_t* lengths, int batch, float length_scale,
                             int32_t* durations_out, int64_t* frame_lengths_out, int32_t* frame_index,
                             int64_t frame_capacity) {
  if (batch <= 0 || !logw_h || !lengths) {
    last_error = "invalid argument: empty batch or null logw/lengths";
    return PIPER_B200_ERR_INVALID_ARGUMENT;
  }
  try {
    CUDA_OK(cudaSetDevice(device));
    invalidate_results();
    Exec ex;
    ex.st = stream;
    std::vector<int> us(batch), ul(batch);
    int64_t n = 0;
    for (int b = 0; b < batch; ++b) {
      if (lengths[b] <= 0) {
        last_error = "invalid argument: non-positive length";
        return PIPER_B200_ERR_INVALID_ARGUMENT;
      }
      us[b] = (int)n;
      ul[b] = (int)lengths[b];
      n += lengths[b];
    }
    const int R = (int)n;
    const int gap = 1;
    front.reserve((size_t)R * 16 + (size_t)batch * 64 + 16 * 256);
    front.reset();
    int* d_us = front.alloc<int>(batch);
    int* d_ul = front.alloc<int>(batch);
    float* d_logw = front.alloc<float>(R);
    int* d_dur = front.alloc<int>(R);
    int* d_cum = front.alloc<int>(R);
    int64_t* d_yl = front.alloc<int64_t>(batch);
    int64_t* d_fs = front.alloc<int64_t>(batch + 1);
    CUDA_OK(cudaMemcpyAsync(d_us, us.data(), (size_t)batch * 4, cudaMemcpyHostToDevice, stream));
    CUDA_OK(cudaMemcpyAsync(d_ul, ul.data(), (size_t)batch * 4, cudaMemcpyHostToDevice, stream));
    CUDA_OK(cudaMemcpyAsync(d_logw, logw_h, (size_t)R * 4, cudaMemcpyHostToDevice, stream));
    launch_durations_from_logw(ex, d_logw, length_scale, d_us, d_ul, batch, d_dur, d_cum, d_yl);
    launch_frame_layout(ex, d_yl, batch, gap, d_fs);
    std::vector<int64_t> yl(batch), fs(batch + 1);
    CUDA_OK(cudaMemcpyAsync(yl.data(), d_yl, (size_t)batch * 8, cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaMemcpyAsync(fs.data(), d_fs, (size_t)(batch + 1) * 8, cudaMemcpyDeviceToHost, stream));
    if (durations_out) CUDA_OK(cudaMemcpyAsync(durations_out, d_dur, (size_t)R * 4, cudaMemcpyDeviceToHost, stream));
    CUDA_OK(cudaStreamSynchronize(stream));
    if (frame_lengths_out) std::memcpy(frame_lengths_out, yl.data(), (size_t)batch * 8);
    if (frame_index) {
      const int64_t Rf = fs[batch];
      if (Rf > (int64_t)1 << 30) { last_error = "capacity: too many frames"; return PIPER_B200_ERR_CAPACITY; }
      for (int b = 0; b < batch; ++b)
        if (yl[b] > frame_capacity) {
          last_error = "capacity: frame_capacity " + std::to_string(frame_capacity) + " < " + std::to_string(yl[b]);
          return PIPER_B200_ERR_CAPACITY;
        }
      back.reserve((size_t)Rf * 8 + 1024);
      back.reset();
      int* d_fu = back.alloc<int>(Rf);
      int* d_fsrc = back.alloc<int>(Rf);
      launch_frame_map(ex, d_fs, d_yl, d_cum, d_us, d_ul, batch, (int)Rf, d_fu, d_fsrc);
      std::vector<int> src(Rf);
      CUDA_OK(cudaMemcpyAsync(src.data(), d_fsrc, (size_t)Rf * 4, cudaMemcpyDeviceToHost, stream));
      CUDA_OK(cudaStreamSynchronize(stream));
      for (int b = 0; b < batch; ++b)
        for (int64_t f = 0; f < yl[b]; ++f) {
          const int s = src[fs[b] + f];
          frame_index[(size_t)b * frame_capacity + f] = s < 0 ? -1 : s - us[b];   // index inside the utterance
        }
    }
    CUDA_OK(cudaGetLastError());
    return PIPER_B200_OK;
  } catch (const std::exception& e) {
    last_error = e.what();
    return PIPER_B200_ERR_CUDA;
  }
}

// Receptive half-width of the generator in frames: walk from the waveform back to the latent.  Every stage adds
// the deepest resblock chain ((k-1)/2 * dilation per convolution) at its own rate, a transposed convolution maps a
// need of n output rows to ceil(n / stride) + taps input rows; conv_pre and conv_post add (k-1)/2.
int decoder_halo_frames(const Voice& v) {
  long need = (v.post_k - 1) / 2;
  for (int si = (int)v.stages.size() - 1; si >= 0; --si) {
    const StageW& s = v.stages[(size_t)si];
    long rb = 0;
    for (const ResBlockW& r : s.rbs) {
      long h = 0;
      for (size_t m = 0; m < r.c1.size(); ++m) {
        h += (long)(r.c1[m].k - 1) / 2 * r.c1[m].dil;
        if (!r.c2.empty()) h += (long)(r.c2[m].k - 1) / 2 * r.c2[m].dil;
      }
      rb = std::max(rb, h);
    }
    need += rb;
    need = (need + s.stride - 1) / s.stride + s.up.taps;
  }
  need += (v.dec_pre.k - 1) / 2;
  return (int)need + 2;
}

}  // namespace piper
