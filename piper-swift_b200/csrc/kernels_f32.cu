// fp32 CUDA-core kernels of the Piper-VITS pipeline (sm_100a).
// These cover the stages whose results feed discrete decisions (text encoder, stochastic
// duration predictor, length regulator: SURVEY 8a4-a9) and serve as the parity-mode
// implementation of the flow and HiFi-GAN stages (8a10-a12); the tensor-core versions of
// the latter live in conv_tc.cu.  Each kernel names the ONNX ops / reference kernels it
// replaces; none of them is a translation of the one-thread-per-output Metal kernels.
#include <cuda_fp16.h>

#include "kernels.h"

#include <cstdlib>
#include <stdexcept>

#include <cfloat>
#include <cmath>
#include <cstdio>

namespace piper {

#define KLAUNCH(ex) (++(ex).launches)
// simple kernels: the next kernel may be staged at once, this one starts its work when its predecessor is complete
#define PDL_ENTER() do { PDL_TRIGGER(); PDL_WAIT(); } while (0)

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// f16x3 split of an fp32 value (conv_tc.h, TcW::split == 2): x = hi + lo' * 2^-11
__device__ __forceinline__ void f32_split2(float x, uint16_t& hi, uint16_t& lo) {
  x = fminf(fmaxf(x, -65504.f), 65504.f);
  const __half h = __float2half_rn(x);
  const __half l = __float2half_rn((x - __half2float(h)) * 2048.f);
  hi = __half_as_ushort(h); lo = __half_as_ushort(l);
}
__device__ __forceinline__ float gelu_erf(float x) {
  // exact-erf GELU (ONNX Erf; the ORT-CPU target, not elementwise.metal:292-312's approximation)
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

// ------------------------------------------------------------------------------------------
// Generic fp32 implicit-GEMM convolution (replaces conv1d_f32 / convtranspose1d_f32,
// conv1d.metal:28-71,97-142, and matmul_f32 for 1x1 convs).  64x64 output tile per CTA,
// K chunks of 16 input channels per tap staged through shared memory, 4x4 register
// micro-tile per thread.  Transposed convs run as `phases` polyphase sub-filters
// (blockIdx.z), which removes the `% stride` tap skipping of conv1d.metal:128-137.
// ------------------------------------------------------------------------------------------
constexpr int CBM = 64, CBN = 64, CBK = 16;

__global__ void __launch_bounds__(256) conv_f32_kernel(ConvArgs a) {
  PDL_ENTER();
  __shared__ __align__(16) float As[CBK][CBM];
  __shared__ __align__(16) float Bs[CBK][CBN];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int q0 = blockIdx.x * CBM;
  const int n0 = blockIdx.y * CBN;
  int in_off = a.in_off, out_add = a.out_add;
  const float* wbase = a.w;
  if (a.phases > 1) {
    const int ph = blockIdx.z;
    out_add = ph;
    in_off = (ph + a.ct_pad) / a.phases;
    wbase += (size_t)ph * a.taps * a.cin * a.ldw;
  }
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int arow = tid & 63, akq = tid >> 6;   // A loader: row, float4 index inside the K chunk
  const int bk = tid >> 4, bn4 = tid & 15;     // B loader
  for (int tap = 0; tap < a.taps; ++tap) {
    const int srow = q0 + arow + in_off + tap * a.tap_step;
    const bool row_ok = (q0 + arow) < a.Q && srow >= 0 && srow < a.R_in;
    const float* arow_ptr = a.in + (size_t)(row_ok ? srow : 0) * a.ldin;
    for (int ci0 = 0; ci0 < a.cin; ci0 += CBK) {
      float4 av = make_float4(0.f, 0.f, 0.f, 0.f);
      const int ci = ci0 + akq * 4;
      if (row_ok && ci < a.cin) av = *reinterpret_cast<const float4*>(arow_ptr + ci);
      if (a.in_act) {
        av.x = av.x >= 0.f ? av.x : av.x * a.in_slope;
        av.y = av.y >= 0.f ? av.y : av.y * a.in_slope;
        av.z = av.z >= 0.f ? av.z : av.z * a.in_slope;
        av.w = av.w >= 0.f ? av.w : av.w * a.in_slope;
      }
      float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ci0 + bk < a.cin)
        bv = *reinterpret_cast<const float4*>(wbase + ((size_t)tap * a.cin + ci0 + bk) * a.ldw + n0 + bn4 * 4);
      __syncthreads();
      As[akq * 4 + 0][arow] = av.x;
      As[akq * 4 + 1][arow] = av.y;
      As[akq * 4 + 2][arow] = av.z;
      As[akq * 4 + 3][arow] = av.w;
      *reinterpret_cast<float4*>(&Bs[bk][bn4 * 4]) = bv;
      __syncthreads();
#pragma unroll
      for (int k = 0; k < CBK; ++k) {
        const float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
        const float4 b4 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
        const float ar[4] = {a4.x, a4.y, a4.z, a4.w};
        const float br[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
      }
    }
  }
  // epilogue
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int q = q0 + ty * 4 + i;
    if (q >= a.Q) continue;
    const int orow = q * a.out_mul + out_add;
    int utt = 0;
    if (a.row_utt) utt = a.row_utt[orow / a.valid_div];
    const bool valid = utt >= 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = n0 + tx * 4 + j;
      if (co >= a.cout) continue;
      float v = acc[i][j];
      if (a.bias) v += a.bias[co];
      if (a.uttbias && valid) v += a.uttbias[(size_t)utt * a.ldub + co];
      if (a.relu) v = fmaxf(v, 0.f);
      if (a.res) v += a.res[(size_t)orow * a.ldres + co];
      float* o = a.out + (size_t)orow * a.ldout + co;
      if (a.accum) v += *o;
      if (a.div != 0.f) v = v / a.div;
      *o = valid ? v : 0.f;
    }
  }
}

// Large-problem variant: 128x128 output tile, 8x8 register micro-tile (as two 4-wide halves 64 apart so that
// every shared-memory read is a conflict-free 128-bit load), K chunks of 16 double-buffered through shared
// memory with register prefetch.  The K order (tap, channel) is the same as in conv_f32_kernel, so both
// kernels produce bit-identical fp32 results.
constexpr int LBM = 128, LBK = 16;

template <int BN>   // 128 or 64 output columns per CTA
__global__ void __launch_bounds__(256, 2) conv_f32_big_kernel(ConvArgs a) {
  PDL_ENTER();
  constexpr int NJ = BN / 16;                          // columns per thread (8 or 4)
  __shared__ __align__(16) float As[2][LBK][LBM];
  __shared__ __align__(16) float Bs[2][LBK][BN];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int q0 = blockIdx.x * LBM;
  const int n0 = blockIdx.y * BN;
  int in_off = a.in_off, out_add = a.out_add;
  const float* wbase = a.w;
  if (a.phases > 1) {
    const int ph = blockIdx.z;
    out_add = ph;
    in_off = (ph + a.ct_pad) / a.phases;
    wbase += (size_t)ph * a.taps * a.cin * a.ldw;
  }
  float acc[8][NJ];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < NJ; ++j) acc[i][j] = 0.f;

  // loaders: A = 128 rows x 16 channels (2 float4 per thread), B = 16 channels x BN columns
  const int arow = tid & 127, akq = tid >> 7;          // akq in {0,1}: float4 index 2*akq, 2*akq+1 of the chunk
  constexpr int BN4 = BN / 4;                          // float4 per B row
  constexpr int BH = BN == 128 ? 2 : 1;                // B float4 per thread
  const int bk = tid / BN4, bn4 = tid % BN4;
  const int kchunks = (a.cin + LBK - 1) / LBK;
  const int n_it = a.taps * kchunks;
  float4 pa[2], pb[BH];
  auto fetch = [&](int it) {
    const int tap = it / kchunks, ci0 = (it - tap * kchunks) * LBK;
    const int srow = q0 + arow + in_off + tap * a.tap_step;
    const bool row_ok = (q0 + arow) < a.Q && srow >= 0 && srow < a.R_in;
    const float* ap = a.in + (size_t)(row_ok ? srow : 0) * a.ldin;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int ci = ci0 + (akq * 2 + h) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (row_ok && ci < a.cin) v = *reinterpret_cast<const float4*>(ap + ci);
      if (a.in_act) {
        v.x = v.x >= 0.f ? v.x : v.x * a.in_slope;
        v.y = v.y >= 0.f ? v.y : v.y * a.in_slope;
        v.z = v.z >= 0.f ? v.z : v.z * a.in_slope;
        v.w = v.w >= 0.f ? v.w : v.w * a.in_slope;
      }
      pa[h] = v;
    }
#pragma unroll
    for (int h = 0; h < BH; ++h) {
      const int kk = bk + (256 / BN4) * h;
      float4 w4 = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ci0 + kk < a.cin && n0 + bn4 * 4 < a.ldw)
        w4 = *reinterpret_cast<const float4*>(wbase + ((size_t)tap * a.cin + ci0 + kk) * a.ldw + n0 + bn4 * 4);
      pb[h] = w4;
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int kq = (akq * 2 + h) * 4;
      As[buf][kq + 0][arow] = pa[h].x;
      As[buf][kq + 1][arow] = pa[h].y;
      As[buf][kq + 2][arow] = pa[h].z;
      As[buf][kq + 3][arow] = pa[h].w;
    }
#pragma unroll
    for (int h = 0; h < BH; ++h) *reinterpret_cast<float4*>(&Bs[buf][bk + (256 / BN4) * h][bn4 * 4]) = pb[h];
  };
  fetch(0);
  stash(0);
  __syncthreads();
  for (int it = 0; it < n_it; ++it) {
    const int buf = it & 1;
    if (it + 1 < n_it) fetch(it + 1);
#pragma unroll
    for (int k = 0; k < LBK; ++k) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][64 + ty * 4]);
      const float ar[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      float br[NJ];
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
      br[0] = b0.x; br[1] = b0.y; br[2] = b0.z; br[3] = b0.w;
      if (NJ == 8) {
        const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][k][(BN / 2) + tx * 4]);
        br[NJ - 4] = b1.x; br[NJ - 3] = b1.y; br[NJ - 2] = b1.z; br[NJ - 1] = b1.w;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) acc[i][j] = fmaf(ar[i], br[j], acc[i][j]);
    }
    if (it + 1 < n_it) {
      stash(buf ^ 1);
      __syncthreads();
    }
  }
  // epilogue (same semantics as conv_f32_kernel)
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int q = q0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (q >= a.Q) continue;
    const int orow = q * a.out_mul + out_add;
    int utt = 0;
    if (a.row_utt) utt = a.row_utt[orow / a.valid_div];
    const bool valid = utt >= 0;
#pragma unroll
    for (int j = 0; j < NJ; ++j) {
      const int co = n0 + (j < 4 ? tx * 4 + j : (BN / 2) + tx * 4 + (j - 4));
      if (co >= a.cout) continue;
      float v = acc[i][j];
      if (a.bias) v += a.bias[co];
      if (a.uttbias && valid) v += a.uttbias[(size_t)utt * a.ldub + co];
      if (a.relu) v = fmaxf(v, 0.f);
      if (a.res) v += a.res[(size_t)orow * a.ldres + co];
      float* o = a.out + (size_t)orow * a.ldout + co;
      if (a.accum) v += *o;
      if (a.div != 0.f) v = v / a.div;
      *o = valid ? v : 0.f;
    }
  }
}

// 96 x 192 tile for the text-encoder widths (cout = 192, 576, 768): with ~27.5k rows per step this gives
// 287 row tiles, i.e. ONE wave of 2 CTAs per SM on 148 SMs, and no half-empty column tile at cout = 192
// (the 128 x 128 grid needed 1.45 waves of 75 %-full tiles there).  256 threads, 6 x 12 outputs each;
// same tap-major / channel-ascending accumulation order as the other fp32 kernels.
constexpr int WBM = 96, WBN = 192, WBK = 16;

__global__ void __launch_bounds__(256, 2) conv_f32_w192_kernel(ConvArgs a) {
  PDL_ENTER();
  __shared__ __align__(16) float As[2][WBK][WBM];
  __shared__ __align__(16) float Bs[2][WBK][WBN];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int q0 = blockIdx.x * WBM;
  const int n0 = blockIdx.y * WBN;
  int in_off = a.in_off, out_add = a.out_add;
  const float* wbase = a.w;
  if (a.phases > 1) {
    const int ph = blockIdx.z;
    out_add = ph;
    in_off = (ph + a.ct_pad) / a.phases;
    wbase += (size_t)ph * a.taps * a.cin * a.ldw;
  }
  // accumulators are kept as f32x2 pairs (columns j, j+1): the inner product runs on FFMA2 (fma.rn.f32x2, one
  // instruction = two IEEE fmas with the row value broadcast), which halves the issue slots the FMAs need -
  // the scalar form was issue-bound at ~60 % FMA-pipe utilisation.  Same rounding as fmaf per element.
  unsigned long long acc2[6][6];
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 6; ++j) acc2[i][j] = 0ull;

  // loaders: A = 96 rows x 16 channels = 384 float4 (threads 0..127 take a second one), B = 16 x 192 = 768 float4
  const int kchunks = (a.cin + WBK - 1) / WBK;
  const int n_it = a.taps * kchunks;
  float4 pa[2], pb[3];
  auto fetch = [&](int it) {
    const int tap = it / kchunks, ci0 = (it - tap * kchunks) * WBK;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int f = tid + 256 * h;                      // float4 index: row f / 4, channel quad f % 4
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (f < WBM * 4) {
        const int row = f >> 2, ci = ci0 + (f & 3) * 4;
        const int srow = q0 + row + in_off + tap * a.tap_step;
        if ((q0 + row) < a.Q && srow >= 0 && srow < a.R_in && ci < a.cin)
          v = *reinterpret_cast<const float4*>(a.in + (size_t)srow * a.ldin + ci);
        if (a.in_act) {
          v.x = v.x >= 0.f ? v.x : v.x * a.in_slope;
          v.y = v.y >= 0.f ? v.y : v.y * a.in_slope;
          v.z = v.z >= 0.f ? v.z : v.z * a.in_slope;
          v.w = v.w >= 0.f ? v.w : v.w * a.in_slope;
        }
      }
      pa[h] = v;
    }
#pragma unroll
    for (int h = 0; h < 3; ++h) {
      const int f = tid + 256 * h;                      // float4 index: channel f / 48, column quad f % 48
      const int kk = f / 48, c4 = f - kk * 48;
      float4 w4 = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ci0 + kk < a.cin && n0 + c4 * 4 < a.ldw)
        w4 = *reinterpret_cast<const float4*>(wbase + ((size_t)tap * a.cin + ci0 + kk) * a.ldw + n0 + c4 * 4);
      pb[h] = w4;
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int f = tid + 256 * h;
      if (f < WBM * 4) {
        const int row = f >> 2, kq = (f & 3) * 4;
        As[buf][kq + 0][row] = pa[h].x;
        As[buf][kq + 1][row] = pa[h].y;
        As[buf][kq + 2][row] = pa[h].z;
        As[buf][kq + 3][row] = pa[h].w;
      }
    }
#pragma unroll
    for (int h = 0; h < 3; ++h) {
      const int f = tid + 256 * h;
      const int kk = f / 48, c4 = f - kk * 48;
      *reinterpret_cast<float4*>(&Bs[buf][kk][c4 * 4]) = pb[h];
    }
  };
  fetch(0);
  stash(0);
  __syncthreads();
  for (int it = 0; it < n_it; ++it) {
    const int buf = it & 1;
    if (it + 1 < n_it) fetch(it + 1);
#pragma unroll
    for (int k = 0; k < WBK; ++k) {
      float ar[6];
      unsigned long long br2[6];
#pragma unroll
      for (int g = 0; g < 3; ++g) {                     // rows g*32 + ty*2 + {0,1}; columns g*64 + tx*4 + {0..3}
        const float2 a2 = *reinterpret_cast<const float2*>(&As[buf][k][g * 32 + ty * 2]);
        ar[2 * g] = a2.x; ar[2 * g + 1] = a2.y;
        const ulonglong2 b4 = *reinterpret_cast<const ulonglong2*>(&Bs[buf][k][g * 64 + tx * 4]);
        br2[2 * g] = b4.x; br2[2 * g + 1] = b4.y;
      }
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        unsigned long long a2;
        asm("mov.b64 %0, {%1, %1};" : "=l"(a2) : "f"(ar[i]));
#pragma unroll
        for (int j = 0; j < 6; ++j) asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc2[i][j]) : "l"(a2), "l"(br2[j]));
      }
    }
    if (it + 1 < n_it) {
      stash(buf ^ 1);
      __syncthreads();
    }
  }
  // epilogue (same semantics as conv_f32_kernel)
  float acc[6][12];
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 6; ++j)
      asm("mov.b64 {%0, %1}, %2;" : "=f"(acc[i][2 * j]), "=f"(acc[i][2 * j + 1]) : "l"(acc2[i][j]));
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    const int q = q0 + (i >> 1) * 32 + ty * 2 + (i & 1);
    if (q >= a.Q) continue;
    const int orow = q * a.out_mul + out_add;
    int utt = 0;
    if (a.row_utt) utt = a.row_utt[orow / a.valid_div];
    const bool valid = utt >= 0;
#pragma unroll
    for (int j = 0; j < 12; ++j) {
      const int co = n0 + (j >> 2) * 64 + tx * 4 + (j & 3);
      if (co >= a.cout) continue;
      float v = acc[i][j];
      if (a.bias) v += a.bias[co];
      if (a.uttbias && valid) v += a.uttbias[(size_t)utt * a.ldub + co];
      if (a.relu) v = fmaxf(v, 0.f);
      if (a.res) v += a.res[(size_t)orow * a.ldres + co];
      float* o = a.out + (size_t)orow * a.ldout + co;
      if (a.accum) v += *o;
      if (a.div != 0.f) v = v / a.div;
      *o = valid ? v : 0.f;
    }
  }
}

void launch_conv_f32(Exec& ex, const ConvArgs& a) {
  if (a.Q <= 0) return;
  static const int w192_env = [] { const char* e = getenv("PIPER_B200_F32_W192"); return e ? atoi(e) : 1; }();
  if (w192_env && a.Q >= 4096 && a.cout % WBN == 0 && a.cin % 4 == 0 && a.ldw % 4 == 0) {
    dim3 grid((a.Q + WBM - 1) / WBM, a.cout / WBN, a.phases);
    ProfScope ps(ex, Prof::CONV_F32, 2.0 * a.Q * a.phases * a.taps * a.cin * a.cout);
    launch_pdl(conv_f32_w192_kernel, grid, dim3(256), 0, ex.st, a);
    KLAUNCH(ex);
    return;
  }
  if (a.Q >= 4096 && a.cout >= 96 && a.cin % 4 == 0) {
    const bool wide = a.cout > 64;   // measured: the 128-wide tile wins even for cout = 192 (half-empty second tile)
    const int bn = wide ? 128 : 64;
    dim3 grid((a.Q + LBM - 1) / LBM, (a.cout + bn - 1) / bn, a.phases);
    ProfScope ps(ex, Prof::CONV_F32, 2.0 * a.Q * a.phases * a.taps * a.cin * a.cout);
    if (wide) launch_pdl(conv_f32_big_kernel<128>, grid, dim3(256), 0, ex.st, a);
    else launch_pdl(conv_f32_big_kernel<64>, grid, dim3(256), 0, ex.st, a);
    KLAUNCH(ex);
    return;
  }
  dim3 grid((a.Q + CBM - 1) / CBM, (a.cout + CBN - 1) / CBN, a.phases);
  ProfScope ps(ex, Prof::CONV_F32, 2.0 * a.Q * a.phases * a.taps * a.cin * a.cout);
  launch_pdl(conv_f32_kernel, grid, dim3(256), 0, ex.st, a);
  KLAUNCH(ex);
}

// ------------------------------------------------------------------------------------------
// Channel LayerNorm (the decomposed ReduceMean/Sub/Pow/ReduceMean/Add/Sqrt/Div/Mul/Add chain,
// GraphExecutor.swift:2104-2125 -> reduce.metal:12-25), fused with the preceding residual
// add (x + y), optional exact GELU and the DDSConv residual.  One warp per row, two-pass
// mean / variance with warp shuffles, values kept in registers.
// ------------------------------------------------------------------------------------------
constexpr int LN_MAXV = 8;   // channels <= 256

__global__ void __launch_bounds__(256) layernorm_kernel(LNArgs a) {
  PDL_ENTER();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= a.R) return;
  const size_t base = (size_t)warp * a.C;
  const bool valid = !a.row_utt || a.row_utt[warp] >= 0;
  if (!valid) {
    for (int c = lane; c < a.C; c += 32) {
      a.out[base + c] = 0.f;
      if (a.out_split2) { a.out_split2[2 * base + c] = 0; a.out_split2[2 * base + a.C + c] = 0; }
    }
    return;
  }
  float v[LN_MAXV];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i) {
    const int c = lane + 32 * i;
    v[i] = 0.f;
    if (c < a.C) {
      v[i] = a.x[base + c];
      if (a.y) v[i] += a.y[base + c];
      s += v[i];
    }
  }
  const float mean = warp_sum(s) / (float)a.C;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < a.C) {
      v[i] -= mean;
      sq += v[i] * v[i];
    }
  }
  const float var = warp_sum(sq) / (float)a.C;
  const float denom = sqrtf(var + 1e-5f);
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < a.C) {
      float o = v[i] / denom * a.gamma[c] + a.beta[c];
      if (a.gelu) o = gelu_erf(o);
      if (a.res) o += a.res[base + c];
      a.out[base + c] = o;
      if (a.out_split2) {
        uint16_t hi, lo;
        f32_split2(o, hi, lo);
        a.out_split2[2 * base + c] = hi; a.out_split2[2 * base + a.C + c] = lo;
      }
    }
  }
}

void launch_layernorm(Exec& ex, const LNArgs& a) {
  if (a.R <= 0) return;
  const int warps_per_block = 8;
  launch_pdl(layernorm_kernel, dim3((a.R + warps_per_block - 1) / warps_per_block), dim3(256), 0, ex.st, a);
  KLAUNCH(ex);
}

// ------------------------------------------------------------------------------------------
// Embedding lookup * sqrt(H) (Gather axis 0 + Mul, GraphExecutor.swift:661-666;
// gather.metal:43-58).  ids_packed holds -1 on gap rows.  Out-of-range ids give a zero row,
// the reference's OOB rule.
// ------------------------------------------------------------------------------------------
__global__ void embed_kernel(const int64_t* __restrict__ ids, const float* __restrict__ emb,
                             float* __restrict__ x, int R, int H, int n_vocab, float scale) {
  PDL_ENTER();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)R * H) return;
  const int r = (int)(i / H), c = (int)(i % H);
  int64_t id = ids[r];
  float v = 0.f;
  if (id >= 0 && id < n_vocab) v = emb[id * H + c] * scale;
  x[i] = v;
}

void launch_embed(Exec& ex, const int64_t* ids_packed, const float* emb, float* x, int R, int H,
                  int n_vocab, float scale) {
  const int64_t n = (int64_t)R * H;
  if (n <= 0) return;
  launch_pdl(embed_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, ex.st, ids_packed, emb, x, R, H, n_vocab, scale);
  KLAUNCH(ex);
}

// ------------------------------------------------------------------------------------------
// Relative-position multi-head attention: replaces MatMul QK^T, the
// pad/reshape/slice "skew" of the relative logits, Softmax, MatMul PV and the relative-value
// term.  The relative embeddings are banded (|j-i| <= window), so they are added inside the
// key loop: O(T*9) instead of the exported O(T*(2T-1)) matmuls.  One warp per query row,
// keys streamed through shared memory 32 at a time, online softmax in fp32.
// ------------------------------------------------------------------------------------------
constexpr int ATT_WARPS = 8;
constexpr int ATT_KT = 32;
constexpr int ATT_MAXD = 4;   // head dim <= 128

__global__ void __launch_bounds__(ATT_WARPS * 32) attention_kernel(
    const float* __restrict__ qkv, const float* __restrict__ rel_k, const float* __restrict__ rel_v,
    float* __restrict__ out, const int* __restrict__ utt_start, const int* __restrict__ utt_len,
    int H, int n_heads, int window) {
  extern __shared__ float sm[];
  const int kc = H / n_heads;
  const int ldk = kc + 1;
  float* Ks = sm;                              // [ATT_KT][ldk]
  float* Vs = Ks + ATT_KT * ldk;               // [ATT_KT][ldk]
  float* Qs = Vs + ATT_KT * ldk;               // [ATT_WARPS][kc]
  float* RK = Qs + ATT_WARPS * kc;             // [2w+1][kc]
  float* RV = RK + (2 * window + 1) * kc;      // [2w+1][kc]
  float* QR = RV + (2 * window + 1) * kc;      // [ATT_WARPS][2w+1]
  const int b = blockIdx.z, head = blockIdx.y;
  const int T = utt_len[b];
  const int row0 = utt_start[b];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x * ATT_WARPS + warp;   // query index inside the utterance
  if (blockIdx.x * ATT_WARPS >= T) return;
  const int nrel = 2 * window + 1;
  const int ld = 3 * H;
  for (int t = threadIdx.x; t < nrel * kc; t += blockDim.x) {
    RK[t] = rel_k[t];
    RV[t] = rel_v[t];
  }
  const bool active = i < T;
  const float inv_sqrt = 1.0f / sqrtf((float)kc);
  if (active) {
    const float* q = qkv + (size_t)(row0 + i) * ld + head * kc;
    for (int d = lane; d < kc; d += 32) Qs[warp * kc + d] = q[d] / sqrtf((float)kc);
  }
  (void)inv_sqrt;
  __syncthreads();
  if (active) {
    for (int m = 0; m < nrel; ++m) {
      float s = 0.f;
      for (int d = lane; d < kc; d += 32) s += Qs[warp * kc + d] * RK[m * kc + d];
      s = warp_sum(s);
      if (lane == 0) QR[warp * nrel + m] = s;
    }
  }
  __syncwarp();
  float mrun = -FLT_MAX, lrun = 0.f;
  float acc[ATT_MAXD];
#pragma unroll
  for (int d = 0; d < ATT_MAXD; ++d) acc[d] = 0.f;
  for (int j0 = 0; j0 < T; j0 += ATT_KT) {
    __syncthreads();
    for (int t = threadIdx.x; t < ATT_KT * kc; t += blockDim.x) {
      const int jj = t / kc, d = t % kc;
      float kv = 0.f, vv = 0.f;
      if (j0 + jj < T) {
        const float* rowp = qkv + (size_t)(row0 + j0 + jj) * ld + head * kc + d;
        kv = rowp[H];
        vv = rowp[2 * H];
      }
      Ks[jj * ldk + d] = kv;
      Vs[jj * ldk + d] = vv;
    }
    __syncthreads();
    if (!active) continue;
    const int j = j0 + lane;
    float s = -FLT_MAX;
    if (j < T) {
      s = 0.f;
      const float* qv = Qs + warp * kc;
      const float* kr = Ks + lane * ldk;
      for (int d = 0; d < kc; ++d) s = fmaf(qv[d], kr[d], s);
      const int rel = j - i;
      if (rel >= -window && rel <= window) s += QR[warp * nrel + rel + window];
    }
    const float mnew = fmaxf(mrun, warp_max(s));
    const float p = (j < T) ? expf(s - mnew) : 0.f;
    const float corr = expf(mrun - mnew);
    lrun = lrun * corr + warp_sum(p);
#pragma unroll
    for (int d = 0; d < ATT_MAXD; ++d) acc[d] *= corr;
    mrun = mnew;
    const int jn = min(ATT_KT, T - j0);
    for (int jj = 0; jj < jn; ++jj) {
      const float pj = __shfl_sync(0xffffffffu, p, jj);
      const int rel = j0 + jj - i;
      const bool inwin = rel >= -window && rel <= window;
#pragma unroll
      for (int dd = 0; dd < ATT_MAXD; ++dd) {
        const int d = lane + 32 * dd;
        if (d < kc) {
          float vv = Vs[jj * ldk + d];
          if (inwin) vv += RV[(rel + window) * kc + d];
          acc[dd] = fmaf(pj, vv, acc[dd]);
        }
      }
    }
  }
  if (active) {
    float* o = out + (size_t)(row0 + i) * H + head * kc;
#pragma unroll
    for (int dd = 0; dd < ATT_MAXD; ++dd) {
      const int d = lane + 32 * dd;
      if (d < kc) o[d] = acc[dd] / lrun;
    }
  }
}

// Tiled variant for head dim 96 (every Piper voice: hidden 192, 2 heads): one CTA per (utterance, head,
// 64 queries), 16 x 16 threads.  Per 64-key tile: S = Q K^T as a register-tiled 64 x 64 x 96 product
// (thread (ty,tx) owns queries ty+16r and keys tx+16c), banded relative-key term from the precomputed
// q.rel_k table, online softmax with half-warp shuffles, P through shared memory, then O += P V with the
// thread owning the same 4 queries and columns tx+16e, plus the banded relative-value term.  K and V of an
// utterance are read ceil(T/64) times instead of ceil(T/8) times, and the inner loops are FMA-bound.
constexpr int AT2_Q = 64, AT2_K = 64, AT2_D = 96, AT2_LD = 100, AT2_LDP = 68, AT2_MAXREL = 17;

// S = Q K^T for the RM live query groups (rows ty + 16 r) and CM live key groups (keys tx + 16 c) of a tile: groups
// that lie wholly beyond the utterance are not computed (their scores are masked anyway).  The FMA order of every
// computed element is the same as in the full 4 x 4 form.
// Long utterances: the keys of one (utterance, head, query tile) are split over several CTAs so that a single
// 2000-id utterance fills the GPU (otherwise 64 CTAs); every CTA leaves its unnormalised output, running maximum and
// sum per query row in `part` and attention96_merge_kernel combines them in split order.  The split count depends
// on the utterance's own length only, so an utterance gives the same bits alone and inside a batch.
constexpr int AT2_PW = AT2_D + 2;    // floats per (row, head, split) in the scratch buffer
__host__ __device__ __forceinline__ int att96_splits(int T) { return T < 512 ? 1 : (T / 256 < 8 ? T / 256 : 8); }

template <int RM, int CM>
__device__ __forceinline__ void att96_qk(const float* __restrict__ Qs, const float* __restrict__ Ks, int ty, int tx,
                                         float (&sacc)[4][4]) {
#pragma unroll 2
  for (int d4 = 0; d4 < AT2_D / 4; ++d4) {
    float4 q4[RM], k4[CM];
#pragma unroll
    for (int r = 0; r < RM; ++r) q4[r] = *reinterpret_cast<const float4*>(Qs + (ty + 16 * r) * AT2_LD + d4 * 4);
#pragma unroll
    for (int c = 0; c < CM; ++c) k4[c] = *reinterpret_cast<const float4*>(Ks + (tx + 16 * c) * AT2_LD + d4 * 4);
#pragma unroll
    for (int r = 0; r < RM; ++r)
#pragma unroll
      for (int c = 0; c < CM; ++c) {
        sacc[r][c] = fmaf(q4[r].x, k4[c].x, sacc[r][c]);
        sacc[r][c] = fmaf(q4[r].y, k4[c].y, sacc[r][c]);
        sacc[r][c] = fmaf(q4[r].z, k4[c].z, sacc[r][c]);
        sacc[r][c] = fmaf(q4[r].w, k4[c].w, sacc[r][c]);
      }
  }
}
template <int RM>
__device__ __forceinline__ void att96_qk_c(const float* Qs, const float* Ks, int ty, int tx, int cmax, float (&sacc)[4][4]) {
  switch (cmax) {
    case 1: att96_qk<RM, 1>(Qs, Ks, ty, tx, sacc); break;
    case 2: att96_qk<RM, 2>(Qs, Ks, ty, tx, sacc); break;
    case 3: att96_qk<RM, 3>(Qs, Ks, ty, tx, sacc); break;
    default: att96_qk<RM, 4>(Qs, Ks, ty, tx, sacc); break;
  }
}
// O += P V for the RM live query groups
template <int RM>
__device__ __forceinline__ void att96_pv(const float* __restrict__ Ps, const float* __restrict__ Vs, int ty, int tx, int jn,
                                         float (&o)[4][6]) {
  for (int jj = 0; jj < jn; jj += 4) {           // rows jj..jj+3 beyond jn hold p = 0 and zero V
    float4 p4[RM];
#pragma unroll
    for (int r = 0; r < RM; ++r) p4[r] = *reinterpret_cast<const float4*>(Ps + (ty + 16 * r) * AT2_LDP + jj);
#pragma unroll
    for (int e = 0; e < 6; ++e) {
      const float v0 = Vs[(jj + 0) * AT2_D + tx + 16 * e], v1 = Vs[(jj + 1) * AT2_D + tx + 16 * e];
      const float v2 = Vs[(jj + 2) * AT2_D + tx + 16 * e], v3 = Vs[(jj + 3) * AT2_D + tx + 16 * e];
#pragma unroll
      for (int r = 0; r < RM; ++r) {
        o[r][e] = fmaf(p4[r].x, v0, o[r][e]);
        o[r][e] = fmaf(p4[r].y, v1, o[r][e]);
        o[r][e] = fmaf(p4[r].z, v2, o[r][e]);
        o[r][e] = fmaf(p4[r].w, v3, o[r][e]);
      }
    }
  }
}

__global__ void __launch_bounds__(256, 2) attention96_kernel(
    const float* __restrict__ qkv, const float* __restrict__ rel_k, const float* __restrict__ rel_v,
    float* __restrict__ out, const int* __restrict__ utt_start, const int* __restrict__ utt_len,
    int H, int window, uint16_t* __restrict__ out_split2, int smax, float* __restrict__ part, int skip_upto) {
  extern __shared__ __align__(16) float sm[];
  float* Qs = sm;                               // [64][100]  (scaled by 1/sqrt(96))
  float* Ks = Qs + AT2_Q * AT2_LD;              // [64][100]
  float* Vs = Ks + AT2_K * AT2_LD;              // [64][96]
  float* Ps = Vs + AT2_K * AT2_D;               // [64][68]
  float* RK = Ps + AT2_Q * AT2_LDP;             // [nrel][96]
  float* RV = RK + AT2_MAXREL * AT2_D;          // [nrel][96]
  float* QR = RV + AT2_MAXREL * AT2_D;          // [64][nrel]
  const int b = blockIdx.z, head = blockIdx.y;
  const int T = utt_len[b];
  if (T <= skip_upto) return;                   // short utterances were taken by attention_tc_kernel
  const int sp = (int)blockIdx.x % smax;        // key split handled by this CTA
  const int i0 = ((int)blockIdx.x / smax) * AT2_Q;
  const int S = smax > 1 ? att96_splits(T) : 1;   // smax == 1: no utterance of this call is long enough (or splitting is off)
  if (i0 >= T || sp >= S) return;
  const int n_kt = (T + AT2_K - 1) / AT2_K;
  const int j_beg = (sp * n_kt / S) * AT2_K, j_end = min(T, ((sp + 1) * n_kt / S) * AT2_K);
  const int row0 = utt_start[b];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int nrel = 2 * window + 1;
  const int ld = 3 * H;
  const float qscale = 1.0f / sqrtf((float)AT2_D);
  for (int t = tid; t < nrel * AT2_D; t += 256) {
    RK[t] = rel_k[t];
    RV[t] = rel_v[t];
  }
  for (int f = tid; f < AT2_Q * (AT2_D / 4); f += 256) {
    const int r = f / (AT2_D / 4), d4 = f - r * (AT2_D / 4);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i0 + r < T) {
      v = *reinterpret_cast<const float4*>(qkv + (size_t)(row0 + i0 + r) * ld + head * AT2_D + d4 * 4);
      v.x /= sqrtf((float)AT2_D); v.y /= sqrtf((float)AT2_D); v.z /= sqrtf((float)AT2_D); v.w /= sqrtf((float)AT2_D);
    }
    *reinterpret_cast<float4*>(Qs + r * AT2_LD + d4 * 4) = v;
  }
  (void)qscale;
  __syncthreads();
  for (int t = tid; t < AT2_Q * nrel; t += 256) {   // q_i . rel_k[m]
    const int r = t / nrel, m = t - r * nrel;
    float acc = 0.f;
    for (int d = 0; d < AT2_D; ++d) acc = fmaf(Qs[r * AT2_LD + d], RK[m * AT2_D + d], acc);
    QR[t] = acc;
  }
  const int rmax = min(4, (T - i0 + 15) >> 4);     // live query groups of this CTA
  float mrun[4], lrun[4], o[4][6];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    mrun[r] = -FLT_MAX; lrun[r] = 0.f;
#pragma unroll
    for (int e = 0; e < 6; ++e) o[r][e] = 0.f;
  }
  for (int j0 = j_beg; j0 < j_end; j0 += AT2_K) {
    __syncthreads();                             // previous tile's K, V, P fully consumed (and QR written)
    for (int f = tid; f < AT2_K * (AT2_D / 4); f += 256) {
      const int r = f / (AT2_D / 4), d4 = f - r * (AT2_D / 4);
      float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
      if (j0 + r < T) {
        const float* rowp = qkv + (size_t)(row0 + j0 + r) * ld + head * AT2_D + d4 * 4;
        kv = *reinterpret_cast<const float4*>(rowp + H);
        vv = *reinterpret_cast<const float4*>(rowp + 2 * H);
      }
      *reinterpret_cast<float4*>(Ks + r * AT2_LD + d4 * 4) = kv;
      *reinterpret_cast<float4*>(Vs + r * AT2_D + d4 * 4) = vv;
    }
    __syncthreads();
    // ---- S = Q K^T ----
    float sacc[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) sacc[r][c] = 0.f;
    const int cmax = min(4, (T - j0 + 15) >> 4);   // live key groups of this tile
    switch (rmax) {
      case 1: att96_qk_c<1>(Qs, Ks, ty, tx, cmax, sacc); break;
      case 2: att96_qk_c<2>(Qs, Ks, ty, tx, cmax, sacc); break;
      case 3: att96_qk_c<3>(Qs, Ks, ty, tx, cmax, sacc); break;
      default: att96_qk_c<4>(Qs, Ks, ty, tx, cmax, sacc); break;
    }
    // ---- relative-key band, mask, online softmax ----
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      if (r >= rmax) continue;                   // CTA-uniform: the whole query group lies beyond the utterance
      const int i = i0 + ty + 16 * r;
      float mx = -FLT_MAX;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int j = j0 + tx + 16 * c;
        const int rel = j - i;
        float v = sacc[r][c];
        if (rel >= -window && rel <= window) v += QR[(ty + 16 * r) * nrel + rel + window];
        v = j < T ? v : -FLT_MAX;
        sacc[r][c] = v;
        mx = fmaxf(mx, v);
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      const float mnew = fmaxf(mrun[r], mx);
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float pv = (j0 + tx + 16 * c) < T ? expf(sacc[r][c] - mnew) : 0.f;
        Ps[(ty + 16 * r) * AT2_LDP + tx + 16 * c] = pv;
        sum += pv;
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
      const float corr = expf(mrun[r] - mnew);
      lrun[r] = lrun[r] * corr + sum;
      mrun[r] = mnew;
#pragma unroll
      for (int e = 0; e < 6; ++e) o[r][e] *= corr;
    }
    __syncthreads();
    // ---- O += P V ----
    const int jn = min(AT2_K, T - j0);
    switch (rmax) {
      case 1: att96_pv<1>(Ps, Vs, ty, tx, jn, o); break;
      case 2: att96_pv<2>(Ps, Vs, ty, tx, jn, o); break;
      case 3: att96_pv<3>(Ps, Vs, ty, tx, jn, o); break;
      default: att96_pv<4>(Ps, Vs, ty, tx, jn, o); break;
    }
    // ---- relative-value band ----
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      if (r >= rmax) continue;
      const int i = i0 + ty + 16 * r;
      for (int m = 0; m < nrel; ++m) {
        const int jl = i + m - window - j0;
        if (jl < 0 || jl >= jn) continue;
        const float pv = Ps[(ty + 16 * r) * AT2_LDP + jl];
#pragma unroll
        for (int e = 0; e < 6; ++e) o[r][e] = fmaf(pv, RV[m * AT2_D + tx + 16 * e], o[r][e]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int i = i0 + ty + 16 * r;
    if (i >= T) continue;
    if (S > 1) {        // partial result of this key split
      float* pp = part + (((size_t)(row0 + i) * gridDim.y + head) * smax + sp) * AT2_PW;
#pragma unroll
      for (int e = 0; e < 6; ++e) pp[tx + 16 * e] = o[r][e];
      if (tx == 0) { pp[AT2_D] = mrun[r]; pp[AT2_D + 1] = lrun[r]; }
      continue;
    }
    if (out_split2) {   // f16x3 split operand of the output projection instead of the fp32 tensor
      uint16_t* sp = out_split2 + (size_t)(row0 + i) * 2 * H + head * AT2_D;
#pragma unroll
      for (int e = 0; e < 6; ++e) {
        uint16_t hi, lo;
        f32_split2(o[r][e] / lrun[r], hi, lo);
        sp[tx + 16 * e] = hi; sp[H + tx + 16 * e] = lo;
      }
      continue;
    }
    float* op = out + (size_t)(row0 + i) * H + head * AT2_D;
#pragma unroll
    for (int e = 0; e < 6; ++e) op[tx + 16 * e] = o[r][e] / lrun[r];
  }
}

// one warp per (query row, head) of the utterances that were split
__global__ void __launch_bounds__(256) attention96_merge_kernel(const float* __restrict__ part, float* __restrict__ out,
                                                                uint16_t* __restrict__ out_split2,
                                                                const int* __restrict__ utt_start,
                                                                const int* __restrict__ utt_len, int H, int smax) {
  const int b = blockIdx.z, head = blockIdx.y;
  const int T = utt_len[b];
  const int i = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  const int S = att96_splits(T);
  if (i >= T || S == 1) return;
  const size_t row = (size_t)utt_start[b] + i;
  const float* pp = part + ((row * gridDim.y + head) * smax) * AT2_PW;
  float M = -FLT_MAX;
  for (int s = 0; s < S; ++s) M = fmaxf(M, pp[s * AT2_PW + AT2_D]);
  float L = 0.f, acc[3] = {0.f, 0.f, 0.f};
  for (int s = 0; s < S; ++s) {
    const float w = expf(pp[s * AT2_PW + AT2_D] - M);
    L += pp[s * AT2_PW + AT2_D + 1] * w;
#pragma unroll
    for (int c = 0; c < 3; ++c) acc[c] = fmaf(pp[s * AT2_PW + lane + 32 * c], w, acc[c]);
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float v = acc[c] / L;
    const int col = head * AT2_D + lane + 32 * c;
    if (out_split2) {
      uint16_t hi, lo;
      f32_split2(v, hi, lo);
      out_split2[row * 2 * H + col] = hi; out_split2[row * 2 * H + H + col] = lo;
    } else {
      out[row * H + col] = v;
    }
  }
}

size_t attention_scratch_bytes(int rows, int max_len, int H, int n_heads) {
  const int smax = (H / n_heads == AT2_D) ? att96_splits(max_len) : 1;
  return smax > 1 ? (size_t)rows * n_heads * smax * AT2_PW * sizeof(float) : 0;
}

void launch_attention(Exec& ex, const float* qkv, const float* rel_k, const float* rel_v, float* out,
                      const int* utt_start, const int* utt_len, int batch, int max_len, int H,
                      int n_heads, int window, uint16_t* out_split2, bool* wrote_split, float* scratch, int skip_upto) {
  if (wrote_split) *wrote_split = false;
  if (batch <= 0 || max_len <= 0) return;
  const int kc = H / n_heads;
  const int nrel = 2 * window + 1;
  const size_t smem = sizeof(float) * (2 * ATT_KT * (kc + 1) + ATT_WARPS * kc + 2 * nrel * kc + ATT_WARPS * nrel);
  static const int at2_env = [] { const char* e = getenv("PIPER_B200_ATT96"); return e ? atoi(e) : 1; }();
  if (at2_env && kc == AT2_D && nrel <= AT2_MAXREL && H % 4 == 0) {
    const size_t smem2 = sizeof(float) * (AT2_Q * AT2_LD + AT2_K * AT2_LD + AT2_K * AT2_D + AT2_Q * AT2_LDP +
                                          2 * AT2_MAXREL * AT2_D + AT2_Q * AT2_MAXREL);
    static PerDeviceOnce once;
    once.run([&] {
      if (cudaFuncSetAttribute(attention96_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2) != cudaSuccess)
        throw std::runtime_error("cuda: cannot raise dynamic shared memory limit for attention96_kernel");
    });
    static const int split_env = [] { const char* e = getenv("PIPER_B200_ATT_SPLIT"); return e ? atoi(e) : 1; }();
    // key splits for long utterances (att96_splits): only with a scratch buffer; the split count per utterance is a
    // function of its own length, the grid covers the longest one
    const int smax = (scratch && split_env) ? att96_splits(max_len) : 1;
    if (smax == 1 && att96_splits(max_len) > 1 && split_env)
      throw std::runtime_error("attention: long utterances need the split scratch buffer");
    dim3 grid2((max_len + AT2_Q - 1) / AT2_Q * smax, n_heads, batch);
    attention96_kernel<<<grid2, 256, smem2, ex.st>>>(qkv, rel_k, rel_v, out, utt_start, utt_len, H, window, out_split2,
                                                     smax, scratch, skip_upto);
    KLAUNCH(ex);
    if (smax > 1) {
      dim3 gm((max_len + 7) / 8, n_heads, batch);
      attention96_merge_kernel<<<gm, 256, 0, ex.st>>>(scratch, out, out_split2, utt_start, utt_len, H, smax);
      KLAUNCH(ex);
    }
    if (wrote_split) *wrote_split = out_split2 != nullptr;
    return;
  }
  dim3 grid((max_len + ATT_WARPS - 1) / ATT_WARPS, n_heads, batch);
  attention_kernel<<<grid, ATT_WARPS * 32, smem, ex.st>>>(qkv, rel_k, rel_v, out, utt_start, utt_len, H,
                                                         n_heads, window);
  KLAUNCH(ex);
}

// ------------------------------------------------------------------------------------------
// DDSConv first half: depthwise Conv (groups=C, dilation d) + LayerNorm + GELU.
// w_sep is tap-major [k][C] so that a warp reads consecutive channels.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) dds_sep_ln_gelu_kernel(
    const float* __restrict__ x, const float* __restrict__ w_sep, const float* __restrict__ b_sep,
    const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ out,
    const int* __restrict__ row_utt, int R, int C, int k, int dil, uint16_t* __restrict__ out_split2) {
  PDL_ENTER();
  const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (r >= R) return;
  const size_t base = (size_t)r * C;
  if (row_utt && row_utt[r] < 0) {
    for (int c = lane; c < C; c += 32) {
      if (out_split2) { out_split2[2 * base + c] = 0; out_split2[2 * base + C + c] = 0; }
      else out[base + c] = 0.f;
    }
    return;
  }
  const int pad = (k * dil - dil) / 2;
  float v[LN_MAXV];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i) {
    const int c = lane + 32 * i;
    v[i] = 0.f;
    if (c < C) {
      float a = b_sep[c];
      for (int t = 0; t < k; ++t) {
        const int rr = r + t * dil - pad;
        if (rr >= 0 && rr < R) a = fmaf(w_sep[t * C + c], x[(size_t)rr * C + c], a);
      }
      v[i] = a;
      s += a;
    }
  }
  const float mean = warp_sum(s) / (float)C;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i)
    if (lane + 32 * i < C) {
      v[i] -= mean;
      sq += v[i] * v[i];
    }
  const float denom = sqrtf(warp_sum(sq) / (float)C + 1e-5f);
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i) {
    const int c = lane + 32 * i;
    if (c < C) {
      const float o = gelu_erf(v[i] / denom * gamma[c] + beta[c]);
      if (out_split2) {
        uint16_t hi, lo;
        f32_split2(o, hi, lo);
        out_split2[2 * base + c] = hi; out_split2[2 * base + C + c] = lo;
      } else {
        out[base + c] = o;
      }
    }
  }
}

void launch_dds_sep_ln_gelu(Exec& ex, const float* x, const float* w_sep, const float* b_sep,
                            const float* gamma, const float* beta, float* out, const int* row_utt,
                            int R, int C, int k, int dil, uint16_t* out_split2) {
  if (R <= 0) return;
  launch_pdl(dds_sep_ln_gelu_kernel, dim3((R + 7) / 8), dim3(256), 0, ex.st, x, w_sep, b_sep, gamma, beta, out, row_utt, R, C, k, dil,
             out_split2);
  KLAUNCH(ex);
}

// ConvFlow.pre (1 -> C channels, k=1) fused with DDSConv's `x + g` conditioning add.
__global__ void convflow_pre_kernel(const float* __restrict__ z, const float* __restrict__ w,
                                    const float* __restrict__ b, const float* __restrict__ cond,
                                    float* __restrict__ h, const int* __restrict__ row_utt, int R, int C) {
  PDL_ENTER();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)R * C) return;
  const int r = (int)(i / C), c = (int)(i % C);
  float v = 0.f;
  if (!row_utt || row_utt[r] >= 0) v = fmaf(w[c], z[r], b[c]) + cond[i];
  h[i] = v;
}

void launch_convflow_pre(Exec& ex, const float* z, const float* w, const float* b, const float* cond,
                         float* h, const int* row_utt, int R, int C) {
  const int64_t n = (int64_t)R * C;
  if (n <= 0) return;
  launch_pdl(convflow_pre_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, ex.st, z, w, b, cond, h, row_utt, R, C);
  KLAUNCH(ex);
}

// ------------------------------------------------------------------------------------------
// Inverse rational-quadratic spline with linear tails: replaces the
// Softmax/Softplus/CumSum/Pad/GreaterOrEqual+ReduceSum (searchsorted)/GatherElements/
// Sqrt/Div chain and the NonZero/GatherND/ScatterND masked assignment
// (GraphExecutor.swift:2379-2602; nonzero.metal's atomic compaction is not needed because
// every element is evaluated in place).  One thread per (utterance position); op order
// follows the oracle so that the discrete bin choice matches.
// ------------------------------------------------------------------------------------------
constexpr int SPL_MAXB = 16;

__global__ void spline_inverse_kernel(float* __restrict__ zt, const float* __restrict__ params, int ldp,
                                      const int* __restrict__ row_utt, int R, int nb, float tb,
                                      float inv_sqrt_fc) {
  PDL_ENTER();
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  if (row_utt && row_utt[r] < 0) {
    zt[r] = 0.f;
    return;
  }
  const float x = zt[r];
  if (!(x >= -tb && x <= tb)) return;   // linear tails: identity
  const float* p = params + (size_t)r * ldp;
  const float min_w = 1e-3f, min_h = 1e-3f, min_d = 1e-3f;
  float cw[SPL_MAXB + 1], chh[SPL_MAXB + 1];
  // widths
  {
    float u[SPL_MAXB];
    float mx = -FLT_MAX;
    for (int i = 0; i < nb; ++i) {
      u[i] = p[i] * inv_sqrt_fc;
      mx = fmaxf(mx, u[i]);
    }
    float sum = 0.f;
    for (int i = 0; i < nb; ++i) {
      u[i] = expf(u[i] - mx);
      sum += u[i];
    }
    float run = 0.f;
    cw[0] = -tb;
    for (int i = 0; i < nb; ++i) {
      const float wd = min_w + (1.f - min_w * nb) * (u[i] / sum);
      run += wd;
      cw[i + 1] = (2.f * tb) * run + (-tb);
    }
    cw[nb] = tb;
  }
  {
    float u[SPL_MAXB];
    float mx = -FLT_MAX;
    for (int i = 0; i < nb; ++i) {
      u[i] = p[nb + i] * inv_sqrt_fc;
      mx = fmaxf(mx, u[i]);
    }
    float sum = 0.f;
    for (int i = 0; i < nb; ++i) {
      u[i] = expf(u[i] - mx);
      sum += u[i];
    }
    float run = 0.f;
    chh[0] = -tb;
    for (int i = 0; i < nb; ++i) {
      const float hd = min_h + (1.f - min_h * nb) * (u[i] / sum);
      run += hd;
      chh[i + 1] = (2.f * tb) * run + (-tb);
    }
    chh[nb] = tb;
  }
  // searchsorted over cumheights (last edge + eps)
  int bin = -1;
  for (int i = 0; i <= nb; ++i) {
    const float edge = (i == nb) ? chh[i] + 1e-6f : chh[i];
    bin += (x >= edge) ? 1 : 0;
  }
  bin = min(max(bin, 0), nb - 1);
  const float in_cw = cw[bin], in_w = cw[bin + 1] - cw[bin];
  const float in_ch = chh[bin], in_h = chh[bin + 1] - chh[bin];
  const float delta = in_h / in_w;
  // derivatives: padded with the constant so that the boundary derivative is 1
  const float cst = logf(expf(1.f - min_d) - 1.f);
  auto deriv = [&](int i) {
    const float ud = (i == 0 || i == nb) ? cst : p[2 * nb + i - 1];
    const float sp = (ud > 20.f) ? ud : log1pf(expf(ud));   // torch softplus (threshold 20)
    return min_d + sp;
  };
  const float d0 = deriv(bin), d1 = deriv(bin + 1);
  const float t = x - in_ch;
  const float s2 = d0 + d1 - 2.f * delta;
  const float a = t * s2 + in_h * (delta - d0);
  const float b = in_h * d0 - t * s2;
  const float c = -delta * t;
  const float disc = b * b - 4.f * a * c;
  const float root = (2.f * c) / (-b - sqrtf(disc));
  zt[r] = root * in_w + in_cw;
}

void launch_spline_inverse(Exec& ex, float* z_target, const float* params, int ldp, const int* row_utt,
                           int R, int bins, float tail_bound, float inv_sqrt_filter) {
  if (R <= 0) return;
  launch_pdl(spline_inverse_kernel, dim3((R + 127) / 128), dim3(128), 0, ex.st, z_target, params, ldp, row_utt, R, bins, tail_bound,
             inv_sqrt_filter);
  KLAUNCH(ex);
}

// ------------------------------------------------------------------------------------------
// RandomNormalLike with the reference's generator (elementwise.metal:132-163): xorshift32 x2
// seeded by seedLo ^ (gid*747796405+2891336453), Box-Muller r*cos(theta).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t xorshift32(uint32_t x) {
  x ^= x << 13;
  x ^= x >> 17;
  x ^= x << 5;
  return x;
}
__device__ __forceinline__ float ref_normal(uint32_t seed, uint32_t gid) {
  uint32_t state = seed ^ (gid * 747796405u + 2891336453u);
  state = xorshift32(state);
  const uint32_t u0i = state;
  state = xorshift32(state);
  const uint32_t u1i = state;
  const float u0 = ((float)u0i + 1.0f) / 4294967296.0f;
  const float u1 = ((float)u1i + 1.0f) / 4294967296.0f;
  const float r = sqrtf(-2.0f * logf(u0));
  const float theta = 6.28318530718f * u1;
  return r * cosf(theta);
}

__global__ void random_normal_kernel(float* out, int64_t count, uint32_t seed) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) out[i] = ref_normal(seed, (uint32_t)i);
}
void launch_random_normal(Exec& ex, float* out, int64_t count, uint32_t seed) {
  if (count <= 0) return;
  random_normal_kernel<<<(unsigned)((count + 255) / 256), 256, 0, ex.st>>>(out, count, seed);
  KLAUNCH(ex);
}

__global__ void fill_i32_kernel(int* p, int v, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}
void launch_fill_i32(Exec& ex, int* p, int v, int64_t n) {
  if (n <= 0) return;
  fill_i32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ex.st>>>(p, v, n);
  KLAUNCH(ex);
}

// z[ch][r] = noise * noise_w for the duration predictor ([1,2,T] per utterance, gid = ch*T+t)
__global__ void noise_dp_kernel(float* __restrict__ z, const float* __restrict__ ext,
                                const int* __restrict__ utt_start, const int* __restrict__ utt_len,
                                const int64_t* __restrict__ ext_off, int R, uint32_t seed, float noise_w) {
  const int b = blockIdx.y;
  const int T = utt_len[b], r0 = utt_start[b];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 2 * T; i += gridDim.x * blockDim.x) {
    const int ch = i / T, t = i % T;
    const float n = ext ? ext[ext_off[b] * 2 + i] : ref_normal(seed, (uint32_t)i);
    z[(size_t)ch * R + r0 + t] = n * noise_w;
  }
}
void launch_noise_dp(Exec& ex, float* z, const float* ext, const int* utt_start, const int* utt_len,
                     const int64_t* ext_off, int batch, int R, uint32_t seed, float noise_w) {
  if (batch <= 0) return;
  noise_dp_kernel<<<dim3(4, batch), 256, 0, ex.st>>>(z, ext, utt_start, utt_len, ext_off, R, seed, noise_w);
  KLAUNCH(ex);
}

// ------------------------------------------------------------------------------------------
// Length regulator, integer part: Exp, Mul(length_scale), Ceil, ReduceSum,
// Clip(min 1), Cast f32->i64 (cast.metal:8-29: NaN -> 0, saturate, trunc), CumSum
// (cumsum.metal's O(n^2) loop becomes a block scan).  One CTA per utterance.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int dur_from_w(float w) {
  const float c = ceilf(w);
  if (isnan(c)) return 0;
  if (c >= 1073741824.f) return 1 << 30;   // saturate (capacity check rejects this later)
  if (c <= 0.f) return 0;
  return (int)c;
}

template <bool FROM_LOGW>
__global__ void __launch_bounds__(256) durations_kernel(const float* __restrict__ src, float ea_m,
                                                        float ea_logs, float ea_scale, int use_scale,
                                                        float length_scale,
                                                        const int* __restrict__ utt_start,
                                                        const int* __restrict__ utt_len,
                                                        float* __restrict__ logw_out, int* __restrict__ dur,
                                                        int* __restrict__ cum, int64_t* __restrict__ y_len) {
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
      float lw = src[r0 + t];
      if (!FROM_LOGW) {
        // ElementwiseAffine reverse, channel 0; use_scale: the export folded exp(-logs) into a constant
        lw = (lw - ea_m) * (use_scale ? ea_scale : expf(-ea_logs));
        if (logw_out) logw_out[r0 + t] = lw;
      }
      const float w = expf(lw) * length_scale;   // x_mask == 1 inside the utterance
      d = dur_from_w(w);
      dur[r0 + t] = d;
    }
    int v = d;   // inclusive warp scan
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

void launch_durations(Exec& ex, const float* z0, float ea_m, float ea_logs, float ea_scale, bool use_scale,
                      float length_scale, const int* utt_start, const int* utt_len, int batch, float* logw_out, int* dur,
                      int* cum, int64_t* y_len) {
  if (batch <= 0) return;
  durations_kernel<false><<<batch, 256, 0, ex.st>>>(z0, ea_m, ea_logs, ea_scale, use_scale ? 1 : 0, length_scale, utt_start, utt_len,
                                                    logw_out, dur, cum, y_len);
  KLAUNCH(ex);
}
void launch_durations_from_logw(Exec& ex, const float* logw, float length_scale, const int* utt_start,
                                const int* utt_len, int batch, int* dur, int* cum, int64_t* y_len) {
  if (batch <= 0) return;
  durations_kernel<true><<<batch, 256, 0, ex.st>>>(logw, 0.f, 0.f, 1.f, 0, length_scale, utt_start, utt_len,
                                                   nullptr, dur, cum, y_len);
  KLAUNCH(ex);
}

// frame rows: fstart[b] = sum_{b'<b} (y_len[b'] + gap); fstart[batch] = total rows
__global__ void frame_layout_kernel(const int64_t* __restrict__ y_len, int batch, int gap,
                                    int64_t* __restrict__ fstart) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    int64_t acc = 0;
    for (int b = 0; b < batch; ++b) {
      fstart[b] = acc;
      acc += y_len[b] + gap;
    }
    fstart[batch] = acc;
  }
}
void launch_frame_layout(Exec& ex, const int64_t* y_len, int batch, int gap, int64_t* fstart) {
  frame_layout_kernel<<<1, 32, 0, ex.st>>>(y_len, batch, gap, fstart);
  KLAUNCH(ex);
}

// Range/Less/CumSum path (generate_path) as an index: frame f of utterance b belongs to the
// first phoneme t with cum[t] > f.  Frames nobody covers (sum of durations 0) get -1.
__global__ void frame_map_kernel(const int64_t* __restrict__ fstart, const int64_t* __restrict__ y_len,
                                 const int* __restrict__ cum, const int* __restrict__ utt_start,
                                 const int* __restrict__ utt_len, int batch, int Rf,
                                 int* __restrict__ frow_utt, int* __restrict__ frow_src) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= Rf) return;
  int lo = 0, hi = batch - 1;   // last b with fstart[b] <= r
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (fstart[mid] <= r) lo = mid; else hi = mid - 1;
  }
  const int b = lo;
  const int64_t f = r - fstart[b];
  int utt = -1, src = -1;
  if (f < y_len[b]) {
    utt = b;
    const int T = utt_len[b], r0 = utt_start[b];
    int l = 0, h = T;   // first t with cum[t] > f
    while (l < h) {
      const int mid = (l + h) >> 1;
      if (cum[r0 + mid] > f) h = mid; else l = mid + 1;
    }
    src = l < T ? r0 + l : -1;
  }
  frow_utt[r] = utt;
  frow_src[r] = src;
}
void launch_frame_map(Exec& ex, const int64_t* fstart, const int64_t* y_len, const int* cum,
                      const int* utt_start, const int* utt_len, int batch, int Rf, int* frow_utt,
                      int* frow_src) {
  if (Rf <= 0) return;
  frame_map_kernel<<<(Rf + 255) / 256, 256, 0, ex.st>>>(fstart, y_len, cum, utt_start, utt_len, batch, Rf,
                                                       frow_utt, frow_src);
  KLAUNCH(ex);
}

// ------------------------------------------------------------------------------------------
// Expansion + prior sample: the exported MatMul path x (m_p, logs_p) is a
// gather by frame index (0 FLOP); z_p = m_p + eps * exp(logs_p) * noise_scale.
// eps is either injected ([hidden, stride] per utterance) or the reference generator with
// gid = c*F_b + f (a [1,192,F] tensor per utterance).
// ------------------------------------------------------------------------------------------
__global__ void expand_prior_kernel(const float* __restrict__ stats, const int* __restrict__ frow_utt,
                                    const int* __restrict__ frow_src, const int64_t* __restrict__ fstart,
                                    const int64_t* __restrict__ y_len, const float* __restrict__ ext,
                                    int64_t ext_stride, uint32_t seed, float noise_scale,
                                    float* __restrict__ zp, int Rf, int H) {
  PDL_ENTER();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)Rf * H) return;
  const int r = (int)(i / H), c = (int)(i % H);
  const int b = frow_utt[r];
  float v = 0.f;
  if (b >= 0) {
    const int src = frow_src[r];
    float m = 0.f, logs = 0.f;
    if (src >= 0) {
      m = stats[(size_t)src * 2 * H + c];
      logs = stats[(size_t)src * 2 * H + H + c];
    }
    const int64_t f = r - fstart[b];
    const int64_t F = y_len[b];
    const float eps = ext ? ext[((size_t)b * H + c) * ext_stride + f]
                          : ref_normal(seed, (uint32_t)((int64_t)c * F + f));
    v = m + eps * expf(logs) * noise_scale;
  }
  zp[i] = v;
}
void launch_expand_prior(Exec& ex, const float* stats, const int* frow_utt, const int* frow_src,
                         const int64_t* fstart, const int64_t* y_len, const float* ext_noise,
                         int64_t ext_stride, uint32_t seed, float noise_scale, float* zp, int Rf, int H) {
  const int64_t n = (int64_t)Rf * H;
  if (n <= 0) return;
  launch_pdl(expand_prior_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, ex.st, stats, frow_utt, frow_src, fstart, y_len,
             ext_noise, ext_stride, seed, noise_scale, zp, Rf, H);
  KLAUNCH(ex);
}

// WN gate (fused_add_tanh_sigmoid_multiply): Add cond, Split, Tanh, Sigmoid, Mul
__global__ void wn_gate_kernel(const float* __restrict__ xin, const float* __restrict__ cond, int ldc,
                               const int* __restrict__ row_utt, float* __restrict__ acts, int R, int H) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)R * H) return;
  const int r = (int)(i / H), c = (int)(i % H);
  const int b = row_utt ? row_utt[r] : 0;
  float v = 0.f;
  if (b >= 0) {
    float ta = xin[(size_t)r * 2 * H + c], sa = xin[(size_t)r * 2 * H + H + c];
    if (cond) {
      ta += cond[(size_t)b * ldc + c];
      sa += cond[(size_t)b * ldc + H + c];
    }
    v = tanhf(ta) * (1.0f / (1.0f + expf(-sa)));
  }
  acts[i] = v;
}
void launch_wn_gate(Exec& ex, const float* xin, const float* cond, int ldc, const int* row_utt,
                    float* acts, int R, int H) {
  const int64_t n = (int64_t)R * H;
  if (n <= 0) return;
  wn_gate_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ex.st>>>(xin, cond, ldc, row_utt, acts, R, H);
  KLAUNCH(ex);
}

// conv_post (C -> 1 channel; stock Piper has no bias, forks may: a scalar added before the Tanh) with its preceding
// LeakyRelu and the final Tanh.
constexpr int CP_ROWS = 128;
__global__ void __launch_bounds__(CP_ROWS) conv_post_tanh_kernel(const float* __restrict__ in, int C,
                                                                   const float* __restrict__ w, int k, float slope,
                                                                   const int* __restrict__ row_utt, int valid_div,
                                                                   float* __restrict__ out, int R, float bias) {
  // a CTA stages CP_ROWS + k - 1 input rows (leaky-relu applied) in shared memory with coalesced
  // loads, then every thread reduces one output row over taps x channels (row stride C+1: no bank conflicts)
  extern __shared__ float sm[];
  float* ws = sm;                 // [k][C]
  float* xs = sm + k * C;         // [CP_ROWS + k - 1][C + 1]
  const int pad = (k - 1) / 2;
  const int r0 = blockIdx.x * CP_ROWS;
  for (int i = threadIdx.x; i < k * C; i += blockDim.x) ws[i] = w[i];
  const int nrows = CP_ROWS + k - 1;
  const int c4n = C / 4;
  for (int i = threadIdx.x; i < nrows * c4n; i += blockDim.x) {
    const int rr = i / c4n, c4 = i - rr * c4n;
    const int gr = r0 + rr - pad;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (gr >= 0 && gr < R) v = *reinterpret_cast<const float4*>(in + (size_t)gr * C + c4 * 4);
    float* d = xs + rr * (C + 1) + c4 * 4;
    d[0] = v.x >= 0.f ? v.x : v.x * slope;
    d[1] = v.y >= 0.f ? v.y : v.y * slope;
    d[2] = v.z >= 0.f ? v.z : v.z * slope;
    d[3] = v.w >= 0.f ? v.w : v.w * slope;
  }
  __syncthreads();
  const int r = r0 + threadIdx.x;
  if (r >= R) return;
  if (row_utt && row_utt[r / valid_div] < 0) {
    out[r] = 0.f;
    return;
  }
  float acc = 0.f;
  for (int t = 0; t < k; ++t) {
    const float* xr = xs + (threadIdx.x + t) * (C + 1);
    const float* wp = ws + t * C;
    for (int c = 0; c < C; ++c) acc = fmaf(xr[c], wp[c], acc);
  }
  out[r] = tanhf(acc + bias);
}
// Warp-per-row-block variant for k <= 7 and C = 32 * n (n <= 4): lane c owns channels c, c+32, ..; the taps'
// weights sit in registers, input rows are read straight from global memory (one coalesced 128-byte load per
// row and channel group, ~40 independent loads in flight per lane), 32 rows are reduced at once with a
// butterfly transpose (31 shuffles per 32 rows) and stored with one coalesced write.  No shared memory: the
// staged kernel above is bound by its 2 LDS per FMA; this one by HBM.
constexpr int CPW_ROWS = 32, CPW_MAXK = 7, CPW_MAXN = 4;
template <int N>
__global__ void __launch_bounds__(256) conv_post_tanh_warp_kernel(const float* __restrict__ in, const float* __restrict__ w,
                                                                  int k, float slope, const int* __restrict__ row_utt,
                                                                  int valid_div, float* __restrict__ out, int R, float bias) {
  constexpr int C = 32 * N;
  const int lane = threadIdx.x & 31;
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t r0 = (int64_t)wid * CPW_ROWS;
  if (r0 >= R) return;
  const int pad = (k - 1) / 2;
  float wv[N][CPW_MAXK];
#pragma unroll
  for (int n = 0; n < N; ++n)
#pragma unroll
    for (int t = 0; t < CPW_MAXK; ++t) wv[n][t] = t < k ? w[t * C + n * 32 + lane] : 0.f;
  float v[CPW_ROWS];
#pragma unroll
  for (int i = 0; i < CPW_ROWS; ++i) v[i] = 0.f;
#pragma unroll
  for (int n = 0; n < N; ++n) {
    float x[CPW_ROWS + CPW_MAXK - 1];
#pragma unroll
    for (int i = 0; i < CPW_ROWS + CPW_MAXK - 1; ++i) {
      const int64_t gr = r0 + i - pad;
      float xv = 0.f;
      if (i < CPW_ROWS + k - 1 && gr >= 0 && gr < R) xv = in[gr * C + n * 32 + lane];
      x[i] = xv >= 0.f ? xv : xv * slope;
    }
#pragma unroll
    for (int i = 0; i < CPW_ROWS; ++i)
#pragma unroll
      for (int t = 0; t < CPW_MAXK; ++t) v[i] = fmaf(x[i + t], wv[n][t], v[i]);
  }
  // butterfly transpose-reduce: afterwards lane l holds the channel sum of row r0 + l
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
#pragma unroll
    for (int j = 0; j < s; ++j) {
      const bool up = (lane & s) != 0;
      const float keep = up ? v[j + s] : v[j];
      const float send = up ? v[j] : v[j + s];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  const int64_t r = r0 + lane;
  if (r < R) {
    const bool valid = !row_utt || row_utt[r / valid_div] >= 0;
    out[r] = valid ? tanhf(v[0] + bias) : 0.f;
  }
}

void launch_conv_post_tanh(Exec& ex, const float* in, int C, const float* w, int k, float slope,
                           const int* row_utt, int valid_div, float* out, int R, float bias) {
  if (R <= 0) return;
  static const int cpw_env = [] { const char* e = getenv("PIPER_B200_CONV_POST_WARP"); return e ? atoi(e) : 1; }();
  if (cpw_env && k <= CPW_MAXK && (k & 1) && C % 32 == 0 && C / 32 <= CPW_MAXN) {
    const int warps = (R + CPW_ROWS - 1) / CPW_ROWS;
    const unsigned grid = (unsigned)((warps + 7) / 8);
    switch (C / 32) {
      case 1: conv_post_tanh_warp_kernel<1><<<grid, 256, 0, ex.st>>>(in, w, k, slope, row_utt, valid_div, out, R, bias); break;
      case 2: conv_post_tanh_warp_kernel<2><<<grid, 256, 0, ex.st>>>(in, w, k, slope, row_utt, valid_div, out, R, bias); break;
      case 3: conv_post_tanh_warp_kernel<3><<<grid, 256, 0, ex.st>>>(in, w, k, slope, row_utt, valid_div, out, R, bias); break;
      default: conv_post_tanh_warp_kernel<4><<<grid, 256, 0, ex.st>>>(in, w, k, slope, row_utt, valid_div, out, R, bias); break;
    }
    KLAUNCH(ex);
    return;
  }
  const size_t smem = sizeof(float) * ((size_t)k * C + (size_t)(CP_ROWS + k - 1) * (C + 1));
  conv_post_tanh_kernel<<<(R + CP_ROWS - 1) / CP_ROWS, CP_ROWS, smem, ex.st>>>(in, C, w, k, slope, row_utt,
                                                                              valid_div, out, R, bias);
  KLAUNCH(ex);
}

// speaker conditioning: Gather(emb_g, sid) + 1x1 Conv on a single column; one warp per output
__global__ void cond_gemv_kernel(const float* __restrict__ emb_g, const int64_t* __restrict__ sids,
                                 const float* __restrict__ W, const float* __restrict__ bias,
                                 float* __restrict__ out, int ldo, int gin, int cout, int n_speakers) {
  const int b = blockIdx.y;
  const int co = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (co >= cout) return;
  int64_t sid = sids ? sids[b] : 0;
  if (sid < 0) sid += n_speakers;
  float s = 0.f;
  if (sid >= 0 && sid < n_speakers)
    for (int ci = lane; ci < gin; ci += 32) s = fmaf(W[(size_t)co * gin + ci], emb_g[sid * gin + ci], s);
  s = warp_sum(s);
  if (lane == 0) out[(size_t)b * ldo + co] = s + (bias ? bias[co] : 0.f);
}
void launch_cond_gemv(Exec& ex, const float* emb_g, const int64_t* sids, const float* W, const float* bias,
                      float* out, int ldo, int batch, int gin, int cout, int n_speakers) {
  if (batch <= 0 || cout <= 0) return;
  cond_gemv_kernel<<<dim3((cout + 7) / 8, batch), 256, 0, ex.st>>>(emb_g, sids, W, bias, out, ldo, gin, cout,
                                                                n_speakers);
  KLAUNCH(ex);
}

// packed sample rows -> dense per-utterance concatenation (+ PCM16 by WavFileWriter.swift:20-30:
// Int(clamp(x,-1,1)*32767), truncation toward zero)
__global__ void gather_audio_kernel(const float* __restrict__ packed, const int64_t* __restrict__ fstart,
                                    const int64_t* __restrict__ y_len, const int64_t* __restrict__ sample_off,
                                    int hop, float* __restrict__ dense, int16_t* __restrict__ pcm16) {
  const int b = blockIdx.y;
  const int64_t n = y_len[b] * hop;
  const float* src = packed + fstart[b] * hop;
  const int64_t o = sample_off[b];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = src[i];
    dense[o + i] = v;
    if (pcm16) pcm16[o + i] = (int16_t)(int)(fminf(fmaxf(v, -1.f), 1.f) * 32767.f);
  }
}
__global__ void pcm16_kernel(const float* __restrict__ src, int16_t* __restrict__ dst, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = (int16_t)(int)(fminf(fmaxf(src[i], -1.f), 1.f) * 32767.f);
}
void launch_pcm16(Exec& ex, const float* src, int16_t* dst, int64_t n) {
  if (n <= 0) return;
  int g = (int)((n + 1023) / 1024);
  pcm16_kernel<<<g > 1184 ? 1184 : g, 256, 0, ex.st>>>(src, dst, n);
  KLAUNCH(ex);
}
void launch_gather_audio(Exec& ex, const float* packed, const int64_t* fstart, const int64_t* y_len,
                         const int64_t* sample_off, int batch, int hop, float* dense, int16_t* pcm16,
                         int64_t total_samples) {
  if (batch <= 0 || total_samples <= 0) return;
  int gx = (int)((total_samples / batch + 1023) / 1024);
  gx = gx < 1 ? 1 : (gx > 256 ? 256 : gx);
  gather_audio_kernel<<<dim3(gx, batch), 256, 0, ex.st>>>(packed, fstart, y_len, sample_off, hop, dense, pcm16);
  KLAUNCH(ex);
}

// concatenated ids -> packed rows (-1 on gap rows) + per-row utterance id
__global__ void pack_ids_kernel(const int64_t* __restrict__ ids, const int64_t* __restrict__ id_off,
                                const int* __restrict__ utt_start, const int* __restrict__ utt_len,
                                int64_t* __restrict__ ids_packed, int* __restrict__ row_utt, int n_vocab) {
  const int b = blockIdx.y;
  const int T = utt_len[b], r0 = utt_start[b];
  const int64_t o = id_off[b];
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < T; t += gridDim.x * blockDim.x) {
    int64_t id = ids[o + t];
    if (id < 0) id = (id >= -(int64_t)n_vocab) ? id + n_vocab : INT64_MIN;   // negative-index wrap (gather.metal); -1 stays the gap marker
    ids_packed[r0 + t] = id;
    row_utt[r0 + t] = b;
  }
}
void launch_pack_ids(Exec& ex, const int64_t* ids, const int64_t* id_off, const int* utt_start,
                     const int* utt_len, int batch, int64_t* ids_packed, int* row_utt, int n_vocab) {
  if (batch <= 0) return;
  pack_ids_kernel<<<dim3(4, batch), 256, 0, ex.st>>>(ids, id_off, utt_start, utt_len, ids_packed, row_utt, n_vocab);
  KLAUNCH(ex);
}

}  // namespace piper

namespace piper {
bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("PIPER_B200_PDL"); return !e || atoi(e) != 0; }();
  return on;
}
}  // namespace piper
