// Relative-position self-attention of the text encoder on the tensor cores (tcgen05 / TMEM / TMA, sm_100a), for head
// dimension 96 and utterances of at most 128 phoneme ids (every utterance of BASELINE configs 1-3 and 5; longer ones keep
// attention96_kernel).  Replaces the MatMul QK^T / rel-key MatMul + skew / Softmax / MatMul PV / rel-value MatMul chain of
// one attention layer (GraphExecutor.swift:1862-1929, matmul.metal:22-49, softmax.metal:13-41).
//
// One CTA (256 threads) per (utterance, head); threads i and i + 128 share query row i = TMEM lane i.
//   S = Q K^T   : Q and K arrive as the f16x3 split planes the qkv convolution writes (hi, lo' = (x - hi) * 2^11); three
//                 products  D0 = Qh Kh,  D1 = Qh Kl + Ql Kh  (M = 128 queries, N = 128 keys, K = 96) - fp32-equivalent,
//                 because ceil() of the predicted durations downstream is a discrete decision
//   softmax     : thread i reads its row of S from TMEM, adds the banded relative-key term q_i . rel_k[j - i] (fp32 dot
//                 products, |j - i| <= window), masks j >= T, exact max / exp / sum in fp32
//   O = P V     : P is split like the other operands (hi, lo') and written to shared memory in the K-major 128-byte
//                 swizzle the tensor core reads; V is transposed through registers into the same layout (keys contiguous);
//                 D0 = Ph Vh, D1 = Ph Vl + Pl Vh (M = 128, N = 96, K = 128 keys); the banded relative-value term and the
//                 division by the row sum happen in the output epilogue.
#include "kernels.h"
#include "tc_common.cuh"

#include <cfloat>
#include <cstdio>
#include <cstdlib>

namespace piper {

using namespace tc;

namespace {

constexpr int ATC_T = 128;            // queries = keys per CTA
constexpr int ATC_D = 96;             // head dimension
constexpr int ATC_MAXREL = 17;
constexpr int ATC_QK_BLOCK = ATC_T * 64;        // bytes of one 32-channel block of a Q / K plane (64-byte rows, SWIZZLE_64B)
constexpr int ATC_P_SUB = ATC_T * 128;          // bytes of one 64-key sub-tile of a P plane (128-byte rows, SWIZZLE_128B)
constexpr int ATC_V_SUB = ATC_D * 128;          // bytes of one 64-key sub-tile of a V^T plane
constexpr int ATC_RK_BLOCK = 16 * 64;           // bytes of one 32-channel block of a rel_k plane (16 rows)

struct AttTcParams {
  const uint16_t* qkvs;      // f16 split planes [rows][2][3H]  (q | k | v per plane)
  const float* rel_v;        // [2w+1][96]
  float* out;                // fp32 [rows][H] or null
  uint16_t* out_split2;      // f16 split planes [rows][2][H] or null
  const int* utt_start;
  const int* utt_len;
  int H, window, rows;
  long long* trace;          // debug (PIPER_B200_ATT_TRACE=1): clock64 of CTA (0, 0), thread 0 at the phase boundaries
};

// byte offset of element (row, key j) inside a K-major plane made of 64-key sub-tiles with 128-byte swizzled rows
__device__ __forceinline__ uint32_t sw128_off(int row, int j, int sub_bytes) {
  const int kt = j >> 6, c = (j & 63) >> 3;
  return (uint32_t)(kt * sub_bytes + row * 128 + ((c ^ (row & 7)) << 4) + ((j & 7) << 1));
}

__global__ void __launch_bounds__(2 * ATC_T, 1) attention_tc_kernel(const __grid_constant__ CUtensorMap tmap,
                                                                  const __grid_constant__ CUtensorMap tmap_rk, const AttTcParams p) {
  PDL_TRIGGER();
  const int head = blockIdx.x, b = blockIdx.y;
  const int T = p.utt_len[b];   // written by a copy long before the preceding kernel
  if (T > ATC_T || T <= 0) return;     // long utterances: attention96_kernel (launched with skip_short)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* Qs = smem;                                  // [2 planes][3 blocks][128 x 64 B]   49152
  uint8_t* Ks = Qs + 6 * ATC_QK_BLOCK;                 // 49152
  uint8_t* Ps = smem;                                  // [2 planes][2 sub-tiles][128 x 128 B] 65536, over Q / K once S is done
  uint8_t* Vt = Ks + 6 * ATC_QK_BLOCK;                 // [2 planes][2 sub-tiles][96 x 128 B]  49152
  uint8_t* RKs = Vt + 4 * ATC_V_SUB;                           // rel_k planes: [2 planes][3 blocks][16 x 64 B]  6144
  float* RV = reinterpret_cast<float*>(RKs + 6 * ATC_RK_BLOCK);   // [nrel][96]
  float* QRs = RV + ATC_MAXREL * ATC_D;                        // [2][128][17]  q_i . rel_k[m], one copy per thread of the row
  float* PBs = QRs + 2 * ATC_T * ATC_MAXREL;                   // [128][17]  p[i][i + m - w]
  float* MXs = PBs + ATC_T * ATC_MAXREL;                       // [2][128] row maxima of the two key halves
  float* SMs = MXs + 2 * ATC_T;                                // [2][128] row sums
  uint64_t* bars = reinterpret_cast<uint64_t*>(SMs + 2 * ATC_T);
  uint64_t* full = bars;
  uint64_t* s_full = bars + 1;
  uint64_t* o_full = bars + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);

  const int tid = threadIdx.x, warp = tid >> 5;
  int trace_n = 0;
  auto TR = [&]() { if (p.trace && tid == 0 && blockIdx.x == 0 && blockIdx.y == 0) p.trace[trace_n++] = clock64(); };
  TR();
  const int row0 = p.utt_start[b];
  const int nrel = 2 * p.window + 1;
  const int H = p.H;
  if (tid == 0) {
    prefetch_tmap(&tmap);
    prefetch_tmap(&tmap_rk);
    mbar_init(full, 1);
    mbar_init(s_full, 1);
    mbar_init(o_full, 1);
    fence_barrier_init();
    PDL_WAIT();   // the qkv convolution's output is read from here on
    // Q and K planes of this head (2 planes x 3 blocks of 32 channels, 128 rows from the utterance start) and rel_k: requested
    // at once, so that the loads run under the TMEM allocation and the rest of the set-up
    mbar_expect_tx(full, 12u * ATC_QK_BLOCK + 6u * ATC_RK_BLOCK);
    for (int pl = 0; pl < 2; ++pl)
      for (int kb = 0; kb < 3; ++kb) {
        tma_load_2d(&tmap, full, Qs + (pl * 3 + kb) * ATC_QK_BLOCK, pl * 3 * H + head * ATC_D + kb * 32, row0);
        tma_load_2d(&tmap, full, Ks + (pl * 3 + kb) * ATC_QK_BLOCK, pl * 3 * H + H + head * ATC_D + kb * 32, row0);
        tma_load_2d(&tmap_rk, full, RKs + (pl * 3 + kb) * ATC_RK_BLOCK, pl * ATC_D + kb * 32, 0);
      }
  }
  if (warp == 0) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  for (int t = tid; t < nrel * ATC_D; t += 2 * ATC_T) RV[t] = p.rel_v[t];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  PDL_WAIT();
  TR();   // 1: setup done
  const uint32_t s0_col = 0, s1_col = 128, o0_col = 256, o1_col = 352, r0_col = 448, r1_col = 464;

  // Two threads per query row (threads i and i + 128, warps w and w + 4 share TMEM lanes): `half` splits the relative
  // offsets, the keys of the softmax and the output channels between them - with one warp per scheduler the kernel was
  // bound by the latency of its own instruction stream.
  const int i = tid & (ATC_T - 1), half = tid >> 7;
  const bool qlive = i < T;
  // the V rows of this thread (48 channels of key row i, both planes) are requested now and scattered after the S products
  // have been issued: their latency hides behind the TMA loads
  uint4 vreg[2][ATC_D / 16];
  {
    const uint16_t* vrow = p.qkvs + (size_t)(row0 + i) * 6 * H + 2 * H + head * ATC_D + half * (ATC_D / 2);
#pragma unroll
    for (int pl = 0; pl < 2; ++pl)
#pragma unroll
      for (int c8 = 0; c8 < ATC_D / 16; ++c8) {
        vreg[pl][c8] = make_uint4(0u, 0u, 0u, 0u);
        if (qlive) vreg[pl][c8] = *reinterpret_cast<const uint4*>(vrow + pl * 3 * H + c8 * 8);
      }
  }
#pragma unroll
  for (int m = 0; m < ATC_MAXREL; ++m)
    if ((m & 1) == half) PBs[i * ATC_MAXREL + m] = 0.f;

  TR();   // 2: q . rel_k done
  // ---- S = Q K^T ----
  if (warp == 0) {
    mbar_wait(full, 0u);
    tc_fence_after();
    if (elect_one()) {
      const uint32_t idesc = make_idesc(ATC_T, 1);
      const uint64_t hi = make_smem_desc(0, 32);
      const uint32_t q_lo = smem_u32(Qs) >> 4, k_lo = smem_u32(Ks) >> 4;
      constexpr uint32_t blk16 = ATC_QK_BLOCK >> 4;
      // (Q plane, K plane, accumulator): Qh Kh -> D0;  Qh Kl, Ql Kh -> D1
      const int qa[3] = {0, 0, 1}, kbp[3] = {0, 1, 0};
      const uint32_t acc[3] = {s0_col, s1_col, s1_col};
      for (int pr = 0; pr < 3; ++pr)
        for (int kb = 0; kb < 3; ++kb)
          for (int ks = 0; ks < 2; ++ks)
            umma_f16(tmem_base + acc[pr], hi | (uint64_t)(q_lo + (uint32_t)(qa[pr] * 3 + kb) * blk16 + 2u * ks),
                     hi | (uint64_t)(k_lo + (uint32_t)(kbp[pr] * 3 + kb) * blk16 + 2u * ks), idesc,
                     (pr == 2 || kb > 0 || ks > 0) ? 1u : 0u);
      // q . rel_k on the tensor cores as well: QR[128 x 16] = Q R^T with the same three products (N = 16)
      const uint32_t idesc16 = make_idesc(16, 1);
      const uint32_t r_lo = smem_u32(RKs) >> 4;
      constexpr uint32_t rblk16 = ATC_RK_BLOCK >> 4;
      const uint32_t racc[3] = {r0_col, r1_col, r1_col};
      for (int pr = 0; pr < 3; ++pr)
        for (int kb = 0; kb < 3; ++kb)
          for (int ks = 0; ks < 2; ++ks)
            umma_f16(tmem_base + racc[pr], hi | (uint64_t)(q_lo + (uint32_t)(qa[pr] * 3 + kb) * blk16 + 2u * ks),
                     hi | (uint64_t)(r_lo + (uint32_t)(kbp[pr] * 3 + kb) * rblk16 + 2u * ks), idesc16,
                     (pr == 2 || kb > 0 || ks > 0) ? 1u : 0u);
      umma_commit(s_full);
    }
    __syncwarp();
  }
  TR();   // 3: S products issued
  // ---- V^T while the S products run: thread (j, half) holds 48 channels of key row j (hi and lo' planes, loaded before
  // the first store) and scatters them, keys contiguous, into the swizzled tiles ----
  {
    const int j = i;
#pragma unroll
    for (int pl = 0; pl < 2; ++pl) {
      uint8_t* dst = Vt + pl * 2 * ATC_V_SUB;
#pragma unroll
      for (int c8 = 0; c8 < ATC_D / 16; ++c8) {
        const uint32_t w[4] = {vreg[pl][c8].x, vreg[pl][c8].y, vreg[pl][c8].z, vreg[pl][c8].w};
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int d = half * (ATC_D / 2) + c8 * 8 + e;
          *reinterpret_cast<uint16_t*>(dst + sw128_off(d, j, ATC_V_SUB)) = (uint16_t)(w[e >> 1] >> ((e & 1) * 16));
        }
      }
    }
  }
  fence_proxy_async();          // V^T was written with generic stores, the tensor core reads it through the async proxy
  TR();   // 4: V^T written
  __syncthreads();              // both halves of every row of QRs / PBs are in place
  mbar_wait(s_full, 0u);
  tc_fence_after();
  TR();   // 5: S ready

  // ---- softmax over the row (two passes over TMEM: maximum, then exp / sum / P); half h takes keys [64 h, 64 h + 64) ----
  const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
  // scores are kept in units of log2(e): p = 2^(t - max t) with t = s * log2(e), one exp2f per element instead of expf
  const float inv_sqrt_d = (1.0f / sqrtf((float)ATC_D)) * 1.4426950408889634f;
  {   // q_i . rel_k[m] (scaled like the scores): every thread keeps its own copy of the row's 16 values in shared memory
    uint32_t r0[16], r1[16];
    __syncwarp();
    tmem_ld16_async(tmem_base + lane_sel + r0_col, r0);
    tmem_ld16_async(tmem_base + lane_sel + r1_col, r1);
    tmem_wait_ld();
#pragma unroll
    for (int m = 0; m < 16; ++m)
      QRs[(half * ATC_T + i) * ATC_MAXREL + m] = fmaf(__uint_as_float(r1[m]), 1.0f / 2048.f, __uint_as_float(r0[m])) * inv_sqrt_d;
  }
  const float* qrs = QRs + (half * ATC_T + i) * ATC_MAXREL;
  const int nchunk = (T + 31) >> 5;
  float mx = -FLT_MAX;
  for (int c = 2 * half; c < 2 * half + 2; ++c) {
    if (c >= nchunk) break;
    uint32_t a0[16], a1[16], c0[16], c1[16];
    __syncwarp();
    tmem_ld16_async(tmem_base + lane_sel + s0_col + (uint32_t)(c * 32), a0);
    tmem_ld16_async(tmem_base + lane_sel + s0_col + (uint32_t)(c * 32 + 16), a1);
    tmem_ld16_async(tmem_base + lane_sel + s1_col + (uint32_t)(c * 32), c0);
    tmem_ld16_async(tmem_base + lane_sel + s1_col + (uint32_t)(c * 32 + 16), c1);
    tmem_wait_ld();
#pragma unroll
    for (int e = 0; e < 32; ++e) {
      const int j = c * 32 + e;
      const float d0 = __uint_as_float(e < 16 ? a0[e] : a1[e - 16]), d1 = __uint_as_float(e < 16 ? c0[e] : c1[e - 16]);
      float sc = fmaf(d1, 1.0f / 2048.f, d0) * inv_sqrt_d;
      const int rel = j - i + p.window;
      if (rel >= 0 && rel < nrel) sc += qrs[rel];
      if (j < T) mx = fmaxf(mx, sc);
    }
  }
  MXs[half * ATC_T + i] = mx;
  TR();   // 6: pass 1 done
  tc_fence_before();
  __syncthreads();              // the maxima of both halves; every thread is also past the S products, so P may overwrite Q / K
  mx = fmaxf(MXs[i], MXs[ATC_T + i]);
  float sum = 0.f;
  for (int c = 2 * half; c < 2 * half + 2; ++c) {
    uint32_t ph[16], pl[16];    // 32 keys: 16 packed pairs per plane
    if (c < nchunk) {
      uint32_t a0[16], a1[16], c0[16], c1[16];
      __syncwarp();
      tmem_ld16_async(tmem_base + lane_sel + s0_col + (uint32_t)(c * 32), a0);
      tmem_ld16_async(tmem_base + lane_sel + s0_col + (uint32_t)(c * 32 + 16), a1);
      tmem_ld16_async(tmem_base + lane_sel + s1_col + (uint32_t)(c * 32), c0);
      tmem_ld16_async(tmem_base + lane_sel + s1_col + (uint32_t)(c * 32 + 16), c1);
      tmem_wait_ld();
      float pv[32];
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        const int j = c * 32 + e;
        const float d0 = __uint_as_float(e < 16 ? a0[e] : a1[e - 16]), d1 = __uint_as_float(e < 16 ? c0[e] : c1[e - 16]);
        float sc = fmaf(d1, 1.0f / 2048.f, d0) * inv_sqrt_d;
        const int rel = j - i + p.window;
        const bool inband = rel >= 0 && rel < nrel;
        if (inband) sc += qrs[rel];
        const float pj = (j < T && qlive) ? exp2f(sc - mx) : 0.f;
        if (inband) PBs[i * ATC_MAXREL + rel] = pj;
        sum += pj;
        pv[e] = pj;
      }
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        const __half h0 = __float2half_rn(pv[2 * e]), h1 = __float2half_rn(pv[2 * e + 1]);
        const __half l0 = __float2half_rn((pv[2 * e] - __half2float(h0)) * 2048.f);
        const __half l1 = __float2half_rn((pv[2 * e + 1] - __half2float(h1)) * 2048.f);
        ph[e] = (uint32_t)__half_as_ushort(h0) | ((uint32_t)__half_as_ushort(h1) << 16);
        pl[e] = (uint32_t)__half_as_ushort(l0) | ((uint32_t)__half_as_ushort(l1) << 16);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 16; ++e) { ph[e] = 0u; pl[e] = 0u; }
    }
#pragma unroll
    for (int g = 0; g < 4; ++g) {   // four 16-byte pieces (8 keys each) per plane
      const uint32_t off = sw128_off(i, c * 32 + g * 8, ATC_P_SUB);
      *reinterpret_cast<uint4*>(Ps + off) = make_uint4(ph[4 * g], ph[4 * g + 1], ph[4 * g + 2], ph[4 * g + 3]);
      *reinterpret_cast<uint4*>(Ps + 2 * ATC_P_SUB + off) = make_uint4(pl[4 * g], pl[4 * g + 1], pl[4 * g + 2], pl[4 * g + 3]);
    }
  }
  SMs[half * ATC_T + i] = sum;
  TR();   // 7: pass 2 done
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();

  // ---- O = P V ----
  if (warp == 0) {
    tc_fence_after();
    if (elect_one()) {
      const uint32_t idesc = make_idesc(ATC_D, 1);
      const uint64_t hi = make_smem_desc(0, 64);
      const uint32_t p_lo = smem_u32(Ps) >> 4, v_lo = smem_u32(Vt) >> 4;
      constexpr uint32_t psub16 = ATC_P_SUB >> 4, vsub16 = ATC_V_SUB >> 4;
      const int ksteps = (T + 15) >> 4;            // keys beyond T carry p = 0 and zero V
      const int pa[3] = {0, 0, 1}, vb[3] = {0, 1, 0};
      const uint32_t acc[3] = {o0_col, o1_col, o1_col};
      for (int pr = 0; pr < 3; ++pr)
        for (int ks = 0; ks < ksteps; ++ks) {
          const uint32_t kt = (uint32_t)ks >> 2, kk = (uint32_t)ks & 3u;
          umma_f16(tmem_base + acc[pr], hi | (uint64_t)(p_lo + (uint32_t)(pa[pr] * 2 + kt) * psub16 + 2u * kk),
                   hi | (uint64_t)(v_lo + (uint32_t)(vb[pr] * 2 + kt) * vsub16 + 2u * kk), idesc, (pr == 2 || ks > 0) ? 1u : 0u);
        }
      umma_commit(o_full);
    }
    __syncwarp();
  }
  // the row sum (the addition order is fixed: keys 0..63, then 64..127) and the band of P while the products run
  sum = SMs[i] + SMs[ATC_T + i];
  const float inv = qlive ? 1.0f / sum : 0.f;
  float pband[ATC_MAXREL];
#pragma unroll
  for (int m = 0; m < ATC_MAXREL; ++m) pband[m] = PBs[i * ATC_MAXREL + m];
  TR();   // 8: PV issued, band read
  mbar_wait(o_full, 0u);
  tc_fence_after();
  TR();   // 9: O ready

  // ---- output: (O + sum_m p[i, i + m - w] rel_v[m]) / sum; half h writes channels [48 h, 48 h + 48) ----
  // (every lane runs the warp-wide TMEM loads; only lanes with a query row store)
  for (int c = 3 * half; c < 3 * half + 3; ++c) {
    uint32_t a0[16], c0[16];
    __syncwarp();
    tmem_ld16_async(tmem_base + lane_sel + o0_col + (uint32_t)(c * 16), a0);
    tmem_ld16_async(tmem_base + lane_sel + o1_col + (uint32_t)(c * 16), c0);
    tmem_wait_ld();
    float o[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) o[e] = fmaf(__uint_as_float(c0[e]), 1.0f / 2048.f, __uint_as_float(a0[e]));
#pragma unroll
    for (int m = 0; m < ATC_MAXREL; ++m) {
      if (m < nrel) {
        const float pm = pband[m];
        const float4* rv4 = reinterpret_cast<const float4*>(RV + m * ATC_D + c * 16);
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          const float4 r = rv4[q4];
          o[4 * q4] = fmaf(pm, r.x, o[4 * q4]); o[4 * q4 + 1] = fmaf(pm, r.y, o[4 * q4 + 1]);
          o[4 * q4 + 2] = fmaf(pm, r.z, o[4 * q4 + 2]); o[4 * q4 + 3] = fmaf(pm, r.w, o[4 * q4 + 3]);
        }
      }
    }
#pragma unroll
    for (int e = 0; e < 16; ++e) o[e] *= inv;
    if (qlive && p.out) {
      float4* op = reinterpret_cast<float4*>(p.out + (size_t)(row0 + i) * H + head * ATC_D + c * 16);
#pragma unroll
      for (int q4 = 0; q4 < 4; ++q4) op[q4] = make_float4(o[4 * q4], o[4 * q4 + 1], o[4 * q4 + 2], o[4 * q4 + 3]);
    }
    if (qlive && p.out_split2) {
      uint32_t hh[8], ll[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float x0 = fminf(fmaxf(o[2 * e], -65504.f), 65504.f), x1 = fminf(fmaxf(o[2 * e + 1], -65504.f), 65504.f);
        const __half h0 = __float2half_rn(x0), h1 = __float2half_rn(x1);
        const __half l0 = __float2half_rn((x0 - __half2float(h0)) * 2048.f), l1 = __float2half_rn((x1 - __half2float(h1)) * 2048.f);
        hh[e] = (uint32_t)__half_as_ushort(h0) | ((uint32_t)__half_as_ushort(h1) << 16);
        ll[e] = (uint32_t)__half_as_ushort(l0) | ((uint32_t)__half_as_ushort(l1) << 16);
      }
      uint16_t* sp = p.out_split2 + (size_t)(row0 + i) * 2 * H + head * ATC_D + c * 16;
      reinterpret_cast<uint4*>(sp)[0] = make_uint4(hh[0], hh[1], hh[2], hh[3]);
      reinterpret_cast<uint4*>(sp)[1] = make_uint4(hh[4], hh[5], hh[6], hh[7]);
      reinterpret_cast<uint4*>(sp + H)[0] = make_uint4(ll[0], ll[1], ll[2], ll[3]);
      reinterpret_cast<uint4*>(sp + H)[1] = make_uint4(ll[4], ll[5], ll[6], ll[7]);
    }
  }
  TR();   // 10: output written
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 512);
}

}  // namespace

bool attention_tc_supported(int H, int n_heads, int window) {
  return n_heads > 0 && H / n_heads == ATC_D && 2 * window + 1 <= 16 && H % 8 == 0;
}

void launch_attention_tc(Exec& ex, const uint16_t* qkvs, const void* relk16, const float* rel_v, float* out,
                         uint16_t* out_split2, const int* utt_start, const int* utt_len, int batch, int rows, int H, int n_heads,
                         int window) {
  if (batch <= 0) return;
  if (!attention_tc_supported(H, n_heads, window)) throw std::runtime_error("attention_tc: unsupported head shape");
  constexpr size_t smem = 12 * ATC_QK_BLOCK + 4 * ATC_V_SUB + 6 * ATC_RK_BLOCK + ATC_MAXREL * ATC_D * 4 + 3 * ATC_T * ATC_MAXREL * 4 +
                          4 * ATC_T * 4 + 64 + 1024;
  static PerDeviceOnce once;
  once.run([&] {
    if (cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
      throw std::runtime_error("cuda: cannot raise dynamic shared memory limit for attention_tc_kernel");
  });
  AttTcParams p{};
  p.qkvs = qkvs; p.rel_v = rel_v; p.out = out; p.out_split2 = out_split2;
  p.utt_start = utt_start; p.utt_len = utt_len; p.H = H; p.window = window; p.rows = rows;
  // one map over the split tensor [rows][2 * 3H] f16: boxes of 32 channels x 128 rows, 64-byte swizzle
  const CUtensorMap tm = make_map(qkvs, (uint64_t)6 * H, (uint64_t)rows, (uint64_t)6 * H, 32, ATC_T, 1);
  const CUtensorMap trk = make_map(relk16, (uint64_t)2 * ATC_D, 16, (uint64_t)2 * ATC_D, 32, 16, 1);
  dim3 grid(n_heads, batch);
  static const bool trace_env = [] { const char* e = getenv("PIPER_B200_ATT_TRACE"); return e && atoi(e) != 0; }();
  if (trace_env) {
    if (cudaMalloc(&p.trace, 16 * sizeof(long long)) != cudaSuccess) throw std::runtime_error("attention_tc: trace buffer");
    cudaMemsetAsync(p.trace, 0, 16 * sizeof(long long), ex.st);
  }
  launch_pdl(attention_tc_kernel, grid, dim3(2 * ATC_T), smem, ex.st, tm, trk, p);
  ++ex.launches;
  if (trace_env) {   // debug only: phase boundaries of CTA (0, 0) in cycles since its start
    long long h[16];
    cudaStreamSynchronize(ex.st);
    cudaMemcpy(h, p.trace, sizeof(h), cudaMemcpyDeviceToHost);
    cudaFree(p.trace);
    fprintf(stderr, "attention_tc trace (cycles):");
    for (int k = 1; k < 11; ++k) fprintf(stderr, " %lld", h[k] - h[0]);
    fprintf(stderr, "\n");
  }
}

}  // namespace piper
