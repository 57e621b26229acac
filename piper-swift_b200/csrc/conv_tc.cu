// tcgen05 / TMEM / TMA implicit-GEMM convolution (sm_100a).  See conv_tc.h for the contract.
//
// One CTA computes a 128-row (time) x nt-column (cout) output tile per column group:
//   warp 0      : TMA producer - per pipeline stage one activation box (128 rows x kb channels,
//                 rows shifted by the tap; out-of-range rows are zero-filled by TMA = the conv's
//                 zero padding) and one weight box per group (nt rows x kb channels)
//   warp 1      : TMEM allocation + the single thread issuing tcgen05.mma (M=128, N=nt, K=16)
//   warps 2..5  : epilogue - tcgen05.ld of the fp32 accumulators (each warp its 32-lane quarter),
//                 bias/residual/gate in registers, vectorised global stores
// Pipeline: full[s]/empty[s] mbarriers between TMA and MMA, tmem_full between MMA and epilogue.
#include "conv_tc.h"
#include "tc_common.cuh"

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>

namespace piper {

using namespace tc;

namespace {

constexpr int TC_SBIAS = 1024;   // floats of bias staged in shared memory per column group (persistent kernel)

struct TcKernelParams {
  int Q, taps, tap_step, in_off, kblocks, kb;
  int nt, n_groups, cout_rows;      // cout_rows = weight rows per (phase, tap)
  int phases, ct_pad, out_mul;
  int gate, f16, window, a_col0, relu;
  int phase_cols;                   // merged polyphase ConvTranspose: columns per phase (0 = off)
  int ppt;                          // phases per column tile (nt / phase_cols; a power of two)
  uint32_t tap_mask[8];             // per column tile: taps to run (merged polyphase weights are zero elsewhere)
  int split_kbc;                    // bf16x6 fp32 emulation: K blocks per operand plane (0 = off)
  int split2;                       // f16x3 fp32 emulation (two planes, two accumulators; uses split_kbc too)
  int sp1;                          // f16x3 in one pass: a stage holds A_hi, A_lo', B_hi, B_lo' of one K block
  // persistent kernel, windowed activations: the A operand of a K block is ONE (128 + span)-row window of the
  // group-planar tensor a_planar [cin / 8][a_rows][8] (row r at plane row a_pad + r), bulk-copied plane by plane into a
  // no-swizzle K-major window; every tap is a descriptor row offset into it.  pwin = window rows (0 = off)
  int pwin;
  const uint16_t* a_planar; long long a_rows; int a_pad;
  TcEpi e[2];
  const float* uttbias; int ldub;
  const int* row_utt; int valid_div;
  int valid_shift;                  // log2(valid_div) when it is a power of two, else -1
  uint32_t tmem_cols;
  int nbuf;                         // accumulator buffers in TMEM (persistent kernel): 2, or 4 when 4 tiles fit in 512 columns
  long long* trace;                 // debug (PIPER_B200_TC_TRACE=1): clock64 of CTA 0 at its phase boundaries
  int dbg;                          // timing attribution probe (PIPER_B200_TC_DBG): 1 no residual load, 2 no fp32 store, 4 no 16-bit store, 8 no staging
};

}  // namespace

namespace {

// x = b0 + b1 + b2 with bf16 terms (round to nearest each time; the remainders are exact in fp32)
__device__ __forceinline__ void split3(float x, uint16_t& b0, uint16_t& b1, uint16_t& b2) {
  const __nv_bfloat16 h0 = __float2bfloat16_rn(x);
  const float r1 = x - __bfloat162float(h0);
  const __nv_bfloat16 h1 = __float2bfloat16_rn(r1);
  const float r2 = r1 - __bfloat162float(h1);
  const __nv_bfloat16 h2 = __float2bfloat16_rn(r2);
  b0 = __bfloat16_as_ushort(h0); b1 = __bfloat16_as_ushort(h1); b2 = __bfloat16_as_ushort(h2);
}
__device__ __forceinline__ void split3x4(const float4& x, uint2& t0, uint2& t1, uint2& t2) {
  uint16_t a[4], b[4], c[4];
  split3(x.x, a[0], b[0], c[0]); split3(x.y, a[1], b[1], c[1]);
  split3(x.z, a[2], b[2], c[2]); split3(x.w, a[3], b[3], c[3]);
  t0 = make_uint2((uint32_t)a[0] | ((uint32_t)a[1] << 16), (uint32_t)a[2] | ((uint32_t)a[3] << 16));
  t1 = make_uint2((uint32_t)b[0] | ((uint32_t)b[1] << 16), (uint32_t)b[2] | ((uint32_t)b[3] << 16));
  t2 = make_uint2((uint32_t)c[0] | ((uint32_t)c[1] << 16), (uint32_t)c[2] | ((uint32_t)c[3] << 16));
}

// x = hi + lo' * 2^-11 with fp16 terms (TcW::split == 2)
__device__ __forceinline__ void split2h(float x, uint16_t& hi, uint16_t& lo) {
  x = fminf(fmaxf(x, -65504.f), 65504.f);
  const __half h = __float2half_rn(x);
  const __half l = __float2half_rn((x - __half2float(h)) * 2048.f);
  hi = __half_as_ushort(h); lo = __half_as_ushort(l);
}
__device__ __forceinline__ void split2x4(const float4& x, uint2& t0, uint2& t1) {
  uint16_t a[4], b[4];
  split2h(x.x, a[0], b[0]); split2h(x.y, a[1], b[1]);
  split2h(x.z, a[2], b[2]); split2h(x.w, a[3], b[3]);
  t0 = make_uint2((uint32_t)a[0] | ((uint32_t)a[1] << 16), (uint32_t)a[2] | ((uint32_t)a[3] << 16));
  t1 = make_uint2((uint32_t)b[0] | ((uint32_t)b[1] << 16), (uint32_t)b[2] | ((uint32_t)b[3] << 16));
}

// Explicit shared-space accesses for the epilogue's staging buffers: the dynamic shared-memory base is rounded up
// through an integer cast, after which the compiler no longer knows the address space and emits generic LD / ST
// (long-scoreboard latency on every staging access).
__device__ __forceinline__ float4 lds128(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts128(uint32_t a, const float4& v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ uint4 lds128u(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts128u(uint32_t a, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// (row tile, column tile, phase) of tile t of the persistent kernel's list, phase fastest.  Every role decodes every tile; the
// three 32-bit divisions of the plain form cost ~250 cycles per tile and warp, which shows where a tile's epilogue is a
// few thousand cycles: one phase needs none, a handful of column tiles a compare-subtract loop.
__device__ __forceinline__ void tile_coords(const TcKernelParams& p, int t, int m_tiles, int n_tiles, int& phase, int& mt,
                                            int& ntile) {
  int rest = t;
  phase = 0;
  if (p.phases > 1) { phase = t % p.phases; rest = t / p.phases; }
  if (n_tiles <= 8) {
    ntile = 0;
    while (rest >= m_tiles) { rest -= m_tiles; ++ntile; }
    mt = rest;
  } else {
    ntile = rest / m_tiles;
    mt = rest - ntile * m_tiles;
  }
}

// Epilogue of one 128-row x (n_groups * nt) accumulator tile, run by the four warps of a warpgroup (TMEM lane
// quarter = warp % 4): waits for `bar`, then bias / utterance bias / residual / accumulate / mask and the
// fp32 and 16-bit stores, or the tanh * sigmoid gate.
__device__ __forceinline__ void tc_epilogue(const TcKernelParams& p, uint32_t tmem_base, int m0, int ntile, int phase,
                                          int warp, int lane, uint64_t* bar, uint32_t bar_parity, float* tbuf,
                                            int part, int nparts,     // this warpgroup's share of the column chunks
                                            const float* sbias = nullptr) {   // bias staged in shared memory: [2][TC_SBIAS]
  const int quarter = warp & 3;
  const int q = m0 + quarter * 32 + lane;
  const bool active = q < p.Q;
  const int orow = active ? q * p.out_mul + (p.phases > 1 ? phase : 0) : 0;
  int utt = 0;
  if (active && p.row_utt) utt = p.row_utt[p.valid_shift >= 0 ? (orow >> p.valid_shift) : (orow / p.valid_div)];
  const bool valid = utt >= 0;
  mbar_wait(bar, bar_parity);
  tc_fence_after();
  const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
  const int chunks = p.nt / 16;
  const uint32_t tbs = tbuf ? smem_u32(tbuf) : 0u, sbs = sbias ? smem_u32(sbias) : 0u;
  if (p.dbg & 8) return;
  if (p.gate) {
    const TcEpi& e0 = p.e[0];
    const TcEpi& e1 = p.e[1];
    for (int c = part * chunks / nparts; c < (part + 1) * chunks / nparts; ++c) {
      float va[16], vb[16];
      __syncwarp();
#ifndef PIPER_TC_EPI_SYNC_LD
      {
        uint32_t ta_[16], tb_[16];
        tmem_ld16_async(lane_base + (uint32_t)(c * 16), ta_);
        tmem_ld16_async(lane_base + (uint32_t)(p.nt + c * 16), tb_);
        tmem_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i) { va[i] = __uint_as_float(ta_[i]); vb[i] = __uint_as_float(tb_[i]); }
      }
#else
      tmem_ld16(lane_base + (uint32_t)(c * 16), va);
      tmem_ld16(lane_base + (uint32_t)(p.nt + c * 16), vb);
#endif
      if (!active) continue;
      const int j = ntile * p.nt + c * 16;
      uint32_t pk[8];
      // bias and per-utterance conditioning come in as 16-byte vectors; the gate uses the hardware tanh
      // (tanh.approx, rel. error ~2^-11 - the result is rounded to 16 bits right after):
      // tanh(a) * sigmoid(b) = tanh(a) * (0.5 + 0.5 tanh(b / 2))
      if (e0.bias && sbias) {   // broadcast reads of the staged copy: the global form waited ~25 % of the epilogue on L1 misses
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 x = lds128(sbs + (uint32_t)(j + 4 * i) * 4u), y = lds128(sbs + (uint32_t)(TC_SBIAS + j + 4 * i) * 4u);
          va[4 * i] += x.x; va[4 * i + 1] += x.y; va[4 * i + 2] += x.z; va[4 * i + 3] += x.w;
          vb[4 * i] += y.x; vb[4 * i + 1] += y.y; vb[4 * i + 2] += y.z; vb[4 * i + 3] += y.w;
        }
      } else if (e0.bias) {
        const float4* b0 = reinterpret_cast<const float4*>(e0.bias + j);
        const float4* b1 = reinterpret_cast<const float4*>(e1.bias + j);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 x = __ldg(b0 + i), y = __ldg(b1 + i);
          va[4 * i] += x.x; va[4 * i + 1] += x.y; va[4 * i + 2] += x.z; va[4 * i + 3] += x.w;
          vb[4 * i] += y.x; vb[4 * i + 1] += y.y; vb[4 * i + 2] += y.z; vb[4 * i + 3] += y.w;
        }
      }
      if (p.uttbias && valid) {
        const float4* u0 = reinterpret_cast<const float4*>(p.uttbias + (size_t)utt * p.ldub + e0.ub_off + j);
        const float4* u1 = reinterpret_cast<const float4*>(p.uttbias + (size_t)utt * p.ldub + e1.ub_off + j);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 x = __ldg(u0 + i), y = __ldg(u1 + i);
          va[4 * i] += x.x; va[4 * i + 1] += x.y; va[4 * i + 2] += x.z; va[4 * i + 3] += x.w;
          vb[4 * i] += y.x; vb[4 * i + 1] += y.y; vb[4 * i + 2] += y.z; vb[4 * i + 3] += y.w;
        }
      }
#pragma unroll
      for (int i = 0; i < 16; i += 2) {
        float o[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          float ta, sb;
          asm("tanh.approx.f32 %0, %1;" : "=f"(ta) : "f"(va[i + h]));
          asm("tanh.approx.f32 %0, %1;" : "=f"(sb) : "f"(vb[i + h] * 0.5f));
          o[h] = valid ? ta * fmaf(sb, 0.5f, 0.5f) : 0.f;
        }
        pk[i / 2] = pack2(o[0], o[1], p.f16);
      }
      uint4* dst = reinterpret_cast<uint4*>(static_cast<uint16_t*>(e0.out_act) + (size_t)orow * e0.ld_act + e0.col0 + j);
      dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    }
  } else if (tbuf && (p.nt & 31) == 0) {
    // Coalesced path: 32-column chunks.  Bias / conditioning are added with lane = row (TMEM layout); the
    // chunk then goes through a per-warp 32 x 36 float buffer so that residual loads, accumulation and the
    // fp32 / 16-bit stores run with 8 lanes per row: every warp instruction touches 4 full 128-byte lines
    // instead of 32 lines (the lane = row form made the L1 tag stage the limit of every epilogue).
    const uint32_t act_mask = __ballot_sync(0xffffffffu, active);
    const int sub = lane >> 3, c4 = (lane & 7) * 4;
    const int row_base = m0 + quarter * 32;
    const TcEpi& em = p.e[0];
    if (p.phase_cols && em.out_act && em.act_planar && !em.out_f32 && nparts == 1) {
      // Merged polyphase ConvTranspose -> group-planar 16-bit operand.  This tile holds ppt consecutive phases of
      // each input row q, i.e. ppt consecutive OUTPUT rows (16 bytes each per plane).  Two planes at a time are
      // assembled in the warp's buffer as [plane][q][phase] and written with every lane storing an adjacent
      // 16-byte piece (512 contiguous bytes per instruction); the lane = row form touched 32 lines per store.
      const int Cp = p.phase_cols, ppt = p.ppt, ph0 = ntile * ppt;
      const int pieces = 32 * ppt;                       // per plane
      const int ppt_shift = __ffs(ppt) - 1;
#ifndef PIPER_TC_UPS_SCALAR
      if (sbias && em.bias) {
        // batched form: the TMEM loads of all phases of a 16-column slice are issued back to back and waited for once
        // (the one-load-one-wait form left the eight epilogue warps idle for a TMEM round trip per phase), bias added as
        // packed fp32 pairs, leaky relu taken in the 16-bit type
        const uint32_t vm = valid ? 0xffffffffu : 0u;
        for (int pp = 0; pp < Cp / 16; ++pp) {
          uint64_t b2[8];
          __syncwarp();
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 x = lds128(sbs + (uint32_t)(pp * 16 + 4 * i) * 4u);
            b2[2 * i] = f2pack(x.x, x.y); b2[2 * i + 1] = f2pack(x.z, x.w);
          }
          // (tried: one phase per round with the next round's load in flight during the arithmetic - tcgen05.wait::ld waits
          // for every load of the thread, so the lookahead is one round at most; 36 bytes of spills, ups launches 50 % slower)
          for (int k0 = 0; k0 < ppt; k0 += 2) {   // two phases per round trip (four would spill at 166 registers)
            uint32_t t[2][16];
            tmem_ld16_async(lane_base + (uint32_t)(k0 * Cp + pp * 16), t[0]);
            if (k0 + 1 < ppt) tmem_ld16_async(lane_base + (uint32_t)((k0 + 1) * Cp + pp * 16), t[1]);
            tmem_wait_ld();
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {
              const int k = k0 + kk;
              if (k < ppt) {
                uint32_t pk[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                  const float2 v = f2unpack(f2add(f2pack(__uint_as_float(t[kk][2 * i]), __uint_as_float(t[kk][2 * i + 1])), b2[i]));
                  pk[i] = lrelu_pack2(v.x, v.y, em.act_slope, p.f16) & vm;
                }
                const int slot = k ^ (lane & (ppt - 1));       // spreads the lanes' 16-byte stores over the banks
                sts128u(tbs + (uint32_t)(lane * ppt + slot) * 16u, make_uint4(pk[0], pk[1], pk[2], pk[3]));
                sts128u(tbs + (uint32_t)(pieces + lane * ppt + slot) * 16u, make_uint4(pk[4], pk[5], pk[6], pk[7]));
              }
            }
          }
          __syncwarp();
          for (int pl = 0; pl < 2; ++pl) {
            uint16_t* plane = static_cast<uint16_t*>(em.out_act) +
                              ((size_t)(em.col0 / 8 + pp * 2 + pl) * em.act_rows + em.act_pad) * 8;
            for (int i = lane; i < pieces; i += 32) {
              const int qq = i >> ppt_shift, sl = i & (ppt - 1);   // ppt is a power of two (checked at launch)
              if (!((act_mask >> qq) & 1u)) continue;
              const int k = sl ^ (qq & (ppt - 1));
              const size_t orow_r = (size_t)(row_base + qq) * p.out_mul + ph0 + k;
              if (!(p.dbg & 4)) *reinterpret_cast<uint4*>(plane + orow_r * 8) = lds128u(tbs + (uint32_t)(pl * pieces + i) * 16u);
            }
          }
        }
        return;
      }
#endif
      for (int pp = 0; pp < Cp / 16; ++pp) {
        __syncwarp();
        for (int k = 0; k < ppt; ++k) {
          float v[16];
          tmem_ld16(lane_base + (uint32_t)(k * Cp + pp * 16), v);
          if (em.bias && sbias) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float4 x = lds128(sbs + (uint32_t)(pp * 16 + 4 * i) * 4u);
              v[4 * i] += x.x; v[4 * i + 1] += x.y; v[4 * i + 2] += x.z; v[4 * i + 3] += x.w;
            }
          } else if (em.bias) {
            const float4* b4 = reinterpret_cast<const float4*>(em.bias + pp * 16);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float4 x = __ldg(b4 + i);
              v[4 * i] += x.x; v[4 * i + 1] += x.y; v[4 * i + 2] += x.z; v[4 * i + 3] += x.w;
            }
          }
          uint32_t pk[8];
#pragma unroll
          for (int i = 0; i < 8; ++i)
            pk[i] = valid ? pack2(lrelu(v[2 * i], em.act_slope), lrelu(v[2 * i + 1], em.act_slope), p.f16) : 0u;
          const int slot = k ^ (lane & (ppt - 1));       // spreads the lanes' 16-byte stores over the banks
          sts128u(tbs + (uint32_t)(lane * ppt + slot) * 16u, make_uint4(pk[0], pk[1], pk[2], pk[3]));
          sts128u(tbs + (uint32_t)(pieces + lane * ppt + slot) * 16u, make_uint4(pk[4], pk[5], pk[6], pk[7]));
        }
        __syncwarp();
        for (int pl = 0; pl < 2; ++pl) {
          uint16_t* plane = static_cast<uint16_t*>(em.out_act) +
                            ((size_t)(em.col0 / 8 + pp * 2 + pl) * em.act_rows + em.act_pad) * 8;
          for (int i = lane; i < pieces; i += 32) {
            const int qq = i >> ppt_shift, sl = i & (ppt - 1);
            if (!((act_mask >> qq) & 1u)) continue;
            const int k = sl ^ (qq & (ppt - 1));
            const size_t orow_r = (size_t)(row_base + qq) * p.out_mul + ph0 + k;
            *reinterpret_cast<uint4*>(plane + orow_r * 8) = lds128u(tbs + (uint32_t)(pl * pieces + i) * 16u);
          }
        }
      }
      return;
    }
    const int cpg = p.nt / 32, nunits = (p.split2 ? 1 : p.n_groups) * cpg;
    const int f_begin = part * nunits / nparts, f_end = (part + 1) * nunits / nparts;
    int g = f_begin / cpg, c = f_begin - g * cpg - 1;
    // the column group's epilogue description lives in registers: read through p.e[g] with a run-time g, every field is
    // an indexed constant load with long-scoreboard latency at its point of use (12 % of the epilogue's stall samples)
    TcEpi ec = p.e[g];
    int gc = g;
    for (int f = f_begin; f < f_end; ++f) {
      if (++c == cpg) { c = 0; ++g; }
      if (g != gc) { ec = p.e[g]; gc = g; }
      const TcEpi& e = ec;
      const bool planar = e.out_act && e.act_planar;
      // Lean form for whole tiles of a plain convolution (every row of the warp's 32 inside the tensor, output row =
      // input row): no per-row predicates, strides hoisted, and the residual loads go out before the accumulator is
      // read so that their latency hides behind the TMEM load, bias add and staging.  The general form below runs
      // ~1100 warp instructions per 32 x 32 chunk at one issue per ~9 cycles (ncu), which made every epilogue with
      // fp32 / 16-bit outputs cost 5-14 us per row tile regardless of K.
      // (a planar 16-bit copy is written from the final value there; the general form writes it before the residual,
      // which only suits convolutions without one)
      const bool fast = !p.phase_cols && p.phases == 1 && p.out_mul == 1 && act_mask == 0xffffffffu;
      const bool planar_early = planar && !fast && !e.res && !e.accum;
      float4 rr[8];
      const size_t fcol = (size_t)e.col0 + (size_t)(ntile * p.nt + c * 32 + c4);
      const size_t frow = (size_t)(row_base + sub);
      const bool fres = fast && e.res && !(p.dbg & 1);
      if (fres) {
        const float* rp = e.res + frow * e.ld_res + fcol;
        const size_t sr = (size_t)4 * e.ld_res;
#pragma unroll
        for (int it = 0; it < 8; ++it) rr[it] = *reinterpret_cast<const float4*>(rp + it * sr);
      }
      {
        float v[32];
        __syncwarp();
#ifndef PIPER_TC_EPI_SYNC_LD
        {   // all TMEM loads of the chunk in flight at once, one wait (two to four round trips per chunk otherwise)
          uint32_t tv[32];
          tmem_ld16_async(lane_base + (uint32_t)(g * p.nt + c * 32), tv);
          tmem_ld16_async(lane_base + (uint32_t)(g * p.nt + c * 32 + 16), tv + 16);
          if (p.split2) {   // f16x3: the cross terms sit in the second accumulator, scaled by 2^11
            uint32_t tw[32];
            tmem_ld16_async(lane_base + (uint32_t)(p.nt + c * 32), tw);
            tmem_ld16_async(lane_base + (uint32_t)(p.nt + c * 32 + 16), tw + 16);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = fmaf(__uint_as_float(tw[i]), 1.0f / 2048.f, __uint_as_float(tv[i]));
          } else {
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(tv[i]);
          }
        }
#else
        tmem_ld16(lane_base + (uint32_t)(g * p.nt + c * 32), v);
        tmem_ld16(lane_base + (uint32_t)(g * p.nt + c * 32 + 16), v + 16);
        if (p.split2) {   // f16x3: the cross terms sit in the second accumulator, scaled by 2^11
          float w[32];
          tmem_ld16(lane_base + (uint32_t)(p.nt + c * 32), w);
          tmem_ld16(lane_base + (uint32_t)(p.nt + c * 32 + 16), w + 16);
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = fmaf(w[i], 1.0f / 2048.f, v[i]);
        }
#endif
        int j = ntile * p.nt + c * 32, ph_c = p.phases > 1 ? phase : 0;
        if (p.phase_cols) {   // merged polyphase weights: the chunk's columns are channels j % phase_cols of phase j / phase_cols
          ph_c = j / p.phase_cols;
          j -= ph_c * p.phase_cols;
        }
        const size_t ocol = (size_t)e.col0 + j;
        const int orow_c = active ? q * p.out_mul + ph_c : 0;
        if (e.bias && sbias) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {   // packed fp32 pairs (FADD2): same rounding, half the issue slots
            const float4 x = lds128(sbs + (uint32_t)(g * TC_SBIAS + j + 4 * i) * 4u);
            const float2 a = f2unpack(f2add(f2pack(v[4 * i], v[4 * i + 1]), f2pack(x.x, x.y)));
            const float2 b = f2unpack(f2add(f2pack(v[4 * i + 2], v[4 * i + 3]), f2pack(x.z, x.w)));
            v[4 * i] = a.x; v[4 * i + 1] = a.y; v[4 * i + 2] = b.x; v[4 * i + 3] = b.y;
          }
        } else if (e.bias) {
          const float4* b4 = reinterpret_cast<const float4*>(e.bias + j);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 x = __ldg(b4 + i);
            v[4 * i] += x.x; v[4 * i + 1] += x.y; v[4 * i + 2] += x.z; v[4 * i + 3] += x.w;
          }
        }
        if (p.uttbias && active && valid) {
          const float4* u4 = reinterpret_cast<const float4*>(p.uttbias + (size_t)utt * p.ldub + e.ub_off + j);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 x = __ldg(u4 + i);
            v[4 * i] += x.x; v[4 * i + 1] += x.y; v[4 * i + 2] += x.z; v[4 * i + 3] += x.w;
          }
        }
        if (e.relu) {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
        }
        if (planar_early && active) {   // group-planar operand of the fused MRF stage: 16 bytes per (row, plane)
          uint16_t* base = static_cast<uint16_t*>(e.out_act) + ((size_t)(ocol >> 3) * e.act_rows + e.act_pad + orow_c) * 8;
#pragma unroll
          for (int pl = 0; pl < 4; ++pl) {
            uint32_t pk[4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
              pk[i] = valid ? pack2(lrelu(v[8 * pl + 2 * i], e.act_slope), lrelu(v[8 * pl + 2 * i + 1], e.act_slope), p.f16) : 0u;
            *reinterpret_cast<uint4*>(base + (size_t)pl * e.act_rows * 8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          }
        }
        if (!e.out_f32 && !e.out_split && (planar_early || !e.out_act)) continue;
        if (p.dbg & 16) continue;
        __syncwarp();                                     // the previous chunk's readers are done with tbuf
#pragma unroll
        for (int i = 0; i < 8; ++i)
          sts128(tbs + (uint32_t)(lane * 36 + 4 * i) * 4u, make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]));
        __syncwarp();
        // the row-validity vote waits for the row -> utterance load of this tile: taken here, after the first chunk's
        // accumulator read and staging, not at the top of the tile where nothing could overlap that latency
        const uint32_t valid_mask = __ballot_sync(0xffffffffu, valid);
        if (fast) {
          float4 x[8];
#pragma unroll
          for (int it = 0; it < 8; ++it) x[it] = lds128(tbs + (uint32_t)((it * 4 + sub) * 36 + c4) * 4u);
          if (fres) {
#pragma unroll
            for (int it = 0; it < 8; ++it) {
              const float2 a = f2unpack(f2add(f2pack(x[it].x, x[it].y), f2pack(rr[it].x, rr[it].y)));
              const float2 b = f2unpack(f2add(f2pack(x[it].z, x[it].w), f2pack(rr[it].z, rr[it].w)));
              x[it] = make_float4(a.x, a.y, b.x, b.y);
            }
          }
          float* of = e.out_f32 ? e.out_f32 + frow * e.ld_f32 + fcol : nullptr;
          const size_t sf = (size_t)4 * e.ld_f32;
          if (e.accum) {
#pragma unroll
            for (int it = 0; it < 8; ++it) rr[it] = *reinterpret_cast<const float4*>(of + it * sf);
#pragma unroll
            for (int it = 0; it < 8; ++it) { x[it].x += rr[it].x; x[it].y += rr[it].y; x[it].z += rr[it].z; x[it].w += rr[it].w; }
          }
          if (e.div != 0.f) {
            const float d = e.div;
#pragma unroll
            for (int it = 0; it < 8; ++it) { x[it].x = x[it].x / d; x[it].y = x[it].y / d; x[it].z = x[it].z / d; x[it].w = x[it].w / d; }
          }
          if (valid_mask != 0xffffffffu) {
#pragma unroll
            for (int it = 0; it < 8; ++it)
              if (!((valid_mask >> (it * 4 + sub)) & 1u)) x[it] = make_float4(0.f, 0.f, 0.f, 0.f);
          }
          if (of && !(p.dbg & 2)) {
#pragma unroll
            for (int it = 0; it < 8; ++it) *reinterpret_cast<float4*>(of + it * sf) = x[it];
          }
          if (e.out_split) {
            uint16_t* sp = static_cast<uint16_t*>(e.out_split) + frow * e.ld_split + fcol;
            const size_t ss = (size_t)4 * e.ld_split;
            const size_t sc = (size_t)e.split_c;
            if (p.split2) {
#pragma unroll
              for (int it = 0; it < 8; ++it) {
                uint2 t0, t1;
                split2x4(x[it], t0, t1);
                *reinterpret_cast<uint2*>(sp + it * ss) = t0;
                *reinterpret_cast<uint2*>(sp + it * ss + sc) = t1;
              }
            } else {
#pragma unroll
              for (int it = 0; it < 8; ++it) {
                uint2 t0, t1, t2;
                split3x4(x[it], t0, t1, t2);
                *reinterpret_cast<uint2*>(sp + it * ss) = t0;
                *reinterpret_cast<uint2*>(sp + it * ss + sc) = t1;
                *reinterpret_cast<uint2*>(sp + it * ss + 2 * sc) = t2;
              }
            }
          }
          if (planar && !(p.dbg & 4)) {
            // group-planar copy [C / 8][act_rows][8]: a lane holds half of a 16-byte (row, plane) piece, the four rows of
            // an instruction are adjacent in the plane (64 contiguous bytes per plane and instruction)
            uint16_t* oa = static_cast<uint16_t*>(e.out_act) + ((fcol >> 3) * (size_t)e.act_rows + e.act_pad + frow) * 8 + (fcol & 7);
            const float sl = e.act_slope;
#pragma unroll
            for (int it = 0; it < 8; ++it)
              *reinterpret_cast<uint2*>(oa + it * 32) = make_uint2(pack2(lrelu(x[it].x, sl), lrelu(x[it].y, sl), p.f16),
                                                                   pack2(lrelu(x[it].z, sl), lrelu(x[it].w, sl), p.f16));
          } else if (e.out_act && !(p.dbg & 4)) {
            uint16_t* oa = static_cast<uint16_t*>(e.out_act) + frow * e.ld_act + fcol;
            const size_t sa = (size_t)4 * e.ld_act;
            const float sl = e.act_slope;
#pragma unroll
            for (int it = 0; it < 8; ++it)
              *reinterpret_cast<uint2*>(oa + it * sa) = make_uint2(pack2(lrelu(x[it].x, sl), lrelu(x[it].y, sl), p.f16),
                                                                   pack2(lrelu(x[it].z, sl), lrelu(x[it].w, sl), p.f16));
          }
          continue;
        }
        // all global loads of the chunk are issued before the first use (8-16 independent 16-byte loads per
        // lane in flight): with only eight epilogue warps per SM the epilogue is latency-, not bandwidth-bound
        float4 aa[8];
        const size_t orow0 = (size_t)(row_base + sub) * p.out_mul + ph_c;
        const size_t rstep = (size_t)4 * p.out_mul;
        float* of = e.out_f32 ? e.out_f32 + orow0 * e.ld_f32 + ocol + c4 : nullptr;
        if (e.res && !(p.dbg & 1)) {
          const float* rp = e.res + orow0 * e.ld_res + ocol + c4;
#pragma unroll
          for (int it = 0; it < 8; ++it)
            if ((act_mask >> (it * 4 + sub)) & 1u) rr[it] = *reinterpret_cast<const float4*>(rp + it * rstep * e.ld_res);
        }
        if (e.accum) {
#pragma unroll
          for (int it = 0; it < 8; ++it)
            if ((act_mask >> (it * 4 + sub)) & 1u) aa[it] = *reinterpret_cast<const float4*>(of + it * rstep * e.ld_f32);
        }
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int row = it * 4 + sub;
          if (!((act_mask >> row) & 1u)) continue;
          float4 x = lds128(tbs + (uint32_t)(row * 36 + c4) * 4u);
          if (e.res && !(p.dbg & 1)) { x.x += rr[it].x; x.y += rr[it].y; x.z += rr[it].z; x.w += rr[it].w; }
          if (e.accum) { x.x += aa[it].x; x.y += aa[it].y; x.z += aa[it].z; x.w += aa[it].w; }
          if (e.div != 0.f) { x.x = x.x / e.div; x.y = x.y / e.div; x.z = x.z / e.div; x.w = x.w / e.div; }
          if (!((valid_mask >> row) & 1u)) x = make_float4(0.f, 0.f, 0.f, 0.f);
          if (of && !(p.dbg & 2)) *reinterpret_cast<float4*>(of + it * rstep * e.ld_f32) = x;
          if (e.out_split) {
            uint16_t* sp = static_cast<uint16_t*>(e.out_split) + (orow0 + it * rstep) * e.ld_split + ocol + c4;
            uint2 t0, t1, t2;
            if (p.split2) {
              split2x4(x, t0, t1);
              *reinterpret_cast<uint2*>(sp) = t0;
              *reinterpret_cast<uint2*>(sp + e.split_c) = t1;
            } else {
              split3x4(x, t0, t1, t2);
              *reinterpret_cast<uint2*>(sp) = t0;
              *reinterpret_cast<uint2*>(sp + e.split_c) = t1;
              *reinterpret_cast<uint2*>(sp + 2 * e.split_c) = t2;
            }
          }
          if (planar && !planar_early) {
            const size_t oc = ocol + c4;
            *reinterpret_cast<uint2*>(static_cast<uint16_t*>(e.out_act) +
                                      ((oc >> 3) * (size_t)e.act_rows + e.act_pad + orow0 + it * rstep) * 8 + (oc & 7)) =
                make_uint2(pack2(lrelu(x.x, e.act_slope), lrelu(x.y, e.act_slope), p.f16),
                           pack2(lrelu(x.z, e.act_slope), lrelu(x.w, e.act_slope), p.f16));
          }
          if (e.out_act && !planar && !(p.dbg & 4)) {
            const uint32_t lo = pack2(lrelu(x.x, e.act_slope), lrelu(x.y, e.act_slope), p.f16);
            const uint32_t hi = pack2(lrelu(x.z, e.act_slope), lrelu(x.w, e.act_slope), p.f16);
            *reinterpret_cast<uint2*>(static_cast<uint16_t*>(e.out_act) + (orow0 + it * rstep) * e.ld_act + ocol + c4) =
                make_uint2(lo, hi);
          }
        }
      }
    }
  } else {
    const int nunits = p.n_groups * chunks;
    for (int f = part * nunits / nparts; f < (part + 1) * nunits / nparts; ++f) {
      const int g = f / chunks, c = f - g * chunks;
      const TcEpi& e = p.e[g];
      {
        float v[16];
        __syncwarp();
        tmem_ld16(lane_base + (uint32_t)(g * p.nt + c * 16), v);
        if (!active) continue;
        const int j = ntile * p.nt + c * 16;
        const size_t ocol = (size_t)e.col0 + j;
        if (e.bias) {
          const float4* b4 = reinterpret_cast<const float4*>(e.bias + j);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 x = __ldg(b4 + i);
            v[4 * i] += x.x; v[4 * i + 1] += x.y; v[4 * i + 2] += x.z; v[4 * i + 3] += x.w;
          }
        }
        if (p.uttbias && valid) {
          const float4* u4 = reinterpret_cast<const float4*>(p.uttbias + (size_t)utt * p.ldub + e.ub_off + j);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 x = __ldg(u4 + i);
            v[4 * i] += x.x; v[4 * i + 1] += x.y; v[4 * i + 2] += x.z; v[4 * i + 3] += x.w;
          }
        }
        if (e.res) {
          const float4* r4 = reinterpret_cast<const float4*>(e.res + (size_t)orow * e.ld_res + ocol);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 r = r4[i];
            v[4 * i] += r.x; v[4 * i + 1] += r.y; v[4 * i + 2] += r.z; v[4 * i + 3] += r.w;
          }
        }
        float4* o4 = e.out_f32 ? reinterpret_cast<float4*>(e.out_f32 + (size_t)orow * e.ld_f32 + ocol) : nullptr;
        if (e.accum) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 r = o4[i];
            v[4 * i] += r.x; v[4 * i + 1] += r.y; v[4 * i + 2] += r.z; v[4 * i + 3] += r.w;
          }
        }
        if (e.div != 0.f) {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = v[i] / e.div;
        }
        if (!valid) {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = 0.f;
        }
        if (o4) {
#pragma unroll
          for (int i = 0; i < 4; ++i) o4[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
        if (e.out_act) {
          uint32_t pk[8];
#pragma unroll
          for (int i = 0; i < 8; ++i)
            pk[i] = pack2(lrelu(v[2 * i], e.act_slope), lrelu(v[2 * i + 1], e.act_slope), p.f16);
          if (e.act_planar) {
            uint16_t* base = static_cast<uint16_t*>(e.out_act) + ((size_t)(ocol >> 3) * e.act_rows + e.act_pad + orow) * 8;
            *reinterpret_cast<uint4*>(base) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            *reinterpret_cast<uint4*>(base + (size_t)e.act_rows * 8) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
          } else {
            uint4* dst = reinterpret_cast<uint4*>(static_cast<uint16_t*>(e.out_act) + (size_t)orow * e.ld_act + ocol);
            dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
          }
        }
      }
    }
  }
  }

__global__ void __launch_bounds__(TC_THREADS) conv_tc_kernel(const __grid_constant__ CUtensorMap tmap_a,
                                                             const __grid_constant__ CUtensorMap tmap_b,
                                                             const __grid_constant__ CUtensorMap tmap_ans,
                                                             const TcKernelParams p, const int stages) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int row_bytes = p.kb * 2;
  const int a_bytes = TC_BM * row_bytes;               // one 128-row activation box
  const int b_bytes = p.nt * row_bytes;                // per group
  // window mode: activations live in a 2-slot ring of 256-row windows (one per K block, shared by
  // all taps through descriptor row offsets); the stage ring then holds weights only
  const int a_ring_bytes = p.window ? 2 * 2 * a_bytes : 0;
  const int stage_bytes = (p.window ? 0 : a_bytes) + p.n_groups * b_bytes;
  uint8_t* stage_base = smem + a_ring_bytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(stage_base + (size_t)stages * stage_bytes);
  uint64_t* empty = full + TC_STAGES;
  uint64_t* a_full = empty + TC_STAGES;
  uint64_t* a_empty = a_full + 2;
  uint64_t* tmem_full = a_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  // warp index and TMEM base go through a shuffle so that the compiler can prove them warp-uniform: the
  // role branches become uniform and the MMA descriptors live in uniform registers (without this every
  // tcgen05.mma pays for several R2UR moves - measured 100 vs 55 cycles per issue, tools/micro/umma_bench.cu)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * TC_BM;
  const int ntile = blockIdx.y;
  const int phase = blockIdx.z;
  int in_off = p.in_off;
  if (p.phases > 1) in_off = (phase + p.ct_pad) / p.phases;
  const int span = (p.taps - 1) * p.tap_step;          // row offset of the last tap (may be negative)
  const int win_row0 = m0 + in_off + (span < 0 ? span : 0);

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_b);
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&a_full[s], 1);
      mbar_init(&a_empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc(tmem_slot, p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      if (p.window) {                                   // K block outer, tap inner
        for (int kbi = 0; kbi < p.kblocks; ++kbi) {
          const int slot = kbi & 1;
          mbar_wait(&a_empty[slot], (((uint32_t)kbi >> 1) & 1u) ^ 1u);
          uint8_t* aw = smem + (size_t)slot * 2 * a_bytes;
          mbar_expect_tx(&a_full[slot], (uint32_t)(2 * a_bytes));
          if (p.window == 2) {   // [channel group][row][8 channels], no swizzle
            tma_load_3d(&tmap_ans, &a_full[slot], aw, 0, win_row0, kbi * (p.kb / 8));
          } else {
            tma_load_2d(&tmap_a, &a_full[slot], aw, kbi * p.kb, win_row0);
            tma_load_2d(&tmap_a, &a_full[slot], aw + a_bytes, kbi * p.kb, win_row0 + TC_BM);
          }
          for (int tap = 0; tap < p.taps; ++tap) {
            mbar_wait(&empty[s], ph ^ 1u);
            uint8_t* st = stage_base + (size_t)s * stage_bytes;
            mbar_expect_tx(&full[s], (uint32_t)stage_bytes);
            const int wrow = (phase * p.taps + tap) * p.cout_rows + ntile * p.nt;
            for (int g = 0; g < p.n_groups; ++g)
              tma_load_2d(&tmap_b, &full[s], st + g * b_bytes, kbi * p.kb, wrow + p.e[g].row0);
            if (++s == stages) { s = 0; ph ^= 1u; }
          }
        }
      } else {
        for (int tap = 0; tap < p.taps; ++tap) {
          const int arow = m0 + in_off + tap * p.tap_step;
          const int wrow = (phase * p.taps + tap) * p.cout_rows + ntile * p.nt;
          for (int kbi = 0; kbi < p.kblocks; ++kbi) {
            mbar_wait(&empty[s], ph ^ 1u);
            uint8_t* st = stage_base + (size_t)s * stage_bytes;
            mbar_expect_tx(&full[s], (uint32_t)stage_bytes);
            tma_load_2d(&tmap_a, &full[s], st, p.a_col0 + kbi * p.kb, arow);
            for (int g = 0; g < p.n_groups; ++g)
              tma_load_2d(&tmap_b, &full[s], st + a_bytes + g * b_bytes, kbi * p.kb, wrow + p.e[g].row0);
            if (++s == stages) { s = 0; ph ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: warp-uniform control flow, one elected lane issues =====
    const uint32_t idesc = make_idesc(p.nt, p.f16);
    const int ksteps = p.kb / 16;
    const uint64_t b_hi = make_smem_desc(0, p.kb);
    const uint64_t a_hi = p.window == 2 ? make_smem_desc_ns(0, 2 * TC_BM * 16, 128) : b_hi;
    const uint32_t a_kstep = p.window == 2 ? (uint32_t)(2 * 2 * TC_BM * 16) >> 4 : 2u;   // descriptor units of 16 bytes
    const uint32_t a_tapstep = (uint32_t)(p.tap_step * (p.window == 2 ? 16 : row_bytes) / 16);   // may be negative
    const uint32_t stage_lo = smem_u32(stage_base) >> 4, ring_lo = smem_u32(smem) >> 4;
    const uint32_t stage16 = (uint32_t)stage_bytes >> 4, a16 = (uint32_t)a_bytes >> 4, b16 = (uint32_t)b_bytes >> 4;
    int s = 0; uint32_t ph = 0;
    const int outer = p.window ? p.kblocks : p.taps, inner = p.window ? p.taps : p.kblocks;
    uint32_t first = 1;
    for (int o = 0; o < outer; ++o) {
      uint32_t a_lo = 0;
      if (p.window) {
        mbar_wait(&a_full[o & 1], ((uint32_t)o >> 1) & 1u);
        a_lo = ring_lo + (uint32_t)(o & 1) * 2u * a16 +
               (uint32_t)(-(span < 0 ? span : 0) * (p.window == 2 ? 16 : row_bytes) / 16);
      }
      for (int i = 0; i < inner; ++i) {
        mbar_wait(&full[s], ph);
        tc_fence_after();
        uint32_t al = a_lo, bl = stage_lo + (uint32_t)s * stage16;
        if (!p.window) { al = bl; bl += a16; }
        if (elect_one()) {
          for (int g = 0; g < p.n_groups; ++g)
            for (int k = 0; k < ksteps; ++k)
              umma_f16(tmem_base + (uint32_t)(g * p.nt), a_hi | (uint64_t)(al + a_kstep * (uint32_t)k),
                       b_hi | (uint64_t)(bl + (uint32_t)g * b16 + 2u * (uint32_t)k), idesc, (first && k == 0) ? 0u : 1u);
          umma_commit(&empty[s]);                 // frees the smem slot when these MMAs retire
        }
        first = 0;
        a_lo += a_tapstep;
        if (++s == stages) { s = 0; ph ^= 1u; }
      }
      if (p.window && elect_one()) umma_commit(&a_empty[o & 1]);
    }
    if (elect_one()) umma_commit(tmem_full);
  } else {
    // ===== epilogue: warps 2..5, TMEM lane quarter = warp % 4 =====
    tc_epilogue(p, tmem_base, m0, ntile, phase, warp, lane, tmem_full, 0u, nullptr, 0, 1);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
}

// Persistent variant: one CTA per SM walks the (row tile, column tile, phase) list; the smem ring runs across
// tiles (up to 8 stages) and the accumulator is double-buffered in TMEM, so the epilogue of tile i (two
// warpgroups, one per TMEM buffer) overlaps the TMA/MMA main loop of tile i+1 inside the same CTA.
constexpr int TCP_MAX_STAGES = 8;
constexpr int TCP_WIN_SLOTS = 3;               // activation windows in flight (window mode)
constexpr int TCP_EG = 1;                      // epilogue warpgroups per TMEM buffer (column split); 2 starves registers
// epilogue warpgroups taking tiles round robin.  The epilogues are latency-bound per warp, so a third group was tried
// (-DPIPER_TC_NG=3): 14 warps put four on one SM sub-partition (16384 registers each), i.e. at most 128 registers per
// thread instead of 168 - a 144-register build fails to launch ("too many resources") although 448 x 144 < 65536, and the
// 128-register build (104 bytes of spills) did not finish the GPU suite inside the call's time limit.  Two groups stay.
#ifndef PIPER_TC_NG
#define PIPER_TC_NG 2
#endif
constexpr int TCP_NG = PIPER_TC_NG;
constexpr int TCP_THREADS = 64 + TCP_NG * TCP_EG * 128;   // producer, MMA issuer, epilogue warpgroups

__global__ void __launch_bounds__(TCP_THREADS, 1) conv_tc_persist_kernel(const __grid_constant__ CUtensorMap tmap_a,
                                                                         const __grid_constant__ CUtensorMap tmap_b,
                                                                         const TcKernelParams p, const int stages,
                                                                         const int m_tiles, const int n_tiles,
                                                                         const int total_tiles) {
  PDL_TRIGGER();   // the next kernel of the stream may be staged; this one waits for its predecessor after the set-up (PDL_WAIT below)
  auto TR = [&](int slot) { if (p.trace && blockIdx.x == 0) p.trace[slot] = clock64(); };
  // wait accounting of the trace build (make trace: -DPIPER_TC_WAIT_TRACE): cycles the MMA warp / epilogue group 0 of
  // CTA 0 spent waiting; compiled out of the product build (the counters cost registers in the epilogue)
#ifdef PIPER_TC_WAIT_TRACE
#define TCW_NOW() (p.trace ? clock64() : 0)
#define TCW_ADD(acc, t0) do { if (p.trace) (acc) += clock64() - (t0); } while (0)
#else
#define TCW_NOW() 0
#define TCW_ADD(acc, t0) do { (void)(t0); } while (0)
#endif
  if (threadIdx.x == 0) TR(0);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int row_bytes = p.kb * 2;
  const int a_bytes = TC_BM * row_bytes;
  const int b_bytes = p.nt * row_bytes;
  const int a_tiles = p.pwin ? 0 : (p.sp1 ? 2 : 1);
  const int stage_bytes = a_tiles * a_bytes + p.n_groups * b_bytes;
  const int win_bytes = p.pwin * 16 * (p.kb / 8);      // window mode: [kb / 8 planes][pwin rows][16 bytes]
  const int a_ring_bytes = p.pwin ? TCP_WIN_SLOTS * win_bytes : 0;
  uint8_t* stage_base = smem + a_ring_bytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(stage_base + (size_t)stages * stage_bytes);
  uint64_t* empty = full + TCP_MAX_STAGES;
  uint64_t* acc_full = empty + TCP_MAX_STAGES;
  uint64_t* acc_empty = acc_full + 4;
  uint64_t* a_full = acc_empty + 4;
  uint64_t* a_empty = a_full + TCP_WIN_SLOTS;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(a_empty + TCP_WIN_SLOTS);
  float* tbuf_base = reinterpret_cast<float*>(stage_base + (size_t)stages * stage_bytes + 512);   // one 32 x 36 float tile per epilogue warp
  float* sbias = tbuf_base + TCP_NG * TCP_EG * 4 * 32 * 36;                                      // [2][TC_SBIAS] bias of both column groups
  const int bias_w = p.phase_cols ? p.phase_cols : n_tiles * p.nt;
  const bool bias_staged = bias_w <= TC_SBIAS;
  if (bias_staged)
    for (int g = 0; g < 2 && g < p.n_groups; ++g)
      if (p.e[g].bias)
        for (int i = threadIdx.x; i < bias_w; i += TCP_THREADS) sbias[g * TC_SBIAS + i] = p.e[g].bias[i];
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_b);
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 4; ++s) {
      mbar_init(&acc_full[s], 1);
      mbar_init(&acc_empty[s], 4 * TCP_EG);   // every warp of the warpgroups draining this buffer
    }
    for (int s = 0; s < TCP_WIN_SLOTS; ++s) {
      mbar_init(&a_full[s], 1);
      mbar_init(&a_empty[s], 1);
    }
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc(tmem_slot, p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
  const uint32_t acc_stride = (uint32_t)(p.n_groups * p.nt);
  if (threadIdx.x == 0) TR(1);
  PDL_WAIT();      // everything above touched constants only (bias, tensor maps); activations are read from here on
  if (threadIdx.x == 0) TR(2);

  if (warp == 0) {
    // ===== TMA producer (one thread; a warp-uniform form with an elected issuer was neutral on the large launches and
    // slower on the small ones) =====
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      int as = 0; uint32_t aph = 0;
      bool first_issued = false;   // trace: when the first stage's loads were issued
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        if (p.pwin) {
          // K block outermost: one window per K block, then the weight tiles of its taps
          int phase_, mt, ntile;
          tile_coords(p, t, m_tiles, n_tiles, phase_, mt, ntile);
          const int plane_bytes = p.pwin * 16;
          for (int kbi = 0; kbi < p.kblocks; ++kbi) {
            mbar_wait(&a_empty[as], aph ^ 1u);
            uint8_t* aw = smem + (size_t)as * win_bytes;
            const uint16_t* src = p.a_planar + ((size_t)(kbi * (p.kb / 8)) * p.a_rows + p.a_pad + mt * TC_BM + p.in_off) * 8;
            if (p.dbg & 32) mbar_arrive(&a_full[as]);
            else {
              mbar_expect_tx(&a_full[as], (uint32_t)win_bytes);
              for (int pl = 0; pl < p.kb / 8; ++pl)
                bulk_load(aw + pl * plane_bytes, src + (size_t)pl * p.a_rows * 8, (uint32_t)plane_bytes, &a_full[as]);
            }
            if (++as == TCP_WIN_SLOTS) { as = 0; aph ^= 1u; }
            for (int tap = 0; tap < p.taps; ++tap) {
              mbar_wait(&empty[s], ph ^ 1u);
              uint8_t* st = stage_base + (size_t)s * stage_bytes;
              const int wrow = tap * p.cout_rows + ntile * p.nt;
              if (p.dbg & 32) mbar_arrive(&full[s]);   // timing probe: no loads, stale operands
              else {
                mbar_expect_tx(&full[s], (uint32_t)stage_bytes);
                for (int g = 0; g < p.n_groups; ++g)
                  tma_load_2d(&tmap_b, &full[s], st + g * b_bytes, kbi * p.kb, wrow + p.e[g].row0);
              }
              if (!first_issued) { first_issued = true; TR(7); }
              if (++s == stages) { s = 0; ph ^= 1u; }
            }
          }
          continue;
        }
        // phase fastest: the polyphase outputs of one row tile interleave in memory (row = q * stride + phase), so
        // they are produced back to back and merge in L2 instead of rewriting every line `phases` times
        int phase, mt, ntile;
        tile_coords(p, t, m_tiles, n_tiles, phase, mt, ntile);
        const int m0 = mt * TC_BM;
        const int in_off = p.phases > 1 ? (phase + p.ct_pad) / p.phases : p.in_off;
        // split (fp32-emulation) weights: operand plane outermost, so that every small term of every tap is
        // accumulated before the first large one (the tensor core aligns addends to the largest exponent and
        // drops what falls below its accumulator width)
        const int kb_inner = p.split_kbc ? p.split_kbc : p.kblocks;
        const int n_outer = p.sp1 ? 1 : (p.split2 ? 2 : (p.split_kbc ? 6 : 1));
        for (int pl = 0; pl < n_outer; ++pl) {
          // bf16x6: A planes {0,1,2,0,1,0} against weight planes 0..5; f16x3: A_hi against (B_hi -> D0, B_lo' -> D1),
          // then A_lo' against B_hi -> D1 only
          const int a_plane = p.split2 ? pl : (p.split_kbc ? (int)((0x010210u >> (4 * pl)) & 0xFu) : 0);
          for (int tap = 0; tap < p.taps; ++tap) {
            if (!((p.tap_mask[ntile & 7] >> tap) & 1u)) continue;
            const int arow = m0 + in_off + tap * p.tap_step;
            const int wrow = (phase * p.taps + tap) * p.cout_rows + ntile * p.nt;
            for (int kbi = 0; kbi < kb_inner; ++kbi) {
              mbar_wait(&empty[s], ph ^ 1u);
              uint8_t* st = stage_base + (size_t)s * stage_bytes;
              if (p.sp1) {
                // one pass: both activation planes and both weight planes of this K block (A_hi, B_hi and B_lo' are
                // fetched once instead of twice; the K chain of a tile is half as long)
                if (p.dbg & 32) mbar_arrive(&full[s]);
                else {
                  mbar_expect_tx(&full[s], (uint32_t)(2 * a_bytes + 2 * b_bytes));
                  for (int h = 0; h < 2; ++h) {
                    tma_load_2d(&tmap_a, &full[s], st + h * a_bytes, (h * kb_inner + kbi) * p.kb, arow);
                    tma_load_2d(&tmap_b, &full[s], st + 2 * a_bytes + h * b_bytes, (h * kb_inner + kbi) * p.kb, wrow + p.e[0].row0);
                  }
                }
                if (!first_issued) { first_issued = true; TR(7); }
                if (++s == stages) { s = 0; ph ^= 1u; }
                continue;
              }
              const int g0 = (p.split2 && pl == 1) ? 1 : 0;
              if (p.dbg & 32) mbar_arrive(&full[s]);   // timing probe: no loads, stale operands
              else {
                mbar_expect_tx(&full[s], (uint32_t)(a_bytes + (p.n_groups - g0) * b_bytes));
                tma_load_2d(&tmap_a, &full[s], st, (a_plane * kb_inner + kbi) * p.kb, arow);
                for (int g = g0; g < p.n_groups; ++g) {
                  const int b_plane = p.split2 ? (pl == 0 ? g : 0) : pl;
                  tma_load_2d(&tmap_b, &full[s], st + a_bytes + g * b_bytes, (b_plane * kb_inner + kbi) * p.kb, wrow + p.e[g].row0);
                }
              }
              if (!first_issued) { first_issued = true; TR(7); }
              if (++s == stages) { s = 0; ph ^= 1u; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: warp-uniform control flow, one elected lane issues =====
    const uint32_t idesc = make_idesc(p.nt, p.f16);
    const bool merge2 = p.n_groups == 2 && 2 * p.nt <= 256;
    const uint32_t idesc2 = make_idesc(2 * p.nt, p.f16);
    const int ksteps = p.kb / 16;
    const uint64_t d_hi = make_smem_desc(0, p.kb);
    const uint32_t stage_lo = smem_u32(stage_base) >> 4, ring_lo = smem_u32(smem) >> 4;
    const uint32_t stage16 = (uint32_t)stage_bytes >> 4, a16 = (uint32_t)a_bytes >> 4, b16 = (uint32_t)b_bytes >> 4;
    const uint64_t a_ns = make_smem_desc_ns(0, (uint32_t)p.pwin * 16u, 128);   // planes pwin * 16 bytes apart, 8-row groups 128
    const uint32_t win16 = (uint32_t)win_bytes >> 4, a_kstep_ns = 2u * (uint32_t)p.pwin;   // 16 channels = two planes
    int as = 0; uint32_t aph = 0;
    int s = 0; uint32_t ph = 0, li = 0;
#ifdef PIPER_TC_WAIT_TRACE
    long long w_acc = 0, w_full = 0;   // trace: cycles this warp waited for a drained accumulator / for operands
#endif
    const uint32_t nbuf = (uint32_t)p.nbuf;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++li) {
      int phase_m, mt_m, ntile_m;
      tile_coords(p, t, m_tiles, n_tiles, phase_m, mt_m, ntile_m);
      const int n_it = __popc(p.tap_mask[ntile_m & 7]) * p.kblocks;
      // With four buffers each epilogue warpgroup owns two: the MMAs of its next tile run while it still drains the
      // current one (with two, a group's MMA -> epilogue chain was serial, which small-K tiles cannot hide).
      const uint32_t buf = nbuf == 4 ? (li & 3u) : (li & 1u);
      const uint32_t use = nbuf == 4 ? (li >> 2) : (li >> 1);
      const long long tw0 = TCW_NOW();
      mbar_wait(&acc_empty[buf], (use & 1u) ^ 1u);           // the epilogue drained this accumulator buffer
      tc_fence_after();
      TCW_ADD(w_acc, tw0);
      const uint32_t acc = tmem_base + buf * acc_stride;
      if (p.pwin) {
        bool have_w = false;
        for (int kbi = 0; kbi < p.kblocks; ++kbi) {
          mbar_wait(&a_full[as], aph);
          tc_fence_after();
          const uint32_t a_lo = ring_lo + (uint32_t)as * win16;
          for (int tap = 0; tap < p.taps; ++tap) {
            if (!have_w) {
              const long long tf0 = TCW_NOW();
              mbar_wait(&full[s], ph);
              tc_fence_after();
              TCW_ADD(w_full, tf0);
            }
            if (li == 0 && kbi == 0 && tap == 0 && lane == 0) TR(3);
            const uint32_t al = a_lo + (uint32_t)(tap * p.tap_step), bl = stage_lo + (uint32_t)s * stage16;
            auto issue = [&](int k0, int k1) {
              if (merge2) {
                for (int k = k0; k < k1; ++k)
                  umma_f16(acc, a_ns | (uint64_t)(al + a_kstep_ns * (uint32_t)k), d_hi | (uint64_t)(bl + 2u * (uint32_t)k), idesc2,
                           (kbi == 0 && tap == 0 && k == 0) ? 0u : 1u);
              } else {
                for (int g = 0; g < p.n_groups; ++g)
                  for (int k = k0; k < k1; ++k)
                    umma_f16(acc + (uint32_t)(g * p.nt), a_ns | (uint64_t)(al + a_kstep_ns * (uint32_t)k),
                             d_hi | (uint64_t)(bl + (uint32_t)g * b16 + 2u * (uint32_t)k), idesc,
                             (kbi == 0 && tap == 0 && k == 0) ? 0u : 1u);
              }
            };
#ifdef PIPER_TC_PREWAIT
            const int k_split = ksteps - 1;
#else
            const int k_split = ksteps;
#endif
            if (elect_one()) issue(0, k_split);
            int s1 = s + 1; uint32_t ph1 = ph;
            if (s1 == stages) { s1 = 0; ph1 ^= 1u; }
            have_w = false;
            if (k_split < ksteps && !(kbi + 1 == p.kblocks && tap + 1 == p.taps)) {   // same pre-wait as the plain loop below
              const long long tf0 = TCW_NOW();
              mbar_wait(&full[s1], ph1);
              tc_fence_after();
              TCW_ADD(w_full, tf0);
              have_w = true;
            }
            if (elect_one()) {
              issue(k_split, ksteps);
              umma_commit(&empty[s]);
            }
            s = s1; ph = ph1;
          }
          if (elect_one()) umma_commit(&a_empty[as]);   // the window is free when the MMAs of its last tap retire
          if (++as == TCP_WIN_SLOTS) { as = 0; aph ^= 1u; }
        }
        if (elect_one()) umma_commit(&acc_full[buf]);
        if (li == 0 && lane == 0) TR(4);
        continue;
      }
      // PIPER_TC_PREWAIT (off): take the wait for stage i + 1 before the last K step of stage i is issued.  Tried to
      // close the gap between the stages' MMA groups; it couples the release of slot s to the arrival of slot s + 1
      // (one stage less in flight) and measured 4 % slower over the conv class.
      bool have = false;
      for (int i = 0; i < n_it; ++i) {
        if (!have) {
          const long long tf0 = TCW_NOW();
          mbar_wait(&full[s], ph);
          tc_fence_after();
          TCW_ADD(w_full, tf0);
        }
        if (li == 0 && i == 0 && lane == 0) TR(3);
        const uint32_t al = stage_lo + (uint32_t)s * stage16, bl = al + (p.sp1 ? 2u : 1u) * a16;
        const int g0 = (p.split2 && !p.sp1 && 2 * i >= n_it) ? 1 : 0;   // f16x3 in two passes: the A_lo' half of the K loop feeds D1 only
        auto issue = [&](int k0, int k1) {
          if (p.sp1) {
            // A_hi x [B_hi ; B_lo'] -> D0 | D1 (adjacent weight tiles, adjacent accumulators: one N = 2 nt instruction),
            // then A_lo' x B_hi -> D1
            for (int k = k0; k < k1; ++k) {
              umma_f16(acc, d_hi | (uint64_t)(al + 2u * (uint32_t)k), d_hi | (uint64_t)(bl + 2u * (uint32_t)k), idesc2,
                       (i == 0 && k == 0) ? 0u : 1u);
              umma_f16(acc + (uint32_t)p.nt, d_hi | (uint64_t)(al + a16 + 2u * (uint32_t)k),
                       d_hi | (uint64_t)(bl + 2u * (uint32_t)k), idesc, 1u);
            }
          } else
#ifndef PIPER_TC_NO_GROUP_MERGE
          if (merge2 && g0 == 0) {
            // Both groups read the same A tile, their B tiles are adjacent in the stage and their accumulators adjacent
            // in TMEM: one N = 2*nt instruction fetches A once instead of twice (an N=96 MMA is operand-fetch bound).
            for (int k = k0; k < k1; ++k)
              umma_f16(acc, d_hi | (uint64_t)(al + 2u * (uint32_t)k), d_hi | (uint64_t)(bl + 2u * (uint32_t)k), idesc2,
                       (i == 0 && k == 0) ? 0u : 1u);
          } else
#endif
          for (int g = g0; g < p.n_groups; ++g)
            for (int k = k0; k < k1; ++k)
              umma_f16(acc + (uint32_t)(g * p.nt), d_hi | (uint64_t)(al + 2u * (uint32_t)k),
                       d_hi | (uint64_t)(bl + (uint32_t)g * b16 + 2u * (uint32_t)k), idesc, (i == 0 && k == 0) ? 0u : 1u);
        };
#ifdef PIPER_TC_PREWAIT
        const int k_split = ksteps > 1 ? ksteps - 1 : ksteps;
#else
        const int k_split = ksteps;
#endif
        if (elect_one()) issue(0, k_split);
        int s1 = s + 1; uint32_t ph1 = ph;
        if (s1 == stages) { s1 = 0; ph1 ^= 1u; }
        have = false;
        if (k_split < ksteps && i + 1 < n_it) {
          const long long tf0 = TCW_NOW();
          mbar_wait(&full[s1], ph1);
          tc_fence_after();
          TCW_ADD(w_full, tf0);
          have = true;
        }
        if (elect_one()) {
          issue(k_split, ksteps);
          umma_commit(&empty[s]);                 // frees the smem slot when these MMAs retire
        }
        s = s1; ph = ph1;
      }
      if (elect_one()) umma_commit(&acc_full[buf]);
      if (li == 0 && lane == 0) TR(4);
    }
#ifdef PIPER_TC_WAIT_TRACE
    if (p.trace && blockIdx.x == 0 && lane == 0) { p.trace[8] = w_acc; p.trace[9] = w_full; }
#endif
  } else {
    // ===== epilogue: four warpgroups = (TMEM buffer, i.e. every other tile) x (half of the column chunks) =====
    const uint32_t group = ((uint32_t)(warp - 2) >> 2) % (uint32_t)TCP_NG, half = (uint32_t)(warp - 2) / (4u * TCP_NG);
    // With two accumulator buffers only two groups take tiles: a third would wait on a buffer's acc_full barrier at
    // every third use, and a parity wait that skips two phases passes early (same hazard as conv_mrf.cu's weight ring).
    // With four buffers a group skips one use per visit, and that use has completed before the group's own previous
    // tile did (MMAs retire in issue order), so the wait is sound.
    const uint32_t ng_eff = p.nbuf == 4 ? (uint32_t)TCP_NG : 2u;
    uint32_t li = 0, turn = 0;
#ifdef PIPER_TC_WAIT_TRACE
    long long e_wait = 0, e_busy = 0;   // trace: cycles warp 2 waited for an accumulator / spent in its epilogues
#endif
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++li) {
      const bool mine = turn == group;
      if (++turn == ng_eff) turn = 0;
      if (!mine) continue;
      int phase, mt, ntile;
      tile_coords(p, t, m_tiles, n_tiles, phase, mt, ntile);
      const uint32_t buf = p.nbuf == 4 ? (li & 3u) : (li & 1u);
      const uint32_t use = p.nbuf == 4 ? (li >> 2) : (li >> 1);
#ifdef PIPER_TC_WAIT_TRACE
      long long te0 = 0, te1 = 0;
      if (p.trace) {   // waits do not consume the phase: the one inside tc_epilogue passes at once
        te0 = clock64();
        mbar_wait(&acc_full[buf], use & 1u);
        te1 = clock64();
      }
#endif
      tc_epilogue(p, tmem_base + buf * acc_stride, mt * TC_BM, ntile, phase, warp, lane, &acc_full[buf], use & 1u,
                  tbuf_base + (warp - 2) * (32 * 36), TCP_EG == 1 ? 0 : (int)half, TCP_EG, bias_staged ? sbias : nullptr);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
#ifdef PIPER_TC_WAIT_TRACE
      if (p.trace) { e_wait += te1 - te0; e_busy += clock64() - te1; }
#endif
      if (li == 0 && warp == 2 && lane == 0) TR(5);
    }
#ifdef PIPER_TC_WAIT_TRACE
    if (p.trace && blockIdx.x == 0 && warp == 2 && lane == 0) { p.trace[10] = e_wait; p.trace[11] = e_busy; }
#endif
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
  if (threadIdx.x == 32) TR(6);
}

}  // namespace

// ---- host side -----------------------------------------------------------------------------
namespace {

}  // namespace

// cin only has to be a multiple of the tcgen05 K step: the last K block may be partial (TMA zero-fills the
// channels beyond cin, the packed weights are zero-padded to cin_pad)
bool tc_conv_supported(int cin, int cout) { return cin % 16 == 0 && cout % 16 == 0 && cin >= 32 && cout >= 16; }

void launch_conv_tc(Exec& ex, const TcConvArgs& a) {
  if (a.Q <= 0) return;
  const TcW& w = *a.w;
  if (a.nt % 16 || a.nt < 16 || a.nt > 256 || a.n_groups < 1 || a.n_groups > 2 || a.n_groups * a.nt > 512)
    throw std::runtime_error("conv_tc: bad column tiling");
  TcKernelParams p{};
  static const int dbg_env = [] { const char* e = getenv("PIPER_B200_TC_DBG"); return e ? atoi(e) : 0; }();
  p.dbg = dbg_env;
  p.Q = a.Q; p.taps = w.taps; p.tap_step = a.tap_step; p.in_off = a.in_off;
  p.kb = w.kb; p.kblocks = w.cin_pad / w.kb;
  const int kmul = w.split == 2 ? 2 : (w.split ? 6 : 1), amul = w.split == 2 ? 2 : (w.split ? 3 : 1);
  int n_groups = a.n_groups;
  if (w.split == 2) {
    if (!a.f16 || a.gate || w.phase_cols || a.n_groups != 1 || a.nt > 128 || (a.nt & 31) || w.cin != w.cin_pad || w.phases != 1)
      throw std::runtime_error("conv_tc: f16x3 split weights need fp16 operands, one column group of at most 128 columns and whole K blocks");
    static const int sp1_env = [] { const char* e = getenv("PIPER_B200_SPLIT_1PASS"); return e ? atoi(e) : 1; }();
    p.split_kbc = p.kblocks;
    p.sp1 = sp1_env ? 1 : 0;
    if (!p.sp1) p.kblocks *= 2;
    p.split2 = 1;
    n_groups = 2;   // second accumulator for the cross terms, same weight rows
  } else if (w.split) {
    if (a.f16 || a.gate || w.phase_cols || a.n_groups * a.nt > 256 || w.cin != w.cin_pad)
      throw std::runtime_error("conv_tc: split (fp32-emulation) weights need bf16 operands, whole K blocks and the persistent kernel");
    p.split_kbc = p.kblocks;
    p.kblocks *= 6;
  }
  for (int g = 0; g < a.n_groups; ++g)
    if ((a.e[g].relu || a.e[g].out_split) && (a.nt & 31))
      throw std::runtime_error("conv_tc: relu / split outputs need 32-column tiles");
  p.nt = a.nt; p.n_groups = n_groups; p.cout_rows = w.cout;
  p.phases = w.phases; p.ct_pad = w.ct_pad; p.out_mul = a.out_mul;
  p.gate = a.gate; p.f16 = a.f16;
  p.e[0] = a.e[0]; p.e[1] = w.split == 2 ? a.e[0] : a.e[1];
  p.uttbias = a.uttbias; p.ldub = a.ldub;
  p.row_utt = a.row_utt; p.valid_div = a.valid_div < 1 ? 1 : a.valid_div;
  p.valid_shift = -1;
  if ((p.valid_div & (p.valid_div - 1)) == 0) { p.valid_shift = 0; while ((1 << p.valid_shift) < p.valid_div) ++p.valid_shift; }
  uint32_t cols = 32;
  while ((int)cols < a.n_groups * a.nt) cols <<= 1;
  p.tmem_cols = cols;
  p.phase_cols = w.phase_cols;
  for (int t = 0; t < 8; ++t) p.tap_mask[t] = 0xFFFFFFFFu >> (32 - (w.taps < 32 ? w.taps : 32));
  if (w.phase_cols) {   // merged polyphase ConvTranspose: only the persistent kernel's coalesced epilogue maps columns to phases
    const int ppt_ = a.nt / w.phase_cols;
    if (a.n_groups != 1 || (a.nt & 31) || w.phase_cols % 32 || a.nt % w.phase_cols || a.n_tiles > 8 || a.gate ||
        (ppt_ & (ppt_ - 1)) || ppt_ > 4)   // ppt: power of two, [2 planes][32][ppt] 16-byte pieces must fit the warp buffer
      throw std::runtime_error("conv_tc: bad tiling for merged polyphase weights");
    const int ppt = a.nt / w.phase_cols;   // phases per column tile
    p.ppt = ppt;
    for (int t = 0; t < a.n_tiles; ++t) {
      uint32_t m = 0;
      for (int ph = t * ppt; ph < (t + 1) * ppt && ph < 16; ++ph) m |= w.phase_tapmask[ph];
      p.tap_mask[t] = m;
    }
  }

  const CUtensorMap ta = make_map(a.in, (uint64_t)a.cin * amul, (uint64_t)a.R_in, (uint64_t)a.ldin, w.kb, TC_BM, a.f16);
  static const int persist_env = [] { const char* e = getenv("PIPER_B200_TC_PERSIST"); return e ? atoi(e) : 1; }();
  if (w.phase_cols && !(persist_env && a.n_groups * a.nt <= 256))
    throw std::runtime_error("conv_tc: merged polyphase weights need the persistent kernel");
  if ((persist_env || w.split) && n_groups * a.nt <= 256) {
    static int num_sms = 0;
    static PerDeviceOnce once;
    once.run([&] {
      int dev = 0;
      cudaGetDevice(&dev);
      if (cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || num_sms <= 0)
        throw std::runtime_error("cuda: cannot query the SM count");
      if (cudaFuncSetAttribute(conv_tc_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess)
        throw std::runtime_error("cuda: cannot raise dynamic shared memory limit for conv_tc_persist_kernel");
    });
    const CUtensorMap tbp = make_map(w.w, (uint64_t)w.cin_pad * kmul, (uint64_t)w.phases * w.taps * w.cout,
                                     (uint64_t)w.cin_pad * kmul, w.kb, a.nt, a.f16);
    p.window = 0;
    static const int nbuf_env = [] { const char* e = getenv("PIPER_B200_TC_NBUF"); return e ? atoi(e) : 4; }();
    p.nbuf = (nbuf_env == 4 && 4 * n_groups * a.nt <= 512) ? 4 : 2;
    uint32_t pc = 32;
    while ((int)pc < p.nbuf * n_groups * a.nt) pc <<= 1;
    p.tmem_cols = pc;
    // windowed activations (see TcKernelParams::pwin): plain multi-tap convolutions whose input exists group-planar
    static const int pwin_env = [] { const char* e = getenv("PIPER_B200_TC_PWIN"); return e ? atoi(e) : 1; }();
    const int span = (w.taps - 1) * a.tap_step;
    if (a.in_planar) {
      if (!pwin_env || w.taps < 2 || w.split || w.phase_cols || w.phases != 1 || w.kb != 64 || w.cin != w.cin_pad || a.tap_step < 1 ||
          span > 64 || a.in_pad + a.in_off < 0 || a.out_mul != 1)
        throw std::runtime_error("conv_tc: a group-planar input needs a plain multi-tap convolution with whole 64-channel K blocks");
      p.pwin = TC_BM + ((span + 7) & ~7);
      p.a_planar = static_cast<const uint16_t*>(a.in_planar);
      p.a_rows = a.in_rows; p.a_pad = a.in_pad;
      if ((int64_t)a.in_pad + a.in_off + (int64_t)((a.Q + TC_BM - 1) / TC_BM) * TC_BM + (p.pwin - TC_BM) > a.in_rows)
        throw std::runtime_error("conv_tc: group-planar input has too few pad rows behind the signal");
    }
    const int ring = p.pwin ? TCP_WIN_SLOTS * p.pwin * 16 * (w.kb / 8) : 0;
    const int sb = (p.pwin ? 0 : (p.sp1 ? 2 : 1) * TC_BM * w.kb * 2) + n_groups * a.nt * w.kb * 2;
    constexpr int TBUF_BYTES = TCP_NG * TCP_EG * 4 * 32 * 36 * 4;
    constexpr int SBIAS_BYTES = 2 * TC_SBIAS * 4;
    int st = (TC_SMEM_MAX - 3072 - TBUF_BYTES - SBIAS_BYTES - ring) / sb;   // 1 KB static + alignment slack + barriers
    if (st > TCP_MAX_STAGES) st = TCP_MAX_STAGES;
    if (st < 2) throw std::runtime_error("conv_tc: tile does not fit in shared memory");
    const int m_tiles = (a.Q + TC_BM - 1) / TC_BM;
    const int total = m_tiles * a.n_tiles * w.phases;
    const int grid = total < num_sms ? total : num_sms;
    const size_t smem = (size_t)ring + (size_t)st * sb + 1024 + 512 + TBUF_BYTES + SBIAS_BYTES;
    double flops = 2.0 * a.Q * w.phases * w.taps * w.cin * (double)(a.n_groups * a.nt * a.n_tiles);
    if (w.phase_cols) {   // algorithmic work only: the zero blocks of the merged weights do not count
      int used = 0;
      for (int ph = 0; ph < w.merged_phases; ++ph) used += __builtin_popcount(w.phase_tapmask[ph]);
      flops = 2.0 * a.Q * w.cin * (double)w.phase_cols * used;
    }
    ProfScope ps(ex, Prof::CONV_TC, flops);
    static const bool trace_env = [] { const char* e = getenv("PIPER_B200_TC_TRACE"); return e && atoi(e) != 0; }();
    if (trace_env) {
      if (cudaMalloc(&p.trace, 12 * sizeof(long long)) != cudaSuccess) throw std::runtime_error("conv_tc: trace buffer");
      cudaMemsetAsync(p.trace, 0, 12 * sizeof(long long), ex.st);
    }
    launch_pdl(conv_tc_persist_kernel, dim3(grid), dim3(TCP_THREADS), smem, ex.st, ta, tbp, p, st, m_tiles, a.n_tiles, total);
    if (trace_env) {   // debug only: CTA 0's phase boundaries in cycles since its start
      long long h[12];
      cudaStreamSynchronize(ex.st);
      cudaMemcpy(h, p.trace, sizeof(h), cudaMemcpyDeviceToHost);
      cudaFree(p.trace);
      fprintf(stderr, "conv_tc trace Q %d cin %d nt %d tiles %d grid %d stages %d: setup %lld pdl_wait %lld first_issue %lld first_full %lld mma_done %lld epi_done %lld exit %lld | mma warp waited: accumulator %lld operands %lld | epilogue group 0: waited %lld busy %lld\n",
              a.Q, w.cin, a.nt, total, grid, st, h[1] - h[0], h[2] - h[0], h[7] - h[0], h[3] - h[0], h[4] - h[0], h[5] - h[0], h[6] - h[0], h[8], h[9],
              h[10], h[11]);
    }
    if (const cudaError_t le = cudaPeekAtLastError(); le != cudaSuccess) {
      cudaFuncAttributes fa{};
      cudaFuncGetAttributes(&fa, conv_tc_persist_kernel);
      throw std::runtime_error(std::string("cuda: conv_tc_persist_kernel launch failed: ") + cudaGetErrorString(le) + " (threads " +
                               std::to_string(TCP_THREADS) + ", dynamic smem " + std::to_string(smem) + ", static smem " +
                               std::to_string(fa.sharedSizeBytes) + ", registers " + std::to_string(fa.numRegs) + ", stages " +
                               std::to_string(st) + ", stage bytes " + std::to_string(sb) + ")");
    }
    ++ex.launches;
    return;
  }
  if (a.in_planar) throw std::runtime_error("conv_tc: a group-planar input needs the persistent kernel");
  const CUtensorMap tb = make_map(w.w, (uint64_t)w.cin_pad, (uint64_t)w.phases * w.taps * w.cout, (uint64_t)w.cin_pad,
                                  w.kb, a.nt, a.f16);
  const int span = (w.taps - 1) * (a.tap_step < 0 ? -a.tap_step : a.tap_step);
  static const int window_env = [] { const char* e = getenv("PIPER_B200_TC_WINDOW"); return e ? atoi(e) : 1; }();
  p.window = (a.window_mode != 0 && window_env != 0 && w.taps > 1 && span <= TC_BM) ? window_env : 0;
  if (p.window == 1 && a.tap_step % 8 != 0) p.window = 0;   // swizzled windows: only atom-aligned shifts are fast
  if (p.window == 2 && (a.ldin % 8 != 0 || a.cin % 8 != 0)) p.window = 0;
  const int a_ring = p.window ? 4 * TC_BM * w.kb * 2 : 0;
  const int stage_bytes = (p.window ? 0 : TC_BM * w.kb * 2) + a.n_groups * a.nt * w.kb * 2;
  int stages = TC_STAGES;
  while (stages > 2 && a_ring + stages * stage_bytes + 2048 > TC_SMEM_MAX) --stages;
  // two CTAs per SM (one's epilogue under the other's main loop) beat a deeper pipeline
  static const int min_stages = [] { const char* e = getenv("PIPER_B200_TC_MINSTAGES"); return e ? atoi(e) : 2; }();
  while (stages > min_stages && 2 * (a_ring + stages * stage_bytes + 2048) > TC_SMEM_MAX) --stages;
  const int n_it = w.taps * p.kblocks;
  if (stages > n_it) stages = n_it < 2 ? 2 : n_it;
  const size_t smem = (size_t)a_ring + (size_t)stages * stage_bytes + 1024 + 256;
  if (smem > (size_t)TC_SMEM_MAX) throw std::runtime_error("conv_tc: tile does not fit in shared memory");
  static PerDeviceOnce once_tile;
  once_tile.run([&] {
    if (cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess)
      throw std::runtime_error("cuda: cannot raise dynamic shared memory limit for conv_tc_kernel");
  });
  CUtensorMap tans = ta;
  if (p.window == 2) tans = make_map_ns(a.in, (uint64_t)a.cin, (uint64_t)a.R_in, (uint64_t)a.ldin, 2 * TC_BM, w.kb / 8, a.f16);
  dim3 grid((a.Q + TC_BM - 1) / TC_BM, a.n_tiles, w.phases);
  ProfScope ps(ex, Prof::CONV_TC, 2.0 * a.Q * w.phases * w.taps * w.cin * (double)(a.n_groups * a.nt * a.n_tiles));
  conv_tc_kernel<<<grid, TC_THREADS, smem, ex.st>>>(ta, tb, tans, p, stages);
  ++ex.launches;
}

// ---- weight repack + activation cast ---------------------------------------------------------
namespace {
__global__ void pack_tc_weights_kernel(const float* __restrict__ w, int ldw, int cin, int cout, uint16_t* __restrict__ dst,
                                       int cout_total, int row_off, int cin_pad, int f16, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int ci = (int)(i % cin_pad);
  const int co = (int)((i / cin_pad) % cout);
  const int pt = (int)(i / ((int64_t)cin_pad * cout));
  float v = 0.f;
  if (ci < cin) v = w[((size_t)pt * cin + ci) * ldw + co];
  uint16_t bits;
  if (f16) {
    const __half h = __float2half_rn(v);
    bits = *reinterpret_cast<const uint16_t*>(&h);
  } else {
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    bits = *reinterpret_cast<const uint16_t*>(&h);
  }
  dst[((size_t)pt * cout_total + row_off + co) * cin_pad + ci] = bits;
}

// weights of a split convolution: row (pt, co) holds the planes [B2 B1 B0 B1 B0 B0] of cin_pad channels each
__global__ void pack_tc_weights_split_kernel(const float* __restrict__ w, int ldw, int cin, int cout,
                                             uint16_t* __restrict__ dst, int cout_total, int row_off, int cin_pad, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int ci = (int)(i % cin_pad);
  const int co = (int)((i / cin_pad) % cout);
  const int pt = (int)(i / ((int64_t)cin_pad * cout));
  float v = 0.f;
  if (ci < cin) v = w[((size_t)pt * cin + ci) * ldw + co];
  uint16_t b[3];
  split3(v, b[0], b[1], b[2]);
  uint16_t* row = dst + ((size_t)pt * cout_total + row_off + co) * (size_t)(6 * cin_pad) + ci;
  row[0] = b[2]; row[cin_pad] = b[1]; row[2 * cin_pad] = b[0];
  row[3 * cin_pad] = b[1]; row[4 * cin_pad] = b[0]; row[5 * cin_pad] = b[0];
}

__global__ void split3_kernel(const float* __restrict__ src, int ldsrc, int64_t R, int C4, uint16_t* __restrict__ dst, int cpad) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= R * C4) return;
  const int64_t r = i / C4;
  const int c = (int)(i - r * C4) * 4;
  const float4 x = *reinterpret_cast<const float4*>(src + r * ldsrc + c);
  uint2 t0, t1, t2;
  split3x4(x, t0, t1, t2);
  uint16_t* o = dst + r * (int64_t)(3 * cpad) + c;
  *reinterpret_cast<uint2*>(o) = t0;
  *reinterpret_cast<uint2*>(o + cpad) = t1;
  *reinterpret_cast<uint2*>(o + 2 * cpad) = t2;
}

__global__ void pack_tc_weights_split2_kernel(const float* __restrict__ w, int ldw, int cin, int cout,
                                              uint16_t* __restrict__ dst, int cout_total, int row_off, int cin_pad, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int ci = (int)(i % cin_pad);
  const int co = (int)((i / cin_pad) % cout);
  const int pt = (int)(i / ((int64_t)cin_pad * cout));
  float v = 0.f;
  if (ci < cin) v = w[((size_t)pt * cin + ci) * ldw + co];
  uint16_t hi, lo;
  split2h(v, hi, lo);
  uint16_t* row = dst + ((size_t)pt * cout_total + row_off + co) * (size_t)(2 * cin_pad) + ci;
  row[0] = hi; row[cin_pad] = lo;
}

__global__ void split2_kernel(const float* __restrict__ src, int ldsrc, int64_t R, int C4, uint16_t* __restrict__ dst, int cpad) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= R * C4) return;
  const int64_t r = i / C4;
  const int c = (int)(i - r * C4) * 4;
  const float4 x = *reinterpret_cast<const float4*>(src + r * ldsrc + c);
  uint2 t0, t1;
  split2x4(x, t0, t1);
  uint16_t* o = dst + r * (int64_t)(2 * cpad) + c;
  *reinterpret_cast<uint2*>(o) = t0;
  *reinterpret_cast<uint2*>(o + cpad) = t1;
}

__global__ void to_act16_kernel(const float* __restrict__ src, uint32_t* __restrict__ dst, int64_t n2, float slope,
                                int f16) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n2) return;
  const float2 v = reinterpret_cast<const float2*>(src)[i];
  dst[i] = pack2(lrelu(v.x, slope), lrelu(v.y, slope), f16);
}
}  // namespace

void launch_pack_tc_weights(Exec& ex, const float* w_f32, int ldw, int phases_taps, int cin, int cout, void* dst,
                            int cout_total, int row_off, int cin_pad, int f16) {
  const int64_t n = (int64_t)phases_taps * cout * cin_pad;
  if (n <= 0) return;
  pack_tc_weights_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ex.st>>>(w_f32, ldw, cin, cout,
                                                                        static_cast<uint16_t*>(dst), cout_total, row_off,
                                                                        cin_pad, f16, n);
  ++ex.launches;
}

void launch_pack_tc_weights_split(Exec& ex, const float* w_f32, int ldw, int phases_taps, int cin, int cout, void* dst,
                                  int cout_total, int row_off, int cin_pad) {
  const int64_t n = (int64_t)phases_taps * cout * cin_pad;
  if (n <= 0) return;
  pack_tc_weights_split_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ex.st>>>(w_f32, ldw, cin, cout,
                                                                              static_cast<uint16_t*>(dst), cout_total,
                                                                              row_off, cin_pad, n);
  ++ex.launches;
}

void launch_pack_tc_weights_split2(Exec& ex, const float* w_f32, int ldw, int phases_taps, int cin, int cout, void* dst,
                                   int cout_total, int row_off, int cin_pad) {
  const int64_t n = (int64_t)phases_taps * cout * cin_pad;
  if (n <= 0) return;
  pack_tc_weights_split2_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ex.st>>>(w_f32, ldw, cin, cout,
                                                                               static_cast<uint16_t*>(dst), cout_total,
                                                                               row_off, cin_pad, n);
  ++ex.launches;
}

void launch_split2(Exec& ex, const float* src, int ldsrc, int64_t R, int C, void* dst, int cpad) {
  const int64_t n = R * (C / 4);
  if (n <= 0) return;
  split2_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ex.st>>>(src, ldsrc, R, C / 4, static_cast<uint16_t*>(dst), cpad);
  ++ex.launches;
}

__global__ void fold_1x1_kernel(const float* __restrict__ A, int lda, const float* __restrict__ ba,
                                const float* __restrict__ B, int ldb, int n, int k, int cout, float* __restrict__ out,
                                int ldo, int row_off, float* __restrict__ bias_out) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < n * cout) {
    const int ci = idx / cout, co = idx - ci * cout;
    double acc = 0.0;
    for (int m = 0; m < k; ++m) acc += (double)A[(size_t)ci * lda + m] * (double)B[(size_t)m * ldb + co];
    out[(size_t)(row_off + ci) * ldo + co] = (float)acc;
  } else if (bias_out && idx < (n + 1) * cout) {
    const int co = idx - n * cout;
    double acc = 0.0;
    for (int m = 0; m < k; ++m) acc += (double)(ba ? ba[m] : 0.f) * (double)B[(size_t)m * ldb + co];
    bias_out[co] += (float)acc;
  }
}

void launch_fold_1x1(Exec& ex, const float* A, int lda, const float* ba, const float* B, int ldb, int n, int k, int cout,
                     float* out, int ldo, int row_off, float* bias_out) {
  const int total = (n + 1) * cout;
  fold_1x1_kernel<<<(total + 255) / 256, 256, 0, ex.st>>>(A, lda, ba, B, ldb, n, k, cout, out, ldo, row_off, bias_out);
}

void launch_split3(Exec& ex, const float* src, int ldsrc, int64_t R, int C, void* dst, int cpad) {
  if (C % 4 || ldsrc % 4 || cpad % 4) throw std::runtime_error("split3: channel counts must be multiples of 4");
  const int64_t n = R * (C / 4);
  if (n <= 0) return;
  split3_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ex.st>>>(src, ldsrc, R, C / 4, static_cast<uint16_t*>(dst), cpad);
  ++ex.launches;
}

void launch_to_act16(Exec& ex, const float* src, void* dst, int64_t n, float slope, int f16) {
  const int64_t n2 = n / 2;
  if (n2 <= 0) return;
  to_act16_kernel<<<(unsigned)((n2 + 255) / 256), 256, 0, ex.st>>>(src, static_cast<uint32_t*>(dst), n2, slope, f16);
  ++ex.launches;
}

}  // namespace piper
