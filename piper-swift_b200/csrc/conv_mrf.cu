// Fused MRF stage (see conv_mrf.h).  Persistent kernel, one CTA per SM, 320 threads:
//   warp 0     : TMA producer - the 16-bit lrelu(x) window ((m+1) boxes of 128 rows, halo included)
//                and the weight taps (ring of C x C tiles; resident for the whole kernel when they fit)
//   warp 1     : TMEM allocation + the single thread issuing tcgen05.mma.  conv1 of a resblock
//                accumulates tile i into D_i (fresh), conv2 accumulates into S_i on top of sum_r y1_r;
//                tap shifts are descriptor row offsets into the swizzled operand windows
//   warps 2..9 : two epilogue warpgroups (tiles alternate).  After conv1: y1 = D + x + b1 (fp32),
//                S += y1 (TMEM read-modify-write), lrelu(y1)*mask -> 16-bit -> swizzled smem operand
//                of conv2.  After the last resblock: out = (S + sum b2) / nk -> global.
// A window of W = 128*m rows yields V = W - 2*HALO valid rows (HALO = max_r pad0_r + pad1_r):
// rows closer than HALO to a window edge see truncated receptive fields and are recomputed by
// the neighbouring window.
#include "conv_mrf.h"

#include <algorithm>
#include <cstdlib>

#include "tc_common.cuh"

namespace piper {

using namespace tc;

namespace {

constexpr int MRF_G = 64;          // guard rows on both sides of the operand windows (>= largest tap shift)
constexpr int MRF_THREADS = 320;
constexpr int MRF_MAX_M = 8;
constexpr int MRF_MAX_RING = 32;

struct MrfParams {
  int R, C, nk, m, V, halo, n_windows;
  int k[MRF_MAX_RB], d0[MRF_MAX_RB], d1[MRF_MAX_RB];
  int tap_base[MRF_MAX_RB][2];
  const float* b1[MRF_MAX_RB];
  const float* b2[MRF_MAX_RB];
  const float* x;
  const uint16_t* x_act; int64_t act_rows; int act_pad;   // group-planar 16-bit lrelu(x): [C/8][act_rows][8]
  const int* row_utt; int valid_div;
  void* out_act; float act_slope;
  float* out_f32;
  int f16, ring, resident;
  uint32_t tmem_cols;
};

template <int C>
__global__ void __launch_bounds__(MRF_THREADS, 1) mrf_tc_kernel(const __grid_constant__ CUtensorMap tmap_w,
                                                                const MrfParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int m = p.m;
  // operand windows are group-planar without swizzle: [C/8 groups][(m+1)*128 rows][16 bytes]; a tap shift is
  // a 16-byte-granular descriptor offset (row-shifted descriptors into SWIZZLED tiles are ~100x slower
  // unless the shift is a multiple of the 8-row swizzle atom - measured, profiles/r01_tc_align_probe.log)
  constexpr int KS = C / 16;                          // tcgen05 K steps per tap
  constexpr int NP = C / 8;                           // planes
  constexpr int tap_bytes = C * C * 2;
  const int win_rows = (m + 1) * 128;                 // W + 2*G rows
  const int plane_bytes = win_rows * 16;
  const int win_bytes = NP * plane_bytes;
  uint8_t* X0 = smem;
  uint8_t* Y1 = smem + win_bytes;
  uint8_t* WR = Y1 + win_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(WR + (size_t)p.ring * tap_bytes);
  uint64_t* x_full = bars;
  uint64_t* x_empty = bars + 1;
  uint64_t* w_full = bars + 2;
  uint64_t* w_empty = w_full + MRF_MAX_RING;
  uint64_t* d_full = w_empty + MRF_MAX_RING;
  uint64_t* d_empty = d_full + MRF_MAX_M;
  uint64_t* y_ready = d_empty + MRF_MAX_M;
  uint64_t* s_full = y_ready + MRF_MAX_M;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_full + MRF_MAX_M);
  float* sbias = reinterpret_cast<float*>(tmem_slot + 4);   // [nk][C] b1 then [C] sum_r b2

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // zero the conv2 operand window once (its guard rows are never written afterwards)
  for (int i = threadIdx.x; i < win_bytes / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(Y1)[i] = make_uint4(0u, 0u, 0u, 0u);
  for (int i = threadIdx.x; i < (p.nk + 1) * C; i += blockDim.x) {
    const int r = i / C, c = i - r * C;
    float v = 0.f;
    if (r < p.nk) v = p.b1[r][c];
    else for (int q = 0; q < p.nk; ++q) v += p.b2[q][c];
    sbias[i] = v;
  }
  fence_proxy_async();
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_w);
    mbar_init(x_full, 1);
    mbar_init(x_empty, 1);
    for (int s = 0; s < p.ring; ++s) {
      mbar_init(&w_full[s], 1);
      mbar_init(&w_empty[s], 1);
    }
    for (int i = 0; i < m; ++i) {
      mbar_init(&d_full[i], 1);
      mbar_init(&d_empty[i], 4);
      mbar_init(&y_ready[i], 4);
      mbar_init(&s_full[i], 1);
    }
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc(tmem_slot, p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t d_col = 0, s_col = (uint32_t)(m * C);

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      uint32_t slot = 0, wph = 0;
      int it = 0;
      for (int win = blockIdx.x; win < p.n_windows; win += gridDim.x, ++it) {
        const int w0 = win * p.V - p.halo;
        mbar_wait(x_empty, ((uint32_t)it & 1u) ^ 1u);
        mbar_expect_tx(x_full, (uint32_t)win_bytes);
        for (int g = 0; g < NP; ++g)
          bulk_load(X0 + (size_t)g * plane_bytes,
                    p.x_act + ((size_t)g * p.act_rows + (size_t)(p.act_pad + w0 - MRF_G)) * 8, (uint32_t)plane_bytes, x_full);
        if (!p.resident || it == 0) {
          for (int r = 0; r < p.nk; ++r)
            for (int j = 0; j < 2; ++j)
              for (int t = 0; t < p.k[r]; ++t) {
                mbar_wait(&w_empty[slot], wph ^ 1u);
                mbar_expect_tx(&w_full[slot], (uint32_t)tap_bytes);
                tma_load_2d(&tmap_w, &w_full[slot], WR + (size_t)slot * tap_bytes, 0, (p.tap_base[r][j] + t) * C);
                if (++slot == (uint32_t)p.ring) { slot = 0; wph ^= 1u; }
              }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the whole warp runs the (warp-uniform) control flow, one elected lane issues =====
    const uint32_t idesc = make_idesc(C, p.f16);
    const uint32_t plane16 = (uint32_t)plane_bytes >> 4;                 // descriptor address units are 16 bytes:
    const uint64_t a_hi = make_smem_desc_ns(0, (uint32_t)plane_bytes, 128);   // one operand row = one unit
    const uint64_t b_hi = make_smem_desc(0, C);
    const uint32_t x0_lo = (smem_u32(X0) >> 4) + MRF_G, y1_lo = (smem_u32(Y1) >> 4) + MRF_G;
    const uint32_t wr_lo = smem_u32(WR) >> 4;
    constexpr uint32_t tap16 = tap_bytes >> 4;
    const uint32_t ring = (uint32_t)p.ring;
    uint32_t slot = 0, wph = 0;      // ring position of the next convolution's first tap (streaming mode)
    int it = 0;
    for (int win = blockIdx.x; win < p.n_windows; win += gridDim.x, ++it) {
      mbar_wait(x_full, (uint32_t)it & 1u);
      const bool wait_w = !p.resident || it == 0;
      for (int r = 0; r < p.nk; ++r) {
        const uint32_t n = (uint32_t)(it * p.nk + r);
        const int k = p.k[r];
        for (int j = 0; j < 2; ++j) {
          const int dil = j ? p.d1[r] : p.d0[r];
          const uint32_t op_lo = (j ? y1_lo : x0_lo) - (uint32_t)(((k - 1) / 2) * dil);
          const uint32_t slot0 = p.resident ? (uint32_t)p.tap_base[r][j] : slot;
          const uint32_t wph0 = wph;
          uint32_t s_end = slot0, ph_end = wph0;
          for (int i = 0; i < m; ++i) {
            if (j == 0) {
              mbar_wait(&d_empty[i], (n & 1u) ^ 1u);           // epilogue of the previous resblock drained D_i
            } else {
              if (i == 0) mbar_wait(&y_ready[0], n & 1u);      // conv2 operand rows of tiles i-1, i, i+1
              if (i + 1 < m) mbar_wait(&y_ready[i + 1], n & 1u);
            }
            tc_fence_after();
            const uint32_t acc = tmem_base + (j ? s_col : d_col) + (uint32_t)(i * C);
            uint32_t a_lo = op_lo + (uint32_t)(128 * i);
            uint32_t s = slot0, ph = wph0;
            for (int t = 0; t < k; ++t) {
              if (i == 0 && wait_w) {
                mbar_wait(&w_full[s], ph);
                tc_fence_after();
              }
              const uint32_t b_lo = wr_lo + s * tap16;
              if (elect_one()) {
#pragma unroll
                for (int ks = 0; ks < KS; ++ks)   // K step of 16 channels = two planes of A, 32 bytes of B
                  umma_f16(acc, a_hi | (uint64_t)(a_lo + (uint32_t)ks * 2u * plane16), b_hi | (uint64_t)(b_lo + 2u * ks),
                           idesc, (j == 1 || t > 0 || ks > 0) ? 1u : 0u);
              }
              a_lo += (uint32_t)dil;
              if (++s == ring) { s = 0; ph ^= 1u; }
            }
            if (elect_one()) umma_commit(j ? &s_full[i] : &d_full[i]);
            s_end = s; ph_end = ph;
          }
          if (!p.resident) {
            uint32_t s = slot0;
            for (int t = 0; t < k; ++t) {
              if (elect_one()) umma_commit(&w_empty[s]);
              if (++s == ring) s = 0;
            }
            slot = s_end; wph = ph_end;
          }
          if (j == 0 && r == p.nk - 1 && elect_one()) umma_commit(x_empty);   // the lrelu(x) window may be refilled
        }
      }
    }
  } else {
    // ===== epilogue warpgroups: warps 2..5 (group 0) and 6..9 (group 1) =====
    const int group = (warp - 2) >> 2;
    const int quarter = warp & 3;                        // TMEM lane quarter this warp may access
    const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
    const float inv_nk = 1.0f / (float)p.nk;
    int it = 0;
    for (int win = blockIdx.x; win < p.n_windows; win += gridDim.x, ++it) {
      const int w0 = win * p.V - p.halo;
      for (int r = 0; r < p.nk; ++r) {
        const uint32_t n = (uint32_t)(it * p.nk + r);
        const float* b1 = sbias + r * C;
        for (int i = group; i < m; i += 2) {
          const int row_in_win = 128 * i + quarter * 32 + lane;
          const int grow = w0 + row_in_win;
          const bool inr = grow >= 0 && grow < p.R;
          // the fp32 residual row does not depend on the MMAs: fetch it before waiting for them
          float4 xr[C / 4];
          if (inr) {
            const float4* x4 = reinterpret_cast<const float4*>(p.x + (size_t)grow * C);
#pragma unroll
            for (int q4 = 0; q4 < C / 4; ++q4) xr[q4] = x4[q4];
          } else {
#pragma unroll
            for (int q4 = 0; q4 < C / 4; ++q4) xr[q4] = make_float4(0.f, 0.f, 0.f, 0.f);
          }
          bool valid = inr;
          if (inr && p.row_utt) valid = p.row_utt[grow / p.valid_div] >= 0;
          if (n > 0) {   // conv2 of the previous resblock must have finished with S_i and with Y1 rows of tile i
            for (int ii = (i > 0 ? i - 1 : 0); ii <= (i + 1 < m ? i + 1 : m - 1); ++ii)
              mbar_wait(&s_full[ii], (n - 1u) & 1u);
          }
          mbar_wait(&d_full[i], n & 1u);
          tc_fence_after();
          const uint32_t row_off = (uint32_t)((MRF_G + row_in_win) * 16);
#pragma unroll
          for (int c0 = 0; c0 < C; c0 += 16) {
            float d[16], s[16];
            __syncwarp();
            tmem_ld16(tmem_base + lane_sel + d_col + (uint32_t)(i * C + c0), d);
#pragma unroll
            for (int e = 0; e < 16; ++e) d[e] += b1[c0 + e];
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) {
              const float4 xv = xr[c0 / 4 + q4];
              d[4 * q4] += xv.x; d[4 * q4 + 1] += xv.y; d[4 * q4 + 2] += xv.z; d[4 * q4 + 3] += xv.w;
            }
            if (r > 0) {
              tmem_ld16(tmem_base + lane_sel + s_col + (uint32_t)(i * C + c0), s);
#pragma unroll
              for (int e = 0; e < 16; ++e) s[e] += d[e];
              tmem_st16(tmem_base + lane_sel + s_col + (uint32_t)(i * C + c0), s);
            } else {
              tmem_st16(tmem_base + lane_sel + s_col + (uint32_t)(i * C + c0), d);
            }
            uint32_t pk[8];
#pragma unroll
            for (int e = 0; e < 8; ++e)
              pk[e] = valid ? pack2(lrelu(d[2 * e], 0.1f), lrelu(d[2 * e + 1], 0.1f), p.f16) : 0u;
            uint8_t* y = Y1 + (size_t)(c0 >> 3) * plane_bytes + row_off;   // planes c0/8 and c0/8 + 1
            *reinterpret_cast<uint4*>(y) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            *reinterpret_cast<uint4*>(y + plane_bytes) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
          }
          tmem_wait_st();
          fence_proxy_async();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(&y_ready[i]);
            mbar_arrive(&d_empty[i]);
          }
        }
      }
      // ---- stage output: (S + sum_r b2_r) / nk over the window's valid rows ----
      const uint32_t nl = (uint32_t)(it * p.nk + p.nk - 1);
      const float* b2s = sbias + p.nk * C;
      for (int i = group; i < m; i += 2) {
        mbar_wait(&s_full[i], nl & 1u);
        tc_fence_after();
        const int row_in_win = 128 * i + quarter * 32 + lane;
        const int grow = w0 + row_in_win;
        const bool owned = row_in_win >= p.halo && row_in_win < p.halo + p.V && grow >= 0 && grow < p.R;
        bool valid = owned;
        if (owned && p.row_utt) valid = p.row_utt[grow / p.valid_div] >= 0;
#pragma unroll
        for (int c0 = 0; c0 < C; c0 += 16) {
          float s[16];
          __syncwarp();
          tmem_ld16(tmem_base + lane_sel + s_col + (uint32_t)(i * C + c0), s);
          if (owned) {
#pragma unroll
            for (int e = 0; e < 16; ++e) s[e] = valid ? (s[e] + b2s[c0 + e]) * inv_nk : 0.f;
            if (p.out_f32) {
              float4* o4 = reinterpret_cast<float4*>(p.out_f32 + (size_t)grow * C + c0);
#pragma unroll
              for (int q4 = 0; q4 < 4; ++q4) o4[q4] = make_float4(s[4 * q4], s[4 * q4 + 1], s[4 * q4 + 2], s[4 * q4 + 3]);
            }
            if (p.out_act) {
              uint32_t pk[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) pk[e] = pack2(lrelu(s[2 * e], p.act_slope), lrelu(s[2 * e + 1], p.act_slope), p.f16);
              uint4* dst = reinterpret_cast<uint4*>(static_cast<uint16_t*>(p.out_act) + (size_t)grow * C + c0);
              dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
              dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            }
          }
        }
        tc_fence_before();
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, p.tmem_cols);
}

}  // namespace

bool mrf_tc_supported(int C, int nk, const int* k, const int* d0, const int* d1) {
  if (C != 32 && C != 64) return false;
  if (nk < 1 || nk > MRF_MAX_RB) return false;
  int taps = 0;
  for (int r = 0; r < nk; ++r) {
    if (k[r] < 1 || (k[r] & 1) == 0) return false;
    const int h = (k[r] - 1) / 2;
    if (h * d0[r] > MRF_G || h * d1[r] > MRF_G) return false;
    if (h * (d0[r] + d1[r]) > 60) return false;          // keeps V = 128*m - 2*halo positive for m >= 2
    taps += 2 * k[r];
  }
  return taps > 0;
}

void launch_mrf_tc(Exec& ex, const MrfArgs& a) {
  if (a.R <= 0) return;
  if (!mrf_tc_supported(a.C, a.nk, a.k, a.d0, a.d1)) throw std::runtime_error("conv_mrf: unsupported stage shape");
  static int num_sms = 0;
  if (!num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || num_sms <= 0)
      throw std::runtime_error("cuda: cannot query the SM count");
    if (cudaFuncSetAttribute(mrf_tc_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess ||
        cudaFuncSetAttribute(mrf_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_MAX) != cudaSuccess)
      throw std::runtime_error("cuda: cannot raise dynamic shared memory limit for mrf_tc_kernel");
  }
  MrfParams p{};
  p.R = a.R; p.C = a.C; p.nk = a.nk;
  int halo = 0, kmax = 0;
  for (int r = 0; r < a.nk; ++r) {
    p.k[r] = a.k[r]; p.d0[r] = a.d0[r]; p.d1[r] = a.d1[r];
    p.tap_base[r][0] = a.tap_base[r][0]; p.tap_base[r][1] = a.tap_base[r][1];
    p.b1[r] = a.b1[r]; p.b2[r] = a.b2[r];
    halo = std::max(halo, (a.k[r] - 1) / 2 * (a.d0[r] + a.d1[r]));
    kmax = std::max(kmax, a.k[r]);
  }
  p.halo = halo;
  const int row_bytes = a.C * 2, tap_bytes = a.C * a.C * 2;
  if (a.act_pad < MRF_G + 64 || a.act_rows < (int64_t)a.R + a.act_pad + 128 * (MRF_MAX_M + 1))
    throw std::runtime_error("conv_mrf: planar operand needs zero pad rows around the signal");
  // window height: TMEM holds D and S for every tile (2*m*C columns <= 512); smem holds two operand
  // windows of (m+1)*128 rows plus at least kmax weight taps
  int m = std::min(MRF_MAX_M, 512 / (2 * a.C));
  auto fits = [&](int mm) { return 2 * (mm + 1) * 128 * row_bytes + kmax * tap_bytes + 4096 <= TC_SMEM_MAX; };
  while (m > 2 && !fits(m)) --m;
  // small inputs: prefer more, shorter windows so that every SM has one
  while (m > 2 && (a.R + (128 * m - 2 * halo) - 1) / (128 * m - 2 * halo) < num_sms) --m;
  static const int m_env = [] { const char* e = getenv("PIPER_B200_MRF_M"); return e ? atoi(e) : 0; }();
  if (m_env >= 2 && m_env <= MRF_MAX_M && 2 * m_env * a.C <= 512 && fits(m_env)) m = m_env;
  if (!fits(m)) throw std::runtime_error("conv_mrf: stage does not fit in shared memory");
  p.m = m;
  p.V = 128 * m - 2 * halo;
  p.n_windows = (a.R + p.V - 1) / p.V;
  const int win_bytes = (m + 1) * 128 * row_bytes;
  int ring = (TC_SMEM_MAX - 4096 - 2 * win_bytes) / tap_bytes;
  if (ring > MRF_MAX_RING) ring = MRF_MAX_RING;
  p.resident = a.total_taps <= ring ? 1 : 0;
  if (p.resident) ring = a.total_taps;
  p.ring = ring;
  p.x_act = static_cast<const uint16_t*>(a.x_act); p.act_rows = a.act_rows; p.act_pad = a.act_pad;
  p.x = a.x; p.row_utt = a.row_utt; p.valid_div = a.valid_div < 1 ? 1 : a.valid_div;
  p.out_act = a.out_act; p.act_slope = a.act_slope; p.out_f32 = a.out_f32;
  p.f16 = a.f16;
  uint32_t cols = 32;
  while ((int)cols < 2 * m * a.C) cols <<= 1;
  p.tmem_cols = cols;
  const CUtensorMap tw = make_map(a.w, (uint64_t)a.C, (uint64_t)a.total_taps * a.C, (uint64_t)a.C, a.C, a.C, a.f16);
  const size_t smem = (size_t)2 * win_bytes + (size_t)ring * tap_bytes + 1024 + 3072;
  if (smem > (size_t)TC_SMEM_MAX) throw std::runtime_error("conv_mrf: shared memory budget exceeded");
  const int grid = std::min(p.n_windows, num_sms);
  ProfScope ps(ex, Prof::MRF, 2.0 * a.R * a.C * a.C * (double)a.total_taps);
  if (a.C == 32) mrf_tc_kernel<32><<<grid, MRF_THREADS, smem, ex.st>>>(tw, p);
  else mrf_tc_kernel<64><<<grid, MRF_THREADS, smem, ex.st>>>(tw, p);
  ++ex.launches;
}

}  // namespace piper
