// Tensor-core backward of the window-12 attention (bf16, window-aligned tensors, analytic SW-MSA mask): the ODA / timm
// Swin-L/384 sibling of SURVEY.md §8f row 3 and the window-12 rows of the microbench sweep.  Round 1 served this shape on
// the CUDA cores at 0.013 of the HBM roofline (2.4 ms per stage-0 launch).
//
// N = 144 tokens = 9 x 16: the window fits the warp-level m16n16k16 bf16 MMA (`mma.sync`, nvcuda::wmma) without any
// padding, which the 128-lane tcgen05 accumulator does not (144 = 128 + 16 or 72 + 72: 44 % of the tile wasted, and the
// S + dP + dQ/dK/dV accumulators of one unit already fill tensor memory).  This is therefore a
// warp-MMA kernel on purpose; it is not the tcgen05 path and is labelled as such in DESIGN.md.
//
// One CTA = one job = (head, contiguous range of windows), 9 warps; warp w owns the 16-row strip w of every product:
//   S strip  = Q_w K^T              (9 tiles x 2 k-steps)  -> fp32 strip in shared memory (warp-private)
//   P        = exp2(S c + bias (+ mask) - lse)              -> bf16 rows of the P image
//   dP strip = dO_w V^T                                     -> the same fp32 strip
//   delta    = sum_j P dP (warp shuffles), dS = P (dP - delta) -> bf16 rows of the dS image
//   -- CTA barrier: the products below read every strip --
//   dV_w = (P^T)_w dO     dK_w = scale (dS^T)_w Q     dQ_w = scale dS_w K      (9 k-steps x 2 tiles each)
//   d_table bins += dS by relative position: one thread per bin gathers its diagonal (no atomics; the bins live in
//   shared memory for the whole job and are added to global memory once per CTA)
// q, k, v, dO rows are gathered from / dq, dk, dv rows scattered to the image-order tensors with the cyclic shift folded
// into the token coordinates (newcrf_layers.py:220-227 / swin_transformer.py:212-219 and their inverses).
#include <mma.h>

#include "swin_common.cuh"

namespace swin {
namespace mma12 {

using namespace nvcuda;

constexpr int N = 144, WS = 12, HD = 32, NT = 23 * 23;
constexpr int kWarps = 9, kThreads = kWarps * 32;
constexpr int kTileLd = 40;            // bf16 elements per row of the q / k / v / dO tiles (80 B: conflict-free ldmatrix rows)
constexpr int kImgLd = 152;            // bf16 elements per row of the P / dS images
constexpr int kStripLd = 148;          // fp32 elements per row of a warp's S / dP strip
constexpr float kLog2e = 1.4426950408889634f;

struct Smem {
  __nv_bfloat16 q[N * kTileLd], k[N * kTileLd], v[N * kTileLd], d_o[N * kTileLd];
  __nv_bfloat16 p[N * kImgLd], ds[N * kImgLd];
  float strip[kWarps][16 * kStripLd];
  float bias[NT + 3], bins[NT + 3];
  float lse2[N];
  unsigned char region[N];
};

// token t of window (wy, wx) -> pixel of the (window-aligned) image, cyclic shift folded in (compile-time window size)
__device__ __forceinline__ int64_t token_pixel(const Problem& p, int b, int wy, int wx, int t) {
  const int i = t / WS, j = t - i * WS;
  int y = wy * WS + i + p.shift, x = wx * WS + j + p.shift;
  if (y >= p.H) y -= p.H;
  if (x >= p.W) x -= p.W;
  return ((int64_t)b * p.H + y) * p.W + x;
}

__global__ void __launch_bounds__(kThreads, 1) mma12_bwd_kernel(const Problem p, int n_win, int win_per_job) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int h = (int)blockIdx.x % p.nH;
  const int w0 = ((int)blockIdx.x / p.nH) * win_per_job, w1 = min(w0 + win_per_job, n_win);
  const float c1 = p.scale * kLog2e;

  for (int i = tid; i < NT; i += kThreads) {
    sm.bias[i] = p.table[i * p.nH + h] * kLog2e;
    sm.bins[i] = 0.f;
  }
  float* strip = sm.strip[warp];
  const int r0 = warp * 16;

  // The four tiles of a window (144 rows x 64 B each = 2304 16-byte chunks = 8 per thread) and its lse row are fetched
  // into REGISTERS one window ahead, so that the global-load latency hides behind the previous window's math.
  uint4 pre[8];
  float lse_pre = 0.f;
  auto fetch = [&](int w) {
    const int wx = w % p.nWx, wy = (w / p.nWx) % p.nWy, b = w / (p.nWx * p.nWy);
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = tid + u * kThreads, ch = i & 3, t = (i >> 2) % N, which = i / (4 * N);
      const int64_t pix = token_pixel(p, b, wy, wx, t);
      const __nv_bfloat16* src = which == 0   ? static_cast<const __nv_bfloat16*>(p.q) + pix * p.q_ps
                                 : which == 1 ? static_cast<const __nv_bfloat16*>(p.k) + pix * p.k_ps
                                 : which == 2 ? static_cast<const __nv_bfloat16*>(p.v) + pix * p.v_ps
                                              : static_cast<const __nv_bfloat16*>(p.d_o) + pix * p.C;
      pre[u] = __ldg(reinterpret_cast<const uint4*>(src + h * HD + ch * 8));
    }
    if (tid < N) lse_pre = __ldg(&p.lse[token_pixel(p, b, wy, wx, tid) * p.nH + h]);
  };
  if (w0 < w1) fetch(w0);

  for (int w = w0; w < w1; ++w) {
    const int wx = w % p.nWx, wy = (w / p.nWx) % p.nWy, b = w / (p.nWx * p.nWy);
    const bool masked = p.shift > 0 && (wy == p.nWy - 1 || wx == p.nWx - 1);
    __syncthreads();   // the previous window's images, tiles and bins gather are done
    // ---- the prefetched tiles, lse and the SW-MSA region ids -> shared memory; then the next window's fetch is issued ----
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = tid + u * kThreads, ch = i & 3, t = (i >> 2) % N, which = i / (4 * N);
      __nv_bfloat16* dst = (which == 0 ? sm.q : which == 1 ? sm.k : which == 2 ? sm.v : sm.d_o) + t * kTileLd + ch * 8;
      *reinterpret_cast<uint4*>(dst) = pre[u];
    }
    if (tid < N) {
      sm.lse2[tid] = lse_pre * kLog2e;
      sm.region[tid] = masked ? (unsigned char)region_id(p, wy, wx, tid) : 0;
    }
    __syncthreads();
    if (w + 1 < w1) fetch(w + 1);

    // ---- S strip = Q_w K^T ----
    {
      wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::row_major> a[2];
      wmma::load_matrix_sync(a[0], sm.q + r0 * kTileLd, kTileLd);
      wmma::load_matrix_sync(a[1], sm.q + r0 * kTileLd + 16, kTileLd);
#pragma unroll
      for (int jt = 0; jt < 9; ++jt) {
        wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc;
        wmma::fill_fragment(acc, 0.f);
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::col_major> bk;   // B[d][j] = K[j][d]
          wmma::load_matrix_sync(bk, sm.k + jt * 16 * kTileLd + kk * 16, kTileLd);
          wmma::mma_sync(acc, a[kk], bk, acc);
        }
        wmma::store_matrix_sync(strip + jt * 16, acc, kStripLd, wmma::mem_row_major);
      }
    }
    __syncwarp();
    // ---- P = exp2(S c + bias (+ mask) - lse) -> bf16 P rows; kept in the strip as fp32 for dS ----
    int cj[5], rg[5];   // this lane's columns j = lane + 32 u: relative-index offset jy * 23 + jx and SW-MSA region
#pragma unroll
    for (int u = 0; u < 5; ++u) {
      const int j = min(lane + 32 * u, N - 1), jy = j / WS;
      cj[u] = jy * (2 * WS - 1) + (j - jy * WS);
      rg[u] = sm.region[j];
    }
    for (int rr = 0; rr < 16; ++rr) {
      const int i = r0 + rr, iy = i / WS;
      const float* brow = sm.bias + iy * (2 * WS - 1) + (i - iy * WS) + (WS - 1) * 2 * WS;   // index = brow_off - cj
      const float nl = -sm.lse2[i];
      const int ri = sm.region[i];
#pragma unroll
      for (int u = 0; u < 5; ++u) {
        const int j = lane + 32 * u;
        if (j < N) {
          float t = fmaf(strip[rr * kStripLd + j], c1, brow[-cj[u]] + nl);
          if (masked && rg[u] != ri) t += -100.0f * kLog2e;
          float pr;
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(pr) : "f"(t));
          sm.p[i * kImgLd + j] = __float2bfloat16_rn(pr);
        }
      }
    }
    __syncwarp();
    // ---- dP strip = dO_w V^T ----
    {
      wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::row_major> a[2];
      wmma::load_matrix_sync(a[0], sm.d_o + r0 * kTileLd, kTileLd);
      wmma::load_matrix_sync(a[1], sm.d_o + r0 * kTileLd + 16, kTileLd);
#pragma unroll
      for (int jt = 0; jt < 9; ++jt) {
        wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc;
        wmma::fill_fragment(acc, 0.f);
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::col_major> bv;
          wmma::load_matrix_sync(bv, sm.v + jt * 16 * kTileLd + kk * 16, kTileLd);
          wmma::mma_sync(acc, a[kk], bv, acc);
        }
        wmma::store_matrix_sync(strip + jt * 16, acc, kStripLd, wmma::mem_row_major);
      }
    }
    __syncwarp();
    // ---- delta = sum_j P dP ; dS = P (dP - delta) -> bf16 dS rows ----
    for (int rr = 0; rr < 16; ++rr) {
      const int i = r0 + rr;
      float pv[5], dv[5], acc = 0.f;
#pragma unroll
      for (int u = 0; u < 5; ++u) {
        const int j = lane + 32 * u;
        pv[u] = j < N ? __bfloat162float(sm.p[i * kImgLd + j]) : 0.f;
        dv[u] = j < N ? strip[rr * kStripLd + j] : 0.f;
        acc = fmaf(pv[u], dv[u], acc);
      }
#pragma unroll
      for (int off = 16; off; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
#pragma unroll
      for (int u = 0; u < 5; ++u) {
        const int j = lane + 32 * u;
        if (j < N) sm.ds[i * kImgLd + j] = __float2bfloat16_rn(pv[u] * (dv[u] - acc));
      }
    }
    __syncthreads();   // every strip of P and dS is written

    // ---- dV_w = (P^T)_w dO, dK_w = (dS^T)_w Q, dQ_w = dS_w K: 16 x 32 each, staged in the warp's strip ----
#pragma unroll
    for (int which = 0; which < 3; ++which) {
      const __nv_bfloat16* img = which == 0 ? sm.p : sm.ds;
      const __nv_bfloat16* rhs = which == 0 ? sm.d_o : which == 1 ? sm.q : sm.k;
      wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc[2];
      wmma::fill_fragment(acc[0], 0.f);
      wmma::fill_fragment(acc[1], 0.f);
#pragma unroll
      for (int kt = 0; kt < 9; ++kt) {
        wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::row_major> b0, b1;
        wmma::load_matrix_sync(b0, rhs + kt * 16 * kTileLd, kTileLd);
        wmma::load_matrix_sync(b1, rhs + kt * 16 * kTileLd + 16, kTileLd);
        if (which < 2) {   // A[j][i] = img[i][j]: column-major view of the image, rows = this warp's 16 keys
          wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::col_major> a;
          wmma::load_matrix_sync(a, img + kt * 16 * kImgLd + r0, kImgLd);
          wmma::mma_sync(acc[0], a, b0, acc[0]);
          wmma::mma_sync(acc[1], a, b1, acc[1]);
        } else {
          wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::row_major> a;
          wmma::load_matrix_sync(a, img + r0 * kImgLd + kt * 16, kImgLd);
          wmma::mma_sync(acc[0], a, b0, acc[0]);
          wmma::mma_sync(acc[1], a, b1, acc[1]);
        }
      }
      wmma::store_matrix_sync(strip + which * 32, acc[0], kStripLd, wmma::mem_row_major);
      wmma::store_matrix_sync(strip + which * 32 + 16, acc[1], kStripLd, wmma::mem_row_major);
    }
    __syncwarp();
    // rows of dV | dK | dQ (fp32, columns [0,32) [32,64) [64,96) of the strip) -> bf16 -> the tokens' pixels
    for (int i = lane; i < 16 * 3 * 4; i += 32) {
      const int ch = i & 3, which = (i >> 2) % 3, rr = i / 12;
      const int64_t pix = token_pixel(p, b, wy, wx, r0 + rr);
      const float sc = which == 0 ? 1.0f : p.scale;
      const float* src = strip + rr * kStripLd + which * 32 + ch * 8;
      __align__(16) __nv_bfloat16 out[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) out[e] = __float2bfloat16_rn(src[e] * sc);
      __nv_bfloat16* dst = which == 0   ? static_cast<__nv_bfloat16*>(p.dv) + pix * p.v_ps
                           : which == 1 ? static_cast<__nv_bfloat16*>(p.dk) + pix * p.k_ps
                                        : static_cast<__nv_bfloat16*>(p.dq) + pix * p.q_ps;
      *reinterpret_cast<uint4*>(dst + h * HD + ch * 8) = *reinterpret_cast<const uint4*>(out);
    }
    // ---- bias-table gradient: bin (di, dj) += sum over keys (cy, cx) of dS[(cy + di, cx + dj)][(cy, cx)], in two steps.
    // (1) Two threads per (query row qy, key row ky): each takes 6 of the 12 query columns of that 12 x 12 block of dS and
    //     sums it along its 23 x-diagonals in registers (fully unrolled: no divergence, no atomics).
    __syncthreads();   // every warp has written its output rows: the strips are free and become the diagonal sums,
    float* diag = &sm.strip[0][0];   // [half][query row qy][key row ky][24]: dS summed over the x diagonals of a 12 x 12 block
    {
      const int pair = tid >> 1, half = tid & 1, qy = pair / WS, ky = pair - qy * WS;
      float dg[17];   // local diagonals e = qx6 - kx in [-11, 5], stored at e + 11
#pragma unroll
      for (int d = 0; d < 17; ++d) dg[d] = 0.f;
      const __nv_bfloat16* blk = sm.ds + (qy * WS + half * 6) * kImgLd + ky * WS;
#pragma unroll
      for (int qx6 = 0; qx6 < 6; ++qx6) {
#pragma unroll
        for (int kx = 0; kx < WS; kx += 2) {
          const float2 v2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(blk + qx6 * kImgLd + kx));
          dg[qx6 - kx + 11] += v2.x;
          dg[qx6 - (kx + 1) + 11] += v2.y;
        }
      }
      // this thread's query columns are qx = half * 6 + qx6, so the true diagonal is dj = e + 6 half
      float* out = diag + ((half * N + pair) * 24);
#pragma unroll
      for (int d = 0; d < 17; ++d) out[d] = dg[d];
    }
    __syncthreads();
    // (2) one thread per bin: <= 12 key rows x the two halves
    for (int bin = tid; bin < NT; bin += kThreads) {
      const int di = bin / (2 * WS - 1) - (WS - 1), dj = bin % (2 * WS - 1) - (WS - 1);
      float sacc = 0.f;
      for (int ky = max(0, -di); ky < min(WS, WS - di); ++ky) {
        const int pair = (ky + di) * WS + ky;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int d = dj - half * 6 + 11;           // index into that half's 17 local diagonals
          if (d >= 0 && d < 17) sacc += diag[(half * N + pair) * 24 + d];
        }
      }
      sm.bins[bin] += sacc;
    }
  }
  __syncthreads();
  for (int i = tid; i < NT; i += kThreads) atomicAdd(&p.d_table[i * p.nH + h], sm.bins[i]);
}

}  // namespace mma12

bool mma12_backward_supported(const Problem& p, int dtype) {
  if (dtype != SWIN_ATTN_BF16 || p.ws != 12 || p.mask != nullptr || p.dropout_p > 0.0f) return false;
  if (p.H != p.Hp || p.W != p.Wp) return false;           // window-aligned tensors (what the op differentiates)
  if ((p.q_ps | p.k_ps | p.v_ps | p.C) & 7) return false;
  auto mis = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) != 0; };
  return !(mis(p.q) || mis(p.k) || mis(p.v) || mis(p.d_o) || mis(p.dq) || mis(p.dk) || mis(p.dv));
}

cudaError_t launch_mma12_backward(const Problem& p, cudaStream_t st) {
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int smem = (int)sizeof(mma12::Smem);
  cudaError_t e = cudaFuncSetAttribute(mma12::mma12_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  const int n_win = p.B * p.nWy * p.nWx;
  int chunks = sms / p.nH;
  if (chunks < 1) chunks = 1;
  if (chunks > n_win) chunks = n_win;
  const int per = (n_win + chunks - 1) / chunks;
  chunks = (n_win + per - 1) / per;
  TimedRegion timed(st);
  e = cudaMemsetAsync(p.d_table, 0, sizeof(float) * mma12::NT * p.nH, st);
  if (e != cudaSuccess) return e;
  mma12::mma12_bwd_kernel<<<chunks * p.nH, mma12::kThreads, smem, st>>>(p, n_win, per);
  count_launch();
  return cudaGetLastError();
}

}  // namespace swin
