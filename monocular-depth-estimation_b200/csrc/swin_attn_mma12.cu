// Warp-level tensor-core kernels of the window-12 attention (bf16, analytic SW-MSA mask, window-aligned or un-padded tensors):
// the backward (first) and the forward (second half of this file) of the ODA / timm Swin-L/384 sibling of SURVEY.md §8f row 3,
// of the window-12 NewCRFs variants (NewCRFDepth "large12", NewCRFDepth.py:26) and of the window-12 rows of the microbench
// sweep.  Round 1 served the backward on the CUDA cores at 0.013 of the HBM roofline (2.4 ms per stage-0 launch).
//
// N = 144 tokens = 9 x 16: the window fits the warp-level m16n8k16 bf16 MMA (`mma.sync`) without any padding, which the
// 128-lane tcgen05 accumulator does not (144 = 128 + 16 or 72 + 72: 44 % of the tile wasted, and the S + dP + dQ/dK/dV
// accumulators of one unit already fill tensor memory).  This is therefore a warp-MMA kernel on
// purpose; it is not the tcgen05 path and is labelled as such in DESIGN.md.
//
// One CTA = one job = (head, contiguous range of windows), 18 warps.  Two warps share the 16-row strip s = warp / 2 of
// the query-side products and split its key columns (n-tiles 0-9 / 10-17); everything a strip needs stays in REGISTERS
// (v1 went through fp32 strips in shared memory with wmma load / store: 47 k warp instructions per window, 459 us on
// the ODA stage-0 grid; v2, 9 warps with whole strips in registers: 21 k, 214 us, latency-bound at 2.25 warps per
// scheduler; v3 is this one):
//   delta    = rowsum(dO . O) of the strip's rows, from the dO / O tiles (so no row sum over the other warp's columns)
//   per n-tile (8 keys): S = Q_s K^T and dP = dO_s V^T (two k-steps each) -> P = exp2(S c + bias (+ mask) - lse),
//              dS = P (dP - delta) -> bf16 pairs into the P / dS images; the dS pairs stay in registers: the
//              accumulator layout of two adjacent n-tiles is the A-fragment layout of one k-step of dQ_s = scale dS_s K
//   dQ       = the two warps' partial products, summed through shared memory (lane-major, conflict-free)
//   -- CTA barrier: the products below read every strip --
//   warp 2s: dV_s = (P^T)_s dO     warp 2s+1: dK_s = scale (dS^T)_s Q     (A fragments by ldmatrix.trans from the images)
//   d_table bins += dS by relative position: one thread per bin gathers its diagonal (no atomics; the bins live in
//   shared memory for the whole job and are added to global memory once per CTA)
// Three CTA barriers per window (tiles landed / images written / diagonal sums written) and one 64-thread barrier per strip.
// q, k, v, dO, o rows are gathered (cp.async, one window ahead, two tile buffers; one token per thread) from / dq, dk, dv
// rows scattered to the image-order tensors with the cyclic shift folded into the token coordinates
// (newcrf_layers.py:220-227 / swin_transformer.py:212-219 and their inverses).
#include <type_traits>

#include "swin_common.cuh"

namespace swin {
namespace mma12 {

constexpr int N = 144, WS = 12, HD = 32, NT = 23 * 23;
constexpr int kWarps = 18, kThreads = kWarps * 32;
constexpr int kImgLd = 152;            // bf16 elements per row of the P / dS images (304 B: conflict-free ldmatrix rows)
constexpr int kTile = N * HD;          // q / k / v / dO / o tiles: 64-byte rows, 16-byte chunks XOR-swizzled by the row pair
constexpr int kDiagLd = 2 * N + 1;     // diagonal sums [17][2 N + 1]: conflict-free for the writers (thread-major) and the bins
constexpr float kLog2e = 1.4426950408889634f;

struct Smem {
  __nv_bfloat16 tile[2][5][kTile];     // q, dO, k, v, o of the current and the next window
  __nv_bfloat16 p[N * kImgLd], ds[N * kImgLd];
  float2 bias2[NT + 1];                // {b[NT-1-k], b[NT-2-k]} log2e: the biases of an (even, odd) key pair in one 8-byte load
  float bins[NT + 3];
  float lse[2][N];
  float dqx[9][16][32];                // dQ partial of the odd warp of a strip: [accumulator register][lane]
  float diag[17 * kDiagLd];            // dS summed over the x diagonals of each (query row, key row) block, per half
  float padacc[2][HD];                 // un-padded tensors: dV / dK summed over the pad tokens of this head (-> d_vpad / d_kpad)
};
// byte offset of 16-byte chunk c of row r in a tile: 8 consecutive rows at one logical chunk (ldmatrix) and 2 rows x 4
// chunks (cp.async by 8 threads) both cover all 32 banks
__device__ __forceinline__ int toff(int r, int c) { return r * 64 + ((c ^ ((r >> 1) & 3)) << 4); }

__device__ __forceinline__ uint32_t s_u32(const void* ptr) { return (uint32_t)__cvta_generic_to_shared(ptr); }
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
// D (16 x 8, fp32) += A (16 x 16, row) B (16 x 8, col), bf16 operands
__device__ __forceinline__ void mma_bf16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void* src, int bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// SW-MSA mask of one query row against the key pairs of a thread, as six bits: bit 3 kh + c3 is set when keys with
// (window row >= 6) == kh in the n-tile class c3 = nt % 3 lie in another region than the row.  Keys j = 8 nt + 2 t4 (+ 1):
// their window row (8 nt) / 12 (+ 1 for nt % 3 == 1, t4 >= 2, which never crosses row 6) is >= 6 iff (8 nt) / 12 >= 6, a
// compile-time fact of nt; their window column is >= 6 iff t4 == 3 (c3 = 0), t4 < 2 (c3 = 1), t4 >= 1 (c3 = 2).
__device__ __forceinline__ int mask_bits(bool hm, bool wm, bool my_h, bool my_w, int t4) {
  const bool kw[3] = {t4 == 3, t4 < 2, t4 >= 1};
  int bits = 0;
#pragma unroll
  for (int kh = 0; kh < 2; ++kh)
#pragma unroll
    for (int c3 = 0; c3 < 3; ++c3)
      if ((hm && (kh != 0) != my_h) || (wm && kw[c3] != my_w)) bits |= 1 << (3 * kh + c3);
  return bits;
}
__host__ __device__ constexpr int mask_bit_of(int nt) { return 3 * ((8 * nt) / 12 >= 6 ? 1 : 0) + nt % 3; }

// Token (ty, tx) of window (wy, wx) of image b -> its pixel, cyclic shift folded in; the image is H x W, the windows tile
// the padded Hp x Wp grid: false for a pad token; `lpix` is the pixel in the padded grid, which is how lse is laid out.
__device__ __forceinline__ bool token_pixel_padded(const Problem& p, int b, int wy, int wx, int ty, int tx, int64_t& pix, int64_t& lpix) {
  int y = wy * WS + ty + p.shift, x = wx * WS + tx + p.shift;
  if (y >= p.Hp) y -= p.Hp;
  if (x >= p.Wp) x -= p.Wp;
  pix = ((int64_t)b * p.H + y) * p.W + x;
  lpix = ((int64_t)b * p.Hp + y) * p.Wp + x;
  return y < p.H && x < p.W;
}
struct Win {
  int b, wy, wx;
  __device__ __forceinline__ void next(const Problem& p) {
    if (++wx == p.nWx) {
      wx = 0;
      if (++wy == p.nWy) {
        wy = 0;
        ++b;
      }
    }
  }
};
__device__ __forceinline__ void pair_barrier(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }

__global__ void __launch_bounds__(kThreads, 1) mma12_bwd_kernel(const Problem p, int n_win, int win_per_job) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int strip = warp >> 1, kh = warp & 1;                  // kh: which part of the strip's key columns / which of dV, dK
  const int g = lane >> 2, t4 = lane & 3;                      // accumulator rows g / g + 8, columns 2 t4, 2 t4 + 1 of an n-tile
  const int h = (int)blockIdx.x % p.nH;
  const int w0 = ((int)blockIdx.x / p.nH) * win_per_job, w1 = min(w0 + win_per_job, n_win);
  const float c1 = p.scale * kLog2e;
  const float2 c1c1 = make_float2(c1, c1);

  for (int i = tid; i < NT; i += kThreads) {
    const float b0 = p.table[(NT - 1 - i) * p.nH + h] * kLog2e;
    const float b1 = i + 1 < NT ? p.table[(NT - 2 - i) * p.nH + h] * kLog2e : 0.f;
    sm.bias2[i] = make_float2(b0, b1);
    sm.bins[i] = 0.f;
  }
  if (tid < 2 * HD) (&sm.padacc[0][0])[tid] = 0.f;
  const int r0 = strip * 16;
  // ldmatrix lane addressing.  "A pattern": matrices (rows 0-7, cols 0-7) (rows 8-15, cols 0-7) (rows 0-7, cols 8-15)
  // (rows 8-15, cols 8-15); "T pattern" swaps the middle two.
  const int la_row = (lane & 7) + ((lane >> 3) & 1) * 8;
  const int lt_row = (lane & 7) + ((lane >> 4) & 1) * 8, lt_col = ((lane >> 3) & 1) * 8;
  // tile offsets (bytes) of this lane's ldmatrix row: A pattern at channels [0,16) / [16,32); B of S / dP (8 keys x 32
  // channels = the four k-blocks of one n-tile).  Row bases are multiples of 8 rows, which the swizzle does not see.
  const int offa0 = toff(la_row, (lane >> 4) & 1), offa1 = toff(la_row, 2 + ((lane >> 4) & 1));
  const int offb = toff(lane & 7, lane >> 3);
  // this thread's rows: window coordinates and the row part of the reversed bias index, k = rb + (23 jy + jx)
  const int i0 = r0 + g, i1 = i0 + 8;
  const int iy0 = i0 / WS, ix0 = i0 - iy0 * WS, iy1 = i1 / WS, ix1 = i1 - iy1 * WS;
  const int rb0 = 264 - 23 * iy0 - ix0, rb1 = 264 - 23 * iy1 - ix1;
  const bool my_h0 = iy0 >= 6, my_w0 = ix0 >= 6, my_h1 = iy1 >= 6, my_w1 = ix1 >= 6;
  // keys j = 8 nt + 2 t4 (+ 1) of n-tile nt lie in window row (8 nt) / 12 + (nt % 3 == 1 && t4 >= 2): the bias index
  // rb + 23 jy + jx = rb + j + 11 jy is a per-thread base (two variants) plus a compile-time constant; same for the images
  const int hi = t4 >> 1;
  const float2 *pb0a = sm.bias2 + rb0 + 2 * t4, *pb0b = pb0a + 11 * hi, *pb1a = sm.bias2 + rb1 + 2 * t4, *pb1b = pb1a + 11 * hi;
  __nv_bfloat16 *pimg0 = sm.p + i0 * kImgLd + 2 * t4, *pimg1 = sm.p + i1 * kImgLd + 2 * t4;
  __nv_bfloat16 *simg0 = sm.ds + i0 * kImgLd + 2 * t4, *simg1 = sm.ds + i1 * kImgLd + 2 * t4;

  // The five tiles of a window (144 rows x 64 B each: one token per four threads, one 16-byte chunk of each tensor per
  // thread) and its lse row are fetched with cp.async one window ahead, into the other tile buffer.
  const int ft = tid >> 2, fch = tid & 3, fty = ft / WS, ftx = ft - fty * WS;
  // Un-padded tensors (swin_attn_backward_unpadded): a pad token has no pixel.  Its q, dO, o rows are zero (its output row
  // was cropped, so no gradient arrives and it sends none: dS = 0), its k / v rows are the pad vectors; lse is laid out
  // on the padded grid by the forward.
  const bool padded = p.H != p.Hp || p.W != p.Wp;
  const __nv_bfloat16* kpad = static_cast<const __nv_bfloat16*>(p.k_pad);
  const __nv_bfloat16* vpad = static_cast<const __nv_bfloat16*>(p.v_pad);
  auto fetch = [&](const Win& wn, int buf) {
    int64_t pix, lpix;
    const bool ok = token_pixel_padded(p, wn.b, wn.wy, wn.wx, fty, ftx, pix, lpix);
    const int off = h * HD + fch * 8;
    const uint32_t dst = s_u32(sm.tile[buf][0]) + toff(ft, fch);
    if (ok) {
      cp_async16(dst, static_cast<const __nv_bfloat16*>(p.q) + pix * p.q_ps + off);
      cp_async16(dst + 2 * kTile, static_cast<const __nv_bfloat16*>(p.d_o) + pix * p.C + off);
      cp_async16(dst + 4 * kTile, static_cast<const __nv_bfloat16*>(p.k) + pix * p.k_ps + off);
      cp_async16(dst + 6 * kTile, static_cast<const __nv_bfloat16*>(p.v) + pix * p.v_ps + off);
      cp_async16(dst + 8 * kTile, static_cast<const __nv_bfloat16*>(p.o_in) + pix * p.C + off);
    } else {
      cp_async16_zfill(dst, p.q, 0);
      cp_async16_zfill(dst + 2 * kTile, p.q, 0);
      cp_async16_zfill(dst + 4 * kTile, kpad ? static_cast<const void*>(kpad + off) : p.q, kpad ? 16 : 0);
      cp_async16_zfill(dst + 6 * kTile, vpad ? static_cast<const void*>(vpad + off) : p.q, vpad ? 16 : 0);
      cp_async16_zfill(dst + 8 * kTile, p.q, 0);
    }
    // lse of a pad query row: +inf, so that its probabilities are exactly 0 whatever the forward left there
    if (fch == 0) {
      if (ok) cp_async4(s_u32(&sm.lse[buf][ft]), &p.lse[lpix * p.nH + h]);
      else sm.lse[buf][ft] = INFINITY;
    }
    cp_async_commit();
  };
  Win cur, nxt;
  cur.wx = w0 % p.nWx;
  cur.wy = (w0 / p.nWx) % p.nWy;
  cur.b = w0 / (p.nWx * p.nWy);
  nxt = cur;
  if (w0 < w1) fetch(cur, 0);
  float padreg[4][2] = {};   // un-padded tensors: this thread's share of dV (even warps) / dK (odd warps) summed over pad tokens

  for (int w = w0; w < w1; ++w, cur = nxt) {
    const int buf = (w - w0) & 1;
    // SW-MSA mask (swin_transformer.py:361-380): regions split at window row / column 6 in the last window row / column
    const bool hm = p.shift > 0 && cur.wy == p.nWy - 1, wm = p.shift > 0 && cur.wx == p.nWx - 1;
    const bool masked = hm || wm;
    cp_async_wait_all();
    __syncthreads();   // this window's tiles have landed; the previous window's images, tiles and bins gather are done
    nxt.next(p);
    if (w + 1 < w1) fetch(nxt, buf ^ 1);
    const uint32_t tq = s_u32(sm.tile[buf][0]), td = s_u32(sm.tile[buf][1]), tk_ = s_u32(sm.tile[buf][2]), tv = s_u32(sm.tile[buf][3]);
    const unsigned char *tdp = reinterpret_cast<const unsigned char*>(sm.tile[buf][1]), *top = reinterpret_cast<const unsigned char*>(sm.tile[buf][4]);
    const float nl0 = -sm.lse[buf][i0] * kLog2e, nl1 = -sm.lse[buf][i1] * kLog2e;
    int64_t pix0, pix1, lp_unused;
    const bool ok0 = token_pixel_padded(p, cur.b, cur.wy, cur.wx, iy0, ix0, pix0, lp_unused);
    const bool ok1 = token_pixel_padded(p, cur.b, cur.wy, cur.wx, iy1, ix1, pix1, lp_unused);

    // ---- delta = rowsum(dO . O) of rows i0, i1: 8 channels per lane of the quad ----
    float d0, d1;
    {
      auto dot8 = [](uint4 x, uint4 y) {
        const __nv_bfloat162* xa = reinterpret_cast<const __nv_bfloat162*>(&x);
        const __nv_bfloat162* ya = reinterpret_cast<const __nv_bfloat162*>(&y);
        float acc = 0.f;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 xf = __bfloat1622float2(xa[e]), yf = __bfloat1622float2(ya[e]);
          acc = fmaf(xf.x, yf.x, fmaf(xf.y, yf.y, acc));
        }
        return acc;
      };
      d0 = dot8(*reinterpret_cast<const uint4*>(tdp + toff(i0, t4)), *reinterpret_cast<const uint4*>(top + toff(i0, t4)));
      d1 = dot8(*reinterpret_cast<const uint4*>(tdp + toff(i1, t4)), *reinterpret_cast<const uint4*>(top + toff(i1, t4)));
      d0 += __shfl_xor_sync(0xffffffffu, d0, 1);
      d0 += __shfl_xor_sync(0xffffffffu, d0, 2);
      d1 += __shfl_xor_sync(0xffffffffu, d1, 1);
      d1 += __shfl_xor_sync(0xffffffffu, d1, 2);
    }
    // 16 x 32 fp32 accumulator (four n-tiles) -> bf16 -> rows i0 / i1 of an image-order tensor
    // Rows of pad tokens are not stored; with `pad_sum` (dV, dK only) they are added up instead, per thread over all the
    // windows of the job: the gradient of the pad vector.
    auto put = [&](float (&acc)[4][4], void* base, int64_t ps, float scl, bool pad_sum) {
      __nv_bfloat16* d0p = static_cast<__nv_bfloat16*>(base) + pix0 * ps + h * HD + 2 * t4;
      __nv_bfloat16* d1p = static_cast<__nv_bfloat16*>(base) + pix1 * ps + h * HD + 2 * t4;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        if (ok0) *reinterpret_cast<uint32_t*>(d0p + nt * 8) = pack2(acc[nt][0] * scl, acc[nt][1] * scl);
        if (ok1) *reinterpret_cast<uint32_t*>(d1p + nt * 8) = pack2(acc[nt][2] * scl, acc[nt][3] * scl);
      }
      if (pad_sum) {
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          if (!ok0) {
            padreg[nt][0] += acc[nt][0];
            padreg[nt][1] += acc[nt][1];
          }
          if (!ok1) {
            padreg[nt][0] += acc[nt][2];
            padreg[nt][1] += acc[nt][3];
          }
        }
      }
    };
    // The query side of the strip, specialised on which half of the key columns this warp takes: every n-tile index is
    // then a compile-time constant (bias offsets, image addresses, tile addresses fold into immediates).
    const float2 nd0 = make_float2(-d0, -d0), nd1 = make_float2(-d1, -d1);
    const int mb0 = masked ? mask_bits(hm, wm, my_h0, my_w0, t4) : 0, mb1 = masked ? mask_bits(hm, wm, my_h1, my_w1, t4) : 0;
    auto query_side = [&](auto KH) {
      constexpr int kN0 = decltype(KH)::value ? 10 : 0, kCnt = decltype(KH)::value ? 8 : 10;
      // ---- per n-tile: S, dP -> P, dS (bf16 pairs) ----
      uint32_t pk[kCnt][2];   // dS pairs [n-tile][row g / g + 8]: the A fragments of dQ
      {
        uint32_t aq[2][4], ad[2][4];
        ldsm_x4(aq[0], tq + r0 * 64 + offa0);
        ldsm_x4(aq[1], tq + r0 * 64 + offa1);
        ldsm_x4(ad[0], td + r0 * 64 + offa0);
        ldsm_x4(ad[1], td + r0 * 64 + offa1);
#pragma unroll
        for (int nn = 0; nn < kCnt; ++nn) {
          const int nt = kN0 + nn;
          uint32_t bk[4], bv[4];
          ldsm_x4(bk, tk_ + nt * 512 + offb);
          ldsm_x4(bv, tv + nt * 512 + offb);
          float sc[4] = {0.f, 0.f, 0.f, 0.f}, dp[4] = {0.f, 0.f, 0.f, 0.f};
          mma_bf16(sc, aq[0][0], aq[0][1], aq[0][2], aq[0][3], bk[0], bk[1]);
          mma_bf16(dp, ad[0][0], ad[0][1], ad[0][2], ad[0][3], bv[0], bv[1]);
          mma_bf16(sc, aq[1][0], aq[1][1], aq[1][2], aq[1][3], bk[2], bk[3]);
          mma_bf16(dp, ad[1][0], ad[1][1], ad[1][2], ad[1][3], bv[2], bv[3]);
          // keys j = 8 nt + 2 t4, j + 1 (same window row): bias index = per-thread base + compile-time constant
          const int cc = 8 * nt + 11 * ((8 * nt) / WS);
          const float2 bb0 = (nt % 3 == 1 ? pb0b : pb0a)[cc], bb1 = (nt % 3 == 1 ? pb1b : pb1a)[cc];
          float m0 = nl0, m1 = nl1;
          if (masked) {   // warp-uniform; both keys of the pair lie in the same region
            if (mb0 & (1 << mask_bit_of(nt))) m0 += -100.0f * kLog2e;
            if (mb1 & (1 << mask_bit_of(nt))) m1 += -100.0f * kLog2e;
          }
          // packed fp32x2 arithmetic: the two keys of a pair share every operand layout
          const float2 e0 = __ffma2_rn(make_float2(sc[0], sc[1]), c1c1, __fadd2_rn(bb0, make_float2(m0, m0)));
          const float2 e1 = __ffma2_rn(make_float2(sc[2], sc[3]), c1c1, __fadd2_rn(bb1, make_float2(m1, m1)));
          const uint32_t p0 = pack2(ex2(e0.x), ex2(e0.y));
          const uint32_t p1 = pack2(ex2(e1.x), ex2(e1.y));
          // dS from the bf16-rounded probabilities, as the products below see them
          const float2 g0 = __fmul2_rn(make_float2(__uint_as_float(p0 << 16), __uint_as_float(p0 & 0xffff0000u)),
                                       __fadd2_rn(make_float2(dp[0], dp[1]), nd0));
          const float2 g1 = __fmul2_rn(make_float2(__uint_as_float(p1 << 16), __uint_as_float(p1 & 0xffff0000u)),
                                       __fadd2_rn(make_float2(dp[2], dp[3]), nd1));
          const uint32_t s0 = pack2(g0.x, g0.y);
          const uint32_t s1 = pack2(g1.x, g1.y);
          *reinterpret_cast<uint32_t*>(pimg0 + nt * 8) = p0;
          *reinterpret_cast<uint32_t*>(pimg1 + nt * 8) = p1;
          *reinterpret_cast<uint32_t*>(simg0 + nt * 8) = s0;
          *reinterpret_cast<uint32_t*>(simg1 + nt * 8) = s1;
          pk[nn][0] = s0;
          pk[nn][1] = s1;
        }
      }
      // ---- dQ_s = scale dS_s K over this warp's keys: A from the registers, B = K (keys x channels) by transposing loads ----
      float acc[4][4];
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
#pragma unroll
      for (int kk = 0; kk < kCnt / 2; ++kk) {
        const int krow = (kN0 + 2 * kk) * 8;
        uint32_t b0[4], b1[4];
        ldsm_x4_t(b0, tk_ + krow * 64 + offa0);
        ldsm_x4_t(b1, tk_ + krow * 64 + offa1);
        const uint32_t a0 = pk[2 * kk][0], a1 = pk[2 * kk][1], a2 = pk[2 * kk + 1][0], a3 = pk[2 * kk + 1][1];
        mma_bf16(acc[0], a0, a1, a2, a3, b0[0], b0[1]);
        mma_bf16(acc[1], a0, a1, a2, a3, b0[2], b0[3]);
        mma_bf16(acc[2], a0, a1, a2, a3, b1[0], b1[1]);
        mma_bf16(acc[3], a0, a1, a2, a3, b1[2], b1[3]);
      }
      // the odd warp hands its partial sum to the even warp of the strip
      if (decltype(KH)::value) {
#pragma unroll
        for (int e = 0; e < 16; ++e) sm.dqx[strip][e][lane] = acc[e >> 2][e & 3];
      }
      pair_barrier(1 + strip);
      if (!decltype(KH)::value) {
#pragma unroll
        for (int e = 0; e < 16; ++e) acc[e >> 2][e & 3] += sm.dqx[strip][e][lane];
        put(acc, p.dq, p.q_ps, p.scale, false);
      }
    };
    if (kh) query_side(std::integral_constant<int, 1>{});
    else query_side(std::integral_constant<int, 0>{});
    __syncthreads();   // every strip of P and dS is written

    // ---- even warp: dV_s = (P^T)_s dO; odd warp: dK_s = scale (dS^T)_s Q.  A = the image, transposed on load ----
    {
      const __nv_bfloat16* img = kh == 0 ? sm.p : sm.ds;
      const uint32_t rhs = kh == 0 ? td : tq;
      float acc[4][4];
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
#pragma unroll
      for (int kt = 0; kt < 9; ++kt) {
        uint32_t a[4], b0[4], b1[4];
        ldsm_x4_t(a, s_u32(img + (kt * 16 + lt_row) * kImgLd + r0 + lt_col));
        ldsm_x4_t(b0, rhs + kt * 1024 + offa0);
        ldsm_x4_t(b1, rhs + kt * 1024 + offa1);
        mma_bf16(acc[0], a[0], a[1], a[2], a[3], b0[0], b0[1]);
        mma_bf16(acc[1], a[0], a[1], a[2], a[3], b0[2], b0[3]);
        mma_bf16(acc[2], a[0], a[1], a[2], a[3], b1[0], b1[1]);
        mma_bf16(acc[3], a[0], a[1], a[2], a[3], b1[2], b1[3]);
      }
      if (kh == 0) put(acc, p.dv, p.v_ps, 1.0f, padded);
      else put(acc, p.dk, p.k_ps, p.scale, padded);
    }
    // ---- bias-table gradient: bin (di, dj) += sum over keys (cy, cx) of dS[(cy + di, cx + dj)][(cy, cx)], in two steps.
    // (1) Two threads per (query row qy, key row ky): each takes 6 of the 12 query columns of that 12 x 12 block of dS and
    //     sums it along its 23 x-diagonals in registers (fully unrolled: no divergence, no atomics).
    //     The sums have their own buffer, so no barrier separates this from dV / dK.
    float* diag = sm.diag;   // [local diagonal][2 (query row qy, key row ky) + half]
    if (tid < 2 * N) {
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
#pragma unroll
      for (int d = 0; d < 17; ++d) diag[d * kDiagLd + tid] = dg[d];
    }
    __syncthreads();
    // (2) one thread per bin: <= 12 key rows x the two halves
    if (tid < NT) {
      const int bin = tid;
      const int di = bin / (2 * WS - 1) - (WS - 1), dj = bin % (2 * WS - 1) - (WS - 1);
      float sacc = 0.f;
      for (int ky = max(0, -di); ky < min(WS, WS - di); ++ky) {
        const int pair = (ky + di) * WS + ky;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int d = dj - half * 6 + 11;           // index into that half's 17 local diagonals
          if (d >= 0 && d < 17) sacc += diag[d * kDiagLd + 2 * pair + half];
        }
      }
      sm.bins[bin] += sacc;
    }
  }
  __syncthreads();
  for (int i = tid; i < NT; i += kThreads) atomicAdd(&p.d_table[i * p.nH + h], sm.bins[i]);
  if (padded) {
    // over the eight row groups of the warp, then the nine warps of a kind through shared memory, then one atomic per channel
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        float v = padreg[nt][e];
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        if (g == 0) atomicAdd(&sm.padacc[kh][nt * 8 + 2 * t4 + e], v * (kh ? p.scale : 1.0f));
      }
    }
    __syncthreads();
    if (tid < 2 * HD) {
      float* dst = tid < HD ? p.d_vpad : p.d_kpad;
      if (dst != nullptr) atomicAdd(&dst[h * HD + (tid & (HD - 1))], (&sm.padacc[0][0])[tid]);
    }
  }
}

}  // namespace mma12

// =====================================================================================================================
// Forward for window 12 on the same warp-MMA building blocks (FlashAttention-2 form).  The tcgen05 forward of
// swin_attn_tc12.cu is bound by its softmax threads (72 of 128 accumulator lanes carry rows, one sub-partition idles, two
// passes over tensor memory); here every lane of every warp carries work, P goes from the S accumulators straight into
// the A fragments of P V, and a CTA needs 60 KB of shared memory, so two CTAs (18 warps) share an SM.
//   warp w = query strip w (16 rows); keys in three chunks of 48 with the running maximum / sum of the online softmax
//   (exp2 domain; the SW-MSA mask is -100, never -inf, so no row is empty);  O is rescaled per chunk and divided by the
//   sum of the bf16-rounded probabilities at the end.
// Padding (H or W not a multiple of 12) costs nothing in the math: the gather writes k_pad / v_pad into the K / V rows of
// pad tokens (the reference pads before the qkv Linear, model/NewCRFs/swin_transformer.py:204-209; window 12 = NewCRFDepth version "large12", NewCRFDepth.py:26), zero-fills their Q rows and the
// epilogue skips their output rows.
namespace mma12 {

constexpr int kFwdWarps = 9, kFwdThreads = kFwdWarps * 32;
struct SmemF {
  __nv_bfloat16 tile[2][3][kTile];     // q, k, v of the current and the next window
  float2 bias2[NT + 1];
};

__global__ void __launch_bounds__(kFwdThreads, 2) mma12_fwd_kernel(const Problem p, int n_win, int win_per_job) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  SmemF& sm = *reinterpret_cast<SmemF*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t4 = lane & 3;
  const int h = (int)blockIdx.x % p.nH;
  const int w0 = ((int)blockIdx.x / p.nH) * win_per_job, w1 = min(w0 + win_per_job, n_win);
  const float c1 = p.scale * kLog2e;
  const float2 c1c1 = make_float2(c1, c1);
  for (int i = tid; i < NT; i += kFwdThreads) {
    const float b0 = p.table[(NT - 1 - i) * p.nH + h] * kLog2e;
    const float b1 = i + 1 < NT ? p.table[(NT - 2 - i) * p.nH + h] * kLog2e : 0.f;
    sm.bias2[i] = make_float2(b0, b1);
  }
  const int r0 = warp * 16;
  const int la_row = (lane & 7) + ((lane >> 3) & 1) * 8;
  const int offa0 = toff(la_row, (lane >> 4) & 1), offa1 = toff(la_row, 2 + ((lane >> 4) & 1));
  const int offb = toff(lane & 7, lane >> 3);
  const int i0 = r0 + g, i1 = i0 + 8;
  const int iy0 = i0 / WS, ix0 = i0 - iy0 * WS, iy1 = i1 / WS, ix1 = i1 - iy1 * WS;
  const int rb0 = 264 - 23 * iy0 - ix0, rb1 = 264 - 23 * iy1 - ix1;
  const bool my_h0 = iy0 >= 6, my_w0 = ix0 >= 6, my_h1 = iy1 >= 6, my_w1 = ix1 >= 6;
  // Keys j = 8 nt + 2 t4 (+ 1) of n-tile nt: their window row is jy = (8 nt) / 12 + (nt % 3 == 1 && t4 >= 2), so the bias
  // index rb + 23 jy + jx = rb + j + 11 jy is a per-thread base (two variants) plus a compile-time constant.
  const int hi = t4 >> 1;
  const float2 *pb0a = sm.bias2 + rb0 + 2 * t4, *pb0b = pb0a + 11 * hi, *pb1a = sm.bias2 + rb1 + 2 * t4, *pb1b = pb1a + 11 * hi;
  const bool padded = p.H != p.Hp || p.W != p.Wp;

  // token (ty, tx) of a window -> un-rolled coordinates in the PADDED image
  auto coords = [&](const Win& wn, int ty, int tx, int& y, int& x) {
    y = wn.wy * WS + ty + p.shift;
    x = wn.wx * WS + tx + p.shift;
    if (y >= p.Hp) y -= p.Hp;
    if (x >= p.Wp) x -= p.Wp;
  };
  // gather: two tokens per thread (ft, ft + 72), one 16-byte chunk of q, k, v each
  const int ft = tid >> 2, fch = tid & 3;
  const int fty0 = ft / WS, ftx0 = ft - fty0 * WS, fty1 = (ft + 72) / WS, ftx1 = (ft + 72) - fty1 * WS;
  const int off = h * HD + fch * 8;
  const __nv_bfloat16* kpad = static_cast<const __nv_bfloat16*>(p.k_pad);
  const __nv_bfloat16* vpad = static_cast<const __nv_bfloat16*>(p.v_pad);
  auto fetch = [&](const Win& wn, int buf) {
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int tk = ft + 72 * u;
      int y, x;
      coords(wn, u ? fty1 : fty0, u ? ftx1 : ftx0, y, x);
      const uint32_t dst = s_u32(sm.tile[buf][0]) + toff(tk, fch);
      if (!padded || (y < p.H && x < p.W)) {
        const int64_t pix = ((int64_t)wn.b * p.H + y) * p.W + x;
        cp_async16(dst, static_cast<const __nv_bfloat16*>(p.q) + pix * p.q_ps + off);
        cp_async16(dst + 2 * kTile, static_cast<const __nv_bfloat16*>(p.k) + pix * p.k_ps + off);
        cp_async16(dst + 4 * kTile, static_cast<const __nv_bfloat16*>(p.v) + pix * p.v_ps + off);
      } else {   // pad token: q = 0 (its output row is cropped), k / v = the pad vectors (or 0 without them)
        cp_async16_zfill(dst, p.q, 0);
        cp_async16_zfill(dst + 2 * kTile, kpad ? static_cast<const void*>(kpad + off) : p.k, kpad ? 16 : 0);
        cp_async16_zfill(dst + 4 * kTile, vpad ? static_cast<const void*>(vpad + off) : p.v, vpad ? 16 : 0);
      }
    }
    cp_async_commit();
  };
  Win cur, nxt;
  cur.wx = w0 % p.nWx;
  cur.wy = (w0 / p.nWx) % p.nWy;
  cur.b = w0 / (p.nWx * p.nWy);
  nxt = cur;
  if (w0 < w1) fetch(cur, 0);

  for (int w = w0; w < w1; ++w, cur = nxt) {
    const int buf = (w - w0) & 1;
    const bool hm = p.shift > 0 && cur.wy == p.nWy - 1, wm = p.shift > 0 && cur.wx == p.nWx - 1;
    const bool masked = hm || wm;
    cp_async_wait_all();
    __syncthreads();   // this window's tiles have landed; every warp is done with the other buffer
    nxt.next(p);
    if (w + 1 < w1) fetch(nxt, buf ^ 1);
    const uint32_t tq = s_u32(sm.tile[buf][0]), tk_ = s_u32(sm.tile[buf][1]), tv = s_u32(sm.tile[buf][2]);

    const int mb0 = masked ? mask_bits(hm, wm, my_h0, my_w0, t4) : 0, mb1 = masked ? mask_bits(hm, wm, my_h1, my_w1, t4) : 0;
    uint32_t aq[2][4];
    ldsm_x4(aq[0], tq + r0 * 64 + offa0);
    ldsm_x4(aq[1], tq + r0 * 64 + offa1);
    float o[5][4];   // four n-tiles of O and a ones tile: its columns are the row sum of the bf16 probabilities
#pragma unroll
    for (int nt = 0; nt < 5; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float2 t[6][2];   // [n-tile][row g / g + 8]: the logits of a key pair
      float cm0 = -INFINITY, cm1 = -INFINITY;
#pragma unroll
      for (int nn = 0; nn < 6; ++nn) {
        const int nt = c * 6 + nn;
        uint32_t bk[4];
        ldsm_x4(bk, tk_ + nt * 512 + offb);
        float sc[4] = {0.f, 0.f, 0.f, 0.f};
        mma_bf16(sc, aq[0][0], aq[0][1], aq[0][2], aq[0][3], bk[0], bk[1]);
        mma_bf16(sc, aq[1][0], aq[1][1], aq[1][2], aq[1][3], bk[2], bk[3]);
        const int cc = 8 * nt + 11 * ((8 * nt) / WS);   // compile-time after unrolling
        const float2 bb0 = (nt % 3 == 1 ? pb0b : pb0a)[cc], bb1 = (nt % 3 == 1 ? pb1b : pb1a)[cc];
        float2 b0 = bb0, b1 = bb1;
        if (masked) {   // warp-uniform
          if (mb0 & (1 << mask_bit_of(nt))) b0 = __fadd2_rn(b0, make_float2(-100.0f * kLog2e, -100.0f * kLog2e));
          if (mb1 & (1 << mask_bit_of(nt))) b1 = __fadd2_rn(b1, make_float2(-100.0f * kLog2e, -100.0f * kLog2e));
        }
        // packed fp32x2 arithmetic: the two keys of a pair share every operand layout
        t[nn][0] = __ffma2_rn(make_float2(sc[0], sc[1]), c1c1, b0);
        t[nn][1] = __ffma2_rn(make_float2(sc[2], sc[3]), c1c1, b1);
        cm0 = fmaxf(cm0, fmaxf(t[nn][0].x, t[nn][0].y));
        cm1 = fmaxf(cm1, fmaxf(t[nn][1].x, t[nn][1].y));
      }
      cm0 = fmaxf(cm0, __shfl_xor_sync(0xffffffffu, cm0, 1));
      cm0 = fmaxf(cm0, __shfl_xor_sync(0xffffffffu, cm0, 2));
      cm1 = fmaxf(cm1, __shfl_xor_sync(0xffffffffu, cm1, 1));
      cm1 = fmaxf(cm1, __shfl_xor_sync(0xffffffffu, cm1, 2));
      const float n0 = fmaxf(mx0, cm0), n1 = fmaxf(mx1, cm1);
      const float al0 = ex2(mx0 - n0), al1 = ex2(mx1 - n1);   // exp2(-inf) = 0 for the first chunk
      mx0 = n0;
      mx1 = n1;
      const float2 a00 = make_float2(al0, al0), a11 = make_float2(al1, al1);
#pragma unroll
      for (int nt = 0; nt < 5; ++nt) {
        const float2 u0 = __fmul2_rn(make_float2(o[nt][0], o[nt][1]), a00), u1 = __fmul2_rn(make_float2(o[nt][2], o[nt][3]), a11);
        o[nt][0] = u0.x;
        o[nt][1] = u0.y;
        o[nt][2] = u1.x;
        o[nt][3] = u1.y;
      }
      const float2 nn0 = make_float2(-n0, -n0), nn1 = make_float2(-n1, -n1);
      uint32_t pa[6][2];
#pragma unroll
      for (int nn = 0; nn < 6; ++nn) {
        const float2 e0 = __fadd2_rn(t[nn][0], nn0), e1 = __fadd2_rn(t[nn][1], nn1);
        pa[nn][0] = pack2(ex2(e0.x), ex2(e0.y));
        pa[nn][1] = pack2(ex2(e1.x), ex2(e1.y));
      }
#pragma unroll
      for (int kk = 0; kk < 3; ++kk) {
        const int krow = (c * 6 + 2 * kk) * 8;
        uint32_t b0[4], b1[4];
        ldsm_x4_t(b0, tv + krow * 64 + offa0);
        ldsm_x4_t(b1, tv + krow * 64 + offa1);
        const uint32_t a0 = pa[2 * kk][0], a1 = pa[2 * kk][1], a2 = pa[2 * kk + 1][0], a3 = pa[2 * kk + 1][1];
        mma_bf16(o[0], a0, a1, a2, a3, b0[0], b0[1]);
        mma_bf16(o[1], a0, a1, a2, a3, b0[2], b0[3]);
        mma_bf16(o[2], a0, a1, a2, a3, b1[0], b1[1]);
        mma_bf16(o[3], a0, a1, a2, a3, b1[2], b1[3]);
        mma_bf16(o[4], a0, a1, a2, a3, 0x3f803f80u, 0x3f803f80u);   // B = ones (bf16 pairs): the normaliser, from the MMA
      }
    }
    const float l0 = o[4][0], l1 = o[4][2];
    // ---- epilogue: rows i0 / i1 -> their pixels (cropped when they are pad tokens); lse on the padded grid ----
    int y0, x0, y1, x1;
    coords(cur, iy0, ix0, y0, x0);
    coords(cur, iy1, ix1, y1, x1);
    const float inv0 = 1.0f / l0, inv1 = 1.0f / l1;
    if (y0 < p.H && x0 < p.W) {
      __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(p.o) + (((int64_t)cur.b * p.H + y0) * p.W + x0) * p.o_ps + h * HD + 2 * t4;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) *reinterpret_cast<uint32_t*>(dst + nt * 8) = pack2(o[nt][0] * inv0, o[nt][1] * inv0);
    }
    if (y1 < p.H && x1 < p.W) {
      __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(p.o) + (((int64_t)cur.b * p.H + y1) * p.W + x1) * p.o_ps + h * HD + 2 * t4;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) *reinterpret_cast<uint32_t*>(dst + nt * 8) = pack2(o[nt][2] * inv1, o[nt][3] * inv1);
    }
    if (p.lse != nullptr && t4 == 0) {
      p.lse[(((int64_t)cur.b * p.Hp + y0) * p.Wp + x0) * p.nH + h] = (mx0 + log2f(l0)) * 0.6931471805599453f;
      p.lse[(((int64_t)cur.b * p.Hp + y1) * p.Wp + x1) * p.nH + h] = (mx1 + log2f(l1)) * 0.6931471805599453f;
    }
  }
}

}  // namespace mma12

bool mma12_forward_supported(const Problem& p, int dtype) {
  if (dtype != SWIN_ATTN_BF16 || p.ws != 12 || p.mask != nullptr || p.dropout_p > 0.0f) return false;
  if (p.shift != 0 && p.shift != 6) return false;
  if (p.H < p.shift || p.W < p.shift) return false;   // the pad rows would wrap around the rolled image
  if ((p.q_ps | p.k_ps | p.v_ps | p.o_ps) & 7) return false;
  auto mis = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) != 0; };
  return !(mis(p.q) || mis(p.k) || mis(p.v) || mis(p.o) || mis(p.k_pad) || mis(p.v_pad));
}

cudaError_t launch_mma12_forward(const Problem& p, cudaStream_t st) {
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int smem = (int)sizeof(mma12::SmemF);
  cudaError_t e = cudaFuncSetAttribute(mma12::mma12_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  const int n_win = p.B * p.nWy * p.nWx;
  int chunks = 2 * sms / p.nH;   // two CTAs per SM
  if (chunks < 1) chunks = 1;
  if (chunks > n_win) chunks = n_win;
  const int per = (n_win + chunks - 1) / chunks;
  chunks = (n_win + per - 1) / per;
  TimedRegion timed(st);
  mma12::mma12_fwd_kernel<<<chunks * p.nH, mma12::kFwdThreads, smem, st>>>(p, n_win, per);
  count_launch();
  return cudaGetLastError();
}

bool mma12_backward_supported(const Problem& p, int dtype) {
  if (dtype != SWIN_ATTN_BF16 || p.ws != 12 || p.mask != nullptr || p.dropout_p > 0.0f) return false;
  if (p.shift != 0 && p.shift != 6) return false;
  if (p.H < p.shift || p.W < p.shift) return false;        // un-padded tensors: the pad rows would wrap around the rolled image
  if ((p.q_ps | p.k_ps | p.v_ps | p.C) & 7) return false;
  auto mis = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) != 0; };
  return p.o_in != nullptr && !(mis(p.q) || mis(p.k) || mis(p.v) || mis(p.d_o) || mis(p.o_in) || mis(p.dq) || mis(p.dk) || mis(p.dv) ||
                                mis(p.k_pad) || mis(p.v_pad));
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
  if (e == cudaSuccess && p.d_kpad != nullptr) e = cudaMemsetAsync(p.d_kpad, 0, sizeof(float) * p.C, st);
  if (e == cudaSuccess && p.d_vpad != nullptr) e = cudaMemsetAsync(p.d_vpad, 0, sizeof(float) * p.C, st);
  if (e != cudaSuccess) return e;
  mma12::mma12_bwd_kernel<<<chunks * p.nH, mma12::kThreads, smem, st>>>(p, n_win, per);
  count_launch();
  return cudaGetLastError();
}

}  // namespace swin
