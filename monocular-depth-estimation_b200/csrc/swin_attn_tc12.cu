// tcgen05 forward of the shifted-window attention for window 12 (bf16, shift 0 / 6): the Swin-L/384
// encoder of the reference's ODA family (model/ODA/oda_swin_transformer.py) and the ws-12 rows of the
// microbench sweep.  Same building blocks as swin_attn_tc.cu (job grid with one head per CTA, TMA window
// gather / scatter, S and PV as tcgen05 MMAs with TMEM accumulators, one softmax thread per query row),
// re-tiled for N = 144 tokens:
//
//   tile  = one (window, head): Q, K, V are 144 x 32 bf16 (9216 B each), one TMA box (12 x 12) or the
//           2 / 4 parts of a shifted window (6 x 12, 6 x 6).
//   unit  = HALF a tile: 72 query rows against all 144 keys.  144 = 128 + 16 does not map onto M = 128
//           accumulators; two M = 128 MMAs whose A operand starts at row 0 / row 72 of the Q tile (rows past
//           72 are computed and ignored) do, and keep the accumulator of a unit at 144 TMEM columns, so
//           three units are in flight (3 x 144 columns).
//   S     = Q[72h .. 72h+127] K^T          M128 N144 K32 (2 instructions)
//   softmax thread (rows 0 .. 71 of the unit): pass 1 reads the 144 raw scores once for their maximum;
//           m = max * scale*log2e + (largest bias of the head) bounds every logit from above, so
//           p = exp2(t - m) <= 1 needs no exact row maximum.  Pass 2 re-reads S from TMEM in the three
//           column blocks of the P image (64 + 64 + 16), adds bias / mask, exponentiates, accumulates the
//           row sum and writes bf16 P rows (128B-swizzled K-major blocks).
//   O     = P[72 x 144] V                  M128 N32 K144 (9 instructions: 4 + 4 + 1 K-steps over the blocks)
//   epilogue: O row / rowsum -> bf16 -> staging -> TMA tensor store of the half window (12 x 6, or 6 x 12 /
//           two 6 x 6 boxes for shifted blocks).
//
// Padding (H or W not a multiple of 12: every stage of the window-12 encoder, NewCRFDepth "large12", on a 352 x 1216 or 480 x 640 input): TMA zero-fills the
// out-of-image part of a load box and clips it from a store box.  The reference pads BEFORE the qkv Linear
// (model/NewCRFs/swin_transformer.py:204-209; window 12 = NewCRFDepth version "large12", NewCRFDepth.py:26), so pad tokens carry the Linear bias: k_pad goes -- once per CTA -- into the
// spare row 144 of every K tile and v_pad into the spare rows 144 / 145 of every V tile (tiles are 160 rows; the
// boxes write 144).  With padding S is N = 160 wide (column 144 = q . k_pad, which the softmax substitutes for the
// columns of pad tokens) and the PV product takes a tenth K-step whose P columns 144 / 145 hold the hi / lo bf16
// halves of the row's summed probability over pad tokens.  Window-aligned launches run the 144 / 9-step shapes.
//
// Not covered here (served by the CUDA-core kernels): explicit masks, fp32; the backward is swin_attn_mma12.cu.
#include "swin_attn_tc.cuh"

namespace swin {
namespace tc12 {

using namespace tc;   // PTX wrappers only; the window-7 constants are shadowed below

constexpr int WS12 = 12;
constexpr int N12 = 144;
constexpr int kHalf = 72;            // query rows per unit
constexpr int kSplit = 6;            // shift = 6: a shifted window splits 6 | 6 along each axis
constexpr int kTile12 = N12 * 64;    // 9216 B: one tensor of one tile (what the TMA boxes write)
constexpr int kRowsKV = 160;         // rows of a K / V tile: 144 tokens + the pad-vector rows (144, 145) + zero rows
constexpr int kTileKV = kRowsKV * 64;
constexpr int kStage12 = kTile12 + 2 * kTileKV;
constexpr int kStages12 = 4;
constexpr int kSlots12 = 3;
constexpr int kThreads12 = 128 + 128 * kSlots12;
constexpr int kFill12 = kThreads12 - 64;
constexpr int kPBlock = kHalf * 128;             // 72 rows x 64 bf16 columns
constexpr int kPSlot = 3 * kPBlock;              // column blocks [0,64) [64,128) [128,144)
constexpr int kSlotCols12 = 160;          // S accumulator of a unit: 144 keys (+ 16 with padding); O in [0, 32)
constexpr int kTable12 = 23 * 23;                // (2 ws - 1)^2

// tensor maps per tensor: [0] 12x12 (x,y), [1] 6x6, [2] 6x12 (6 wide, 12 tall), [3] 12x6 (12 wide, 6 tall)
struct TmaMaps12 {
  CUtensorMap q[3], k[3], v[3];   // loads use [0..2]
  CUtensorMap o[4];               // stores use [1..3]
};
constexpr int kBoxX12[4] = {12, 6, 6, 12};
constexpr int kBoxY12[4] = {12, 6, 12, 6};

struct SmemMap12 {
  uint32_t stages, slots, bias, bars, tmem_ptr, total;
};
__host__ __device__ inline SmemMap12 smem_map_12() {
  SmemMap12 m;
  m.stages = 0;
  m.slots = m.stages + kStages12 * kStage12;
  // the M = 128 A operand of the PV product reads 128 rows from a 72-row block: 56 rows of slack after the last
  m.bias = m.slots + kSlots12 * kPSlot + 56 * 128;
  m.bars = m.bias + (kTable12 + 16) * 4;         // bias[529], per-warp maxima [14]
  m.bars = (m.bars + 7) & ~7u;
  m.tmem_ptr = m.bars + 8 * (2 * kStages12 + 4 * kSlots12);
  m.total = m.tmem_ptr + 16;
  return m;
}

// Row r of a tile <-> token (i, j) of the window.  Natural order for un-shifted blocks; for shifted blocks the
// TMA parts follow one another: rows [0,36) i<6,j<6  [36,72) i>=6,j<6  [72,108) i<6,j>=6  [108,144) i>=6,j>=6,
// so that a unit (72 rows) is the left / right half of the window.
template <bool kPerm>
__host__ __device__ constexpr int row_i12(int r) {
  if (!kPerm) return r / WS12;
  return r < 36 ? r / 6 : r < 72 ? 6 + (r - 36) / 6 : r < 108 ? (r - 72) / 6 : 6 + (r - 108) / 6;
}
template <bool kPerm>
__host__ __device__ constexpr int row_j12(int r) {
  if (!kPerm) return r % WS12;
  return r < 72 ? r % 6 : 6 + r % 6;
}

// bits [lo, hi) clipped to the 12 rows / columns of a window
__device__ __forceinline__ int span_bits(int lo, int hi) {
  lo = min(max(lo, 0), WS12);
  hi = min(max(hi, 0), WS12);
  return ((1 << hi) - 1) & ~((1 << lo) - 1);
}

// One window of one tensor, global -> shared: 1 box (un-shifted) or the 2 / 4 parts of a shifted window.
template <bool kShifted>
__device__ __forceinline__ void load_window12(const CUtensorMap* maps, uint32_t dst, uint32_t bar, const Problem& p, int b, int wy,
                                              int wx, int h) {
  const int ch = h * kHeadDim;
  if (!kShifted) {
    tma_load_4d(dst, maps, bar, ch, wx * WS12, wy * WS12, b);
  } else {
    const int xa = wx * WS12 + kSplit, ya = wy * WS12 + kSplit;
    const int xb = wx == p.nWx - 1 ? 0 : wx * WS12 + WS12, yb = wy == p.nWy - 1 ? 0 : wy * WS12 + WS12;
    if (wy == p.nWy - 1) {
      tma_load_4d(dst, maps + 1, bar, ch, xa, ya, b);
      tma_load_4d(dst + 36 * 64, maps + 1, bar, ch, xa, yb, b);
      tma_load_4d(dst + 72 * 64, maps + 1, bar, ch, xb, ya, b);
      tma_load_4d(dst + 108 * 64, maps + 1, bar, ch, xb, yb, b);
    } else {
      tma_load_4d(dst, maps + 2, bar, ch, xa, ya, b);
      tma_load_4d(dst + 72 * 64, maps + 2, bar, ch, xb, ya, b);
    }
  }
}

// One HALF window (72 staged rows) of the output, shared -> global.
template <bool kShifted>
__device__ __forceinline__ void store_half12(const CUtensorMap* maps, uint32_t src, const Problem& p, int b, int wy, int wx, int h,
                                             int half) {
  const int ch = h * kHeadDim;
  if (!kShifted) {
    tma_store_4d(maps + 3, src, ch, wx * WS12, wy * WS12 + half * 6, b);   // window rows [6 half, 6 half + 6)
  } else {
    const int ya = wy * WS12 + kSplit, yb = wy == p.nWy - 1 ? 0 : wy * WS12 + WS12;
    const int x = half == 0 ? wx * WS12 + kSplit : (wx == p.nWx - 1 ? 0 : wx * WS12 + WS12);   // window columns [6 half, ..)
    if (wy == p.nWy - 1) {
      tma_store_4d(maps + 1, src, ch, x, ya, b);
      tma_store_4d(maps + 1, src + 36 * 64, ch, x, yb, b);
    } else {
      tma_store_4d(maps + 2, src, ch, x, ya, b);
    }
  }
}

template <bool kShifted>
__global__ void __launch_bounds__(kThreads12, 1)
tc12_fwd_kernel(const Problem p, int n_win, int win_per_job, int has_pad, const __grid_constant__ TmaMaps12 maps) {
  constexpr bool kPerm = kShifted;
  extern __shared__ __align__(1024) unsigned char smem[];
  const SmemMap12 map = smem_map_12();
  const uint32_t sbase = smem_u32(smem);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int h = (int)blockIdx.x % p.nH;
  const int w0 = ((int)blockIdx.x / p.nH) * win_per_job;
  const int w1 = min(w0 + win_per_job, n_win);
  const int n_tiles = w1 - w0;
  const int n_units = 2 * n_tiles;

  const uint32_t bar_full = sbase + map.bars;              // [kStages12] TMA -> S MMA
  const uint32_t bar_empty = bar_full + 8 * kStages12;     // [kStages12] PV MMA of the second half -> producer
  const uint32_t bar_s = bar_empty + 8 * kStages12;        // [kSlots12]  S ready
  const uint32_t bar_p = bar_s + 8 * kSlots12;             // [kSlots12]  P written
  const uint32_t bar_o = bar_p + 8 * kSlots12;             // [kSlots12]  O ready
  const uint32_t bar_r = bar_o + 8 * kSlots12;             // [kSlots12]  O read, slot reusable

  if (tid == 0) {
    for (int i = 0; i < kStages12; ++i) {
      mbar_init(bar_full + 8 * i, 1);
      mbar_init(bar_empty + 8 * i, 1);
    }
    for (int i = 0; i < kSlots12; ++i) {
      mbar_init(bar_s + 8 * i, 1);
      mbar_init(bar_p + 8 * i, 128);
      mbar_init(bar_o + 8 * i, 1);
      mbar_init(bar_r + 8 * i, 128);
    }
    fence_barrier_init();
  }
  __syncthreads();
  float* bias = reinterpret_cast<float*>(smem + map.bias);
  if (warp != 0 && warp != 3) {
    const int t14 = tid - (warp < 3 ? 32 : 64);
    if (warp == 1) tmem_alloc(sbase + map.tmem_ptr, 512);
    float bmax = -INFINITY;
    for (int i = t14; i < kTable12; i += kFill12) {
      const float v = p.table[i * p.nH + h] * kLog2e;
      bias[i] = v;
      bmax = fmaxf(bmax, v);
    }
    // largest bias of the head (upper bound of every logit's bias term): one partial per warp, combined by the readers
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) bmax = fmaxf(bmax, __shfl_xor_sync(0xffffffffu, bmax, o));
    if (lane == 0) bias[kTable12 + (t14 >> 5)] = bmax;
    if (has_pad) {
      // spare rows 144 .. 159 of every K / V tile (64B-swizzled rows, never written by the TMA boxes): row 144 = the
      // head's pad vector (K and V), row 145 = the pad vector again (V: the lo half of the pad probability), rest zero
      for (int i = t14; i < kStages12 * 2 * 16 * 4; i += kFill12) {
        const int chunk = i & 3, rr = (i >> 2) & 15, is_v = (i >> 6) & 1, st = i >> 7;
        const int rowx = N12 + rr;
        const void* vecp = is_v ? p.v_pad : p.k_pad;
        const bool vec_row = rr == 0 || (is_v && rr == 1);
        const uint4 val = vec_row && vecp != nullptr ? reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(vecp) + h * kHeadDim)[chunk]
                                                     : make_uint4(0u, 0u, 0u, 0u);
        unsigned char* tile = smem + map.stages + st * kStage12 + kTile12 + is_v * kTileKV;
        *reinterpret_cast<uint4*>(tile + rowx * 64 + ((chunk ^ ((rowx >> 1) & 3)) << 4)) = val;
      }
    }
    fence_proxy_async();
    tc_fence_before();
    named_barrier(6, kFill12);
    tc_fence_after();
  } else if (lane == 0) {
    for (int i = kShifted ? 1 : 0; i < (kShifted ? 3 : 1); ++i) {
      tma_prefetch_desc(&maps.q[i]);
      tma_prefetch_desc(&maps.k[i]);
      tma_prefetch_desc(&maps.v[i]);
    }
  }
  const uint32_t tmem_base = warp != 0 && warp != 3 ? *reinterpret_cast<volatile uint32_t*>(smem + map.tmem_ptr) : 0u;

  if (warp == 0) {
    // ===== TMA producer: one tile (q, k, v of a window) per stage ==============================
    WinIter it;
    it.init(p, w0);
    for (int n = 0; n < n_tiles; ++n, it.advance(p, 1)) {
      const int s = n % kStages12;
      const uint32_t st = sbase + map.stages + s * kStage12, bar = bar_full + 8 * s;
      mbar_wait(bar_empty + 8 * s, ((n / kStages12) & 1) ^ 1);
      if (elect_one()) {
        mbar_expect_tx(bar, 3 * kTile12);
        load_window12<kShifted>(maps.q, st, bar, p, it.b, it.wy, it.wx, h);
        load_window12<kShifted>(maps.k, st + kTile12, bar, p, it.b, it.wy, it.wx, h);
        load_window12<kShifted>(maps.v, st + kTile12 + kTileKV, bar, p, it.b, it.wy, it.wx, h);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===== S = Q_half K^T issuer ================================================================
    const uint32_t idesc_s = has_pad ? instr_desc(128, kRowsKV, 0) : instr_desc(128, N12, 0);
    for (int u = 0; u < n_units; ++u) {
      const int n = u >> 1, half = u & 1, s = n % kStages12, a = u % kSlots12;
      if (half == 0) mbar_wait(bar_full + 8 * s, (n / kStages12) & 1);
      mbar_wait(bar_r + 8 * a, ((u / kSlots12) & 1) ^ 1);
      tc_fence_after();
      const uint32_t st = sbase + map.stages + s * kStage12;
      const uint64_t qd = smem_desc(st + half * kHalf * 64, 16, 512, 4);
      const uint64_t kd = smem_desc(st + kTile12, 16, 512, 4);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 2; ++k) umma_ss(tmem_base + a * kSlotCols12, qd + 2 * k, kd + 2 * k, idesc_s, k);
        umma_commit(bar_s + 8 * a);
      }
      __syncwarp();
    }
  } else if (warp == 2) {
    // ===== O = P V issuer (9 K-steps over the three column blocks of P; a tenth for the pad columns) ==
    constexpr uint32_t idesc_o = instr_desc(128, 32, 1);
    const int ksteps = has_pad ? 10 : 9;
    for (int u = 0; u < n_units; ++u) {
      const int n = u >> 1, half = u & 1, s = n % kStages12, a = u % kSlots12;
      mbar_wait(bar_p + 8 * a, (u / kSlots12) & 1);
      tc_fence_after();
      const uint32_t vt = sbase + map.stages + s * kStage12 + kTile12 + kTileKV;
      const uint32_t pb = sbase + map.slots + a * kPSlot;
      const uint64_t vd = smem_desc(vt, 16, 512, 4);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 10; ++k)
          if (k < ksteps) umma_ss(tmem_base + a * kSlotCols12, smem_desc(pb + (k >> 2) * kPBlock, 16, 1024, 2) + 2 * (k & 3), vd + 64 * k, idesc_o, k);
        umma_commit(bar_o + 8 * a);
        if (half == 1) umma_commit(bar_empty + 8 * s);
      }
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ===== softmax / epilogue warpgroups =======================================
    const int g = (warp >> 2) - 1;          // slot owned by this warpgroup
    const int q = warp & 3;                 // TMEM sub-partition: lanes 32q .. 32q+31
    const int row = q * 32 + lane;          // row of the unit's S / O held by this thread (valid below 72)
    const bool row_ok = row < kHalf;
    const bool warp_on = q * 32 < kHalf;    // warp 3 of a warpgroup holds no valid row
    const int rowc = row_ok ? row : kHalf - 1;
    const int ri_h[2] = {row_i12<kPerm>(rowc), row_i12<kPerm>(kHalf + rowc)};
    const int rj_h[2] = {row_j12<kPerm>(rowc), row_j12<kPerm>(kHalf + rowc)};
    float bmax = bias[kTable12];
#pragma unroll
    for (int i = 1; i < kFill12 / 32; ++i) bmax = fmaxf(bmax, bias[kTable12 + i]);
    const float c1s = p.scale * kLog2e;
    const float2 c1 = make_float2(c1s, c1s);
    const uint32_t sl = sbase + map.slots + g * kPSlot;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + g * kSlotCols12;
    WinIter it;
    it.init(p, w0);
    int cur = 0;

    int m = 0;
    for (int u = g; u < n_units; u += kSlots12, ++m) {
      const int n = u >> 1, half = u & 1;
      it.advance(p, n - cur);
      cur = n;
      const int ri = ri_h[half], rj = rj_h[half];
      const float* brow = bias + ri * 23 + rj + 11 * 23 + 11;   // relative_position_index[row, col] = .. - (ci*23 + cj)
      // SW-MSA mask (swin_transformer.py:361-380): -100 where the region ids of query and key differ.  Regions split
      // at window row / column 6 in the last window row / column; nm[a][b] = (mask ? -100 log2e : 0) - m for keys
      // with (ci >= 6) == a, (cj >= 6) == b.
      const bool hm = kShifted && it.wy == p.nWy - 1, wm = kShifted && it.wx == p.nWx - 1;
      const bool my_h = ri >= kSplit, my_w = rj >= kSplit;
      // window rows / columns of this window that are pad tokens (bit i of pi: window row i): the pad rows [H, Hp) of
      // the image are rows [H - shift, Hp - shift) of the rolled image
      int pi = 0, pj = 0;
      if (has_pad) {
        pi = span_bits(p.H - p.shift - it.wy * WS12, p.Hp - p.shift - it.wy * WS12);
        pj = span_bits(p.W - p.shift - it.wx * WS12, p.Wp - p.shift - it.wx * WS12);
      }
      const bool wpad = (pi | pj) != 0;   // warp-uniform

      mbar_wait(bar_s + 8 * g, m & 1);
      tc_fence_after();
      // the TMA store of this warpgroup's previous unit must have finished reading the staging rows (same bytes as
      // P block 0) before P is overwritten
      if (q == 0 && elect_one()) bulk_wait_read();
      named_barrier(1 + g, 128);
      float l = 0.f, m_ub = 0.f;
      if (warp_on) {
        // pass 1: maximum of the raw scores
        // (batches of 64 + 64 + 16 columns, four independent 3-input max chains per batch)
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int blk = 0; blk < 3; ++blk) {
          const int cols = blk < 2 ? 64 : 16;
          uint32_t r[64];
#pragma unroll
          for (int c = 0; c < 4; ++c)
            if (16 * c < cols) TMEM_LD_X16(taddr + 64 * blk + 16 * c, (r + 16 * c));
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 64; e += 2)
            if (e < cols) mx4[(e >> 1) & 3] = fmaxf(mx4[(e >> 1) & 3], fmaxf(__uint_as_float(r[e]), __uint_as_float(r[e + 1])));
        }
        float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        // K rows of pad tokens are zero (TMA fill): their raw score is q . k_pad = S column 144.  (The zeros those
        // columns held in the maximum above only loosen the bound.)
        uint32_t s_pad = 0u;
        if (wpad) {
          TMEM_LD_X1(taddr + N12, s_pad);
          tmem_ld_wait();
          mx = fmaxf(mx, __uint_as_float(s_pad));
        }
        float pp_pad = 0.f;
        m_ub = fmaf(mx, c1s, bmax);
        float nm[2][2];
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
          for (int b = 0; b < 2; ++b)
            nm[a][b] = (((hm && (a != (int)my_h)) || (wm && (b != (int)my_w))) ? -100.0f * kLog2e : 0.0f) - m_ub;
        // pass 2: the three column blocks of P
        float2 l2 = make_float2(0.f, 0.f);
#pragma unroll
        for (int blk = 0; blk < 3; ++blk) {
          constexpr int kMaxCols = 64;
          const int cols = blk < 2 ? 64 : 16;
          uint32_t r[kMaxCols];
#pragma unroll
          for (int c = 0; c < kMaxCols / 16; ++c)
            if (16 * c < cols) TMEM_LD_X16(taddr + 64 * blk + 16 * c, (r + 16 * c));
          tmem_ld_wait();
          if (wpad) {
#pragma unroll
            for (int e = 0; e < kMaxCols; ++e)
              if (e < cols && (((pi >> row_i12<kPerm>(64 * blk + e)) | (pj >> row_j12<kPerm>(64 * blk + e))) & 1)) r[e] = s_pad;
          }
          uint32_t w[kMaxCols / 2];
#pragma unroll
          for (int e = 0; e < kMaxCols / 2; ++e) {
            if (2 * e < cols) {
              const int j0 = 64 * blk + 2 * e, j1 = j0 + 1;
              const int ci0 = row_i12<kPerm>(j0), cj0 = row_j12<kPerm>(j0), ci1 = row_i12<kPerm>(j1), cj1 = row_j12<kPerm>(j1);
              const float2 b2 = __fadd2_rn(make_float2(brow[-(ci0 * 23 + cj0)], brow[-(ci1 * 23 + cj1)]),
                                           make_float2(nm[ci0 >= kSplit][cj0 >= kSplit], nm[ci1 >= kSplit][cj1 >= kSplit]));
              const float2 t = __ffma2_rn(make_float2(__uint_as_float(r[2 * e]), __uint_as_float(r[2 * e + 1])), c1, b2);
              const float2 pp = make_float2(fast_exp2(t.x), fast_exp2(t.y));
              l2 = __fadd2_rn(l2, pp);
              w[e] = pack_bf16(pp.x, pp.y);
            }
          }
          if (wpad) {   // V rows of pad tokens are zero: their share of O is (their summed probability) x v_pad, below
#pragma unroll
            for (int e = 0; e < kMaxCols; ++e)
              if (e < cols && (((pi >> row_i12<kPerm>(64 * blk + e)) | (pj >> row_j12<kPerm>(64 * blk + e))) & 1))
                pp_pad += __uint_as_float((e & 1) ? (w[e >> 1] & 0xffff0000u) : (w[e >> 1] << 16));
          }
          if (row_ok) {
            const uint32_t p_row = sl + blk * kPBlock + row * 128;
#pragma unroll
            for (int ch = 0; ch < kMaxCols / 8; ++ch) {
              if (8 * ch < cols) {
                const uint32_t dst = p_row + ((ch ^ (row & 7)) << 4);
                asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(w[4 * ch]), "r"(w[4 * ch + 1]), "r"(w[4 * ch + 2]),
                             "r"(w[4 * ch + 3]) : "memory");
              }
            }
          }
        }
        l = l2.x + l2.y;
        if (has_pad && row_ok) {
          // P columns 144 / 145 = hi / lo bf16 halves of the pad probability (x V rows 144 / 145 = v_pad: the sum enters
          // with 16 mantissa bits), columns 146 .. 159 zero: chunks 2 and 3 of the row in column block 2
          const float hi = __uint_as_float(pack_bf16(pp_pad, 0.f) << 16);
          const uint32_t p_row = sl + 2 * kPBlock + row * 128;
          const uint32_t w8 = pack_bf16(hi, pp_pad - hi), z = 0u;
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%2,%2};" ::"r"(p_row + ((2 ^ (row & 7)) << 4)), "r"(w8), "r"(z) : "memory");
          asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(p_row + ((3 ^ (row & 7)) << 4)), "r"(z) : "memory");
        }
      }
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(bar_p + 8 * g);

      mbar_wait(bar_o + 8 * g, m & 1);
      tc_fence_after();
      uint32_t o[32];
      if (warp_on) {
        TMEM_LD_X16(taddr, (o + 0));
        TMEM_LD_X16(taddr + 16, (o + 16));
        tmem_ld_wait();
      }
      tc_fence_before();
      mbar_arrive(bar_r + 8 * g);

      if (warp_on && row_ok) {
        const float inv1 = 1.0f / l;
        const float2 inv = make_float2(inv1, inv1);
        const uint32_t o_row = sl + row * 64;
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          uint32_t v[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 s2 = __fmul2_rn(make_float2(__uint_as_float(o[8 * ch + 2 * e]), __uint_as_float(o[8 * ch + 2 * e + 1])), inv);
            v[e] = pack_bf16(s2.x, s2.y);
          }
          const uint32_t dst = o_row + (uint32_t)((ch ^ ((row >> 1) & 3)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]) : "memory");
        }
        if (p.lse != nullptr) {
          int y = it.wy * WS12 + ri + p.shift, x = it.wx * WS12 + rj + p.shift;
          if (y >= p.Hp) y -= p.Hp;
          if (x >= p.Wp) x -= p.Wp;
          p.lse[((int64_t)(it.b * p.Hp + y) * p.Wp + x) * p.nH + h] = (m_ub + log2f(l)) * 0.6931471805599453f;
        }
      }
      fence_proxy_async();
      named_barrier(1 + g, 128);
      if (q == 0 && elect_one()) {
        store_half12<kShifted>(maps.o, sl, p, it.b, it.wy, it.wx, h, half);
        bulk_commit();
      }
      __syncwarp();
    }
    if (q == 0 && elect_one()) bulk_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

}  // namespace tc12

// ---- host side ------------------------------------------------------------------
bool tc12_forward_supported(const Problem& p, int dtype) {
  if (dtype != SWIN_ATTN_BF16 || p.ws != 12 || p.mask != nullptr || p.dropout_p > 0.0f) return false;
  if (p.shift != 0 && p.shift != tc12::kSplit) return false;
  if (p.H < p.shift || p.W < p.shift) return false;   // the pad rows would wrap around the rolled image
  if ((p.q_ps | p.k_ps | p.v_ps | p.o_ps) & 7) return false;
  auto mis = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) != 0; };
  if (mis(p.q) || mis(p.k) || mis(p.v) || mis(p.o) || mis(p.k_pad) || mis(p.v_pad)) return false;
  return tc::tma_available();
}

cudaError_t launch_tc12_forward(const Problem& p, cudaStream_t st, const char** err) {
  tc12::TmaMaps12 maps;
  memset(&maps, 0, sizeof(maps));
  const bool shifted = p.shift != 0;
  bool ok = true;
  for (int i = 0; i < 3; ++i) {
    const int bx = tc12::kBoxX12[i], by = tc12::kBoxY12[i];
    ok = ok && tc::encode_window_map(&maps.q[i], p.q, p, p.q_ps, bx, by) && tc::encode_window_map(&maps.k[i], p.k, p, p.k_ps, bx, by) &&
         tc::encode_window_map(&maps.v[i], p.v, p, p.v_ps, bx, by);
  }
  for (int i = 1; i < 4; ++i) ok = ok && tc::encode_window_map(&maps.o[i], p.o, p, p.o_ps, tc12::kBoxX12[i], tc12::kBoxY12[i]);
  if (!ok) {
    if (err) *err = "cuTensorMapEncodeTiled failed";
    return cudaErrorInvalidValue;
  }
  const tc12::SmemMap12 map = tc12::smem_map_12();
  auto kernel = shifted ? tc12::tc12_fwd_kernel<true> : tc12::tc12_fwd_kernel<false>;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)map.total);
  if (e != cudaSuccess) {
    if (err) *err = "cudaFuncSetAttribute(max dynamic smem) failed";
    return e;
  }
  const int has_pad = (p.H != p.Hp || p.W != p.Wp) && (p.k_pad != nullptr || p.v_pad != nullptr) ? 1 : 0;
  int n_win, per, grid;
  tc::job_grid(p, tc::sm_count(), &n_win, &per, &grid);
  {
    TimedRegion timed(st);
    kernel<<<grid, tc12::kThreads12, map.total, st>>>(p, n_win, per, has_pad, maps);
  }
  count_launch();
  return cudaGetLastError();
}

}  // namespace swin
