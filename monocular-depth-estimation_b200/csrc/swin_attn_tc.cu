// tcgen05 forward kernel + launcher; helpers and the design notes are in swin_attn_tc.cuh.
#include "swin_attn_tc.cuh"

namespace swin {
namespace tc {

// stage layout (bytes): q0 0, q1 4096 | k0 8192, k1 8192+3584 | v0 16384, v1 20480 | ones 24576
__device__ __forceinline__ uint32_t q_off(int t) { return t * kTileBytes; }
__device__ __forceinline__ uint32_t k_off(int t) { return 2 * kTileBytes + t * kKTile1; }
__device__ __forceinline__ uint32_t v_off(int t) { return 4 * kTileBytes + t * kTileBytes; }

// One window of one tensor: 1 box (un-shifted) or the 4 parts of a shifted window.
template <bool kShifted>
__device__ __forceinline__ void load_window(const CUtensorMap* maps, uint32_t dst, uint32_t bar, const Problem& p,
                                            const TileCoord& c) {
  const int ch = c.h * kHeadDim;
  if (!kShifted) {
    tma_load_4d(dst, maps, bar, ch, c.wx * WS, c.wy * WS, c.b);
  } else {
    // token (i, j) sits at ((wy*7 + i + 3) mod Hp, (wx*7 + j + 3) mod Wp): rows/cols [0,4) never wrap,
    // rows/cols [4,7) start at wy*7 + 7, which wraps to 0 in the last window row/column.
    const int xa = c.wx * WS + kShift, ya = c.wy * WS + kShift;
    const int xb = c.wx == p.nWx - 1 ? 0 : c.wx * WS + WS, yb = c.wy == p.nWy - 1 ? 0 : c.wy * WS + WS;
    if (c.wy == p.nWy - 1) {
      tma_load_4d(dst, maps + 1, bar, ch, xa, ya, c.b);             // i<4,  j<4  -> rows [0,16)
      tma_load_4d(dst + 16 * 64, maps + 2, bar, ch, xa, yb, c.b);   // i>=4, j<4  -> rows [16,28)
      tma_load_4d(dst + 28 * 64, maps + 3, bar, ch, xb, ya, c.b);   // i<4,  j>=4 -> rows [28,40)
      tma_load_4d(dst + 40 * 64, maps + 4, bar, ch, xb, yb, c.b);   // i>=4, j>=4 -> rows [40,49)
    } else {                                                        // the 7 rows are contiguous in y
      tma_load_4d(dst, maps + 5, bar, ch, xa, ya, c.b);             // j<4  -> rows [0,28)
      tma_load_4d(dst + 28 * 64, maps + 6, bar, ch, xb, ya, c.b);   // j>=4 -> rows [28,49)
    }
  }
}

template <bool kPerm>
struct Patterns {
  uint64_t rowp[WS], colp[WS];   // tokens (as tile-row bits) of window row i / window column j
};
template <bool kPerm>
__device__ __forceinline__ uint64_t pattern_row(int i) {
  uint64_t m = 0;
#pragma unroll
  for (int j = 0; j < WS; ++j) m |= 1ull << token_row<kPerm>(i, j);
  return m;
}
template <bool kPerm>
__device__ __forceinline__ uint64_t pattern_col(int j) {
  uint64_t m = 0;
#pragma unroll
  for (int i = 0; i < WS; ++i) m |= 1ull << token_row<kPerm>(i, j);
  return m;
}

template <bool kShifted>
__global__ void __launch_bounds__(kThreads, 1)
tc_fwd_kernel(const Problem p, int n_tiles, int has_pad, const __grid_constant__ TmaMaps maps) {
  constexpr bool kPerm = kShifted;
  extern __shared__ __align__(1024) unsigned char smem[];
  const SmemMap map = smem_map(p.nH);
  const uint32_t sbase = smem_u32(smem);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform for the compiler as well
  const int n_pairs = (n_tiles + 1) >> 1;
  const int n_local = (n_pairs - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  const uint32_t bar_full = sbase + map.bars;              // [kStages] TMA (+ pad helper) -> MMA
  const uint32_t bar_empty = bar_full + 8 * kStages;       // [kStages] MMA -> producer
  const uint32_t bar_s = bar_empty + 8 * kStages;          // [kSlots]  S ready
  const uint32_t bar_p = bar_s + 8 * kSlots;               // [kSlots]  P written
  const uint32_t bar_o = bar_p + 8 * kSlots;               // [kSlots]  O ready
  const uint32_t bar_r = bar_o + 8 * kSlots;               // [kSlots]  O read, slot reusable

  // ---- one-time setup -------------------------------------------------------
  {
    // Rows the TMA boxes never write: K rows 49-55 feed S columns that are never read unless the pad helper
    // filled row 49; V rows 49-63 are multiplied by P columns that are 0 and must therefore be finite.
    for (int i = tid; i < kStages * 2 * 15 * 4; i += kThreads) {
      const int chunk = i & 3, rr = (i >> 2) % 15, t = ((i >> 2) / 15) & 1, s = (i >> 2) / 30;
      const uint32_t st = map.stages + s * kStageBytes;
      *reinterpret_cast<uint4*>(smem + st + v_off(t) + (49 + rr) * 64 + chunk * 16) = make_uint4(0u, 0u, 0u, 0u);
      if (rr < 7) *reinterpret_cast<uint4*>(smem + st + k_off(t) + (49 + rr) * 64 + chunk * 16) = make_uint4(0u, 0u, 0u, 0u);
    }
    // ones tile of every stage: element 0 of each key row is 1.0 (bf16), so column 64 of O = sum_j P[i, j]
    for (int i = tid; i < kStages * 64 * 4; i += kThreads) {
      const int chunk = i & 3, rr = (i >> 2) & 63, s = i >> 8;
      const uint4 v = make_uint4(chunk == 0 ? 0x00003f80u : 0u, 0u, 0u, 0u);
      *reinterpret_cast<uint4*>(smem + map.stages + s * kStageBytes + 6 * kTileBytes + rr * 64 + ((chunk ^ ((rr >> 1) & 3)) << 4)) = v;
    }
    float* bias = reinterpret_cast<float*>(smem + map.bias);
    for (int i = tid; i < p.nH * 169; i += kThreads) {
      const int e = i / p.nH, h = i - e * p.nH;
      bias[h * 169 + e] = p.table[i] * kLog2e;
    }
    for (int i = tid; i < 2 * p.nH * 4; i += kThreads) {   // 16-byte chunks of k_pad | v_pad
      const int isv = i >= p.nH * 4, j = i - isv * p.nH * 4;
      const __nv_bfloat16* src = static_cast<const __nv_bfloat16*>(isv ? p.v_pad : p.k_pad);
      *reinterpret_cast<uint4*>(smem + map.padv + i * 16) =
          src != nullptr ? *reinterpret_cast<const uint4*>(src + j * 8) : make_uint4(0u, 0u, 0u, 0u);
    }
    if (tid == 0) {
      for (int i = 0; i < kStages; ++i) {
        mbar_init(bar_full + 8 * i, has_pad ? 4 : 2);
        mbar_init(bar_empty + 8 * i, 1);
      }
      for (int i = 0; i < kSlots; ++i) {
        mbar_init(bar_s + 8 * i, 1);
        mbar_init(bar_p + 8 * i, 128);
        mbar_init(bar_o + 8 * i, 1);
        mbar_init(bar_r + 8 * i, 128);
      }
      fence_barrier_init();
      for (int i = kShifted ? 1 : 0; i < (kShifted ? 7 : 1); ++i) {
        tma_prefetch_desc(&maps.q[i]);
        tma_prefetch_desc(&maps.k[i]);
        tma_prefetch_desc(&maps.v[i]);
      }
    }
    if (warp == 1) tmem_alloc(sbase + map.tmem_ptr, 512);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + map.tmem_ptr);

  if (warp == 0 || warp == 3) {
    // ===== TMA producers: warp 0 -> tile 0 of every pair, warp 3 -> tile 1 =====================
    // lane 0 issues the TMA boxes; with padding, lanes 8-15 copy k_pad / v_pad of the tile's head into the
    // spare row 49 of the K / V tile (see swin_attn_tc.cuh).
    const int t = warp == 0 ? 0 : 1;
    const int isv = (lane >> 2) & 1, chunk = lane & 3;
    TileIter it;
    it.init(p, 2 * (int)blockIdx.x + t, 2 * (int)gridDim.x);
    for (int n = 0; n < n_local; ++n, it.next(p)) {
      const int s = n % kStages;
      const uint32_t st = sbase + map.stages + s * kStageBytes, bar = bar_full + 8 * s;
      const TileCoord c = it.coord(n_tiles);
      mbar_wait(bar_empty + 8 * s, ((n / kStages) & 1) ^ 1);
      if (elect_one()) {
        if (t == 0) TC_TRACE(p, n, 0);
        if (c.valid) {
          mbar_expect_tx(bar, 3 * kBoxBytes);
          load_window<kShifted>(maps.q, st + q_off(t), bar, p, c);
          load_window<kShifted>(maps.k, st + k_off(t), bar, p, c);
          load_window<kShifted>(maps.v, st + v_off(t), bar, p, c);
        } else {
          mbar_arrive(bar);
        }
        if (t == 0) TC_TRACE(p, n, 1);
      }
      if (has_pad) {
        __syncwarp();
        if (lane >= 8 && lane < 16) {
          const uint4 v = *reinterpret_cast<const uint4*>(smem + map.padv + ((isv * p.nH + c.h) * 4 + chunk) * 16);
          const uint32_t dst = st + (isv ? v_off(t) : k_off(t)) + 49 * 64 + chunk * 16;
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
          fence_proxy_async();
        }
        __syncwarp();
        if (lane == 8) mbar_arrive(bar);
      }
    }
  } else if (warp == 1) {
    // ===== S = Q K^T issuer (one thread, 2 instructions per pair) ===============================
    constexpr uint32_t idesc_s = instr_desc(128, 112, 0);
    for (int n = 0; n < n_local; ++n) {
      const int s = n % kStages, a = n % kSlots;
      mbar_wait(bar_full + 8 * s, (n / kStages) & 1);
      mbar_wait(bar_r + 8 * a, ((n / kSlots) & 1) ^ 1);
      tc_fence_after();
      const uint32_t st = sbase + map.stages + s * kStageBytes;
      const uint32_t d = tmem_base + a * kSlotCols;
      const uint64_t qd = smem_desc(st + q_off(0), 16, 512, 4);   // 128 rows: q0 | q1
      const uint64_t kd = smem_desc(st + k_off(0), 16, 512, 4);   // 112 rows: k0 (56) | k1 (56)
      if (elect_one()) {
        TC_TRACE(p, n, 3);
#pragma unroll
        for (int k = 0; k < 2; ++k) umma_ss(d, qd + 2 * k, kd + 2 * k, idesc_s, k);
        umma_commit(bar_s + 8 * a);
        TC_TRACE(p, n, 4);
      }
      __syncwarp();
    }
  } else if (warp == 2) {
    // ===== O = P V issuer (one thread, 4 instructions per pair) =================================
    constexpr uint32_t idesc_o = instr_desc(128, 96, 1);
    for (int m = 0; m < n_local; ++m) {
      const int s = m % kStages, a = m % kSlots;
      mbar_wait(bar_p + 8 * a, (m / kSlots) & 1);
      tc_fence_after();
      const uint32_t st = sbase + map.stages + s * kStageBytes;
      const uint32_t d = tmem_base + a * kSlotCols;                // O overwrites the S columns
      const uint64_t pd = smem_desc(sbase + map.pslots + a * 2 * kPTileBytes, 16, 1024, 2);  // 128 rows: p0 | p1
      const uint64_t vd = smem_desc(st + v_off(0), kTileBytes, 512, 4);   // N = 96: v0 | v1 | ones, one tile apart
      if (elect_one()) {
        TC_TRACE(p, m, 5);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ss(d, pd + 2 * k, vd + 64 * k, idesc_o, k);
        umma_commit(bar_o + 8 * a);
        umma_commit(bar_empty + 8 * s);
        TC_TRACE(p, m, 6);
      }
      __syncwarp();
    }
  } else if (warp >= 4) {
    // ===== softmax / epilogue warpgroups =======================================
    const int g = (warp >> 2) - 1;          // slot owned by this warpgroup
    const int q = warp & 3;                 // TMEM sub-partition: lanes 32q .. 32q+31
    const int sel = q >> 1;                 // which tile of the pair (rows 0-63 / 64-127 of the accumulator)
    const int row = (q & 1) * 32 + lane;    // row of this tile's S / O held by this thread
    const int rowc = row < N ? row : N - 1;
    const int ri = row_i<kPerm>(rowc), rj = row_j<kPerm>(rowc);
    const int base_i = ri * 13 + rj + 84;   // relative_position_index[row, col] = base_i - (ci*13 + cj)
    const float* bias = reinterpret_cast<const float*>(smem + map.bias);
    const float2 c1 = make_float2(p.scale * kLog2e, p.scale * kLog2e);
    const uint64_t all = (1ull << N) - 1;
    // SW-MSA regions (newcrf_layers.py:335-345) as tile-row bit sets: window rows / columns >= ws - shift
    uint64_t hb = 0, wb = 0;
    if (kShifted) {
#pragma unroll
      for (int i = WS - kShift; i < WS; ++i) hb |= pattern_row<kPerm>(i);
#pragma unroll
      for (int j = WS - kShift; j < WS; ++j) wb |= pattern_col<kPerm>(j);
    }
    // Padding tokens (y >= H or x >= W) can only occur in the last two window rows / columns: their
    // tile-row bit sets, for window row nWy-2 ([0]) and nWy-1 ([1]), likewise for columns.
    uint64_t pad_y[2] = {0, 0}, pad_x[2] = {0, 0};
    if (has_pad) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
#pragma unroll
        for (int i = 0; i < WS; ++i) {
          int yy = (p.nWy - 2 + e) * WS + i + p.shift, xx = (p.nWx - 2 + e) * WS + i + p.shift;
          if (yy >= p.Hp) yy -= p.Hp;
          if (xx >= p.Wp) xx -= p.Wp;
          if (yy >= p.H && p.nWy - 2 + e >= 0) pad_y[e] |= pattern_row<kPerm>(i);
          if (xx >= p.W && p.nWx - 2 + e >= 0) pad_x[e] |= pattern_col<kPerm>(i);
        }
      }
    }
    const uint32_t p_row = sbase + map.pslots + (g * 2 + sel) * kPTileBytes + row * 128;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + g * kSlotCols;
    const uint32_t s_addr = taddr + sel * 56, o_addr = taddr + sel * 32;   // this tile's columns of S and of O
    TileIter it;
    it.init(p, 2 * ((int)blockIdx.x + g * (int)gridDim.x) + sel, 2 * kSlots * (int)gridDim.x);
    __nv_bfloat16* out = static_cast<__nv_bfloat16*>(p.o);

    int m = 0;
    for (int n = g; n < n_local; n += kSlots, ++m, it.next(p)) {
      const TileCoord c = it.coord(n_tiles);
      uint64_t masked = 0, padded = 0;
      if (kShifted) {
        const uint64_t hm = c.wy == p.nWy - 1 ? hb : 0, wm = c.wx == p.nWx - 1 ? wb : 0;
        masked = (((hm >> rowc) & 1) ? (~hm & all) : hm) | (((wm >> rowc) & 1) ? (~wm & all) : wm);
      }
      if (has_pad)
        padded = (c.wy == p.nWy - 1 ? pad_y[1] : c.wy == p.nWy - 2 ? pad_y[0] : 0) |
                 (c.wx == p.nWx - 1 ? pad_x[1] : c.wx == p.nWx - 2 ? pad_x[0] : 0);
      // every row of a tile with a mask / padding has masked columns, so this is warp-uniform
      const bool special = (kShifted && (c.wy == p.nWy - 1 || c.wx == p.nWx - 1)) || padded != 0;
      const float* brow = bias + c.h * 169 + base_i;

      const bool tr = q == 0 && lane == 0;   // timeline of tile 0, row 0
      if (tr) TC_TRACE(p, n, 7);
      mbar_wait(bar_s + 8 * g, m & 1);
      tc_fence_after();
      if (tr) TC_TRACE(p, n, 8);
      uint32_t r[N + 15];
      TMEM_LD_X16(s_addr, (r + 0));
      TMEM_LD_X16(s_addr + 16, (r + 16));
      TMEM_LD_X16(s_addr + 32, (r + 32));
      TMEM_LD_X2(s_addr + 48, r[48], r[49]);   // column 49 = q . k_pad (spare K row, see the producers)
      tmem_ld_wait();
      if (tr) TC_TRACE(p, n, 9);

      // pass 1: t = S * scale*log2e + bias*log2e (+ mask), row max.  Packed fp32x2 FMAs.
      float mx = -INFINITY;
      if (!special) {
#pragma unroll
        for (int j = 0; j < N - 1; j += 2) {
          const int cj0 = row_i<kPerm>(j) * 13 + row_j<kPerm>(j), cj1 = row_i<kPerm>(j + 1) * 13 + row_j<kPerm>(j + 1);
          const float2 t = __ffma2_rn(make_float2(__uint_as_float(r[j]), __uint_as_float(r[j + 1])), c1,
                                      make_float2(brow[-cj0], brow[-cj1]));
          r[j] = __float_as_uint(t.x);
          r[j + 1] = __float_as_uint(t.y);
          mx = fmaxf(mx, fmaxf(t.x, t.y));
        }
        {
          const int cj = row_i<kPerm>(N - 1) * 13 + row_j<kPerm>(N - 1);
          const float t = fmaf(__uint_as_float(r[N - 1]), c1.x, brow[-cj]);
          r[N - 1] = __float_as_uint(t);
          mx = fmaxf(mx, t);
        }
      } else {
        const float a49 = __uint_as_float(r[49]);
#pragma unroll
        for (int j = 0; j < N; ++j) {
          const int cj = row_i<kPerm>(j) * 13 + row_j<kPerm>(j);
          const float acc = ((padded >> j) & 1) ? a49 : __uint_as_float(r[j]);
          float t = fmaf(acc, c1.x, brow[-cj]);
          if ((masked >> j) & 1) t += -100.0f * kLog2e;
          r[j] = __float_as_uint(t);
          mx = fmaxf(mx, t);
        }
      }
      r[49] = __float_as_uint(-INFINITY);   // exp2 -> 0
      // pass 2: p = exp2(t - max) evaluated as bf16x2 (P feeds a bf16 MMA anyway); the row sum comes out of
      // the PV MMA through the ones column, i.e. it is the sum of exactly the rounded values in P.
      const float2 nmx = make_float2(-mx, -mx);
      uint32_t w[25];
#pragma unroll
      for (int e = 0; e < 25; ++e) {
        const float2 x = __fadd2_rn(make_float2(__uint_as_float(r[2 * e]), __uint_as_float(r[2 * e + 1])), nmx);
        w[e] = ex2_bf16x2(pack_bf16(x.x, x.y));
      }
      if (special && padded != 0) {
        // pad columns: their V rows are zero (TMA fill); move their probability mass to column 49 (x v_pad)
        float pad_sum = 0.f;
#pragma unroll
        for (int j = 0; j < N; ++j) {
          if ((padded >> j) & 1) {
            pad_sum += __uint_as_float((j & 1) ? (w[j >> 1] & 0xffff0000u) : (w[j >> 1] << 16));
            w[j >> 1] &= (j & 1) ? 0x0000ffffu : 0xffff0000u;
          }
        }
        w[24] = (w[24] & 0x0000ffffu) | (pack_bf16(0.f, pad_sum) & 0xffff0000u);
      }
      // P row (64 bf16, columns >= 50 zero) into the 128B-swizzled A tile of the PV MMA
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        const uint32_t dst = p_row + ((ch ^ (row & 7)) << 4);
        const uint32_t w0 = 4 * ch < 25 ? w[4 * ch < 25 ? 4 * ch : 0] : 0u, w1 = 4 * ch + 1 < 25 ? w[4 * ch + 1 < 25 ? 4 * ch + 1 : 0] : 0u;
        const uint32_t w2 = 4 * ch + 2 < 25 ? w[4 * ch + 2 < 25 ? 4 * ch + 2 : 0] : 0u, w3 = 4 * ch + 3 < 25 ? w[4 * ch + 3 < 25 ? 4 * ch + 3 : 0] : 0u;
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
      }
      if (tr) TC_TRACE(p, n, 10);
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(bar_p + 8 * g);
      if (tr) TC_TRACE(p, n, 11);

      // un-shifted pixel of this thread's token (computed while the PV MMA runs)
      int y = c.wy * WS + ri + p.shift, x = c.wx * WS + rj + p.shift;
      if (y >= p.Hp) y -= p.Hp;
      if (x >= p.Wp) x -= p.Wp;

      mbar_wait(bar_o + 8 * g, m & 1);
      tc_fence_after();
      if (tr) TC_TRACE(p, n, 12);
      uint32_t o[32];
      TMEM_LD_X16(o_addr, (o + 0));
      TMEM_LD_X16(o_addr + 16, (o + 16));
      const float l = __uint_as_float(tmem_ld_x1(taddr + 64));   // row sum: P x ones column
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(bar_r + 8 * g);
      if (tr) TC_TRACE(p, n, 13);

      if (c.valid && row < N) {
        const float inv1 = 1.0f / l;
        const float2 inv = make_float2(inv1, inv1);
        if (y < p.H && x < p.W) {
          __nv_bfloat16* dst = out + ((int64_t)(c.b * p.H + y) * p.W + x) * p.o_ps + c.h * kHeadDim;
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            uint32_t v[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 s2 = __fmul2_rn(make_float2(__uint_as_float(o[8 * ch + 2 * e]), __uint_as_float(o[8 * ch + 2 * e + 1])), inv);
              v[e] = pack_bf16(s2.x, s2.y);
            }
            *reinterpret_cast<uint4*>(dst + 8 * ch) = make_uint4(v[0], v[1], v[2], v[3]);
          }
        }
        if (p.lse != nullptr)
          p.lse[((int64_t)(c.b * p.Hp + y) * p.Wp + x) * p.nH + c.h] = (mx + log2f(l)) * 0.6931471805599453f;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

}  // namespace tc

// ---- host side ------------------------------------------------------------------
namespace {

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime's driver entry point: no link-time dependency on libcuda.
EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      ptr = nullptr;
    return reinterpret_cast<EncodeTiledFn>(ptr);
  }();
  return fn;
}

// Tensor map over an image-order (B, H, W, pixel_stride) bf16 tensor whose base pointer already includes
// the channel offset of q / k / v; box = (32 channels, bx pixels, by rows, 1 image), 64B swizzle,
// out-of-bounds (the padding the reference adds with F.pad) zero-filled.
}  // namespace

bool tc::encode_window_map(CUtensorMap* out, const void* base, const Problem& p, int64_t pix_stride, int bx, int by) {
  if (encode_fn() == nullptr) return false;
  // the channel extent is what remains of the pixel after this tensor's channel offset; use C: boxes of
  // head h start at channel 32 h and never leave [0, C)
  const cuuint64_t dims[4] = {(cuuint64_t)p.C, (cuuint64_t)p.W, (cuuint64_t)p.H, (cuuint64_t)p.B};
  const cuuint64_t strides[3] = {(cuuint64_t)pix_stride * 2, (cuuint64_t)pix_stride * 2 * p.W,
                                 (cuuint64_t)pix_stride * 2 * p.W * p.H};
  const cuuint32_t box[4] = {(cuuint32_t)kHeadDim, (cuuint32_t)bx, (cuuint32_t)by, 1u};
  const cuuint32_t estr[4] = {1u, 1u, 1u, 1u};
  return encode_fn()(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool tc::tma_available() { return encode_fn() != nullptr; }

bool tc_forward_supported(const Problem& p, int dtype) {
  if (dtype != SWIN_ATTN_BF16 || p.ws != 7 || p.mask != nullptr) return false;
  if (p.shift != 0 && p.shift != tc::kShift) return false;
  // TMA: 16-byte aligned base pointers and strides; vector stores of o
  if ((p.q_ps | p.k_ps | p.v_ps | p.o_ps) & 7) return false;
  auto mis = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) != 0; };
  if (mis(p.q) || mis(p.k) || mis(p.v) || mis(p.o) || mis(p.k_pad) || mis(p.v_pad)) return false;
  if (encode_fn() == nullptr) return false;
  return tc::smem_map(p.nH).total <= 227 * 1024;
}

cudaError_t launch_tc_forward(const Problem& p, cudaStream_t st, const char** err) {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  tc::TmaMaps maps;
  memset(&maps, 0, sizeof(maps));
  const bool shifted = p.shift != 0;
  const int a = tc::WS - tc::kShift, s = tc::kShift;
  const int bx[7] = {tc::WS, a, a, s, s, a, s}, by[7] = {tc::WS, a, s, a, s, tc::WS, tc::WS};
  for (int i = shifted ? 1 : 0; i < (shifted ? 7 : 1); ++i) {
    if (!tc::encode_window_map(&maps.q[i], p.q, p, p.q_ps, bx[i], by[i]) || !tc::encode_window_map(&maps.k[i], p.k, p, p.k_ps, bx[i], by[i]) ||
        !tc::encode_window_map(&maps.v[i], p.v, p, p.v_ps, bx[i], by[i])) {
      if (err) *err = "cuTensorMapEncodeTiled failed";
      return cudaErrorInvalidValue;
    }
  }
  const tc::SmemMap map = tc::smem_map(p.nH);
  auto kernel = shifted ? tc::tc_fwd_kernel<true> : tc::tc_fwd_kernel<false>;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)map.total);
  if (e != cudaSuccess) {
    if (err) *err = "cudaFuncSetAttribute(max dynamic smem) failed";
    return e;
  }
  const int has_pad = (p.H != p.Hp || p.W != p.Wp) && (p.k_pad != nullptr || p.v_pad != nullptr) ? 1 : 0;
  const int n_tiles = p.B * p.nWy * p.nWx * p.nH;
  const int n_pairs = (n_tiles + 1) / 2;
  const int grid = n_pairs < sms ? n_pairs : sms;
  kernel<<<grid, tc::kThreads, map.total, st>>>(p, n_tiles, has_pad, maps);
  count_launch();
  return cudaGetLastError();
}

}  // namespace swin
