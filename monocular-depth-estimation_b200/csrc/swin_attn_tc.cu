// tcgen05 forward of the shifted-window attention (bf16, ws 7, shift 0 / 3) for sm_100a.
// PTX wrappers, the window -> TMA box mapping and the token order are in swin_attn_tc.cuh.
//
// Tile = one (window, head): Q, K, V are 49 x 32 bf16.  One CTA = one JOB = (head h, contiguous
// range of windows); consecutive CTAs are consecutive heads over the SAME windows, so the 128-byte
// lines two heads share are fetched at about the same time.  Two windows of the job form a PAIR
// stacked along M: tile 0 owns TMEM lanes 0-63, tile 1 lanes 64-127, and one `tcgen05.ld.32x32b`
// hands every softmax thread one complete row of S (no cross-lane reduction).
//
//   producers (warps 0 / 3, one elected lane): window partition, cyclic shift, right/bottom zero
//       padding are TMA tensor-map addressing over the image-order (B,H,W,channels) tensors
//       (swin_attn_tc.cuh::move_window); two rings: 5 stages of (q0 q1 | k0 k1), 7 of (v0 v1 | ones).
//   pad tokens carry the Linear bias in the reference (newcrf_layers.py:212-217) while TMA zero-fills
//       them.  All pad tokens of a head share ONE vector and the head is fixed per CTA: once a K / V
//       tile has landed, the warp that issues its MMA overwrites the rows of the tile's pad tokens with
//       k_pad / v_pad (write_pad_rows), so both MMAs see the reference's operands and the softmax has no
//       padding code path (an earlier version substituted q . k_pad per column in registers: ~620 extra
//       instructions per row on half of the stage-2 tiles).
//   MMA issuers (warps 1 / 2, one elected lane):
//       S = [Q0;Q1] [K0;K1]^T   (M128 N112 K32)   tile t: rows 64t.., columns 56t..
//       O = [P0;P1] [V0|V1|1]   (M128 N96  K64, A = P in TMEM, V MN-major)   tile t: rows 64t.., columns 32t..;
//           column 64 multiplies a constant ones tile and holds the softmax row sums
//   softmax warpgroups (4 x 4 warps, one accumulator slot each): scale + relative position bias (the
//       head's bias expanded to a [row][col] matrix in smem, read with 128-bit loads) + analytic SW-MSA
//       mask, exact row max, exp2, P as bf16 written with tcgen05.st into the slot's spare TMEM columns
//       (the PV MMA takes its A operand from tensor memory: P never touches shared memory); then the
//       O row x 1/rowsum -> bf16 -> 64B-swizzled staging tile -> TMA tensor store with the same boxes as
//       the loads (window reverse, inverse shift and crop are tensor-map addressing as well; out-of-bounds
//       = padding tokens are clipped by the TMA unit).
#include <cstdlib>

#include "swin_attn_tc.cuh"

namespace swin {
namespace tcf {

using namespace tc;

// Two rings: Q/K tiles live only until the S MMA has consumed them, V tiles until the PV MMA has.
constexpr int kQKStages = 5;     // (q0 q1 | k0 k1): pairs prefetched ahead of the S MMA
constexpr int kVStages = 7;      // (v0 v1 | ones): 4 pairs inside the softmax warpgroups + 3 prefetched
constexpr int kSlotsF = 4;
constexpr int kThreadsF = 128 + 128 * kSlotsF;
constexpr int kFillF = kThreadsF - 64;         // threads of the non-producer warps
constexpr int kQKBytes = 4 * kTileBytes;
constexpr int kVBytes = 3 * kTileBytes;
// One accumulator slot = 128 TMEM columns: S in [0, 112); O (PV output, N = 96) overwrites [0, 96) once S has been
// read; P (bf16, two per column: 64 K-values = 32 columns) lives in [96, 128) and is the A operand of the PV MMA,
// read straight from tensor memory -- P never touches shared memory.  O goes through a 64B-swizzled staging tile per
// (slot, tile) and a TMA tensor store (a direct store from registers was measured: every STG.128 of a warp touches 32
// different 64-byte rows, ~2000 LSU wavefront-cycles per pair against ~500 for the staging stores).
constexpr int kSlotColsF = 128;
constexpr int kPColF = 96;
constexpr int kBiasStride = 52;                // floats per row of the expanded bias matrix (49 used; 16-byte aligned rows)

__device__ __forceinline__ uint32_t q_off(int t) { return t * kTileBytes; }
__device__ __forceinline__ uint32_t k_off(int t) { return 2 * kTileBytes + t * kKTile1; }   // in a QK stage
__device__ __forceinline__ uint32_t v_off(int t) { return t * kTileBytes; }                    // in a V stage

struct TmaMapsF {
  CUtensorMap q[7], k[7], v[7];   // loads
  CUtensorMap o[7];               // stores
};

struct SmemMapF {
  uint32_t qk, v, stage, bias, pads, bars, tmem_ptr, total;
};
__host__ __device__ inline SmemMapF smem_map_f() {
  SmemMapF m;
  m.qk = 0;
  m.v = m.qk + kQKStages * kQKBytes;
  m.stage = m.v + kVStages * kVBytes;                       // O staging: one 64-row x 64 B tile per (slot, tile of the pair)
  m.bias = m.stage + kSlotsF * 2 * kTileBytes;              // bias[row][col] * log2e of this CTA's head, tile order, fp32
  m.pads = m.bias + ((N * kBiasStride * 4 + 127) & ~127);   // PadSets: y[0], y[1], x[0], x[1] (uint64 tile-row bit sets)
  m.bars = m.pads + 32;
  m.tmem_ptr = m.bars + 8 * (2 * kQKStages + 2 * kVStages + 4 * kSlotsF);
  m.total = m.tmem_ptr + 16;
  return m;
}

template <bool kShifted, bool kDrop>
__global__ void __launch_bounds__(kThreadsF, 1)
tc_fwd_kernel(const Problem p, int n_win, int win_per_job, int has_pad, const __grid_constant__ TmaMapsF maps) {
  constexpr bool kPerm = kShifted;
  extern __shared__ __align__(1024) unsigned char smem[];
  const SmemMapF map = smem_map_f();
  const uint32_t sbase = smem_u32(smem);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform for the compiler as well
  const int h = (int)blockIdx.x % p.nH;
  const int w0 = ((int)blockIdx.x / p.nH) * win_per_job;
  const int w1 = min(w0 + win_per_job, n_win);
  const int n_pairs = (w1 - w0 + 1) >> 1;
  if (tid == 0) TC_TRACE(p, 0, 14);   // kernel entry
  if (tid == 0) TC_TRACE_CTA(p, 0);

  const uint32_t bar_qk_full = sbase + map.bars;               // [kQKStages] TMA -> S MMA
  const uint32_t bar_qk_empty = bar_qk_full + 8 * kQKStages;   // [kQKStages] S MMA -> producers
  const uint32_t bar_v_full = bar_qk_empty + 8 * kQKStages;    // [kVStages]  TMA -> PV MMA
  const uint32_t bar_v_empty = bar_v_full + 8 * kVStages;      // [kVStages]  PV MMA -> producers
  const uint32_t bar_s = bar_v_empty + 8 * kVStages;           // [kSlotsF]  S ready
  const uint32_t bar_p = bar_s + 8 * kSlotsF;              // [kSlotsF]  P written
  const uint32_t bar_o = bar_p + 8 * kSlotsF;              // [kSlotsF]  O ready
  const uint32_t bar_r = bar_o + 8 * kSlotsF;              // [kSlotsF]  O read, slot reusable

  // ---- one-time setup -------------------------------------------------------
  // Everything up to pdl_wait() touches only kernel parameters and shared memory, so under programmatic dependent
  // launch it overlaps the tail of the previous kernel in the stream.  The tensor-map descriptors are prefetched
  // first (their fetch is on the critical path of the first TMA load), 36 threads initialise one mbarrier each;
  // after one CTA barrier the producers start loading right away while the other warps fill the bytes TMA never
  // writes, stage the bias and allocate TMEM, then meet on a named barrier (the MMA issuers are among them, so no
  // MMA reads a tile before the fills are complete).
  if ((warp == 0 || warp == 3) && lane == 0) {
    for (int i = kShifted ? 1 : 0; i < (kShifted ? 7 : 1); ++i) {
      if (warp == 0) tma_prefetch_desc(&maps.q[i]);
      if (warp == 0) tma_prefetch_desc(&maps.k[i]);
      if (warp == 3) tma_prefetch_desc(&maps.v[i]);
      if (warp == 3) tma_prefetch_desc(&maps.o[i]);
    }
  }
  if (warp == 1) {
    constexpr int kBars = 2 * kQKStages + 2 * kVStages + 4 * kSlotsF;
    static_assert(kBars <= 64, "two barriers per lane at most");
    for (int i = lane; i < kBars; i += 32) {
      // counts: qk_full 2, qk_empty 1, v_full 2, v_empty 1, s 1, p 128, o 1, r 128 (layout of the bar_* addresses above)
      uint32_t count;
      if (i < kQKStages) count = 2;
      else if (i < 2 * kQKStages) count = 1;
      else if (i < 2 * kQKStages + kVStages) count = 2;
      else if (i < 2 * kQKStages + 2 * kVStages) count = 1;
      else {
        const int k = (i - 2 * kQKStages - 2 * kVStages) / kSlotsF;
        count = (k == 1 || k == 3) ? 128 : 1;
      }
      mbar_init(bar_qk_full + 8 * i, count);
    }
    fence_barrier_init();
  }
  __syncthreads();
  if (warp == 0 || warp == 3) pdl_wait();   // the producers read q / k / v: the previous kernel's output
  if (warp != 0 && warp != 3) {
    const int t14 = tid - (warp < 3 ? 32 : 64);   // 0 .. kFillF-1 over the non-producer warps
    // Global loads first, so that their latency overlaps the fills: the head's relative-position bias (a parameter)
    // expanded to bias[row][col] in TILE order (newcrf_layers.py:124-127: table[index[i, j], h]), pre-multiplied by
    // log2(e).  Every softmax thread owns one row and reads it with 128-bit loads: 13 instead of 49 shared-memory
    // instructions per row and pair.
    constexpr int kBiasIters = (N * kBiasStride + kFillF - 1) / kFillF;
    float bias_v[kBiasIters];
#pragma unroll
    for (int it = 0; it < kBiasIters; ++it) {
      const int idx = t14 + it * kFillF;
      const int r = idx / kBiasStride, c = idx - r * kBiasStride;
      bias_v[it] = 0.f;
      if (r < N && c < N) {
        const int ti = (row_i<kPerm>(r) - row_i<kPerm>(c) + WS - 1) * (2 * WS - 1) + (row_j<kPerm>(r) - row_j<kPerm>(c) + WS - 1);
        bias_v[it] = p.table[ti * p.nH + h] * kLog2e;
      }
    }
    if (warp == 1) tmem_alloc(sbase + map.tmem_ptr, 512);
    // Rows 49-63 of the K / V tiles (TMA boxes write rows 0-48 only) must be finite: K rows 49-55 feed S
    // columns that are never read, V rows 49-63 are multiplied by P columns that are 0.  K tile 1 starts at row 56 of
    // K tile 0.
    for (int i = t14; i < kQKStages * 2 * 7 * 4; i += kFillF) {          // K rows 49-55
      const int chunk = i & 3, rr = (i >> 2) % 7, t = ((i >> 2) / 7) & 1, st = (i >> 2) / 14;
      const int r50 = 49 + rr;
      *reinterpret_cast<uint4*>(smem + map.qk + st * kQKBytes + k_off(t) + r50 * 64 + ((chunk ^ ((r50 >> 1) & 3)) << 4)) =
          make_uint4(0u, 0u, 0u, 0u);
    }
    for (int i = t14; i < kVStages * 2 * 15 * 4; i += kFillF) {          // V rows 49-63
      const int chunk = i & 3, rr = (i >> 2) % 15, t = ((i >> 2) / 15) & 1, st = (i >> 2) / 30;
      const int r50 = 49 + rr;
      *reinterpret_cast<uint4*>(smem + map.v + st * kVBytes + v_off(t) + r50 * 64 + ((chunk ^ ((r50 >> 1) & 3)) << 4)) =
          make_uint4(0u, 0u, 0u, 0u);
    }
    // ones tile of every V stage (64B-swizzled, MN-major B of the PV MMA): element 0 of each key row is 1.0
    // (bf16), so column 64 of O = sum_j P[i, j] -- except the spare key rows 49 / 50 (see the pad vectors below)
    for (int i = t14; i < kVStages * 64 * 4; i += kFillF) {
      const int chunk = i & 3, rr = (i >> 2) & 63, st = i >> 8;
      *reinterpret_cast<uint4*>(smem + map.v + st * kVBytes + 2 * kTileBytes + rr * 64 + ((chunk ^ ((rr >> 1) & 3)) << 4)) =
          make_uint4(chunk == 0 && rr != N && rr != N + 1 ? 0x00003f80u : 0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int it = 0; it < kBiasIters; ++it) {
      const int idx = t14 + it * kFillF;
      if (idx < N * kBiasStride) reinterpret_cast<float*>(smem + map.bias)[idx] = bias_v[it];
    }
    fence_proxy_async();
    tc_fence_before();
    named_barrier(9, kFillF);
    tc_fence_after();
    pdl_wait();   // the softmax warps write o (and lse); the pad vectors may come from a cast kernel: after the dependency
    if (has_pad) {
      // Pad tokens carry the projection bias (the reference pads before the Linear) while TMA zero-fills them.  All
      // pad tokens of a head share ONE vector and the head is fixed per CTA, so the vectors go -- once -- into the
      // spare row 49 of every K / V tile (rows 49+ are never written by the TMA boxes):
      //   S column 49 = q . k_pad: the softmax substitutes it for the columns of pad tokens (whose K rows are 0);
      //   P columns 49 / 50 = hi / lo bf16 halves of the row's summed probability over pad tokens (their V rows are 0),
      //   times V rows 49 / 50 = v_pad -- two halves, so that the sum enters with 16 mantissa bits like the fp32
      //   accumulation of the individual products would.
      // The ones tile has 0 in rows 49 / 50, so the row sum is not affected.  No per-pair work is left in the MMA issuers.
      constexpr int kKRows = kQKStages * 2 * 4, kVRows = kVStages * 2 * 2 * 4;   // x 4 chunks of 16 B; V: rows 49 and 50
      if (t14 < kKRows + kVRows) {
        const bool is_k = t14 < kKRows;
        const int i = is_k ? t14 : t14 - kKRows;
        const int chunk = i & 3, t = (i >> 2) & 1;
        const int st = is_k ? i >> 3 : i >> 4, rowx = is_k ? N : N + ((i >> 3) & 1);
        const void* vecp = is_k ? p.k_pad : p.v_pad;
        const uint4 val = vecp == nullptr ? make_uint4(0u, 0u, 0u, 0u)
                                          : reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(vecp) + h * kHeadDim)[chunk];
        unsigned char* tile = is_k ? smem + map.qk + st * kQKBytes + k_off(t) : smem + map.v + st * kVBytes + v_off(t);
        *reinterpret_cast<uint4*>(tile + rowx * 64 + ((chunk ^ ((rowx >> 1) & 3)) << 4)) = val;
      }
      if (t14 == kFillF - 1) {
        PadSets<kPerm> pads;
        pads.init(p);
        uint64_t* dst = reinterpret_cast<uint64_t*>(smem + map.pads);
        dst[0] = pads.y[0]; dst[1] = pads.y[1]; dst[2] = pads.x[0]; dst[3] = pads.x[1];
      }
      fence_proxy_async();
      named_barrier(9, kFillF);
    }
  }
  pdl_launch_dependents();   // the next PDL kernel may begin its own prologue; its pdl_wait still waits for this grid
  const uint32_t tmem_base = warp != 0 && warp != 3 ? *reinterpret_cast<volatile uint32_t*>(smem + map.tmem_ptr) : 0u;

  if (warp == 0 || warp == 3) {
    // ===== TMA producers: warp 0 -> tile 0 of every pair, warp 3 -> tile 1 =====================
    const int t = warp == 0 ? 0 : 1;
    WinIter it;
    it.init(p, w0 + t);
    for (int n = 0; n < n_pairs; ++n, it.advance(p, 2)) {
      const int s = n % kQKStages, sv = n % kVStages;
      const uint32_t st = sbase + map.qk + s * kQKBytes, bar = bar_qk_full + 8 * s;
      const uint32_t stv = sbase + map.v + sv * kVBytes, barv = bar_v_full + 8 * sv;
      const bool valid = it.w < w1;
      mbar_wait(bar_qk_empty + 8 * s, ((n / kQKStages) & 1) ^ 1);
      if (elect_one()) {
        if (t == 0) TC_TRACE(p, n, 0);
        if (valid) {
          mbar_expect_tx(bar, 2 * kBoxBytes);
          move_window<kShifted, false>(maps.q, st + q_off(t), bar, p, it.b, it.wy, it.wx, h);
          move_window<kShifted, false>(maps.k, st + k_off(t), bar, p, it.b, it.wy, it.wx, h);
        } else {
          mbar_arrive(bar);
        }
      }
      __syncwarp();
      mbar_wait(bar_v_empty + 8 * sv, ((n / kVStages) & 1) ^ 1);
      if (elect_one()) {
        if (valid) {
          mbar_expect_tx(barv, kBoxBytes);
          move_window<kShifted, false>(maps.v, stv + v_off(t), barv, p, it.b, it.wy, it.wx, h);
        } else {
          mbar_arrive(barv);
        }
        if (t == 0) TC_TRACE(p, n, 1);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ===== S = Q K^T issuer (2 instructions per pair) ===========================================
    constexpr uint32_t idesc_s = instr_desc(128, 112, 0);
    // The tiles of pair n + 1 are waited for right after pair n has been issued, so that nothing but the slot
    // hand-over (bar_r) sits between a free accumulator and the S MMA.
    auto prepare = [&](int n) {
      const int s = n % kQKStages;
      mbar_wait(bar_qk_full + 8 * s, (n / kQKStages) & 1);
    };
    if (n_pairs > 0) prepare(0);
    for (int n = 0; n < n_pairs; ++n) {
      const int s = n % kQKStages, a = n % kSlotsF;
      const uint32_t st = sbase + map.qk + s * kQKBytes;
      mbar_wait(bar_r + 8 * a, ((n / kSlotsF) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d = tmem_base + a * kSlotColsF;
      const uint64_t qd = smem_desc(st + q_off(0), 16, 512, 4);   // 128 rows: q0 | q1
      const uint64_t kd = smem_desc(st + k_off(0), 16, 512, 4);   // 112 rows: k0 (56) | k1 (56)
      if (elect_one()) {
        TC_TRACE(p, n, 3);
#pragma unroll
        for (int k = 0; k < 2; ++k) umma_ss(d, qd + 2 * k, kd + 2 * k, idesc_s, k);
        umma_commit(bar_s + 8 * a);
        umma_commit(bar_qk_empty + 8 * s);
        TC_TRACE(p, n, 4);
      }
      __syncwarp();
      if (n + 1 < n_pairs) prepare(n + 1);
    }
  } else if (warp == 2) {
    // ===== O = P V issuer (4 instructions per pair) =============================================
    constexpr uint32_t idesc_o = instr_desc(128, 96, 1);
    // As in the S issuer.  (An earlier form wrote the V rows of pad tokens here, per pair: 330-800 cycles between
    // "P ready" and the MMA on the critical path of every pair -- the PV issuer set the pace of the whole CTA.)
    auto prepare = [&](int m) {
      const int s = m % kVStages;
      if (elect_one()) TC_TRACE2(p, m, 8);
      mbar_wait(bar_v_full + 8 * s, (m / kVStages) & 1);
      if (elect_one()) TC_TRACE2(p, m, 9);
    };
    if (n_pairs > 0) prepare(0);
    for (int m = 0; m < n_pairs; ++m) {
      const int s = m % kVStages, a = m % kSlotsF;
      const uint32_t st = sbase + map.v + s * kVBytes;
      if (elect_one()) TC_TRACE2(p, m, 10);
      mbar_wait(bar_p + 8 * a, (m / kSlotsF) & 1);
      tc_fence_after();
      const uint32_t d = tmem_base + a * kSlotColsF;               // O overwrites the S columns
      const uint32_t pa = d + kPColF;                              // A = P in tensor memory: 128 lanes x 64 bf16 (32 columns)
      const uint64_t vd = smem_desc(st + v_off(0), kTileBytes, 512, 4);   // N = 96: v0 | v1 | ones, one tile apart
      if (elect_one()) {
        TC_TRACE(p, m, 5);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ts(d, pa + 8 * k, vd + 64 * k, idesc_o, k);   // K = 16 per step: 8 columns of A
        TC_TRACE2(p, m, 11);
        umma_commit(bar_o + 8 * a);
        umma_commit(bar_v_empty + 8 * s);
        TC_TRACE(p, m, 6);
      }
      __syncwarp();
      if (m + 1 < n_pairs) prepare(m + 1);
    }
  } else {
    // ===== softmax / epilogue warpgroups =======================================
    const int g = (warp >> 2) - 1;          // slot owned by this warpgroup
    const int q = warp & 3;                 // TMEM sub-partition: lanes 32q .. 32q+31
    const int sel = q >> 1;                 // which tile of the pair (rows 0-63 / 64-127 of the accumulator)
    const int row = (q & 1) * 32 + lane;    // row of this tile's S / O held by this thread
    const int rowc = row < N ? row : N - 1;
    const int ri = row_i<kPerm>(rowc), rj = row_j<kPerm>(rowc);
    // this thread's row of the expanded bias matrix (tile order, * log2e): 13 x 128-bit loads per pair
    const float4* brow = reinterpret_cast<const float4*>(smem + map.bias + rowc * kBiasStride * 4);
    const float c1s = p.scale * kLog2e;
    const float2 c1 = make_float2(c1s, c1s);
    const uint64_t all = (1ull << N) - 1;
    // SW-MSA regions (newcrf_layers.py:335-345) as tile-row bit sets: window rows / columns >= ws - shift
    uint64_t hb = 0, wb = 0;
    if (kShifted) {
#pragma unroll
      for (int i = WS - kShift; i < WS; ++i) hb |= rows_of_window_row<kPerm>(i);
#pragma unroll
      for (int j = WS - kShift; j < WS; ++j) wb |= rows_of_window_col<kPerm>(j);
    }
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + g * kSlotColsF;
    const uint32_t s_addr = taddr + sel * 56, o_addr = taddr + sel * 32;   // this tile's columns of S and of O
    const uint32_t p_addr = taddr + kPColF;                                // this row's 32 columns of P
    // staging tile of this thread's output row; it couples only the two warps of a tile (64-thread named barriers)
    const uint32_t o_tile = sbase + map.stage + (2 * g + sel) * kTileBytes;
    const uint32_t o_row = o_tile + row * 64;
    const int tile_bar = 1 + 2 * g + sel;   // named barriers 1 .. 8: (slot, tile); 9 is the setup barrier
    const bool issuer = (q & 1) == 0;       // warps 0 and 2 of the warpgroup store tile 0 / tile 1
    WinIter it;
    it.init(p, w0 + 2 * g + sel);

    int m = 0;
    for (int n = g; n < n_pairs; n += kSlotsF, ++m, it.advance(p, 2 * kSlotsF)) {
      const bool valid = it.w < w1;
      // tile rows (= columns) of this window's pad tokens; they occur in the last two window rows / columns only
      uint64_t padm = 0;
      if (has_pad) {
        const uint64_t* ps = reinterpret_cast<const uint64_t*>(smem + map.pads);
        if (it.wy >= p.nWy - 2) padm |= ps[it.wy - (p.nWy - 2)];
        if (it.wx >= p.nWx - 2) padm |= ps[2 + it.wx - (p.nWx - 2)];
      }
      uint64_t masked = 0;
      if (kShifted) {
        const uint64_t hm = it.wy == p.nWy - 1 ? hb : 0, wm = it.wx == p.nWx - 1 ? wb : 0;
        masked = (((hm >> rowc) & 1) ? (~hm & all) : hm) | (((wm >> rowc) & 1) ? (~wm & all) : wm);
      }
      // every row of a tile with a mask has masked columns, so this is warp-uniform
      const bool special = kShifted && (it.wy == p.nWy - 1 || it.wx == p.nWx - 1);

      const bool tr = q == 0 && lane == 0;   // timeline of tile 0, row 0
      if (tr) TC_TRACE(p, n, 7);
      mbar_wait(bar_s + 8 * g, m & 1);
      tc_fence_after();
      if (tr) TC_TRACE(p, n, 8);
      uint32_t r[N + 15];
      TMEM_LD_X16(s_addr, (r + 0));
      TMEM_LD_X16(s_addr + 16, (r + 16));
      TMEM_LD_X16(s_addr + 32, (r + 32));
      TMEM_LD_X2(s_addr + 48, r[48], r[49]);   // column 49 belongs to the zero rows and is not used
      tmem_ld_wait();
      if (tr) TC_TRACE(p, n, 9);
      if (padm != 0) {   // warp-uniform.  K rows of pad tokens are zero (TMA fill): their logit is q . k_pad = S column 49
#pragma unroll
        for (int j = 0; j < N; ++j)
          if ((padm >> j) & 1) r[j] = r[N];
      }

      // pass 1: t = S * scale*log2e + bias*log2e (+ mask), exact row max.  Packed fp32x2 FMAs, four independent max
      // chains (a serial one is 25 deep).  (A single-pass form that exponentiates against an upper bound of the logits,
      // c1 max S + max bias, was measured in round 2: no faster, and it costs accuracy -- the dominant probability of a
      // row is then no longer exactly 1.0 in bf16: profiles/r2_softmax_bound_vs_exact.md.)
      float mx;
      {
        constexpr float kMasked = -100.0f * kLog2e;   // the SW-MSA mask value of the reference (-100, not -inf)
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int c4 = 0; c4 < 12; ++c4) {       // columns 4 c4 .. 4 c4 + 3
          const float4 b4 = brow[c4];
          float2 t0 = __ffma2_rn(make_float2(__uint_as_float(r[4 * c4]), __uint_as_float(r[4 * c4 + 1])), c1, make_float2(b4.x, b4.y));
          float2 t1 = __ffma2_rn(make_float2(__uint_as_float(r[4 * c4 + 2]), __uint_as_float(r[4 * c4 + 3])), c1, make_float2(b4.z, b4.w));
          if (special) {   // warp-uniform: tiles of the last window row / column of a shifted block
            if ((masked >> (4 * c4)) & 1) t0.x += kMasked;
            if ((masked >> (4 * c4 + 1)) & 1) t0.y += kMasked;
            if ((masked >> (4 * c4 + 2)) & 1) t1.x += kMasked;
            if ((masked >> (4 * c4 + 3)) & 1) t1.y += kMasked;
          }
          r[4 * c4] = __float_as_uint(t0.x);
          r[4 * c4 + 1] = __float_as_uint(t0.y);
          r[4 * c4 + 2] = __float_as_uint(t1.x);
          r[4 * c4 + 3] = __float_as_uint(t1.y);
          m4[c4 & 3] = fmaxf(m4[c4 & 3], fmaxf(fmaxf(t0.x, t0.y), fmaxf(t1.x, t1.y)));
        }
        {
          float t = fmaf(__uint_as_float(r[N - 1]), c1s, brow[12].x);
          if (special && ((masked >> (N - 1)) & 1)) t += kMasked;
          r[N - 1] = __float_as_uint(t);
          m4[0] = fmaxf(m4[0], t);
        }
        mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      }
      if (tr) TC_TRACE2(p, n, 0);
      // pass 2: p = exp2(t - max), rounded to bf16 for the PV MMA; the row sum comes out of that MMA through the
      // ones column, i.e. it is the sum of exactly the rounded values in P.
      const float2 nmx = make_float2(-mx, -mx);
      uint32_t w[32];   // P row: 64 bf16 (columns >= 49 zero), two per register = the 32 TMEM columns of this lane
#pragma unroll
      for (int e = 0; e < 24; ++e) {
        const float2 x = __fadd2_rn(make_float2(__uint_as_float(r[2 * e]), __uint_as_float(r[2 * e + 1])), nmx);
        w[e] = pack_bf16(fast_exp2(x.x), fast_exp2(x.y));
      }
      w[24] = pack_bf16(fast_exp2(__uint_as_float(r[N - 1]) - mx), 0.f);
#pragma unroll
      for (int e = 25; e < 32; ++e) w[e] = 0u;
      // Attention dropout (nn.Dropout on the softmax output, newcrf_layers.py:137 / swin_transformer.py:139; training
      // only): the probabilities are dropped in registers between the two MMAs with the counter-hash mask the CUDA-core
      // kernels and both backward kernels regenerate (swin_common.cuh), keyed by (window, head, query token, key token).
      // The ones column would then sum the DROPPED row, so the normaliser -- the sum of the un-dropped bf16 values -- is
      // taken here, and 1 / (1 - p) goes into the output scale.
      float lsum = 0.f, keep_scale = 1.f;
      if (kDrop) {
        const Dropout drop = make_dropout(p, it.w, h, N);
        keep_scale = drop.keep_scale;
        const uint64_t rbase = drop.base + (uint64_t)(ri * WS + rj) * N;
        float l4[2] = {0.f, 0.f};
#pragma unroll
        for (int e = 0; e < 25; ++e) {
          const float p0 = __uint_as_float(w[e] << 16), p1 = __uint_as_float(w[e] & 0xffff0000u);
          l4[0] += p0;
          l4[1] += p1;
          const int c0 = row_i<kPerm>(2 * e) * WS + row_j<kPerm>(2 * e);
          if (dropout_hash(drop.seed, rbase + c0) < drop.threshold) w[e] &= 0xffff0000u;
          if (2 * e + 1 < N) {
            const int c1k = row_i<kPerm>(2 * e + 1) * WS + row_j<kPerm>(2 * e + 1);
            if (dropout_hash(drop.seed, rbase + c1k) < drop.threshold) w[e] &= 0x0000ffffu;
          }
        }
        lsum = l4[0] + l4[1];
      }
      if (padm != 0) {   // V rows of pad tokens are zero: their share of O is (their summed probability) x v_pad = P cols 49 / 50 x V rows 49 / 50
        float pp = 0.f;
#pragma unroll
        for (int j = 0; j < N; ++j)
          if ((padm >> j) & 1) pp += __uint_as_float((j & 1) ? (w[j >> 1] & 0xffff0000u) : (w[j >> 1] << 16));
        const uint32_t hi = pack_bf16(0.f, pp) & 0xffff0000u;                         // column 49
        w[24] = (w[24] & 0x0000ffffu) | hi;
        w[25] = pack_bf16(pp - __uint_as_float(hi), 0.f);                            // column 50: what the first half lost
      }
      if (tr) TC_TRACE2(p, n, 1);
      // P row -> this lane's columns [96, 128) of the slot: the A operand of the PV MMA, read from tensor memory.  The
      // S columns it overlaps ([96, 112)) are dead for this lane: its S row is in registers (tmem_ld_wait above), and
      // every thread touches its own lane only.
      TMEM_ST_X16(p_addr, (w + 0));
      TMEM_ST_X16(p_addr + 16, (w + 16));
      tmem_st_wait();
      if (tr) TC_TRACE(p, n, 10);
      tc_fence_before();
      mbar_arrive(bar_p + 8 * g);
      if (tr) TC_TRACE(p, n, 11);
      // while the PV MMA runs: the TMA store of this tile's previous pair must have finished reading the staging tile
      if (issuer && elect_one()) bulk_wait_read();
      named_barrier(tile_bar, 64);

      mbar_wait(bar_o + 8 * g, m & 1);
      tc_fence_after();
      if (tr) TC_TRACE(p, n, 12);
      uint32_t o[32];
      TMEM_LD_X16(o_addr, (o + 0));
      TMEM_LD_X16(o_addr + 16, (o + 16));
      float l = __uint_as_float(tmem_ld_x1(taddr + 64));   // row sum: P x ones column
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(bar_r + 8 * g);
      if (tr) TC_TRACE(p, n, 13);

      // O row / rowsum -> bf16 -> 64B-swizzled staging tile -> TMA tensor store with the boxes of the loads: window
      // reverse, inverse shift and the final crop (newcrf_layers.py:239-249) are tensor-map addressing; out-of-bounds
      // elements (padding tokens) are clipped by the TMA unit.
      if (kDrop) l = lsum;                   // the ones column summed the dropped row
      const float inv1 = kDrop ? keep_scale / l : 1.0f / l;
      const float2 inv = make_float2(inv1, inv1);
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
      if (tr) TC_TRACE2(p, n, 4);
      fence_proxy_async();
      named_barrier(tile_bar, 64);
      if (tr) TC_TRACE2(p, n, 5);
      if (issuer && valid && elect_one()) {
        move_window<kShifted, true>(maps.o, o_tile, 0, p, it.b, it.wy, it.wx, h);
        bulk_commit();
      }
      if (p.lse != nullptr && valid && row < N) {
        int y = it.wy * WS + ri + p.shift, x = it.wx * WS + rj + p.shift;
        if (y >= p.Hp) y -= p.Hp;
        if (x >= p.Wp) x -= p.Wp;
        p.lse[((int64_t)(it.b * p.Hp + y) * p.Wp + x) * p.nH + h] = (mx + log2f(l)) * 0.6931471805599453f;
      }
      __syncwarp();
      if (tr) TC_TRACE2(p, n, 6);
    }
    if (issuer && elect_one()) bulk_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
  if (tid == 0) TC_TRACE(p, 0, 15);   // kernel exit
  if (tid == 0) TC_TRACE_CTA(p, 1);
}

}  // namespace tcf

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

}  // namespace

// Tensor map over an image-order (B, H, W, pixel_stride) bf16 tensor whose base pointer already includes
// the channel offset of q / k / v / o; box = (32 channels, bx pixels, by rows, 1 image), 64B swizzle.
// Loads: out-of-bounds elements (the padding the reference adds with F.pad) are zero-filled.  Stores:
// out-of-bounds elements are not written (the reference's final crop).
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

// Tensor map over a row-major (rows, cols) bf16 matrix with `pitch_bytes` between rows; box = (box_cols, box_rows),
// 128B swizzle for box_cols = 64, 64B swizzle for box_cols = 32.  Out-of-bounds rows load as zeros and are not stored.
bool tc::encode_matrix_map(CUtensorMap* out, const void* base, int64_t rows, int64_t cols, int64_t pitch_bytes, int box_cols, int box_rows) {
  if (encode_fn() == nullptr) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)pitch_bytes};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  return encode_fn()(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, box_cols == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Tensor map over (batch, rows, width) bf16 rows `row_stride` elements apart (batches rows * row_stride apart); box =
// (box_cols, box_rows, 1), 128B swizzle for 64-column boxes.  Rows beyond `rows` load as zeros.
bool tc::encode_rows_map(CUtensorMap* out, const void* base, int64_t batch, int64_t rows, int64_t width, int64_t row_stride, int box_cols,
                         int box_rows) {
  if (encode_fn() == nullptr) return false;
  const cuuint64_t dims[4] = {(cuuint64_t)width, (cuuint64_t)rows, (cuuint64_t)batch, 1u};
  const cuuint64_t strides[3] = {(cuuint64_t)row_stride * 2, (cuuint64_t)row_stride * 2 * rows, (cuuint64_t)row_stride * 2 * rows * batch};
  const cuuint32_t box[4] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1u, 1u};
  const cuuint32_t estr[4] = {1u, 1u, 1u, 1u};
  return encode_fn()(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, box_cols == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// jobs = heads x window chunks, at most one wave of CTAs; chunks hold an even number of windows
void tc::job_grid(const Problem& p, int sms, int* n_win, int* win_per_job, int* grid) {
  *n_win = p.B * p.nWy * p.nWx;
  int chunks = sms / p.nH;
  if (chunks < 1) chunks = 1;
  int per = (*n_win + chunks - 1) / chunks;
  per += per & 1;
  chunks = (*n_win + per - 1) / per;
  *win_per_job = per;
  *grid = chunks * p.nH;
}

// Programmatic dependent launch: the prologue of a kernel launched with the attribute (everything in front of its
// griddepcontrol.wait) overlaps the tail of the previous kernel in the stream, also inside captured CUDA graphs.
bool tc::pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("SWIN_PDL");
    return e == nullptr || e[0] != '0';
  }();
  return on;
}

int tc::sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return sms;
}

bool tc_forward_supported(const Problem& p, int dtype) {
  if (dtype != SWIN_ATTN_BF16 || p.ws != 7 || p.mask != nullptr) return false;
  if (p.shift != 0 && p.shift != tc::kShift) return false;
  // TMA: 16-byte aligned base pointers and strides
  if ((p.q_ps | p.k_ps | p.v_ps | p.o_ps) & 7) return false;
  auto mis = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) != 0; };
  if (mis(p.q) || mis(p.k) || mis(p.v) || mis(p.o) || mis(p.k_pad) || mis(p.v_pad)) return false;
  return tc::tma_available();
}

cudaError_t launch_tc_forward(const Problem& p, cudaStream_t st, const char** err) {
  tcf::TmaMapsF maps;
  memset(&maps, 0, sizeof(maps));
  const bool shifted = p.shift != 0;
  for (int i = shifted ? 1 : 0; i < (shifted ? 7 : 1); ++i) {
    const int bx = tc::kBoxX[i], by = tc::kBoxY[i];
    if (!tc::encode_window_map(&maps.q[i], p.q, p, p.q_ps, bx, by) || !tc::encode_window_map(&maps.k[i], p.k, p, p.k_ps, bx, by) ||
        !tc::encode_window_map(&maps.v[i], p.v, p, p.v_ps, bx, by) || !tc::encode_window_map(&maps.o[i], p.o, p, p.o_ps, bx, by)) {
      if (err) *err = "cuTensorMapEncodeTiled failed";
      return cudaErrorInvalidValue;
    }
  }
  const tcf::SmemMapF map = tcf::smem_map_f();
  const bool drop = p.dropout_p > 0.0f;
  auto kernel = shifted ? (drop ? tcf::tc_fwd_kernel<true, true> : tcf::tc_fwd_kernel<true, false>)
                        : (drop ? tcf::tc_fwd_kernel<false, true> : tcf::tc_fwd_kernel<false, false>);
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)map.total);
  if (e != cudaSuccess) {
    if (err) *err = "cudaFuncSetAttribute(max dynamic smem) failed";
    return e;
  }
  const int has_pad = (p.H != p.Hp || p.W != p.Wp) && (p.k_pad != nullptr || p.v_pad != nullptr) ? 1 : 0;
  int n_win, per, grid;
  tc::job_grid(p, tc::sm_count(), &n_win, &per, &grid);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(tcf::kThreadsF);
  cfg.dynamicSmemBytes = map.total;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = tc::pdl_enabled() ? 1 : 0;
  {
    TimedRegion timed(st);
    e = cudaLaunchKernelEx(&cfg, kernel, p, n_win, per, has_pad, maps);
  }
  count_launch();
  return e != cudaSuccess ? e : cudaGetLastError();
}

}  // namespace swin
