// tcgen05 backward of the shifted-window attention (bf16, ws 7, shift 0 / 3) for sm_100a.
//
// Recomputes the probabilities from (q, k, lse) instead of storing them (the reference's autograd
// keeps S-sized tensors per block).  Per (window, head) tile, N = 49, d = 32:
//
//   S  = Q K^T          dP = dO V^T                      (tensor pipe, phase 1)
//   P  = exp2(S c + bias - lse)    delta_i = sum_j P_ij dP_ij   (= rowsum(dO o O); O is never read)
//   dS = P o (dP - delta)                                 (registers, one thread per query row)
//   dV = P^T dO    dK = scale dS^T Q    dQ = scale dS K   (tensor pipe, phase 2)
//   dTable[rel_index(i, j), h] += dS_ij                   (tensor pipe: T += dS I, see below)
//
// HBM traffic per tile: read q, k, v, dO (4 x 49 x 64 B) + lse, write dq, dk, dv (3 x 49 x 64 B).
//
// Work decomposition: one CTA = one JOB = (head h, contiguous range of windows).  Consecutive CTAs
// are consecutive heads over the SAME windows, so the 128-byte lines two heads share are fetched
// at about the same time.  Two windows of the job form a PAIR stacked along M (as in the forward):
// tile t owns TMEM lanes 64t .. 64t+63.
//
//   phase 1 (M128 N112 K32, twice):  S  = [Q0;Q1]  [K0;K1]^T     dP = [dO0;dO1] [V0;V1]^T
//   phase 2 (M128 N64  K64, 4 times): every product pairs an A operand built from the bf16 P / dS
//       rows the softmax threads wrote ONCE (row i, 128 B, 128B swizzle) with a B operand that is one
//       of the TMA-landed tiles read MN-major:
//         dV = [P0^T ;P1^T ] [dO0|dO1]    A = P  image read MN-major (M = key j, K = query i)
//         dK = [dS0^T;dS1^T] [Q0 |Q1 ]    A = dS image read MN-major
//         dQ = [dS0  ;dS1  ] [K0 |K1 ]    A = dS image read K-major  (M = query i, K = key j)
//         T += [dS0  ;dS1  ] I            A = dS image read K-major, B = 64x64 identity
//       T (128 lanes x 64 columns of TMEM) accumulates sum_windows dS for the job's head over the whole
//       CTA lifetime; at the end the rows go through shared memory once, one thread per relative position sums
//       its entries, and the 169 sums are added to d_bias_table with 169 global atomics per CTA.
//       The off-diagonal blocks of every stacked product are computed and ignored.
//   epilogue: dQ/dK/dV rows TMEM -> registers -> bf16 -> 64B-swizzled staging tiles (they reuse the
//       P / dS bytes of the slot) -> TMA tensor stores with the same boxes as the loads, so the
//       window reverse and the inverse cyclic shift are tensor-map addressing as well.
//
// Warps: 0 / 3 TMA producers (tile 0 / tile 1), 1 phase-1 MMA issuer, 2 phase-2 MMA issuer,
// 4-11 two softmax / gradient warpgroups (one accumulator slot each).  The producers also issue the TMA stores.
#include <utility>

#include "swin_attn_tc.cuh"

namespace swin {
namespace tcb {

using namespace tc;

// timeline of CTA 0 (tools/tc_trace.py --bwd; trace build only): event e of local pair n at trace[12288 + n * 16 + e]
#ifdef SWIN_TC_TRACE
#define TCB_TRACE(p, n, e)                                                                          \
  do {                                                                                              \
    if ((p).trace != nullptr && blockIdx.x == 0 && (n) < 240) (p).trace[12288 + (n) * 16 + (e)] = clock64(); \
  } while (0)
#else
#define TCB_TRACE(p, n, e) do { } while (0)
#endif

constexpr int kStagesB = 4;
constexpr int kSlotsB = 2;
constexpr int kThreadsB = 128 + 128 * kSlotsB;
constexpr int kStageBytesB = 8 * kTileBytes;   // q0 q1 | k0 k1 | v0 v1 | do0 do1
constexpr int kSlotBytes = 4 * kPTileBytes;    // P0 P1 | dS0 dS1 ; reused as dq0 dq1 dk0 dk1 dv0 dv1 staging
constexpr int kSlotColsB = 224;                // S 0..111, dP 112..223 ; then dQ 0..63, dK 64..127, dV 128..191
constexpr int kDtCol = 448;                    // T accumulator: columns 448..511
constexpr int kLoadBytes = 4 * kBoxBytes;      // per tile: q, k, v, dO

__device__ __forceinline__ uint32_t bq_off(int t) { return t * kTileBytes; }
__device__ __forceinline__ uint32_t bk_off(int t) { return 2 * kTileBytes + t * kKTile1; }
__device__ __forceinline__ uint32_t bv_off(int t) { return 4 * kTileBytes + t * kKTile1; }
__device__ __forceinline__ uint32_t bdo_off(int t) { return 6 * kTileBytes + t * kTileBytes; }

struct TmaMapsB {
  CUtensorMap q[7], k[7], v[7], d_o[7];   // loads
  CUtensorMap dq[7], dk[7], dv[7];        // stores
};

struct SmemMapB {
  uint32_t stages, slots, ident, bias, padacc, bars, tmem_ptr, total;
};
__host__ __device__ inline SmemMapB smem_map_b() {
  SmemMapB m;
  m.stages = 0;
  m.slots = m.stages + kStagesB * kStageBytesB;
  m.ident = m.slots + kSlotsB * kSlotBytes;
  m.bias = m.ident + kPTileBytes;
  m.padacc = (m.bias + 176 * 4 + 15) & ~15u;   // [slot][tile][dK | dV][32] fp32: sums over padding tokens (un-padded mode)
  m.bars = m.padacc + kSlotsB * 2 * 64 * 4;
  m.tmem_ptr = m.bars + 8 * (2 * kStagesB + 6 * kSlotsB);
  m.total = m.tmem_ptr + 16;
  return m;
}

template <bool kShifted, bool kDrop>
__global__ void __launch_bounds__(kThreadsB, 1)
tc_bwd_kernel(const Problem p, int n_win, int win_per_job, int has_pad, const __grid_constant__ TmaMapsB maps) {
  constexpr bool kPerm = kShifted;
  extern __shared__ __align__(1024) unsigned char smem[];
  const SmemMapB map = smem_map_b();
  const uint32_t sbase = smem_u32(smem);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int h = (int)blockIdx.x % p.nH;
  const int w0 = ((int)blockIdx.x / p.nH) * win_per_job;
  const int w1 = min(w0 + win_per_job, n_win);
  const int n_pairs = (w1 - w0 + 1) >> 1;

  const uint32_t bar_full = sbase + map.bars;              // [kStagesB] TMA -> phase-1 MMA
  const uint32_t bar_empty = bar_full + 8 * kStagesB;      // [kStagesB] phase-2 MMA -> producers
  const uint32_t bar_s = bar_empty + 8 * kStagesB;         // [kSlotsB]  S, dP ready
  const uint32_t bar_p = bar_s + 8 * kSlotsB;              // [kSlotsB]  P, dS written
  const uint32_t bar_o = bar_p + 8 * kSlotsB;              // [kSlotsB]  dQ, dK, dV ready
  const uint32_t bar_r = bar_o + 8 * kSlotsB;              // [kSlotsB]  outputs read, slot reusable
  const uint32_t bar_g = bar_r + 8 * kSlotsB;              // [kSlotsB]  dQ, dK, dV staged (softmax -> producers)
  const uint32_t bar_f = bar_g + 8 * kSlotsB;              // [kSlotsB]  staging tiles read by the TMA stores

  if (tid == 0) TCB_TRACE(p, 239, 0);                             // kernel entry
  // ---- one-time setup -------------------------------------------------------
  // As in the forward: mbarrier init, one CTA barrier, then the producers start loading while the other 10
  // warps fill what TMA never writes, stage the bias, build the identity and allocate TMEM.
  if (tid == 0) {
    for (int i = 0; i < kStagesB; ++i) {
      mbar_init(bar_full + 8 * i, 2);
      mbar_init(bar_empty + 8 * i, 1);
    }
    for (int i = 0; i < kSlotsB; ++i) {
      mbar_init(bar_s + 8 * i, 1);
      mbar_init(bar_p + 8 * i, 128);
      mbar_init(bar_o + 8 * i, 1);
      mbar_init(bar_r + 8 * i, 128);
      mbar_init(bar_g + 8 * i, 128);
      mbar_init(bar_f + 8 * i, 2);
    }
    fence_barrier_init();
  }
  __syncthreads();
  if (tid == 0) TCB_TRACE(p, 239, 1);                             // barriers initialised
  // Un-padded mode (H, W not multiples of the window): TMA zero-fills the padding tokens, which carry the projection
  // bias in the reference; their K / V rows are overwritten with k_pad / v_pad once the tiles have landed (phase-1
  // issuer).  The vectors are read here, ahead of the fills, so that the (cold) loads overlap the set-up.
  const bool fix = has_pad && (p.k_pad != nullptr || p.v_pad != nullptr);
  uint4 kvec[4] = {}, vvec[4] = {};
  if (warp == 1 && fix) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      if (p.k_pad != nullptr) kvec[c] = __ldg(reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p.k_pad) + h * kHeadDim) + c);
      if (p.v_pad != nullptr) vvec[c] = __ldg(reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(p.v_pad) + h * kHeadDim) + c);
    }
  }
  if (warp != 0 && warp != 3) {
    constexpr int kFill = kThreadsB - 64;
    const int t10 = tid - (warp < 3 ? 32 : 64);   // 0 .. 319 over the 10 non-producer warps
    const float bias_v = t10 < 169 ? p.table[t10 * p.nH + h] * kLog2e : 0.f;
    if (warp == 1) tmem_alloc(sbase + map.tmem_ptr, 512);
    if (tid == 32) TCB_TRACE(p, 239, 2);                          // TMEM allocated
    // Rows 49-63 of every tile (the TMA boxes write rows 0-48) must stay finite: as the K dimension of a
    // phase-2 product they are multiplied by zero columns of P / dS.  K / V tile 1 start at row 56 of tile 0.
    for (int i = t10; i < kStagesB * 8 * 15 * 4; i += kFill) {
      const int chunk = i & 3, rr = (i >> 2) % 15, t8 = ((i >> 2) / 15) & 7, s = (i >> 2) / 120;
      const int kind = t8 >> 1, t = t8 & 1;       // 0 q, 1 k, 2 v, 3 dO
      if ((kind == 1 || kind == 2) && t == 0 && rr >= 7) continue;
      const uint32_t base = kind == 0 ? bq_off(t) : kind == 1 ? bk_off(t) : kind == 2 ? bv_off(t) : bdo_off(t);
      *reinterpret_cast<uint4*>(smem + map.stages + s * kStageBytesB + base + (49 + rr) * 64 + (chunk << 4)) = make_uint4(0u, 0u, 0u, 0u);
    }
    // identity, K-major 128B-swizzled 64 x 64 bf16: element (n, n) = 1
    for (int i = t10; i < 64 * 8; i += kFill) {
      const int n = i >> 3, chunk = i & 7;
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      if (chunk == (n >> 3)) {
        const uint32_t one = (n & 1) ? 0x3f800000u : 0x00003f80u;
        const int wsel = (n & 7) >> 1;
        v.x = wsel == 0 ? one : 0u; v.y = wsel == 1 ? one : 0u; v.z = wsel == 2 ? one : 0u; v.w = wsel == 3 ? one : 0u;
      }
      *reinterpret_cast<uint4*>(smem + map.ident + n * 128 + ((chunk ^ (n & 7)) << 4)) = v;
    }
    if (t10 < 176) {
      reinterpret_cast<float*>(smem + map.bias)[t10] = bias_v;
    }
    if (t10 < kSlotsB * 2 * 64) reinterpret_cast<float*>(smem + map.padacc)[t10] = 0.f;
    fence_proxy_async();
    tc_fence_before();
    if (tid == 32) TCB_TRACE(p, 239, 3);                          // fills done
    named_barrier(5, kFill);
    tc_fence_after();
    if (tid == 32) TCB_TRACE(p, 239, 4);                          // set-up barrier passed
  } else if (lane == 0) {
    for (int i = kShifted ? 1 : 0; i < (kShifted ? 7 : 1); ++i) {
      tma_prefetch_desc(&maps.q[i]); tma_prefetch_desc(&maps.k[i]); tma_prefetch_desc(&maps.v[i]);
      tma_prefetch_desc(&maps.d_o[i]);
    }
  }
  const uint32_t tmem_base = warp != 0 && warp != 3 ? *reinterpret_cast<volatile uint32_t*>(smem + map.tmem_ptr) : 0u;

  if (warp == 0 || warp == 3) {
    // ===== TMA producers: warp 0 -> tile 0 of every pair, warp 3 -> tile 1 =====================
    const int t = warp == 0 ? 0 : 1;
    // They also issue the TMA STORES of their tile: the store of pair e is issued in iteration e + kStagesB, right
    // after the loads that refill the stage pair e's phase 2 has just released -- its epilogue finishes a few hundred
    // cycles later, so the wait on bar_g is short and the softmax warpgroups never issue or wait for a store.
    WinIter it, its;
    it.init(p, w0 + t);
    its = it;
    for (int n = 0; n < n_pairs + kStagesB; ++n) {
      if (n < n_pairs) {
        const int s = n % kStagesB;
        const uint32_t st = sbase + map.stages + s * kStageBytesB, bar = bar_full + 8 * s;
        mbar_wait(bar_empty + 8 * s, ((n / kStagesB) & 1) ^ 1);
        if (it.w < w1) {
          if (elect_one()) {
            if (t == 0) TCB_TRACE(p, n, 0);                       // producer: stage free, loads issued
            mbar_expect_tx(bar, kLoadBytes);
            move_window<kShifted, false>(maps.q, st + bq_off(t), bar, p, it.b, it.wy, it.wx, h);
            move_window<kShifted, false>(maps.k, st + bk_off(t), bar, p, it.b, it.wy, it.wx, h);
            move_window<kShifted, false>(maps.d_o, st + bdo_off(t), bar, p, it.b, it.wy, it.wx, h);
            move_window<kShifted, false>(maps.v, st + bv_off(t), bar, p, it.b, it.wy, it.wx, h);
          }
        } else {
          // Half-empty last pair of the job.  Rows 0-7 of K tile 1 are rows 56-63 of the K dimension of the dQ
          // product of tile 0 (multiplied by zero columns of dS): they must be finite even if never loaded.
          *reinterpret_cast<uint4*>(smem + map.stages + s * kStageBytesB + bk_off(t) + lane * 16) = make_uint4(0u, 0u, 0u, 0u);
          fence_proxy_async();
          __syncwarp();
          if (elect_one()) mbar_arrive(bar);
        }
        __syncwarp();
        it.advance(p, 2);
      }
      if (n >= kStagesB) {
        const int e = n - kStagesB, a = e % kSlotsB;
        mbar_wait(bar_g + 8 * a, (e / kSlotsB) & 1);            // dQ / dK / dV of pair e are in the staging tiles
        if (elect_one()) {
          if (its.w < w1) {
            const uint32_t src = sbase + map.slots + a * kSlotBytes + t * kTileBytes;
            move_window<kShifted, true>(maps.dq, src, 0, p, its.b, its.wy, its.wx, h);
            move_window<kShifted, true>(maps.dk, src + 2 * kTileBytes, 0, p, its.b, its.wy, its.wx, h);
            move_window<kShifted, true>(maps.dv, src + 4 * kTileBytes, 0, p, its.b, its.wy, its.wx, h);
            bulk_commit();
          }
          bulk_wait_read();                  // the stores have read the staging tiles:
          mbar_arrive(bar_f + 8 * a);        // P / dS of pair e + 2 may be written over them
          if (t == 0) TCB_TRACE(p, e, 14);                        // stores issued and read
        }
        __syncwarp();
        its.advance(p, 2);
      }
    }
    if (elect_one()) bulk_wait_all();
  } else if (warp == 1) {
    // ===== phase 1: S = Q K^T, dP = dO V^T ======================================================
    constexpr uint32_t idesc = instr_desc(128, 112, 0);
    PadSets<kPerm> pads;
    WinIter it;
    it.init(p, w0);
    if (fix) pads.init(p);
    for (int n = 0; n < n_pairs; ++n) {
      const int s = n % kStagesB, a = n % kSlotsB;
      const uint32_t st = sbase + map.stages + s * kStageBytesB;
      if (elect_one()) TCB_TRACE(p, n, 1);                      // phase 1: loop top
      mbar_wait(bar_full + 8 * s, (n / kStagesB) & 1);
      if (elect_one()) TCB_TRACE(p, n, 2);                      // phase 1: tiles landed
      if (fix) {
        const uint64_t pad0 = pads.of(p, it.wy, it.wx);
        it.advance(p, 1);
        const uint64_t pad1 = it.w < w1 ? pads.of(p, it.wy, it.wx) : 0;
        it.advance(p, 1);
        if (pad0 | pad1) {
          write_pad_rows(st + bk_off(0), pad0, kvec, lane);
          write_pad_rows(st + bk_off(1), pad1, kvec, lane);
          write_pad_rows(st + bv_off(0), pad0, vvec, lane);
          write_pad_rows(st + bv_off(1), pad1, vvec, lane);
          fence_proxy_async();
        }
        __syncwarp();
      }
      if (elect_one()) TCB_TRACE(p, n, 3);                      // phase 1: pads fixed, waiting for the slot
      mbar_wait(bar_r + 8 * a, ((n / kSlotsB) & 1) ^ 1);
      tc_fence_after();
      const uint32_t d = tmem_base + a * kSlotColsB;
      const uint64_t qd = smem_desc(st + bq_off(0), 16, 512, 4), kd = smem_desc(st + bk_off(0), 16, 512, 4);
      const uint64_t gd = smem_desc(st + bdo_off(0), 16, 512, 4), vd = smem_desc(st + bv_off(0), 16, 512, 4);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 2; ++k) umma_ss(d, qd + 2 * k, kd + 2 * k, idesc, k);
#pragma unroll
        for (int k = 0; k < 2; ++k) umma_ss(d + 112, gd + 2 * k, vd + 2 * k, idesc, k);
        umma_commit(bar_s + 8 * a);
        TCB_TRACE(p, n, 4);                                     // phase 1 issued
      }
      __syncwarp();
    }
  } else if (warp == 2) {
    // ===== phase 2: dV, dK, dQ and the bias-gradient accumulator ================================
    constexpr uint32_t idesc_tt = instr_desc(128, 64, 1, 1);   // A and B MN-major
    constexpr uint32_t idesc_nt = instr_desc(128, 64, 1, 0);   // A K-major, B MN-major
    constexpr uint32_t idesc_nn = instr_desc(128, 64, 0, 0);   // both K-major
    const uint64_t id_k = smem_desc(sbase + map.ident, 16, 1024, 2);
    for (int m = 0; m < n_pairs; ++m) {
      const int s = m % kStagesB, a = m % kSlotsB;
      mbar_wait(bar_p + 8 * a, (m / kSlotsB) & 1);
      tc_fence_after();
      if (elect_one()) TCB_TRACE(p, m, 5);                      // phase 2: P / dS seen
      const uint32_t st = sbase + map.stages + s * kStageBytesB;
      const uint32_t sl = sbase + map.slots + a * kSlotBytes;
      const uint32_t d = tmem_base + a * kSlotColsB;
      const uint64_t p_mn = smem_desc(sl, 2 * kPTileBytes / 2, 1024, 2);                    // M blocks (tiles) 8 KB apart
      const uint64_t ds_mn = smem_desc(sl + 2 * kPTileBytes, 2 * kPTileBytes / 2, 1024, 2);
      const uint64_t ds_k = smem_desc(sl + 2 * kPTileBytes, 16, 1024, 2);
      const uint64_t g_mn = smem_desc(st + bdo_off(0), kTileBytes, 512, 4);                 // N blocks: dO0 | dO1
      const uint64_t q_mn = smem_desc(st + bq_off(0), kTileBytes, 512, 4);                  // Q0 | Q1
      const uint64_t k_mn = smem_desc(st + bk_off(0), kKTile1, 512, 4);                     // K0 | K1
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ss(d + 128, p_mn + 128 * k, g_mn + 64 * k, idesc_tt, k);   // dV
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ss(d + 64, ds_mn + 128 * k, q_mn + 64 * k, idesc_tt, k);   // dK
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ss(d, ds_k + 2 * k, k_mn + 64 * k, idesc_nt, k);           // dQ
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ss(tmem_base + kDtCol, ds_k + 2 * k, id_k + 2 * k, idesc_nn, (m | k) != 0);
        umma_commit(bar_o + 8 * a);
        umma_commit(bar_empty + 8 * s);
        TCB_TRACE(p, m, 6);                                     // phase 2 issued
      }
      __syncwarp();
    }
  } else {
    // ===== softmax / gradient warpgroups: warpgroup g owns accumulator slot g, i.e. pairs g, g + 2, ... =====
    const int g = (warp >> 2) - 1;
    const int q = warp & 3;                 // TMEM sub-partition: lanes 32q .. 32q+31
    const int sel = q >> 1;                 // tile of the pair
    const int row = (q & 1) * 32 + lane;    // row of this tile's S / dP / dQ / dK / dV held by this thread
    const int rowc = row < N ? row : N - 1;
    const int ri = row_i<kPerm>(rowc), rj = row_j<kPerm>(rowc);
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    const int base_i = ri * 13 + rj + 84;   // relative_position_index[row, col] = base_i - (ci*13 + cj)
    const float* brow = reinterpret_cast<const float*>(smem + map.bias) + base_i;
    const float c1s = p.scale * kLog2e;
    const float2 c1 = make_float2(c1s, c1s);
    const uint64_t all = (1ull << N) - 1;
    uint64_t hb = 0, wb = 0;                // SW-MSA regions as tile-row bit sets (newcrf_layers.py:335-345)
    if (kShifted) {
#pragma unroll
      for (int i = WS - kShift; i < WS; ++i) hb |= rows_of_window_row<kPerm>(i);
#pragma unroll
      for (int j = WS - kShift; j < WS; ++j) wb |= rows_of_window_col<kPerm>(j);
    }
    PadSets<kPerm> pads;
    if (has_pad) pads.init(p);
    // one owner thread (row 49) per slot and tile: plain read-modify-write, no atomics
    float4* padacc = reinterpret_cast<float4*>(smem + map.padacc) + (g * 2 + sel) * 16;
    const uint32_t sl = sbase + map.slots + g * kSlotBytes;
    const uint32_t p_row = sl + sel * kPTileBytes + row * 128;
    const uint32_t ds_row = p_row + 2 * kPTileBytes;
    const uint32_t stage_row = sl + sel * kTileBytes + row * 64;   // + 2 * kTileBytes per output tensor
    const uint32_t taddr = lane_base + g * kSlotColsB;
    const uint32_t s_addr = taddr + sel * 56, dp_addr = taddr + 112 + sel * 56;
    const bool tr = q == 0 && lane == 0;
    WinIter it;
    it.init(p, w0 + 2 * g + sel);
    // Per-pair state of this thread's row, fetched one pair AHEAD (in the idle window while phase 2 of the current pair
    // runs): the lse read is a global load of about 1500 cycles under load.  Rows that must produce P = dS = 0 get
    // lse = +inf: rows 49-63 of the tile, and the padding tokens as QUERY rows (un-padded mode; their output was
    // cropped, the forward's lse there is unspecified).  As KEY columns the padding tokens have their dS / P mass
    // summed into column 49, so that row 49 of the dK / dV products is the gradient of k_pad / v_pad (rows 49+ of the
    // K, Q, dO tiles are zero).
    struct RowState {
      bool valid;
      uint64_t padded;
      float lse;
    };
    auto fetch = [&](const WinIter& w) {
      RowState s;
      s.valid = w.w < w1;
      int y = w.wy * WS + ri + p.shift, x = w.wx * WS + rj + p.shift;
      if (y >= p.Hp) y -= p.Hp;
      if (x >= p.Wp) x -= p.Wp;
      s.padded = has_pad && s.valid ? pads.of(p, w.wy, w.wx) : 0;
      s.lse = INFINITY;
      if (s.valid && row < N && !((s.padded >> rowc) & 1)) s.lse = __ldg(&p.lse[((int64_t)(w.b * p.Hp + y) * p.Wp + x) * p.nH + h]);
      return s;
    };
    RowState nxt = {false, 0, INFINITY};
    if (g < n_pairs) nxt = fetch(it);

    int m = 0;
    for (int n = g; n < n_pairs; n += kSlotsB, ++m) {
      const RowState cur = nxt;
      const bool valid = cur.valid;
      const uint64_t padded = cur.padded;
      const float lse2 = cur.lse * kLog2e;
      uint64_t masked = 0;
      if (kShifted) {
        const uint64_t hm = it.wy == p.nWy - 1 ? hb : 0, wm = it.wx == p.nWx - 1 ? wb : 0;
        masked = (((hm >> rowc) & 1) ? (~hm & all) : hm) | (((wm >> rowc) & 1) ? (~wm & all) : wm);
      }
      const bool special = kShifted && (it.wy == p.nWy - 1 || it.wx == p.nWx - 1);   // warp-uniform

      if (tr) TCB_TRACE(p, n, 7);                               // softmax: waiting for S / dP
      mbar_wait(bar_s + 8 * g, m & 1);
      tc_fence_after();
      if (tr) TCB_TRACE(p, n, 8);                               // S / dP ready
      uint32_t r[N + 15], d[N + 15];
      TMEM_LD_X16(s_addr, (r + 0));
      TMEM_LD_X16(s_addr + 16, (r + 16));
      TMEM_LD_X16(s_addr + 32, (r + 32));
      TMEM_LD_X1(s_addr + 48, r[48]);
      TMEM_LD_X16(dp_addr, (d + 0));
      TMEM_LD_X16(dp_addr + 16, (d + 16));
      TMEM_LD_X16(dp_addr + 32, (d + 32));
      TMEM_LD_X1(dp_addr + 48, d[48]);
      tmem_ld_wait();
      if (tr) TCB_TRACE(p, n, 9);                               // loaded

      // P = exp2(S * scale*log2e + bias*log2e (+ mask) - lse*log2e)
      const float2 nl = make_float2(-lse2, -lse2);
      if (!special) {
#pragma unroll
        for (int j = 0; j < N - 1; j += 2) {
          const int cj0 = row_i<kPerm>(j) * 13 + row_j<kPerm>(j), cj1 = row_i<kPerm>(j + 1) * 13 + row_j<kPerm>(j + 1);
          const float2 b2 = __fadd2_rn(make_float2(brow[-cj0], brow[-cj1]), nl);
          const float2 t = __ffma2_rn(make_float2(__uint_as_float(r[j]), __uint_as_float(r[j + 1])), c1, b2);
          r[j] = __float_as_uint(fast_exp2(t.x));
          r[j + 1] = __float_as_uint(fast_exp2(t.y));
        }
        const int cj = row_i<kPerm>(N - 1) * 13 + row_j<kPerm>(N - 1);
        r[N - 1] = __float_as_uint(fast_exp2(fmaf(__uint_as_float(r[N - 1]), c1s, brow[-cj] - lse2)));
      } else {
#pragma unroll
        for (int j = 0; j < N; ++j) {
          const int cj = row_i<kPerm>(j) * 13 + row_j<kPerm>(j);
          float t = fmaf(__uint_as_float(r[j]), c1s, brow[-cj] - lse2);
          if ((masked >> j) & 1) t += -100.0f * kLog2e;
          r[j] = __float_as_uint(fast_exp2(t));
        }
      }
      // Attention dropout (training with attn_drop > 0): the forward multiplied P by keep / (1 - p) before the value
      // product, so dP is multiplied by the same factor here (the mask is regenerated from the counter hash,
      // swin_common.cuh) and dV is taken with the DROPPED probabilities; delta and dS then follow unchanged.
      uint64_t km = ~0ull;
      float ks = 1.f;
      if (kDrop) {
        const Dropout drop = make_dropout(p, it.w, h, N);
        ks = drop.keep_scale;
        const uint64_t rbase = drop.base + (uint64_t)(ri * WS + rj) * N;
        km = 0;
#pragma unroll
        for (int j = 0; j < N; ++j) {
          const bool keep = dropout_hash(drop.seed, rbase + row_i<kPerm>(j) * WS + row_j<kPerm>(j)) >= drop.threshold;
          km |= (uint64_t)keep << j;
          d[j] = keep ? __float_as_uint(__uint_as_float(d[j]) * ks) : 0u;
        }
      }
      auto kept = [&](int j) { return kDrop ? (((km >> j) & 1) ? ks : 0.f) : 1.f; };   // factor on P for the dV product
      // delta = sum_j P_ij dP_ij ; dS = P (dP - delta)
      float2 acc = make_float2(0.f, 0.f), acc_b = make_float2(0.f, 0.f);
#pragma unroll
      for (int j = 0; j < N - 1; j += 4) {
        acc = __ffma2_rn(make_float2(__uint_as_float(r[j]), __uint_as_float(r[j + 1])),
                         make_float2(__uint_as_float(d[j]), __uint_as_float(d[j + 1])), acc);
        acc_b = __ffma2_rn(make_float2(__uint_as_float(r[j + 2]), __uint_as_float(r[j + 3])),
                           make_float2(__uint_as_float(d[j + 2]), __uint_as_float(d[j + 3])), acc_b);
      }
      const float delta = fmaf(__uint_as_float(r[N - 1]), __uint_as_float(d[N - 1]), (acc.x + acc.y) + (acc_b.x + acc_b.y));
      const float2 nd = make_float2(-delta, -delta);
      uint32_t wp[25], wd[25];
#pragma unroll
      for (int e = 0; e < 24; ++e) {
        const float2 pp = make_float2(__uint_as_float(r[2 * e]), __uint_as_float(r[2 * e + 1]));
        const float2 ds = __fmul2_rn(pp, __fadd2_rn(make_float2(__uint_as_float(d[2 * e]), __uint_as_float(d[2 * e + 1])), nd));
        wp[e] = kDrop ? pack_bf16(pp.x * kept(2 * e), pp.y * kept(2 * e + 1)) : pack_bf16(pp.x, pp.y);
        wd[e] = pack_bf16(ds.x, ds.y);
      }
      {
        const float pl = __uint_as_float(r[N - 1]);
        const float dl = pl * (__uint_as_float(d[N - 1]) - delta);
        float u_pad = 0.f, s_pad = 0.f;
        if (padded != 0) {   // warp-uniform
#pragma unroll
          for (int j = 0; j < N; ++j) {
            if ((padded >> j) & 1) {
              const float pj = __uint_as_float(r[j]);
              u_pad += pj * kept(j);
              s_pad = fmaf(pj, __uint_as_float(d[j]) - delta, s_pad);
            }
          }
        }
        wp[24] = pack_bf16(pl * kept(N - 1), u_pad);
        wd[24] = pack_bf16(dl, s_pad);
      }
      if (!valid) {   // missing second tile of the job's last pair (warp-uniform): its stage bytes are not data
#pragma unroll
        for (int e = 0; e < 25; ++e) wp[e] = wd[e] = 0u;
      }
      if (tr) TCB_TRACE(p, n, 10);                              // math done
      // the staging tiles of this slot (same bytes as P / dS) must have been read by the TMA stores of pair n - 2
      mbar_wait(bar_f + 8 * g, (m & 1) ^ 1);
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        const uint32_t off = (uint32_t)((ch ^ (row & 7)) << 4);
        auto word = [&](const uint32_t* w, int i) { return i < 25 ? w[i < 25 ? i : 0] : 0u; };
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(p_row + off), "r"(word(wp, 4 * ch)), "r"(word(wp, 4 * ch + 1)),
                     "r"(word(wp, 4 * ch + 2)), "r"(word(wp, 4 * ch + 3)) : "memory");
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(ds_row + off), "r"(word(wd, 4 * ch)), "r"(word(wd, 4 * ch + 1)),
                     "r"(word(wd, 4 * ch + 2)), "r"(word(wd, 4 * ch + 3)) : "memory");
      }
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(bar_p + 8 * g);
      if (tr) TCB_TRACE(p, n, 11);                              // P / dS handed over

      // idle window (phase 2 runs): state of this warpgroup's next pair
      WinIter itn = it;
      itn.advance(p, 2 * kSlotsB);
      if (n + kSlotsB < n_pairs) nxt = fetch(itn);

      mbar_wait(bar_o + 8 * g, m & 1);
      tc_fence_after();
      if (tr) TCB_TRACE(p, n, 12);                              // dQ / dK / dV ready
      // dQ (x scale), dK (x scale), dV rows -> bf16 -> 64B-swizzled staging tiles
#pragma unroll
      for (int t3 = 0; t3 < 3; ++t3) {
        uint32_t o[32];
        const uint32_t oaddr = taddr + t3 * 64 + sel * 32;
        TMEM_LD_X16(oaddr, (o + 0));
        TMEM_LD_X16(oaddr + 16, (o + 16));
        tmem_ld_wait();
        const float sc = t3 < 2 ? p.scale : 1.0f;
        const float2 sc2 = make_float2(sc, sc);
        if (t3 > 0 && row == N && padded != 0) {   // row 49 of dK / dV = this tile's gradient of k_pad / v_pad
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            float4 a = padacc[(t3 - 1) * 8 + c];
            a.x = fmaf(__uint_as_float(o[4 * c]), sc, a.x);
            a.y = fmaf(__uint_as_float(o[4 * c + 1]), sc, a.y);
            a.z = fmaf(__uint_as_float(o[4 * c + 2]), sc, a.z);
            a.w = fmaf(__uint_as_float(o[4 * c + 3]), sc, a.w);
            padacc[(t3 - 1) * 8 + c] = a;
          }
        }
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          uint32_t v[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 s2 = __fmul2_rn(make_float2(__uint_as_float(o[8 * ch + 2 * e]), __uint_as_float(o[8 * ch + 2 * e + 1])), sc2);
            v[e] = pack_bf16(s2.x, s2.y);
          }
          const uint32_t dst = stage_row + t3 * 2 * kTileBytes + (uint32_t)((ch ^ ((row >> 1) & 3)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]) : "memory");
        }
      }
      tc_fence_before();
      mbar_arrive(bar_r + 8 * g);                               // TMEM slot free for phase 1 of pair n + 2
      if (tr) TCB_TRACE(p, n, 13);                              // outputs read, slot released
      fence_proxy_async();
      mbar_arrive(bar_g + 8 * g);                               // staged: the producers issue the TMA stores
      it = itn;
    }
    if (tid == 128) TCB_TRACE(p, 239, 5);                       // warpgroup 0: last pair staged

    // ---- bias-table gradient: T[row i, col j] = sum over the job's windows of dS -------------------
    // The phase-2 MMAs complete in issue order, so once the warpgroup that owns the LAST pair has seen its bar_o, T is
    // final and every load has been consumed (the stage buffers are free); the barrier over both warpgroups publishes
    // that, and every padacc update, to all 256 threads.
    named_barrier(3, 256);
    if (tid == 128) TCB_TRACE(p, 239, 6);                         // both warpgroups done
    if (n_pairs > 0) {
      // T rows -> shared memory in WINDOW-TOKEN order [tile][query token][key token] (lane stride 49 words), then one
      // thread per relative position (di, dj) sums the (7 - |di|)(7 - |dj|) entries of both tiles that share it:
      // element ((ci+di)*7 + cj+dj, ci*7 + cj) = 49 (7 di + dj) + 350 ci + 50 cj.  (fp32 shared atomics are CAS loops:
      // the 49-per-thread scatter of the first version took 5 us per launch.)
      float* ts = reinterpret_cast<float*>(smem + map.stages);
      if (g == 0) {
        tc_fence_after();
        uint32_t tt[N + 15];
        const uint32_t t_addr = lane_base + kDtCol;
        TMEM_LD_X16(t_addr, (tt + 0));
        TMEM_LD_X16(t_addr + 16, (tt + 16));
        TMEM_LD_X16(t_addr + 32, (tt + 32));
        TMEM_LD_X1(t_addr + 48, tt[48]);
        tmem_ld_wait();
        if (row < N) {
          float* dst = ts + (sel * N + ri * WS + rj) * N;
#pragma unroll
          for (int j = 0; j < N; ++j) dst[row_i<kPerm>(j) * WS + row_j<kPerm>(j)] = __uint_as_float(tt[j]);
        }
      }
      named_barrier(4, 256);
      const int b = tid - 128;
      if (b < 169) {
        const int di = b / 13 - 6, dj = b % 13 - 6;   // bin = (ri - ci + 6) * 13 + (rj - cj + 6)
        const float* src = ts + N * (WS * di + dj);
        float s0 = 0.f, s1 = 0.f;
#pragma unroll
        for (int ci = 0; ci < WS; ++ci) {
#pragma unroll
          for (int cj = 0; cj < WS; ++cj) {
            if ((unsigned)(ci + di) < (unsigned)WS && (unsigned)(cj + dj) < (unsigned)WS) {
              s0 += src[350 * ci + 50 * cj];
              s1 += src[N * N + 350 * ci + 50 * cj];
            }
          }
        }
        atomicAdd(&p.d_table[b * p.nH + h], s0 + s1);
      }
      if (has_pad && b < 64) {
        float* dst = b < 32 ? p.d_kpad : p.d_vpad;
        const float* acc = reinterpret_cast<const float*>(smem + map.padacc) + b;
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < kSlotsB * 2; ++i) sum += acc[i * 64];
        if (dst != nullptr) atomicAdd(&dst[h * kHeadDim + (b & 31)], sum);
      }
      if (tid == 128) TCB_TRACE(p, 239, 7);                       // bias-table / pad-vector gradients added
    }
  }

  tc_fence_before();
  __syncthreads();
  if (tid == 0) TCB_TRACE(p, 239, 8);                             // kernel exit
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

}  // namespace tcb

// ---- host side ------------------------------------------------------------------
bool tc_backward_supported(const Problem& p, int dtype) {
  if (dtype != SWIN_ATTN_BF16 || p.ws != 7 || p.mask != nullptr) return false;
  if (p.shift != 0 && p.shift != tc::kShift) return false;
  // un-padded tensors (H, W not multiples of the window) are served here and only here: TMA zero fill + pad vectors
  if ((p.q_ps | p.k_ps | p.v_ps | p.C) & 7) return false;
  auto mis = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) != 0; };
  if (mis(p.q) || mis(p.k) || mis(p.v) || mis(p.d_o) || mis(p.dq) || mis(p.dk) || mis(p.dv) || mis(p.k_pad) || mis(p.v_pad)) return false;
  return tc::tma_available();
}

cudaError_t launch_tc_backward(const Problem& p, cudaStream_t st, const char** err) {
  tcb::TmaMapsB maps;
  memset(&maps, 0, sizeof(maps));
  const bool shifted = p.shift != 0;
  for (int i = shifted ? 1 : 0; i < (shifted ? 7 : 1); ++i) {
    const int bx = tc::kBoxX[i], by = tc::kBoxY[i];
    const bool ok = tc::encode_window_map(&maps.q[i], p.q, p, p.q_ps, bx, by) && tc::encode_window_map(&maps.k[i], p.k, p, p.k_ps, bx, by) &&
                    tc::encode_window_map(&maps.v[i], p.v, p, p.v_ps, bx, by) && tc::encode_window_map(&maps.d_o[i], p.d_o, p, p.C, bx, by) &&
                    tc::encode_window_map(&maps.dq[i], p.dq, p, p.q_ps, bx, by) && tc::encode_window_map(&maps.dk[i], p.dk, p, p.k_ps, bx, by) &&
                    tc::encode_window_map(&maps.dv[i], p.dv, p, p.v_ps, bx, by);
    if (!ok) {
      if (err) *err = "cuTensorMapEncodeTiled failed";
      return cudaErrorInvalidValue;
    }
  }
  const tcb::SmemMapB map = tcb::smem_map_b();
  const bool drop = p.dropout_p > 0.0f;
  auto kernel = shifted ? (drop ? tcb::tc_bwd_kernel<true, true> : tcb::tc_bwd_kernel<true, false>)
                        : (drop ? tcb::tc_bwd_kernel<false, true> : tcb::tc_bwd_kernel<false, false>);
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)map.total);
  if (e != cudaSuccess) {
    if (err) *err = "cudaFuncSetAttribute(max dynamic smem) failed";
    return e;
  }
  int n_win, per, grid;
  tc::job_grid(p, tc::sm_count(), &n_win, &per, &grid);
  TimedRegion timed(st);   // the memsets are device work of the op: inside the timed region
  // d_bias_table, d_k_pad and d_v_pad are accumulated with atomics and start at zero.  Buffers that touch (the torch
  // binding lays them out back to back: k_pad | v_pad | table) are cleared by ONE memset instead of three launches.
  struct Span { char* lo; char* hi; } spans[3];
  int n_spans = 0;
  auto add_span = [&](float* ptr, size_t count) {
    if (ptr != nullptr) spans[n_spans++] = Span{reinterpret_cast<char*>(ptr), reinterpret_cast<char*>(ptr + count)};
  };
  add_span(p.d_table, (size_t)169 * p.nH);
  add_span(p.d_kpad, (size_t)p.C);
  add_span(p.d_vpad, (size_t)p.C);
  for (int i = 1; i < n_spans; ++i)
    for (int j = i; j > 0 && spans[j].lo < spans[j - 1].lo; --j) std::swap(spans[j], spans[j - 1]);
  for (int i = 0; i < n_spans;) {
    int j = i + 1;
    while (j < n_spans && spans[j].lo == spans[j - 1].hi) ++j;
    if ((e = cudaMemsetAsync(spans[i].lo, 0, (size_t)(spans[j - 1].hi - spans[i].lo), st)) != cudaSuccess) return e;
    i = j;
  }
  const int has_pad = (p.H != p.Hp || p.W != p.Wp) ? 1 : 0;
  kernel<<<grid, tcb::kThreadsB, map.total, st>>>(p, n_win, per, has_pad, maps);
  count_launch();
  return cudaGetLastError();
}

}  // namespace swin
