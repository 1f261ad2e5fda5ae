// Shared pieces of the tcgen05 shifted-window attention kernels for sm_100a (forward: swin_attn_tc.cu,
// backward: swin_attn_tc_bwd.cu): PTX wrappers (mbarrier, TMA tensor loads / stores, tcgen05 MMA / TMEM
// load, descriptors), the window -> TMA box mapping that folds window partition, cyclic shift, padding,
// window reverse and crop into tensor-map addressing, and the token order of a tile.
#pragma once
#include <cuda.h>

#include <cstring>

#include "swin_common.cuh"

namespace swin {
namespace tc {

constexpr int WS = 7;
constexpr int N = 49;
constexpr int kShift = 3;           // the only non-zero shift served here (ws // 2, as the reference uses)
constexpr int kTileBytes = 64 * 64;       // 64 rows x 32 bf16
constexpr int kKTile1 = 56 * 64;          // K tile 1 follows the 56 MMA rows of K tile 0
constexpr int kPTileBytes = 64 * 128;     // 64 rows x 64 bf16
constexpr int kBoxBytes = N * 64;         // bytes one tile of one tensor receives from TMA
constexpr float kLog2e = 1.4426950408889634f;

// Tensor maps of one launch, per tensor: [0] = whole window (7x7); [1..4] = the four parts of a shifted
// window that wraps in y, (x extent, y extent) = (4,4) (4,3) (3,4) (3,3), stored at rows 0/16/28/40;
// [5..6] = the two parts (4,7) (3,7) of a shifted window that does not wrap in y, stored at rows 0/28
// (same token order).  kBoxX / kBoxY below hold the extents.

// ---- PTX wrappers -----------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity), "r"(0x989680u)   // suspend-time hint: sleep in hardware instead of spinning
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}

// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread.
__device__ __forceinline__ void umma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc),
      "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]: the A operand (M rows = TMEM lanes, K elements packed two bf16 per 32-bit column,
// K-major) is read from tensor memory, so a P matrix produced in registers never touches shared memory.
__device__ __forceinline__ void umma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc),
      "r"(accumulate)
      : "memory");
}
// mbarrier arrives once every previously issued tcgen05.mma of this thread has completed.
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout): start address >> 4 in
// [0,14), leading byte offset >> 4 in [16,30), stride byte offset >> 4 in [32,46), version = 1 in
// [46,48), swizzle mode in [61,64) (2 = 128B, 4 = 64B).
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout) {
  return (uint64_t)((addr & 0x3FFFF) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
         (1ull << 46) | ((uint64_t)layout << 61);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): D = F32 (1 << 4), A = B = BF16 (1 << 7, 1 << 10),
// a_major bit 15 / b_major bit 16 (1 = MN-major), N >> 3 in [17,23), M >> 4 in [24,29).
__host__ __device__ constexpr uint32_t instr_desc(int m, int n, int b_mn_major, int a_mn_major = 0) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

#define TMEM_LD_X16(taddr, r)                                                                                      \
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];" \
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),  \
                 "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]) \
               : "r"(taddr))
__device__ __forceinline__ uint32_t tmem_ld_x1(uint32_t taddr) {
  uint32_t r;
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r) : "r"(taddr));
  return r;
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// registers -> 16 consecutive 32-bit columns of this thread's TMEM lane (lane = 32 * (warp % 4) + lane id)
#define TMEM_ST_X16(taddr, r)                                                                                      \
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" \
               ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), \
                 "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])     \
               : "memory")
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
// exp2 of two bf16 at once (one MUFU op for two probabilities; P is consumed as bf16 by the PV MMA)
__device__ __forceinline__ uint32_t ex2_bf16x2(uint32_t x) {
  uint32_t y;
  asm("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(y) : "r"(x));
  return y;
}
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// One elected lane of a fully converged warp.  Issuing TMA / tcgen05 instructions from `if (elect_one())`
// inside warp-uniform control flow lets the compiler keep their operands in uniform registers; issuing them
// from `if (lane == 0)` costs a R2UR broadcast loop (~100 cycles) per instruction.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}

// ---- TMA ------------------------------------------------------------------------
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// box of tensor map `map` at (channel, x, y, image) -> shared memory, completion on `bar`
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
// shared memory box -> tensor map `map` at (channel, x, y, image); out-of-bounds elements are not written
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(map), "r"(src),
               "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all bulk groups of this thread have finished READING their shared-memory source
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void named_barrier(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
#define TMEM_LD_X2(taddr, r0, r1) \
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(taddr))

// Debug timelines (tools/tc_trace.py; see swin_attn_debug_set_trace).  Compiled in only with -DSWIN_TC_TRACE
// (csrc/build.sh with SWIN_TC_TRACE=1): even predicated off, the stamps cost ~35 issue slots per softmax warp and pair.
#ifdef SWIN_TC_TRACE
// event e of local pair n (< 256) of CTA 0; the buffer holds 8192 int64
#define TC_TRACE(p, n, e)                                                        \
  do {                                                                           \
    if ((p).trace != nullptr && blockIdx.x == 0) (p).trace[(n) * 16 + (e)] = clock64(); \
  } while (0)

// every CTA: globaltimer (ns) at entry (which = 0) / exit (1) in trace[4096 + 2 * blockIdx.x + which]
#define TC_TRACE_CTA(p, which)                                                   \
  do {                                                                           \
    if ((p).trace != nullptr) {                                                  \
      unsigned long long t_;                                                     \
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_));                      \
      (p).trace[4096 + 2 * blockIdx.x + (which)] = (long long)t_;                \
    }                                                                            \
  } while (0)
// second table (softmax sub-phases): event e of local pair n at trace[8192 + n * 16 + e]
#define TC_TRACE2(p, n, e)                                                       \
  do {                                                                           \
    if ((p).trace != nullptr && blockIdx.x == 0) (p).trace[8192 + (n) * 16 + (e)] = clock64(); \
  } while (0)
#else
#define TC_TRACE(p, n, e) do { } while (0)
#define TC_TRACE_CTA(p, which) do { } while (0)
#define TC_TRACE2(p, n, e) do { } while (0)
#endif

// Programmatic dependent launch (PDL).  `pdl_wait` blocks until every grid this launch depends on has completed and
// flushed its memory: everything that reads the previous kernel's output (or writes global memory) comes after it;
// the prologue in front of it (mbarrier init, TMEM allocation, shared-memory fills, parameter staging, tensor-map
// prefetch) overlaps the previous kernel's tail.  `pdl_launch_dependents` lets the next PDL-launched kernel start
// its own prologue while this grid is still running.  Both are no-ops when the launch carries no PDL attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---- tiles and token order --------------------------------------------------------
// Row r of a tile <-> token (i, j) of the window.  Natural order for un-shifted blocks;
// for shifted blocks the four TMA parts follow one another:
//   rows [0,16): i<4, j<4   [16,28): i>=4, j<4   [28,40): i<4, j>=4   [40,49): i>=4, j>=4
template <bool kPerm>
__host__ __device__ constexpr int row_i(int r) {
  if (!kPerm) return r / WS;
  return r < 16 ? r / 4 : r < 28 ? 4 + (r - 16) / 4 : r < 40 ? (r - 28) / 3 : 4 + (r - 40) / 3;
}
template <bool kPerm>
__host__ __device__ constexpr int row_j(int r) {
  if (!kPerm) return r % WS;
  return r < 16 ? r % 4 : r < 28 ? (r - 16) % 4 : r < 40 ? 4 + (r - 28) % 3 : 4 + (r - 40) % 3;
}
template <bool kPerm>
__host__ __device__ constexpr int token_row(int i, int j) {
  if (!kPerm) return i * WS + j;
  return j < 4 ? (i < 4 ? i * 4 + j : 16 + (i - 4) * 4 + j) : (i < 4 ? 28 + i * 3 + (j - 4) : 40 + (i - 4) * 3 + (j - 4));
}

// Window w of the job -> (image, window row, window column); advance by a few windows without divisions.
struct WinIter {
  int b, wy, wx, w;
  __device__ __forceinline__ void init(const Problem& p, int first) {
    w = first;
    wx = first % p.nWx;
    const int r = first / p.nWx;
    wy = r % p.nWy;
    b = r / p.nWy;
  }
  __device__ __forceinline__ void advance(const Problem& p, int step) {
    w += step;
    wx += step;
    while (wx >= p.nWx) { wx -= p.nWx; ++wy; }
    while (wy >= p.nWy) { wy -= p.nWy; ++b; }
  }
};

// One window of one tensor, global -> shared (kStore = false) or shared -> global (kStore = true):
// 1 box (un-shifted) or the 2 / 4 parts of a shifted window (see swin_attn_tc.cu::load_window).
template <bool kShifted, bool kStore>
__device__ __forceinline__ void move_window(const CUtensorMap* maps, uint32_t smem_addr, uint32_t bar, const Problem& p,
                                            int b, int wy, int wx, int h) {
  const int ch = h * kHeadDim;
  auto go = [&](int map, uint32_t row, int x, int y) {
    if (kStore) tma_store_4d(maps + map, smem_addr + row * 64, ch, x, y, b);
    else tma_load_4d(smem_addr + row * 64, maps + map, bar, ch, x, y, b);
  };
  if (!kShifted) {
    go(0, 0, wx * WS, wy * WS);
  } else {
    const int xa = wx * WS + kShift, ya = wy * WS + kShift;
    const int xb = wx == p.nWx - 1 ? 0 : wx * WS + WS, yb = wy == p.nWy - 1 ? 0 : wy * WS + WS;
    if (wy == p.nWy - 1) {
      go(1, 0, xa, ya);
      go(2, 16, xa, yb);
      go(3, 28, xb, ya);
      go(4, 40, xb, yb);
    } else {
      go(5, 0, xa, ya);
      go(6, 28, xb, ya);
    }
  }
}

template <bool kPerm>
__device__ __forceinline__ uint64_t rows_of_window_row(int i) {
  uint64_t m = 0;
#pragma unroll
  for (int j = 0; j < WS; ++j) m |= 1ull << token_row<kPerm>(i, j);
  return m;
}
template <bool kPerm>
__device__ __forceinline__ uint64_t rows_of_window_col(int j) {
  uint64_t m = 0;
#pragma unroll
  for (int i = 0; i < WS; ++i) m |= 1ull << token_row<kPerm>(i, j);
  return m;
}

// Padding tokens (y >= H or x >= W) can only occur in the last two window rows / columns: their tile-row bit sets,
// for window row nWy-2 ([0]) and nWy-1 ([1]), likewise for columns.
template <bool kPerm>
struct PadSets {
  uint64_t y[2], x[2];
  __device__ __forceinline__ void init(const Problem& p) {
    y[0] = y[1] = x[0] = x[1] = 0;
#pragma unroll
    for (int e = 0; e < 2; ++e) {
#pragma unroll
      for (int i = 0; i < WS; ++i) {
        int yy = (p.nWy - 2 + e) * WS + i + p.shift, xx = (p.nWx - 2 + e) * WS + i + p.shift;
        if (yy >= p.Hp) yy -= p.Hp;
        if (xx >= p.Wp) xx -= p.Wp;
        if (yy >= p.H && p.nWy - 2 + e >= 0) y[e] |= rows_of_window_row<kPerm>(i);
        if (xx >= p.W && p.nWx - 2 + e >= 0) x[e] |= rows_of_window_col<kPerm>(i);
      }
    }
  }
  __device__ __forceinline__ uint64_t of(const Problem& p, int wy, int wx) const {
    return (wy == p.nWy - 1 ? y[1] : wy == p.nWy - 2 ? y[0] : 0) | (wx == p.nWx - 1 ? x[1] : wx == p.nWx - 2 ? x[0] : 0);
  }
};

// Pad tokens carry the projection bias (the reference pads before the Linear), but TMA zero-fills them: after a
// K / V tile has landed, the issuing warp overwrites the rows of its pad tokens with the head's pad vector, so the
// S and PV MMAs see exactly the reference's operands and the softmax needs no pad handling at all.  `vec` holds the
// 64 bytes of the vector (4 x 16 B); rows are 64 B with the 64B swizzle of the TMA box.
__device__ __forceinline__ void write_pad_rows(uint32_t tile, uint64_t padded, const uint4 (&vec)[4], int lane) {
#pragma unroll
  for (int r0 = 0; r0 < 64; r0 += 32) {
    const int r = r0 + lane;
    if (r < N && ((padded >> r) & 1)) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const uint32_t dst = tile + r * 64 + ((c ^ ((r >> 1) & 3)) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(vec[c].x), "r"(vec[c].y), "r"(vec[c].z), "r"(vec[c].w)
                     : "memory");
      }
    }
  }
}

#define TMEM_LD_X1(taddr, r0) asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r0) : "r"(taddr))

// Host: tensor map over an image-order (B, H, W, pix_stride) bf16 tensor, box (32 ch, bx, by, 1), 64B swizzle
// (swin_attn_tc.cu).  The seven box shapes of a launch, indexed like TmaMaps.
bool encode_window_map(CUtensorMap* out, const void* base, const Problem& p, int64_t pix_stride, int bx, int by);
bool tma_available();
bool encode_matrix_map(CUtensorMap* out, const void* base, int64_t rows, int64_t cols, int64_t pitch_bytes, int box_cols, int box_rows);
bool encode_rows_map(CUtensorMap* out, const void* base, int64_t batch, int64_t rows, int64_t width, int64_t row_stride, int box_cols,
                     int box_rows);
void job_grid(const Problem& p, int sms, int* n_win, int* win_per_job, int* grid);
int sm_count();
bool pdl_enabled();   // programmatic dependent launch on the tcgen05 / LayerNorm launches (env SWIN_PDL=0 turns it off)
constexpr int kBoxX[7] = {WS, WS - kShift, WS - kShift, kShift, kShift, WS - kShift, kShift};
constexpr int kBoxY[7] = {WS, WS - kShift, kShift, WS - kShift, kShift, WS, WS};

}  // namespace tc
}  // namespace swin
