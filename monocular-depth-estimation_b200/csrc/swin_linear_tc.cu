// Per-token Linear of the blocks with its activation folded into the epilogue, tcgen05 / TMEM / TMA, sm_100a.
//
//   y = act(x W^T + b)      x (M, K) bf16 tokens, W (N, K) bf16 = nn.Linear.weight, b (N) bf16, y (M, N) bf16
//   y = r + (x W^T + b)     r (M, N) bf16: the residual stream (kEpi == kEpiResidual)
//
// Replaces `Mlp.fc1` followed by `nn.GELU()` (newcrf_layers.py:9-27; the same Mlp in swin_transformer.py) for
// inference under bf16 autocast: the hidden tensor (M x 4C) is written once instead of written, read and written
// again (the per-token Linears as tcgen05 GEMMs with fused epilogues).  The residual form replaces
// `attn.proj` / `Mlp.fc2` followed by `shortcut + drop_path(x)` (swin_transformer.py:240-244, newcrf_layers.py:251-255)
// where the stream is bf16 (every stage behind the first PatchMerging and the decoder levels under autocast): the branch
// output is never written, the stream is read and written once by the GEMM.  Numerics are the
// reference's step by step: fp32 accumulation, bias added in fp32, rounded to bf16 (what the cuBLASLt bias epilogue
// stores), GELU evaluated in fp32 on that bf16 value and rounded to bf16 (what nn.GELU does on a bf16 tensor); the
// residual is added in fp32 to that bf16 value and the sum rounded to bf16 (what `bf16 + bf16` does in torch).
//
// Persistent kernel, one CTA per SM, tiles of 128 tokens x 256 outputs, K in blocks of 64 (one 128B-swizzle atom):
//   warp 0      TMA producer: x tile (128 x 64) and W tile (256 x 64) per K block into a 4-stage ring
//   warp 1      MMA issuer: 4 x tcgen05.mma M128 N256 K16 per K block, accumulators double-buffered in TMEM (2 x 256 cols)
//   warp 2      TMEM allocation
//   warp 3      bias of the tile's 256 outputs -> fp32 in shared memory (double-buffered with the accumulators)
//   warps 4-19  epilogue (the GELU is ~16 instructions per element, 2 of them MUFU: it needs all four schedulers busy):
//               warp = (TMEM lane quarter, 64-column quarter of the tile); tcgen05.ld 16 columns at a time, + bias, bf16
//               round, GELU, bf16 round, 64B-swizzled staging of 32 x 32 boxes in shared memory, TMA tensor store (rows
//               beyond M are clipped by the tensor map)
// The epilogue of tile t overlaps the MMAs of tile t + 1.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "swin_attn_tc.cuh"
#include "swin_gelu.cuh"

namespace swin {
namespace tcl {

using namespace tc;

constexpr int BM = 128, BN = 256, BK = 64;
constexpr int kABytes = BM * BK * 2;                      // 16 KB
// One CTA per tile: W tile 256 x 64 (32 KB), 4 stages of 48 KB.  CTA pair (cta_group::2, M = 256 over two SMs): each CTA
// stages its own 128 tokens and HALF of the W tile (128 x 64, 16 KB), 6 stages of 32 KB -- 2/3 of the shared-memory
// fill per MMA cycle, which is what limited the one-CTA kernel on the tensor-bound shapes.
template <int kCtas> struct Ring {
  static constexpr int kBRows = BN / kCtas;
  static constexpr int kBBytes = kBRows * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = kCtas == 1 ? 4 : 6;
  static_assert(kStages * kStageBytes == 192 * 1024, "ring size");
};
constexpr int kMaxStages = 6;
constexpr int kStagingOff = 192 * 1024;
constexpr int kStagingWarp = 32 * 64;                     // 32 rows x 32 bf16 per epilogue warp
constexpr int kEpiWarps = 16;
constexpr int kBiasOff = kStagingOff + kEpiWarps * kStagingWarp;
constexpr int kBarOff = kBiasOff + 2 * BN * 4;
constexpr int kSmemBytes = kBarOff + 256;
constexpr int kThreads = 128 + kEpiWarps * 32;
static_assert(kSmemBytes <= 232448, "shared memory budget");
constexpr uint32_t kPeerMask = 0xFEFFFFFFu;   // shared::cluster address of the same offset in the pair's leader CTA (rank 0)

struct Maps {
  CUtensorMap x, w, y;
};

__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
               "l"(map), "r"(bar), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map), "r"(src), "r"(c0), "r"(c1)
               : "memory");
}

// ---- CTA-pair (cta_group::2) forms ----
// TMA load issued by either CTA of the pair into its own shared memory; the bytes are counted on the LEADER's mbarrier
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar & kPeerMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_ss_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc),
      "r"(accumulate)
      : "memory");
}
// the mbarrier at this offset in BOTH CTAs of the pair arrives once the pair's MMAs issued so far have completed
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}
// arrive on the mbarrier at this offset in the pair's leader CTA (from either CTA)
__device__ __forceinline__ void mbar_arrive_leader(uint32_t bar) {
  // default semantics (release at CTA scope): the TMEM reads it orders are fenced by tcgen05.fence::before_thread_sync;
  // .release.cluster costs a cluster-scope memory barrier per arrival (7 % of the stall samples in ncu)
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar & kPeerMask) : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t dst_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_cta_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}

#define TMEM_LD_X32(taddr, r)                                                                                          \
  asm volatile(                                                                                                        \
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                        \
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];" \
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),  \
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),     \
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),    \
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])                                                                       \
      : "r"(taddr))

constexpr int kEpiNone = 0, kEpiGelu = 1, kEpiResidual = 2, kEpiGeluBoth = 3;   // Both: y = GELU(h) AND h (training)

template <int kEpi, int kCtas>
__global__ void __launch_bounds__(kThreads, 1)
tc_linear_kernel(const __grid_constant__ Maps maps, const __nv_bfloat16* __restrict__ bias, const __nv_bfloat16* __restrict__ residual,
                 __nv_bfloat16* __restrict__ pre_act, int M, int N, int K) {
  constexpr bool kGelu = kEpi == kEpiGelu || kEpi == kEpiGeluBoth;
  constexpr bool kPair = kCtas == 2;
  constexpr int kStages = Ring<kCtas>::kStages, kStageBytes = Ring<kCtas>::kStageBytes;
  extern __shared__ __align__(1024) unsigned char smem[];
  const uint32_t sbase = smem_u32(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // CTA pair: rank 0 (the leader) issues the MMAs of the pair's 256-token tile; both CTAs load, both run the epilogue of
  // their own 128 tokens.  The persistent loop walks pair tiles with the pair index.
  const int rank = kPair ? (int)cluster_cta_rank() : 0;
  const int worker = kPair ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int workers = kPair ? (int)(gridDim.x >> 1) : (int)gridDim.x;

  const uint32_t bar_full = sbase + kBarOff;                  // [kStages]  (pair: the leader's counts both CTAs' bytes)
  const uint32_t bar_empty = bar_full + 8 * kMaxStages;       // [kStages]
  const uint32_t bar_acc_full = bar_empty + 8 * kMaxStages;   // [2]
  const uint32_t bar_acc_empty = bar_acc_full + 16;           // [2]  (pair: the leader's, kCtas * kEpiWarps arrivals)
  const uint32_t bar_bias_full = bar_acc_empty + 16;          // [2]
  const uint32_t bar_bias_empty = bar_bias_full + 16;         // [2]
  const uint32_t tmem_slot = bar_bias_empty + 16;
  float* bias_s = reinterpret_cast<float*>(smem + kBiasOff);   // [2][BN]

  const int n_blocks = N / BN, m_blocks = (M + kCtas * BM - 1) / (kCtas * BM), k_blocks = K / BK;
  const int tiles = n_blocks * m_blocks;
  auto tile_m0 = [&](int t) { return (t / n_blocks) * (kCtas * BM) + rank * BM; };   // first token of THIS CTA's rows
  auto tile_n0 = [&](int t) { return (t % n_blocks) * BN; };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.x);
    tma_prefetch_desc(&maps.w);
    tma_prefetch_desc(&maps.y);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(bar_acc_full + 8 * a, 1);
      mbar_init(bar_acc_empty + 8 * a, kCtas * kEpiWarps);
      mbar_init(bar_bias_full + 8 * a, 1);
      mbar_init(bar_bias_empty + 8 * a, kEpiWarps);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    if constexpr (kPair) tmem_alloc_pair(tmem_slot, 512);
    else tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  if constexpr (kPair) cluster_sync_all();   // the peer's barriers are initialised before anything signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + (tmem_slot - sbase));

  if (warp == 0) {
    // ---- TMA producer ----
    int s = 0;
    uint32_t ph = 0;
    for (int t = worker; t < tiles; t += workers) {
      const int m0 = tile_m0(t), n0 = tile_n0(t);
      for (int kb = 0; kb < k_blocks; ++kb) {
        mbar_wait(bar_empty + 8 * s, ph ^ 1);
        if (elect_one()) {
          const uint32_t st = sbase + s * kStageBytes;
          if constexpr (kPair) {
            if (rank == 0) mbar_expect_tx(bar_full + 8 * s, 2 * kStageBytes);
            tma_load_2d_pair(st, &maps.x, bar_full + 8 * s, kb * BK, m0);
            tma_load_2d_pair(st + kABytes, &maps.w, bar_full + 8 * s, kb * BK, n0 + rank * (BN / 2));
          } else {
            mbar_expect_tx(bar_full + 8 * s, kStageBytes);
            tma_load_2d(st, &maps.x, bar_full + 8 * s, kb * BK, m0);
            tma_load_2d(st + kABytes, &maps.w, bar_full + 8 * s, kb * BK, n0);
          }
        }
        __syncwarp();
        if (++s == kStages) {
          s = 0;
          ph ^= 1;
        }
      }
    }
  } else if (warp == 1 && rank == 0) {
    // ---- MMA issuer (pair: the leader only; M = 256 spans both CTAs' tokens, each CTA supplies half of the W tile) ----
    constexpr uint32_t idesc = instr_desc(kCtas * BM, BN, 0);
    int s = 0, a = 0;
    uint32_t ph = 0, aph = 0;
    for (int t = worker; t < tiles; t += workers) {
      mbar_wait(bar_acc_empty + 8 * a, aph ^ 1);
      tc_fence_after();
      for (int kb = 0; kb < k_blocks; ++kb) {
        mbar_wait(bar_full + 8 * s, ph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t st = sbase + s * kStageBytes;
          const uint64_t ad = smem_desc(st, 16, 1024, 2);
          const uint64_t bd = smem_desc(st + kABytes, 16, 1024, 2);
          if constexpr (kPair) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) umma_ss_pair(tmem_base + a * BN, ad + 2 * k, bd + 2 * k, idesc, (kb | k) != 0);
            umma_commit_pair(bar_empty + 8 * s);
            if (kb == k_blocks - 1) umma_commit_pair(bar_acc_full + 8 * a);
          } else {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) umma_ss(tmem_base + a * BN, ad + 2 * k, bd + 2 * k, idesc, (kb | k) != 0);
            umma_commit(bar_empty + 8 * s);
            if (kb == k_blocks - 1) umma_commit(bar_acc_full + 8 * a);
          }
        }
        __syncwarp();
        if (++s == kStages) {
          s = 0;
          ph ^= 1;
        }
      }
      a ^= 1;
      if (a == 0) aph ^= 1;
    }
  } else if (warp == 3) {
    // ---- bias loader: keeps the epilogue warps free of CTA-wide barriers ----
    int a = 0;
    uint32_t aph = 0;
    for (int t = worker; t < tiles; t += workers) {
      const int n0 = tile_n0(t);
      mbar_wait(bar_bias_empty + 8 * a, aph ^ 1);   // the epilogue of the tile that used this buffer is done with it
      const uint4 raw = __ldg(reinterpret_cast<const uint4*>(bias + n0) + lane);
      const uint32_t* rw = reinterpret_cast<const uint32_t*>(&raw);
      float4* dst = reinterpret_cast<float4*>(bias_s + a * BN + lane * 8);
      const float2 f0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rw[0]));
      const float2 f1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rw[1]));
      const float2 f2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rw[2]));
      const float2 f3 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rw[3]));
      dst[0] = make_float4(f0.x, f0.y, f1.x, f1.y);
      dst[1] = make_float4(f2.x, f2.y, f3.x, f3.y);
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_bias_full + 8 * a);
      a ^= 1;
      if (a == 0) aph ^= 1;
    }
  } else if (warp >= 4) {
    // ---- epilogue ----
    const int e = warp - 4, q = warp & 3, quarter = e >> 2;   // TMEM lane quarter of this warp, 64-column quarter of the tile
    const uint32_t staging = sbase + kStagingOff + e * kStagingWarp;
    const uint32_t row_off = lane * 64, sw = (lane >> 1) & 3;
    int a = 0;
    uint32_t aph = 0;
    for (int t = worker; t < tiles; t += workers) {
      const int m0 = tile_m0(t), n0 = tile_n0(t);
      mbar_wait(bar_bias_full + 8 * a, aph);
      mbar_wait(bar_acc_full + 8 * a, aph);
      tc_fence_after();
      const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + a * BN + quarter * 64;
      const bool rows_live = m0 + q * 32 < M;
#pragma unroll 1
      for (int box = 0; box < 2; ++box) {
        if (lane == 0) bulk_wait_read();   // the previous store of this warp has read the staging buffer
        __syncwarp();
        // residual: this lane's row, the box's 32 columns (64 contiguous bytes), issued ahead of the TMEM load
        uint4 res[4];
        if constexpr (kEpi == kEpiResidual) {
          const int row = m0 + q * 32 + lane;
          if (row < M) {
            const uint4* rp = reinterpret_cast<const uint4*>(residual + (int64_t)row * N + n0 + quarter * 64 + box * 32);
#pragma unroll
            for (int i = 0; i < 4; ++i) res[i] = __ldg(rp + i);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) res[i] = make_uint4(0u, 0u, 0u, 0u);
          }
        }
#pragma unroll
        for (int sub = 0; sub < 2; ++sub) {
          uint32_t r[16];
          TMEM_LD_X16(tacc + box * 32 + sub * 16, r);
          tmem_ld_wait();
          const float4* b4 = reinterpret_cast<const float4*>(bias_s + a * BN + quarter * 64 + box * 32 + sub * 16);
          uint32_t w[8];
          [[maybe_unused]] uint32_t hw[8];   // kEpiGeluBoth: the bf16 Linear output h, kept for the GELU backward
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 b = b4[i];
            const float2 v0 = __fadd2_rn(make_float2(__uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1])), make_float2(b.x, b.y));
            const float2 v1 = __fadd2_rn(make_float2(__uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3])), make_float2(b.z, b.w));
            uint32_t h0 = pack_bf16(v0.x, v0.y), h1 = pack_bf16(v1.x, v1.y);
            if constexpr (kEpi == kEpiGeluBoth) {
              hw[2 * i] = h0;
              hw[2 * i + 1] = h1;
            }
            if (kGelu) {
              const float2 g0 = gelu_value2<false>(__bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&h0)), make_float2(0.f, 0.f));
              const float2 g1 = gelu_value2<false>(__bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&h1)), make_float2(0.f, 0.f));
              h0 = pack_bf16(g0.x, g0.y);
              h1 = pack_bf16(g1.x, g1.y);
            }
            if constexpr (kEpi == kEpiResidual) {
              // columns sub * 16 + 4 i .. + 3 of the box = 32-bit words 2 i, 2 i + 1 of res[2 sub + (i >> 1)]
              const uint32_t* rw = reinterpret_cast<const uint32_t*>(&res[2 * sub + (i >> 1)]) + 2 * (i & 1);
              const float2 a0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&h0));
              const float2 a1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&h1));
              const float2 r0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rw[0]));
              const float2 r1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rw[1]));
              h0 = pack_bf16(r0.x + a0.x, r0.y + a0.y);
              h1 = pack_bf16(r1.x + a1.x, r1.y + a1.y);
            }
            w[2 * i] = h0;
            w[2 * i + 1] = h1;
          }
          if constexpr (kEpi == kEpiGeluBoth) {
            // h goes straight from registers to global memory: this lane's row, 16 columns = one full 32-byte sector
            const int row = m0 + q * 32 + lane;
            if (row < M) {
              uint4* hp = reinterpret_cast<uint4*>(pre_act + (int64_t)row * N + n0 + quarter * 64 + box * 32 + sub * 16);
              hp[0] = make_uint4(hw[0], hw[1], hw[2], hw[3]);
              hp[1] = make_uint4(hw[4], hw[5], hw[6], hw[7]);
            }
          }
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const uint32_t chunk = (uint32_t)(sub * 2 + j) ^ sw;
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(staging + row_off + chunk * 16), "r"(w[4 * j]), "r"(w[4 * j + 1]),
                         "r"(w[4 * j + 2]), "r"(w[4 * j + 3])
                         : "memory");
          }
        }
        if (box == 1) {
          // the accumulator buffer is in registers / shared memory: hand it back to the MMA issuer
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(bar_bias_empty + 8 * a);
            if constexpr (kPair) mbar_arrive_leader(bar_acc_empty + 8 * a);
            else mbar_arrive(bar_acc_empty + 8 * a);
          }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0 && rows_live) {
          tma_store_2d(&maps.y, staging, n0 + quarter * 64 + box * 32, m0 + q * 32);
          bulk_commit();
        }
      }
      a ^= 1;
      if (a == 0) aph ^= 1;
    }
    if (lane == 0) bulk_wait_all();
  }

  tc_fence_before();
  if constexpr (kPair) cluster_sync_all();   // neither CTA leaves (or frees TMEM) while the pair's MMAs / arrivals are in flight
  else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if constexpr (kPair) tmem_dealloc_pair(tmem_base, 512);
    else tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace tcl

bool tc_linear_supported(int64_t rows, int in_features, int out_features) {
  return rows > 0 && rows < (1ll << 31) - tcl::BM && in_features >= tcl::BK && in_features % tcl::BK == 0 && out_features >= tcl::BN &&
         out_features % tcl::BN == 0 && tc::tma_available();
}

// -1: automatic (CTA pairs on the tensor-bound shapes), 0: one CTA per tile, 1: CTA pairs.  Debug / microbench hook.
static int g_pair_mode = -1;
void set_tc_linear_pair_mode(int mode) { g_pair_mode = mode; }

template <int kEpi, int kCtas>
static cudaError_t launch_tc_linear_as(const tcl::Maps& maps, const void* bias, const void* residual, void* pre_act, int64_t rows,
                                       int in_features, int out_features, cudaStream_t st, const char** err) {
  auto kernel = tcl::tc_linear_kernel<kEpi, kCtas>;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tcl::kSmemBytes);
  if (e != cudaSuccess) {
    if (err) *err = "cudaFuncSetAttribute(max dynamic smem) failed";
    return e;
  }
  const int64_t tiles = ((rows + kCtas * tcl::BM - 1) / (kCtas * tcl::BM)) * (out_features / tcl::BN);
  const int workers = tc::sm_count() / kCtas;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(kCtas * (tiles < workers ? tiles : workers)));
  cfg.blockDim = dim3(tcl::kThreads);
  cfg.dynamicSmemBytes = tcl::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCtas;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = kCtas > 1 ? 1 : 0;
  {
    TimedRegion timed(st);
    e = cudaLaunchKernelEx(&cfg, kernel, maps, static_cast<const __nv_bfloat16*>(bias), static_cast<const __nv_bfloat16*>(residual),
                           static_cast<__nv_bfloat16*>(pre_act), (int)rows, out_features, in_features);
  }
  if (e != cudaSuccess) return e;
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_tc_linear(const void* x, const void* weight, const void* bias, const void* residual, void* pre_act, void* y, int64_t rows,
                             int in_features, int out_features, int activation, cudaStream_t st, const char** err) {
  // CTA pairs (cta_group::2) halve the W fill per CTA; they win from K = 192 up (profiles/r1_micro_linear_pair.jsonl)
  // unless the coarser 256-token tiles cost an extra round of the persistent loop (3 344 x 1024 x 4096: 224 pair tiles on
  // 74 pairs = 4 rounds against 3 rounds of 432 tiles on 148 CTAs).
  bool pair = g_pair_mode > 0;
  if (g_pair_mode < 0) {
    const int sms = tc::sm_count();
    const int64_t nb = out_features / tcl::BN;
    const int64_t tiles1 = ((rows + tcl::BM - 1) / tcl::BM) * nb, tiles2 = ((rows + 2 * tcl::BM - 1) / (2 * tcl::BM)) * nb;
    const int64_t rounds1 = (tiles1 + sms - 1) / sms, rounds2 = (tiles2 + sms / 2 - 1) / (sms / 2);
    pair = in_features >= 192 && sms >= 2 && rounds2 <= rounds1;
  }
  tcl::Maps maps;
  memset(&maps, 0, sizeof(maps));
  if (!tc::encode_matrix_map(&maps.x, x, rows, in_features, (int64_t)in_features * 2, tcl::BK, tcl::BM) ||
      !tc::encode_matrix_map(&maps.w, weight, out_features, in_features, (int64_t)in_features * 2, tcl::BK, pair ? tcl::BN / 2 : tcl::BN) ||
      !tc::encode_matrix_map(&maps.y, y, rows, out_features, (int64_t)out_features * 2, 32, 32)) {
    if (err) *err = "cuTensorMapEncodeTiled failed";
    return cudaErrorInvalidValue;
  }
#define SWIN_TCL_LAUNCH(EPI)                                                                                                   \
  (pair ? launch_tc_linear_as<EPI, 2>(maps, bias, residual, pre_act, rows, in_features, out_features, st, err)                 \
        : launch_tc_linear_as<EPI, 1>(maps, bias, residual, pre_act, rows, in_features, out_features, st, err))
  if (residual != nullptr) return SWIN_TCL_LAUNCH(tcl::kEpiResidual);
  if (pre_act != nullptr) return SWIN_TCL_LAUNCH(tcl::kEpiGeluBoth);
  if (activation) return SWIN_TCL_LAUNCH(tcl::kEpiGelu);
  return SWIN_TCL_LAUNCH(tcl::kEpiNone);
#undef SWIN_TCL_LAUNCH
}

}  // namespace swin
