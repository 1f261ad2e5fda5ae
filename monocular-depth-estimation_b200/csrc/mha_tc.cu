// tcgen05 form of the global attention that returns its map (mha_global.cu has the CUDA-core / fp32 form and the
// description of the op): bf16, head dims 64, up to 448 tokens -- the Depthformer decoder's shapes (418 tokens at
// 352x1216, 300 at 480x640; model/Depthformer/self_attention.py:70-80).
//
// One CTA = 128 query rows of one (image, head); everything of the CTA is resident at once, nothing is pipelined
// because nothing repeats:
//   TMA:      Q block (128 x 64), K and V (S x 64, 64-row boxes, rows >= S zero-filled) -> 128B-swizzled smem
//   S MMA:    S = Q K^T as M128 x N(<= 256 + <= 192) x K64 into TMEM columns [0, 448)
//   softmax:  one thread per row, three passes over its TMEM lane: max; e = exp2((s - max) c) written back in place with
//             the row sum; p = e / sum -> (a) fp32 to a per-warp staging tile and from there to the map with coalesced
//             128-byte stores, (b) bf16 pairs back into TMEM columns [0, 224) -- in place: chunk c of P lands on columns
//             the thread has already consumed
//   PV MMA:   O = P V with P read FROM TENSOR MEMORY (TS form), V MN-major from smem, into TMEM columns [448, 512)
//   epilogue: O row -> bf16 -> global
// The map (B nH S^2 fp32) is 50x the q / k / v bytes, so the kernel is bound by writing it.
#include "swin_attn_tc.cuh"

namespace swin {
namespace mhatc {

using namespace tc;

constexpr int kHd = 64;
constexpr int kRowsM = 128;
constexpr int kMaxSeq = 448;
constexpr int kBoxRows = 64;                      // TMA box: 64 rows x 64 bf16 = 8 KB
constexpr int kBoxBytesM = kBoxRows * kHd * 2;
constexpr int kThreadsM = 192;                    // warps 0-3 softmax / epilogue (TMEM lanes 32 w ..), 4 TMA, 5 MMA
constexpr int kOCol = 448;

struct MapsM {
  CUtensorMap q, k, v;
};
struct ArgsM {
  int S, Sk, s16, nH;                             // S query rows, Sk keys, s16 = Sk rounded up to 16
  float c1;                                       // scale * log2(e)
  __nv_bfloat16* o;
  int64_t o_rs;
  float* attn;
  // optional ordered bias (see mha_global.cu): logit(i, j) += idx_bias[tok_idx[i] - tok_idx[j] + n_emb - 1, head]
  const int* tok_idx;
  const float* idx_bias;
  int n_emb;
};
constexpr int kMaxEmb = 256;                      // 2 n_emb - 1 <= 511 bias entries per head in shared memory
struct SmemM {
  uint32_t q, k, v, stage, idx, bias, bars, tmem_ptr, total;
};
__host__ __device__ inline SmemM smem_m() {
  SmemM m;
  m.q = 0;
  m.k = m.q + 2 * kBoxBytesM;
  m.v = m.k + 7 * kBoxBytesM;
  m.stage = m.v + 7 * kBoxBytesM;                 // [4 warps][32 rows][33] fp32
  m.idx = m.stage + 4 * 32 * 33 * 4;              // [kMaxSeq] token indices of the image
  m.bias = m.idx + kMaxSeq * 4;                   // [2 kMaxEmb] bias column of the head, * log2(e)
  m.bars = m.bias + 2 * kMaxEmb * 4;
  m.tmem_ptr = m.bars + 8 * 8;
  m.total = m.tmem_ptr + 16;
  return m;
}

#define TMEM_LD_X32(taddr, r)              \
  do {                                     \
    TMEM_LD_X16((taddr), ((r) + 0));       \
    TMEM_LD_X16((taddr) + 16, ((r) + 16)); \
  } while (0)

template <bool kBias>
__global__ void __launch_bounds__(kThreadsM, 1) mha_tc_kernel(const ArgsM a, const __grid_constant__ MapsM maps) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const SmemM map = smem_m();
  const uint32_t sbase = smem_u32(smem);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int row0 = blockIdx.x * kRowsM, h = blockIdx.y, b = blockIdx.z;
  const uint32_t bar_qk = sbase + map.bars, bar_v = bar_qk + 8, bar_s = bar_qk + 16, bar_p = bar_qk + 24, bar_o = bar_qk + 32;
  const int n_box = (a.s16 + kBoxRows - 1) / kBoxRows;

  if (tid == 0) {
    mbar_init(bar_qk, 1);
    mbar_init(bar_v, 1);
    mbar_init(bar_s, 1);
    mbar_init(bar_p, 128);
    mbar_init(bar_o, 1);
    fence_barrier_init();
  }
  if (warp == 5) tmem_alloc(sbase + map.tmem_ptr, 512);
  if (kBias && warp < 4) {
    int* sidx = reinterpret_cast<int*>(smem + map.idx);
    float* sbias = reinterpret_cast<float*>(smem + map.bias);
    for (int i = tid; i < kMaxSeq; i += 128) sidx[i] = i < a.Sk ? a.tok_idx[(int64_t)b * a.Sk + i] : 0;
    for (int i = tid; i < 2 * a.n_emb - 1; i += 128) sbias[i] = a.idx_bias[(int64_t)i * a.nH + h] * 1.4426950408889634f;
  }
  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&maps.q);
    tma_prefetch_desc(&maps.k);
    tma_prefetch_desc(&maps.v);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + map.tmem_ptr);

  if (warp == 4) {
    if (elect_one()) {
      mbar_expect_tx(bar_qk, (2 + n_box) * kBoxBytesM);
      for (int i = 0; i < 2; ++i) tma_load_4d(sbase + map.q + i * kBoxBytesM, &maps.q, bar_qk, h * kHd, row0 + i * kBoxRows, b, 0);
      for (int i = 0; i < n_box; ++i) tma_load_4d(sbase + map.k + i * kBoxBytesM, &maps.k, bar_qk, h * kHd, i * kBoxRows, b, 0);
      mbar_expect_tx(bar_v, n_box * kBoxBytesM);
      for (int i = 0; i < n_box; ++i) tma_load_4d(sbase + map.v + i * kBoxBytesM, &maps.v, bar_v, h * kHd, i * kBoxRows, b, 0);
    }
    __syncwarp();
  } else if (warp == 5) {
    const int n0 = a.s16 < 256 ? a.s16 : 256, n1 = a.s16 - n0;
    mbar_wait(bar_qk, 0);
    tc_fence_after();
    if (elect_one()) {
      const uint64_t qd = smem_desc(sbase + map.q, 16, 1024, 2);             // K-major, 128-byte rows, 128B swizzle
      const uint64_t kd0 = smem_desc(sbase + map.k, 16, 1024, 2);
      const uint64_t kd1 = smem_desc(sbase + map.k + 256 * kHd * 2, 16, 1024, 2);
      const uint32_t id0 = instr_desc(kRowsM, n0, 0);
#pragma unroll
      for (int k = 0; k < 4; ++k) umma_ss(tmem_base, qd + 2 * k, kd0 + 2 * k, id0, k);
      if (n1 > 0) {
        const uint32_t id1 = instr_desc(kRowsM, n1, 0);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_ss(tmem_base + 256, qd + 2 * k, kd1 + 2 * k, id1, k);
      }
      umma_commit(bar_s);
    }
    __syncwarp();
    mbar_wait(bar_v, 0);
    mbar_wait(bar_p, 0);
    tc_fence_after();
    if (elect_one()) {
      const uint64_t vd = smem_desc(sbase + map.v, 16, 1024, 2);             // MN-major B: 8 key rows x 128 B per atom
      constexpr uint32_t ido = instr_desc(kRowsM, kHd, 1);
      for (int k = 0; k < a.s16 / 16; ++k) umma_ts(tmem_base + kOCol, tmem_base + 8 * k, vd + 128 * k, ido, k);
      umma_commit(bar_o);
    }
    __syncwarp();
  } else {
    const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16);
    const int n_chunks = (a.s16 + 31) / 32;
    float* stage = reinterpret_cast<float*>(smem + map.stage) + warp * 32 * 33;
    mbar_wait(bar_s, 0);
    tc_fence_after();
    uint32_t r[32];
    // ordered bias of this thread's row: t(j) = s c1 + brow[-idx[j]] (already in the log2 domain)
    const int* sidx = reinterpret_cast<const int*>(smem + map.idx);
    const float* brow = reinterpret_cast<const float*>(smem + map.bias) +
                        (kBias ? sidx[min(row0 + warp * 32 + lane, a.S - 1)] + a.n_emb - 1 : 0);
    // pass 1: row maximum of the logits (log2 domain)
    float m = -INFINITY;
    for (int c = 0; c < n_chunks; ++c) {
      TMEM_LD_X32(taddr + 32 * c, r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (32 * c + j < a.Sk) m = fmaxf(m, kBias ? fmaf(__uint_as_float(r[j]), a.c1, brow[-sidx[32 * c + j]]) : __uint_as_float(r[j]) * a.c1);
    }
    // pass 2: e = exp2(t - max), written back in place; row sum
    const float mc = m;
    float sum = 0.f;
    for (int c = 0; c < n_chunks; ++c) {
      TMEM_LD_X32(taddr + 32 * c, r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float bj = kBias && 32 * c + j < a.Sk ? brow[-sidx[32 * c + j]] : 0.f;
        const float e = 32 * c + j < a.Sk ? fast_exp2(fmaf(__uint_as_float(r[j]), a.c1, bj - mc)) : 0.f;
        sum += e;
        r[j] = __float_as_uint(e);
      }
      TMEM_ST_X16(taddr + 32 * c, (r + 0));
      TMEM_ST_X16(taddr + 32 * c + 16, (r + 16));
    }
    tmem_st_wait();
    // pass 3: p = e / sum -> the map (through the warp's staging tile: coalesced rows) and bf16 pairs back into TMEM
    const float inv = 1.f / sum;
    float* attn_bh = a.attn != nullptr ? a.attn + ((int64_t)b * a.nH + h) * a.S * a.Sk : nullptr;
    for (int c = 0; c < n_chunks; ++c) {
      TMEM_LD_X32(taddr + 32 * c, r);
      tmem_ld_wait();
      uint32_t w[16];
#pragma unroll
      for (int j = 0; j < 32; j += 2) {
        const float p0 = __uint_as_float(r[j]) * inv, p1 = __uint_as_float(r[j + 1]) * inv;
        stage[lane * 33 + j] = p0;
        stage[lane * 33 + j + 1] = p1;
        w[j >> 1] = pack_bf16(p0, p1);
      }
      TMEM_ST_X16(taddr + 16 * c, w);             // columns [16 c, 16 c + 16): S chunks this thread has consumed
      __syncwarp();
      if (attn_bh != nullptr) {
        const int col = 32 * c + lane;
        for (int rr = 0; rr < 32; ++rr) {
          const int row = row0 + warp * 32 + rr;
          if (row < a.S && col < a.Sk) attn_bh[(int64_t)row * a.Sk + col] = stage[rr * 33 + lane];
        }
      }
      __syncwarp();
    }
    tmem_st_wait();
    tc_fence_before();
    mbar_arrive(bar_p);

    mbar_wait(bar_o, 0);
    tc_fence_after();
    const int row = row0 + warp * 32 + lane;
    uint32_t oo[64];
    TMEM_LD_X32(taddr + kOCol, oo);
    TMEM_LD_X32(taddr + kOCol + 32, (oo + 32));
    tmem_ld_wait();
    if (row < a.S) {
      uint4* dst = reinterpret_cast<uint4*>(a.o + ((int64_t)b * a.S + row) * a.o_rs + h * kHd);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        uint4 v;
        v.x = pack_bf16(__uint_as_float(oo[8 * c]), __uint_as_float(oo[8 * c + 1]));
        v.y = pack_bf16(__uint_as_float(oo[8 * c + 2]), __uint_as_float(oo[8 * c + 3]));
        v.z = pack_bf16(__uint_as_float(oo[8 * c + 4]), __uint_as_float(oo[8 * c + 5]));
        v.w = pack_bf16(__uint_as_float(oo[8 * c + 6]), __uint_as_float(oo[8 * c + 7]));
        dst[c] = v;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_dealloc(tmem_base, 512);
}

}  // namespace mhatc

bool mha_tc_supported(int dtype, int seq, int kv_seq, int dq, int dv, int64_t q_rs, int64_t k_rs, int64_t v_rs, int64_t o_rs, const void* q,
                      const void* k, const void* v, const void* o) {
  if (dtype != SWIN_ATTN_BF16 || dq != mhatc::kHd || dv != mhatc::kHd || seq < 1 || kv_seq < 16 || kv_seq > mhatc::kMaxSeq) return false;
  if ((q_rs | k_rs | v_rs | o_rs) & 7) return false;
  auto mis = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) != 0; };
  if (mis(q) || mis(k) || mis(v) || mis(o)) return false;
  return tc::tma_available();
}

cudaError_t launch_mha_tc_forward(int batch, int seq, int kv_seq, int heads, int64_t q_rs, int64_t k_rs, int64_t v_rs, int64_t o_rs, float scale,
                                  const void* q, const void* k, const void* v, void* o, float* attn, const int* tok_idx,
                                  const float* idx_bias, int n_emb, cudaStream_t st, const char** err) {
  mhatc::MapsM maps;
  memset(&maps, 0, sizeof(maps));
  const int width = heads * mhatc::kHd;
  if (!tc::encode_rows_map(&maps.q, q, batch, seq, width, q_rs, mhatc::kHd, mhatc::kBoxRows) ||
      !tc::encode_rows_map(&maps.k, k, batch, kv_seq, width, k_rs, mhatc::kHd, mhatc::kBoxRows) ||
      !tc::encode_rows_map(&maps.v, v, batch, kv_seq, width, v_rs, mhatc::kHd, mhatc::kBoxRows)) {
    if (err) *err = "cuTensorMapEncodeTiled failed";
    return cudaErrorInvalidValue;
  }
  const mhatc::SmemM map = mhatc::smem_m();
  const bool biased = tok_idx != nullptr;
  auto kernel = biased ? mhatc::mha_tc_kernel<true> : mhatc::mha_tc_kernel<false>;
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)map.total);
  if (e != cudaSuccess) {
    if (err) *err = "cudaFuncSetAttribute(max dynamic smem) failed";
    return e;
  }
  mhatc::ArgsM a;
  a.S = seq; a.Sk = kv_seq; a.s16 = (kv_seq + 15) / 16 * 16; a.nH = heads;
  a.c1 = scale * 1.4426950408889634f;
  a.o = static_cast<__nv_bfloat16*>(o); a.o_rs = o_rs; a.attn = attn;
  a.tok_idx = tok_idx; a.idx_bias = idx_bias; a.n_emb = biased ? n_emb : 0;
  const dim3 grid((seq + mhatc::kRowsM - 1) / mhatc::kRowsM, heads, batch);
  TimedRegion timed(st);
  kernel<<<grid, mhatc::kThreadsM, map.total, st>>>(a, maps);
  count_launch();
  return cudaGetLastError();
}

}  // namespace swin
