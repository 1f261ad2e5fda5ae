// tcgen05 forward kernel + launcher; helpers and the design notes are in swin_attn_tc.cuh.
#include "swin_attn_tc.cuh"

namespace swin {
namespace tc {

__global__ void __launch_bounds__(kThreads, 1) tc_fwd_kernel(const Problem p, int n_tiles) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const SmemMap map = smem_map(p.nH);
  const uint32_t sbase = smem_u32(smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n_pairs = (n_tiles + 1) >> 1;
  const int n_local = (n_pairs - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  const uint32_t bar_full = sbase + map.bars;              // [kStages] producers -> MMA
  const uint32_t bar_empty = bar_full + 8 * kStages;       // [kStages] MMA -> producers
  const uint32_t bar_s = bar_empty + 8 * kStages;          // [kSlots]  S ready
  const uint32_t bar_p = bar_s + 8 * kSlots;               // [kSlots]  P written
  const uint32_t bar_o = bar_p + 8 * kSlots;               // [kSlots]  O ready
  const uint32_t bar_r = bar_o + 8 * kSlots;               // [kSlots]  O read, slot reusable

  // ---- one-time setup -------------------------------------------------------
  {
    uint4 z = make_uint4(0u, 0u, 0u, 0u);
    uint4* s4 = reinterpret_cast<uint4*>(smem);
    for (int i = tid; i < (int)(map.bias / 16); i += kThreads) s4[i] = z;   // tiles + P slots (pad rows must be 0)
    float* bias = reinterpret_cast<float*>(smem + map.bias);
    for (int i = tid; i < p.nH * 169; i += kThreads) {
      int h = i / 169, e = i - h * 169;
      bias[i] = p.table[e * p.nH + h] * kLog2e;
    }
    if (tid == 0) {
      for (int i = 0; i < kStages; ++i) {
        mbar_init(bar_full + 8 * i, kProducerWarps * 32);
        mbar_init(bar_empty + 8 * i, 1);
      }
      for (int i = 0; i < kSlots; ++i) {
        mbar_init(bar_s + 8 * i, 1);
        mbar_init(bar_p + 8 * i, 128);
        mbar_init(bar_o + 8 * i, 1);
        mbar_init(bar_r + 8 * i, 128);
      }
      fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(sbase + map.tmem_ptr, 512);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
  }
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + map.tmem_ptr);

  if (warp == 0 || warp == 2 || warp == 3) {
    // ===== producers: warp 0 -> q, warp 2 -> k, warp 3 -> v =====================
    const int which = warp == 0 ? 0 : warp - 1;
    const __nv_bfloat16* src = static_cast<const __nv_bfloat16*>(which == 0 ? p.q : which == 1 ? p.k : p.v);
    const __nv_bfloat16* pad = static_cast<const __nv_bfloat16*>(which == 0 ? p.q_pad : which == 1 ? p.k_pad : p.v_pad);
    const int64_t ps = which == 0 ? p.q_ps : which == 1 ? p.k_ps : p.v_ps;
    for (int n = 0; n < n_local; ++n) {
      const int s = n % kStages;
      const int pair = blockIdx.x + n * gridDim.x;
      mbar_wait(bar_empty + 8 * s, ((n / kStages) & 1) ^ 1);
      TileCoord c0 = tile_coord(p, 2 * pair, n_tiles), c1 = tile_coord(p, 2 * pair + 1, n_tiles);
      if (warp == 0 && lane == 0) TC_TRACE(p, n, 0);
      produce_pair(p, src, pad, ps, sbase + map.stages + s * kStageBytes + which * 2 * kTileBytes, c0, c1, lane);
      cp_async_commit();
      if (warp == 0 && lane == 0) TC_TRACE(p, n, 1);
      if (n >= 2) {
        cp_async_wait<2>();
        fence_proxy_async();
        mbar_arrive(bar_full + 8 * ((n - 2) % kStages));
      }
    }
    if (n_local >= 2) {
      cp_async_wait<1>();
      fence_proxy_async();
      mbar_arrive(bar_full + 8 * ((n_local - 2) % kStages));
    }
    if (n_local >= 1) {
      cp_async_wait<0>();
      fence_proxy_async();
      mbar_arrive(bar_full + 8 * ((n_local - 1) % kStages));
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) ==============================================
    if (lane == 0) {
      constexpr uint32_t idesc_s = instr_desc(64, 56, 0);
      constexpr uint32_t idesc_o = instr_desc(64, 32, 1);
      for (int n = 0; n <= n_local; ++n) {
        if (n < n_local) {
          const int s = n % kStages, a = n % kSlots;
          mbar_wait(bar_full + 8 * s, (n / kStages) & 1);
          TC_TRACE(p, n, 2);
          mbar_wait(bar_r + 8 * a, ((n / kSlots) & 1) ^ 1);
          TC_TRACE(p, n, 3);
          tc_fence_after();
          const uint32_t st = sbase + map.stages + s * kStageBytes;
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            const uint32_t d = tmem_base + ((uint32_t)(t * 16) << 16) + a * kSlotCols;
            const uint64_t qd = smem_desc(st + t * kTileBytes, 16, 512, 4);
            const uint64_t kd = smem_desc(st + (2 + t) * kTileBytes, 16, 512, 4);
#pragma unroll
            for (int k = 0; k < 2; ++k) umma_ss(d, qd + 2 * k, kd + 2 * k, idesc_s, k);
          }
          umma_commit(bar_s + 8 * a);
          TC_TRACE(p, n, 4);
        }
        if (n >= 1) {
          const int m = n - 1, s = m % kStages, a = m % kSlots;
          mbar_wait(bar_p + 8 * a, (m / kSlots) & 1);
          TC_TRACE(p, m, 5);
          tc_fence_after();
          const uint32_t st = sbase + map.stages + s * kStageBytes;
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            const uint32_t d = tmem_base + ((uint32_t)(t * 16) << 16) + a * kSlotCols + 64;
            const uint64_t pd = smem_desc(sbase + map.pslots + (a * 2 + t) * kPTileBytes, 16, 1024, 2);
            const uint64_t vd = smem_desc(st + (4 + t) * kTileBytes, 16, 512, 4);
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_ss(d, pd + 2 * k, vd + 64 * k, idesc_o, k);
          }
          umma_commit(bar_o + 8 * a);
          umma_commit(bar_empty + 8 * s);
          TC_TRACE(p, m, 6);
        }
      }
    }
    __syncwarp();
  } else {
    // ===== softmax / epilogue warpgroups =======================================
    const int g = (warp >> 2) - 1;          // slot owned by this warpgroup
    const int q = warp & 3;                 // TMEM sub-partition
    const int sel = lane >> 4;              // which tile of the pair
    const int row = q * 16 + (lane & 15);   // row of S / O held by this thread
    const int rowc = row < N ? row : N - 1;
    const int ri = rowc / WS, rj = rowc - ri * WS;
    const int base_i = ri * 13 + rj + 84;
    const float* bias = reinterpret_cast<const float*>(smem + map.bias);
    const float c1 = p.scale * kLog2e;
    // tokens whose window-row (hb) / window-column (wb) index is >= ws - shift (SW-MSA regions)
    uint64_t hb = 0, wb = 0;
    if (p.shift > 0) {
      for (int t = 0; t < N; ++t) {
        if (t / WS >= WS - p.shift) hb |= 1ull << t;
        if (t % WS >= WS - p.shift) wb |= 1ull << t;
      }
    }
    const uint64_t all = (1ull << N) - 1;
    const uint32_t p_row = sbase + map.pslots + (g * 2 + sel) * kPTileBytes + row * 128;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + g * kSlotCols;
    __nv_bfloat16* out = static_cast<__nv_bfloat16*>(p.o);

    int m = 0;
    for (int n = g; n < n_local; n += kSlots, ++m) {
      const int pair = blockIdx.x + n * gridDim.x;
      const TileCoord c = tile_coord(p, 2 * pair + sel, n_tiles);
      int y, x;
      token_pixel(p, c, rowc, y, x);
      uint64_t masked = 0;
      if (p.shift > 0) {
        uint64_t hm = c.wy == p.nWy - 1 ? hb : 0, wm = c.wx == p.nWx - 1 ? wb : 0;
        masked = (((hm >> rowc) & 1) ? (~hm & all) : hm) | (((wm >> rowc) & 1) ? (~wm & all) : wm);
      }
      const float* brow = bias + c.h * 169 + base_i;

      const bool tr = q == 0 && lane == 0;
      if (tr) TC_TRACE(p, n, 7);
      mbar_wait(bar_s + 8 * g, m & 1);
      tc_fence_after();
      if (tr) TC_TRACE(p, n, 8);
      uint32_t r[N + 15];
      TMEM_LD_X16(taddr, (r + 0));
      TMEM_LD_X16(taddr + 16, (r + 16));
      TMEM_LD_X16(taddr + 32, (r + 32));
      r[48] = tmem_ld_x1(taddr + 48);
      tmem_ld_wait();
      if (tr) TC_TRACE(p, n, 9);

      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < N; ++j) {
        const int cj = (j / WS) * 13 + (j % WS);
        float t = fmaf(__uint_as_float(r[j]), c1, brow[-cj]);
        if (p.shift > 0 && ((masked >> j) & 1)) t += -100.0f * kLog2e;
        r[j] = __float_as_uint(t);
        mx = fmaxf(mx, t);
      }
      float l = 0.f;
#pragma unroll
      for (int j = 0; j < N; ++j) {
        float e = fast_exp2(__uint_as_float(r[j]) - mx);
        l += e;
        r[j] = __float_as_uint(e);
      }
      // P row (64 bf16, columns >= 49 zero) into the 128B-swizzled A tile of the PV MMA
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        uint32_t w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int j0 = ch * 8 + 2 * e;
          float lo = j0 < N ? __uint_as_float(r[j0 < N ? j0 : 0]) : 0.f;
          float hi = j0 + 1 < N ? __uint_as_float(r[j0 + 1 < N ? j0 + 1 : 0]) : 0.f;
          w[e] = pack_bf16(lo, hi);
        }
        const uint32_t dst = p_row + ((ch ^ (row & 7)) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
      }
      if (tr) TC_TRACE(p, n, 10);
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(bar_p + 8 * g);
      if (tr) TC_TRACE(p, n, 11);

      mbar_wait(bar_o + 8 * g, m & 1);
      tc_fence_after();
      if (tr) TC_TRACE(p, n, 12);
      uint32_t o[32];
      TMEM_LD_X16(taddr + 64, (o + 0));
      TMEM_LD_X16(taddr + 80, (o + 16));
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(bar_r + 8 * g);
      if (tr) TC_TRACE(p, n, 13);

      if (c.valid && row < N) {
        const float inv = 1.0f / l;
        if (y < p.H && x < p.W) {
          __nv_bfloat16* dst = out + ((int64_t)(c.b * p.H + y) * p.W + x) * p.o_ps + c.h * kHeadDim;
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            uint4 v;
            v.x = pack_bf16(__uint_as_float(o[8 * ch + 0]) * inv, __uint_as_float(o[8 * ch + 1]) * inv);
            v.y = pack_bf16(__uint_as_float(o[8 * ch + 2]) * inv, __uint_as_float(o[8 * ch + 3]) * inv);
            v.z = pack_bf16(__uint_as_float(o[8 * ch + 4]) * inv, __uint_as_float(o[8 * ch + 5]) * inv);
            v.w = pack_bf16(__uint_as_float(o[8 * ch + 6]) * inv, __uint_as_float(o[8 * ch + 7]) * inv);
            *reinterpret_cast<uint4*>(dst + 8 * ch) = v;
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

bool tc_forward_supported(const Problem& p, int dtype) {
  if (dtype != SWIN_ATTN_BF16 || p.ws != 7 || p.mask != nullptr) return false;
  // 16-byte cp.async / vector stores: pixel strides multiples of 8 bf16, pointers 16B aligned
  if ((p.q_ps | p.k_ps | p.v_ps | p.o_ps) & 7) return false;
  auto mis = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) != 0; };
  if (mis(p.q) || mis(p.k) || mis(p.v) || mis(p.o) || mis(p.q_pad) || mis(p.k_pad) || mis(p.v_pad)) return false;
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
  const tc::SmemMap map = tc::smem_map(p.nH);
  cudaError_t e = cudaFuncSetAttribute(tc::tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)map.total);
  if (e != cudaSuccess) {
    if (err) *err = "cudaFuncSetAttribute(max dynamic smem) failed";
    return e;
  }
  const int n_tiles = p.B * p.nWy * p.nWx * p.nH;
  const int n_pairs = (n_tiles + 1) / 2;
  const int grid = n_pairs < sms ? n_pairs : sms;
  tc::tc_fwd_kernel<<<grid, tc::kThreads, map.total, st>>>(p, n_tiles);
  count_launch();
  return cudaGetLastError();
}

}  // namespace swin
