// Shifted-window attention, CUDA-core (FFMA) kernels: the fp32-exact path
// (north_star: <=1e-5 vs the reference in fp32 rules out TF32/bf16 MMAs,
// SURVEY.md F6), the backward pass, and the shapes the tcgen05 kernel does
// not cover.  One CTA works on one (window, head) tile at a time:
//
//   forward : S = scale * Q K^T + bias[rel_index] + shift_mask ; P = softmax(S)
//             O = P V ; lse = logsumexp(S)
//   backward: P recomputed from (Q, K, lse);  dP = dO V^T ; dS = P o (dP - delta)
//             dQ = scale dS K ; dK = scale dS^T Q ; dV = P^T dO ;
//             dTable[rel_index] += dS   (reduced in shared memory per CTA)
//
// Pad / roll / partition / reverse / crop are address arithmetic
// (swin_common.cuh::token_coord); no intermediate tensor is materialised.
#include "swin_common.cuh"

namespace swin {

template <int WS>
struct Tile {
  static constexpr int N = WS * WS;
  static constexpr int NC = (N + 31) / 32;       // key-column chunks per lane
  static constexpr int NCOL = NC * 32;
  static constexpr int NG = (N + 3) / 4;         // groups of 4 rows
  static constexpr int TS = NCOL + 4;            // stride of transposed [32][TS] arrays
  static constexpr int NT = (2 * WS - 1) * (2 * WS - 1);
};

constexpr int kThreads = 128;
constexpr int kWarps = kThreads / 32;

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void decode_item(const Problem& p, int win, int& b, int& wy, int& wx) {
  wx = win % p.nWx;
  int r = win / p.nWx;
  wy = r % p.nWy;
  b = r / p.nWy;
}

// ---------------------------------------------------------------------------
// forward
// ---------------------------------------------------------------------------
template <int WS>
struct FwdSmem {
  using TL = Tile<WS>;
  float qt[32 * TL::TS];
  float kt[32 * TL::TS];
  float vs[TL::N * 32];
  float4 ps[kWarps * TL::NCOL];
  float bias[TL::NT];
  int reg[TL::N];
  int pix[TL::N];
  int lpix[TL::N];
};

template <typename T, int WS>
__global__ void __launch_bounds__(kThreads) simt_fwd_kernel(const Problem p) {
  using TL = Tile<WS>;
  constexpr int N = TL::N, NC = TL::NC, TS = TL::TS;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FwdSmem<WS>& sm = *reinterpret_cast<FwdSmem<WS>*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int item = blockIdx.x;
  const int h = item % p.nH;
  int b, wy, wx;
  decode_item(p, item / p.nH, b, wy, wx);

  for (int t = tid; t < N; t += kThreads) {
    int y, x;
    token_coord(p, wy, wx, t, y, x);
    sm.pix[t] = (y < p.H && x < p.W) ? (b * p.H + y) * p.W + x : -1;
    sm.lpix[t] = (b * p.Hp + y) * p.Wp + x;
    sm.reg[t] = p.shift > 0 ? region_id(p, wy, wx, t) : 0;
  }
  for (int i = tid; i < TL::NT; i += kThreads) sm.bias[i] = p.table[i * p.nH + h];
  // zero the padding columns of the transposed tiles
  for (int i = tid; i < 32 * (TS - N); i += kThreads) {
    int d = i / (TS - N), c = N + i % (TS - N);
    sm.qt[d * TS + c] = 0.f;
    sm.kt[d * TS + c] = 0.f;
  }
  __syncthreads();

  const T* q = static_cast<const T*>(p.q);
  const T* k = static_cast<const T*>(p.k);
  const T* v = static_cast<const T*>(p.v);
  const T* qpad = static_cast<const T*>(p.q_pad);
  const T* kpad = static_cast<const T*>(p.k_pad);
  const T* vpad = static_cast<const T*>(p.v_pad);
  for (int idx = tid; idx < N * 8; idx += kThreads) {
    int t = idx >> 3, ch = (idx & 7) * 4;
    int px = sm.pix[t];
    int co = h * kHeadDim + ch;
    float4 q4 = make_float4(0.f, 0.f, 0.f, 0.f), k4 = q4, v4 = q4;
    if (px >= 0) {
      q4 = Elem<T>::load4(q + (int64_t)px * p.q_ps + co);
      k4 = Elem<T>::load4(k + (int64_t)px * p.k_ps + co);
      v4 = Elem<T>::load4(v + (int64_t)px * p.v_ps + co);
    } else {
      if (qpad) q4 = Elem<T>::load4(qpad + co);
      if (kpad) k4 = Elem<T>::load4(kpad + co);
      if (vpad) v4 = Elem<T>::load4(vpad + co);
    }
    sm.qt[(ch + 0) * TS + t] = q4.x; sm.qt[(ch + 1) * TS + t] = q4.y;
    sm.qt[(ch + 2) * TS + t] = q4.z; sm.qt[(ch + 3) * TS + t] = q4.w;
    sm.kt[(ch + 0) * TS + t] = k4.x; sm.kt[(ch + 1) * TS + t] = k4.y;
    sm.kt[(ch + 2) * TS + t] = k4.z; sm.kt[(ch + 3) * TS + t] = k4.w;
    *reinterpret_cast<float4*>(&sm.vs[t * 32 + ch]) = v4;
  }
  __syncthreads();

  int cy[NC], cx[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    int j = lane + 32 * c;
    cy[c] = j / WS;
    cx[c] = j - cy[c] * WS;
  }
  float4* ps = sm.ps + warp * TL::NCOL;
  T* o = static_cast<T*>(p.o);
  const float* mask_w = p.mask ? p.mask + (int64_t)((item / p.nH) % p.mask_nw) * N * N : nullptr;
  const Dropout drop = make_dropout(p, item / p.nH, h, N);

  for (int g = warp; g < TL::NG; g += kWarps) {
    float acc[4][NC];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < NC; ++c) acc[r][c] = 0.f;
#pragma unroll 8
    for (int d = 0; d < 32; ++d) {
      float4 qv = *reinterpret_cast<const float4*>(&sm.qt[d * TS + 4 * g]);
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        float kv = sm.kt[d * TS + lane + 32 * c];
        acc[0][c] = fmaf(qv.x, kv, acc[0][c]);
        acc[1][c] = fmaf(qv.y, kv, acc[1][c]);
        acc[2][c] = fmaf(qv.z, kv, acc[2][c]);
        acc[3][c] = fmaf(qv.w, kv, acc[3][c]);
      }
    }
    float inv[4], lse[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      int row = min(4 * g + r, N - 1);
      int ry = row / WS, rx = row - ry * WS;
      int rreg = sm.reg[row];
      float s[NC], m = -INFINITY;
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        int j = lane + 32 * c;
        if (j < N) {
          int bi = (ry - cy[c] + WS - 1) * (2 * WS - 1) + (rx - cx[c] + WS - 1);
          float val = fmaf(acc[r][c], p.scale, sm.bias[bi]);
          if (mask_w != nullptr) val += mask_w[row * N + j];
          else if (sm.reg[j] != rreg) val += -100.0f;
          s[c] = val;
        } else {
          s[c] = -INFINITY;
        }
        m = fmaxf(m, s[c]);
      }
      m = warp_max(m);
      float l = 0.f;
      float pr[NC];
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        pr[c] = expf(s[c] - m);
        l += pr[c];
      }
      l = warp_sum(l);
      inv[r] = 1.0f / l;
      lse[r] = m + logf(l);
#pragma unroll
      for (int c = 0; c < NC; ++c)
        reinterpret_cast<float*>(&ps[lane + 32 * c])[r] = pr[c] * drop.scale_of(N, row, min(lane + 32 * c, N - 1));
    }
    __syncwarp();
    float o0 = 0.f, o1 = 0.f, o2 = 0.f, o3 = 0.f;
#pragma unroll 7
    for (int j = 0; j < N; ++j) {
      float4 pv = ps[j];
      float vv = sm.vs[j * 32 + lane];
      o0 = fmaf(pv.x, vv, o0);
      o1 = fmaf(pv.y, vv, o1);
      o2 = fmaf(pv.z, vv, o2);
      o3 = fmaf(pv.w, vv, o3);
    }
    float ov[4] = {o0 * inv[0], o1 * inv[1], o2 * inv[2], o3 * inv[3]};
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      int row = 4 * g + r;
      if (row < N) {
        int px = sm.pix[row];
        if (px >= 0) Elem<T>::store1(o + (int64_t)px * p.o_ps + h * kHeadDim + lane, ov[r]);
        if (p.lse != nullptr && lane == 0) p.lse[(int64_t)sm.lpix[row] * p.nH + h] = lse[r];
      }
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// backward
// ---------------------------------------------------------------------------
template <int WS>
struct BwdSmem {
  using TL = Tile<WS>;
  float qt[32 * TL::TS];
  float kt[32 * TL::TS];
  float vt[32 * TL::TS];
  float dot[32 * TL::TS];
  float qs[TL::N * 32];
  float ks[TL::N * 32];
  float dos[TL::N * 32];
  float4 p1[kWarps * TL::NCOL];
  float4 p2[kWarps * TL::NCOL];
  float bias[TL::NT];
  float dtab[TL::NT];
  float lse[TL::NCOL];
  float delta[TL::NCOL];
  int reg[TL::NCOL];
  int pix[TL::N];
};

// Persistent: CTA c handles the contiguous item range [c*chunk, (c+1)*chunk) of
// the head-major item list (item = h * nWin + win), so the shared-memory
// d_table accumulator is flushed to global memory only when the head changes.
template <typename T, int WS>
__global__ void __launch_bounds__(kThreads) simt_bwd_kernel(const Problem p, int n_win, int chunk) {
  using TL = Tile<WS>;
  constexpr int N = TL::N, NC = TL::NC, TS = TL::TS, NCOL = TL::NCOL;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BwdSmem<WS>& sm = *reinterpret_cast<BwdSmem<WS>*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n_items = n_win * p.nH;
  const int first = blockIdx.x * chunk;
  const int last = min(first + chunk, n_items);
  if (first >= last) return;

  for (int i = tid; i < TL::NT; i += kThreads) sm.dtab[i] = 0.f;
  for (int i = tid; i < 32 * (TS - N); i += kThreads) {
    int d = i / (TS - N), c = N + i % (TS - N);
    sm.qt[d * TS + c] = 0.f; sm.kt[d * TS + c] = 0.f;
    sm.vt[d * TS + c] = 0.f; sm.dot[d * TS + c] = 0.f;
  }
  for (int i = N + tid; i < NCOL; i += kThreads) { sm.lse[i] = 0.f; sm.delta[i] = 0.f; sm.reg[i] = 0; }

  const T* q = static_cast<const T*>(p.q);
  const T* k = static_cast<const T*>(p.k);
  const T* v = static_cast<const T*>(p.v);
  const T* o = static_cast<const T*>(p.o_in);
  const T* d_o = static_cast<const T*>(p.d_o);
  T* dq = static_cast<T*>(p.dq);
  T* dk = static_cast<T*>(p.dk);
  T* dv = static_cast<T*>(p.dv);

  int cy[NC], cx[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    int j = min(lane + 32 * c, N - 1);
    cy[c] = j / WS;
    cx[c] = j - cy[c] * WS;
  }
  float4* p1 = sm.p1 + warp * NCOL;
  float4* p2 = sm.p2 + warp * NCOL;
  int cur_h = first / n_win;

  for (int item = first; item < last; ++item) {
    const int h = item / n_win;
    int b, wy, wx;
    decode_item(p, item - h * n_win, b, wy, wx);
    const float* mask_w = p.mask ? p.mask + (int64_t)((item - h * n_win) % p.mask_nw) * N * N : nullptr;
    const Dropout drop = make_dropout(p, item - h * n_win, h, N);
    __syncthreads();  // previous item fully consumed
    if (h != cur_h) {
      for (int i = tid; i < TL::NT; i += kThreads) {
        atomicAdd(&p.d_table[i * p.nH + cur_h], sm.dtab[i]);
        sm.dtab[i] = 0.f;
      }
      cur_h = h;
    }
    if (item == first || h != (item - 1) / n_win)
      for (int i = tid; i < TL::NT; i += kThreads) sm.bias[i] = p.table[i * p.nH + h];
    for (int t = tid; t < N; t += kThreads) {
      int y, x;
      token_coord(p, wy, wx, t, y, x);
      int px = (b * p.H + y) * p.W + x;   // backward requires H == Hp, W == Wp
      sm.pix[t] = px;
      sm.reg[t] = p.shift > 0 ? region_id(p, wy, wx, t) : 0;
      sm.lse[t] = p.lse[(int64_t)px * p.nH + h];
    }
    __syncthreads();
    constexpr int kLoadIters = (N * 8 + kThreads - 1) / kThreads;
    for (int it = 0; it < kLoadIters; ++it) {
      int idx = tid + it * kThreads;
      bool act = idx < N * 8;
      int t = act ? idx >> 3 : 0, ch = (idx & 7) * 4;
      int64_t px = sm.pix[t];
      int co = h * kHeadDim + ch;
      float4 q4 = Elem<T>::load4(q + px * p.q_ps + co);
      float4 k4 = Elem<T>::load4(k + px * p.k_ps + co);
      float4 v4 = Elem<T>::load4(v + px * p.v_ps + co);
      float4 g4 = Elem<T>::load4(d_o + px * p.C + co);
      float4 o4 = Elem<T>::load4(o + px * p.C + co);
      float part = g4.x * o4.x + g4.y * o4.y + g4.z * o4.z + g4.w * o4.w;
      part += __shfl_xor_sync(0xffffffffu, part, 1);
      part += __shfl_xor_sync(0xffffffffu, part, 2);
      part += __shfl_xor_sync(0xffffffffu, part, 4);
      if (act) {
        if ((idx & 7) == 0) sm.delta[t] = part;
        sm.qt[(ch + 0) * TS + t] = q4.x; sm.qt[(ch + 1) * TS + t] = q4.y;
        sm.qt[(ch + 2) * TS + t] = q4.z; sm.qt[(ch + 3) * TS + t] = q4.w;
        sm.kt[(ch + 0) * TS + t] = k4.x; sm.kt[(ch + 1) * TS + t] = k4.y;
        sm.kt[(ch + 2) * TS + t] = k4.z; sm.kt[(ch + 3) * TS + t] = k4.w;
        sm.vt[(ch + 0) * TS + t] = v4.x; sm.vt[(ch + 1) * TS + t] = v4.y;
        sm.vt[(ch + 2) * TS + t] = v4.z; sm.vt[(ch + 3) * TS + t] = v4.w;
        sm.dot[(ch + 0) * TS + t] = g4.x; sm.dot[(ch + 1) * TS + t] = g4.y;
        sm.dot[(ch + 2) * TS + t] = g4.z; sm.dot[(ch + 3) * TS + t] = g4.w;
        *reinterpret_cast<float4*>(&sm.qs[t * 32 + ch]) = q4;
        *reinterpret_cast<float4*>(&sm.ks[t * 32 + ch]) = k4;
        *reinterpret_cast<float4*>(&sm.dos[t * 32 + ch]) = g4;
      }
    }
    __syncthreads();

    // ---- phase A: 4 query rows per warp step, lanes over key columns -> dQ, dTable
    for (int g = warp; g < TL::NG; g += kWarps) {
      float as[4][NC], ad[4][NC];
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < NC; ++c) { as[r][c] = 0.f; ad[r][c] = 0.f; }
#pragma unroll 4
      for (int d = 0; d < 32; ++d) {
        float4 qv = *reinterpret_cast<const float4*>(&sm.qt[d * TS + 4 * g]);
        float4 gv = *reinterpret_cast<const float4*>(&sm.dot[d * TS + 4 * g]);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          float kv = sm.kt[d * TS + lane + 32 * c];
          float vv = sm.vt[d * TS + lane + 32 * c];
          as[0][c] = fmaf(qv.x, kv, as[0][c]); as[1][c] = fmaf(qv.y, kv, as[1][c]);
          as[2][c] = fmaf(qv.z, kv, as[2][c]); as[3][c] = fmaf(qv.w, kv, as[3][c]);
          ad[0][c] = fmaf(gv.x, vv, ad[0][c]); ad[1][c] = fmaf(gv.y, vv, ad[1][c]);
          ad[2][c] = fmaf(gv.z, vv, ad[2][c]); ad[3][c] = fmaf(gv.w, vv, ad[3][c]);
        }
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        int row = min(4 * g + r, N - 1);
        bool rv = 4 * g + r < N;
        int ry = row / WS, rx = row - ry * WS;
        int rreg = sm.reg[row];
        float rl = sm.lse[row], rd = sm.delta[row];
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          int j = lane + 32 * c;
          float ds = 0.f;
          if (rv && j < N) {
            int bi = (ry - cy[c] + WS - 1) * (2 * WS - 1) + (rx - cx[c] + WS - 1);
            float s = fmaf(as[r][c], p.scale, sm.bias[bi]);
            if (mask_w != nullptr) s += mask_w[row * N + j];
            else if (sm.reg[j] != rreg) s += -100.0f;
            float pr = expf(s - rl);
            ds = pr * (ad[r][c] * drop.scale_of(N, row, j) - rd);
            atomicAdd(&sm.dtab[bi], ds);
          }
          reinterpret_cast<float*>(&p1[j])[r] = ds;
        }
      }
      __syncwarp();
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll 7
      for (int j = 0; j < N; ++j) {
        float4 dsv = p1[j];
        float kv = sm.ks[j * 32 + lane];
        a0 = fmaf(dsv.x, kv, a0); a1 = fmaf(dsv.y, kv, a1);
        a2 = fmaf(dsv.z, kv, a2); a3 = fmaf(dsv.w, kv, a3);
      }
      float outv[4] = {a0 * p.scale, a1 * p.scale, a2 * p.scale, a3 * p.scale};
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        int row = 4 * g + r;
        if (row < N) Elem<T>::store1(dq + (int64_t)sm.pix[row] * p.q_ps + h * kHeadDim + lane, outv[r]);
      }
      __syncwarp();
    }

    // ---- phase B: 4 key columns per warp step, lanes over query rows -> dK, dV
    for (int g = warp; g < TL::NG; g += kWarps) {
      float as[4][NC], ad[4][NC];
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < NC; ++c) { as[r][c] = 0.f; ad[r][c] = 0.f; }
#pragma unroll 4
      for (int d = 0; d < 32; ++d) {
        float4 kv = *reinterpret_cast<const float4*>(&sm.kt[d * TS + 4 * g]);
        float4 vv = *reinterpret_cast<const float4*>(&sm.vt[d * TS + 4 * g]);
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          float qv = sm.qt[d * TS + lane + 32 * c];
          float gv = sm.dot[d * TS + lane + 32 * c];
          as[0][c] = fmaf(kv.x, qv, as[0][c]); as[1][c] = fmaf(kv.y, qv, as[1][c]);
          as[2][c] = fmaf(kv.z, qv, as[2][c]); as[3][c] = fmaf(kv.w, qv, as[3][c]);
          ad[0][c] = fmaf(vv.x, gv, ad[0][c]); ad[1][c] = fmaf(vv.y, gv, ad[1][c]);
          ad[2][c] = fmaf(vv.z, gv, ad[2][c]); ad[3][c] = fmaf(vv.w, gv, ad[3][c]);
        }
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        int col = min(4 * g + r, N - 1);
        bool cv = 4 * g + r < N;
        int ky = col / WS, kx = col - ky * WS;
        int creg = sm.reg[col];
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          int i = lane + 32 * c;
          float pr = 0.f, ds = 0.f;
          if (cv && i < N) {
            int bi = (cy[c] - ky + WS - 1) * (2 * WS - 1) + (cx[c] - kx + WS - 1);
            float s = fmaf(as[r][c], p.scale, sm.bias[bi]);
            if (mask_w != nullptr) s += mask_w[i * N + col];
            else if (sm.reg[i] != creg) s += -100.0f;
            pr = expf(s - sm.lse[i]);
            const float keep = drop.scale_of(N, i, col);
            ds = pr * (ad[r][c] * keep - sm.delta[i]);
            pr *= keep;   // dV = (dropout(P))^T dO
          }
          reinterpret_cast<float*>(&p1[i])[r] = pr;
          reinterpret_cast<float*>(&p2[i])[r] = ds;
        }
      }
      __syncwarp();
      float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f, k0 = 0.f, k1 = 0.f, k2 = 0.f, k3 = 0.f;
#pragma unroll 7
      for (int i = 0; i < N; ++i) {
        float4 pv = p1[i], dsv = p2[i];
        float gv = sm.dos[i * 32 + lane], qv = sm.qs[i * 32 + lane];
        v0 = fmaf(pv.x, gv, v0); v1 = fmaf(pv.y, gv, v1); v2 = fmaf(pv.z, gv, v2); v3 = fmaf(pv.w, gv, v3);
        k0 = fmaf(dsv.x, qv, k0); k1 = fmaf(dsv.y, qv, k1); k2 = fmaf(dsv.z, qv, k2); k3 = fmaf(dsv.w, qv, k3);
      }
      float dvv[4] = {v0, v1, v2, v3};
      float dkv[4] = {k0 * p.scale, k1 * p.scale, k2 * p.scale, k3 * p.scale};
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        int col = 4 * g + r;
        if (col < N) {
          int64_t px = sm.pix[col];
          Elem<T>::store1(dv + px * p.v_ps + h * kHeadDim + lane, dvv[r]);
          Elem<T>::store1(dk + px * p.k_ps + h * kHeadDim + lane, dkv[r]);
        }
      }
      __syncwarp();
    }
  }
  __syncthreads();
  for (int i = tid; i < TL::NT; i += kThreads) atomicAdd(&p.d_table[i * p.nH + cur_h], sm.dtab[i]);
}

// ---------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------
template <typename T, int WS>
static cudaError_t launch_fwd(const Problem& p, cudaStream_t st) {
  size_t smem = sizeof(FwdSmem<WS>);
  cudaError_t e = cudaFuncSetAttribute(simt_fwd_kernel<T, WS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int items = p.B * p.nWy * p.nWx * p.nH;
  {
    TimedRegion timed(st);
    simt_fwd_kernel<T, WS><<<items, kThreads, smem, st>>>(p);
  }
  count_launch();
  return cudaGetLastError();
}

template <typename T, int WS>
static cudaError_t launch_bwd(const Problem& p, cudaStream_t st) {
  size_t smem = sizeof(BwdSmem<WS>);
  cudaError_t e = cudaFuncSetAttribute(simt_bwd_kernel<T, WS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int dev = 0, sms = 148, occ = 1;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, simt_bwd_kernel<T, WS>, kThreads, smem);
  if (occ < 1) occ = 1;
  int n_win = p.B * p.nWy * p.nWx;
  int items = n_win * p.nH;
  // several waves per SM so the tail is short, but few enough flushes of d_table
  int ctas = sms * occ * 4;
  if (ctas > items) ctas = items;
  int chunk = (items + ctas - 1) / ctas;
  ctas = (items + chunk - 1) / chunk;
  TimedRegion timed(st);
  e = cudaMemsetAsync(p.d_table, 0, sizeof(float) * Tile<WS>::NT * p.nH, st);
  if (e != cudaSuccess) return e;
  simt_bwd_kernel<T, WS><<<ctas, kThreads, smem, st>>>(p, n_win, chunk);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_simt_forward(const Problem& p, int dtype, cudaStream_t st) {
  if (p.ws == 7) return dtype == SWIN_ATTN_F32 ? launch_fwd<float, 7>(p, st) : launch_fwd<__nv_bfloat16, 7>(p, st);
  if (p.ws == 12) return dtype == SWIN_ATTN_F32 ? launch_fwd<float, 12>(p, st) : launch_fwd<__nv_bfloat16, 12>(p, st);
  return cudaErrorInvalidValue;
}

cudaError_t launch_simt_backward(const Problem& p, int dtype, cudaStream_t st) {
  if (p.ws == 7) return dtype == SWIN_ATTN_F32 ? launch_bwd<float, 7>(p, st) : launch_bwd<__nv_bfloat16, 7>(p, st);
  if (p.ws == 12) return dtype == SWIN_ATTN_F32 ? launch_bwd<float, 12>(p, st) : launch_bwd<__nv_bfloat16, 12>(p, st);
  return cudaErrorInvalidValue;
}

}  // namespace swin
