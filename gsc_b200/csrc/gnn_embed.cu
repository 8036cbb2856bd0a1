// Fused GNN embedder: every GATv2Conv(aggr='mean') application of GNNEmbedder.forward plus the
// global_mean_pool in ONE launch (reference src/rlsp/agents/models.py:37-53; PyG 2.3.0 GATv2Conv
// semantics restated in SURVEY.md section 3.5 and oracle/gsc_oracle.py).
//
// Mapping.  A CTA owns a chunk of G consecutive graphs (G * max_nodes <= cap nodes) and keeps
// their node states in shared memory for the whole forward: two [cap][HP] fp32 arrays (xl = lin_l
// output that neighbours gather, hx = lin_r output / layer output) next to the transposed,
// zero-padded weights of every conv.  One thread per node:
//   phase A  xl_i = Wl h_i + bl, xr_i = Wr h_i + br          (weights broadcast from smem, float4)
//   phase B  online segment-softmax over the CSR row of i, gathering xl_j rows with 128-bit LDS,
//            out_i = (1/deg_i) * sum_j alpha_ij xl_j + bias   (+ ReLU)
// and finally one thread per (graph, channel) for the mean-pool.  HBM traffic is the compulsory
// 4nF + 4(e+n) + 4(n+1) bytes in and 4H bytes out per graph (+ the saved activations when training).
#include "common.cuh"
#include "tc_ptx.cuh"
#include <math.h>
#include <stdlib.h>

namespace gsc {

constexpr int kGnnThreads = 256;
constexpr float kNegSlope = 0.2f;  // GATv2Conv default negative_slope

struct GnnArgs {
  const float* x;
  const int32_t* rowptr;
  const int32_t* col;
  const int32_t* graph_ptr;
  const float* params;
  float* emb;
  float* saved;           // fwd: written (nullable); bwd: read
  const float* d_emb;     // bwd only
  float* partials;        // bwd only: [gridDim.x][param_floats]
  int64_t Nt;
  int64_t emb_ld;         // row stride of emb (fwd) / d_emb (bwd), floats
  int B, in_dim, H, num_convs, num_apps, G, cap;
  int app_conv[GSC_MAX_APPS];
  int app_relu[GSC_MAX_APPS];
};

template <int HP>
struct ConvLayout {
  // smem floats of one conv: WlT[HP][HP] (k-major: WlT[k*HP + c] = Wl[c][k]), WrT, bl, br, att, bias
  static constexpr int kWl = 0, kWr = HP * HP, kBl = 2 * HP * HP, kBr = kBl + HP, kAtt = kBr + HP,
                       kBias = kAtt + HP, kFloats = kBias + HP;
};

__host__ __device__ inline int64_t conv_param_floats(int in, int H) { return 4LL * H + 2LL * H * in; }
__host__ __device__ inline int64_t conv_param_offset(int conv, int in_dim, int H) {
  return conv == 0 ? 0 : conv_param_floats(in_dim, H) + (int64_t)(conv - 1) * conv_param_floats(H, H);
}

// Global (state-dict order: att, bias, Wl[H,in], bl, Wr[H,in], br) -> padded transposed smem copy.
template <int HP>
__device__ void load_conv_weights(float* __restrict__ dst, const float* __restrict__ src, int in, int H) {
  using L = ConvLayout<HP>;
  for (int t = threadIdx.x; t < L::kFloats; t += blockDim.x) dst[t] = 0.f;
  __syncthreads();
  const float* att = src;
  const float* bias = src + H;
  const float* Wl = bias + H;
  const float* bl = Wl + (int64_t)H * in;
  const float* Wr = bl + H;
  const float* br = Wr + (int64_t)H * in;
  for (int t = threadIdx.x; t < H * in; t += blockDim.x) {
    int c = t / in, k = t - c * in;
    dst[L::kWl + k * HP + c] = Wl[t];
    dst[L::kWr + k * HP + c] = Wr[t];
  }
  for (int t = threadIdx.x; t < H; t += blockDim.x) {
    dst[L::kBl + t] = bl[t];
    dst[L::kBr + t] = br[t];
    dst[L::kAtt + t] = att[t];
    dst[L::kBias + t] = bias[t];
  }
}

// xl = Wl h + bl, xr = Wr h + br for one node; weights broadcast from smem.
template <int HP>
__device__ __forceinline__ void node_transform(const float* __restrict__ w, const float (&hin)[HP],
                                               int d_in, float (&xl)[HP], float (&xr)[HP]) {
  using L = ConvLayout<HP>;
#pragma unroll
  for (int c4 = 0; c4 < HP / 4; ++c4) {
    float4 a = reinterpret_cast<const float4*>(w + L::kBl)[c4];
    float4 b = reinterpret_cast<const float4*>(w + L::kBr)[c4];
    xl[4 * c4] = a.x; xl[4 * c4 + 1] = a.y; xl[4 * c4 + 2] = a.z; xl[4 * c4 + 3] = a.w;
    xr[4 * c4] = b.x; xr[4 * c4 + 1] = b.y; xr[4 * c4 + 2] = b.z; xr[4 * c4 + 3] = b.w;
  }
#pragma unroll
  for (int k = 0; k < HP; ++k) {
    if (k < d_in) {
      const float hk = hin[k];
      const float4* wl = reinterpret_cast<const float4*>(w + L::kWl + k * HP);
      const float4* wr = reinterpret_cast<const float4*>(w + L::kWr + k * HP);
#pragma unroll
      for (int c4 = 0; c4 < HP / 4; ++c4) {
        float4 a = wl[c4], b = wr[c4];
        xl[4 * c4] = fmaf(hk, a.x, xl[4 * c4]);
        xl[4 * c4 + 1] = fmaf(hk, a.y, xl[4 * c4 + 1]);
        xl[4 * c4 + 2] = fmaf(hk, a.z, xl[4 * c4 + 2]);
        xl[4 * c4 + 3] = fmaf(hk, a.w, xl[4 * c4 + 3]);
        xr[4 * c4] = fmaf(hk, b.x, xr[4 * c4]);
        xr[4 * c4 + 1] = fmaf(hk, b.y, xr[4 * c4 + 1]);
        xr[4 * c4 + 2] = fmaf(hk, b.z, xr[4 * c4 + 2]);
        xr[4 * c4 + 3] = fmaf(hk, b.w, xr[4 * c4 + 3]);
      }
    }
  }
}

template <int HP>
__device__ __forceinline__ void load_row(const float* __restrict__ row, float (&v)[HP]) {
#pragma unroll
  for (int c4 = 0; c4 < HP / 4; ++c4) {
    float4 a = reinterpret_cast<const float4*>(row)[c4];
    v[4 * c4] = a.x; v[4 * c4 + 1] = a.y; v[4 * c4 + 2] = a.z; v[4 * c4 + 3] = a.w;
  }
}
template <int HP>
__device__ __forceinline__ void store_row(float* __restrict__ row, const float (&v)[HP]) {
#pragma unroll
  for (int c4 = 0; c4 < HP / 4; ++c4)
    reinterpret_cast<float4*>(row)[c4] = make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
}

// Input row of application `a` for node (nbase + i): x for the encoder, else the smem state.
template <int HP>
__device__ __forceinline__ void load_app_input(const GnnArgs& p, int a, int64_t node, const float* hx_row,
                                               float (&hin)[HP]) {
  if (a == 0) {
#pragma unroll
    for (int k = 0; k < HP; ++k) hin[k] = (k < p.in_dim) ? p.x[node * p.in_dim + k] : 0.f;
  } else {
    load_row<HP>(hx_row, hin);
  }
}

template <int HP>
__global__ void __launch_bounds__(kGnnThreads, 2) gnn_fwd_kernel(const GnnArgs p) {
  using L = ConvLayout<HP>;
  extern __shared__ float4 smem4[];
  float* wsm = reinterpret_cast<float*>(smem4);
  float* xl_s = wsm + p.num_convs * L::kFloats;
  float* hx_s = xl_s + (size_t)p.cap * HP;

  for (int cv = 0; cv < p.num_convs; ++cv)
    load_conv_weights<HP>(wsm + cv * L::kFloats, p.params + conv_param_offset(cv, p.in_dim, p.H),
                          cv == 0 ? p.in_dim : p.H, p.H);
  __syncthreads();

  const int num_chunks = (p.B + p.G - 1) / p.G;
  for (int chunk = blockIdx.x; chunk < num_chunks; chunk += gridDim.x) {
    const int g0 = chunk * p.G, g1 = min(p.B, g0 + p.G);
    const int nbase = p.graph_ptr[g0];
    const int nc = p.graph_ptr[g1] - nbase;

    for (int a = 0; a < p.num_apps; ++a) {
      const int cv = p.app_conv[a];
      const float* w = wsm + cv * L::kFloats;
      const int d_in = (cv == 0) ? p.in_dim : p.H;
      // ---- phase A: per-node linear transforms ----
      for (int i = threadIdx.x; i < nc; i += kGnnThreads) {
        float hin[HP], xl[HP], xr[HP];
        load_app_input<HP>(p, a, (int64_t)nbase + i, hx_s + (size_t)i * HP, hin);
        node_transform<HP>(w, hin, d_in, xl, xr);
        store_row<HP>(xl_s + (size_t)i * HP, xl);
        store_row<HP>(hx_s + (size_t)i * HP, xr);
      }
      __syncthreads();
      // ---- phase B: attention scores, online segment softmax, mean aggregation ----
      for (int i = threadIdx.x; i < nc; i += kGnnThreads) {
        float xr[HP], att[HP], acc[HP];
        load_row<HP>(hx_s + (size_t)i * HP, xr);
        load_row<HP>(w + L::kAtt, att);
#pragma unroll
        for (int c = 0; c < HP; ++c) acc[c] = 0.f;
        float m = -INFINITY, l = 0.f;
        const int e0 = p.rowptr[nbase + i], e1 = p.rowptr[nbase + i + 1];
        for (int e = e0; e < e1; ++e) {
          const int j = p.col[e] - nbase;
          float v[HP];
          load_row<HP>(xl_s + (size_t)j * HP, v);
          float s = 0.f;
#pragma unroll
          for (int c = 0; c < HP; ++c) {
            float t = v[c] + xr[c];
            t = fmaxf(t, kNegSlope * t);
            s = fmaf(att[c], t, s);
          }
          const float m_new = fmaxf(m, s);
          const float sc = expf(m - m_new);
          const float pe = expf(s - m_new);
          l = fmaf(l, sc, pe);
#pragma unroll
          for (int c = 0; c < HP; ++c) acc[c] = fmaf(acc[c], sc, pe * v[c]);
          m = m_new;
        }
        const float inv = 1.f / ((l + 1e-16f) * (float)(e1 - e0));
        float bias[HP];
        load_row<HP>(w + L::kBias, bias);
        const bool relu = p.app_relu[a] != 0;
#pragma unroll
        for (int c = 0; c < HP; ++c) {
          float o = fmaf(acc[c], inv, bias[c]);
          acc[c] = relu ? fmaxf(o, 0.f) : o;
        }
        store_row<HP>(hx_s + (size_t)i * HP, acc);
        if (p.saved != nullptr)
          store_row<HP>(p.saved + ((size_t)a * p.Nt + nbase + i) * HP, acc);
      }
      __syncthreads();
    }
    // ---- global_mean_pool: one thread per (graph, channel) ----
    for (int t = threadIdx.x; t < (g1 - g0) * HP; t += kGnnThreads) {
      const int g = g0 + t / HP, c = t % HP;
      if (c < p.H) {
        const int n0 = p.graph_ptr[g] - nbase, n1 = p.graph_ptr[g + 1] - nbase;
        float s = 0.f;
        for (int i = n0; i < n1; ++i) s += hx_s[(size_t)i * HP + c];
        p.emb[(size_t)g * p.emb_ld + c] = (n1 > n0) ? s / (float)(n1 - n0) : 0.f;
      }
    }
    __syncthreads();
  }
}

// ================================ tcgen05 forward ===========================================
// Same fused forward, with the per-node linear transforms (the bulk of the FLOPs: 2*2*H*d_in per
// node and application) moved to the 5th-gen tensor cores while staying fp32-faithful:
//   * node rows are split h = h_hi + h_lo (both exactly representable in TF32) and written by their
//     owner thread into two shared-memory A operands; the weights [Wl;Wr] are split the same way
//     once per CTA into two B operands;
//   * one elected thread issues tcgen05.mma kind::tf32 for A_hi*B_hi + A_hi*B_lo + A_lo*B_hi
//     (3xTF32, relative error ~2^-21) into a 128-lane x 2H-column TMEM accumulator per 128 nodes;
//   * TMEM lane i belongs to node i, so tcgen05.ld hands every thread exactly its own xl|xr row --
//     the thread-per-node layout that the attention/softmax/aggregation phase wants.
// Operands use the no-swizzle K-major core-matrix layout (8 rows x 16 B), which lets a thread store
// its row with 128-bit conflict-free st.shared and needs no padding for K = H.
template <int HP>
struct TcL {
  static constexpr int kN = 2 * HP;                       // accumulator columns: xl | xr
  static constexpr int kChunks = HP / 4;                  // 16-byte chunks per operand row
  static constexpr int kGroupBytes = kChunks * 128;       // 8 rows x kChunks core matrices
  static constexpr int kATile = 16 * kGroupBytes;         // 128-row A operand (== 128*HP*4 bytes)
  static constexpr int kB = (kN / 8) * kGroupBytes;       // kN-row B operand
  static constexpr int kVecFloats = 4 * HP;               // bl, br, att, bias
  static constexpr int kConvBytes = 2 * kB + kVecFloats * 4;
  __device__ __forceinline__ static int off(int r, int c) { return (r >> 3) * kGroupBytes + c * 128 + (r & 7) * 16; }
};

template <int HP>
__device__ __forceinline__ void store_split_row(uint8_t* a_hi, uint8_t* a_lo, int row, const float (&h)[HP]) {
  using T = TcL<HP>;
#pragma unroll
  for (int c4 = 0; c4 < HP / 4; ++c4) {
    float4 hi, lo;
    hi.x = to_tf32(h[4 * c4]);     lo.x = to_tf32(h[4 * c4] - hi.x);
    hi.y = to_tf32(h[4 * c4 + 1]); lo.y = to_tf32(h[4 * c4 + 1] - hi.y);
    hi.z = to_tf32(h[4 * c4 + 2]); lo.z = to_tf32(h[4 * c4 + 2] - hi.z);
    hi.w = to_tf32(h[4 * c4 + 3]); lo.w = to_tf32(h[4 * c4 + 3] - hi.w);
    *reinterpret_cast<float4*>(a_hi + T::off(row, c4)) = hi;
    *reinterpret_cast<float4*>(a_lo + T::off(row, c4)) = lo;
  }
}

// HP consecutive accumulator columns of this thread's TMEM lane.
template <int HP>
__device__ __forceinline__ void tmem_load_row(uint32_t taddr, float (&v)[HP]) {
  uint32_t r[HP / 8][8];
#pragma unroll
  for (int q = 0; q < HP / 8; ++q) tc_ld8_nowait(taddr + 8 * q, r[q]);
  tc_ld_wait();
#pragma unroll
  for (int q = 0; q < HP / 8; ++q)
#pragma unroll
    for (int e = 0; e < 8; ++e) v[8 * q + e] = __uint_as_float(r[q][e]);
}

template <int HP>
__global__ void __launch_bounds__(kGnnThreads, 2) gnn_fwd_tc_kernel(const GnnArgs p, const int tmem_cols) {
  using T = TcL<HP>;
  extern __shared__ float4 smem4[];
  uint8_t* sm = reinterpret_cast<uint8_t*>(smem4);
  const int tiles = p.cap / 128;
  uint8_t* convs = sm;
  uint8_t* a_hi = convs + p.num_convs * T::kConvBytes;
  uint8_t* a_lo = a_hi + tiles * T::kATile;
  float* xl_s = reinterpret_cast<float*>(a_lo + tiles * T::kATile);
  uint64_t* bar_ptr = reinterpret_cast<uint64_t*>(xl_s + (size_t)p.cap * HP);
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar_ptr + 1);
  const uint32_t bar = smem_u32(bar_ptr);
  const int tid = threadIdx.x, warp = tid >> 5;

  // ---- weights -> split B operands + small vectors ----
  for (int cv = 0; cv < p.num_convs; ++cv) {
    const int in = cv == 0 ? p.in_dim : p.H, H = p.H;
    const float* src = p.params + conv_param_offset(cv, p.in_dim, p.H);
    const float* att = src;
    const float* bias = src + H;
    const float* Wl = bias + H;
    const float* bl = Wl + (int64_t)H * in;
    const float* Wr = bl + H;
    const float* br = Wr + (int64_t)H * in;
    uint8_t* b_hi = convs + cv * T::kConvBytes;
    uint8_t* b_lo = b_hi + T::kB;
    float* vec = reinterpret_cast<float*>(b_lo + T::kB);
    for (int t = tid; t < T::kN * HP; t += kGnnThreads) {
      const int n = t / HP, k = t - n * HP, c = n % HP;
      float val = 0.f;
      if (c < H && k < in) val = (n >= HP ? Wr : Wl)[c * in + k];
      const float hi = to_tf32(val), lo = to_tf32(val - hi);
      const int byte = T::off(n, k >> 2) + (k & 3) * 4;
      *reinterpret_cast<float*>(b_hi + byte) = hi;
      *reinterpret_cast<float*>(b_lo + byte) = lo;
    }
    for (int t = tid; t < HP; t += kGnnThreads) {
      const bool ok = t < H;
      vec[t] = ok ? bl[t] : 0.f;
      vec[HP + t] = ok ? br[t] : 0.f;
      vec[2 * HP + t] = ok ? att[t] : 0.f;
      vec[3 * HP + t] = ok ? bias[t] : 0.f;
    }
  }
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    __syncwarp();
    tmem_alloc(smem_u32(tslot), (uint32_t)tmem_cols);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tslot;
  const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;     // TMEM lane quadrant of this warp
  // instruction descriptor: D=f32, A=B=tf32, both K-major, M=128, N=2*HP
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(T::kN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  uint32_t parity = 0;
  // optional phase timers (debug): partials[0..7] += cycles of {setup, x rows, mma, xl store, phase B, pool}
  long long tprev = clock64();
  const bool dbg = p.partials != nullptr && blockIdx.x == 0 && tid == 0;
#define GSC_TICK(slot)                                                                  \
  do {                                                                                  \
    if (dbg) { long long tn = clock64(); p.partials[slot] += (float)(tn - tprev); tprev = tn; } \
  } while (0)
  GSC_TICK(0);

  const int num_chunks = (p.B + p.G - 1) / p.G;
  for (int chunk = blockIdx.x; chunk < num_chunks; chunk += gridDim.x) {
    const int g0 = chunk * p.G, g1 = min(p.B, g0 + p.G);
    const int nbase = p.graph_ptr[g0];
    const int nc = p.graph_ptr[g1] - nbase;
    const int warp_first = tid & ~31;                                // slot loops must be warp-uniform (tcgen05.ld)

    // encoder input rows: x -> split operand rows
    for (int i = tid; i < nc; i += kGnnThreads) {
      float h[HP];
#pragma unroll
      for (int k = 0; k < HP; ++k) h[k] = (k < p.in_dim) ? p.x[((int64_t)nbase + i) * p.in_dim + k] : 0.f;
      store_split_row<HP>(a_hi + (i >> 7) * T::kATile, a_lo + (i >> 7) * T::kATile, i & 127, h);
    }
    GSC_TICK(1);

    for (int a = 0; a < p.num_apps; ++a) {
      const int cv = p.app_conv[a];
      const uint8_t* b_hi = convs + cv * T::kConvBytes;
      const float* vec = reinterpret_cast<const float*>(b_hi + 2 * T::kB);
      const int ksteps = ((cv == 0 ? p.in_dim : p.H) + 7) / 8;
      const bool last = a == p.num_apps - 1;

      // ---- phase A on the tensor cores ----
      fence_proxy_async_smem();          // operand rows were written with st.shared
      tc_fence_before();
      __syncthreads();
      if (tid == 0) {
        tc_fence_after();
        const int used_tiles = (nc + 127) >> 7;
        for (int t = 0; t < used_tiles; ++t) {
          const uint32_t ah = smem_u32(a_hi + t * T::kATile), al = smem_u32(a_lo + t * T::kATile);
          const uint32_t bh = smem_u32(b_hi), blo = bh + T::kB;
          const uint32_t d = tmem_base + (uint32_t)(t * T::kN);
          for (int s = 0; s < ksteps; ++s) {
            const uint32_t ko = s * 256;   // 8 k = two 16-byte chunks = 2 core matrices of 128 B
            tc_mma_tf32(d, smem_desc(ah + ko, 128, T::kGroupBytes, 0), smem_desc(bh + ko, 128, T::kGroupBytes, 0), idesc, s > 0);
            tc_mma_tf32(d, smem_desc(ah + ko, 128, T::kGroupBytes, 0), smem_desc(blo + ko, 128, T::kGroupBytes, 0), idesc, 1);
            tc_mma_tf32(d, smem_desc(al + ko, 128, T::kGroupBytes, 0), smem_desc(bh + ko, 128, T::kGroupBytes, 0), idesc, 1);
          }
        }
        tc_commit(bar);
      }
      mbar_wait(bar, parity);
      parity ^= 1u;
      tc_fence_after();
      GSC_TICK(2);
      // xl rows -> shared memory (neighbours gather them)
      for (int i0 = warp_first; i0 < nc; i0 += kGnnThreads) {
        const int i = i0 + (tid & 31);
        float xl[HP];
        tmem_load_row<HP>(tmem_base + lane_base + (uint32_t)((i >> 7) * T::kN), xl);
        if (i < nc) {
#pragma unroll
          for (int c = 0; c < HP; ++c) xl[c] += vec[c];
          store_row<HP>(xl_s + (size_t)i * HP, xl);
        }
      }
      __syncthreads();
      GSC_TICK(3);

      // ---- phase B: attention scores, online segment softmax, mean aggregation ----
      for (int i0 = warp_first; i0 < nc; i0 += kGnnThreads) {
        const int i = i0 + (tid & 31);
        float xr[HP], att[HP], acc[HP];
        tmem_load_row<HP>(tmem_base + lane_base + (uint32_t)((i >> 7) * T::kN + HP), xr);
        if (i < nc) {
          load_row<HP>(vec + 2 * HP, att);
#pragma unroll
          for (int c = 0; c < HP; ++c) { xr[c] += vec[HP + c]; acc[c] = 0.f; }
          float m = -INFINITY, l = 0.f;
          const int e0 = p.rowptr[nbase + i], e1 = p.rowptr[nbase + i + 1];
          for (int e = e0; e < e1; ++e) {
            const int j = p.col[e] - nbase;
            float v[HP];
            load_row<HP>(xl_s + (size_t)j * HP, v);
            float s = 0.f;
#pragma unroll
            for (int c = 0; c < HP; ++c) {
              float t = v[c] + xr[c];
              t = fmaxf(t, kNegSlope * t);
              s = fmaf(att[c], t, s);
            }
            const float m_new = fmaxf(m, s);
            const float sc = expf(m - m_new);
            const float pe = expf(s - m_new);
            l = fmaf(l, sc, pe);
#pragma unroll
            for (int c = 0; c < HP; ++c) acc[c] = fmaf(acc[c], sc, pe * v[c]);
            m = m_new;
          }
          const float inv = 1.f / ((l + 1e-16f) * (float)(e1 - e0));
          const bool relu = p.app_relu[a] != 0;
#pragma unroll
          for (int c = 0; c < HP; ++c) {
            const float o = fmaf(acc[c], inv, vec[3 * HP + c]);
            acc[c] = relu ? fmaxf(o, 0.f) : o;
          }
          if (p.saved != nullptr) store_row<HP>(p.saved + ((size_t)a * p.Nt + nbase + i) * HP, acc);
          if (last) store_row<HP>(reinterpret_cast<float*>(a_lo) + (size_t)i * HP, acc);     // pool input
          else store_split_row<HP>(a_hi + (i >> 7) * T::kATile, a_lo + (i >> 7) * T::kATile, i & 127, acc);
        }
      }
      GSC_TICK(4);
    }
    __syncthreads();
    // ---- global_mean_pool ----
    const float* hx = reinterpret_cast<const float*>(a_lo);
    for (int t = tid; t < (g1 - g0) * HP; t += kGnnThreads) {
      const int g = g0 + t / HP, c = t % HP;
      if (c < p.H) {
        const int n0 = p.graph_ptr[g] - nbase, n1 = p.graph_ptr[g + 1] - nbase;
        float s = 0.f;
        for (int i = n0; i < n1; ++i) s += hx[(size_t)i * HP + c];
        p.emb[(size_t)g * p.emb_ld + c] = (n1 > n0) ? s / (float)(n1 - n0) : 0.f;
      }
    }
    __syncthreads();
    GSC_TICK(5);
  }
#undef GSC_TICK

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)tmem_cols);
  }
}

// ================================ backward ==================================================
// Per chunk and per application (last to first) the CTA recomputes xl, xr and the attention from
// the saved layer inputs, then back-propagates:
//   g_i      = d_out_i / deg_i                       d_bias   += d_out_i
//   d_alpha  = g_i . xl_j                            d_xl_j   += alpha_ij g_i        (smem atomics)
//   d_s      = alpha_ij (d_alpha_ij - sum_j' alpha_ij' d_alpha_ij')
//   d_u      = d_s * att * lrelu'(xl_j + xr_i)       d_att    += d_s * lrelu(u)
//   d_xl_j  += d_u ; d_xr_i += d_u
//   d_W{l,r} += d_x{l,r}_i (x) h_in_i ; d_b{l,r} += d_x{l,r}_i ; d_h_in_i = Wl^T d_xl_i + Wr^T d_xr_i
// Weight gradients are accumulated per CTA in shared memory over all its chunks and applications,
// written to partials[blockIdx.x][*] and summed in a fixed order by gnn_bwd_reduce_kernel.
template <int HP>
__global__ void __launch_bounds__(kGnnThreads, 1) gnn_bwd_kernel(const GnnArgs p) {
  using L = ConvLayout<HP>;
  extern __shared__ float4 smem4[];
  float* wsm = reinterpret_cast<float*>(smem4);             // weights   [num_convs][kFloats]
  float* gsm = wsm + p.num_convs * L::kFloats;              // gradients [num_convs][kFloats]
  float* xl_s = gsm + p.num_convs * L::kFloats;             // [cap][HP] xl, later d_h_in scratch
  float* xr_s = xl_s + (size_t)p.cap * HP;                  // [cap][HP] xr, later d_xr
  float* dxl_s = xr_s + (size_t)p.cap * HP;                 // [cap][HP] d_xl (atomic target)
  float* dout_s = dxl_s + (size_t)p.cap * HP;               // [cap][HP] d_out of the current app
  float* hin_s = dout_s + (size_t)p.cap * HP;               // [cap][HP] input of the current app

  for (int cv = 0; cv < p.num_convs; ++cv)
    load_conv_weights<HP>(wsm + cv * L::kFloats, p.params + conv_param_offset(cv, p.in_dim, p.H),
                          cv == 0 ? p.in_dim : p.H, p.H);
  for (int t = threadIdx.x; t < p.num_convs * L::kFloats; t += kGnnThreads) gsm[t] = 0.f;
  __syncthreads();

  const int num_chunks = (p.B + p.G - 1) / p.G;
  for (int chunk = blockIdx.x; chunk < num_chunks; chunk += gridDim.x) {
    const int g0 = chunk * p.G, g1 = min(p.B, g0 + p.G);
    const int nbase = p.graph_ptr[g0];
    const int nc = p.graph_ptr[g1] - nbase;

    // d_out of the last application = pool backward: d_emb[g] / n_g (times ReLU' if it had one).
    for (int t = threadIdx.x; t < (g1 - g0) * HP; t += kGnnThreads) {
      const int g = g0 + t / HP, c = t % HP;
      const int n0 = p.graph_ptr[g] - nbase, n1 = p.graph_ptr[g + 1] - nbase;
      const float v = (c < p.H && n1 > n0) ? p.d_emb[(size_t)g * p.emb_ld + c] / (float)(n1 - n0) : 0.f;
      for (int i = n0; i < n1; ++i) dout_s[(size_t)i * HP + c] = v;
    }
    __syncthreads();

    for (int a = p.num_apps - 1; a >= 0; --a) {
      const int cv = p.app_conv[a];
      const float* w = wsm + cv * L::kFloats;
      float* gw = gsm + cv * L::kFloats;
      const int d_in = (cv == 0) ? p.in_dim : p.H;

      // ---- recompute phase A; apply ReLU' of this application to d_out; clear d_xl ----
      for (int i = threadIdx.x; i < nc; i += kGnnThreads) {
        float hin[HP], xl[HP], xr[HP];
        const int64_t node = (int64_t)nbase + i;
        if (a == 0) {
#pragma unroll
          for (int k = 0; k < HP; ++k) hin[k] = (k < p.in_dim) ? p.x[node * p.in_dim + k] : 0.f;
        } else {
          load_row<HP>(p.saved + ((size_t)(a - 1) * p.Nt + node) * HP, hin);
        }
        node_transform<HP>(w, hin, d_in, xl, xr);
        store_row<HP>(xl_s + (size_t)i * HP, xl);
        store_row<HP>(xr_s + (size_t)i * HP, xr);
        store_row<HP>(hin_s + (size_t)i * HP, hin);
        if (p.app_relu[a]) {
          float out[HP], d[HP];
          load_row<HP>(p.saved + ((size_t)a * p.Nt + node) * HP, out);
          load_row<HP>(dout_s + (size_t)i * HP, d);
#pragma unroll
          for (int c = 0; c < HP; ++c) d[c] = out[c] > 0.f ? d[c] : 0.f;
          store_row<HP>(dout_s + (size_t)i * HP, d);
        }
#pragma unroll
        for (int c = 0; c < HP; ++c) hin[c] = 0.f;
        store_row<HP>(dxl_s + (size_t)i * HP, hin);
      }
      __syncthreads();

      // ---- phase B backward: one thread per target node ----
      float datt[HP], dbias[HP];
#pragma unroll
      for (int c = 0; c < HP; ++c) { datt[c] = 0.f; dbias[c] = 0.f; }
      for (int i = threadIdx.x; i < nc; i += kGnnThreads) {
        float xr[HP], att[HP], g[HP], dxr[HP];
        load_row<HP>(xr_s + (size_t)i * HP, xr);
        load_row<HP>(w + L::kAtt, att);
        load_row<HP>(dout_s + (size_t)i * HP, g);
        const int e0 = p.rowptr[nbase + i], e1 = p.rowptr[nbase + i + 1];
        const float inv_deg = 1.f / (float)(e1 - e0);
#pragma unroll
        for (int c = 0; c < HP; ++c) { dbias[c] += g[c]; g[c] *= inv_deg; dxr[c] = 0.f; }
        // pass 1: softmax statistics and sum_j alpha_ij d_alpha_ij
        float m = -INFINITY, l = 0.f, dot = 0.f;
        for (int e = e0; e < e1; ++e) {
          const int j = p.col[e] - nbase;
          float v[HP];
          load_row<HP>(xl_s + (size_t)j * HP, v);
          float s = 0.f, da = 0.f;
#pragma unroll
          for (int c = 0; c < HP; ++c) {
            float t = v[c] + xr[c];
            t = fmaxf(t, kNegSlope * t);
            s = fmaf(att[c], t, s);
            da = fmaf(g[c], v[c], da);
          }
          const float m_new = fmaxf(m, s);
          const float sc = expf(m - m_new);
          const float pe = expf(s - m_new);
          l = fmaf(l, sc, pe);
          dot = fmaf(dot, sc, pe * da);
          m = m_new;
        }
        const float inv_l = 1.f / (l + 1e-16f);
        dot *= inv_l;
        // pass 2: per-edge gradients
        for (int e = e0; e < e1; ++e) {
          const int j = p.col[e] - nbase;
          float v[HP];
          load_row<HP>(xl_s + (size_t)j * HP, v);
          float s = 0.f, da = 0.f;
#pragma unroll
          for (int c = 0; c < HP; ++c) {
            float t = v[c] + xr[c];
            t = fmaxf(t, kNegSlope * t);
            s = fmaf(att[c], t, s);
            da = fmaf(g[c], v[c], da);
          }
          const float alpha = expf(s - m) * inv_l;
          const float ds = alpha * (da - dot);
          float* dj = dxl_s + (size_t)j * HP;
#pragma unroll
          for (int c = 0; c < HP; ++c) {
            const float u = v[c] + xr[c];
            const float lu = fmaxf(u, kNegSlope * u);
            datt[c] = fmaf(ds, lu, datt[c]);
            const float du = ds * att[c] * (u > 0.f ? 1.f : kNegSlope);
            dxr[c] += du;
            atomicAdd(dj + c, fmaf(alpha, g[c], du));
          }
        }
        store_row<HP>(xr_s + (size_t)i * HP, dxr);  // own row: xr no longer needed
      }
      // block-reduce d_att / d_bias into the CTA gradient accumulator
#pragma unroll
      for (int c = 0; c < HP; ++c) {
        const float sa = warp_sum(datt[c]), sb = warp_sum(dbias[c]);
        if ((threadIdx.x & 31) == 0) {
          atomicAdd(gw + L::kAtt + c, sa);
          atomicAdd(gw + L::kBias + c, sb);
        }
      }
      __syncthreads();

      // ---- transform backward ----
      // (1) weight gradients: thread t owns pairs (k, c) of Wl and Wr; sums over the chunk's nodes.
      for (int t = threadIdx.x; t < d_in * HP; t += kGnnThreads) {
        const int k = t / HP, c = t - k * HP;
        float sl = 0.f, sr = 0.f;
        for (int i = 0; i < nc; ++i) {
          const float h = hin_s[(size_t)i * HP + k];
          sl = fmaf(dxl_s[(size_t)i * HP + c], h, sl);
          sr = fmaf(xr_s[(size_t)i * HP + c], h, sr);
        }
        gw[L::kWl + k * HP + c] += sl;
        gw[L::kWr + k * HP + c] += sr;
      }
      for (int t = threadIdx.x; t < 2 * HP; t += kGnnThreads) {
        const int c = t % HP;
        const float* src = (t < HP) ? dxl_s : xr_s;
        float s = 0.f;
        for (int i = 0; i < nc; ++i) s += src[(size_t)i * HP + c];
        gw[(t < HP ? L::kBl : L::kBr) + c] += s;
      }
      // (2) d_h_in_i = Wl^T d_xl_i + Wr^T d_xr_i  -> becomes d_out of application a-1
      if (a > 0) {
        for (int i = threadIdx.x; i < nc; i += kGnnThreads) {
          float dl[HP], dr[HP], dh[HP];
          load_row<HP>(dxl_s + (size_t)i * HP, dl);
          load_row<HP>(xr_s + (size_t)i * HP, dr);
#pragma unroll
          for (int k = 0; k < HP; ++k) {
            float s = 0.f;
            if (k < d_in) {
              const float4* wl = reinterpret_cast<const float4*>(w + L::kWl + k * HP);
              const float4* wr = reinterpret_cast<const float4*>(w + L::kWr + k * HP);
#pragma unroll
              for (int c4 = 0; c4 < HP / 4; ++c4) {
                const float4 a4 = wl[c4], b4 = wr[c4];
                s = fmaf(a4.x, dl[4 * c4], s); s = fmaf(a4.y, dl[4 * c4 + 1], s);
                s = fmaf(a4.z, dl[4 * c4 + 2], s); s = fmaf(a4.w, dl[4 * c4 + 3], s);
                s = fmaf(b4.x, dr[4 * c4], s); s = fmaf(b4.y, dr[4 * c4 + 1], s);
                s = fmaf(b4.z, dr[4 * c4 + 2], s); s = fmaf(b4.w, dr[4 * c4 + 3], s);
              }
            }
            dh[k] = s;
          }
          store_row<HP>(dout_s + (size_t)i * HP, dh);
        }
      }
      __syncthreads();
    }
  }

  // CTA partial -> workspace, in the flat state-dict layout of `params`.
  const int64_t P = conv_param_offset(p.num_convs, p.in_dim, p.H);
  float* out = p.partials + (size_t)blockIdx.x * P;
  for (int cv = 0; cv < p.num_convs; ++cv) {
    const int in = (cv == 0) ? p.in_dim : p.H;
    const int H = p.H;
    const float* gw = gsm + cv * L::kFloats;
    float* o = out + conv_param_offset(cv, p.in_dim, p.H);
    for (int t = threadIdx.x; t < H; t += kGnnThreads) {
      o[t] = gw[L::kAtt + t];
      o[H + t] = gw[L::kBias + t];
      o[2 * H + (int64_t)H * in + t] = gw[L::kBl + t];
      o[3 * H + 2 * (int64_t)H * in + t] = gw[L::kBr + t];
    }
    for (int t = threadIdx.x; t < H * in; t += kGnnThreads) {
      const int c = t / in, k = t - c * in;
      o[2 * H + t] = gw[L::kWl + k * HP + c];
      o[3 * H + (int64_t)H * in + t] = gw[L::kWr + k * HP + c];
    }
  }
}

__global__ void gnn_bwd_reduce_kernel(const float* __restrict__ partials, int num_partials, int64_t P,
                                      float* __restrict__ d_params) {
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < P;
       t += (int64_t)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int b = 0; b < num_partials; ++b) s += partials[(size_t)b * P + t];
    d_params[t] = s;
  }
}

// ------------------------------- host side ----------------------------------------------------
static int pick_hp(int H) { return H <= 8 ? 8 : H <= 16 ? 16 : H <= 24 ? 24 : 32; }

static int build_args(GnnArgs& a, const float* x, int64_t Nt, int in_dim, const int32_t* rowptr,
                      const int32_t* col, const int32_t* graph_ptr, int64_t B, int max_nodes,
                      const float* params, int H, int num_layers, int num_iter) {
  GSC_REQUIRE(H >= 1 && H <= GSC_MAX_HIDDEN, "gnn: hidden=%d outside 1..%d", H, GSC_MAX_HIDDEN);
  GSC_REQUIRE(in_dim >= 1 && in_dim <= pick_hp(H), "gnn: in_dim=%d must be in 1..%d", in_dim, pick_hp(H));
  GSC_REQUIRE(num_layers >= 1 && (num_layers == 1 || num_iter >= 1), "gnn: bad num_layers/num_iter");
  GSC_REQUIRE(B >= 1 && B < INT32_MAX && Nt >= 0, "gnn: bad sizes");
  GSC_REQUIRE(x && rowptr && col && graph_ptr && params, "gnn: null pointer");
  GSC_REQUIRE(max_nodes >= 1, "gnn: max_nodes must be >= 1");
  a.x = x; a.rowptr = rowptr; a.col = col; a.graph_ptr = graph_ptr; a.params = params;
  a.Nt = Nt; a.B = (int)B; a.in_dim = in_dim; a.H = H; a.num_convs = num_layers;
  int n = 0;
  a.app_conv[n] = 0; a.app_relu[n] = 1; ++n;           // encoder + ReLU (models.py:38-40)
  if (num_layers > 1) {
    for (int it = 0; it < num_iter; ++it)
      for (int i = 0; i < num_layers - 1; ++i) {
        GSC_REQUIRE(n < GSC_MAX_APPS, "gnn: more than %d GATv2 applications", GSC_MAX_APPS);
        const bool last = (i == num_layers - 2) && (it == num_iter - 1);   // models.py:50-53
        a.app_conv[n] = i + 1; a.app_relu[n] = last ? 0 : 1; ++n;
      }
  }
  a.num_apps = n;
  a.cap = max_nodes <= kGnnThreads ? kGnnThreads : (max_nodes + 31) / 32 * 32;
  a.G = a.cap / max_nodes;
  if (a.G < 1) a.G = 1;
  return GSC_OK;
}

template <int HP>
static int launch_fwd(const GnnArgs& a, cudaStream_t st) {
  const size_t smem = ((size_t)a.num_convs * ConvLayout<HP>::kFloats + 2 * (size_t)a.cap * HP) * 4;
  GSC_REQUIRE(smem <= 227 * 1024, "gnn fwd: max_nodes=%d needs %zu B of shared memory (> 227 KB)", a.cap, smem);
  GSC_CUDA_TRY(cudaFuncSetAttribute(gnn_fwd_kernel<HP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int chunks = (a.B + a.G - 1) / a.G;
  const int grid = chunks < 2 * kNumSMs ? chunks : 2 * kNumSMs;
  gnn_fwd_kernel<HP><<<grid, kGnnThreads, smem, st>>>(a);
  GSC_LAUNCH_CHECK("gnn_fwd_kernel");
  return GSC_OK;
}

// tcgen05 forward: usable when the TMEM accumulators (2*HP columns per 128 nodes) and the operand
// tiles of the largest graph fit; otherwise the FMA kernel above runs.
template <int HP>
static int launch_fwd_tc(GnnArgs a, int max_nodes, cudaStream_t st, bool* done) {
  using T = TcL<HP>;
  *done = false;
  const int cap = max_nodes <= kGnnThreads ? kGnnThreads : (max_nodes + 127) / 128 * 128;
  const int tiles = cap / 128;
  const size_t smem = (size_t)a.num_convs * T::kConvBytes + 2 * (size_t)tiles * T::kATile + (size_t)cap * HP * 4 + 16;
  int cols = 32;
  while (cols < tiles * T::kN) cols *= 2;
  if (cols > 512 || smem > 200 * 1024) return GSC_OK;
  a.cap = cap;
  a.G = cap / max_nodes;
  if (a.G < 1) a.G = 1;
  GSC_CUDA_TRY(cudaFuncSetAttribute(gnn_fwd_tc_kernel<HP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int chunks = (a.B + a.G - 1) / a.G;
  const int per_sm = (2 * smem <= 220 * 1024 && 2 * cols <= 512) ? 2 : 1;
  const int grid = chunks < per_sm * kNumSMs ? chunks : per_sm * kNumSMs;
  gnn_fwd_tc_kernel<HP><<<grid, kGnnThreads, smem, st>>>(a, cols);
  GSC_LAUNCH_CHECK("gnn_fwd_tc_kernel");
  *done = true;
  return GSC_OK;
}

static bool gnn_use_tc() {
  static int mode = -1;
  if (mode < 0) {
    const char* e = getenv("GSC_GNN_FMA");
    mode = (e && e[0] == '1') ? 0 : 1;
  }
  return mode == 1;
}

static int bwd_grid(int B, int G) {
  const int chunks = (B + G - 1) / G;
  return chunks < kNumSMs ? chunks : kNumSMs;
}

template <int HP>
static int launch_bwd(const GnnArgs& a, float* d_params, cudaStream_t st) {
  const size_t smem = (2 * (size_t)a.num_convs * ConvLayout<HP>::kFloats + 5 * (size_t)a.cap * HP) * 4;
  GSC_REQUIRE(smem <= 227 * 1024, "gnn bwd: max_nodes=%d needs %zu B of shared memory (> 227 KB)", a.cap, smem);
  GSC_CUDA_TRY(cudaFuncSetAttribute(gnn_bwd_kernel<HP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int grid = bwd_grid(a.B, a.G);
  gnn_bwd_kernel<HP><<<grid, kGnnThreads, smem, st>>>(a);
  GSC_LAUNCH_CHECK("gnn_bwd_kernel");
  const int64_t P = conv_param_offset(a.num_convs, a.in_dim, a.H);
  gnn_bwd_reduce_kernel<<<(int)ceil_div(P, 256), 256, 0, st>>>(a.partials, grid, P, d_params);
  GSC_LAUNCH_CHECK("gnn_bwd_reduce_kernel");
  return GSC_OK;
}

}  // namespace gsc

using namespace gsc;

static float* g_gnn_debug = nullptr;
// Debug hook (not part of the public header): device buffer of >= 8 floats that block 0 of the
// tcgen05 forward fills with per-phase cycle counts; nullptr disables it.
extern "C" void gsc_debug_set_gnn_timers(float* dev_buf) { g_gnn_debug = dev_buf; }

extern "C" int64_t gsc_gnn_param_floats(int in_dim, int hidden, int num_layers) {
  return conv_param_offset(num_layers, in_dim, hidden);
}

extern "C" int64_t gsc_gnn_saved_floats(int64_t num_nodes, int hidden, int num_layers, int num_iter) {
  const int apps = 1 + (num_layers > 1 ? num_iter * (num_layers - 1) : 0);
  return (int64_t)apps * num_nodes * pick_hp(hidden);
}

extern "C" int gsc_gnn_embed_fwd(const float* x, int64_t Nt, int in_dim, const int32_t* rowptr,
                                 const int32_t* col, const int32_t* graph_ptr, int64_t B, int max_nodes,
                                 const float* params, int H, int num_layers, int num_iter, float* emb,
                                 int64_t emb_ld, float* saved, void* stream) {
  GnnArgs a{};
  int rc = build_args(a, x, Nt, in_dim, rowptr, col, graph_ptr, B, max_nodes, params, H, num_layers, num_iter);
  if (rc) return rc;
  GSC_REQUIRE(emb && emb_ld >= H, "gnn fwd: null emb or emb_ld < hidden");
  a.emb = emb; a.saved = saved; a.emb_ld = emb_ld;
  a.partials = g_gnn_debug;
  cudaStream_t st = (cudaStream_t)stream;
  if (gnn_use_tc()) {
    bool done = false;
    switch (pick_hp(H)) {
      case 8: rc = launch_fwd_tc<8>(a, max_nodes, st, &done); break;
      case 16: rc = launch_fwd_tc<16>(a, max_nodes, st, &done); break;
      case 24: rc = launch_fwd_tc<24>(a, max_nodes, st, &done); break;
      default: rc = launch_fwd_tc<32>(a, max_nodes, st, &done); break;
    }
    if (rc || done) return rc;
  }
  switch (pick_hp(H)) {
    case 8: return launch_fwd<8>(a, st);
    case 16: return launch_fwd<16>(a, st);
    case 24: return launch_fwd<24>(a, st);
    default: return launch_fwd<32>(a, st);
  }
}

extern "C" size_t gsc_gnn_bwd_workspace_bytes(int64_t B, int in_dim, int H, int num_layers) {
  (void)B;
  return (size_t)kNumSMs * (size_t)conv_param_offset(num_layers, in_dim, H) * 4;
}

extern "C" int gsc_gnn_embed_bwd(const float* x, int64_t Nt, int in_dim, const int32_t* rowptr,
                                 const int32_t* col, const int32_t* graph_ptr, int64_t B, int max_nodes,
                                 const float* params, int H, int num_layers, int num_iter,
                                 const float* saved, const float* d_emb, int64_t d_emb_ld, float* d_params,
                                 void* workspace, size_t workspace_bytes, void* stream) {
  GnnArgs a{};
  int rc = build_args(a, x, Nt, in_dim, rowptr, col, graph_ptr, B, max_nodes, params, H, num_layers, num_iter);
  if (rc) return rc;
  GSC_REQUIRE(saved && d_emb && d_params && workspace && d_emb_ld >= H, "gnn bwd: null pointer or d_emb_ld < hidden");
  a.emb_ld = d_emb_ld;
  if (workspace_bytes < gsc_gnn_bwd_workspace_bytes(B, in_dim, H, num_layers))
    return fail(GSC_EWORKSPACE, "gnn bwd: workspace %zu < %zu", workspace_bytes,
                gsc_gnn_bwd_workspace_bytes(B, in_dim, H, num_layers));
  a.saved = const_cast<float*>(saved); a.d_emb = d_emb; a.partials = (float*)workspace;
  cudaStream_t st = (cudaStream_t)stream;
  switch (pick_hp(H)) {
    case 8: return launch_bwd<8>(a, d_params, st);
    case 16: return launch_bwd<16>(a, d_params, st);
    case 24: return launch_bwd<24>(a, d_params, st);
    default: return launch_bwd<32>(a, d_params, st);
  }
}
