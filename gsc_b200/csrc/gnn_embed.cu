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
constexpr int kColPerNode = 8;     // CSR entries staged in shared memory per node slot (else read from global)
constexpr int kBwdEdgesPerNode = 7;  // gather backward: per-edge arrays (score, gradient, reverse index) per node slot
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
  float* rows_ws;         // bwd, large graphs only: [gridDim.x][5][cap][HP] node-row scratch in global memory
  float* dx;              // bwd, tcgen05 kernel: [num_apps][Nt][2*HP] = d_xl | d_xr of every node, read by gnn_wgrad_kernel
  float* dbg;             // optional phase timers (debug hook), nullptr in production
  const int32_t* stats;   // bwd: packer stats (stats[3] == 0 <=> symmetric graph), nullable
  const int32_t* flags;   // bwd: flags[0] != 0 <=> some chunk's edges exceed the gather kernel's shared memory
  int gather_kernel_launched;   // bwd: the out-edge gather kernel handles symmetric graphs
  const uint16_t* prep;         // bwd gather kernel: per chunk [perm | rev] (gnn_chunk_prep_kernel)
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
    if (nc > p.cap) __trap();     // a graph larger than the max_nodes the launch was sized for: fail loudly

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
// leaky_relu(t, 0.2) = 0.6 t + 0.4 |t|, so the GATv2 score of edge j -> i splits into
//   s_ij = att . lrelu(xl_j + xr_i) = 0.6 att.xl_j + 0.6 att.xr_i + sum_c (0.4 att_c) |xl_j,c + xr_i,c|
// with the first two terms per-node scalars: two instructions per channel and edge instead of four.
// Sums run in four independent accumulators (the kernels have few warps to hide FMA latency).
template <int HP>
__device__ __forceinline__ float dot4(const float (&a)[HP], const float (&b)[HP]) {
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
  for (int c = 0; c < HP; c += 4) {
    s0 = fmaf(a[c], b[c], s0);
    s1 = fmaf(a[c + 1], b[c + 1], s1);
    s2 = fmaf(a[c + 2], b[c + 2], s2);
    s3 = fmaf(a[c + 3], b[c + 3], s3);
  }
  return (s0 + s1) + (s2 + s3);
}
template <int HP>
__device__ __forceinline__ float abs_score(const float (&v)[HP], const float (&xr)[HP], const float (&b)[HP]) {
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
  for (int c = 0; c < HP; c += 4) {
    s0 = fmaf(b[c], fabsf(v[c] + xr[c]), s0);
    s1 = fmaf(b[c + 1], fabsf(v[c + 1] + xr[c + 1]), s1);
    s2 = fmaf(b[c + 2], fabsf(v[c + 2] + xr[c + 2]), s2);
    s3 = fmaf(b[c + 3], fabsf(v[c + 3] + xr[c + 3]), s3);
  }
  return (s0 + s1) + (s2 + s3);
}

template <int HP>
struct TcL {
  static constexpr int kN = 2 * HP;                       // accumulator columns: xl | xr
  static constexpr int kChunks = HP / 4;                  // 16-byte chunks per operand row
  static constexpr int kGroupBytes = kChunks * 128;       // 8 rows x kChunks core matrices
  static constexpr int kATile = 16 * kGroupBytes;         // 128-row A operand (== 128*HP*4 bytes)
  static constexpr int kB = (kN / 8) * kGroupBytes;       // kN-row B operand
  static constexpr int kVecFloats = 4 * HP;               // bl, br, att, bias
  static constexpr int kConvBytes = 2 * kB + kVecFloats * 4;
  static constexpr int kXS = HP + 4;                      // xl row stride: 16 B-aligned, bank-conflict free;
                                                          // slot [HP] holds 0.6 att.xl_j
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
  int* rp_s = reinterpret_cast<int*>(xl_s + (size_t)p.cap * T::kXS);          // [cap + 4] chunk-local CSR row offsets
  int* col_s = rp_s + p.cap + 4;                                              // [kColPerNode * cap] chunk-local sources
  uint64_t* bar_ptr = reinterpret_cast<uint64_t*>(col_s + kColPerNode * p.cap);
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
  const bool dbg = p.dbg != nullptr && blockIdx.x == 0 && tid == 0;
#define GSC_TICK(slot)                                                                  \
  do {                                                                                  \
    if (dbg) { long long tn = clock64(); p.dbg[slot] += (float)(tn - tprev); tprev = tn; } \
  } while (0)
  GSC_TICK(0);

  const int num_chunks = (p.B + p.G - 1) / p.G;
  for (int chunk = blockIdx.x; chunk < num_chunks; chunk += gridDim.x) {
    const int g0 = chunk * p.G, g1 = min(p.B, g0 + p.G);
    const int nbase = p.graph_ptr[g0];
    const int nc = p.graph_ptr[g1] - nbase;
    if (nc > p.cap) __trap();     // a graph larger than the max_nodes the launch was sized for: fail loudly
    const int warp_first = tid & ~31;                                // slot loops must be warp-uniform (tcgen05.ld)
    // the chunk's CSR rows are walked by every application: stage them in shared memory once
    const int ebase = p.rowptr[nbase];
    const int ne = p.rowptr[nbase + nc] - ebase;
    const bool staged = ne <= kColPerNode * p.cap;
    for (int i = tid; i <= nc; i += kGnnThreads) rp_s[i] = p.rowptr[nbase + i] - ebase;
    if (staged)
      for (int e = tid; e < ne; e += kGnnThreads) col_s[e] = p.col[ebase + e] - nbase;
    const int* colp = staged ? col_s : p.col + ebase;
    const int coff = staged ? 0 : nbase;

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
          float att[HP];
          load_row<HP>(vec + 2 * HP, att);
#pragma unroll
          for (int c = 0; c < HP; ++c) xl[c] += vec[c];
          store_row<HP>(xl_s + (size_t)i * T::kXS, xl);
          xl_s[(size_t)i * T::kXS + HP] = 0.6f * dot4<HP>(att, xl);
        }
      }
      __syncthreads();
      GSC_TICK(3);

      // ---- phase B: attention scores, online segment softmax, mean aggregation ----
      for (int i0 = warp_first; i0 < nc; i0 += kGnnThreads) {
        const int i = i0 + (tid & 31);
        float xr[HP], b[HP], acc[HP];
        tmem_load_row<HP>(tmem_base + lane_base + (uint32_t)((i >> 7) * T::kN + HP), xr);
        if (i < nc) {
          load_row<HP>(vec + 2 * HP, b);
#pragma unroll
          for (int c = 0; c < HP; ++c) { xr[c] += vec[HP + c]; acc[c] = 0.f; }
          const float q_i = 0.6f * dot4<HP>(b, xr);
#pragma unroll
          for (int c = 0; c < HP; ++c) b[c] *= 0.4f;
          float m = -INFINITY, l = 0.f;
          const int e0 = rp_s[i], e1 = rp_s[i + 1];
          for (int e = e0; e < e1; ++e) {
            const int j = colp[e] - coff;
            float v[HP];
            load_row<HP>(xl_s + (size_t)j * T::kXS, v);
            const float s = xl_s[(size_t)j * T::kXS + HP] + q_i + abs_score<HP>(v, xr, b);
            const float m_new = fmaxf(m, s);
            const float sc = __expf(m - m_new);
            const float pe = __expf(s - m_new);
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
    // few large graphs: eight lanes per (graph, channel), each adds every 8th node row, then a fixed
    // shuffle tree; many small graphs: one thread per (graph, channel)
    const int pool_items = (g1 - g0) * HP;
    const int parts = pool_items * 8 <= 2 * kGnnThreads ? 8 : 1;
    const int pshift = parts == 8 ? 3 : 0;
    for (int base = 0; base < pool_items * parts; base += kGnnThreads) {
      const int pr = (base + tid) >> pshift, part = tid & (parts - 1);
      const bool ok = pr < pool_items;
      const int g = g0 + (ok ? pr / HP : 0), c = pr % HP;
      const int n0 = p.graph_ptr[g] - nbase, n1 = p.graph_ptr[g + 1] - nbase;
      float s = 0.f;
      if (ok)
        for (int i = n0 + part; i < n1; i += parts) s += hx[(size_t)i * HP + c];
      if (parts == 8) {
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 4);
      }
      if (ok && part == 0 && c < p.H) p.emb[(size_t)g * p.emb_ld + c] = (n1 > n0) ? s / (float)(n1 - n0) : 0.f;
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
// kGlobalRows: graphs whose five node-row arrays do not fit in shared memory (e.g. 1000-node
// topologies) keep them in a per-CTA slice of the workspace instead (L2-resident; the scatter then
// uses native global float reductions).  Same code, same summation order per node.
template <int HP, bool kGlobalRows>
__global__ void __launch_bounds__(kGnnThreads, 1) gnn_bwd_kernel(const GnnArgs p) {
  using L = ConvLayout<HP>;
  extern __shared__ float4 smem4[];
  float* wsm = reinterpret_cast<float*>(smem4);             // weights   [num_convs][kFloats]
  float* gsm = wsm + p.num_convs * L::kFloats;              // gradients [num_convs][kFloats]
  float* xl_s = kGlobalRows ? p.rows_ws + (size_t)blockIdx.x * 5 * p.cap * HP
                            : gsm + p.num_convs * L::kFloats;   // [cap][HP] xl, later d_h_in scratch
  float* xr_s = xl_s + (size_t)p.cap * HP;                  // [cap][HP] xr, later d_xr
  float* dxl_s = xr_s + (size_t)p.cap * HP;                 // [cap][HP] d_xl (atomic target)
  float* dout_s = dxl_s + (size_t)p.cap * HP;               // [cap][HP] d_out of the current app
  float* hin_s = dout_s + (size_t)p.cap * HP;               // [cap][HP] input of the current app
  // symmetric batches are handled by gnn_bwd_tc_kernel (out-edge gather, no atomics)
  if (p.gather_kernel_launched && p.stats != nullptr && p.stats[3] == 0 && p.flags[0] == 0) return;

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
    if (nc > p.cap) __trap();     // a graph larger than the max_nodes the launch was sized for: fail loudly

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

// ================================ tcgen05 backward ==========================================
// Backward with all three GEMM-shaped pieces on the tensor cores (3xTF32, fp32-faithful):
//   (R) recompute   [xl|xr]      = h_in  x [Wl;Wr]^T          M=128 nodes, N=2HP,   K=d_in
//   (H) input grad  d_h_in       = [dxl|dxr] x [Wl;Wr]        M=128 nodes, N=HP',   K=2HP
// d_h_in stays in TMEM (lane = node) until the next application's edge phase consumes it.  The
// edge phase (softmax backward, d_att, scatter of d_xl with shared-memory atomics) is one thread
// per node.  The weight gradient
//   (W) [dWl|dbl ; dWr|dbr] += [dxl|dxr]^T x [h_in|1]       (reduction over nodes)
// contracts over the node index, i.e. both operands would be MN-major; for 32-bit operands the
// tensor core only takes that through the 32-byte-atom swizzled image, a second copy of both
// operands that does not fit next to the K-major ones.  It is not part of this kernel: the
// [dxl|dxr] rows of every application are written to the workspace (they are in registers anyway
// when the (H) operand is built) and gnn_wgrad_kernel reduces them over the nodes afterwards --
// a streaming pass with a 4 x 8 register tile per thread instead of accumulators that this
// kernel would have to carry (and spill) across all of its phases.
template <int HP>
struct TcB {
  static constexpr int kN1 = 2 * HP;                               // (R) columns: xl | xr
  static constexpr int kNH = (HP + 15) / 16 * 16;                  // (H) columns (N % 16 == 0 for M = 128)
  static constexpr int kHChunks = HP / 4;                          // h row: HP floats
  static constexpr int kHGroup = kHChunks * 128;
  static constexpr int kHTile = 16 * kHGroup;
  static constexpr int kDChunks = 2 * HP / 4;                      // dx row: dxl | dxr
  static constexpr int kDGroup = kDChunks * 128;
  static constexpr int kDTile = 16 * kDGroup;
  static constexpr int kB1Group = (HP / 4) * 128;                  // B1: 2HP rows x HP k
  static constexpr int kB1 = (kN1 / 8) * kB1Group;
  static constexpr int kB2Group = kDChunks * 128;                  // B2: kNH rows x 2HP k
  static constexpr int kB2 = (kNH / 8) * kB2Group;
  static constexpr int kVecFloats = 4 * HP;                        // bl, br, att, pad
  static constexpr int kSlotFloats = (kGnnThreads / 32) * 2 * HP;  // per-warp private d_att | d_bias sums
  static constexpr int kConvBytes = 2 * kB1 + 2 * kB2 + (kVecFloats + kSlotFloats) * 4;
  static constexpr int kMaxConvs = 4;
  __device__ __forceinline__ static int hoff(int r, int c) { return (r >> 3) * kHGroup + c * 128 + (r & 7) * 16; }
  __device__ __forceinline__ static int doff(int r, int c) { return (r >> 3) * kDGroup + c * 128 + (r & 7) * 16; }
};

// ds times the LeakyReLU branch sign of u: +ds for u > 0, -ds for u < 0 (one LOP3; u = +0 counts as
// positive, which only the zero-padded channels ever hit, and those are multiplied by zero later).
__device__ __forceinline__ float signed_by(float ds, float u) {
  return __int_as_float(__float_as_int(ds) ^ (__float_as_int(u) & (int)0x80000000));
}

__device__ __forceinline__ void split4(const float4 v, float4& hi, float4& lo) {
  hi.x = to_tf32(v.x); lo.x = to_tf32(v.x - hi.x);
  hi.y = to_tf32(v.y); lo.y = to_tf32(v.y - hi.y);
  hi.z = to_tf32(v.z); lo.z = to_tf32(v.z - hi.z);
  hi.w = to_tf32(v.w); lo.w = to_tf32(v.w - hi.w);
}

template <int HP>
__global__ void __launch_bounds__(kGnnThreads, 1) gnn_bwd_tc_kernel(const GnnArgs p, const int tmem_cols) {
  using T = TcB<HP>;
  extern __shared__ float4 smem4[];
  uint8_t* sm = reinterpret_cast<uint8_t*>(smem4);
  const int tiles = p.cap / 128;
  uint8_t* convs = sm;
  uint8_t* h_hi = convs + p.num_convs * T::kConvBytes;
  uint8_t* h_lo = h_hi + tiles * T::kHTile;
  uint8_t* d_hi = h_lo + tiles * T::kHTile;
  uint8_t* d_lo = d_hi + tiles * T::kDTile;
  uint8_t* tail = d_lo + tiles * T::kDTile;
  // Needs every edge's reverse in the same CSR and every chunk's edges in shared memory: other
  // batches go to gnn_bwd_kernel instead (p.flags[0] is set by gnn_chunk_edges_kernel).
  if (p.stats == nullptr || p.stats[3] != 0 || p.flags[0] != 0) return;
  // scratch that lives inside the (idle) dx operand buffers during the edge phase
  float* xl_s = reinterpret_cast<float*>(d_lo);                               // [cap][HP] xl, later d_xl
  float* g_s = reinterpret_cast<float*>(d_lo) + (size_t)p.cap * HP;           // [cap][HP] d_out / deg
  float* xr_s = reinterpret_cast<float*>(d_hi);                               // [cap][HP]
  float* dxr_raw = reinterpret_cast<float*>(d_hi) + (size_t)p.cap * HP;       // [cap][HP]
  const int emax = kBwdEdgesPerNode * p.cap;
  int* gid_s = reinterpret_cast<int*>(tail);                                  // [cap] graph of every node
  float2* pq_s = reinterpret_cast<float2*>(gid_s + p.cap);                    // [cap] 0.6 att.xl_j, 0.6 att.xr_i
  int* rp_s = reinterpret_cast<int*>(pq_s + p.cap);                           // [cap + 4] chunk-local CSR row offsets
  float* es_s = reinterpret_cast<float*>(rp_s + p.cap + 4);                   // [emax] per edge: score, then alpha
  float* eda_s = es_s + emax;                                                 // [emax] per edge: g_i.xl_j, then d_score
  uint16_t* col_s = reinterpret_cast<uint16_t*>(eda_s + emax);                // [emax] chunk-local source of every edge
  uint16_t* rev_s = col_s + emax;                                             // [emax] position of the reverse edge
  uint16_t* perm_s = rev_s + emax;                                            // [cap] nodes by descending degree
  int* hist_s = reinterpret_cast<int*>(perm_s + p.cap);                       // [64] degree histogram / cursors
  uint64_t* bar_ptr = reinterpret_cast<uint64_t*>(hist_s + 64);
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar_ptr + 1);
  const uint32_t bar = smem_u32(bar_ptr);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  // ---- weights: B1 = [Wl;Wr] rows (K = d_in), B2 = [Wl;Wr]^T rows (K = 2HP), split hi/lo ----
  for (int cv = 0; cv < p.num_convs; ++cv) {
    const int in = cv == 0 ? p.in_dim : p.H, H = p.H;
    const float* src = p.params + conv_param_offset(cv, p.in_dim, p.H);
    const float* att = src;
    const float* Wl = src + 2 * H;
    const float* bl = Wl + (int64_t)H * in;
    const float* Wr = bl + H;
    const float* br = Wr + (int64_t)H * in;
    uint8_t* b1_hi = convs + cv * T::kConvBytes;
    uint8_t* b1_lo = b1_hi + T::kB1;
    uint8_t* b2_hi = b1_lo + T::kB1;
    uint8_t* b2_lo = b2_hi + T::kB2;
    float* vec = reinterpret_cast<float*>(b2_lo + T::kB2);
    for (int t = tid; t < T::kN1 * HP; t += kGnnThreads) {           // B1[n][k], n = channel (xl | xr)
      const int n = t / HP, k = t - n * HP, c = n % HP;
      float val = 0.f;
      if (c < H && k < in) val = (n >= HP ? Wr : Wl)[c * in + k];
      const float hi = to_tf32(val), lo = to_tf32(val - hi);
      const int byte = (n >> 3) * T::kB1Group + (k >> 2) * 128 + (n & 7) * 16 + (k & 3) * 4;
      *reinterpret_cast<float*>(b1_hi + byte) = hi;
      *reinterpret_cast<float*>(b1_lo + byte) = lo;
    }
    for (int t = tid; t < T::kNH * 2 * HP; t += kGnnThreads) {        // B2[n = k_in][kk = channel]
      const int n = t / (2 * HP), kk = t - n * 2 * HP, c = kk % HP;
      float val = 0.f;
      if (c < H && n < in) val = (kk >= HP ? Wr : Wl)[c * in + n];
      const float hi = to_tf32(val), lo = to_tf32(val - hi);
      const int byte = (n >> 3) * T::kB2Group + (kk >> 2) * 128 + (n & 7) * 16 + (kk & 3) * 4;
      *reinterpret_cast<float*>(b2_hi + byte) = hi;
      *reinterpret_cast<float*>(b2_lo + byte) = lo;
    }
    for (int t = tid; t < HP; t += kGnnThreads) {
      const bool ok = t < H;
      vec[t] = ok ? bl[t] : 0.f;
      vec[HP + t] = ok ? br[t] : 0.f;
      vec[2 * HP + t] = ok ? att[t] : 0.f;
      vec[3 * HP + t] = 0.f;
    }
    for (int t = tid; t < T::kSlotFloats; t += kGnnThreads) vec[4 * HP + t] = 0.f;   // d_att | d_bias slots
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
  const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
  const uint32_t col_d1 = 0, col_d2 = (uint32_t)(tiles * T::kN1);
  // instruction descriptors
  const uint32_t idesc_base = (1u << 4) | (2u << 7) | (2u << 10);
  const uint32_t idesc_r = idesc_base | ((uint32_t)(T::kN1 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  const uint32_t idesc_h = idesc_base | ((uint32_t)(T::kNH >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  uint32_t parity = 0;
  long long tprev = clock64();
  const bool dbg = p.dbg != nullptr && blockIdx.x == 0 && tid == 0;
#define GSC_TICK(slot)                                                                  \
  do {                                                                                  \
    if (dbg) { long long tn = clock64(); p.dbg[slot] += (float)(tn - tprev); tprev = tn; } \
  } while (0)

  const int num_chunks = (p.B + p.G - 1) / p.G;
  for (int chunk = blockIdx.x; chunk < num_chunks; chunk += gridDim.x) {
    const int g0 = chunk * p.G, g1 = min(p.B, g0 + p.G);
    const int nbase = p.graph_ptr[g0];
    const int nc = p.graph_ptr[g1] - nbase;
    if (nc > p.cap) __trap();     // a graph larger than the max_nodes the launch was sized for: fail loudly
    const int warp_first = tid & ~31;
    const int used_tiles = (nc + 127) >> 7;
    GSC_TICK(0);
    for (int g = g0 + warp; g < g1; g += kGnnThreads / 32)                  // node -> graph map
      for (int i = p.graph_ptr[g] - nbase + lane; i < p.graph_ptr[g + 1] - nbase; i += 32) gid_s[i] = g;
    // CSR rows of the chunk -> shared memory (walked 3x per application), and a degree-sorted node
    // order for the edge passes: lanes of a warp then run rows of similar length instead of all
    // waiting for the longest one.
    const int ebase = p.rowptr[nbase];
    const int ne = p.rowptr[nbase + nc] - ebase;            // <= emax (checked by gnn_chunk_edges_kernel)
    for (int i = tid; i <= nc; i += kGnnThreads) rp_s[i] = p.rowptr[nbase + i] - ebase;
    for (int e = tid; e < ne; e += kGnnThreads) col_s[e] = (uint16_t)(p.col[ebase + e] - nbase);
    if (tid < 64) hist_s[tid] = 0;
    __syncthreads();
    for (int i = tid; i < nc; i += kGnnThreads) atomicAdd(hist_s + 63 - min(rp_s[i + 1] - rp_s[i], 63), 1);
    __syncthreads();
    if (warp == 0) {                                   // exclusive scan of the 64 buckets
      int a0 = hist_s[2 * lane], a1 = hist_s[2 * lane + 1];
      int incl = a0 + a1;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      hist_s[2 * lane] = incl - a0 - a1;
      hist_s[2 * lane + 1] = incl - a1;
    }
    __syncthreads();
    for (int i = tid; i < nc; i += kGnnThreads)
      perm_s[atomicAdd(hist_s + 63 - min(rp_s[i + 1] - rp_s[i], 63), 1)] = (uint16_t)i;
    // Reverse-edge index: the edge j -> i stored at position e of row i has its reverse i -> j in row j
    // (the graph is symmetric); parallel edges pair up in order of occurrence.  The source pass reads
    // the attention weight and score gradient the target pass left at that position.
    for (int i = tid; i < nc; i += kGnnThreads) {
      const int e0 = rp_s[i], e1 = rp_s[i + 1];
      for (int e = e0; e < e1; ++e) {
        const int j = col_s[e];
        int k = 0;
        for (int q = e0; q < e; ++q) k += (col_s[q] == j);
        int r = e;
        for (int q = rp_s[j]; q < rp_s[j + 1]; ++q)
          if (col_s[q] == i && k-- == 0) { r = q; break; }
        rev_s[e] = (uint16_t)r;
      }
    }

    for (int a = p.num_apps - 1; a >= 0; --a) {
      const int cv = p.app_conv[a];
      const int d_in = cv == 0 ? p.in_dim : p.H;
      uint8_t* b1_hi = convs + cv * T::kConvBytes;
      uint8_t* b2_hi = b1_hi + 2 * T::kB1;
      float* vec = reinterpret_cast<float*>(b2_hi + 2 * T::kB2);
      const bool last_app = a == p.num_apps - 1;

      // ---- 1. layer input rows -> split h operand ----
      for (int i = tid; i < nc; i += kGnnThreads) {
        const int64_t node = (int64_t)nbase + i;
        uint8_t* hh = h_hi + (i >> 7) * T::kHTile;
        uint8_t* hl = h_lo + (i >> 7) * T::kHTile;
        const int r = i & 127;
#pragma unroll
        for (int c4 = 0; c4 < HP / 4; ++c4) {
          float4 v;
          if (a == 0) {
            v.x = (4 * c4 + 0 < p.in_dim) ? p.x[node * p.in_dim + 4 * c4 + 0] : 0.f;
            v.y = (4 * c4 + 1 < p.in_dim) ? p.x[node * p.in_dim + 4 * c4 + 1] : 0.f;
            v.z = (4 * c4 + 2 < p.in_dim) ? p.x[node * p.in_dim + 4 * c4 + 2] : 0.f;
            v.w = (4 * c4 + 3 < p.in_dim) ? p.x[node * p.in_dim + 4 * c4 + 3] : 0.f;
          } else {
            v = reinterpret_cast<const float4*>(p.saved + ((size_t)(a - 1) * p.Nt + node) * HP)[c4];
          }
          float4 hi, lo;
          split4(v, hi, lo);
          *reinterpret_cast<float4*>(hh + T::hoff(r, c4)) = hi;
          *reinterpret_cast<float4*>(hl + T::hoff(r, c4)) = lo;
        }
      }
      GSC_TICK(1);
      // ---- 2. (R) on the tensor cores ----
      fence_proxy_async_smem();
      tc_fence_before();
      __syncthreads();
      if (tid == 0) {
        tc_fence_after();
        const int ksteps = (d_in + 7) / 8;
        for (int t = 0; t < used_tiles; ++t) {
          const uint32_t ah = smem_u32(h_hi + t * T::kHTile), al = smem_u32(h_lo + t * T::kHTile);
          const uint32_t bh = smem_u32(b1_hi), blo = bh + T::kB1;
          const uint32_t d = tmem_base + col_d1 + (uint32_t)(t * T::kN1);
          for (int s = 0; s < ksteps; ++s) {
            const uint32_t ko = s * 256;
            tc_mma_tf32(d, smem_desc(ah + ko, 128, T::kHGroup, 0), smem_desc(bh + ko, 128, T::kB1Group, 0), idesc_r, s > 0);
            tc_mma_tf32(d, smem_desc(ah + ko, 128, T::kHGroup, 0), smem_desc(blo + ko, 128, T::kB1Group, 0), idesc_r, 1);
            tc_mma_tf32(d, smem_desc(al + ko, 128, T::kHGroup, 0), smem_desc(bh + ko, 128, T::kB1Group, 0), idesc_r, 1);
          }
        }
        tc_commit(bar);
      }
      mbar_wait(bar, parity);
      parity ^= 1u;
      tc_fence_after();
      GSC_TICK(2);
      // ---- 3. xl, xr, g = d_out/deg rows -> smem (own rows); per-node score scalars ----
      {
      float dbias[HP];
#pragma unroll
      for (int c = 0; c < HP; ++c) dbias[c] = 0.f;
      for (int i0 = warp_first; i0 < nc; i0 += kGnnThreads) {
        const int i = i0 + lane;
        float xl[HP], xr[HP], g[HP];
        const uint32_t t1 = tmem_base + lane_base + col_d1 + (uint32_t)((i >> 7) * T::kN1);
        tmem_load_row<HP>(t1, xl);
        tmem_load_row<HP>(t1 + HP, xr);
        if (!last_app) tmem_load_row<HP>(tmem_base + lane_base + col_d2 + (uint32_t)((i >> 7) * T::kNH), g);
        if (i < nc) {
          const int64_t node = (int64_t)nbase + i;
          if (last_app) {           // pool backward: d_emb[g] / n_g
            const int gg = gid_s[i];
            const float inv_n = 1.f / (float)(p.graph_ptr[gg + 1] - p.graph_ptr[gg]);
#pragma unroll
            for (int c = 0; c < HP; ++c) g[c] = (c < p.H) ? p.d_emb[(size_t)gg * p.emb_ld + c] * inv_n : 0.f;
          }
          if (p.app_relu[a]) {
            float out[HP];
            load_row<HP>(p.saved + ((size_t)a * p.Nt + node) * HP, out);
#pragma unroll
            for (int c = 0; c < HP; ++c) g[c] = out[c] > 0.f ? g[c] : 0.f;
          }
          const float inv_deg = 1.f / (float)(rp_s[i + 1] - rp_s[i]);
          float att[HP];
          load_row<HP>(vec + 2 * HP, att);
#pragma unroll
          for (int c = 0; c < HP; ++c) {
            xl[c] += vec[c];
            xr[c] += vec[HP + c];
            dbias[c] += g[c];
            g[c] *= inv_deg;
          }
          store_row<HP>(xl_s + (size_t)i * HP, xl);
          store_row<HP>(xr_s + (size_t)i * HP, xr);
          store_row<HP>(g_s + (size_t)i * HP, g);
          pq_s[i] = make_float2(0.6f * dot4<HP>(att, xl), 0.6f * dot4<HP>(att, xr));   // score scalars p_j, q_i
        }
      }
#pragma unroll
      for (int c = 0; c < HP; ++c) {
        const float sb = warp_sum(dbias[c]);
        if (lane == (c & 31)) vec[4 * HP + (warp * 2 + 1) * HP + c] += sb;      // warp-private slot
      }
      }
      __syncthreads();
      GSC_TICK(3);

      // ---- 4T. target side (gather over in-edges): softmax statistics, attention weight alpha_e and
      //      score gradient ds_e of every edge (left in es_s / eda_s for the source pass), d_xr.
      //      With lrelu(u) = (0.6 + 0.4 sgn u) u and w_c = sum_e ds_e (0.6 + 0.4 sgn u_e,c):
      //        d_xr_i,c = att_c w_in_i,c        d_xl_j,c = sum_e alpha_e g_i,c + att_c w_out_j,c
      //        d_att_c  = sum_i xr_i,c w_in_i,c + sum_j xl_j,c w_out_j,c
      //      so the only per-edge, per-channel work of the two gradient loops is the sign sum. ----
      float datt[HP];
#pragma unroll
      for (int c = 0; c < HP; ++c) datt[c] = 0.f;
      for (int t = tid; t < nc; t += kGnnThreads) {
        const int i = perm_s[t];
        float xr[HP], sg[HP];
        load_row<HP>(xr_s + (size_t)i * HP, xr);
        const float q_i = pq_s[i].y;
        const int e0 = rp_s[i], e1 = rp_s[i + 1];
        float m = -INFINITY, l = 0.f, dot = 0.f;
        {
          float g[HP], b[HP];
          load_row<HP>(g_s + (size_t)i * HP, g);
          load_row<HP>(vec + 2 * HP, b);
#pragma unroll
          for (int c = 0; c < HP; ++c) b[c] *= 0.4f;
          for (int e = e0; e < e1; ++e) {
            const int j = col_s[e];
            float v[HP];
            load_row<HP>(xl_s + (size_t)j * HP, v);
            const float sc_e = pq_s[j].x + q_i + abs_score<HP>(v, xr, b);
            const float da = dot4<HP>(g, v);
            es_s[e] = sc_e;
            eda_s[e] = da;
            const float m_new = fmaxf(m, sc_e);
            const float sc = __expf(m - m_new);
            const float pe = __expf(sc_e - m_new);
            l = fmaf(l, sc, pe);
            dot = fmaf(dot, sc, pe * da);
            m = m_new;
          }
        }
        const float inv_l = 1.f / (l + 1e-16f);
        dot *= inv_l;
        float ds_sum = 0.f;
#pragma unroll
        for (int c = 0; c < HP; ++c) sg[c] = 0.f;
        for (int e = e0; e < e1; ++e) {
          const int j = col_s[e];
          float v[HP];
          load_row<HP>(xl_s + (size_t)j * HP, v);
          const float alpha = __expf(es_s[e] - m) * inv_l;
          const float ds = alpha * (eda_s[e] - dot);
          es_s[e] = alpha;
          eda_s[e] = ds;
          ds_sum += ds;
#pragma unroll
          for (int c = 0; c < HP; ++c) sg[c] += signed_by(ds, v[c] + xr[c]);
        }
        const float base = 0.6f * ds_sum;
#pragma unroll
        for (int c = 0; c < HP; ++c) {
          const float w = fmaf(0.4f, sg[c], base);
          datt[c] = fmaf(xr[c], w, datt[c]);
          sg[c] = vec[2 * HP + c] * w;
        }
        store_row<HP>(dxr_raw + (size_t)i * HP, sg);
      }
      __syncthreads();
      GSC_TICK(8);
      // ---- 4S. source side: d_xl_j over the out-edges j -> i; the graph is symmetric, so they are the
      //      entries of j's own CSR row and rev_s points at what the target pass computed for them ----
      for (int t = tid; t < nc; t += kGnnThreads) {
        const int j = perm_s[t];
        float v[HP], dxl[HP], sg[HP];
        load_row<HP>(xl_s + (size_t)j * HP, v);
#pragma unroll
        for (int c = 0; c < HP; ++c) { dxl[c] = 0.f; sg[c] = 0.f; }
        float ds_sum = 0.f;
        const int e0 = rp_s[j], e1 = rp_s[j + 1];
        for (int e = e0; e < e1; ++e) {
          const int i = col_s[e], r = rev_s[e];
          const float alpha = es_s[r], ds = eda_s[r];
          float xr[HP], g[HP];
          load_row<HP>(xr_s + (size_t)i * HP, xr);
          load_row<HP>(g_s + (size_t)i * HP, g);
          ds_sum += ds;
#pragma unroll
          for (int c = 0; c < HP; ++c) {
            dxl[c] = fmaf(alpha, g[c], dxl[c]);
            sg[c] += signed_by(ds, v[c] + xr[c]);
          }
        }
        const float base = 0.6f * ds_sum;
#pragma unroll
        for (int c = 0; c < HP; ++c) {
          const float w = fmaf(0.4f, sg[c], base);
          datt[c] = fmaf(v[c], w, datt[c]);
          dxl[c] = fmaf(vec[2 * HP + c], w, dxl[c]);
        }
        store_row<HP>(xl_s + (size_t)j * HP, dxl);       // own row: xl_j is not needed any more
      }
#pragma unroll
      for (int c = 0; c < HP; ++c) {
        const float sa = warp_sum(datt[c]);
        if (lane == (c & 31)) vec[4 * HP + (warp * 2) * HP + c] += sa;          // warp-private slot
      }
      __syncthreads();
      GSC_TICK(4);
      // ---- 5. [dxl | dxr] rows -> workspace (for gnn_wgrad_kernel) and -> split dx operand of (H)
      //      (the operand buffers overlap the scratch rows: read, sync, write) ----
      {
        float dl0[HP], dr0[HP], dl1[HP], dr1[HP];      // up to 2 node slots per thread (cap <= 512)
        const int i_a = tid, i_b = tid + kGnnThreads;
        float* dxa = p.dx + ((size_t)a * p.Nt + nbase) * (2 * HP);
        if (i_a < nc) {
          load_row<HP>(xl_s + (size_t)i_a * HP, dl0); load_row<HP>(dxr_raw + (size_t)i_a * HP, dr0);
          store_row<HP>(dxa + (size_t)i_a * 2 * HP, dl0); store_row<HP>(dxa + (size_t)i_a * 2 * HP + HP, dr0);
        }
        if (i_b < nc) {
          load_row<HP>(xl_s + (size_t)i_b * HP, dl1); load_row<HP>(dxr_raw + (size_t)i_b * HP, dr1);
          store_row<HP>(dxa + (size_t)i_b * 2 * HP, dl1); store_row<HP>(dxa + (size_t)i_b * 2 * HP + HP, dr1);
        }
        __syncthreads();
        auto put = [&](int i, const float (&dl)[HP], const float (&dr)[HP]) {
          uint8_t* dh = d_hi + (i >> 7) * T::kDTile;
          uint8_t* dlo = d_lo + (i >> 7) * T::kDTile;
          const int r = i & 127;
#pragma unroll
          for (int c4 = 0; c4 < HP / 4; ++c4) {
            float4 vl = make_float4(0.f, 0.f, 0.f, 0.f), vr = vl, hi, lo;
            if (i < nc) {
              vl = make_float4(dl[4 * c4], dl[4 * c4 + 1], dl[4 * c4 + 2], dl[4 * c4 + 3]);
              vr = make_float4(dr[4 * c4], dr[4 * c4 + 1], dr[4 * c4 + 2], dr[4 * c4 + 3]);
            }
            split4(vl, hi, lo);
            *reinterpret_cast<float4*>(dh + T::doff(r, c4)) = hi;
            *reinterpret_cast<float4*>(dlo + T::doff(r, c4)) = lo;
            split4(vr, hi, lo);
            *reinterpret_cast<float4*>(dh + T::doff(r, HP / 4 + c4)) = hi;
            *reinterpret_cast<float4*>(dlo + T::doff(r, HP / 4 + c4)) = lo;
          }
        };
        if (i_a < used_tiles * 128) put(i_a, dl0, dr0);
        if (i_b < used_tiles * 128) put(i_b, dl1, dr1);
      }
      GSC_TICK(6);
      // ---- 6. (H) on the tensor cores ----
      fence_proxy_async_smem();
      tc_fence_before();
      __syncthreads();
      if (tid == 0) {
        tc_fence_after();
        for (int t = 0; t < used_tiles; ++t) {
          const uint32_t xh = smem_u32(d_hi + t * T::kDTile), xlo = smem_u32(d_lo + t * T::kDTile);
          if (a > 0) {
            const uint32_t bh = smem_u32(b2_hi), blo = bh + T::kB2;
            const uint32_t d = tmem_base + col_d2 + (uint32_t)(t * T::kNH);
            for (int s = 0; s < 2 * HP / 8; ++s) {
              const uint32_t ko = s * 256;
              tc_mma_tf32(d, smem_desc(xh + ko, 128, T::kDGroup, 0), smem_desc(bh + ko, 128, T::kB2Group, 0), idesc_h, s > 0);
              tc_mma_tf32(d, smem_desc(xh + ko, 128, T::kDGroup, 0), smem_desc(blo + ko, 128, T::kB2Group, 0), idesc_h, 1);
              tc_mma_tf32(d, smem_desc(xlo + ko, 128, T::kDGroup, 0), smem_desc(bh + ko, 128, T::kB2Group, 0), idesc_h, 1);
            }
          }
        }
        tc_commit(bar);
      }
      mbar_wait(bar, parity);
      parity ^= 1u;
      tc_fence_after();
      GSC_TICK(7);
    }
  }
#undef GSC_TICK

  // ---- CTA partial gradients -> workspace, in the flat state-dict layout of `params` ----
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const int64_t P = conv_param_offset(p.num_convs, p.in_dim, p.H);
  float* out = p.partials + (size_t)blockIdx.x * P;
  for (int cv = 0; cv < p.num_convs; ++cv) {          // att and conv-bias entries; the rest is gnn_wgrad_kernel's
    const int H = p.H;
    float* o = out + conv_param_offset(cv, p.in_dim, p.H);
    const float* vec = reinterpret_cast<const float*>(convs + cv * T::kConvBytes + 2 * T::kB1 + 2 * T::kB2);
    for (int t = tid; t < H; t += kGnnThreads) {
      float sa = 0.f, sb = 0.f;
      for (int w = 0; w < kGnnThreads / 32; ++w) {
        sa += vec[4 * HP + (w * 2) * HP + t];
        sb += vec[4 * HP + (w * 2 + 1) * HP + t];
      }
      o[t] = sa;
      o[H + t] = sb;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)tmem_cols);
  }
}

// ================================ gather backward, warp-level node transforms ===============
// The same gather algorithm as gnn_bwd_tc_kernel (target pass over in-edges, source pass over the
// out-edges of a symmetric graph through the reverse-edge index, no floating-point atomics), built
// for occupancy instead of tensor cores: the edge passes are chains of dependent shared-memory
// gathers and short FMA sequences, i.e. latency, and what hides latency is more resident warps.
//   * CTA = 128 threads = 128 node slots (one thread owns one node for a whole chunk of graphs);
//   * shared memory per CTA: two row arrays (xl, gathered by the target pass; g = d_out/deg, gathered by
//     the source pass), per edge the (alpha, ds) pair and ONE WORD OF SIGN BITS -- the LeakyReLU branch of
//     every channel of u = xl_j + xr_i, which is all that the two gradient loops need of xr and xl -- the
//     chunk's CSR with its reverse index and one c-major copy of every conv's weights: ~50 KB, so FOUR
//     CTAs (16 warps, 128 registers per thread) share an SM where the tcgen05 variant, with its operand
//     images, fits one; 1024 graphs then take two rounds of 592 resident CTAs instead of three of 444;
//   * the d_xl / d_xr / d_h gradients and the own xr row stay in the owner's registers; the MMA operand
//     rows are staged in the row arrays themselves, in the window where nobody gathers from them;
//   * the node transforms ([32 nodes of a warp] x [HP] x [2HP] and back) run as warp-level 3xTF32
//     mma.sync m16n8k8 through a small warp-private tile: ~110 instructions per warp instead of the
//     ~1400 of the FMA form, and none of the operand staging, TMEM round trips and CTA barriers that
//     tcgen05 needs for such small products;
//   * weight gradients: [dxl|dxr] rows go to the workspace for gnn_wgrad_kernel, as before.
// Used for symmetric batches whose largest graph has at most 128 nodes; larger graphs take the
// tcgen05 kernel (<= 384 nodes) or the scatter kernel.
#ifndef GSC_EDGE_UNROLL
#define GSC_EDGE_UNROLL 2
#endif
constexpr int kGbEdgeUnroll = GSC_EDGE_UNROLL;   // edge loops of the gather kernel
constexpr int kGbThreads = 128;
constexpr int kGbEdgesPerNode = 6;       // per-edge arrays per node slot (self-loops included)
constexpr size_t kBwdFlagBytes = 256;    // kernel-selection flags, between the partials and the node-row scratch
constexpr int kGbCtasPerSm = 4;          // resident CTAs per SM the shared-memory layout and register budget are sized for

template <int HP>
struct GbL {
  static constexpr int kConvFloats = 2 * HP * HP + 4 * HP;     // Wl[c][k], Wr[c][k], bl, br, att, 0.4 att
  static constexpr int kSlotFloats = (kGbThreads / 32) * 2 * HP;   // per-warp d_att | d_bias sums, per conv
  // Row stride of the two gathered row arrays: HP + 4 floats.  An odd number of 16-byte units spreads the
  // rows of a random gather over all eight 16-byte bank groups (HP = 24 alone reaches only four of them),
  // makes the MMA fragment loads conflict-free, and the pad of an xl row holds its score scalar p_j.
  static constexpr int kRS = HP + 4;
  __host__ __device__ static constexpr size_t bytes(int num_convs) {
    return (size_t)num_convs * (kConvFloats + kSlotFloats) * 4 + 2 * (size_t)kGbThreads * kRS * 4 /*xl, g rows*/ +
           ((size_t)kGbThreads + 4) * 4 /*rp*/ +
           (size_t)kGbEdgesPerNode * kGbThreads * (4 + 4 + 4 + 2 + 2) + (size_t)kGbThreads * 2 /*perm*/;
  }
};

// Column sums of a [32 lanes x N] register matrix, added to dst[0..N).  Instead of one 5-step butterfly
// per column (10 N instructions), every step with an even column count exchanges HALF of the columns
// with the partner lane (it keeps the other half), so the matrix shrinks as the lanes combine:
// N = 24: 12 + 6 + 3 exchanges, then two plain butterfly steps on 3 columns = 27 shuffles instead of 120.
template <int N, int MASK>
__device__ __forceinline__ void warp_colsum_step(float (&v)[N], int lane, int base, int writers, float* __restrict__ dst) {
  if constexpr (MASK == 0) {
    if ((lane & writers) == 0) {
#pragma unroll
      for (int i = 0; i < N; ++i) dst[base + i] += v[i];
    }
  } else if constexpr (N % 2 == 0) {
    const bool up = (lane & MASK) != 0;
    float h[N / 2];
#pragma unroll
    for (int i = 0; i < N / 2; ++i) {
      const float send = up ? v[i] : v[i + N / 2];
      const float keep = up ? v[i + N / 2] : v[i];
      h[i] = keep + __shfl_xor_sync(0xffffffffu, send, MASK);
    }
    warp_colsum_step<N / 2, MASK / 2>(h, lane, base + (up ? N / 2 : 0), writers, dst);
  } else {
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] += __shfl_xor_sync(0xffffffffu, v[i], MASK);
    warp_colsum_step<N, MASK / 2>(v, lane, base, writers | MASK, dst);      // replicated over this lane bit
  }
}
template <int N>
__device__ __forceinline__ void warp_colsum_add(float (&v)[N], int lane, float* __restrict__ dst) {
  warp_colsum_step<N, 16>(v, lane, 0, 0, dst);
}

// Warp-level tensor-core MMA (m16n8k8, TF32 operands, fp32 accumulate).  The node transforms of the
// gather kernel are [32 nodes] x [HP] x [2HP] per warp: far too small to stage for tcgen05, but as
// 3xTF32 mma.sync they cost ~110 instructions per warp where the FMA form costs ~1400.
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// x = hi + lo with hi = the 19 high bits (what the tensor core reads), lo = x - hi (exact), truncated likewise
__device__ __forceinline__ void tf32_split(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x) & 0xffffe000u;
  lo = __float_as_uint(x - __uint_as_float(hi)) & 0xffffe000u;
}
// acc[mt][nt] += A[32 x 8*KS] (warp tile in shared memory, row stride TS) x B, with B(k, n) = bmat[n*bn + k*bk]
// (any strided view of a weight matrix in shared memory), as 3xTF32: hi*hi + hi*lo + lo*hi.
template <int KS, int NT, int TS>
__device__ __forceinline__ void warp_gemm_3xtf32(float (&acc)[2][NT][4], const float* __restrict__ tile,
                                                 const float* __restrict__ bmat, int bn, int bk, int lane) {
  const int gid = lane >> 2, tig = lane & 3;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) {
    uint32_t ah[2][4], al[2][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
      const float* r0 = tile + (16 * mt + gid) * TS + 8 * ks + tig;
      tf32_split(r0[0], ah[mt][0], al[mt][0]);
      tf32_split(r0[8 * TS], ah[mt][1], al[mt][1]);
      tf32_split(r0[4], ah[mt][2], al[mt][2]);
      tf32_split(r0[8 * TS + 4], ah[mt][3], al[mt][3]);
    }
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const float* bp = bmat + (8 * nt + gid) * bn + (8 * ks + tig) * bk;
      uint32_t bh0, bl0, bh1, bl1;
      tf32_split(bp[0], bh0, bl0);
      tf32_split(bp[4 * bk], bh1, bl1);
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        mma_tf32(acc[mt][nt], ah[mt], bl0, bl1);
        mma_tf32(acc[mt][nt], al[mt], bh0, bh1);
        mma_tf32(acc[mt][nt], ah[mt], bh0, bh1);
      }
    }
  }
}

template <int HP>
__global__ void __launch_bounds__(kGbThreads, 4) gnn_bwd_gather_kernel(const GnnArgs p) {
  using L = GbL<HP>;
  constexpr int cap = kGbThreads, emax = kGbEdgesPerNode * kGbThreads;
  extern __shared__ float4 smem4[];
  float* wsm = reinterpret_cast<float*>(smem4);                     // [num_convs][kConvFloats]
  float* slot_s = wsm + p.num_convs * L::kConvFloats;               // [num_convs][warps][2HP]
  constexpr int kRS = L::kRS;
  float* xl_s = slot_s + p.num_convs * L::kSlotFloats;              // [cap][kRS] xl rows by node (+ 0.6 att.xl_j in the pad); step 4: MMA staging by slot
  float* g_s = xl_s + cap * kRS;                                    // [cap][kRS] g rows by node; step 1: MMA staging by slot
  int* rp_s = reinterpret_cast<int*>(g_s + cap * kRS);              // [cap + 4]
  float* es_s = reinterpret_cast<float*>(rp_s + cap + 4);           // [emax] score, then alpha
  float* eda_s = es_s + emax;                                       // [emax] g_i.xl_j, then d_score
  uint32_t* sb_s = reinterpret_cast<uint32_t*>(eda_s + emax);       // [emax] bit c: sign of xl_j,c + xr_i,c (LeakyReLU branch)
  uint16_t* col_s = reinterpret_cast<uint16_t*>(sb_s + emax);       // [emax]
  uint16_t* rev_s = col_s + emax;                                   // [emax]
  uint16_t* perm_s = rev_s + emax;                                  // [cap]
  if (p.stats == nullptr || p.stats[3] != 0 || p.flags[0] != 0) return;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int gid = lane >> 2, tig = lane & 3;

  // ---- weights, c-major and zero padded: Wl[c][k] at wsm[c*HP + k], Wr behind it, then bl, br, att, 0.4 att ----
  for (int cv = 0; cv < p.num_convs; ++cv) {
    const int in = cv == 0 ? p.in_dim : p.H, H = p.H;
    const float* src = p.params + conv_param_offset(cv, p.in_dim, p.H);
    const float* att = src;
    const float* Wl = src + 2 * H;
    const float* bl = Wl + (int64_t)H * in;
    const float* Wr = bl + H;
    const float* br = Wr + (int64_t)H * in;
    float* w = wsm + cv * L::kConvFloats;
    for (int t = tid; t < 2 * HP * HP; t += kGbThreads) {
      const int r = t >= HP * HP, tt = t - r * HP * HP, c = tt / HP, k = tt - c * HP;
      w[t] = (c < H && k < in) ? (r ? Wr : Wl)[c * in + k] : 0.f;
    }
    for (int t = tid; t < HP; t += kGbThreads) {
      const bool ok = t < H;
      w[2 * HP * HP + t] = ok ? bl[t] : 0.f;
      w[2 * HP * HP + HP + t] = ok ? br[t] : 0.f;
      w[2 * HP * HP + 2 * HP + t] = ok ? att[t] : 0.f;
      w[2 * HP * HP + 3 * HP + t] = ok ? 0.4f * att[t] : 0.f;
    }
    for (int t = tid; t < L::kSlotFloats; t += kGbThreads) slot_s[cv * L::kSlotFloats + t] = 0.f;
  }
  __syncthreads();

  const int num_chunks = (p.B + p.G - 1) / p.G;
  for (int chunk = blockIdx.x; chunk < num_chunks; chunk += gridDim.x) {
    const int g0 = chunk * p.G, g1 = min(p.B, g0 + p.G);
    const int nbase = p.graph_ptr[g0];
    const int nc = p.graph_ptr[g1] - nbase;
    if (nc > cap) __trap();     // a graph larger than the max_nodes the launch was sized for: fail loudly
    // chunk CSR -> shared memory; degree-sorted node order and reverse-edge index from gnn_chunk_prep_kernel
    const int ebase = p.rowptr[nbase];
    const int ne = p.rowptr[nbase + nc] - ebase;            // <= emax (checked by gnn_chunk_prep_kernel)
    const uint16_t* prep = p.prep + (size_t)chunk * (cap + emax);
    for (int i = tid; i <= nc; i += kGbThreads) rp_s[i] = p.rowptr[nbase + i] - ebase;
    for (int e = tid; e < ne; e += kGbThreads) {
      col_s[e] = (uint16_t)(p.col[ebase + e] - nbase);
      rev_s[e] = prep[cap + e];
    }
    if (tid < nc) perm_s[tid] = prep[tid];
    __syncthreads();

    // this thread's node for the whole chunk (rows of similar length share a warp)
    const bool active = tid < nc;
    const int i = active ? perm_s[tid] : 0;
    const int64_t node = (int64_t)nbase + i;
    const int e0 = active ? rp_s[i] : 0, e1 = active ? rp_s[i + 1] : 0;
    const float inv_deg = active ? 1.f / (float)(e1 - e0) : 0.f;
    float g[HP];                                         // d_out of the current application for node i
#pragma unroll
    for (int c = 0; c < HP; ++c) g[c] = 0.f;
    if (active) {                                        // pool backward: d_emb[graph] / n_graph
      int gg = g0;
      while (gg + 1 < g1 && p.graph_ptr[gg + 1] - nbase <= i) ++gg;
      const float inv_n = 1.f / (float)(p.graph_ptr[gg + 1] - p.graph_ptr[gg]);
#pragma unroll
      for (int c = 0; c < HP; ++c) g[c] = (c < p.H) ? p.d_emb[(size_t)gg * p.emb_ld + c] * inv_n : 0.f;
    }

    for (int a = p.num_apps - 1; a >= 0; --a) {
      const int cv = p.app_conv[a];
      const float* w = wsm + cv * L::kConvFloats;
      const float* vec = w + 2 * HP * HP;                // bl | br | att
      float* slot = slot_s + cv * L::kSlotFloats + warp * 2 * HP;
      // ---- 1. recompute [xl | xr] = h_in [Wl ; Wr]^T on the warp's 32 node slots (3xTF32 mma.sync).  The own
      //         input row is staged in this warp's slot rows of g_s (nobody reads g rows between the last
      //         barrier and the next one); B fragments come straight from the c-major weights; the xl
      //         fragments (+ bias) go to the xl rows of their nodes, the xr fragments return to their owner
      //         through the same slot rows ----
      float xr[HP];
      {
        float* stage_rows = g_s + warp * 32 * kRS;
        {
          float hin[HP];
#pragma unroll
          for (int k = 0; k < HP; ++k) hin[k] = 0.f;
          if (active) {
            if (a == 0) {
#pragma unroll
              for (int k = 0; k < HP; ++k) hin[k] = (k < p.in_dim) ? p.x[node * p.in_dim + k] : 0.f;
            } else {
              load_row<HP>(p.saved + ((size_t)(a - 1) * p.Nt + node) * HP, hin);
            }
          }
          store_row<HP>(stage_rows + lane * kRS, hin);
        }
        __syncwarp();
        float acc[2][2 * HP / 8][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int nt = 0; nt < 2 * HP / 8; ++nt)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[mt][nt][q] = 0.f;
        if (a == 0 && p.in_dim <= 8) warp_gemm_3xtf32<1, 2 * HP / 8, kRS>(acc, stage_rows, w, HP, 1, lane);
        else warp_gemm_3xtf32<HP / 8, 2 * HP / 8, kRS>(acc, stage_rows, w, HP, 1, lane);
        __syncwarp();                                    // every lane is done reading the staged rows
        // rows (16 mt + gid) and (+ 8), channels 8 nt + 2 tig (+ 1): nt < HP/8 is xl, the rest xr
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int hrow = 0; hrow < 2; ++hrow) {
            const int sl = 16 * mt + 8 * hrow + gid;
            const int slot_i = warp * 32 + sl;
            const bool live = slot_i < nc;
            const int ni = live ? perm_s[slot_i] : 0;
            float pl = 0.f;
#pragma unroll
            for (int nt = 0; nt < 2 * HP / 8; ++nt) {
              const int ch = (8 * nt + 2 * tig) % HP;
              const bool is_r = nt >= HP / 8;
              const float v0 = acc[mt][nt][2 * hrow] + vec[(is_r ? HP : 0) + ch];
              const float v1 = acc[mt][nt][2 * hrow + 1] + vec[(is_r ? HP : 0) + ch + 1];
              if (is_r) {
                *reinterpret_cast<float2*>(stage_rows + sl * kRS + ch) = make_float2(v0, v1);
              } else {
                pl = fmaf(vec[2 * HP + ch], v0, fmaf(vec[2 * HP + ch + 1], v1, pl));
                if (live) *reinterpret_cast<float2*>(xl_s + ni * kRS + ch) = make_float2(v0, v1);
              }
            }
            pl += __shfl_xor_sync(0xffffffffu, pl, 1);
            pl += __shfl_xor_sync(0xffffffffu, pl, 2);
            if (live && tig == 0) xl_s[ni * kRS + HP] = 0.6f * pl;                  // score scalar p_j
          }
        __syncwarp();
        load_row<HP>(stage_rows + lane * kRS, xr);
        // g = relu'(out) d_out / deg of the own node; conv-bias gradient
        float dbias[HP];
#pragma unroll
        for (int c = 0; c < HP; ++c) dbias[c] = 0.f;
        if (active) {
          if (p.app_relu[a]) {
            float out[HP];
            load_row<HP>(p.saved + ((size_t)a * p.Nt + node) * HP, out);
#pragma unroll
            for (int c = 0; c < HP; ++c) g[c] = out[c] > 0.f ? g[c] : 0.f;
          }
#pragma unroll
          for (int c = 0; c < HP; ++c) { dbias[c] = g[c]; g[c] *= inv_deg; }
        }
        warp_colsum_add<HP>(dbias, lane, slot + HP);      // warp-private slot
      }
      __syncthreads();

      // ---- 2. target side: per in-edge score, g.xl_j and the LeakyReLU branch signs (one word per edge);
      //         softmax statistics; alpha_e, ds_e left per edge; d_xr.  With lrelu(u) = (0.6 + 0.4 sgn u) u and
      //         w_c = sum_e ds_e (0.6 + 0.4 sgn u_e,c):  d_xr = att o w_in,  d_xl = sum alpha g + att o w_out,
      //         d_att = sum_i xr_i o w_in,i + sum_j xl_j o w_out,j ----
      float datt[HP], vown[HP];
      float* dxa = p.dx + ((size_t)a * p.Nt + node) * (2 * HP);      // [dxl | dxr] row of the own node (gnn_wgrad_kernel)
#pragma unroll
      for (int c = 0; c < HP; ++c) { datt[c] = 0.f; vown[c] = 0.f; }
      if (active) {
        store_row<HP>(g_s + i * kRS, g);                  // read by the neighbours' source passes (after the next barrier)
        float sg[HP];
        const float q_i = 0.6f * dot4<HP>(*reinterpret_cast<const float(*)[HP]>(vec + 2 * HP), xr);
        float m = -INFINITY;
#pragma unroll 1
        for (int e = e0; e < e1; ++e) {
          const int j = col_s[e];
          float v[HP];
          load_row<HP>(xl_s + j * kRS, v);
          float s0 = 0.f, s1 = 0.f, d0 = 0.f, d1 = 0.f;
          uint32_t bits = 0;
#pragma unroll
          for (int c = 0; c < HP; c += 2) {
            const float t0 = v[c] + xr[c], t1 = v[c + 1] + xr[c + 1];
            s0 = fmaf(vec[3 * HP + c], fabsf(t0), s0);
            s1 = fmaf(vec[3 * HP + c + 1], fabsf(t1), s1);
            d0 = fmaf(g[c], v[c], d0);
            d1 = fmaf(g[c + 1], v[c + 1], d1);
            if (t0 < 0.f) bits |= 1u << c;
            if (t1 < 0.f) bits |= 2u << c;
          }
          const float sc_e = xl_s[j * kRS + HP] + q_i + (s0 + s1);
          es_s[e] = sc_e;
          eda_s[e] = d0 + d1;
          sb_s[e] = bits;
          m = fmaxf(m, sc_e);
        }
        float l = 0.f, dot = 0.f;
        for (int e = e0; e < e1; ++e) {
          const float pe = __expf(es_s[e] - m);
          es_s[e] = pe;
          l += pe;
          dot = fmaf(pe, eda_s[e], dot);
        }
        const float inv_l = 1.f / (l + 1e-16f);
        dot *= inv_l;
        float ds_sum = 0.f;
#pragma unroll
        for (int c = 0; c < HP; ++c) sg[c] = 0.f;
#pragma unroll 1
        for (int e = e0; e < e1; ++e) {
          const float alpha = es_s[e] * inv_l;
          const float ds = alpha * (eda_s[e] - dot);
          const uint32_t bits = sb_s[e];
          es_s[e] = alpha;
          eda_s[e] = ds;
          ds_sum += ds;
#pragma unroll
          for (int c = 0; c < HP; ++c)
            if (bits & (1u << c)) sg[c] += ds;           // sum of ds over the edges on the negative branch
        }
#pragma unroll
        for (int c = 0; c < HP; ++c) {
          const float wv = fmaf(-0.8f, sg[c], ds_sum);   // sum_e ds_e (0.6 + 0.4 sgn) = ds_sum - 0.8 sum_neg
          datt[c] = xr[c] * wv;
          sg[c] = vec[2 * HP + c] * wv;                  // d_xr
        }
        store_row<HP>(dxa + HP, sg);                     // read back in step 4 (registers are short at 4 CTAs per SM)
        load_row<HP>(xl_s + i * kRS, vown);               // own xl row for the source pass: xl_s is free after the barrier
      }
      warp_colsum_add<HP>(datt, lane, slot);             // target half of d_att
#pragma unroll
      for (int c = 0; c < HP; ++c) datt[c] = 0.f;
      __syncthreads();

      // ---- 3. source side: d_xl over the out-edges = this node's own CSR row of a symmetric graph; what the
      //         target pass left for edge j -> i (alpha, ds, branch signs) is found through the reverse index ----
      float dxl[HP];
#pragma unroll
      for (int c = 0; c < HP; ++c) dxl[c] = 0.f;
      if (active) {
        float sg[HP];
#pragma unroll
        for (int c = 0; c < HP; ++c) sg[c] = 0.f;
        float ds_sum = 0.f;
#pragma unroll 1
        for (int e = e0; e < e1; ++e) {
          const int ti = col_s[e], r = rev_s[e];
          const float alpha = es_s[r], ds = eda_s[r];
          const uint32_t bits = sb_s[r];
          float gi[HP];
          load_row<HP>(g_s + ti * kRS, gi);
          ds_sum += ds;
#pragma unroll
          for (int c = 0; c < HP; ++c) {
            dxl[c] = fmaf(alpha, gi[c], dxl[c]);
            if (bits & (1u << c)) sg[c] += ds;
          }
        }
#pragma unroll
        for (int c = 0; c < HP; ++c) {
          const float wv = fmaf(-0.8f, sg[c], ds_sum);
          datt[c] = vown[c] * wv;
          dxl[c] = fmaf(vec[2 * HP + c], wv, dxl[c]);
        }
        store_row<HP>(dxa, dxl);
      }
      __syncwarp();                                      // slot[] is updated by different lanes in the two halves
      warp_colsum_add<HP>(datt, lane, slot);
      // ---- 4. d_h_in = d_xl Wl + d_xr Wr -> d_out of application a-1: two [32 x HP] x [HP x HP] products; the
      //         operand rows are staged in this warp's slot rows of xl_s (no xl row is read after the last
      //         barrier), and the result rows return to their owners the same way ----
      if (a > 0) {
        float* stage_rows = xl_s + warp * 32 * kRS;
        float acc[2][HP / 8][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int nt = 0; nt < HP / 8; ++nt)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[mt][nt][q] = 0.f;
        store_row<HP>(stage_rows + lane * kRS, dxl);
        __syncwarp();
        warp_gemm_3xtf32<HP / 8, HP / 8, kRS>(acc, stage_rows, w, 1, HP, lane);               // B(k = c, n) = Wl[c][n]
        __syncwarp();
        {
          float dxr[HP];
#pragma unroll
          for (int c = 0; c < HP; ++c) dxr[c] = 0.f;
          if (active) {                                  // this thread's own store of step 2, from L2
#pragma unroll
            for (int c4 = 0; c4 < HP / 4; ++c4) {
              const float4 t4 = __ldcg(reinterpret_cast<const float4*>(dxa + HP) + c4);
              dxr[4 * c4] = t4.x; dxr[4 * c4 + 1] = t4.y; dxr[4 * c4 + 2] = t4.z; dxr[4 * c4 + 3] = t4.w;
            }
          }
          store_row<HP>(stage_rows + lane * kRS, dxr);
        }
        __syncwarp();
        warp_gemm_3xtf32<HP / 8, HP / 8, kRS>(acc, stage_rows, w + HP * HP, 1, HP, lane);     // ... + d_xr Wr
        __syncwarp();
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int nt = 0; nt < HP / 8; ++nt) {
            *reinterpret_cast<float2*>(stage_rows + (16 * mt + gid) * kRS + 8 * nt + 2 * tig) = make_float2(acc[mt][nt][0], acc[mt][nt][1]);
            *reinterpret_cast<float2*>(stage_rows + (16 * mt + 8 + gid) * kRS + 8 * nt + 2 * tig) = make_float2(acc[mt][nt][2], acc[mt][nt][3]);
          }
        __syncwarp();
        load_row<HP>(stage_rows + lane * kRS, g);
      }
      __syncthreads();        // the rows are rewritten by the next application / chunk
    }
  }

  // ---- CTA partial sums of d_att and the conv bias -> partial row (the rest is gnn_wgrad_kernel's) ----
  __syncthreads();
  const int64_t P = conv_param_offset(p.num_convs, p.in_dim, p.H);
  float* out = p.partials + (size_t)blockIdx.x * P;
  for (int cv = 0; cv < p.num_convs; ++cv) {
    float* o = out + conv_param_offset(cv, p.in_dim, p.H);
    const float* sl = slot_s + cv * L::kSlotFloats;
    for (int t = tid; t < p.H; t += kGbThreads) {
      float sa = 0.f, sb = 0.f;
      for (int wv = 0; wv < kGbThreads / 32; ++wv) {
        sa += sl[wv * 2 * HP + t];
        sb += sl[wv * 2 * HP + HP + t];
      }
      o[t] = sa;
      o[p.H + t] = sb;
    }
  }
}

// (W) of the gather backward kernels: for every conv, [dWl|dbl ; dWr|dbr] = sum over its applications and over
// all nodes of [dxl|dxr]_i (x) [h_in_i | 1], i.e. the product dX^T [2HP x nodes] x [H_in | 1] [nodes x (HP+1)]
// with the reduction over the nodes.  Streaming kernel: CTA b owns the node slice [b*Nt/grid, (b+1)*Nt/grid)
// and walks it in tiles of 64 nodes staged in shared memory (next tile prefetched into registers); each of
// the eight warps takes one 8-node k-step of a tile and issues 3xTF32 mma.sync m16n8k8 for all 2HP/16 x
// (HP+8)/8 output tiles (the A operand is dX read transposed out of the staged rows).  The warps' accumulators are
// added in a fixed order and the CTA's sums go to the W / bl / br entries of partial row b
// (gnn_bwd_reduce_kernel adds the rows).  Algorithmic bytes: 4 * (2HP + d_in) per node and application.
template <int HP>
struct WgL {
  static constexpr int kCh = 2 * HP;                           // channels: dxl | dxr (padded to a multiple of 16 below)
  static constexpr int kMT = (kCh + 15) / 16;                  // m16 tiles over the channels
  static constexpr int kNT = (HP + 8) / 8;                     // n8 tiles over [h | 1 | 0...]
  static constexpr int pad_stride(int n) { return (n + 8) % 16 == 8 ? n + 8 : n + 16; }   // result % 32 is 8 or 24
  static constexpr int kDS = pad_stride(16 * kMT);             // dx row stride in smem: conflict-free transposed A reads
  static constexpr int kHS = pad_stride(8 * kNT);              // h row stride in smem: likewise for the B reads
  static constexpr int kTile = 64;                             // nodes per staged tile = 8 warps x one k-step
  static constexpr int kTileF4 = kTile * (kCh + HP) / 4;       // float4 per staged tile (global side)
  static constexpr int kPre = (kTileF4 + kGnnThreads - 1) / kGnnThreads;
  static constexpr int kRedStride = 8 * kNT + 1;
  static constexpr int kSmemFloats = 2 * kTile * (kDS + kHS) + (kGnnThreads / 32) * 16 * kMT * kRedStride;
};

template <int HP>
__global__ void __launch_bounds__(kGnnThreads) gnn_wgrad_kernel(const GnnArgs p) {
  using L = WgL<HP>;
  constexpr int kCh = L::kCh, kMT = L::kMT, kNT = L::kNT, kDS = L::kDS, kHS = L::kHS, kTile = L::kTile;
  extern __shared__ float4 smem4[];
  float* dx_s = reinterpret_cast<float*>(smem4);               // [2][kTile][kDS]
  float* h_s = dx_s + 2 * kTile * kDS;                         // [2][kTile][kHS]
  float* red_s = h_s + 2 * kTile * kHS;                        // [warps][16 kMT][kRedStride]
  if (p.stats == nullptr || p.stats[3] != 0 || p.flags[0] != 0) return;     // the scatter kernel ran instead
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, gid = lane >> 2, tig = lane & 3;
  const int64_t lo = p.Nt * blockIdx.x / gridDim.x, hi = p.Nt * (blockIdx.x + 1) / gridDim.x;
  const int64_t P = conv_param_offset(p.num_convs, p.in_dim, p.H);
  float* out = p.partials + (size_t)blockIdx.x * P;
  // constant parts of the staged rows: dx pad channels = 0; h columns HP.. = 1, 0, 0, ...
  for (int t = tid; t < 2 * kTile * (kDS - kCh); t += kGnnThreads) dx_s[(t / (kDS - kCh)) * kDS + kCh + t % (kDS - kCh)] = 0.f;
  for (int t = tid; t < 2 * kTile * (kHS - HP); t += kGnnThreads) h_s[(t / (kHS - HP)) * kHS + HP + t % (kHS - HP)] = (t % (kHS - HP)) == 0 ? 1.f : 0.f;

  for (int cv = 0; cv < p.num_convs; ++cv) {
    const int d_in = cv == 0 ? p.in_dim : p.H;
    float acc[kMT][kNT][4];
#pragma unroll
    for (int mt = 0; mt < kMT; ++mt)
#pragma unroll
      for (int nt = 0; nt < kNT; ++nt)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[mt][nt][q] = 0.f;
    for (int a = 0; a < p.num_apps; ++a) {
      if (p.app_conv[a] != cv) continue;
      const float* dxa = p.dx + (size_t)a * p.Nt * kCh;
      const float* ha = a == 0 ? nullptr : p.saved + (size_t)(a - 1) * p.Nt * HP;
      // staged tile, float4 index e on the global side: e < kTile*kCh/4 -> dx rows, else h rows
      float4 pre[L::kPre];
      auto fetch = [&](int64_t n0) {
#pragma unroll
        for (int r = 0; r < L::kPre; ++r) {
          const int e = tid + r * kGnnThreads;
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          if (e < kTile * kCh / 4) {
            if (n0 + e / (kCh / 4) < hi) v = reinterpret_cast<const float4*>(dxa + (size_t)n0 * kCh)[e];
          } else if (e < L::kTileF4) {
            const int eh = e - kTile * kCh / 4, i = eh / (HP / 4), q = eh % (HP / 4);
            if (n0 + i < hi) {
              if (ha != nullptr) {
                v = reinterpret_cast<const float4*>(ha + (size_t)(n0 + i) * HP)[q];
              } else {
                const float* xr = p.x + (size_t)(n0 + i) * p.in_dim;
                v.x = 4 * q + 0 < p.in_dim ? xr[4 * q + 0] : 0.f;
                v.y = 4 * q + 1 < p.in_dim ? xr[4 * q + 1] : 0.f;
                v.z = 4 * q + 2 < p.in_dim ? xr[4 * q + 2] : 0.f;
                v.w = 4 * q + 3 < p.in_dim ? xr[4 * q + 3] : 0.f;
              }
            }
          }
          pre[r] = v;
        }
      };
      auto stash = [&](int buf) {
#pragma unroll
        for (int r = 0; r < L::kPre; ++r) {
          const int e = tid + r * kGnnThreads;
          if (e < kTile * kCh / 4) {
            const int i = e / (kCh / 4), q = e % (kCh / 4);
            *reinterpret_cast<float4*>(dx_s + (buf * kTile + i) * kDS + 4 * q) = pre[r];
          } else if (e < L::kTileF4) {
            const int eh = e - kTile * kCh / 4, i = eh / (HP / 4), q = eh % (HP / 4);
            *reinterpret_cast<float4*>(h_s + (buf * kTile + i) * kHS + 4 * q) = pre[r];
          }
        }
      };
      int buf = 0;
      fetch(lo);
      __syncthreads();                                         // previous readers of buffer 0 are done
      stash(0);
      __syncthreads();
      for (int64_t n0 = lo; n0 < hi; n0 += kTile, buf ^= 1) {
        const bool more = n0 + kTile < hi;
        if (more) fetch(n0 + kTile);
        {                                                      // this warp's k-step: nodes 8*warp .. 8*warp+7 of the tile
          const float* dt = dx_s + (buf * kTile + 8 * warp) * kDS;     // rows past the slice are zero
          const float* ht = h_s + (buf * kTile + 8 * warp) * kHS;
          uint32_t bh[kNT][2], bl[kNT][2];
#pragma unroll
          for (int nt = 0; nt < kNT; ++nt) {                   // B(k = node, n = input column)
            tf32_split(ht[tig * kHS + 8 * nt + gid], bh[nt][0], bl[nt][0]);
            tf32_split(ht[(tig + 4) * kHS + 8 * nt + gid], bh[nt][1], bl[nt][1]);
          }
#pragma unroll
          for (int mt = 0; mt < kMT; ++mt) {                   // A(row = channel, col = node): dX read transposed
            uint32_t ah[4], al[4];
            tf32_split(dt[tig * kDS + 16 * mt + gid], ah[0], al[0]);
            tf32_split(dt[tig * kDS + 16 * mt + gid + 8], ah[1], al[1]);
            tf32_split(dt[(tig + 4) * kDS + 16 * mt + gid], ah[2], al[2]);
            tf32_split(dt[(tig + 4) * kDS + 16 * mt + gid + 8], ah[3], al[3]);
#pragma unroll
            for (int nt = 0; nt < kNT; ++nt) {
              mma_tf32(acc[mt][nt], ah, bl[nt][0], bl[nt][1]);
              mma_tf32(acc[mt][nt], al, bh[nt][0], bh[nt][1]);
              mma_tf32(acc[mt][nt], ah, bh[nt][0], bh[nt][1]);
            }
          }
        }
        if (more) stash(buf ^ 1);                              // nobody reads that buffer in this iteration
        __syncthreads();
      }
    }
    // warps -> CTA sums (fixed order) -> partial row
    {
      float* rw = red_s + warp * 16 * kMT * L::kRedStride;
#pragma unroll
      for (int mt = 0; mt < kMT; ++mt)
#pragma unroll
        for (int nt = 0; nt < kNT; ++nt) {
          rw[(16 * mt + gid) * L::kRedStride + 8 * nt + 2 * tig] = acc[mt][nt][0];
          rw[(16 * mt + gid) * L::kRedStride + 8 * nt + 2 * tig + 1] = acc[mt][nt][1];
          rw[(16 * mt + gid + 8) * L::kRedStride + 8 * nt + 2 * tig] = acc[mt][nt][2];
          rw[(16 * mt + gid + 8) * L::kRedStride + 8 * nt + 2 * tig + 1] = acc[mt][nt][3];
        }
    }
    __syncthreads();
    const int H = p.H;
    float* o = out + conv_param_offset(cv, p.in_dim, p.H);
    for (int t = tid; t < kCh * (d_in + 1); t += kGnnThreads) {
      const int ch = t / (d_in + 1), kk = t % (d_in + 1);
      const int is_r = ch >= HP, c = ch - is_r * HP;
      if (c >= H) continue;
      const int k = kk == d_in ? HP : kk;                      // last entry: the bias column
      float sum = 0.f;
#pragma unroll
      for (int w = 0; w < kGnnThreads / 32; ++w) sum += red_s[(w * 16 * kMT + ch) * L::kRedStride + k];
      const int64_t base = 2 * H + (is_r ? (int64_t)H * d_in + H : 0);
      if (kk == d_in) o[base + (int64_t)H * d_in + c] = sum;   // lin_{l,r}.bias
      else o[base + (int64_t)c * d_in + kk] = sum;             // lin_{l,r}.weight[c][k]
    }
    __syncthreads();
  }
}

// flags[0] := 1 if some chunk of G graphs holds more edges (self-loops included) than the gather
// backward can stage in shared memory; both backward kernels read it to decide which of them runs.
__global__ void gnn_chunk_edges_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ graph_ptr,
                                       int B, int G, int emax, int32_t* __restrict__ flags) {
  const int num_chunks = (B + G - 1) / G;
  for (int ch = blockIdx.x * blockDim.x + threadIdx.x; ch < num_chunks; ch += gridDim.x * blockDim.x) {
    const int g0 = ch * G, g1 = min(B, g0 + G);
    if (rowptr[graph_ptr[g1]] - rowptr[graph_ptr[g0]] > emax) flags[0] = 1;
  }
}

// Per-chunk index work of the gather backward, which depends on the topology only, done once per launch
// instead of once per resident CTA pass: for every chunk of G graphs
//   perm[slot] : the chunk's nodes by descending degree (stable: ties by node index, so the slot of a
//                node -- and with it every fp32 summation order in the gather kernel -- is deterministic);
//   rev[e]     : for CSR entry e = (j -> i) of row i, the entry of row j that holds (i -> j) (the k-th
//                duplicate maps to the k-th duplicate), i.e. where the target pass of i leaves what the
//                source pass of j needs.
// Layout: prep[chunk][kGbThreads + emax] uint16.  flags[0] := 1 if a chunk does not fit (the scatter
// kernel then handles the batch).
__global__ void __launch_bounds__(128)
gnn_chunk_prep_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                      const int32_t* __restrict__ graph_ptr, int B, int G, int cap, int emax,
                      int32_t* __restrict__ flags, uint16_t* __restrict__ prep) {
  extern __shared__ int prep_smem[];
  int* rp_s = prep_smem;                                           // [cap + 1]
  int* cnt_s = rp_s + cap + 1;                                     // [4 warps][65] nodes per degree bucket
  int* start_s = cnt_s + 4 * 65;                                   // [64]
  uint16_t* col_s = reinterpret_cast<uint16_t*>(start_s + 64);     // [emax]
  const int tid = threadIdx.x;
  const int num_chunks = (B + G - 1) / G;
  for (int chunk = blockIdx.x; chunk < num_chunks; chunk += gridDim.x) {
    const int g0 = chunk * G, g1 = min(B, g0 + G);
    const int nbase = graph_ptr[g0], nc = graph_ptr[g1] - nbase;
    const int ebase = rowptr[nbase], ne = rowptr[nbase + nc] - ebase;
    if (nc > cap || ne > emax) {                                   // uniform per CTA
      if (tid == 0) flags[0] = 1;
      continue;
    }
    uint16_t* out = prep + (size_t)chunk * (cap + emax);
    for (int i = tid; i <= nc; i += blockDim.x) rp_s[i] = rowptr[nbase + i] - ebase;
    for (int e = tid; e < ne; e += blockDim.x) col_s[e] = (uint16_t)(col[ebase + e] - nbase);
    __syncthreads();
    // stable counting sort by descending degree (bucket 63 - min(deg, 63)): per-warp bucket counts from
    // match.any, bucket starts from a scan over the 64 buckets, position = start + earlier warps + earlier lanes
    const int warp = tid >> 5, lane = tid & 31;
    const int i = tid;                                             // cap == blockDim.x: one node per thread
    const int bkt = i < nc ? 63 - min(rp_s[i + 1] - rp_s[i], 63) : 64;
    for (int t = tid; t < 4 * 65; t += blockDim.x) cnt_s[t] = 0;
    __syncthreads();
    const uint32_t same = __match_any_sync(0xffffffffu, bkt);
    if ((same & ((1u << lane) - 1u)) == 0) cnt_s[warp * 65 + bkt] = __popc(same);     // first lane of its group
    __syncthreads();
    if (warp == 0) {                                               // exclusive scan of the bucket totals (2 per lane)
      int t0 = 0, t1 = 0;
      for (int w = 0; w < 4; ++w) { t0 += cnt_s[w * 65 + 2 * lane]; t1 += cnt_s[w * 65 + 2 * lane + 1]; }
      int incl = t0 + t1;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      start_s[2 * lane] = incl - t0 - t1;
      start_s[2 * lane + 1] = incl - t1;
    }
    __syncthreads();
    if (i < nc) {
      int pos = start_s[bkt] + __popc(same & ((1u << lane) - 1u));
      for (int w = 0; w < warp; ++w) pos += cnt_s[w * 65 + bkt];
      out[pos] = (uint16_t)i;
      const int e0 = rp_s[i], e1 = rp_s[i + 1];
      for (int e = e0; e < e1; ++e) {
        const int j = col_s[e];
        int k = 0;
        for (int q = e0; q < e; ++q) k += (col_s[q] == j);
        int r = e;
        for (int q = rp_s[j]; q < rp_s[j + 1]; ++q)
          if (col_s[q] == i && k-- == 0) { r = q; break; }
        out[cap + e] = (uint16_t)r;
      }
    }
    __syncthreads();
  }
}

// d_params[t] = sum of the CTA partials, in a fixed order: block = 32 parameters x 8 partial-lanes,
// lane r adds the partials r, r + 8, ... (independent loads in flight), then the lanes are added.
// Sums the partial rows in a fixed order.  When the gather kernel ran (same run-time condition as in
// the kernels themselves), the W / lin-bias entries were written by gnn_wgrad_kernel, which uses
// `rows_w` CTAs (a streaming kernel wants long node slices), the att / bias entries by the gather
// kernel's `num_partials` CTAs.
constexpr int kRedLanes = 32;     // partial-row lanes per parameter: 32 parameters x 32 lanes per block
__global__ void __launch_bounds__(32 * kRedLanes)
gnn_bwd_reduce_kernel(const float* __restrict__ partials, int num_partials, int rows_w, int64_t P, float* __restrict__ d_params,
                      const int32_t* __restrict__ stats, const int32_t* __restrict__ flags, int num_convs, int in_dim, int H) {
  __shared__ float red[kRedLanes][33];
  const int c = threadIdx.x & 31, r = threadIdx.x >> 5;
  const int64_t t = (int64_t)blockIdx.x * 32 + c;
  int rows = num_partials;
  if (rows_w < num_partials && stats != nullptr && stats[3] == 0 && flags[0] == 0) {
    for (int cv = 0; cv < num_convs; ++cv) {
      const int64_t o = conv_param_offset(cv, in_dim, H);
      if (t >= o + 2 * H && t < conv_param_offset(cv + 1, in_dim, H)) rows = rows_w;
    }
  }
  // lane r adds the rows r, r + 32, ... (two independent chains); the lanes are then added in lane order
  float s0 = 0.f, s1 = 0.f;
  if (t < P) {
    int b = r;
    for (; b + kRedLanes < rows; b += 2 * kRedLanes) {
      s0 += partials[(size_t)b * P + t];
      s1 += partials[(size_t)(b + kRedLanes) * P + t];
    }
    if (b < rows) s0 += partials[(size_t)b * P + t];
  }
  red[r][c] = s0 + s1;
  __syncthreads();
  if (r == 0 && t < P) {
    float tot = 0.f;
#pragma unroll
    for (int k = 0; k < kRedLanes; ++k) tot += red[k][c];
    d_params[t] = tot;
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

// Graphs per CTA: as many as fit the node capacity, but never so many that a small batch leaves SMs
// idle (B = 100 graphs of 100 nodes would otherwise occupy 50 of the 148 SMs).
static int graphs_per_cta(int cap, int max_nodes, int B, int ctas_per_sm) {
  int G = cap / max_nodes;
  if (G < 1) G = 1;
  const int spread = (int)ceil_div(B, (int64_t)kNumSMs * ctas_per_sm);
  if (G > spread) G = spread < 1 ? 1 : spread;
  return G;
}

// tcgen05 forward: usable when the TMEM accumulators (2*HP columns per 128 nodes) and the operand
// tiles of the largest graph fit; otherwise the FMA kernel above runs.
template <int HP>
static int launch_fwd_tc(GnnArgs a, int max_nodes, cudaStream_t st, bool* done) {
  using T = TcL<HP>;
  *done = false;
  const int cap = max_nodes <= kGnnThreads ? kGnnThreads : (max_nodes + 127) / 128 * 128;
  const int tiles = cap / 128;
  const size_t smem = (size_t)a.num_convs * T::kConvBytes + 2 * (size_t)tiles * T::kATile + (size_t)cap * T::kXS * 4 +
                      ((size_t)cap + 4 + (size_t)kColPerNode * cap) * 4 + 16;
  int cols = 32;
  while (cols < tiles * T::kN) cols *= 2;
  if (cols > 512 || smem > 200 * 1024) return GSC_OK;
  a.cap = cap;
  const int per_sm = (2 * smem <= 220 * 1024 && 2 * cols <= 512) ? 2 : 1;
  a.G = graphs_per_cta(cap, max_nodes, a.B, per_sm);
  GSC_CUDA_TRY(cudaFuncSetAttribute(gnn_fwd_tc_kernel<HP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int chunks = (a.B + a.G - 1) / a.G;
  const int grid = chunks < per_sm * kNumSMs ? chunks : per_sm * kNumSMs;
  gnn_fwd_tc_kernel<HP><<<grid, kGnnThreads, smem, st>>>(a, cols);
  GSC_LAUNCH_CHECK("gnn_fwd_tc_kernel");
  *done = true;
  return GSC_OK;
}

// GSC_GNN_BWD=1 keeps the tcgen05 gather backward for graphs that the FMA gather kernel could take (experiments)
static int gnn_bwd_variant() {
  static int mode = -1;
  if (mode < 0) {
    const char* e = getenv("GSC_GNN_BWD");
    mode = e ? atoi(e) : 0;
  }
  return mode;
}

static bool gnn_use_tc() {
  static int mode = -1;
  if (mode < 0) {
    const char* e = getenv("GSC_GNN_FMA");
    mode = (e && e[0] == '1') ? 0 : 1;
  }
  return mode == 1;
}

// The backward kernels fill every SM they run on (registers and shared memory), so nothing else can
// start beside them; under data parallelism a few SMs are left to the NCCL kernels that reduce the
// dense-layer gradients meanwhile (gsc_set_sm_reserve).
static int bwd_grid(int B, int G) {
  const int chunks = (B + G - 1) / G;
  const int sms = kNumSMs - sm_reserve();
  return chunks < sms ? chunks : sms;
}

// Launch plan of the tcgen05 / gather backward (symmetric batches): feasible when the two-slot
// node mapping, the operand images and the TMEM accumulators fit.
template <int HP>
static bool plan_bwd_tc(const GnnArgs& a, int max_nodes, int* cap, int* cols, size_t* smem) {
  using T = TcB<HP>;
  *cap = max_nodes <= kGnnThreads ? kGnnThreads : (max_nodes + 127) / 128 * 128;
  const int tiles = *cap / 128;
  if (*cap > 384) return false;     // two node slots per thread; the atomic kernel must fit as well
  const size_t emax = (size_t)kBwdEdgesPerNode * *cap;
  *smem = (size_t)a.num_convs * T::kConvBytes + 2 * (size_t)tiles * T::kHTile + 2 * (size_t)tiles * T::kDTile +
          (size_t)*cap * (4 + 8 + 4 + 2) + 16 + emax * (4 + 4 + 2 + 2) + 64 * 4 + 16;
  const int need_cols = tiles * (T::kN1 + T::kNH);
  *cols = 32;
  while (*cols < need_cols) *cols *= 2;
  return *cols <= 512 && *smem <= 227 * 1024 && a.num_convs <= T::kMaxConvs;
}

template <int HP>
static int launch_bwd_both(GnnArgs a, int max_nodes, float* d_params, const void* prep_ext, cudaStream_t st) {
  // atomic (any graph) kernel geometry
  size_t smem_f = (2 * (size_t)a.num_convs * ConvLayout<HP>::kFloats + 5 * (size_t)a.cap * HP) * 4;
  const bool global_rows = smem_f > 227 * 1024;
  if (global_rows) smem_f = 2 * (size_t)a.num_convs * ConvLayout<HP>::kFloats * 4;
  const int grid_f = bwd_grid(a.B, a.G);
  int cap_t = 0, cols = 0;
  size_t smem_t = 0;
  const bool tc = gnn_use_tc() && a.stats != nullptr && plan_bwd_tc<HP>(a, max_nodes, &cap_t, &cols, &smem_t);
  int grid = grid_f, rows_w = 0;
  const size_t smem_g = GbL<HP>::bytes(a.num_convs);
  // resident CTAs per SM: four when the shared-memory layout allows (two convs at HP = 24), else fewer
  int gb_ctas = (int)((227 * 1024) / (smem_g + 1024));
  if (gb_ctas > kGbCtasPerSm) gb_ctas = kGbCtasPerSm;
  const bool gather_fma = gnn_bwd_variant() != 1 && a.stats != nullptr && max_nodes <= kGbThreads && gb_ctas >= 1;
  // the chunk index work (gsc_gnn_bwd_prep) may come with the batch: [flag word, 256 B][perm | rev per chunk]
  const bool ext = prep_ext != nullptr && gather_fma;
  if (ext) {
    a.flags = static_cast<const int32_t*>(prep_ext);
    a.prep = reinterpret_cast<const uint16_t*>(static_cast<const uint8_t*>(prep_ext) + kBwdFlagBytes);
  } else {
    GSC_CUDA_TRY(cudaMemsetAsync(const_cast<int32_t*>(a.flags), 0, sizeof(int32_t), st));
  }
  if (gather_fma) {
    GnnArgs t = a;
    t.cap = kGbThreads;
    t.G = graphs_per_cta(kGbThreads, max_nodes, t.B, gb_ctas);
    const int chunks_t = (t.B + t.G - 1) / t.G;
    if (!ext) {
      constexpr int emax = kGbEdgesPerNode * kGbThreads;
      const int smem_p = (kGbThreads + 1 + 4 * 65 + 64) * 4 + emax * 2;
      gnn_chunk_prep_kernel<<<chunks_t < 8 * kNumSMs ? chunks_t : 8 * kNumSMs, 128, smem_p, st>>>(
          t.rowptr, t.col, t.graph_ptr, t.B, t.G, kGbThreads, emax, const_cast<int32_t*>(a.flags),
          const_cast<uint16_t*>(a.prep));
      GSC_LAUNCH_CHECK("gnn_chunk_prep_kernel");
    }
    int grid_g = gb_ctas * (kNumSMs - sm_reserve());
    if (chunks_t < grid_g) grid_g = chunks_t;
    grid = grid_g > grid_f ? grid_g : grid_f;       // one partial row per CTA of either kernel
    GSC_CUDA_TRY(cudaFuncSetAttribute(gnn_bwd_gather_kernel<HP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_g));
    gnn_bwd_gather_kernel<HP><<<grid, kGbThreads, smem_g, st>>>(t);       // returns at once unless symmetric
    GSC_LAUNCH_CHECK("gnn_bwd_gather_kernel");
    const int smem_w = WgL<HP>::kSmemFloats * 4;
    GSC_CUDA_TRY(cudaFuncSetAttribute(gnn_wgrad_kernel<HP>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_w));
    static const int wg_ctas = [] { const char* e = getenv("GSC_GNN_WGRAD_CTAS"); return e ? atoi(e) : 2; }();
    rows_w = wg_ctas > 0 ? min(grid, wg_ctas * kNumSMs) : grid;
    gnn_wgrad_kernel<HP><<<rows_w, kGnnThreads, smem_w, st>>>(t);         // its weight gradients (same condition)
    GSC_LAUNCH_CHECK("gnn_wgrad_kernel");
    a.gather_kernel_launched = 1;
  } else if (tc) {
    GnnArgs t = a;
    t.cap = cap_t;
    t.G = graphs_per_cta(cap_t, max_nodes, t.B, 1);
    const int chunks_t = (t.B + t.G - 1) / t.G;
    gnn_chunk_edges_kernel<<<(chunks_t + 255) / 256, 256, 0, st>>>(t.rowptr, t.graph_ptr, t.B, t.G,
                                                                   kBwdEdgesPerNode * cap_t, const_cast<int32_t*>(a.flags));
    GSC_LAUNCH_CHECK("gnn_chunk_edges_kernel");
    const int grid_t = bwd_grid(t.B, t.G);
    grid = grid_t > grid_f ? grid_t : grid_f;      // one partial per CTA of either kernel
    GSC_CUDA_TRY(cudaFuncSetAttribute(gnn_bwd_tc_kernel<HP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_t));
    gnn_bwd_tc_kernel<HP><<<grid, kGnnThreads, smem_t, st>>>(t, cols);   // returns at once unless symmetric
    GSC_LAUNCH_CHECK("gnn_bwd_tc_kernel");
    const int smem_w = WgL<HP>::kSmemFloats * 4;
    GSC_CUDA_TRY(cudaFuncSetAttribute(gnn_wgrad_kernel<HP>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_w));
    gnn_wgrad_kernel<HP><<<grid, kGnnThreads, smem_w, st>>>(t);           // its weight gradients (same condition)
    GSC_LAUNCH_CHECK("gnn_wgrad_kernel");
    a.gather_kernel_launched = 1;
  }
  if (global_rows) {
    GSC_CUDA_TRY(cudaFuncSetAttribute(gnn_bwd_kernel<HP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_f));
    gnn_bwd_kernel<HP, true><<<grid, kGnnThreads, smem_f, st>>>(a);
  } else {
    GSC_CUDA_TRY(cudaFuncSetAttribute(gnn_bwd_kernel<HP, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_f));
    gnn_bwd_kernel<HP, false><<<grid, kGnnThreads, smem_f, st>>>(a);     // returns at once if the other one ran
  }
  GSC_LAUNCH_CHECK("gnn_bwd_kernel");
  const int64_t P = conv_param_offset(a.num_convs, a.in_dim, a.H);
  if (rows_w == 0) rows_w = grid;
  gnn_bwd_reduce_kernel<<<(int)ceil_div(P, 32), 32 * kRedLanes, 0, st>>>(a.partials, grid, rows_w, P, d_params, a.stats, a.flags,
                                                              a.num_convs, a.in_dim, a.H);
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
  a.dbg = g_gnn_debug;
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

// Partial gradients of every CTA, plus (graphs too large for shared memory) the node-row scratch.
static size_t bwd_partials_bytes(int in_dim, int H, int num_layers) {   // one row per CTA, up to 4 CTAs per SM
  return ((size_t)4 * kNumSMs * (size_t)conv_param_offset(num_layers, in_dim, H) * 4 + 255) / 256 * 256;
}
static bool bwd_rows_in_global(int max_nodes, int H, int num_layers) {
  const int hp = pick_hp(H);
  const size_t cap = max_nodes <= kGnnThreads ? kGnnThreads : ((size_t)max_nodes + 31) / 32 * 32;
  const size_t conv_floats = hp == 8 ? ConvLayout<8>::kFloats : hp == 16 ? ConvLayout<16>::kFloats
                           : hp == 24 ? ConvLayout<24>::kFloats : ConvLayout<32>::kFloats;
  return (2 * (size_t)num_layers * conv_floats + 5 * cap * hp) * 4 > 227 * 1024;
}

static int num_apps_of(int num_layers, int num_iter) { return 1 + (num_layers > 1 ? num_iter * (num_layers - 1) : 0); }
// [d_xl | d_xr] rows of every application (tcgen05 backward -> gnn_wgrad_kernel)
static size_t bwd_dx_bytes(int64_t Nt, int H, int num_layers, int num_iter) {
  return ((size_t)num_apps_of(num_layers, num_iter) * (size_t)Nt * 2 * pick_hp(H) * 4 + 255) / 256 * 256;
}

// [perm | rev] of every chunk of the gather backward (at most one chunk per graph)
static size_t bwd_prep_bytes(int64_t B, int max_nodes) {
  if (max_nodes > kGbThreads) return 0;
  return ((size_t)(B > 0 ? B : 1) * (kGbThreads + kGbEdgesPerNode * kGbThreads) * 2 + 255) / 256 * 256;
}

extern "C" size_t gsc_gnn_bwd_workspace_bytes(int64_t Nt, int64_t B, int max_nodes, int in_dim, int H, int num_layers,
                                              int num_iter) {
  if (Nt < 0 || max_nodes < 1 || H < 1 || H > GSC_MAX_HIDDEN || num_layers < 1 || num_iter < 0) return 0;
  size_t bytes = bwd_partials_bytes(in_dim, H, num_layers) + kBwdFlagBytes + bwd_prep_bytes(B, max_nodes) +
                 bwd_dx_bytes(Nt, H, num_layers, num_iter);
  if (bwd_rows_in_global(max_nodes, H, num_layers)) {
    const size_t cap = ((size_t)max_nodes + 31) / 32 * 32;
    bytes += (size_t)kNumSMs * 5 * cap * pick_hp(H) * 4;
  }
  return bytes;
}

static int gnn_bwd_impl(const float* x, int64_t Nt, int in_dim, const int32_t* rowptr,
                        const int32_t* col, const int32_t* graph_ptr, int64_t B, int max_nodes,
                        const float* params, int H, int num_layers, int num_iter,
                        const float* saved, const float* d_emb, int64_t d_emb_ld, float* d_params,
                        const int32_t* pack_stats, void* workspace, size_t workspace_bytes,
                        const void* prep_ext, void* stream) {
  GnnArgs a{};
  int rc = build_args(a, x, Nt, in_dim, rowptr, col, graph_ptr, B, max_nodes, params, H, num_layers, num_iter);
  if (rc) return rc;
  GSC_REQUIRE(saved && d_emb && d_params && workspace && d_emb_ld >= H, "gnn bwd: null pointer or d_emb_ld < hidden");
  a.emb_ld = d_emb_ld;
  if (workspace_bytes < gsc_gnn_bwd_workspace_bytes(Nt, B, max_nodes, in_dim, H, num_layers, num_iter))
    return fail(GSC_EWORKSPACE, "gnn bwd: workspace %zu < %zu", workspace_bytes,
                gsc_gnn_bwd_workspace_bytes(Nt, B, max_nodes, in_dim, H, num_layers, num_iter));
  a.saved = const_cast<float*>(saved); a.d_emb = d_emb; a.partials = (float*)workspace;
  a.flags = reinterpret_cast<const int32_t*>(static_cast<uint8_t*>(workspace) + bwd_partials_bytes(in_dim, H, num_layers));
  a.prep = reinterpret_cast<const uint16_t*>(static_cast<uint8_t*>(workspace) + bwd_partials_bytes(in_dim, H, num_layers) + kBwdFlagBytes);
  a.dx = reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) + bwd_partials_bytes(in_dim, H, num_layers) + kBwdFlagBytes +
                                  bwd_prep_bytes(B, max_nodes));
  a.rows_ws = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(a.dx) + bwd_dx_bytes(Nt, H, num_layers, num_iter));
  a.dbg = g_gnn_debug;
  a.stats = pack_stats;
  cudaStream_t st = (cudaStream_t)stream;
  switch (pick_hp(H)) {
    case 8: return launch_bwd_both<8>(a, max_nodes, d_params, prep_ext, st);
    case 16: return launch_bwd_both<16>(a, max_nodes, d_params, prep_ext, st);
    case 24: return launch_bwd_both<24>(a, max_nodes, d_params, prep_ext, st);
    default: return launch_bwd_both<32>(a, max_nodes, d_params, prep_ext, st);
  }
}

extern "C" int gsc_gnn_embed_bwd(const float* x, int64_t Nt, int in_dim, const int32_t* rowptr,
                                 const int32_t* col, const int32_t* graph_ptr, int64_t B, int max_nodes,
                                 const float* params, int H, int num_layers, int num_iter,
                                 const float* saved, const float* d_emb, int64_t d_emb_ld, float* d_params,
                                 const int32_t* pack_stats, void* workspace, size_t workspace_bytes,
                                 void* stream) {
  return gnn_bwd_impl(x, Nt, in_dim, rowptr, col, graph_ptr, B, max_nodes, params, H, num_layers, num_iter, saved,
                      d_emb, d_emb_ld, d_params, pack_stats, workspace, workspace_bytes, nullptr, stream);
}

extern "C" int gsc_gnn_embed_bwd_prepared(const float* x, int64_t Nt, int in_dim, const int32_t* rowptr,
                                          const int32_t* col, const int32_t* graph_ptr, int64_t B, int max_nodes,
                                          const float* params, int H, int num_layers, int num_iter,
                                          const float* saved, const float* d_emb, int64_t d_emb_ld, float* d_params,
                                          const int32_t* pack_stats, void* workspace, size_t workspace_bytes,
                                          const void* prep, void* stream) {
  return gnn_bwd_impl(x, Nt, in_dim, rowptr, col, graph_ptr, B, max_nodes, params, H, num_layers, num_iter, saved,
                      d_emb, d_emb_ld, d_params, pack_stats, workspace, workspace_bytes, prep, stream);
}

// Geometry of the gather backward for (hidden, num_layers, max_nodes, B): resident CTAs per SM and graphs per chunk
static bool gather_geometry(int H, int num_layers, int max_nodes, int64_t B, int* G) {
  if (max_nodes > kGbThreads || max_nodes < 1 || num_layers < 1 || H < 1 || H > GSC_MAX_HIDDEN) return false;
  const int hp = pick_hp(H);
  const size_t smem_g = hp == 8 ? GbL<8>::bytes(num_layers) : hp == 16 ? GbL<16>::bytes(num_layers)
                      : hp == 24 ? GbL<24>::bytes(num_layers) : GbL<32>::bytes(num_layers);
  int ctas = (int)((227 * 1024) / (smem_g + 1024));
  if (ctas > kGbCtasPerSm) ctas = kGbCtasPerSm;
  if (ctas < 1) return false;
  *G = graphs_per_cta(kGbThreads, max_nodes, (int)B, ctas);
  return true;
}

extern "C" size_t gsc_gnn_bwd_prep_bytes(int64_t B, int max_nodes) {
  const size_t body = bwd_prep_bytes(B, max_nodes);
  return body ? kBwdFlagBytes + body : 0;
}

extern "C" int gsc_gnn_bwd_prep(const int32_t* rowptr, const int32_t* col, const int32_t* graph_ptr, int64_t B,
                                int max_nodes, int H, int num_layers, void* prep, size_t prep_bytes, void* stream) {
  GSC_REQUIRE(rowptr && col && graph_ptr && prep && B >= 1 && B < INT32_MAX, "gsc_gnn_bwd_prep: bad arguments");
  int G = 1;
  GSC_REQUIRE(gather_geometry(H, num_layers, max_nodes, B, &G),
              "gsc_gnn_bwd_prep: no gather backward for max_nodes=%d hidden=%d layers=%d", max_nodes, H, num_layers);
  if (prep_bytes < gsc_gnn_bwd_prep_bytes(B, max_nodes))
    return fail(GSC_EWORKSPACE, "gsc_gnn_bwd_prep: buffer %zu < %zu", prep_bytes, gsc_gnn_bwd_prep_bytes(B, max_nodes));
  cudaStream_t st = (cudaStream_t)stream;
  GSC_CUDA_TRY(cudaMemsetAsync(prep, 0, sizeof(int32_t), st));
  constexpr int emax = kGbEdgesPerNode * kGbThreads;
  const int smem_p = (kGbThreads + 1 + 4 * 65 + 64) * 4 + emax * 2;
  const int chunks = (int)((B + G - 1) / G);
  gnn_chunk_prep_kernel<<<chunks < 8 * kNumSMs ? chunks : 8 * kNumSMs, 128, smem_p, st>>>(
      rowptr, col, graph_ptr, (int)B, G, kGbThreads, emax, static_cast<int32_t*>(prep),
      reinterpret_cast<uint16_t*>(static_cast<uint8_t*>(prep) + kBwdFlagBytes));
  GSC_LAUNCH_CHECK("gnn_chunk_prep_kernel");
  return GSC_OK;
}
