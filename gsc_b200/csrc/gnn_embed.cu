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
#include <math.h>

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
  cudaStream_t st = (cudaStream_t)stream;
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
