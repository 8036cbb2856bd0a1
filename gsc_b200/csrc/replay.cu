// Device-resident replay storage helpers and the structured action mask.
//
// The reference keeps its replay buffer in host numpy/object arrays and re-collates a minibatch with
// Batch.from_data_list on every gradient step (reference src/rlsp/agents/buffer.py:36-89); every
// observation also carries a dense mask[1, A] that is a pure function of its node count
// (mask[:n, :, :, :n] = 1, reference src/rlsp/envs/simulator_wrapper.py:302-306).  Here transitions
// live in HBM in padded struct-of-arrays form, a minibatch is gathered by index on the device
// straight into the Batch layout the packer reads, and masks are generated from n instead of being
// stored or copied.
//
// All kernels are flat HBM streaming: coalesced 4/8/16-byte accesses, grid a multiple of the SM count.
#include "common.cuh"

namespace gsc {

// mask[b, ((i*cs + q)*n_max + j)] = (i < n_b) && (j < n_b)
__global__ void build_mask_kernel(const int32_t* __restrict__ n, int64_t B, int n_max, int cs,
                                  float* __restrict__ mask) {
  const int64_t A = (int64_t)n_max * cs * n_max;
  const int64_t total4 = B * A / 4;                       // A % 4 == 0 checked by the host
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total4;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = t * 4;
    const int64_t b = e / A;
    const int64_t k = e - b * A;
    const int nb = n[b];
    float v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int64_t kk = k + u;
      const int i = (int)(kk / ((int64_t)cs * n_max));
      const int j = (int)(kk % n_max);
      v[u] = (i < nb && j < nb) ? 1.f : 0.f;
    }
    reinterpret_cast<float4*>(mask)[t] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

__global__ void build_mask_scalar_kernel(const int32_t* __restrict__ n, int64_t B, int n_max, int cs,
                                         float* __restrict__ mask) {
  const int64_t A = (int64_t)n_max * cs * n_max;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < B * A;
       e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = e / A, k = e - b * A;
    const int nb = n[b];
    mask[e] = ((int)(k / ((int64_t)cs * n_max)) < nb && (int)(k % n_max) < nb) ? 1.f : 0.f;
  }
}

// One CTA per sampled transition: padded storage slot idx[b] -> Batch layout at node_ptr[b] / edge_ptr[b].
__global__ void __launch_bounds__(128)
replay_gather_kernel(const float* __restrict__ sx, const int32_t* __restrict__ sei, const int32_t* __restrict__ sn,
                     const int32_t* __restrict__ sec, const int64_t* __restrict__ idx,
                     const int32_t* __restrict__ node_ptr, const int32_t* __restrict__ edge_ptr, int64_t B,
                     int n_max, int F, int e_max, int64_t e_total, float* __restrict__ x_out,
                     int64_t* __restrict__ ei_out, int64_t* __restrict__ batch_out, int32_t* __restrict__ n_out) {
  for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
    const int64_t slot = idx[b];
    const int nb = sn[slot], eb = sec[slot];
    const int np = node_ptr[b], ep = edge_ptr[b];
    if (threadIdx.x == 0 && n_out) n_out[b] = nb;
    const float* xs = sx + slot * (int64_t)n_max * F;
    for (int t = threadIdx.x; t < nb * F; t += blockDim.x) x_out[(int64_t)np * F + t] = xs[t];
    for (int t = threadIdx.x; t < nb; t += blockDim.x) batch_out[np + t] = b;
    const int32_t* es = sei + slot * 2 * (int64_t)e_max;
    for (int t = threadIdx.x; t < eb; t += blockDim.x) {
      ei_out[ep + t] = (int64_t)es[t] + np;
      ei_out[e_total + ep + t] = (int64_t)es[e_max + t] + np;
    }
  }
}

// out[b, :] = src[idx[b], :]   (actions [cap, A] -> [B, A]; also rewards/dones with A == 1)
__global__ void gather_rows_kernel(const float* __restrict__ src, const int64_t* __restrict__ idx, int64_t B,
                                   int64_t A, float* __restrict__ out) {
  if ((A & 3) == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
    const int64_t A4 = A >> 2, total = B * A4;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * blockDim.x) {
      const int64_t b = t / A4, k = t - b * A4;
      reinterpret_cast<float4*>(out)[t] = reinterpret_cast<const float4*>(src + idx[b] * A)[k];
    }
  } else {
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < B * A;
         t += (int64_t)gridDim.x * blockDim.x) {
      const int64_t b = t / A, k = t - b * A;
      out[t] = src[idx[b] * A + k];
    }
  }
}

static inline int stream_grid(int64_t n, int threads) {
  int64_t b = ceil_div(n > 0 ? n : 1, threads);
  const int64_t cap = (int64_t)kNumSMs * 8;
  return (int)(b < cap ? b : cap);
}

}  // namespace gsc

using namespace gsc;

extern "C" int gsc_build_mask(const int32_t* n, int64_t B, int n_max, int groups_per_node, float* mask,
                              void* stream) {
  GSC_REQUIRE(n && mask && B >= 0 && n_max >= 1 && groups_per_node >= 1, "gsc_build_mask: bad arguments");
  if (B == 0) return GSC_OK;
  const int64_t A = (int64_t)n_max * groups_per_node * n_max;
  if ((A & 3) == 0 && (reinterpret_cast<uintptr_t>(mask) & 15) == 0)
    build_mask_kernel<<<stream_grid(B * A / 4, 256), 256, 0, (cudaStream_t)stream>>>(n, B, n_max, groups_per_node, mask);
  else
    build_mask_scalar_kernel<<<stream_grid(B * A, 256), 256, 0, (cudaStream_t)stream>>>(n, B, n_max, groups_per_node, mask);
  GSC_LAUNCH_CHECK("build_mask_kernel");
  return GSC_OK;
}

extern "C" int gsc_replay_gather(const float* store_x, const int32_t* store_ei, const int32_t* store_n,
                                 const int32_t* store_ecount, const int64_t* idx, const int32_t* node_ptr,
                                 const int32_t* edge_ptr, int64_t B, int n_max, int num_feat, int e_max,
                                 int64_t num_edges_total, float* x_out, int64_t* edge_index_out,
                                 int64_t* batch_out, int32_t* n_out, void* stream) {
  GSC_REQUIRE(store_x && store_ei && store_n && store_ecount && idx && node_ptr && edge_ptr && x_out &&
                  (edge_index_out || num_edges_total == 0) && batch_out && B >= 0,
              "gsc_replay_gather: bad arguments");
  if (B == 0) return GSC_OK;
  const int grid = (int)(B < 8LL * kNumSMs ? B : 8LL * kNumSMs);
  replay_gather_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(store_x, store_ei, store_n, store_ecount, idx,
                                                               node_ptr, edge_ptr, B, n_max, num_feat, e_max,
                                                               num_edges_total, x_out, edge_index_out, batch_out, n_out);
  GSC_LAUNCH_CHECK("replay_gather_kernel");
  return GSC_OK;
}

extern "C" int gsc_gather_rows(const float* src, const int64_t* idx, int64_t B, int64_t A, float* out, void* stream) {
  GSC_REQUIRE(src && idx && out && B >= 0 && A >= 1, "gsc_gather_rows: bad arguments");
  if (B == 0) return GSC_OK;
  gather_rows_kernel<<<stream_grid(B * A / 4 + 1, 256), 256, 0, (cudaStream_t)stream>>>(src, idx, B, A, out);
  GSC_LAUNCH_CHECK("gather_rows_kernel");
  return GSC_OK;
}
