// Node features of a batch of network states.
//
// The reference builds every observation on the host: deepcopy of the networkx graph, python loops
// over the nodes and from_networkx (reference src/rlsp/envs/simulator_wrapper.py:237-308).  The
// arithmetic of its three node features is three per-state normalisations (:255-294):
//   ingress_traffic : clip(t / (max(t) + 1e-3), 0, 1)
//   node_load       : u = (cap == 0 ? 1 : usage / cap);  clip(u / (max(u) + 1e-3), 0, 1)
//   node_cap        : clip(cap / (max(cap) + 1e-3), 0, 1)
// with float32 arrays, the maximum taken per state, and `max + 1e-3` evaluated in double and rounded
// back to float32 (numpy 1.23: float32 scalar + Python float -> float64, float32 array / that scalar
// -> float32).  Here the raw per-node quantities of many states (parallel rollouts) sit in device
// arrays and one launch writes x[Nt, 3] in the Batch layout the packer and the embedder read.
// One warp per state; bit-exact with the oracle restatement.
#include "common.cuh"

namespace gsc {

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// usage / cap is a Python-float (double) division in the reference, stored into a float32 array (:277-281)
__device__ __forceinline__ float utilisation(float usage, float cap) {
  return cap == 0.f ? 1.f : (float)((double)usage / (double)cap);
}

// float32(max + 1e-3 in double): the divisor numpy 1.23 uses
__device__ __forceinline__ float norm_denominator(float mx) { return (float)((double)mx + 1e-3); }

__device__ __forceinline__ float clip01(float v) { return fminf(fmaxf(v, 0.f), 1.f); }

__global__ void __launch_bounds__(256)
node_features_kernel(const float* __restrict__ ingress, const float* __restrict__ usage, const float* __restrict__ cap,
                     const int32_t* __restrict__ graph_ptr, int64_t B, float* __restrict__ x) {
  const int lane = threadIdx.x & 31;
  const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t g = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); g < B; g += warps) {
    const int n0 = graph_ptr[g], n1 = graph_ptr[g + 1];
    float m_in = -INFINITY, m_ut = -INFINITY, m_cap = -INFINITY;
    for (int i = n0 + lane; i < n1; i += 32) {
      m_in = fmaxf(m_in, ingress[i]);
      m_ut = fmaxf(m_ut, utilisation(usage[i], cap[i]));
      m_cap = fmaxf(m_cap, cap[i]);
    }
    const float d_in = norm_denominator(warp_max(m_in));
    const float d_ut = norm_denominator(warp_max(m_ut));
    const float d_cap = norm_denominator(warp_max(m_cap));
    for (int i = n0 + lane; i < n1; i += 32) {
      x[(int64_t)i * 3 + 0] = clip01(ingress[i] / d_in);
      x[(int64_t)i * 3 + 1] = clip01(utilisation(usage[i], cap[i]) / d_ut);
      x[(int64_t)i * 3 + 2] = clip01(cap[i] / d_cap);
    }
  }
}

}  // namespace gsc

using namespace gsc;

extern "C" int gsc_node_features(const float* ingress, const float* usage, const float* cap,
                                 const int32_t* graph_ptr, int64_t B, float* x, void* stream) {
  GSC_REQUIRE(ingress && usage && cap && graph_ptr && x && B >= 0, "gsc_node_features: bad arguments");
  if (B == 0) return GSC_OK;
  int64_t blocks = ceil_div(B, 8);
  if (blocks > 8LL * kNumSMs) blocks = 8LL * kNumSMs;
  node_features_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(ingress, usage, cap, graph_ptr, B, x);
  GSC_LAUNCH_CHECK("node_features_kernel");
  return GSC_OK;
}
