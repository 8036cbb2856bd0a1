// Dense layer with ONE output unit (the critic's head, reference src/rlsp/agents/models.py:119-122 with
// critic_hidden_layer_nodes = [64] -> Linear(64, 1)), forward and backward in one launch each.
//
// As a GEMM this layer is [B x 1 x K]: one column of output.  Routed through gsc_gemm it cost a SIMT
// launch in the forward pass and, in the backward pass, a dgrad launch, a split-K wgrad launch with its
// finalize launch and a column-sum launch for the bias -- five ~5-10 us launches on the critical path
// of every critic evaluation for 128 KFLOP of work.  Here: one warp per row.
//
//   forward : q[i] = sum_k h[i,k] w[k] + b
//   backward: d_h[i,k] = dq[i] w[k] (h[i,k] > 0 ? 1 : 0 when the input is a ReLU output)
//             dW[k]   = sum_i dq[i] h[i,k]        db = sum_i dq[i]
//   The row-block partials of dW / db are added in block order by the block that finishes last
//   (device counter, reset by that block): deterministic, no second launch.
#include "common.cuh"

namespace gsc {

constexpr int kN1Threads = 256;                  // 8 rows per CTA pass
constexpr int kN1MaxK = 1024;
constexpr int kN1MaxBlocks = 256;

__global__ void __launch_bounds__(kN1Threads)
linear_n1_fwd_kernel(const float* __restrict__ h, int64_t ld_h, int64_t M, int K, const float* __restrict__ w,
                     const float* __restrict__ bias, float* __restrict__ out, int64_t ld_out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float b = bias ? bias[0] : 0.f;
  for (int64_t i = (int64_t)blockIdx.x * (kN1Threads / 32) + warp; i < M; i += (int64_t)gridDim.x * (kN1Threads / 32)) {
    const float* row = h + i * ld_h;
    float s = 0.f;
    for (int k = lane; k < K; k += 32) s = fmaf(row[k], w[k], s);
    s = warp_sum(s);
    if (lane == 0) out[i * ld_out] = s + b;
  }
}

// partials: [gridDim.x][K + 1] (dW | db) ; counter: one int32, zero before the first launch ever
__global__ void __launch_bounds__(kN1Threads)
linear_n1_bwd_kernel(const float* __restrict__ dq, int64_t ld_dq, const float* __restrict__ h, int64_t ld_h, int64_t M,
                     int K, const float* __restrict__ w, int gate, float* __restrict__ d_h, int64_t ld_dh,
                     float* __restrict__ dW, float* __restrict__ db, float* __restrict__ partials,
                     int32_t* __restrict__ counter) {
  __shared__ float acc_s[kN1Threads / 32][kN1MaxK + 1];
  __shared__ int last_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool want_w = dW != nullptr || db != nullptr;
  if (want_w)
    for (int k = lane; k <= K; k += 32) acc_s[warp][k] = 0.f;
  float dbs = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * (kN1Threads / 32) + warp; i < M; i += (int64_t)gridDim.x * (kN1Threads / 32)) {
    const float g = dq[i * ld_dq];
    const float* row = h + i * ld_h;
    for (int k = lane; k < K; k += 32) {
      const float hv = row[k];
      if (d_h) d_h[i * ld_dh + k] = (!gate || hv > 0.f) ? g * w[k] : 0.f;
      if (want_w) acc_s[warp][k] = fmaf(g, hv, acc_s[warp][k]);     // lane-private column k
    }
    dbs += g;                                                        // same value in every lane
  }
  if (!want_w) return;
  if (lane == 0) acc_s[warp][K] = dbs;
  __syncthreads();
  float* mine = partials + (size_t)blockIdx.x * (K + 1);
  for (int k = threadIdx.x; k <= K; k += kN1Threads) {
    float s = 0.f;
#pragma unroll
    for (int wv = 0; wv < kN1Threads / 32; ++wv) s += acc_s[wv][k];
    mine[k] = s;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last_s = (atomicAdd(counter, 1) == (int)gridDim.x - 1);
  __syncthreads();
  if (!last_s) return;
  __threadfence();
  for (int k = threadIdx.x; k <= K; k += kN1Threads) {
    float s = 0.f;
    for (int b = 0; b < (int)gridDim.x; ++b) s += __ldcg(partials + (size_t)b * (K + 1) + k);
    if (k < K) { if (dW) dW[k] = s; }
    else if (db) db[0] = s;
  }
  if (threadIdx.x == 0) *counter = 0;
}

static int n1_grid(int64_t M) {
  int64_t b = ceil_div(M > 0 ? M : 1, 4 * (kN1Threads / 32));     // >= 4 rows per warp
  if (b > kN1MaxBlocks) b = kN1MaxBlocks;
  return (int)(b < 1 ? 1 : b);
}

}  // namespace gsc

using namespace gsc;

extern "C" int gsc_linear_n1_fwd(const float* h, int64_t ld_h, int64_t M, int K, const float* w, const float* bias,
                                 float* out, int64_t ld_out, void* stream) {
  GSC_REQUIRE(h && w && out && M >= 0 && K >= 1 && ld_h >= K && ld_out >= 1, "gsc_linear_n1_fwd: bad arguments");
  if (M == 0) return GSC_OK;
  linear_n1_fwd_kernel<<<n1_grid(M), kN1Threads, 0, (cudaStream_t)stream>>>(h, ld_h, M, K, w, bias, out, ld_out);
  GSC_LAUNCH_CHECK("linear_n1_fwd_kernel");
  return GSC_OK;
}

extern "C" size_t gsc_linear_n1_bwd_workspace_bytes(int K) {
  return align_up((size_t)kN1MaxBlocks * (size_t)(K + 1) * 4, 256) + 256;
}

extern "C" int gsc_linear_n1_bwd(const float* dq, int64_t ld_dq, const float* h, int64_t ld_h, int64_t M, int K,
                                 const float* w, int relu_input, float* d_h, int64_t ld_dh, float* dW, float* db,
                                 void* workspace, size_t workspace_bytes, void* stream) {
  GSC_REQUIRE(dq && h && w && M >= 0 && K >= 1 && K <= kN1MaxK && ld_h >= K && ld_dq >= 1 && (!d_h || ld_dh >= K),
              "gsc_linear_n1_bwd: bad arguments (K <= %d)", kN1MaxK);
  const bool want_w = dW != nullptr || db != nullptr;
  if (want_w) {
    GSC_REQUIRE(workspace != nullptr, "gsc_linear_n1_bwd: weight gradients need the workspace");
    if (workspace_bytes < gsc_linear_n1_bwd_workspace_bytes(K))
      return fail(GSC_EWORKSPACE, "gsc_linear_n1_bwd: workspace %zu < %zu", workspace_bytes,
                  gsc_linear_n1_bwd_workspace_bytes(K));
  }
  if (M == 0) {
    if (dW) GSC_CUDA_TRY(cudaMemsetAsync(dW, 0, (size_t)K * 4, (cudaStream_t)stream));
    if (db) GSC_CUDA_TRY(cudaMemsetAsync(db, 0, 4, (cudaStream_t)stream));
    return GSC_OK;
  }
  // workspace: [counter (256 B, must read zero at the first launch: see the header)] [partials]
  int32_t* counter = static_cast<int32_t*>(workspace);
  float* partials = workspace ? reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) + 256) : nullptr;
  linear_n1_bwd_kernel<<<n1_grid(M), kN1Threads, 0, (cudaStream_t)stream>>>(dq, ld_dq, h, ld_h, M, K, w, relu_input, d_h,
                                                                           ld_dh, dW, db, partials, counter);
  GSC_LAUNCH_CHECK("linear_n1_bwd_kernel");
  return GSC_OK;
}
