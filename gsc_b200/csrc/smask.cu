// Structured-mask shortcuts for the first dense layer.
//
// Every observation's mask is mask[:n, :, :, :n] = 1 over the scheduling shape (N_max, G, N_max)
// (reference src/rlsp/envs/simulator_wrapper.py:302-306), i.e. a function of the node count n only.
// For such batches the mask column block of Linear(cat[emb, mask(, a)]) (reference
// src/rlsp/agents/models.py:90-91,150) needs neither the [B, A] mask nor a GEMM:
//   forward : mask_b . W1m[o, :]  = T[n_b][o],  T[n][o] = sum over {(i,g,j): i<n, j<n} of W1m[o, (i*G+g)*N + j]
//   wgrad   : dW1m[o, k]          = S[max(i(k), j(k))][o],  S[t][o] = sum over {b: n_b > t} of dY[b, o]
// The table costs one read of W1m per weight update (instead of reading the [B, A] mask and W1m
// per forward), the weight gradient is one coalesced write of W1m's shape.  Exact fp32 sums.
#include "common.cuh"

namespace gsc {

// One CTA per output row o.  Warp w owns the rows i = w, w + warps, ... of the N x G x N block: it
// sums the G groups of row i (coalesced 128 B loads) into C[i][j] on the fly and splits the row at the
// diagonal: the j <= i part is a warp sum that belongs to the "shell" D[i], the j > i part is added to
// the warp's private column accumulators (D[j], j > i).  T[n] = sum_{k < n} D[k] (block prefix sum).
// No atomics, fixed summation order.
constexpr int kTableThreads = 512;

__global__ void __launch_bounds__(kTableThreads)
mask_table_kernel(const float* __restrict__ W, int64_t ldw, int n_max, int groups, float* __restrict__ table,
                  int64_t ld_table) {
  extern __shared__ float smem[];
  constexpr int kWarps = kTableThreads / 32;
  float* colpart = smem;                                   // [kWarps][n_max]
  float* D = smem + (size_t)kWarps * n_max;                // [n_max] row parts, then the shells
  float* scan = D + n_max;                                 // [kWarps] block-scan carries
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float* w = W + (int64_t)blockIdx.x * ldw;
  float* mycol = colpart + (size_t)warp * n_max;
  for (int j = lane; j < n_max; j += 32) mycol[j] = 0.f;
#pragma unroll 2                                            // two rows' loads in flight: the kernel is latency-bound
  for (int i = warp; i < n_max; i += kWarps) {
    const float* wi = w + (int64_t)i * groups * n_max;
    float rowsum = 0.f;
    for (int j0 = 0; j0 < n_max; j0 += 128) {              // 4 independent column groups per pass
      float c[4] = {0.f, 0.f, 0.f, 0.f};
      for (int g = 0; g < groups; ++g) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int j = j0 + 32 * u + lane;
          if (j < n_max) c[u] += wi[(int64_t)g * n_max + j];
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int j = j0 + 32 * u + lane;
        if (j < n_max) {
          if (j <= i) rowsum += c[u];
          else mycol[j] += c[u];
        }
      }
    }
    rowsum = warp_sum(rowsum);
    if (lane == 0) D[i] = rowsum;
  }
  __syncthreads();
  // shell k = row part + column parts of all warps; inclusive block scan over k (n_max may exceed the block)
  float carry = 0.f;
  for (int k0 = 0; k0 < n_max; k0 += kTableThreads) {
    const int k = k0 + threadIdx.x;
    float v = 0.f;
    if (k < n_max) {
      v = D[k];
      for (int q = 0; q < kWarps; ++q) v += colpart[(size_t)q * n_max + k];
    }
    float x = v;                                           // warp inclusive scan
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const float y = __shfl_up_sync(0xffffffffu, x, d);
      if (lane >= d) x += y;
    }
    if (lane == 31) scan[warp] = x;
    __syncthreads();
    float base = carry;
    for (int q = 0; q < warp; ++q) base += scan[q];
    if (k < n_max) table[(int64_t)(k + 1) * ld_table + blockIdx.x] = base + x;
    float total = 0.f;
    for (int q = 0; q < kWarps; ++q) total += scan[q];
    carry += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) table[blockIdx.x] = 0.f;
}

// S_part[z][o][t] = sum over {b in row slice z: n_b > t} dY[b, o],  t = 0..n_max-1.
// A CTA owns 32 columns and one slice of the batch; warp w adds its rows b = w (mod warps) into
// private shared-memory bins[w][n_b][lane] (B additions per column instead of B * n_max, no atomics);
// the bins are summed over the warps and suffix-summed over n.  The slices are added by the expand
// kernel in a fixed order, so the result is bit-reproducible.
__global__ void mask_wgrad_bins_kernel(const float* __restrict__ dY, int64_t ldy, const int32_t* __restrict__ n,
                                       int64_t B, int64_t N1, int n_max, int64_t rows_per_slice,
                                       float* __restrict__ S_part) {
  extern __shared__ float bins[];        // [warps][n_max + 1][32]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
  const int64_t o = (int64_t)blockIdx.x * 32 + lane;
  const int64_t b0 = (int64_t)blockIdx.y * rows_per_slice;
  const int64_t b1 = b0 + rows_per_slice < B ? b0 + rows_per_slice : B;
  float* mine = bins + (size_t)warp * (n_max + 1) * 32;
  for (int t = 0; t <= n_max; ++t) mine[t * 32 + lane] = 0.f;
  if (o < N1)
    for (int64_t b = b0 + warp; b < b1; b += warps) {
      int nb = n[b];
      nb = nb < 0 ? 0 : (nb > n_max ? n_max : nb);
      mine[nb * 32 + lane] += dY[b * ldy + o];
    }
  __syncthreads();
  // sum over the warps: thread (warp, lane) takes the bins t = warp, warp + warps, ...
  for (int t = warp; t <= n_max; t += warps) {
    float v = 0.f;
    for (int w = 0; w < warps; ++w) v += bins[((size_t)w * (n_max + 1) + t) * 32 + lane];
    bins[(size_t)t * 32 + lane] = v;            // into warp 0's bins (its own slot t is read before it is written)
  }
  __syncthreads();
  if (warp == 0 && o < N1) {
    float* out = S_part + ((int64_t)blockIdx.y * N1 + o) * n_max;
    float run = 0.f;
    for (int t = n_max - 1; t >= 0; --t) {      // S[t] = sum_{n > t} bins[n]
      run += bins[(size_t)(t + 1) * 32 + lane];
      out[t] = run;
    }
  }
}

// dW[o, (i*G + g)*N + j] = S[o][max(i, j)]: blockIdx.x = o, blockIdx.y strides over the i*G + g rows.
__global__ void __launch_bounds__(256)
mask_wgrad_expand_kernel(const float* __restrict__ S_part, int slices, int64_t N1, int n_max, int groups,
                         float* __restrict__ dW, int64_t ldw) {
  extern __shared__ float s_col[];                         // S[o][:] = sum of the batch slices
  for (int t = threadIdx.x; t < n_max; t += blockDim.x) {
    float v = 0.f;
    for (int z = 0; z < slices; ++z) v += S_part[((int64_t)z * N1 + blockIdx.x) * n_max + t];
    s_col[t] = v;
  }
  __syncthreads();
  float* out = dW + (int64_t)blockIdx.x * ldw;
  const int rows = n_max * groups;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
  for (int r = blockIdx.y * warps + warp; r < rows; r += gridDim.y * warps) {
    const int i = r / groups;
    const float si = s_col[i];
    float* dst = out + (int64_t)r * n_max;
    for (int j = lane; j < n_max; j += 32) dst[j] = j <= i ? si : s_col[j];
  }
}

}  // namespace gsc

using namespace gsc;

extern "C" int gsc_mask_table(const float* W, int64_t ldw, int64_t N1, int n_max, int groups, float* table,
                              int64_t ld_table, void* stream) {
  GSC_REQUIRE(W && table && N1 >= 1 && n_max >= 1 && groups >= 1 && ld_table >= N1, "gsc_mask_table: bad arguments");
  const size_t smem = ((size_t)(kTableThreads / 32 + 1) * n_max + kTableThreads / 32) * 4;
  GSC_REQUIRE(smem <= 200 * 1024, "gsc_mask_table: n_max=%d needs %zu B of shared memory", n_max, smem);
  if (smem > 48 * 1024)
    GSC_CUDA_TRY(cudaFuncSetAttribute(mask_table_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  mask_table_kernel<<<(int)N1, kTableThreads, smem, (cudaStream_t)stream>>>(W, ldw, n_max, groups, table, ld_table);
  GSC_LAUNCH_CHECK("mask_table_kernel");
  return GSC_OK;
}

// Batch slices of the binning pass: enough CTAs to cover the SMs, at least 64 rows each.
static int mask_wgrad_slices(int64_t B, int64_t N1) {
  int64_t z = ceil_div((int64_t)kNumSMs, ceil_div(N1, 32));
  const int64_t by_rows = ceil_div(B > 0 ? B : 1, 64);
  if (z > by_rows) z = by_rows;
  if (z > 32) z = 32;
  return z < 1 ? 1 : (int)z;
}

extern "C" size_t gsc_mask_wgrad_workspace_bytes(int64_t N1, int n_max) {
  return (size_t)32 * (size_t)N1 * (size_t)n_max * 4;       // up to 32 batch slices of S
}

extern "C" int gsc_mask_wgrad(const float* dY, int64_t ldy, const int32_t* n, int64_t B, int64_t N1, int n_max,
                              int groups, float* dW, int64_t ldw, void* workspace, size_t workspace_bytes,
                              void* stream) {
  GSC_REQUIRE(dY && n && dW && workspace && B >= 0 && N1 >= 1 && n_max >= 1 && groups >= 1,
              "gsc_mask_wgrad: bad arguments");
  GSC_REQUIRE(n_max <= 48 * 1024 / 4, "gsc_mask_wgrad: n_max=%d too large", n_max);
  if (workspace_bytes < gsc_mask_wgrad_workspace_bytes(N1, n_max))
    return fail(GSC_EWORKSPACE, "gsc_mask_wgrad: workspace %zu < %zu", workspace_bytes,
                gsc_mask_wgrad_workspace_bytes(N1, n_max));
  int warps = (int)((200 * 1024) / ((size_t)(n_max + 1) * 32 * 4));
  if (warps > 8) warps = 8;
  GSC_REQUIRE(warps >= 1, "gsc_mask_wgrad: n_max=%d too large", n_max);
  const size_t smem = (size_t)warps * (n_max + 1) * 32 * 4;
  if (smem > 48 * 1024)
    GSC_CUDA_TRY(cudaFuncSetAttribute(mask_wgrad_bins_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  float* S = (float*)workspace;
  const int slices = mask_wgrad_slices(B, N1);
  const int64_t rows_per_slice = ceil_div(B > 0 ? B : 1, (int64_t)slices);
  const dim3 gb((unsigned)ceil_div(N1, 32), (unsigned)slices);
  mask_wgrad_bins_kernel<<<gb, warps * 32, smem, (cudaStream_t)stream>>>(dY, ldy, n, B, N1, n_max, rows_per_slice, S);
  GSC_LAUNCH_CHECK("mask_wgrad_bins_kernel");
  int64_t ysplit = ceil_div(4LL * kNumSMs, N1);             // ~4 CTAs per SM in total
  const int64_t row_groups = ceil_div((int64_t)n_max * groups, 8);
  if (ysplit > row_groups) ysplit = row_groups;
  if (ysplit < 1) ysplit = 1;
  const dim3 ge((unsigned)N1, (unsigned)ysplit);
  mask_wgrad_expand_kernel<<<ge, 256, (size_t)n_max * 4, (cudaStream_t)stream>>>(S, slices, N1, n_max, groups, dW, ldw);
  GSC_LAUNCH_CHECK("mask_wgrad_expand_kernel");
  return GSC_OK;
}
