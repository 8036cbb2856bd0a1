// Dense layers of the PyG MLP (reference src/rlsp/agents/models.py:80-83,119-122) -- exact-fp32
// path.  One strided, K-segmented GEMM serves forward (Linear over the never-materialised
// cat([emb, mask, a])), dgrad and wgrad; see include/gsc_b200.h for the operand conventions.
//
// Kernel: 64x64 output tile per CTA, BK = 16, 256 threads x (4x4) register tile, operands staged
// through shared memory ([k][m] / [k][n] so the inner product reads 128-bit rows), next tile
// prefetched into registers while the current one is multiplied.  Shapes with few output tiles
// and a long reduction (Linear(22+2A -> 64), dgrad over A) are split along K across CTAs into
// the caller's workspace and reduced in a fixed order by gemm_finalize_kernel (deterministic).
// Algorithmic bytes: every A and B element once per tile row/column, C once.
#include "common.cuh"

namespace gsc {

// gemm_tc.cu
bool gemm_tc_supported(const gsc_gemm_args* a, int64_t M, int64_t N, bool* a_mn, bool* b_mn);
int gemm_tc_launch(const gsc_gemm_args* a, int64_t M, int64_t N, float* C, int64_t ldc, void* workspace,
                   size_t workspace_bytes, bool a_mn, bool b_mn, bool x3, cudaStream_t st, int* used_splits);

constexpr int kBM = 64, kBN = 64, kBK = 16, kGemmThreads = 256, kPad = 4;

struct GemmDev {
  const float* A[3];
  const float* B[3];
  int64_t K[3], a_rs[3], a_cs[3], b_rs[3], b_cs[3];
  int tile_end[3];       // exclusive prefix of k-tiles per segment
  int num_seg, flags;
  const float* bias;
  const float* mul;
  const float* gate;
  int64_t ld_mul, ld_gate;
  const float* row_table;
  const int32_t* row_n;
  int64_t ld_table;
  int n_max, groups;
  int64_t M, N, ldc;
  float* C;
  float* partial;        // [splits][M][N] when splits > 1
  int tiles_total, tiles_per_split, splits;
  const float* dot_w;    // optional one-unit head on the finished rows (gsc_gemm_args::dot_*)
  const float* dot_bias;
  float* dot_out;
};

__device__ __forceinline__ float apply_epilogue(const GemmDev& g, float v, int64_t i, int64_t j) {
  if (g.flags & GSC_EPI_ROWTABLE) v += g.row_table[(int64_t)g.row_n[i] * g.ld_table + j];
  if (g.flags & GSC_EPI_BIAS) v += g.bias[j];
  if (g.flags & GSC_EPI_RELU) v = fmaxf(v, 0.f);
  if (g.flags & GSC_EPI_MUL) v *= g.mul[i * g.ld_mul + j];
  if (g.flags & GSC_EPI_SMASK) {
    const int nb = g.row_n[i];
    const uint32_t ju = (uint32_t)j;            // N < 2^31 is checked on entry
    const int src = (int)(ju / ((uint32_t)g.groups * (uint32_t)g.n_max)), dst = (int)(ju % (uint32_t)g.n_max);
    v = (src < nb && dst < nb) ? v : 0.f;
  }
  if (g.flags & GSC_EPI_GATE) v = g.gate[i * g.ld_gate + j] > 0.f ? v : 0.f;
  if (g.flags & GSC_EPI_CLAMP1) v = fminf(fmaxf(v, -1.f), 1.f);
  return v;
}

__global__ void __launch_bounds__(kGemmThreads) gemm_simt_kernel(const GemmDev g) {
  __shared__ __align__(16) float As[kBK][kBM + kPad];
  __shared__ __align__(16) float Bs[kBK][kBN + kPad];
  const int tid = threadIdx.x;
  const int64_t m0 = (int64_t)blockIdx.y * kBM, n0 = (int64_t)blockIdx.x * kBN;
  const int t_begin = blockIdx.z * g.tiles_per_split;
  const int t_end = min(g.tiles_total, t_begin + g.tiles_per_split);
  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads, 4x4 outputs each

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  float ra[4], rb[4];
  auto load_regs = [&](int t) {
    int s = 0;
    while (s < g.num_seg - 1 && t >= g.tile_end[s]) ++s;
    const int64_t k0 = (int64_t)(t - (s ? g.tile_end[s - 1] : 0)) * kBK;
    const int64_t Ks = g.K[s];
    const float* Ap = g.A[s];
    const float* Bp = g.B[s];
    const int64_t ars = g.a_rs[s], acs = g.a_cs[s], brs = g.b_rs[s], bcs = g.b_cs[s];
    if (acs == 1) {          // k-contiguous rows: threads run along k
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int k = tid & 15, i = (tid >> 4) + 16 * r;
        const int64_t gi = m0 + i, gk = k0 + k;
        ra[r] = (gi < g.M && gk < Ks) ? Ap[gi * ars + gk] : 0.f;
      }
    } else {                 // m-contiguous: threads run along m
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int i = tid & 63, k = (tid >> 6) + 4 * r;
        const int64_t gi = m0 + i, gk = k0 + k;
        ra[r] = (gi < g.M && gk < Ks) ? Ap[gi * ars + gk * acs] : 0.f;
      }
    }
    if (brs == 1) {          // k-contiguous columns (weight[N, K] used as B^T)
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int k = tid & 15, j = (tid >> 4) + 16 * r;
        const int64_t gj = n0 + j, gk = k0 + k;
        rb[r] = (gj < g.N && gk < Ks) ? Bp[gk + gj * bcs] : 0.f;
      }
    } else {                 // n-contiguous
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int j = tid & 63, k = (tid >> 6) + 4 * r;
        const int64_t gj = n0 + j, gk = k0 + k;
        rb[r] = (gj < g.N && gk < Ks) ? Bp[gk * brs + gj * bcs] : 0.f;
      }
    }
    return s;
  };
  auto store_smem = [&](int s) {
    if (g.a_cs[s] == 1) {
#pragma unroll
      for (int r = 0; r < 4; ++r) As[tid & 15][(tid >> 4) + 16 * r] = ra[r];
    } else {
#pragma unroll
      for (int r = 0; r < 4; ++r) As[(tid >> 6) + 4 * r][tid & 63] = ra[r];
    }
    if (g.b_rs[s] == 1) {
#pragma unroll
      for (int r = 0; r < 4; ++r) Bs[tid & 15][(tid >> 4) + 16 * r] = rb[r];
    } else {
#pragma unroll
      for (int r = 0; r < 4; ++r) Bs[(tid >> 6) + 4 * r][tid & 63] = rb[r];
    }
  };

  if (t_begin < t_end) {
    int seg = load_regs(t_begin);
    for (int t = t_begin; t < t_end; ++t) {
      store_smem(seg);
      __syncthreads();
      if (t + 1 < t_end) seg = load_regs(t + 1);
#pragma unroll
      for (int k = 0; k < kBK; ++k) {
        const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
        const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
        const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
      }
      __syncthreads();
    }
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t gi = m0 + ty * 4 + i;
    if (gi >= g.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t gj = n0 + tx * 4 + j;
      if (gj >= g.N) continue;
      if (g.splits == 1) g.C[gi * g.ldc + gj] = apply_epilogue(g, acc[i][j], gi, gj);
      else g.partial[((int64_t)blockIdx.z * g.M + gi) * g.N + gj] = acc[i][j];
    }
  }
}

__global__ void gemm_finalize_kernel(const GemmDev g) {
  const int64_t total = g.M * g.N;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int z = 0; z < g.splits; ++z) s += g.partial[(int64_t)z * total + t];
    const int64_t i = t / g.N, j = t - i * g.N;
    g.C[i * g.ldc + j] = apply_epilogue(g, s, i, j);
  }
}

// Split-K finalize with a one-unit head on top: one warp per row adds the partials of the row, applies the
// epilogue, stores C and folds the finished values into dot_out[i] = sum_j C[i, j] dot_w[j] + dot_bias[0]
// (same summation order as gemm_finalize_kernel followed by linear_n1_fwd_kernel: lane-strided FMA chain,
// then the shuffle tree).
__global__ void __launch_bounds__(256) gemm_finalize_rowdot_kernel(const GemmDev g) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t total = g.M * g.N;
  const float b = g.dot_bias ? g.dot_bias[0] : 0.f;
  for (int64_t i = warp; i < g.M; i += warps) {
    float dot = 0.f;
    for (int64_t j = lane; j < g.N; j += 32) {
      const int64_t t = i * g.N + j;
      float s = 0.f;
      for (int z = 0; z < g.splits; ++z) s += g.partial[(int64_t)z * total + t];
      const float v = apply_epilogue(g, s, i, j);
      g.C[i * g.ldc + j] = v;
      dot = fmaf(v, g.dot_w[j], dot);
    }
    dot = warp_sum(dot);
    if (lane == 0) g.dot_out[i] = dot + b;
  }
}

__global__ void __launch_bounds__(1024) colsum_kernel(const float* __restrict__ X, int64_t ld, int64_t M, int64_t N,
                                                      float* __restrict__ out) {
  // block = 32 columns x 32 row-lanes; every thread keeps four independent partial sums (the loop is
  // latency-bound otherwise), the 32 row-lanes are reduced through shared memory in a fixed order.
  __shared__ float red[32][33];
  const int c = threadIdx.x & 31, r = threadIdx.x >> 5;
  const int64_t j = (int64_t)blockIdx.x * 32 + c;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if (j < N) {
    int64_t i = r;
    for (; i + 96 < M; i += 128) {
      s0 += X[i * ld + j];
      s1 += X[(i + 32) * ld + j];
      s2 += X[(i + 64) * ld + j];
      s3 += X[(i + 96) * ld + j];
    }
    for (; i < M; i += 32) s0 += X[i * ld + j];
  }
  red[r][c] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (r == 0 && j < N) {
    float tot = 0.f;
#pragma unroll
    for (int k = 0; k < 32; ++k) tot += red[k][c];
    out[j] = tot;
  }
}

// Upper bound on the number of K-splits for this shape (monotone in tiles_total).
static int64_t max_splits(int64_t M, int64_t N, int tiles_total) {
  const int64_t tiles_mn = ceil_div(M, kBM) * ceil_div(N, kBN);
  int64_t want = ceil_div(4 * kNumSMs, tiles_mn);
  int64_t max_by_k = tiles_total / 8;       // at least 8 k-tiles (128 k) per split
  if (max_by_k < 1) max_by_k = 1;
  if (want > max_by_k) want = max_by_k;
  return want < 1 ? 1 : want;
}

static void plan_split(int64_t M, int64_t N, int tiles_total, int* splits, int* tps) {
  const int64_t want = max_splits(M, N, tiles_total);
  int t = (int)ceil_div(tiles_total, want);
  if (t < 1) t = 1;
  *tps = t;
  *splits = (int)ceil_div(tiles_total > 0 ? tiles_total : 1, t);
}

}  // namespace gsc

using namespace gsc;

extern "C" size_t gsc_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K_total) {
  // upper bound over any segmentation of K_total into <= 3 segments
  if (M <= 0 || N <= 0) return 0;
  const int tiles = (int)(ceil_div(K_total > 0 ? K_total : 1, kBK) + 2);
  int64_t s = max_splits(M, N, tiles);
  // tensor-core path: 128x128 tiles, up to one split per SM, >= 4 k-blocks of 32 per split
  const int64_t tc_tiles = ceil_div(M, 128) * ceil_div(N, 128);
  int64_t tc = tc_tiles >= kNumSMs ? 1 : ceil_div(kNumSMs, tc_tiles);
  const int64_t tc_by_k = (ceil_div(K_total > 0 ? K_total : 1, 32) + 2) / 4;
  if (tc > tc_by_k) tc = tc_by_k;
  if (tc > s) s = tc;
  return s > 1 ? (size_t)s * (size_t)M * (size_t)N * 4 : 0;
}

extern "C" int gsc_gemm(const gsc_gemm_args* a, int64_t M, int64_t N, float* C, int64_t ldc,
                        void* workspace, size_t workspace_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  GSC_REQUIRE(a && C, "gsc_gemm: null args/C");
  GSC_REQUIRE(M >= 0 && N >= 0 && a->num_seg >= 1 && a->num_seg <= 3, "gsc_gemm: bad M/N/num_seg");
  if (M == 0 || N == 0) return GSC_OK;
  GSC_REQUIRE(ceil_div(M, kBM) <= 65535, "gsc_gemm: M=%lld too large for the grid", (long long)M);
  GemmDev g{};
  int tiles = 0;
  for (int s = 0; s < a->num_seg; ++s) {
    GSC_REQUIRE(a->K[s] >= 0 && (a->K[s] == 0 || (a->A[s] && a->B[s])), "gsc_gemm: segment %d invalid", s);
    g.A[s] = a->A[s]; g.B[s] = a->B[s]; g.K[s] = a->K[s];
    g.a_rs[s] = a->a_rs[s]; g.a_cs[s] = a->a_cs[s]; g.b_rs[s] = a->b_rs[s]; g.b_cs[s] = a->b_cs[s];
    tiles += (int)ceil_div(a->K[s], kBK);
    g.tile_end[s] = tiles;
  }
  g.num_seg = a->num_seg; g.flags = a->flags;
  g.bias = a->bias; g.mul = a->mul; g.gate = a->gate; g.ld_mul = a->ld_mul; g.ld_gate = a->ld_gate;
  g.row_table = a->row_table; g.row_n = a->row_n; g.ld_table = a->ld_table; g.n_max = a->n_max; g.groups = a->groups;
  GSC_REQUIRE(!(g.flags & GSC_EPI_ROWTABLE) || (g.row_table && g.row_n), "gsc_gemm: GSC_EPI_ROWTABLE without table/row_n");
  GSC_REQUIRE(!(g.flags & GSC_EPI_SMASK) || (g.row_n && g.n_max > 0 && g.groups > 0), "gsc_gemm: GSC_EPI_SMASK without row_n/shape");
  GSC_REQUIRE(!(g.flags & GSC_EPI_SMASK) || N < (int64_t)1 << 31, "gsc_gemm: GSC_EPI_SMASK needs N < 2^31");
  GSC_REQUIRE(!(g.flags & GSC_EPI_BIAS) || g.bias, "gsc_gemm: GSC_EPI_BIAS without bias");
  GSC_REQUIRE(!(g.flags & GSC_EPI_MUL) || g.mul, "gsc_gemm: GSC_EPI_MUL without mul");
  GSC_REQUIRE(!(g.flags & GSC_EPI_GATE) || g.gate, "gsc_gemm: GSC_EPI_GATE without gate");
  g.M = M; g.N = N; g.ldc = ldc; g.C = C;
  g.dot_w = a->dot_w; g.dot_bias = a->dot_bias; g.dot_out = a->dot_out;
  GSC_REQUIRE(!g.dot_out || (g.dot_w && N <= 1024), "gsc_gemm: dot_out needs dot_w and N <= 1024");
  // the reduction of split-K partials (with the optional head), or the head alone behind an unsplit GEMM
  auto finish = [&](int splits) -> int {
    if (splits > 1) {
      if (g.dot_out) {
        int64_t blocks = ceil_div(M, 8);
        if (blocks > 8 * kNumSMs) blocks = 8 * kNumSMs;
        gemm_finalize_rowdot_kernel<<<(int)blocks, 256, 0, st>>>(g);
        GSC_LAUNCH_CHECK("gemm_finalize_rowdot_kernel");
      } else {
        int64_t blocks = ceil_div(M * N, 256);
        if (blocks > 8 * kNumSMs) blocks = 8 * kNumSMs;
        gemm_finalize_kernel<<<(int)blocks, 256, 0, st>>>(g);
        GSC_LAUNCH_CHECK("gemm_finalize_kernel");
      }
    } else if (g.dot_out) {
      return gsc_linear_n1_fwd(C, ldc, M, (int)N, g.dot_w, g.dot_bias, g.dot_out, 1, (void*)st);
    }
    return GSC_OK;
  };
  if (a->flags & (GSC_EPI_TF32 | GSC_EPI_TF32X3)) {
    bool a_mn = false, b_mn = false;
    // tiny problems stay on the exact path: the tensor-core pipeline has a fixed fill cost
    if ((double)M * (double)N * (double)tiles * kBK >= 4.0e6 && gemm_tc_supported(a, M, N, &a_mn, &b_mn)) {
      int splits = 1;
      int rc = gemm_tc_launch(a, M, N, C, ldc, workspace, workspace_bytes, a_mn, b_mn,
                              (a->flags & GSC_EPI_TF32X3) != 0, st, &splits);
      if (rc) return rc;
      g.splits = splits;
      g.partial = (float*)workspace;
      return finish(splits);
    }
  }
  g.tiles_total = tiles;
  plan_split(M, N, tiles, &g.splits, &g.tiles_per_split);
  if (g.splits > 1) {
    const size_t need = (size_t)g.splits * (size_t)M * (size_t)N * 4;
    if (workspace == nullptr || workspace_bytes < need)
      return fail(GSC_EWORKSPACE, "gsc_gemm: workspace %zu < %zu (M=%lld N=%lld splits=%d)",
                  workspace_bytes, need, (long long)M, (long long)N, g.splits);
    g.partial = (float*)workspace;
  }
  dim3 grid((unsigned)ceil_div(N, kBN), (unsigned)ceil_div(M, kBM), (unsigned)g.splits);
  gemm_simt_kernel<<<grid, kGemmThreads, 0, st>>>(g);
  GSC_LAUNCH_CHECK("gemm_simt_kernel");
  return finish(g.splits);
}

extern "C" int gsc_colsum(const float* X, int64_t ld, int64_t M, int64_t N, float* out, void* stream) {
  GSC_REQUIRE(X && out && M >= 0 && N >= 1, "gsc_colsum: bad arguments");
  colsum_kernel<<<(int)ceil_div(N, 32), 1024, 0, (cudaStream_t)stream>>>(X, ld, M, N, out);
  GSC_LAUNCH_CHECK("colsum_kernel");
  return GSC_OK;
}
