// Batch packer: int64 COO (PyG Batch layout) -> int32 CSR-by-target with one self-loop per node.
//
// Replaces the graph part of Batch.from_data_list (reference src/rlsp/agents/buffer.py:65-81) and
// the remove_self_loops + add_self_loops that GATv2Conv.forward repeats on every call (reference
// src/rlsp/agents/models.py:38,47 -> PyG 2.3.0).  Done once per minibatch, on the device.
//
// HBM traffic: reads 16 B/edge (twice: count + fill) + 8 B/node, writes 4 B/edge + 4 B/node.
// All kernels are flat, coalesced, grid-stride; the row sort makes the neighbour order (and with
// it every downstream fp32 sum) independent of the atomic fill order.
#include "common.cuh"
#include <mutex>

namespace gsc {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 4;
constexpr int kScanTile = kScanThreads * kScanItems;  // 1024 rows per scan block

__device__ __forceinline__ void pack_graph_ptr_body(const int64_t* __restrict__ batch, int64_t Nt, int64_t B,
                                                    int32_t* __restrict__ graph_ptr, int32_t* __restrict__ stats) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (batch == nullptr) {
    if (i == 0) { graph_ptr[0] = 0; for (int64_t g = 1; g <= B; ++g) graph_ptr[g] = (int32_t)Nt; }
    return;
  }
  if (Nt == 0) {
    for (; i <= B; i += (int64_t)gridDim.x * blockDim.x) graph_ptr[i] = 0;
    return;
  }
  for (; i < Nt; i += (int64_t)gridDim.x * blockDim.x) {
    int64_t b = batch[i];
    int64_t prev = (i > 0) ? batch[i - 1] : -1;
    if (b < 0 || b >= B) { atomicOr(stats + 2, 2); continue; }
    if (b < prev) { atomicOr(stats + 2, 4); continue; }
    for (int64_t g = prev + 1; g <= b; ++g) graph_ptr[g] = (int32_t)i;  // empty graphs too
    if (i == Nt - 1)
      for (int64_t g = b + 1; g <= B; ++g) graph_ptr[g] = (int32_t)Nt;
  }
}
__global__ void pack_init_kernel(int32_t* __restrict__ deg, int64_t Nt, int32_t* __restrict__ stats) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) { stats[0] = 0; stats[1] = 0; stats[2] = 0; stats[3] = 0; }
  for (; i < Nt; i += (int64_t)gridDim.x * blockDim.x) deg[i] = 1;  // the self-loop
}

// In-degrees; the same launch derives graph_ptr from the (sorted) batch vector -- both only need the
// stats words zeroed by pack_init_kernel.
__global__ void pack_count_kernel(const int64_t* __restrict__ ei, int64_t E, int64_t Nt,
                                  const int64_t* __restrict__ batch, int32_t* __restrict__ deg,
                                  int32_t* __restrict__ stats, int64_t B, int32_t* __restrict__ graph_ptr) {
  pack_graph_ptr_body(batch, Nt, B, graph_ptr, stats);
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E;
       e += (int64_t)gridDim.x * blockDim.x) {
    int64_t s = ei[e], d = ei[E + e];
    if (s < 0 || s >= Nt || d < 0 || d >= Nt) { atomicOr(stats + 2, 1); continue; }
    if (batch != nullptr && batch[s] != batch[d]) { atomicOr(stats + 2, 8); continue; }
    if (s != d) atomicAdd(deg + d, 1);
  }
}

// Block-local exclusive scan of deg -> rowptr, block totals -> blocksum.
__global__ void __launch_bounds__(kScanThreads)
pack_scan_local_kernel(const int32_t* __restrict__ deg, int64_t Nt, int32_t* __restrict__ rowptr,
                       int32_t* __restrict__ blocksum) {
  __shared__ int32_t warp_tot[kScanThreads / 32];
  int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  int32_t v[kScanItems];
  int32_t tsum = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    v[k] = (base + k < Nt) ? deg[base + k] : 0;
    tsum += v[k];
  }
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int32_t incl = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  int32_t woff = 0;
  for (int w = 0; w < warp; ++w) woff += warp_tot[w];
  int32_t run = woff + incl - tsum;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    if (base + k < Nt) rowptr[base + k] = run;
    run += v[k];
  }
  if (threadIdx.x == kScanThreads - 1) blocksum[blockIdx.x] = run;
}

// Single block: exclusive scan of the block totals in place; total -> blocksum[nb].
__global__ void __launch_bounds__(1024) pack_scan_blocks_kernel(int32_t* __restrict__ blocksum, int nb) {
  __shared__ int32_t warp_tot[32];
  __shared__ int32_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int start = 0; start < nb; start += 1024) {
    int i = start + threadIdx.x;
    int32_t v = (i < nb) ? blocksum[i] : 0;
    int32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    int32_t woff = 0;
    for (int w = 0; w < warp; ++w) woff += warp_tot[w];
    int32_t carry = carry_s;
    if (i < nb) blocksum[i] = carry + woff + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = carry + woff + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) blocksum[nb] = carry_s;
}

// Small batches (nb <= kFusedScanBlocks scan tiles): one launch instead of three.  Block b adds up the tile
// totals in front of it itself (nb loads per block), finishes its tile of rowptr, zeroes the cursors;
// block 0 also writes the grand total, and every block folds its share of the graphs into max_nodes.
constexpr int kFusedScanBlocks = 2048;
__global__ void __launch_bounds__(kScanThreads)
pack_scan_finish_kernel(int32_t* __restrict__ rowptr, int32_t* __restrict__ cursor, const int32_t* __restrict__ blocksum,
                        int64_t Nt, int nb, int32_t* __restrict__ stats, const int32_t* __restrict__ graph_ptr, int64_t B) {
  __shared__ int32_t red[kScanThreads / 32];
  __shared__ int32_t offs_s[2];
  int32_t before = 0, all = 0;
  for (int j = threadIdx.x; j < nb; j += kScanThreads) {
    const int32_t v = blocksum[j];
    all += v;
    if (j < (int)blockIdx.x) before += v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    before += __shfl_xor_sync(0xffffffffu, before, o);
    all += __shfl_xor_sync(0xffffffffu, all, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) red[warp] = before;
  __syncthreads();
  if (threadIdx.x == 0) { int32_t t = 0; for (int w = 0; w < kScanThreads / 32; ++w) t += red[w]; offs_s[0] = t; }
  __syncthreads();
  if (lane == 0) red[warp] = all;
  __syncthreads();
  if (threadIdx.x == 0) { int32_t t = 0; for (int w = 0; w < kScanThreads / 32; ++w) t += red[w]; offs_s[1] = t; }
  __syncthreads();
  const int32_t off = offs_s[0];
  if (blockIdx.x == 0 && threadIdx.x == 0) { rowptr[Nt] = offs_s[1]; stats[0] = offs_s[1]; }
  const int64_t base = (int64_t)blockIdx.x * kScanTile;
  for (int k = threadIdx.x; k < kScanTile; k += kScanThreads)
    if (base + k < Nt) { rowptr[base + k] += off; cursor[base + k] = 0; }
  int32_t m = 0;
  for (int64_t g = (int64_t)blockIdx.x * kScanThreads + threadIdx.x; g < B; g += (int64_t)gridDim.x * kScanThreads)
    m = max(m, graph_ptr[g + 1] - graph_ptr[g]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (lane == 0 && m > 0) atomicMax(stats + 1, m);
}

// rowptr += block offset; cursor (the deg array) := 0; rowptr[Nt] and stats[0] := total.
__global__ void pack_scan_add_kernel(int32_t* __restrict__ rowptr, int32_t* __restrict__ cursor,
                                     const int32_t* __restrict__ blocksum, int64_t Nt, int nb,
                                     int32_t* __restrict__ stats) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) { rowptr[Nt] = blocksum[nb]; stats[0] = blocksum[nb]; }
  for (; i < Nt; i += (int64_t)gridDim.x * blockDim.x) {
    rowptr[i] += blocksum[i / kScanTile];
    cursor[i] = 0;
  }
}

__global__ void pack_fill_kernel(const int64_t* __restrict__ ei, int64_t E, int64_t Nt,
                                 const int64_t* __restrict__ batch, const int32_t* __restrict__ rowptr,
                                 int32_t* __restrict__ cursor, int32_t* __restrict__ col) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E;
       e += (int64_t)gridDim.x * blockDim.x) {
    int64_t s = ei[e], d = ei[E + e];
    if (s < 0 || s >= Nt || d < 0 || d >= Nt || s == d) continue;
    if (batch != nullptr && batch[s] != batch[d]) continue;
    int pos = atomicAdd(cursor + d, 1);
    col[rowptr[d] + pos] = (int32_t)s;
  }
}

// One thread per row: sort the neighbour sources ascending, put the self-loop last.  Rows of up to 16
// neighbours are sorted in registers (two dependent global round trips per row instead of one per
// compare/swap: this kernel is pure latency).
__global__ void pack_rows_kernel(const int32_t* __restrict__ rowptr, int32_t* __restrict__ col, int64_t Nt) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < Nt;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int lo = rowptr[i], hi = rowptr[i + 1] - 1;
    const int n = hi - lo;
    if (n <= 16) {
      int32_t v[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) v[k] = k < n ? col[lo + k] : INT32_MAX;
#pragma unroll
      for (int pass = 0; pass < 16; ++pass) {            // odd-even transposition sort, fully unrolled
        if (pass >= n) break;
#pragma unroll
        for (int k = pass & 1; k + 1 < 16; k += 2) {
          const int32_t a = v[k], b = v[k + 1];
          v[k] = min(a, b);
          v[k + 1] = max(a, b);
        }
      }
#pragma unroll
      for (int k = 0; k < 16; ++k)
        if (k < n) col[lo + k] = v[k];
    } else {
      for (int a = lo + 1; a < hi; ++a) {
        int32_t key = col[a];
        int b = a - 1;
        while (b >= lo && col[b] > key) { col[b + 1] = col[b]; --b; }
        col[b + 1] = key;
      }
    }
    col[hi] = (int32_t)i;
  }
}

// stats[3] := 1 unless the packed graph is symmetric (every edge s->d has as many reverse edges
// d->s).  The embedder backward uses the flag to read a node's out-edges from the same CSR instead of
// scattering with atomics.  One thread per INPUT edge (s, d): with the rows sorted, the multiplicity of
// s in row d and of d in row s are two short binary searches -- a chain of ~8 dependent loads per
// thread, where one thread per row walked all its neighbours one after the other.
__device__ __forceinline__ int pack_count_in_row(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                                                 int64_t row, int32_t key) {
  int l = rowptr[row];
  const int hi = rowptr[row + 1] - 1;                             // self-loop excluded
  int r = hi;
  while (l < r) { const int m = (l + r) >> 1; if (col[m] < key) l = m + 1; else r = m; }
  int cnt = 0;
  while (l + cnt < hi && col[l + cnt] == key) ++cnt;
  return cnt;
}
__global__ void pack_symmetry_kernel(const int64_t* __restrict__ ei, int64_t E, int64_t Nt,
                                     const int64_t* __restrict__ batch, const int32_t* __restrict__ rowptr,
                                     const int32_t* __restrict__ col, int32_t* __restrict__ stats) {
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E;
       e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t s = ei[e], d = ei[E + e];
    if (s < 0 || s >= Nt || d < 0 || d >= Nt || s == d) continue;           // dropped by the packer
    if (batch != nullptr && batch[s] != batch[d]) continue;
    if (pack_count_in_row(rowptr, col, d, (int32_t)s) != pack_count_in_row(rowptr, col, s, (int32_t)d))
      atomicOr(stats + 3, 1);
  }
}

__global__ void pack_graph_ptr_kernel(const int64_t* __restrict__ batch, int64_t Nt, int64_t B,
                                      int32_t* __restrict__ graph_ptr, int32_t* __restrict__ stats) {
  pack_graph_ptr_body(batch, Nt, B, graph_ptr, stats);
}

__global__ void pack_max_nodes_kernel(const int32_t* __restrict__ graph_ptr, int64_t B,
                                      int32_t* __restrict__ stats) {
  int32_t m = 0;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < B;
       g += (int64_t)gridDim.x * blockDim.x)
    m = max(m, graph_ptr[g + 1] - graph_ptr[g]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(stats + 1, m);
}

static inline int flat_grid(int64_t n, int threads) {
  int64_t b = ceil_div(n > 0 ? n : 1, threads);
  int64_t cap = (int64_t)kNumSMs * 16;
  return (int)(b < cap ? b : cap);
}

}  // namespace gsc

using namespace gsc;

extern "C" size_t gsc_pack_csr_workspace_bytes(int64_t num_nodes, int64_t num_edges) {
  (void)num_edges;
  int64_t nb = ceil_div(num_nodes > 0 ? num_nodes : 1, kScanTile);
  return align_up((size_t)(num_nodes > 0 ? num_nodes : 1) * 4, 256) + align_up((size_t)(nb + 1) * 4, 256);
}

extern "C" int gsc_pack_csr(const int64_t* edge_index, int64_t E, const int64_t* batch, int64_t Nt,
                            int64_t B, int32_t* rowptr, int32_t* col, int32_t* graph_ptr,
                            int32_t* stats, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t st = (cudaStream_t)stream_;
  GSC_REQUIRE(Nt >= 0 && E >= 0 && B >= 1, "gsc_pack_csr: bad sizes Nt=%lld E=%lld B=%lld",
              (long long)Nt, (long long)E, (long long)B);
  GSC_REQUIRE(E + Nt < (int64_t)INT32_MAX, "gsc_pack_csr: E+Nt=%lld exceeds the int32 CSR range",
              (long long)(E + Nt));
  GSC_REQUIRE(rowptr && col && graph_ptr && stats && workspace, "gsc_pack_csr: null output pointer");
  GSC_REQUIRE(E == 0 || edge_index, "gsc_pack_csr: null edge_index");
  if (workspace_bytes < gsc_pack_csr_workspace_bytes(Nt, E))
    return fail(GSC_EWORKSPACE, "gsc_pack_csr: workspace %zu < %zu", workspace_bytes,
                gsc_pack_csr_workspace_bytes(Nt, E));
  int32_t* deg = (int32_t*)workspace;
  int nb = (int)ceil_div(Nt > 0 ? Nt : 1, kScanTile);
  int32_t* blocksum = (int32_t*)((char*)workspace + align_up((size_t)(Nt > 0 ? Nt : 1) * 4, 256));

  // 7 launches for batches of up to kFusedScanBlocks * 1024 nodes (10 beyond): init | in-degrees + graph_ptr |
  // tile scan | offsets + cursors + max_nodes | fill | row sort | symmetry
  pack_init_kernel<<<flat_grid(Nt, 256), 256, 0, st>>>(deg, Nt, stats);
  GSC_LAUNCH_CHECK("pack_init_kernel");
  if (E > 0) {
    const int64_t work = E > Nt ? E : Nt;
    pack_count_kernel<<<flat_grid(work, 256), 256, 0, st>>>(edge_index, E, Nt, batch, deg, stats, B, graph_ptr);
    GSC_LAUNCH_CHECK("pack_count_kernel");
  } else {
    int64_t gp_n = (batch == nullptr) ? 1 : (Nt > 0 ? Nt : B + 1);
    pack_graph_ptr_kernel<<<flat_grid(gp_n, 256), 256, 0, st>>>(batch, Nt, B, graph_ptr, stats);
    GSC_LAUNCH_CHECK("pack_graph_ptr_kernel");
  }
  pack_scan_local_kernel<<<nb, kScanThreads, 0, st>>>(deg, Nt, rowptr, blocksum);
  GSC_LAUNCH_CHECK("pack_scan_local_kernel");
  if (nb <= kFusedScanBlocks) {
    pack_scan_finish_kernel<<<nb, kScanThreads, 0, st>>>(rowptr, deg, blocksum, Nt, nb, stats, graph_ptr, B);
    GSC_LAUNCH_CHECK("pack_scan_finish_kernel");
  } else {
    pack_scan_blocks_kernel<<<1, 1024, 0, st>>>(blocksum, nb);
    GSC_LAUNCH_CHECK("pack_scan_blocks_kernel");
    pack_scan_add_kernel<<<flat_grid(Nt, 256), 256, 0, st>>>(rowptr, deg, blocksum, Nt, nb, stats);
    GSC_LAUNCH_CHECK("pack_scan_add_kernel");
    pack_max_nodes_kernel<<<flat_grid(B, 256), 256, 0, st>>>(graph_ptr, B, stats);
    GSC_LAUNCH_CHECK("pack_max_nodes_kernel");
  }
  if (E > 0) {
    pack_fill_kernel<<<flat_grid(E, 256), 256, 0, st>>>(edge_index, E, Nt, batch, rowptr, deg, col);
    GSC_LAUNCH_CHECK("pack_fill_kernel");
  }
  if (Nt > 0) {
    pack_rows_kernel<<<flat_grid(Nt, 128), 128, 0, st>>>(rowptr, col, Nt);
    GSC_LAUNCH_CHECK("pack_rows_kernel");
  }
  if (E > 0) {
    pack_symmetry_kernel<<<flat_grid(E, 256), 256, 0, st>>>(edge_index, E, Nt, batch, rowptr, col, stats);
    GSC_LAUNCH_CHECK("pack_symmetry_kernel");
  }
  return GSC_OK;
}

// ---------------------------------------------------------------------------------------------
// gsc_pack_csr_noalloc: the packer for callers that cannot provide a workspace.
//
// gsc_pack_csr needs two scratch arrays (per-node degree counters that later serve as fill
// cursors, and the scan's block totals) and expects the caller to allocate them.  Some callers of
// this library cannot allocate device memory at the point where they pack a batch:
//   * the per-environment-step rollout / serving loop (predict() on one small graph), which runs
//     beside a training stream: a cudaMalloc there synchronises the whole device and stalls the
//     update that is in flight;
//   * an integrator binding the C-ABI from a host that was handed a fixed device-memory budget
//     (only its input and output buffers) and owns no caching allocator such as torch's.
// For them the scratch lives in the library image: a file-scope __device__ array, resident on every
// GPU that loads libgsc_b200 from load to exit.  It covers batches of up to
// GSC_PACK_NOALLOC_MAX_NODES nodes (4 B per node + 4 B per 1024 nodes, ~4.2 MB); larger batches
// are rejected with GSC_EINVAL and must use gsc_pack_csr with a caller workspace.
//
// Rules of use:
//   * The first call on each device resolves the array's device address (cudaGetSymbolAddress, with
//     that device current) and caches it; make that first call before any stream capture begins.
//   * Every call on a device shares the one array and finds in it whatever the previous call left.
//     Nothing is assumed about those contents: the packer's own init/scan kernels rewrite every
//     element they later read (degrees are set to 1, block totals are written before they are
//     scanned, cursors are zeroed by the scan), so no separate clearing pass is needed.
//   * Two calls whose kernels overlap on the same GPU (different streams, or different host
//     threads) race on the array and corrupt each other's CSR.  Serialise them, or use gsc_pack_csr.
// No cudaMalloc / cudaFree / allocating library routine is called here.
namespace gsc {

constexpr int64_t kNoAllocMaxNodes = GSC_PACK_NOALLOC_MAX_NODES;
constexpr size_t kNoAllocScratchBytes =
    ((size_t)kNoAllocMaxNodes * 4 + 255) / 256 * 256 +
    (((size_t)(kNoAllocMaxNodes + kScanTile - 1) / kScanTile + 1) * 4 + 255) / 256 * 256;

static __device__ __align__(256) unsigned char g_pack_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
static void* g_pack_scratch_addr[kMaxDevices] = {nullptr};   // device address per ordinal, set once
static std::mutex g_pack_scratch_mu;

}  // namespace gsc

extern "C" int gsc_pack_csr_noalloc(const int64_t* edge_index, int64_t E, const int64_t* batch, int64_t Nt,
                                    int64_t B, int32_t* rowptr, int32_t* col, int32_t* graph_ptr,
                                    int32_t* stats, void* stream_) {
  GSC_REQUIRE(Nt >= 0 && Nt <= kNoAllocMaxNodes,
              "gsc_pack_csr_noalloc: %lld nodes exceed the built-in scratch (max %lld nodes); "
              "use gsc_pack_csr with a caller-provided workspace",
              (long long)Nt, (long long)kNoAllocMaxNodes);
  int dev = -1;
  GSC_CUDA_TRY(cudaGetDevice(&dev));
  GSC_REQUIRE(dev >= 0 && dev < kMaxDevices, "gsc_pack_csr_noalloc: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_pack_scratch_mu);
  void* scratch = g_pack_scratch_addr[dev];
  if (scratch == nullptr) {
    // `g_pack_scratch` named in host code is only a handle; the kernels need its device address.
    GSC_CUDA_TRY(cudaGetSymbolAddress(&scratch, g_pack_scratch));
    GSC_REQUIRE(scratch != nullptr, "gsc_pack_csr_noalloc: cudaGetSymbolAddress returned null");
    g_pack_scratch_addr[dev] = scratch;
  }
  if (gsc_pack_csr_workspace_bytes(Nt, E) > kNoAllocScratchBytes)
    return fail(GSC_EINVAL, "gsc_pack_csr_noalloc: batch needs %zu scratch bytes, built-in array has %zu",
                gsc_pack_csr_workspace_bytes(Nt, E), kNoAllocScratchBytes);
  return gsc_pack_csr(edge_index, E, batch, Nt, B, rowptr, col, graph_ptr, stats, scratch,
                      kNoAllocScratchBytes, stream_);
}
