// Tensor-core path of gsc_gemm (GSC_EPI_TF32): tcgen05.mma kind::tf32, fp32 accumulation in TMEM,
// operands streamed by TMA (128-byte swizzle) through a multi-stage shared-memory ring.
//
// Serves the dense layers of the PyG MLP (reference src/rlsp/agents/models.py:80-83,119-122) in
// all three roles -- forward (both operands K-major), dgrad (B MN-major) and wgrad (both MN-major)
// -- with the same K-segment / epilogue contract as the exact-fp32 SIMT kernel in gemm.cu.
//
// CTA = 320 threads (448 in the 3xTF32 variant), persistent over (m-tile, n-tile, k-split) work items:
//   warp 8/12   TMA producer   one elected lane issues cp.async.bulk.tensor per stage
//   warp 9/13   MMA issuer     one elected lane issues 4 (3xTF32: 12) x tcgen05.mma (K = 8 each) per
//                              32-wide k-block; tcgen05.commit frees the stage / publishes the accumulator
//   warps 8..11 3xTF32 only    split every staged operand tile into hi + lo TF32 images
//   warps 0..7  epilogue       two warps per TMEM lane quadrant, 64 of the tile's 128 columns each
//                              (a lone warp per scheduler cannot hide its own latencies):
//                              tcgen05.ld (32 lanes x 32 columns), bias/ReLU/mask/gate/
//                              clamp.  The mask (or gate) tile is prefetched by TMA into a 64 KB
//                              swizzled buffer, combined in place with the accumulators and
//                              written out with TMA stores: both streams are fully coalesced and
//                              asynchronous (the [B, A] mask-in / action-out traffic is what
//                              bounds these layers).  A direct-store path covers split-K partials
//                              and operands that TMA cannot map.
// Tile 128 x 128, two TMEM accumulator buffers (2 x 128 columns) so the epilogue of tile i overlaps
// the loads and MMAs of tile i+1.  Roofline: HBM for the mask/action/output streams
// (4(MK + KN + MN) algorithmic bytes), tensor pipe (TF32) for large-batch shapes.
#include "common.cuh"
#include "tc_ptx.cuh"
#include <mutex>
#include <stdlib.h>

namespace gsc {

constexpr int kTcBM = 128, kTcBN = 128, kTcBK = 32;          // fp32 elements; BK * 4 B = one 128 B swizzle row
// Shared memory: six 32 KB ring slots.  Launches whose output leaves through the 64 KB TMA epilogue
// tile use four of them as the operand ring and the last two as that tile; split-K launches (direct
// partial stores, long K streams) use all six as the ring: bytes in flight per SM are what bounds a
// K-streaming GEMM (latency x bandwidth), not the tensor pipe.
constexpr int kTcMaxStages = 6;
constexpr int kTcEpiStages = 4;                              // ring depth next to the epilogue tile
constexpr int kTcEpiWarps = 8;                               // two warps per TMEM lane quadrant, 64 columns each
constexpr int kTcThreads = 64 + 32 * kTcEpiWarps;            // 320
constexpr int kTcStageBytes = (kTcBM + kTcBN) * kTcBK * 4;   // 32 KB
constexpr int kTcEpiBytes = kTcBM * kTcBN * 4;                 // 64 KB: epilogue operand in, output tile out (in place)
constexpr int kTcSmemBytes = kTcMaxStages * kTcStageBytes + 1024 /*align*/ + 256 /*barriers*/;
static_assert(kTcEpiStages * kTcStageBytes + kTcEpiBytes == kTcMaxStages * kTcStageBytes, "ring/epilogue split");
constexpr int kTmemCols = 2 * kTcBN;                          // 256
// 3xTF32 (fp32-faithful) variant: four converter warps split every staged operand tile x into
// hi = tf32(x) (written back in place) and lo = tf32(x - hi) (written to a second ring), and the MMA
// warp issues hi*hi + hi*lo + lo*hi.  Seven 32 KB slots: raw ring | lo ring | epilogue tile.
constexpr int kTcX3Slots = 7;
constexpr int kTcX3RawEpi = 3, kTcX3LoEpi = 2;                // next to the 64 KB epilogue tile
constexpr int kTcX3RawSplit = 4, kTcX3LoSplit = 3;            // split-K launches (direct partial stores)
constexpr int kTcConvWarps = 4;
constexpr int kTcThreadsX3 = kTcThreads + 32 * kTcConvWarps;  // 448
constexpr int kTcSmemBytesX3 = kTcX3Slots * kTcStageBytes + 1024 + 256;
constexpr int kTcBarSlots = 8;                                // barrier slots per ring
static_assert(kTcMaxStages <= kTcBarSlots && kTcX3RawSplit <= kTcBarSlots, "barrier slots");

struct TcParams {
  int64_t M, N, ldc, ld_mul, ld_gate;
  float* C;
  float* partial;
  const float* bias;
  const float* mul;
  const float* gate;
  const float* row_table;
  const int32_t* row_n;
  int64_t ld_table;
  int n_max, groups;
  int flags, num_seg;
  int kb_end[3];          // exclusive prefix of 32-wide k-blocks per segment
  int kb_total, kb_per_split, splits, tiles_m, tiles_n;
  int kb_interleave;      // split-K items take every splits-th k-block instead of a contiguous range
  int bn;                 // N extent of the tile: 128, or 64 when N <= 64 (half the MMA and B traffic)
  int stages;             // operand ring depth: kTcEpiStages with the TMA epilogue, kTcMaxStages without
  int lo_stages;          // 3xTF32: depth of the lo ring (0 otherwise)
  int lo_off, lo_slot;    // 3xTF32: byte offset of the lo ring and bytes per slot
  int ebuf_off;           // byte offset of the epilogue tile
  int bar_off;            // byte offset of the barriers
  int a_tmem;             // 3xTF32: the A operand (hi and lo images) lives in tensor memory
  int dbg;                // GSC_TC_DBG experiment bits: 1 no MMAs, 2 no TMA loads, 4 no epilogue body, 8 no conversion
  int epi_tma;            // 1: output tile leaves through shared memory + TMA store (coalesced, asynchronous)
  int epi_kind;           // operand streamed in by TMA for the epilogue: 0 none, 1 mul, 2 gate
  int epi_sf;             // flag combination with a compile-time instance of the TMA epilogue, or -1
};

__device__ __forceinline__ float smask_apply(const TcParams& p, float v, int nb, int64_t j) {
  const uint32_t ju = (uint32_t)j;
  const int src = (int)(ju / ((uint32_t)p.groups * (uint32_t)p.n_max)), dst = (int)(ju % (uint32_t)p.n_max);
  return (src < nb && dst < nb) ? v : 0.f;
}

// Walks consecutive columns of one row of the structured mask without a division per element:
// column k = (src*G + g)*n_max + dst is inside the mask of an nb-node state iff src < nb && dst < nb.
struct SmaskWalk {
  int src, rem, dst, gn, nm;
  __device__ __forceinline__ void seek(const TcParams& p, int64_t col) {
    nm = p.n_max;
    gn = p.groups * p.n_max;
    const uint32_t c = (uint32_t)col;
    src = (int)(c / (uint32_t)gn);
    rem = (int)(c - (uint32_t)src * (uint32_t)gn);
    dst = (int)((uint32_t)rem % (uint32_t)nm);
  }
  __device__ __forceinline__ bool next(int nb) {
    const bool in = src < nb && dst < nb;
    if (++dst == nm) dst = 0;
    if (++rem == gn) { rem = 0; ++src; }
    return in;
  }
};

// Bit q of the result: is column col0 + q inside the structured mask of an nb-node state?  For
// n_max >= 32 a run of 32 columns crosses at most one (source, group) boundary, so two range tests
// replace 32 walks; smaller n_max walks the run.
__device__ __forceinline__ uint32_t low_bits(int k) { return k >= 32 ? 0xffffffffu : (k <= 0 ? 0u : ((1u << k) - 1u)); }
__device__ __forceinline__ uint32_t smask_bits(const TcParams& p, int nb, int col0) {
  const int nm = p.n_max, gn = p.groups * p.n_max;
  int src = (int)((uint32_t)col0 / (uint32_t)gn);
  int rem = col0 - src * gn;
  int dst = (int)((uint32_t)rem % (uint32_t)nm);
  if (nm >= 32) {
    const int first = min(32, nm - dst);                       // columns before the next group boundary
    uint32_t bits = src < nb ? low_bits(min(nb - dst, first)) : 0u;
    if (first < 32) {
      if (rem + first == gn) ++src;                            // the boundary also ends the source row
      if (src < nb) bits |= low_bits(min(nb, 32 - first)) << first;
    }
    return bits;
  }
  uint32_t bits = 0;
#pragma unroll 1
  for (int q = 0; q < 32; ++q) {
    bits |= (uint32_t)(src < nb && dst < nb) << q;
    if (++dst == nm) dst = 0;
    if (++rem == gn) { rem = 0; ++src; }
  }
  return bits;
}

__device__ __forceinline__ float tc_epilogue(const TcParams& p, float v, int64_t i, int64_t j) {
  if (p.flags & GSC_EPI_ROWTABLE) v += p.row_table[(int64_t)p.row_n[i] * p.ld_table + j];
  if (p.flags & GSC_EPI_BIAS) v += p.bias[j];
  if (p.flags & GSC_EPI_RELU) v = fmaxf(v, 0.f);
  if (p.flags & GSC_EPI_MUL) v *= p.mul[i * p.ld_mul + j];
  if (p.flags & GSC_EPI_SMASK) v = smask_apply(p, v, p.row_n[i], j);
  if (p.flags & GSC_EPI_GATE) v = p.gate[i * p.ld_gate + j] > 0.f ? v : 0.f;
  if (p.flags & GSC_EPI_CLAMP1) v = fminf(fmaxf(v, -1.f), 1.f);
  return v;
}

struct WorkItem {
  int m0, n0, split, kb_first, kb_step, kb_count;
};
// Split-K work items walk the k-blocks either as one contiguous range each or INTERLEAVED (item s takes
// k-blocks s, s + splits, ...).  Interleaved, the CTAs that share an m-tile read neighbouring 128-byte
// pieces of the same operand rows at about the same time: for a K-major [1024 x 30000] operand (row
// pitch 120 KB) that is what lets the DRAM controller serve them from open pages instead of
// activating a page per 128 (256 with L2 promotion) bytes.
__device__ __forceinline__ WorkItem decode_work(const TcParams& p, int w) {
  WorkItem it;
  const int tm = w % p.tiles_m;
  const int rest = w / p.tiles_m;
  const int tn = rest % p.tiles_n;
  it.split = rest / p.tiles_n;
  it.m0 = tm * kTcBM;
  it.n0 = tn * p.bn;
  if (p.kb_interleave) {
    it.kb_first = it.split;
    it.kb_step = p.splits;
    it.kb_count = (p.kb_total - it.split + p.splits - 1) / p.splits;
  } else {
    it.kb_first = it.split * p.kb_per_split;
    it.kb_step = 1;
    it.kb_count = min(p.kb_total, it.kb_first + p.kb_per_split) - it.kb_first;
  }
  return it;
}

// lo = x - trunc(x) of U x 2 KB of a staged tile (16 bytes per thread and 2 KB slice), all loads first
template <int U>
__device__ __forceinline__ void lo_image(uint32_t src, uint32_t dst) {
  uint32_t v[U][4];
#pragma unroll
  for (int u = 0; u < U; ++u)
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v[u][0]), "=r"(v[u][1]), "=r"(v[u][2]), "=r"(v[u][3]) : "r"(src + u * 2048));
#pragma unroll
  for (int u = 0; u < U; ++u) {
    uint32_t lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e)
      lo[e] = __float_as_uint(__uint_as_float(v[u][e]) - __uint_as_float(v[u][e] & 0xffffe000u));
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + u * 2048), "r"(lo[0]), "r"(lo[1]), "r"(lo[2]), "r"(lo[3]) : "memory");
  }
}

// TMA epilogue of one 128 x 128 tile: accumulators (+ prefetched mask/gate tile) -> smem tile in place ->
// TMA store.  SF >= 0: the GSC_EPI_* bits of this launch, known at compile time, no operand tile
// (epi_kind == 0); SF = -1: flags read at run time (any combination, mask / gate tile included).
template <int SF>
__device__ __forceinline__ void epi_tma_tile(const TcParams& p, const CUtensorMap* mapC, const WorkItem& it, int as, int iter,
                                             int warp, int lane, uint32_t ebuf, uint32_t tmem_base, uint32_t tempty,
                                             uint32_t efull_bar, uint32_t eempty_bar) {
  constexpr bool kDyn = SF < 0;
  const int quad = warp & 3;
  const int c_begin = (warp >> 2) * (kTcBN / 64);
  const int epi_kind = kDyn ? p.epi_kind : 0;
  if (epi_kind) mbar_wait(efull_bar, (uint32_t)iter & 1u);
  const int lrow = quad * 32 + lane;
  const int64_t grow = (int64_t)it.m0 + lrow;
  const bool f_table = kDyn ? (p.flags & GSC_EPI_ROWTABLE) != 0 : (SF & GSC_EPI_ROWTABLE) != 0;
  const bool f_bias = kDyn ? (p.flags & GSC_EPI_BIAS) != 0 : (SF & GSC_EPI_BIAS) != 0;
  const bool f_relu = kDyn ? (p.flags & GSC_EPI_RELU) != 0 : (SF & GSC_EPI_RELU) != 0;
  const bool f_smask = kDyn ? (p.flags & GSC_EPI_SMASK) != 0 : (SF & GSC_EPI_SMASK) != 0;
  const bool f_clamp = kDyn ? (p.flags & GSC_EPI_CLAMP1) != 0 : (SF & GSC_EPI_CLAMP1) != 0;
  const bool f_gate_direct = kDyn && (p.flags & GSC_EPI_GATE) && p.epi_kind != 2;
  const int nb = ((f_table || f_smask) && grow < p.M) ? p.row_n[grow] : 0;
  const float* trow = f_table ? p.row_table + (int64_t)nb * p.ld_table : nullptr;
  const bool vec_ok = (!f_table || ((reinterpret_cast<uintptr_t>(trow) & 15) == 0)) &&
                      (!f_bias || ((reinterpret_cast<uintptr_t>(p.bias) & 15) == 0));
  const int ncols = (int)p.N;                                  // the tensor-core path requires N < 2^31
#pragma unroll 1
  for (int cc = 0; cc < kTcBN / 64; ++cc) {
    const int c = c_begin + cc;
    const int col0 = it.n0 + c * 32;
    // Columns past N compute garbage that the TMA store clips; only their loads are guarded.
    // Additive per-column terms (bias, per-n table row) are fetched before waiting for TMEM.
    float tadd[32], badd[32];                          // raw loads: consumed only after the TMEM wait
    if (f_table || f_bias) {
      if (vec_ok && col0 + 32 <= ncols) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (f_table) {
            const float4 t4 = *reinterpret_cast<const float4*>(trow + col0 + 4 * q);
            tadd[4 * q] = t4.x; tadd[4 * q + 1] = t4.y; tadd[4 * q + 2] = t4.z; tadd[4 * q + 3] = t4.w;
          }
          if (f_bias) {
            const float4 b4 = *reinterpret_cast<const float4*>(p.bias + col0 + 4 * q);
            badd[4 * q] = b4.x; badd[4 * q + 1] = b4.y; badd[4 * q + 2] = b4.z; badd[4 * q + 3] = b4.w;
          }
        }
      } else {
#pragma unroll
        for (int q = 0; q < 32; ++q) {
          const int cl = min(col0 + q, ncols - 1);
          if (f_table) tadd[q] = trow[cl];
          if (f_bias) badd[q] = p.bias[cl];
        }
      }
    }
    const uint32_t keep = f_smask ? smask_bits(p, nb, col0) : 0xffffffffu;
    float v[32];
    tc_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * kTcBN + c * 32), v);
    if (cc == kTcBN / 64 - 1) {                                 // last TMEM read of this tile:
      tc_fence_before();                                        // the accumulator buffer is free again
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty);
    }
    const uint32_t rbase = ebuf + c * 16384 + lrow * 128;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const uint32_t addr = rbase + (uint32_t)((q ^ (lrow & 7)) << 4);          // 128 B swizzle
      float4 e = make_float4(0.f, 0.f, 0.f, 0.f);
      if (epi_kind) asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(e.x), "=f"(e.y), "=f"(e.z), "=f"(e.w) : "r"(addr));
      const float ev[4] = {e.x, e.y, e.z, e.w};
      float o[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float t = v[4 * q + u];
        if (f_table) t += tadd[4 * q + u];
        if (f_bias) t += badd[4 * q + u];
        if (f_relu) t = fmaxf(t, 0.f);
        if (epi_kind == 1) t *= ev[u];
        if (f_smask) t = (keep & (1u << (4 * q + u))) ? t : 0.f;
        if (epi_kind == 2) t = ev[u] > 0.f ? t : 0.f;
        if (f_gate_direct) {
          const int col = col0 + 4 * q + u;
          if (grow < p.M && col < ncols) t = p.gate[grow * p.ld_gate + col] > 0.f ? t : 0.f;
        }
        if (f_clamp) t = fminf(fmaxf(t, -1.f), 1.f);
        o[u] = t;
      }
      asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(o[0]), "f"(o[1]), "f"(o[2]), "f"(o[3]) : "memory");
    }
    if (epi_kind == 0) {
      // No prefetched operand tile: every 32-column chunk has its own 16 KB buffer and leaves with
      // its own TMA store, which drains while the next chunk is computed.  One named barrier per
      // chunk: before it the issuing thread makes sure the previous store of this column half has
      // been read out of shared memory (its buffer is the next one to be overwritten).
      fence_proxy_async_smem();                                  // st.shared -> visible to the TMA store
      const bool issuer = (warp & 3) == 0 && lane == 0;
      if (issuer) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      asm volatile("bar.sync %0, 128;" ::"r"(1 + (warp >> 2)) : "memory");
      if (issuer && col0 < ncols && !(p.dbg & 32)) {
        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                         reinterpret_cast<uint64_t>(mapC)), "r"(ebuf + c * 16384), "r"(col0), "r"(it.m0) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
  }
  if (epi_kind == 0) return;
  fence_proxy_async_smem();                                    // st.shared -> visible to the TMA store
  // the two column halves (warps 0..3 / 4..7) own their half of the tile buffer, their own named
  // barrier and their own TMA store: one half computes while the other drains
  const int half = warp >> 2;
  asm volatile("bar.sync %0, 128;" ::"r"(1 + half) : "memory");
  if ((warp & 3) == 0 && lane == 0) {
#pragma unroll 1
    for (int c = c_begin; c < c_begin + kTcBN / 64; ++c)
      if (it.n0 + 32 * c < p.N && !(p.dbg & 32))
        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                         reinterpret_cast<uint64_t>(mapC)), "r"(ebuf + c * 16384), "r"(it.n0 + 32 * c), "r"(it.m0) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // smem tile may be overwritten
    if (p.epi_kind) mbar_arrive(eempty_bar);
  }
  asm volatile("bar.sync %0, 128;" ::"r"(1 + half) : "memory");
}

// Warp roles.  The two single-thread control loops (TMA producer, MMA issuer) are pure instruction
// latency: whatever they spend per k-block bounds the whole pipeline (measured: ~90 SASS instructions
// per k-block in either loop cost more than the four MMAs they feed).  They therefore (i) run under
// elect.sync so that the compiler emits straight-line uniform-datapath code, (ii) keep descriptors as
// precomputed 32-bit words that advance by small constants, and (iii) sit in the HIGHEST warp ids,
// which the SM's issue arbiter serves first, ahead of the epilogue and converter warps.
//   warps 0..7    epilogue (TMEM lane quadrant = warp & 3, column half = warp >> 2)
//   warps 8..11   3xTF32 only: hi/lo converters
//   next warp     TMA producer
//   last warp     MMA issuer
template <bool A_MN, bool B_MN, bool X3>
__global__ void __launch_bounds__(X3 ? kTcThreadsX3 : kTcThreads, 1)
gemm_tc_kernel(const __grid_constant__ TcParams p, const __grid_constant__ CUtensorMap mapA0,
               const __grid_constant__ CUtensorMap mapA1, const __grid_constant__ CUtensorMap mapA2,
               const __grid_constant__ CUtensorMap mapB0, const __grid_constant__ CUtensorMap mapB1,
               const __grid_constant__ CUtensorMap mapB2, const __grid_constant__ CUtensorMap mapC,
               const __grid_constant__ CUtensorMap mapE) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;       // 1024 B: swizzle-128B atoms
  const uint32_t ebuf = smem_base + p.ebuf_off;                            // epilogue tile, 4 x [128 rows x 128 B]
  const uint32_t lo_base = smem_base + p.lo_off;                           // 3xTF32: lo ring behind the raw ring
  const uint32_t bar_base = smem_base + p.bar_off;
  constexpr bool kATm = X3;                                                // A-from-TMEM (hi and lo images), either major-ness
  const bool a_tmem = kATm && p.a_tmem;
  const uint32_t tmem_cols = a_tmem ? 512u : (uint32_t)kTmemCols;
  constexpr int kTcStages = kTcBarSlots;                                   // barrier slots per ring
  const int num_stages = p.stages;
  // barriers: full[S], empty[S], lo_full[4], lo_empty[4], tmem_full[2], tmem_empty[2], epilogue
  // operand full/empty, then the TMEM base address slot
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kTcStages + s); };
  auto lofull_bar = [&](int s) { return bar_base + 8u * (2 * kTcStages + s); };
  auto loempty_bar = [&](int s) { return bar_base + 8u * (2 * kTcStages + 4 + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * kTcStages + 8 + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * kTcStages + 10 + s); };
  const uint32_t efull_bar = bar_base + 8u * (2 * kTcStages + 12);
  const uint32_t eempty_bar = bar_base + 8u * (2 * kTcStages + 13);
  const uint32_t tmem_slot = bar_base + 8u * (2 * kTcStages + 14);
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_work = p.tiles_m * p.tiles_n * p.splits;
  constexpr int kProducerWarp = kTcEpiWarps + (X3 ? kTcConvWarps : 0);
  constexpr int kMmaWarp = kProducerWarp + 1;

  if (warp == kMmaWarp && lane == 0) {
    for (int s = 0; s < kTcStages; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 4; ++s) { mbar_init(lofull_bar(s), kTcConvWarps); mbar_init(loempty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), kTcEpiWarps); }
    mbar_init(efull_bar, 1);
    mbar_init(eempty_bar, 2);                     // both column halves release the mask/gate tile
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {   // one warp allocates TMEM (256 columns) and later frees it
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(tmem_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == kProducerWarp) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0, ephase = 0;
      for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
        const WorkItem it = decode_work(p, w);
        const uint32_t stage_tx = (uint32_t)(kTcBM + p.bn) * kTcBK * 4;
#pragma unroll 1
        for (int i = 0, kb = it.kb_first; i < it.kb_count; ++i, kb += it.kb_step) {
          const int s = (kb >= p.kb_end[0]) + (kb >= p.kb_end[1]);          // K segment of this k-block
          const int seg_begin = s == 0 ? 0 : (s == 1 ? p.kb_end[0] : p.kb_end[1]);
          const CUtensorMap* mA = s == 0 ? &mapA0 : (s == 1 ? &mapA1 : &mapA2);
          const CUtensorMap* mB = s == 0 ? &mapB0 : (s == 1 ? &mapB1 : &mapB2);
          const int k0 = (kb - seg_begin) * kTcBK;
          const uint32_t fb = full_bar(stage);
          mbar_wait(empty_bar(stage), phase ^ 1u);
          if (p.dbg & 2) { mbar_arrive(fb); if (++stage == num_stages) { stage = 0; phase ^= 1u; } continue; }
          mbar_expect_tx(fb, stage_tx);
          const uint32_t sa = smem_base + stage * kTcStageBytes;
          const uint32_t sb = sa + kTcBM * kTcBK * 4;
          if (!A_MN) {
            tma_load_2d(sa, mA, fb, k0, it.m0);                                   // box {32 k, 128 m}
          } else {
#pragma unroll
            for (int b = 0; b < kTcBM / 32; ++b)                                  // box {32 m, 32 k}
              tma_load_2d(sa + b * 4096, mA, fb, it.m0 + 32 * b, k0);
          }
          if (!B_MN) {
            tma_load_2d(sb, mB, fb, k0, it.n0);                                   // box {32 k, bn n}
          } else {
#pragma unroll
            for (int b = 0; b < kTcBN / 32; ++b)                                  // box {32 n, 32 k}
              if (32 * b < p.bn) tma_load_2d(sb + b * 4096, mB, fb, it.n0 + 32 * b, k0);
          }
          if (++stage == num_stages) { stage = 0; phase ^= 1u; }
        }
        if (p.epi_tma && p.epi_kind) {             // prefetch the mask / gate tile of this work item
          mbar_wait(eempty_bar, ephase ^ 1u);
          mbar_expect_tx(efull_bar, kTcEpiBytes);
#pragma unroll
          for (int c = 0; c < kTcBN / 32; ++c) tma_load_2d(ebuf + c * 16384, &mapE, efull_bar, it.n0 + 32 * c, it.m0);
          ephase ^= 1u;
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      // instruction descriptor: D=f32, A=B=tf32, M=128, N=bn, major bits
      const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (((A_MN && !a_tmem) ? 1u : 0u) << 15) |
                             ((B_MN ? 1u : 0u) << 16) | ((uint32_t)(p.bn >> 3) << 17) |
                             ((uint32_t)(kTcBM >> 4) << 24);
      // shared-memory descriptors as (low word, high word): the high word is constant per operand, the
      // low word = (address >> 4) | LBO << 16 advances by constants (see smem_desc in tc_ptx.cuh).
      //   K-major : SWIZZLE_128B (layout 2), LBO 16 B, SBO 1024 B; 8 k further = +32 B.
      //   MN-major: SWIZZLE_128B_BASE32B (layout 1), LBO 4096 B, SBO 512 B; 8 k further = +1024 B.
      constexpr uint32_t kHiK = 64u | (1u << 14) | (2u << 29), kHiMN = 32u | (1u << 14) | (1u << 29);
      constexpr uint32_t a_hi = A_MN ? kHiMN : kHiK, b_hi = B_MN ? kHiMN : kHiK;
      constexpr uint32_t a_step = A_MN ? 64u : 2u, b_step = B_MN ? 64u : 2u;
      constexpr uint32_t kStageW = kTcStageBytes >> 4, kOperandW = (kTcBM * kTcBK * 4) >> 4;
      const uint32_t a_raw0 = (smem_base >> 4) | ((A_MN ? 256u : 1u) << 16);
      const uint32_t b_raw0 = ((smem_base >> 4) + kOperandW) | ((B_MN ? 256u : 1u) << 16);
      const uint32_t kLoW = (uint32_t)p.lo_slot >> 4;
      const uint32_t a_lo0 = (lo_base >> 4) | ((A_MN ? 256u : 1u) << 16);
      const uint32_t b_lo0 = ((lo_base >> 4) + (a_tmem ? 0u : kOperandW)) | ((B_MN ? 256u : 1u) << 16);
      const uint32_t tmem_a0 = tmem_base + (uint32_t)kTmemCols;    // A ring in tensor memory: [lo_stages][hi 32 | lo 32] columns
      int stage = 0, lstage = 0;
      uint32_t phase = 0, lphase = 0;
      int iter = 0;
      for (int w = blockIdx.x; w < total_work; w += gridDim.x, ++iter) {
        const WorkItem it = decode_work(p, w);
        const int as = iter & 1;
        mbar_wait(tempty_bar(as), (((uint32_t)iter >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(as * kTcBN);
        uint32_t acc = 0;
#pragma unroll 1
        for (int kb = 0; kb < it.kb_count; ++kb) {
          if (X3) mbar_wait(lofull_bar(lstage), lphase);     // hi (in place) and lo images are ready
          else mbar_wait(full_bar(stage), phase);
          const uint32_t da = a_raw0 + stage * kStageW, db = b_raw0 + stage * kStageW;
          const uint32_t la = a_lo0 + lstage * kLoW, lb = b_lo0 + lstage * kLoW;
          if (kATm && a_tmem) {
            tc_fence_after();                              // the A images were written with tcgen05.st by other warps
            const uint32_t ta = tmem_a0 + (uint32_t)(lstage * 2 * kTcBK);
#pragma unroll
            for (int j = 0; j < kTcBK / 8; ++j) {
              tc_mma_tf32_ts(tmem_d, ta + 8 * j, db + j * b_step, b_hi, idesc, acc);
              acc = 1;
              tc_mma_tf32_ts(tmem_d, ta + 8 * j, lb + j * b_step, b_hi, idesc, 1u);
              tc_mma_tf32_ts(tmem_d, ta + kTcBK + 8 * j, db + j * b_step, b_hi, idesc, 1u);
            }
          } else if (!(p.dbg & 1)) {
#pragma unroll
            for (int j = 0; j < kTcBK / 8; ++j) {
              tc_mma_tf32_w(tmem_d, da + j * a_step, a_hi, db + j * b_step, b_hi, idesc, acc);
              acc = 1;
              if (X3) {
                tc_mma_tf32_w(tmem_d, da + j * a_step, a_hi, lb + j * b_step, b_hi, idesc, 1u);
                tc_mma_tf32_w(tmem_d, la + j * a_step, a_hi, db + j * b_step, b_hi, idesc, 1u);
              }
            }
          }
          tc_commit(empty_bar(stage));            // frees the smem stage once these MMAs retire
          if (++stage == num_stages) { stage = 0; phase ^= 1u; }
          if (X3) {
            tc_commit(loempty_bar(lstage));
            if (++lstage == p.lo_stages) { lstage = 0; lphase ^= 1u; }
          }
        }
        tc_commit(tfull_bar(as));                 // accumulator complete -> epilogue
      }
    }
  } else if (X3 && warp >= kTcEpiWarps) {
    // ===================== converter warps, 3xTF32 only =====================
    // The tensor core reads only the 19 high bits of an fp32 operand (sign, exponent, 10 mantissa bits;
    // checked on the device by tools/tf32_trunc_check.py), so the staged tile itself already is the hi
    // image, hi = trunc(x).  The converters write lo = x - trunc(x) (exact in fp32, at most 13
    // significant bits, of which the tensor core again keeps the top 11) at the same offset of the lo
    // ring.  Dropped: lo*lo and the 2 low bits of lo, together ~2^-21 relative per product.  The split
    // is elementwise, hence layout-agnostic: K-major and MN-major swizzled tiles alike.
    // (GSC_TC_DBG bit 16, experiments only: round-to-nearest split, hi rewritten in place.)
    const int ct = (warp - kTcEpiWarps) * 32 + lane;            // 0..127
    int stage = 0, lstage = 0;
    uint32_t phase = 0, lphase = 0;
    constexpr int kIters = kTcStageBytes / (16 * 32 * kTcConvWarps);   // 16 float4 per thread and stage
    const bool inplace = (p.dbg & 16) != 0;
    const uint32_t rnd = inplace ? 0x1000u : 0u;
    for (int w = blockIdx.x; w < total_work; w += gridDim.x) {
      const WorkItem it = decode_work(p, w);
#pragma unroll 1
      for (int kb = 0; kb < it.kb_count; ++kb) {
        mbar_wait(full_bar(stage), phase);                       // TMA bytes have landed
        mbar_wait(loempty_bar(lstage), lphase ^ 1u);             // the MMAs that read this lo slot retired
        const uint32_t src = smem_base + stage * kTcStageBytes + ct * 16;
        const uint32_t dst = lo_base + lstage * p.lo_slot + ct * 16;
        if (kATm && a_tmem && (p.dbg & 8)) {
          // experiment: no conversion work at all
        } else if (kATm && a_tmem) {
          // A rows -> tensor memory.  Converter warp q may access TMEM lanes 32q..32q+31, so thread (q, lane)
          // takes row 32q + lane of the staged K-major tile (eight swizzled 16-byte chunks), writes
          // hi = trunc(x) to 32 columns and lo = x - hi to the next 32.  B: lo image into the lo ring.
          tc_fence_after();                                      // TMEM slot freed by retired MMAs (mbarrier observed)
          const int q = warp - kTcEpiWarps, r = 32 * q + lane;
          const uint32_t rowp = smem_base + stage * kTcStageBytes + r * 128;
          uint32_t x[32];
          if (!A_MN) {
#pragma unroll
            for (int c = 0; c < 8; ++c)
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(x[4 * c]), "=r"(x[4 * c + 1]), "=r"(x[4 * c + 2]), "=r"(x[4 * c + 3])
                           : "r"(rowp + (uint32_t)((c ^ (r & 7)) << 4)));
          } else {
            // MN-major tile: four boxes of [32 k][32 m], 128-byte rows swizzled in 32-byte atoms with a
            // period of four rows (atom ^= k & 3).  A warp reads one 128-byte row per k: conflict-free.
            const uint32_t colp = smem_base + stage * kTcStageBytes + (uint32_t)q * 4096u;
#pragma unroll
            for (int k = 0; k < 32; ++k)
              asm volatile("ld.shared.b32 %0, [%1];" : "=r"(x[k])
                           : "r"(colp + (uint32_t)(k * 128) + (uint32_t)((lane * 4) ^ ((k & 3) << 5))));
          }
          uint32_t y[32];
#pragma unroll
          for (int e = 0; e < 32; ++e) y[e] = x[e] & 0xffffe000u;
          const uint32_t ta = tmem_base + (uint32_t)kTmemCols + (uint32_t)(lstage * 2 * kTcBK) + ((uint32_t)(32 * q) << 16);
          tc_st32(ta, y);
#pragma unroll
          for (int e = 0; e < 32; ++e) y[e] = __float_as_uint(__uint_as_float(x[e]) - __uint_as_float(y[e]));
          tc_st32(ta + kTcBK, y);
          const uint32_t srcb = src + kTcBM * kTcBK * 4;
          if (p.bn == kTcBN) lo_image<8>(srcb, dst);
          else lo_image<4>(srcb, dst);                          // bn = 64: the B tile is 8 KB
          tc_st_wait();
          tc_fence_before();
        } else if (!(p.dbg & 8)) {
#pragma unroll
          for (int i0 = 0; i0 < kIters; i0 += 8) {               // eight 128-bit loads in flight per thread
            uint32_t v[8][4];
#pragma unroll
            for (int u = 0; u < 8; ++u)
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(v[u][0]), "=r"(v[u][1]), "=r"(v[u][2]), "=r"(v[u][3]) : "r"(src + (i0 + u) * 2048));
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              uint32_t hi[4], lo[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                hi[e] = ((v[u][e] + rnd) & 0xffffe000u);
                const float l = __uint_as_float(v[u][e]) - __uint_as_float(hi[e]);
                lo[e] = inplace ? ((__float_as_uint(l) + 0x1000u) & 0xffffe000u) : __float_as_uint(l);
              }
              if (inplace)
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(src + (i0 + u) * 2048), "r"(hi[0]), "r"(hi[1]), "r"(hi[2]), "r"(hi[3]) : "memory");
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (i0 + u) * 2048), "r"(lo[0]), "r"(lo[1]), "r"(lo[2]), "r"(lo[3]) : "memory");
            }
          }
        }
        fence_proxy_async_smem();                                 // st.shared -> visible to the tensor core
        __syncwarp();
        if (lane == 0) mbar_arrive(lofull_bar(lstage));
        if (++stage == num_stages) { stage = 0; phase ^= 1u; }
        if (++lstage == p.lo_stages) { lstage = 0; lphase ^= 1u; }
      }
    }
  } else {
    // ===================== epilogue warps (0..7) =====================
    const int quad = warp & 3;                    // TMEM lane quadrant this warp may access
    const int c_begin = (warp >> 2) * (kTcBN / 64), c_end = c_begin + kTcBN / 64;   // its 32-column chunks
    int iter = 0;
    for (int w = blockIdx.x; w < total_work; w += gridDim.x, ++iter) {
      const WorkItem it = decode_work(p, w);
      const int as = iter & 1;
      mbar_wait(tfull_bar(as), ((uint32_t)iter >> 1) & 1u);
      tc_fence_after();
      if (p.dbg & 4) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty_bar(as));
        continue;
      }
      if (p.epi_tma) {
        // the flag combinations of the large-output layers have their own compile-time instance: the
        // run-time-flag form costs ~54 instructions per output element, which made these layers
        // issue-bound in their epilogue warps (8 000 cycles per 128 x 128 tile)
        const uint32_t te = tempty_bar(as);
        switch (p.epi_sf) {
          case 0: epi_tma_tile<0>(p, &mapC, it, as, iter, warp, lane, ebuf, tmem_base, te, efull_bar, eempty_bar); break;
          case GSC_EPI_SMASK: epi_tma_tile<GSC_EPI_SMASK>(p, &mapC, it, as, iter, warp, lane, ebuf, tmem_base, te, efull_bar, eempty_bar); break;
          case GSC_EPI_BIAS: epi_tma_tile<GSC_EPI_BIAS>(p, &mapC, it, as, iter, warp, lane, ebuf, tmem_base, te, efull_bar, eempty_bar); break;
          case GSC_EPI_BIAS | GSC_EPI_RELU: epi_tma_tile<GSC_EPI_BIAS | GSC_EPI_RELU>(p, &mapC, it, as, iter, warp, lane, ebuf, tmem_base, te, efull_bar, eempty_bar); break;
          case GSC_EPI_BIAS | GSC_EPI_SMASK: epi_tma_tile<GSC_EPI_BIAS | GSC_EPI_SMASK>(p, &mapC, it, as, iter, warp, lane, ebuf, tmem_base, te, efull_bar, eempty_bar); break;
          case GSC_EPI_BIAS | GSC_EPI_SMASK | GSC_EPI_CLAMP1: epi_tma_tile<GSC_EPI_BIAS | GSC_EPI_SMASK | GSC_EPI_CLAMP1>(p, &mapC, it, as, iter, warp, lane, ebuf, tmem_base, te, efull_bar, eempty_bar); break;
          default: epi_tma_tile<-1>(p, &mapC, it, as, iter, warp, lane, ebuf, tmem_base, te, efull_bar, eempty_bar); break;
        }
        continue;
      }
      const int64_t row = (int64_t)it.m0 + quad * 32 + lane;
#pragma unroll 1
      for (int c = c_begin; c < c_end; ++c) {
        float v[32];
        tc_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * kTcBN + c * 32), v);
        const int64_t col0 = (int64_t)it.n0 + c * 32;
        if (row < p.M && col0 < p.N) {
          if (p.splits == 1) {
            float* dst = p.C + row * p.ldc + col0;
            if (col0 + 32 <= p.N && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                float4 o;
                o.x = tc_epilogue(p, v[4 * q + 0], row, col0 + 4 * q + 0);
                o.y = tc_epilogue(p, v[4 * q + 1], row, col0 + 4 * q + 1);
                o.z = tc_epilogue(p, v[4 * q + 2], row, col0 + 4 * q + 2);
                o.w = tc_epilogue(p, v[4 * q + 3], row, col0 + 4 * q + 3);
                reinterpret_cast<float4*>(dst)[q] = o;
              }
            } else {
#pragma unroll
              for (int q = 0; q < 32; ++q)
                if (col0 + q < p.N) dst[q] = tc_epilogue(p, v[q], row, col0 + q);
            }
          } else {
            float* dst = p.partial + ((int64_t)it.split * p.M + row) * p.N + col0;
            if (col0 + 32 <= p.N && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
              for (int q = 0; q < 8; ++q)
                reinterpret_cast<float4*>(dst)[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
            } else {
#pragma unroll
              for (int q = 0; q < 32; ++q)
                if (col0 + q < p.N) dst[q] = v[q];
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(as));
    }
  }

  if ((warp == 0 || warp == 4) && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols));
  }
}

// ---------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

// 2-D fp32 tensor map: inner extent `inner` (contiguous), `rows` rows of `row_stride` floats.
static int make_map(CUtensorMap* map, const float* base, int64_t inner, int64_t rows, int64_t row_stride,
                    int box_inner, int box_rows, int tf32_round, bool mn_major) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return fail(GSC_ECUDA, "gemm_tc: cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)row_stride * 4};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, tf32_round ? CU_TENSOR_MAP_DATA_TYPE_TFLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                  const_cast<float*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(GSC_ECUDA, "gemm_tc: cuTensorMapEncodeTiled failed (%d) inner=%lld rows=%lld stride=%lld", (int)r,
                (long long)inner, (long long)rows, (long long)row_stride);
  return GSC_OK;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// Can this problem run on the TMA-fed tensor-core kernel?  (16-byte aligned bases, row strides
// that are multiples of 4 floats, a consistent major-ness across segments.)
bool gemm_tc_supported(const gsc_gemm_args* a, int64_t M, int64_t N, bool* a_mn, bool* b_mn) {
  if (M < 1 || N < 8 || N > INT32_MAX - 256 || M > INT32_MAX - 256) return false;   // tile coordinates are ints
  bool amn = false, bmn = false;
  for (int s = 0; s < a->num_seg; ++s) {
    if (a->K[s] <= 0) return false;
    const bool ak = a->a_cs[s] == 1, am = a->a_rs[s] == 1 && !ak;
    const bool bk = a->b_rs[s] == 1, bm = a->b_cs[s] == 1 && !bk;
    if (!(ak || am) || !(bk || bm)) return false;
    if (s == 0) { amn = am; bmn = bm; }
    else if (amn != am || bmn != bm) return false;
    const int64_t a_stride = am ? a->a_cs[s] : a->a_rs[s];
    const int64_t b_stride = bm ? a->b_rs[s] : a->b_cs[s];
    if (a_stride % 4 || b_stride % 4 || a_stride <= 0 || b_stride <= 0) return false;
    if (!aligned16(a->A[s]) || !aligned16(a->B[s])) return false;
  }
  if (amn && !bmn) return false;      // not needed by any layer; not instantiated
  *a_mn = amn; *b_mn = bmn;
  return true;
}

static int tc_round_mode() {
  static int mode = -1;
  if (mode < 0) {
    const char* e = getenv("GSC_TF32_TMA_ROUND");
    mode = (e && e[0] == '0') ? 0 : 1;
  }
  return mode;
}

int gemm_tc_launch(const gsc_gemm_args* a, int64_t M, int64_t N, float* C, int64_t ldc, void* workspace,
                   size_t workspace_bytes, bool a_mn, bool b_mn, bool x3, cudaStream_t st, int* used_splits) {
  TcParams p{};
  static const int bn64_env = [] { const char* e = getenv("GSC_GEMM_BN64"); return (e && e[0] == '0') ? 0 : 1; }();
  const int bn = (N <= 64 && bn64_env) ? 64 : kTcBN;
  CUtensorMap maps[8];
  memset(maps, 0, sizeof(maps));
  int kb = 0;
  const int rnd = x3 ? 0 : tc_round_mode();     // 3xTF32 needs the unrounded fp32 values in shared memory
  for (int s = 0; s < a->num_seg; ++s) {
    int rc;
    if (!a_mn) rc = make_map(&maps[s], a->A[s], a->K[s], M, a->a_rs[s], kTcBK, kTcBM, rnd, false);
    else rc = make_map(&maps[s], a->A[s], M, a->K[s], a->a_cs[s], 32, kTcBK, rnd, true);
    if (rc) return rc;
    if (!b_mn) rc = make_map(&maps[3 + s], a->B[s], a->K[s], N, a->b_cs[s], kTcBK, bn, rnd, false);
    else rc = make_map(&maps[3 + s], a->B[s], N, a->K[s], a->b_rs[s], 32, kTcBK, rnd, true);
    if (rc) return rc;
    kb += (int)ceil_div(a->K[s], kTcBK);
    p.kb_end[s] = kb;
  }
  for (int s = a->num_seg; s < 3; ++s) { maps[s] = maps[0]; maps[3 + s] = maps[3]; p.kb_end[s] = kb; }
  p.M = M; p.N = N; p.ldc = ldc; p.C = C; p.flags = a->flags; p.num_seg = a->num_seg;
  p.bias = a->bias; p.mul = a->mul; p.gate = a->gate; p.ld_mul = a->ld_mul; p.ld_gate = a->ld_gate;
  p.row_table = a->row_table; p.row_n = a->row_n; p.ld_table = a->ld_table; p.n_max = a->n_max; p.groups = a->groups;
  p.kb_total = kb;
  p.tiles_m = (int)ceil_div(M, kTcBM);
  p.tiles_n = (int)ceil_div(N, bn);
  p.bn = bn;
  const int64_t tiles = (int64_t)p.tiles_m * p.tiles_n;
  // split K when the output tiles alone cannot fill the machine; keep >= 4 k-blocks per split
  // (tiles * splits work items on kNumSMs persistent CTAs: never more items than CTAs, or a few CTAs
  // would run two of them and the launch would take two rounds)
  int64_t want = tiles >= kNumSMs ? 1 : kNumSMs / tiles;
  int64_t max_by_k = kb / 4 > 0 ? kb / 4 : 1;
  if (want > max_by_k) want = max_by_k;
  if (workspace == nullptr) want = 1;
  while (want > 1 && (size_t)want * (size_t)M * (size_t)N * 4 > workspace_bytes) --want;
  static const int il_env = [] { const char* e = getenv("GSC_SPLITK_INTERLEAVE"); return (e && e[0] == '0') ? 0 : 1; }();
  p.kb_interleave = il_env;
  p.kb_per_split = (int)ceil_div(kb, want);
  p.splits = p.kb_interleave ? (int)want : (int)ceil_div(kb, p.kb_per_split);
  p.partial = (float*)workspace;
  *used_splits = p.splits;
  // TMA epilogue: needs a TMA-mappable output (and mask/gate operand); split-K partials keep direct stores
  maps[6] = maps[0]; maps[7] = maps[0];
  p.epi_tma = 0; p.epi_kind = 0; p.epi_sf = -1;
  const char* no_tma_epi = getenv("GSC_GEMM_DIRECT_EPILOGUE");
  if (p.splits == 1 && !(no_tma_epi && no_tma_epi[0] == '1') && aligned16(C) && ldc % 4 == 0) {
    const float* e = nullptr; int64_t lde = 0; int kind = 0;
    if (a->flags & GSC_EPI_MUL) { e = a->mul; lde = a->ld_mul; kind = 1; }
    else if (a->flags & GSC_EPI_GATE) { e = a->gate; lde = a->ld_gate; kind = 2; }
    const bool e_ok = kind == 0 || (aligned16(e) && lde % 4 == 0);
    if (e_ok) {
      int rc = make_map(&maps[6], C, N, M, ldc, 32, kTcBM, 0, false);
      if (rc) return rc;
      if (kind) { rc = make_map(&maps[7], e, N, M, lde, 32, kTcBM, 0, false); if (rc) return rc; }
      p.epi_tma = 1; p.epi_kind = kind;
      const int ef = a->flags & ~(GSC_EPI_TF32 | GSC_EPI_TF32X3);
      static const int sf_env = [] { const char* e = getenv("GSC_GEMM_DYNAMIC_EPILOGUE"); return (e && e[0] == '1') ? 0 : 1; }();
      p.epi_sf = (kind == 0 && sf_env && (ef & ~(GSC_EPI_BIAS | GSC_EPI_RELU | GSC_EPI_SMASK | GSC_EPI_CLAMP1)) == 0) ? ef : -1;
    }
  }
  static const int atm_env = [] { const char* e = getenv("GSC_X3_ATMEM"); return (e && e[0] == '0') ? 0 : 1; }();
  static const int atm_mn_env = [] { const char* e = getenv("GSC_X3_ATMEM_MN"); return (e && e[0] == '0') ? 0 : 1; }();
  p.a_tmem = (x3 && atm_env && (!a_mn || atm_mn_env)) ? 1 : 0;
  if (p.a_tmem) {
    // A operand in tensor memory: the lo ring holds B only (16 KB slots), so the raw ring can be deeper
    p.stages = p.epi_tma ? 4 : 5;
    p.lo_stages = p.epi_tma ? 2 : 4;
    p.lo_slot = kTcBN * kTcBK * 4;
    p.lo_off = p.stages * kTcStageBytes;
    p.ebuf_off = p.lo_off + p.lo_stages * p.lo_slot;
    p.bar_off = kTcX3Slots * kTcStageBytes;
  } else if (x3) {
    p.stages = p.epi_tma ? kTcX3RawEpi : kTcX3RawSplit;
    p.lo_stages = p.epi_tma ? kTcX3LoEpi : kTcX3LoSplit;
    p.lo_slot = kTcStageBytes;
    p.lo_off = p.stages * kTcStageBytes;
    p.ebuf_off = (kTcX3RawEpi + kTcX3LoEpi) * kTcStageBytes;
    p.bar_off = kTcX3Slots * kTcStageBytes;
  } else {
    p.stages = p.epi_tma ? kTcEpiStages : kTcMaxStages;
    p.lo_stages = 0;
    p.lo_slot = 0;
    p.lo_off = p.stages * kTcStageBytes;
    p.ebuf_off = kTcEpiStages * kTcStageBytes;
    p.bar_off = kTcMaxStages * kTcStageBytes;
  }
  { const char* e = getenv("GSC_TC_DBG"); p.dbg = e ? atoi(e) : 0; }
  const int64_t work = tiles * p.splits;
  const int grid = (int)(work < kNumSMs ? work : kNumSMs);
  const int smem = x3 ? kTcSmemBytesX3 : kTcSmemBytes;
  const int threads = x3 ? kTcThreadsX3 : kTcThreads;

  auto launch = [&](auto kernel) -> int {
    GSC_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    kernel<<<grid, threads, smem, st>>>(p, maps[0], maps[1], maps[2], maps[3], maps[4], maps[5], maps[6], maps[7]);
    GSC_LAUNCH_CHECK("gemm_tc_kernel");
    return GSC_OK;
  };
  if (x3) {
    if (!a_mn && !b_mn) return launch(gemm_tc_kernel<false, false, true>);
    if (!a_mn && b_mn) return launch(gemm_tc_kernel<false, true, true>);
    return launch(gemm_tc_kernel<true, true, true>);
  }
  if (!a_mn && !b_mn) return launch(gemm_tc_kernel<false, false, false>);
  if (!a_mn && b_mn) return launch(gemm_tc_kernel<false, true, false>);
  return launch(gemm_tc_kernel<true, true, false>);
}

}  // namespace gsc
