// Actor head: exploration step of _choose_action (reference src/rlsp/agents/simple_ddpg.py:190-199)
// and post_process_actions / normalize_scheduling_probabilities (reference :374-395, :338-372).
//
// Numeric types follow what the reference's pinned numpy 1.23.4 does with its scalars: the fp32
// action elements meet Python ints/floats and the fp64 `scheduling_accuracy`, so every sum,
// quotient and round(., 10) is carried out in fp64 on the exactly-widened fp32 inputs; the
// exploration arithmetic is fp32 array math with the fp64 noise added in fp64 and rounded back.
// Sums are sequential in destination order (Python's sum), so one thread owns one
// (source, sfc, sf) group of `group` destinations; nothing is reassociated and no FMA contraction
// is allowed (explicit *_rn intrinsics).
#include "common.cuh"
#include <stdlib.h>

namespace gsc {

constexpr double kSchedThreshold = 0.1;                        // simple_ddpg.py:130
constexpr double kSchedAccuracy = 1.4901161193847656e-08;      // sqrt(eps_f64), simple_ddpg.py:131

// numpy round(x, 10) on an fp64 scalar: rint(x * 1e10) / 1e10.
__device__ __forceinline__ double np_round10(double x) { return __ddiv_rn(rint(__dmul_rn(x, 1e10)), 1e10); }

// One threshold + normalise pass over v[0..n) in place (simple_ddpg.py:385-386 and :338-372).
__device__ void normalize_pass(double* __restrict__ v, int n, double uniform) {
  double total = 0.0;
  int kept = 0;
  for (int k = 0; k < n; ++k)
    if (v[k] >= kSchedThreshold) { total = __dadd_rn(total, v[k]); ++kept; }
  double s = 0.0;
  if (kept == 0) {                      // sum == 0 -> equal split, round(1/len, 10) as Python floats
    for (int k = 0; k < n; ++k) { v[k] = uniform; s = __dadd_rn(s, uniform); }
  } else {
    const bool renorm = fabs(__dsub_rn(1.0, total)) > kSchedAccuracy;
    for (int k = 0; k < n; ++k) {
      double r = (v[k] >= kSchedThreshold) ? v[k] : 0.0;
      if (renorm) r = np_round10(__ddiv_rn(r, total));
      v[k] = r;
      s = __dadd_rn(s, r);
    }
  }
  const double off = __dsub_rn(1.0, s);
  if (off != 0.0) {
    int i = 0;
    while (i < n - 1 && __dadd_rn(v[i], off) < 0.0) ++i;
    v[i] = __dadd_rn(v[i], off);
  }
}

__global__ void __launch_bounds__(128) actor_head_kernel(const float* __restrict__ y, const double* __restrict__ noise,
                                                         int64_t num_groups, int group, float low, float high,
                                                         double uniform, double* __restrict__ out) {
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < num_groups;
       g += (int64_t)gridDim.x * blockDim.x) {
    const float* yi = y + g * group;
    double* o = out + g * group;
    for (int k = 0; k < group; ++k) {
      float a = yi[k];
      if (noise != nullptr) {
        // scale_action, += N(mu, sigma), unscale_action, clip   (models.py:127-144, simple_ddpg.py:192-199)
        const float span = __fsub_rn(high, low);
        float sc = __fsub_rn(__fmul_rn(2.0f, __fdiv_rn(__fsub_rn(a, low), span)), 1.0f);
        sc = (float)__dadd_rn((double)sc, noise[g * group + k]);
        a = __fadd_rn(low, __fmul_rn(__fmul_rn(0.5f, __fadd_rn(sc, 1.0f)), span));
        a = fminf(fmaxf(a, low), high);
      }
      o[k] = (double)a;
    }
    normalize_pass(o, group, uniform);   // "for _ in range(2)" at simple_ddpg.py:384
    normalize_pass(o, group, uniform);
  }
}

}  // namespace gsc

using namespace gsc;

extern "C" int gsc_actor_head(const float* y, const double* noise, int64_t R, int64_t A, int group,
                              float low, float high, double* out, void* stream) {
  GSC_REQUIRE(y && out && R >= 0 && A >= 1 && group >= 1 && A % group == 0,
              "gsc_actor_head: bad arguments (A=%lld must be a multiple of group=%d)", (long long)A, group);
  if (R == 0) return GSC_OK;
  // Python: round(1 / len(input_list), 10) -- correctly rounded decimal, as printf/strtod give.
  char buf[64];
  snprintf(buf, sizeof buf, "%.10f", 1.0 / (double)group);
  const double uniform = strtod(buf, nullptr);
  const int64_t num_groups = R * (A / group);
  int64_t blocks = ceil_div(num_groups, 128);
  if (blocks > 16LL * kNumSMs) blocks = 16LL * kNumSMs;
  actor_head_kernel<<<(int)blocks, 128, 0, (cudaStream_t)stream>>>(y, noise, num_groups, group, low, high,
                                                                   uniform, out);
  GSC_LAUNCH_CHECK("actor_head_kernel");
  return GSC_OK;
}
