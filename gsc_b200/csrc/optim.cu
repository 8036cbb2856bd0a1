// Fused Adam + soft target update, and the TD-target/MSE kernel.
//
// gsc_adam_soft_update: torch 1.13 Adam (reference src/rlsp/agents/simple_ddpg.py:162-163,217,225;
// formula in SURVEY.md section 3.5) followed by target <- tau*p + (1-tau)*target (reference
// simple_ddpg.py:229-234) in ONE pass over a flat parameter buffer.  Pure HBM streaming: reads
// p, g, m, v, target (20 B/param) and writes p, m, v, target (16 B/param); 128-bit accesses,
// grid = a multiple of the SM count.  The reference does the same in ~6 separate passes per tensor.
#include "common.cuh"

namespace gsc {

struct AdamScalars {
  float lr_over_bc1, sqrt_bc2, beta1, beta2, eps, tau;
};

// target <- tau*p + (1-tau)*target with a fixed operation order (one product, one fused multiply-add): the
// fused Adam pass and the stand-alone soft update must agree bit for bit, because data-parallel ranks with a
// sharded optimizer compute the same target element through either of them.
__device__ __forceinline__ float soft_mix(float tau, float p, float t) {
  return __fmaf_rn(tau, p, __fmul_rn(1.f - tau, t));
}

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float* tgt, const AdamScalars& s) {
  m = s.beta1 * m + (1.f - s.beta1) * g;
  v = s.beta2 * v + (1.f - s.beta2) * g * g;
  const float denom = sqrtf(v) / s.sqrt_bc2 + s.eps;
  p = p - s.lr_over_bc1 * (m / denom);
  if (tgt) *tgt = soft_mix(s.tau, p, *tgt);
}

__global__ void __launch_bounds__(256) adam_soft_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                        float* __restrict__ m, float* __restrict__ v,
                                                        float* __restrict__ tgt, int64_t n, AdamScalars s) {
  const int64_t n4 = n >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float4* p4 = reinterpret_cast<float4*>(p);
  const float4* g4 = reinterpret_cast<const float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  float4* t4 = reinterpret_cast<float4*>(tgt);
  for (int64_t i = tid; i < n4; i += stride) {
    float4 pp = p4[i], gg = g4[i], mm = m4[i], vv = v4[i];
    float4 tt = tgt ? t4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    adam_one(pp.x, gg.x, mm.x, vv.x, tgt ? &tt.x : nullptr, s);
    adam_one(pp.y, gg.y, mm.y, vv.y, tgt ? &tt.y : nullptr, s);
    adam_one(pp.z, gg.z, mm.z, vv.z, tgt ? &tt.z : nullptr, s);
    adam_one(pp.w, gg.w, mm.w, vv.w, tgt ? &tt.w : nullptr, s);
    p4[i] = pp; m4[i] = mm; v4[i] = vv;
    if (tgt) t4[i] = tt;
  }
  for (int64_t i = (n4 << 2) + tid; i < n; i += stride)
    adam_one(p[i], g[i], m[i], v[i], tgt ? tgt + i : nullptr, s);
}

// Device-side step counter variant (CUDA-graph replays cannot re-bake a host scalar): every thread
// derives the bias corrections from *step_dev + 1; adam_step_inc_kernel bumps the counter afterwards.
__global__ void __launch_bounds__(256) adam_soft_dev_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                            float* __restrict__ m, float* __restrict__ v,
                                                            float* __restrict__ tgt, int64_t n,
                                                            const int64_t* __restrict__ step_dev, float lr, float beta1,
                                                            float beta2, float eps, float tau) {
  const double step = (double)(*step_dev + 1);
  AdamScalars s;
  s.lr_over_bc1 = (float)((double)lr / (1.0 - pow((double)beta1, step)));
  s.sqrt_bc2 = (float)sqrt(1.0 - pow((double)beta2, step));
  s.beta1 = beta1; s.beta2 = beta2; s.eps = eps; s.tau = tau;
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                     reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(tgt)) & 15) == 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t done = 0;
  if (vec) {
    const int64_t n4 = n >> 2;
    float4* p4 = reinterpret_cast<float4*>(p);
    const float4* g4 = reinterpret_cast<const float4*>(g);
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* v4 = reinterpret_cast<float4*>(v);
    float4* t4 = reinterpret_cast<float4*>(tgt);
    for (int64_t i = tid; i < n4; i += stride) {
      float4 pp = p4[i], gg = g4[i], mm = m4[i], vv = v4[i];
      float4 tt = tgt ? t4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
      adam_one(pp.x, gg.x, mm.x, vv.x, tgt ? &tt.x : nullptr, s);
      adam_one(pp.y, gg.y, mm.y, vv.y, tgt ? &tt.y : nullptr, s);
      adam_one(pp.z, gg.z, mm.z, vv.z, tgt ? &tt.z : nullptr, s);
      adam_one(pp.w, gg.w, mm.w, vv.w, tgt ? &tt.w : nullptr, s);
      p4[i] = pp; m4[i] = mm; v4[i] = vv;
      if (tgt) t4[i] = tt;
    }
    done = n4 << 2;
  }
  for (int64_t i = done + tid; i < n; i += stride) adam_one(p[i], g[i], m[i], v[i], tgt ? tgt + i : nullptr, s);
}

__global__ void adam_step_inc_kernel(int64_t* step_dev) { *step_dev += 1; }

// Scalar variant for buffers that are not 16-byte aligned.
__global__ void __launch_bounds__(256) adam_soft_scalar_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                               float* __restrict__ m, float* __restrict__ v,
                                                               float* __restrict__ tgt, int64_t n, AdamScalars s) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    adam_one(p[i], g[i], m[i], v[i], tgt ? tgt + i : nullptr, s);
}

// Single CTA: B is a minibatch size (<= a few thousand per rank).
__global__ void __launch_bounds__(1024) td_mse_kernel(const float* __restrict__ r, const float* __restrict__ d,
                                                      const float* __restrict__ qn, const float* __restrict__ q,
                                                      float gamma, float scale, int64_t B, float* __restrict__ y,
                                                      float* __restrict__ dq, float* __restrict__ loss) {
  __shared__ float red[32];
  float s = 0.f;
  for (int64_t i = threadIdx.x; i < B; i += blockDim.x) {
    const float yi = r[i] + (1.f - d[i]) * gamma * qn[i];
    const float diff = q[i] - yi;
    if (y) y[i] = yi;
    dq[i] = 2.f * scale * diff;
    s = fmaf(diff, diff, s);
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) loss[0] = t * scale;
  }
}

// out[i,j] = x[i,j] * mul[i,j] (if mul) and zeroed where gate[i,j] <= 0 (if gate).
__global__ void mul_gate_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ mul,
                                int64_t ldm, const float* __restrict__ gate, int64_t ldg, int64_t M, int64_t N,
                                float* __restrict__ out, int64_t ldo) {
  const int64_t total = M * N;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = t / N, j = t - i * N;
    float v = x[i * ldx + j];
    if (mul) v *= mul[i * ldm + j];
    if (gate) v = gate[i * ldg + j] > 0.f ? v : 0.f;
    out[i * ldo + j] = v;
  }
}

}  // namespace gsc

using namespace gsc;

extern "C" int gsc_mul_gate(const float* x, int64_t ldx, const float* mul, int64_t ld_mul, const float* gate,
                            int64_t ld_gate, int64_t M, int64_t N, float* out, int64_t ldo, void* stream) {
  GSC_REQUIRE(x && out && M >= 0 && N >= 0, "gsc_mul_gate: bad arguments");
  if (M * N == 0) return GSC_OK;
  int64_t blocks = ceil_div(M * N, 256);
  if (blocks > 8LL * kNumSMs) blocks = 8LL * kNumSMs;
  mul_gate_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(x, ldx, mul, ld_mul, gate, ld_gate, M, N, out, ldo);
  GSC_LAUNCH_CHECK("mul_gate_kernel");
  return GSC_OK;
}

// target <- tau*p + (1-tau)*target over a flat range (the soft update alone: data-parallel ranks with a
// sharded optimizer apply it to the parameter slices they received from the other ranks)
__global__ void __launch_bounds__(256) soft_update_kernel(const float* __restrict__ p, float* __restrict__ tgt,
                                                          int64_t n, float tau) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(tgt)) & 15) == 0) {
    const int64_t n4 = n >> 2;
    const float4* p4 = reinterpret_cast<const float4*>(p);
    float4* t4 = reinterpret_cast<float4*>(tgt);
    for (int64_t i = tid; i < n4; i += stride) {
      const float4 pp = p4[i];
      float4 tt = t4[i];
      tt.x = soft_mix(tau, pp.x, tt.x); tt.y = soft_mix(tau, pp.y, tt.y);
      tt.z = soft_mix(tau, pp.z, tt.z); tt.w = soft_mix(tau, pp.w, tt.w);
      t4[i] = tt;
    }
    for (int64_t i = (n4 << 2) + tid; i < n; i += stride) tgt[i] = soft_mix(tau, p[i], tgt[i]);
  } else {
    for (int64_t i = tid; i < n; i += stride) tgt[i] = soft_mix(tau, p[i], tgt[i]);
  }
}

extern "C" int gsc_soft_update(const float* p, float* target, int64_t n, float tau, void* stream) {
  GSC_REQUIRE(n >= 0 && (n == 0 || (p && target)), "gsc_soft_update: bad arguments");
  if (n == 0) return GSC_OK;
  int64_t blocks = ceil_div(n, 256 * 4);
  if (blocks > 8 * kNumSMs) blocks = 8 * kNumSMs;
  soft_update_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(p, target, n, tau);
  GSC_LAUNCH_CHECK("soft_update_kernel");
  return GSC_OK;
}

extern "C" int gsc_adam_soft_update(float* p, const float* g, float* m, float* v, float* target, int64_t n,
                                    float lr, float beta1, float beta2, float eps, int64_t step, float tau,
                                    void* stream) {
  GSC_REQUIRE(p && g && m && v && n >= 0 && step >= 1, "gsc_adam_soft_update: bad arguments");
  if (n == 0) return GSC_OK;
  AdamScalars s;
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  s.lr_over_bc1 = (float)((double)lr / bc1);
  s.sqrt_bc2 = (float)sqrt(bc2);
  s.beta1 = beta1; s.beta2 = beta2; s.eps = eps; s.tau = tau;
  int64_t blocks = ceil_div(ceil_div(n, 4), 256);
  const int64_t cap = 8LL * kNumSMs;
  if (blocks > cap) blocks = cap;
  const uintptr_t bits = (uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v | (uintptr_t)target;
  if ((bits & 15) == 0)
    adam_soft_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, target, n, s);
  else
    adam_soft_scalar_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, target, n, s);
  GSC_LAUNCH_CHECK("adam_soft_kernel");
  return GSC_OK;
}

extern "C" int gsc_adam_soft_update_dev(float* p, const float* g, float* m, float* v, float* target, int64_t n,
                                        float lr, float beta1, float beta2, float eps, int64_t* step_dev, float tau,
                                        void* stream) {
  GSC_REQUIRE(p && g && m && v && step_dev && n >= 0, "gsc_adam_soft_update_dev: bad arguments");
  if (n == 0) return GSC_OK;
  int64_t blocks = ceil_div(ceil_div(n, 4), 256);
  const int64_t cap = 8LL * kNumSMs;
  if (blocks > cap) blocks = cap;
  adam_soft_dev_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, target, n, step_dev, lr, beta1, beta2,
                                                                      eps, tau);
  GSC_LAUNCH_CHECK("adam_soft_dev_kernel");
  adam_step_inc_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_dev);
  GSC_LAUNCH_CHECK("adam_step_inc_kernel");
  return GSC_OK;
}

extern "C" int gsc_adam_soft_update_slice(float* p, const float* g, float* m, float* v, float* target, int64_t n,
                                          float lr, float beta1, float beta2, float eps, int64_t* step_dev, float tau,
                                          int bump_step, void* stream) {
  GSC_REQUIRE(p && g && m && v && step_dev && n >= 0, "gsc_adam_soft_update_slice: bad arguments");
  if (n > 0) {
    int64_t blocks = ceil_div(ceil_div(n, 4), 256);
    const int64_t cap = 8LL * kNumSMs;
    if (blocks > cap) blocks = cap;
    adam_soft_dev_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(p, g, m, v, target, n, step_dev, lr, beta1, beta2,
                                                                        eps, tau);
    GSC_LAUNCH_CHECK("adam_soft_dev_kernel");
  }
  if (bump_step) {
    adam_step_inc_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_dev);
    GSC_LAUNCH_CHECK("adam_step_inc_kernel");
  }
  return GSC_OK;
}

extern "C" int gsc_td_mse(const float* rewards, const float* dones, const float* q_next, const float* q,
                          float gamma, float loss_scale, int64_t B, float* y, float* dq, float* loss,
                          void* stream) {
  GSC_REQUIRE(rewards && dones && q_next && q && dq && loss && B >= 1, "gsc_td_mse: bad arguments");
  td_mse_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(rewards, dones, q_next, q, gamma, loss_scale, B, y, dq, loss);
  GSC_LAUNCH_CHECK("td_mse_kernel");
  return GSC_OK;
}
