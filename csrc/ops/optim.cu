// Multi-tensor fused optimizers: one launch updates up to DTB_MT_MAX_TENSORS
// parameters (bf16 or fp32), optionally through fp32 master weights, reading the
// gradient once and writing parameter (+ master, + moments) once.
// Replaces torch.optim.SGD's _foreach_ kernels (SURVEY K13;
// /root/reference/src/distributed_trainer.py:165,200) and adds AdamW.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cmath>
#include <cstdint>

#include "../../include/dtb200.h"
#include "../common/host.h"

namespace dtb {
namespace optim {

template <bool FP32>
__device__ __forceinline__ float load_one(const void* p, long long i) {
  if (FP32) return reinterpret_cast<const float*>(p)[i];
  return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
}
template <bool FP32>
__device__ __forceinline__ void store_one(void* p, long long i, float v) {
  if (FP32) reinterpret_cast<float*>(p)[i] = v;
  else reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16(v);
}

// Hyper-parameters with everything that does not depend on the element folded on the host: the ncu capture of the
// first version (profiles/kernel_zoo_ncu_metrics.csv) had AdamW issue-bound at 184 instructions per element -- three
// IEEE divisions and an IEEE square root each -- at 0.53 of the HBM copy rate.
struct Hyper {
  float lr, momentum, wd;
  float beta1, beta2, omb1, omb2;   // omb = 1 - beta
  float eps, decay, step, rs_bc2;   // decay = 1 - lr*wd, step = lr / bias_corr1, rs_bc2 = 1 / sqrt(bias_corr2)
};

__device__ __forceinline__ float sqrt_approx(float x) {
  float y;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// AdamW as torch.optim.AdamW writes it: p *= 1 - lr*wd; m, v updated; p -= (lr/bc1) * m / (sqrt(v)/sqrt(bc2) + eps).
// sqrt.approx / rcp.approx are within 1-2 ulp, far below the bf16 rounding of the parameter and the fp32 noise of v.
template <bool ADAM>
__device__ __forceinline__ float update(float w, float g, float* m, float* v, const Hyper& h) {
  if (ADAM) {
    *m = fmaf(h.beta1, *m, h.omb1 * g);
    *v = fmaf(h.beta2, *v, h.omb2 * g * g);
    const float denom = fmaf(sqrt_approx(*v), h.rs_bc2, h.eps);
    return fmaf(w, h.decay, -h.step * (*m) * rcp_approx(denom));
  }
  g = fmaf(h.wd, w, g);
  if (m != nullptr) {
    *m = fmaf(h.momentum, *m, g);
    g = *m;
  }
  return fmaf(-h.lr, g, w);
}

template <bool FP32>
__device__ __forceinline__ void load8(const void* p, long long i, float (&f)[8]) {
  if (FP32) {
    const float4* q = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p) + i);
    const float4 a = q[0], b = q[1];
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
    f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
  } else {
    const uint4 t = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(p) + i);
    const uint32_t u[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      f[2 * j] = __uint_as_float(u[j] << 16);
      f[2 * j + 1] = __uint_as_float(u[j] & 0xffff0000u);
    }
  }
}
template <bool FP32>
__device__ __forceinline__ void store8(void* p, long long i, const float (&f)[8]) {
  if (FP32) {
    float4* q = reinterpret_cast<float4*>(reinterpret_cast<float*>(p) + i);
    q[0] = make_float4(f[0], f[1], f[2], f[3]);
    q[1] = make_float4(f[4], f[5], f[6], f[7]);
  } else {
    uint4 t;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&t);
#pragma unroll
    for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
    *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p) + i) = t;
  }
}

// 8 consecutive elements per thread per step: one 16-byte access per bf16 stream, two per fp32 stream, every load of
// the step issued before the first use.
template <bool ADAM, bool P32, bool G32>
__global__ void __launch_bounds__(256) mt_kernel(const dtb_mt_args a, const Hyper h) {
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
  for (int t = 0; t < a.count; ++t) {
    void* p = a.param[t];
    const void* g = a.grad[t];
    float* master = a.master_on ? a.master[t] : nullptr;
    float* m = a.m[t];
    float* v = a.v[t];
    const long long n = a.numel[t];
    const uintptr_t bits = reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                           reinterpret_cast<uintptr_t>(master) | reinterpret_cast<uintptr_t>(m) |
                           reinterpret_cast<uintptr_t>(v);
    const long long nvec = (bits & 15) == 0 ? n / 8 : 0;
    for (long long i = tid; i < nvec; i += nthreads) {
      float w[8], gr[8], mm[8] = {}, vv[8] = {};
      if (master) load8<true>(master, i * 8, w);
      else load8<P32>(p, i * 8, w);
      load8<G32>(g, i * 8, gr);
      if (m) load8<true>(m, i * 8, mm);
      if (ADAM) load8<true>(v, i * 8, vv);
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = update<ADAM>(w[j], gr[j], m ? &mm[j] : nullptr, &vv[j], h);
      if (master) store8<true>(master, i * 8, w);
      store8<P32>(p, i * 8, w);
      if (m) store8<true>(m, i * 8, mm);
      if (ADAM) store8<true>(v, i * 8, vv);
    }
    for (long long i = nvec * 8 + tid; i < n; i += nthreads) {
      float w = master ? master[i] : load_one<P32>(p, i);
      float gr = load_one<G32>(g, i);
      float mm = m ? m[i] : 0.f, vv = ADAM ? v[i] : 0.f;
      w = update<ADAM>(w, gr, m ? &mm : nullptr, &vv, h);
      if (master) master[i] = w;
      store_one<P32>(p, i, w);
      if (m) m[i] = mm;
      if (ADAM) v[i] = vv;
    }
  }
}

template <bool ADAM>
static int launch(const dtb_mt_args* a, const Hyper& h, cudaStream_t s) {
  long long total = 0;
  for (int t = 0; t < a->count; ++t) total += a->numel[t];
  long long blocks = (total / 8 + 255) / 256;
  const long long cap = static_cast<long long>(sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  const int grid = static_cast<int>(blocks);
  if (a->param_fp32 && a->grad_fp32) mt_kernel<ADAM, true, true><<<grid, 256, 0, s>>>(*a, h);
  else if (a->param_fp32) mt_kernel<ADAM, true, false><<<grid, 256, 0, s>>>(*a, h);
  else if (a->grad_fp32) mt_kernel<ADAM, false, true><<<grid, 256, 0, s>>>(*a, h);
  else mt_kernel<ADAM, false, false><<<grid, 256, 0, s>>>(*a, h);
  DTB_LAUNCH_OK();
  return 0;
}

}  // namespace optim
}  // namespace dtb

using namespace dtb;

extern "C" int dtb_multi_tensor_sgd(const dtb_mt_args* args, float lr, float momentum, float weight_decay,
                                    void* stream) {
  DTB_REQUIRE(args && args->count > 0 && args->count <= DTB_MT_MAX_TENSORS, "dtb_multi_tensor_sgd: bad tensor count");
  optim::Hyper h = {lr, momentum, weight_decay, 0.f, 0.f, 1.f, 1.f, 0.f, 1.f, lr, 1.f};
  return optim::launch<false>(args, h, static_cast<cudaStream_t>(stream));
}

extern "C" int dtb_multi_tensor_adamw(const dtb_mt_args* args, float lr, float beta1, float beta2, float eps,
                                      float weight_decay, float bias_corr1, float bias_corr2, void* stream) {
  DTB_REQUIRE(args && args->count > 0 && args->count <= DTB_MT_MAX_TENSORS, "dtb_multi_tensor_adamw: bad tensor count");
  for (int t = 0; t < args->count; ++t)
    DTB_REQUIRE(args->m[t] && args->v[t], "dtb_multi_tensor_adamw: moments missing for tensor %d", t);
  optim::Hyper h = {lr, 0.f, weight_decay, beta1, beta2, 1.f - beta1, 1.f - beta2, eps, 1.f - lr * weight_decay,
                    lr / bias_corr1, 1.f / sqrtf(bias_corr2)};
  return optim::launch<true>(args, h, static_cast<cudaStream_t>(stream));
}
