// Multi-tensor fused optimizers: one launch updates up to DTB_MT_MAX_TENSORS
// parameters (bf16 or fp32), optionally through fp32 master weights, reading the
// gradient once and writing parameter (+ master, + moments) once.
// Replaces torch.optim.SGD's _foreach_ kernels (SURVEY K13;
// /root/reference/src/distributed_trainer.py:165,200) and adds AdamW.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

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

// 4 consecutive elements per thread per step (8 B bf16 / 16 B fp32 accesses).
template <bool FP32>
__device__ __forceinline__ void load4(const void* p, long long i, float (&f)[4]) {
  if (FP32) {
    float4 t = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p) + i);
    f[0] = t.x; f[1] = t.y; f[2] = t.z; f[3] = t.w;
  } else {
    uint2 t = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(p) + i);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
    float2 a = __bfloat1622float2(h[0]), b = __bfloat1622float2(h[1]);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
  }
}
template <bool FP32>
__device__ __forceinline__ void store4(void* p, long long i, const float (&f)[4]) {
  if (FP32) {
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(p) + i) = make_float4(f[0], f[1], f[2], f[3]);
  } else {
    uint2 t;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&t);
    h[0] = __floats2bfloat162_rn(f[0], f[1]);
    h[1] = __floats2bfloat162_rn(f[2], f[3]);
    *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p) + i) = t;
  }
}

struct Hyper {
  float lr, momentum, wd, beta1, beta2, eps, bc1, bc2;
};

template <bool ADAM>
__device__ __forceinline__ float update(float w, float g, float* m, float* v, const Hyper& h) {
  if (ADAM) {
    *m = h.beta1 * (*m) + (1.f - h.beta1) * g;
    *v = h.beta2 * (*v) + (1.f - h.beta2) * g * g;
    const float mhat = *m / h.bc1;
    const float vhat = *v / h.bc2;
    return w * (1.f - h.lr * h.wd) - h.lr * mhat / (sqrtf(vhat) + h.eps);
  }
  g += h.wd * w;
  if (m != nullptr) {
    *m = h.momentum * (*m) + g;
    g = *m;
  }
  return w - h.lr * g;
}

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
    const bool aligned = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g)) & 15) == 0;
    const long long nvec = aligned ? n / 4 : 0;
    for (long long i = tid; i < nvec; i += nthreads) {
      float w[4], gr[4], mm[4] = {0, 0, 0, 0}, vv[4] = {0, 0, 0, 0};
      if (master) load4<true>(master, i * 4, w);
      else load4<P32>(p, i * 4, w);
      load4<G32>(g, i * 4, gr);
      if (m) load4<true>(m, i * 4, mm);
      if (ADAM) load4<true>(v, i * 4, vv);
#pragma unroll
      for (int j = 0; j < 4; ++j) w[j] = update<ADAM>(w[j], gr[j], m ? &mm[j] : nullptr, &vv[j], h);
      if (master) store4<true>(master, i * 4, w);
      store4<P32>(p, i * 4, w);
      if (m) store4<true>(m, i * 4, mm);
      if (ADAM) store4<true>(v, i * 4, vv);
    }
    for (long long i = nvec * 4 + tid; i < n; i += nthreads) {
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
  long long blocks = (total / 4 + 255) / 256;
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
  optim::Hyper h = {lr, momentum, weight_decay, 0.f, 0.f, 0.f, 1.f, 1.f};
  return optim::launch<false>(args, h, static_cast<cudaStream_t>(stream));
}

extern "C" int dtb_multi_tensor_adamw(const dtb_mt_args* args, float lr, float beta1, float beta2, float eps,
                                      float weight_decay, float bias_corr1, float bias_corr2, void* stream) {
  DTB_REQUIRE(args && args->count > 0 && args->count <= DTB_MT_MAX_TENSORS, "dtb_multi_tensor_adamw: bad tensor count");
  for (int t = 0; t < args->count; ++t)
    DTB_REQUIRE(args->m[t] && args->v[t], "dtb_multi_tensor_adamw: moments missing for tensor %d", t);
  optim::Hyper h = {lr, 0.f, weight_decay, beta1, beta2, eps, bias_corr1, bias_corr2};
  return optim::launch<true>(args, h, static_cast<cudaStream_t>(stream));
}
