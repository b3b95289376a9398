// Token + position embedding forward and backward for GPT-2.
//   fwd: out[b,t,:] = wte[tok[b,t],:] + wpe[t,:]
//   bwd: dwte[tok] += dout (bf16x2 atomics: rows repeat inside a batch),
//        dwpe[t] = sum_b dout[b,t,:] (one CTA per position, fixed order).
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../../include/dtb200.h"
#include "../common/host.h"

namespace dtb {
namespace emb {

__global__ void __launch_bounds__(128)
fwd_kernel(const long long* __restrict__ tokens, const __nv_bfloat16* __restrict__ wte,
           const __nv_bfloat16* __restrict__ wpe, __nv_bfloat16* __restrict__ out, int seq, int channels) {
  const int bt = blockIdx.x;
  const int t = bt % seq;
  const long long tok = tokens[bt];
  const uint4* e = reinterpret_cast<const uint4*>(wte + tok * channels);
  const uint4* p = reinterpret_cast<const uint4*>(wpe + static_cast<long long>(t) * channels);
  uint4* o = reinterpret_cast<uint4*>(out + static_cast<long long>(bt) * channels);
  for (int v = threadIdx.x; v < channels / 8; v += blockDim.x) {
    uint4 a = e[v], b = p[v], r;
    const __nv_bfloat162* ha = reinterpret_cast<const __nv_bfloat162*>(&a);
    const __nv_bfloat162* hb = reinterpret_cast<const __nv_bfloat162*>(&b);
    __nv_bfloat162* hr = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) hr[i] = __hadd2(ha[i], hb[i]);
    o[v] = r;
  }
}

__global__ void __launch_bounds__(128)
bwd_wte_kernel(const __nv_bfloat16* __restrict__ dout, const long long* __restrict__ tokens,
               __nv_bfloat16* __restrict__ dwte, int channels) {
  const int bt = blockIdx.x;
  const long long tok = tokens[bt];
  const __nv_bfloat162* g = reinterpret_cast<const __nv_bfloat162*>(dout + static_cast<long long>(bt) * channels);
  __nv_bfloat162* d = reinterpret_cast<__nv_bfloat162*>(dwte + tok * channels);
  for (int v = threadIdx.x; v < channels / 2; v += blockDim.x) atomicAdd(d + v, g[v]);
}

__global__ void __launch_bounds__(128)
bwd_wpe_kernel(const __nv_bfloat16* __restrict__ dout, __nv_bfloat16* __restrict__ dwpe, int batch, int seq,
               int channels) {
  const int t = blockIdx.x;
  for (int v = threadIdx.x; v < channels / 2; v += blockDim.x) {
    float2 acc = make_float2(0.f, 0.f);
    for (int b = 0; b < batch; ++b) {
      const __nv_bfloat162 g =
          reinterpret_cast<const __nv_bfloat162*>(dout + (static_cast<long long>(b) * seq + t) * channels)[v];
      const float2 f = __bfloat1622float2(g);
      acc.x += f.x;
      acc.y += f.y;
    }
    reinterpret_cast<__nv_bfloat162*>(dwpe + static_cast<long long>(t) * channels)[v] = __floats2bfloat162_rn(acc.x, acc.y);
  }
}

}  // namespace emb
}  // namespace dtb

using namespace dtb;

extern "C" int dtb_embedding_fwd(const long long* tokens, const void* wte, const void* wpe, void* out, int batch,
                                 int seq, int channels, void* stream) {
  DTB_REQUIRE(batch > 0 && seq > 0 && channels % 8 == 0, "dtb_embedding_fwd: channels must be a multiple of 8");
  emb::fwd_kernel<<<batch * seq, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      tokens, static_cast<const __nv_bfloat16*>(wte), static_cast<const __nv_bfloat16*>(wpe),
      static_cast<__nv_bfloat16*>(out), seq, channels);
  DTB_LAUNCH_OK();
  return 0;
}

extern "C" int dtb_embedding_bwd(const void* dout, const long long* tokens, void* dwte, void* dwpe, int batch, int seq,
                                 int channels, int accumulate_wte, void* stream) {
  DTB_REQUIRE(batch > 0 && seq > 0 && channels % 8 == 0, "dtb_embedding_bwd: channels must be a multiple of 8");
  DTB_REQUIRE(accumulate_wte, "dtb_embedding_bwd: dwte is accumulated into; zero it first (accumulate_wte=1)");
  auto s = static_cast<cudaStream_t>(stream);
  emb::bwd_wte_kernel<<<batch * seq, 128, 0, s>>>(static_cast<const __nv_bfloat16*>(dout), tokens,
                                                  static_cast<__nv_bfloat16*>(dwte), channels);
  DTB_LAUNCH_OK();
  emb::bwd_wpe_kernel<<<seq, 128, 0, s>>>(static_cast<const __nv_bfloat16*>(dout), static_cast<__nv_bfloat16*>(dwpe),
                                          batch, seq, channels);
  DTB_LAUNCH_OK();
  return 0;
}
