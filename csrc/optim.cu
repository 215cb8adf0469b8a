// Fused optimizer kernels for sm_100a (SURVEY §8 O1 single-rank part, K14–K18 replacement).
//
// The reference's step walks Python lists of parameters: per-parameter overflow checks with a host sync each
// (core/optimizer/loss_scaler.py:23-41), per-parameter unscale + norm (parameter_group.py:483-535), an fp32 copy of the
// gradients (parameter_group.py:346-410), torch.optim.AdamW on the fp32 shard (optimizer.py:179) and a cast back to the
// model dtype (parameter_group.py:423-430).  Here the ZeRO shard is one flat arena, so the whole step is two launches:
//   sb_grad_sumsq : Σg² and a non-finite flag over the flat gradient shard, deterministic two-stage reduction
//   sb_adamw      : unscale · clip (coefficient derived on the device from the reduced norm, no host sync) · AdamW on
//                   the fp32 master/m/v shard · cast to the bf16 model shard · optional gradient zeroing, in one pass.
#include "common.cuh"

namespace sb {
namespace optim {

constexpr int kThreads = 256;
constexpr int kMaxBlocks = 148 * 8;

template <typename G>
SB_DEVICE float ldg(const G* p, size_t i) {
  return (float)p[i];
}

// stage 1: per-block partial sums -> partial[blockIdx], non-finite -> flag
template <typename G>
__global__ void __launch_bounds__(kThreads) sumsq_kernel(const G* __restrict__ g, size_t n, float* __restrict__ partial,
                                                         int* __restrict__ nonfinite) {
  __shared__ float red[kThreads / 32];
  float s = 0.f;
  bool bad = false;
  for (size_t i = blockIdx.x * (size_t)kThreads + threadIdx.x; i < n; i += (size_t)gridDim.x * kThreads) {
    float v = (float)g[i];
    bad |= !isfinite(v);
    s += v * v;
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  if (__syncthreads_or(bad) && threadIdx.x == 0) atomicOr(nonfinite, 1);
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < kThreads / 32; ++i) t += red[i];
    partial[blockIdx.x] = t;
  }
}
// stage 2: fixed-order sum of the partials; out[0] (+)= total
__global__ void sumsq_final_kernel(const float* __restrict__ partial, int parts, float* __restrict__ out, int accumulate) {
  __shared__ float red[kThreads / 32];
  float s = 0.f;
  for (int i = threadIdx.x; i < parts; i += kThreads) s += partial[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < kThreads / 32; ++i) t += red[i];
    out[0] = accumulate ? out[0] + t : t;
  }
}

struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay, bias_corr1, bias_corr2_sqrt;
  float inv_loss_scale;  // gradients are multiplied by this (1 / loss scale)
  float clip;            // max global norm, 0 = off
  int zero_grad;
};

// p32/m/v: fp32 shard; g: gradient shard (G); out: model-dtype shard (P) written with the updated value.
// norm_sq: device scalar with the *global* Σg² of the still-scaled gradients (may be null when clip == 0).
// skip: device flag, non-zero -> leave everything untouched (overflow step).
template <typename T>
struct V4 {
  static SB_DEVICE void load(const T* p, float* f);
  static SB_DEVICE void store(T* p, const float* f);
};
template <>
struct V4<float> {
  static SB_DEVICE void load(const float* p, float* f) {
    float4 v = *reinterpret_cast<const float4*>(p);
    f[0] = v.x, f[1] = v.y, f[2] = v.z, f[3] = v.w;
  }
  static SB_DEVICE void store(float* p, const float* f) { *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]); }
};
template <>
struct V4<__nv_bfloat16> {
  static SB_DEVICE void load(const __nv_bfloat16* p, float* f) {
    uint2 v = *reinterpret_cast<const uint2*>(p);
    float2 a = unpack_bf16(v.x), b = unpack_bf16(v.y);
    f[0] = a.x, f[1] = a.y, f[2] = b.x, f[3] = b.y;
  }
  static SB_DEVICE void store(__nv_bfloat16* p, const float* f) {
    *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]));
  }
};
template <>
struct V4<__half> {
  static SB_DEVICE void load(const __half* p, float* f) {
    uint2 v = *reinterpret_cast<const uint2*>(p);
    const __half2* h = reinterpret_cast<const __half2*>(&v);
    float2 a = __half22float2(h[0]), b = __half22float2(h[1]);
    f[0] = a.x, f[1] = a.y, f[2] = b.x, f[3] = b.y;
  }
  static SB_DEVICE void store(__half* p, const float* f) {
    uint2 v;
    __half2* h = reinterpret_cast<__half2*>(&v);
    h[0] = __floats2half2_rn(f[0], f[1]);
    h[1] = __floats2half2_rn(f[2], f[3]);
    *reinterpret_cast<uint2*>(p) = v;
  }
};

// p32/m/v: fp32 shard; g: gradient shard (G); out: model-dtype shard (P) written with the updated value.
// norm_sq: device scalar with the *global* Σg² of the still-scaled gradients (may be null when clip == 0).
// skip: device flag, non-zero -> leave everything untouched (overflow step).
// Four elements per thread with 128-bit accesses on the fp32 streams; a scalar tail handles n % 4.
// `outs`: where the updated model-dtype values go — the rank's own parameter arena and, under fused ZeRO, the same
// offset of every data-parallel peer's arena (peer-mapped pointers: the all-gather is these stores).
constexpr int kMaxPeers = 8;
template <typename P>
struct OutPtrs {
  P* p[kMaxPeers];
  int n;
};
template <typename G, typename P>
__global__ void __launch_bounds__(kThreads) adamw_kernel(float* __restrict__ p32, float* __restrict__ m,
                                                         float* __restrict__ v, G* __restrict__ g, OutPtrs<P> outs,
                                                         size_t n, AdamArgs a, const float* __restrict__ norm_sq,
                                                         const int* __restrict__ skip) {
  if (skip && *skip) return;
  float gs = a.inv_loss_scale;
  if (a.clip > 0.f && norm_sq) {
    const float norm = sqrtf(*norm_sq) * a.inv_loss_scale;
    const float coef = a.clip / (norm + 1e-6f);
    if (coef < 1.f) gs *= coef;
  }
  const float step_size = a.lr / a.bias_corr1;
  const float decay = 1.f - a.lr * a.weight_decay;
  const float inv_bc2 = 1.f / a.bias_corr2_sqrt;
  const float om1 = 1.f - a.beta1, om2 = 1.f - a.beta2;
  const size_t n4 = n / 4;
  for (size_t i = blockIdx.x * (size_t)kThreads + threadIdx.x; i < n4; i += (size_t)gridDim.x * kThreads) {
    const size_t e = i * 4;
    float gv[4], mv[4], vv[4], pv[4];
    V4<G>::load(g + e, gv);
    V4<float>::load(m + e, mv);
    V4<float>::load(v + e, vv);
    V4<float>::load(p32 + e, pv);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float grad = gv[k] * gs;
      mv[k] = a.beta1 * mv[k] + om1 * grad;
      vv[k] = a.beta2 * vv[k] + om2 * grad * grad;
      pv[k] = pv[k] * decay - step_size * mv[k] / (sqrtf(vv[k]) * inv_bc2 + a.eps);
      gv[k] = 0.f;
    }
    V4<float>::store(m + e, mv);
    V4<float>::store(v + e, vv);
    V4<float>::store(p32 + e, pv);
    for (int r = 0; r < outs.n; ++r) V4<P>::store(outs.p[r] + e, pv);
    if (a.zero_grad) V4<G>::store(g + e, gv);
  }
  if (blockIdx.x == 0) {
    for (size_t i = n4 * 4 + threadIdx.x; i < n; i += kThreads) {
      const float grad = (float)g[i] * gs;
      float mi = a.beta1 * m[i] + om1 * grad, vi = a.beta2 * v[i] + om2 * grad * grad;
      float pi = p32[i] * decay - step_size * mi / (sqrtf(vi) * inv_bc2 + a.eps);
      m[i] = mi, v[i] = vi, p32[i] = pi;
      for (int r = 0; r < outs.n; ++r) outs.p[r][i] = (P)pi;
      if (a.zero_grad) g[i] = (G)0.f;
    }
  }
}

// Fused ZeRO reduce-scatter: own[i] = Σ_r peer_r[i] over this rank's shard, read straight from the data-parallel peers'
// gradient arenas over NVLink (peer-mapped pointers), summed in rank order (identical bits wherever it is computed).
template <typename G>
struct InPtrs {
  const G* p[kMaxPeers];
  int n;
};
template <typename G>
__global__ void __launch_bounds__(kThreads) peer_reduce_kernel(InPtrs<G> ins, G* __restrict__ own, size_t n) {
  const size_t n4 = n / 4;
  for (size_t i = blockIdx.x * (size_t)kThreads + threadIdx.x; i < n4; i += (size_t)gridDim.x * kThreads) {
    const size_t e = i * 4;
    float acc[4], x[4];
    V4<G>::load(ins.p[0] + e, acc);
    for (int r = 1; r < ins.n; ++r) {
      V4<G>::load(ins.p[r] + e, x);
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[k] += x[k];
    }
    V4<G>::store(own + e, acc);
  }
  if (blockIdx.x == 0) {
    for (size_t i = n4 * 4 + threadIdx.x; i < n; i += kThreads) {
      float acc = (float)ins.p[0][i];
      for (int r = 1; r < ins.n; ++r) acc += (float)ins.p[r][i];
      own[i] = (G)acc;
    }
  }
}

// Fused data-parallel all-reduce (no ZeRO): this rank owns one 1/dp chunk of the gradient arena; it reads that chunk from
// every peer's arena (NVLink loads), sums in rank order and stores the sum back into every peer's arena (NVLink stores)
// — reduce-scatter and all-gather in one pass, no intermediate buffer.  Chunks of different ranks are disjoint, so no
// rank ever writes what another rank is still reading; barriers before and after the launch bracket it.
template <typename G>
__global__ void __launch_bounds__(kThreads) peer_allreduce_kernel(OutPtrs<G> io, size_t n) {
  const size_t n4 = n / 4;
  for (size_t i = blockIdx.x * (size_t)kThreads + threadIdx.x; i < n4; i += (size_t)gridDim.x * kThreads) {
    const size_t e = i * 4;
    float acc[4], x[4];
    V4<G>::load(io.p[0] + e, acc);
    for (int r = 1; r < io.n; ++r) {
      V4<G>::load(io.p[r] + e, x);
#pragma unroll
      for (int k = 0; k < 4; ++k) acc[k] += x[k];
    }
    for (int r = 0; r < io.n; ++r) V4<G>::store(io.p[r] + e, acc);
  }
  if (blockIdx.x == 0) {
    for (size_t i = n4 * 4 + threadIdx.x; i < n; i += kThreads) {
      float acc = (float)io.p[0][i];
      for (int r = 1; r < io.n; ++r) acc += (float)io.p[r][i];
      for (int r = 0; r < io.n; ++r) io.p[r][i] = (G)acc;
    }
  }
}

// dst (P) <- src (fp32) cast, used when refreshing the model copy from master weights
template <typename P>
__global__ void cast_kernel(const float* __restrict__ src, P* __restrict__ dst, size_t n) {
  for (size_t i = blockIdx.x * (size_t)kThreads + threadIdx.x; i < n; i += (size_t)gridDim.x * kThreads) dst[i] = (P)src[i];
}

static int blocks(size_t n) {
  size_t b = (n + kThreads - 1) / kThreads;
  return (int)(b < 1 ? 1 : (b > (size_t)kMaxBlocks ? kMaxBlocks : b));
}

}  // namespace optim
}  // namespace sb

using namespace sb::optim;

extern "C" int sb_grad_sumsq_scratch_floats() { return kMaxBlocks; }

// grad dtype: 0 bf16, 1 fp32, 2 fp16.  scratch: >= sb_grad_sumsq_scratch_floats() floats.  out[0] (+)= Σ g²;
// *nonfinite |= any(!isfinite(g)).
extern "C" int sb_grad_sumsq(const void* g, long long n, int dtype, float* scratch, float* out, int accumulate,
                             int* nonfinite, cudaStream_t st) {
  if (n <= 0) {
    if (!accumulate) cudaMemsetAsync(out, 0, 4, st);
    return 0;
  }
  const int b = blocks(n);
  if (dtype == 0)
    sumsq_kernel<__nv_bfloat16><<<b, kThreads, 0, st>>>((const __nv_bfloat16*)g, n, scratch, nonfinite);
  else if (dtype == 1)
    sumsq_kernel<float><<<b, kThreads, 0, st>>>((const float*)g, n, scratch, nonfinite);
  else
    sumsq_kernel<__half><<<b, kThreads, 0, st>>>((const __half*)g, n, scratch, nonfinite);
  sumsq_final_kernel<<<1, kThreads, 0, st>>>(scratch, b, out, accumulate);
  return (int)cudaGetLastError();
}

static int adamw_impl(float* p32, float* m, float* v, void* g, void* out, const unsigned long long* peer_outs,
                      int n_peer_outs, long long n, int grad_dtype, int out_dtype, float lr, float beta1, float beta2,
                      float eps, float weight_decay, int step, float inv_loss_scale, float clip, const float* norm_sq,
                      const int* skip, int zero_grad, cudaStream_t st) {
  if (n <= 0) return 0;
  if (n_peer_outs > kMaxPeers) return -5;
  AdamArgs a;
  a.lr = lr, a.beta1 = beta1, a.beta2 = beta2, a.eps = eps, a.weight_decay = weight_decay;
  a.bias_corr1 = 1.f - powf(beta1, (float)step);
  a.bias_corr2_sqrt = sqrtf(1.f - powf(beta2, (float)step));
  a.inv_loss_scale = inv_loss_scale, a.clip = clip, a.zero_grad = zero_grad;
  // the vector path needs 16-byte aligned shards (arena shards are 128-element aligned)
  if (((uintptr_t)p32 | (uintptr_t)m | (uintptr_t)v | (uintptr_t)g | (uintptr_t)out) & 15) return -4;
  const int b = blocks((size_t)n / 4 + 1);
#define SB_ADAM(GT, PT)                                                                                   \
  do {                                                                                                    \
    OutPtrs<PT> o{};                                                                                      \
    o.n = 0;                                                                                              \
    if (n_peer_outs > 0)                                                                                  \
      for (int r = 0; r < n_peer_outs; ++r) o.p[o.n++] = (PT*)peer_outs[r];                                \
    else if (out)                                                                                         \
      o.p[o.n++] = (PT*)out;                                                                              \
    adamw_kernel<GT, PT><<<b, kThreads, 0, st>>>(p32, m, v, (GT*)g, o, (size_t)n, a, norm_sq, skip);      \
  } while (0)
  if (grad_dtype == 0 && out_dtype == 0) SB_ADAM(__nv_bfloat16, __nv_bfloat16);
  else if (grad_dtype == 1 && out_dtype == 0) SB_ADAM(float, __nv_bfloat16);
  else if (grad_dtype == 1 && out_dtype == 1) SB_ADAM(float, float);
  else if (grad_dtype == 0 && out_dtype == 1) SB_ADAM(__nv_bfloat16, float);
  else if (grad_dtype == 2 && out_dtype == 2) SB_ADAM(__half, __half);
  else if (grad_dtype == 1 && out_dtype == 2) SB_ADAM(float, __half);
  else return -2;
#undef SB_ADAM
  return (int)cudaGetLastError();
}

// grad_dtype / out_dtype: 0 bf16, 1 fp32, 2 fp16 (out may be null)
extern "C" int sb_adamw(float* p32, float* m, float* v, void* g, void* out, long long n, int grad_dtype, int out_dtype,
                        float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                        float inv_loss_scale, float clip, const float* norm_sq, const int* skip, int zero_grad,
                        cudaStream_t st) {
  return adamw_impl(p32, m, v, g, out, nullptr, 0, n, grad_dtype, out_dtype, lr, beta1, beta2, eps, weight_decay, step,
                    inv_loss_scale, clip, norm_sq, skip, zero_grad, st);
}

// Fused ZeRO-1 step, second half: AdamW on this rank's shard, and the updated model-dtype values are stored to the
// shard's offset in EVERY data-parallel rank's parameter arena (`peer_outs`: dp peer-mapped addresses, own included) —
// the parameter all-gather is the kernel's own stores over NVLink.  Reference: AdamW then all_gather
// (core/optimizer/parameter_group.py:423-481).
extern "C" int sb_adamw_zero(float* p32, float* m, float* v, void* g, const unsigned long long* peer_outs, int n_peers,
                             long long n, int grad_dtype, int out_dtype, float lr, float beta1, float beta2, float eps,
                             float weight_decay, int step, float inv_loss_scale, float clip, const float* norm_sq,
                             const int* skip, int zero_grad, cudaStream_t st) {
  if (n_peers < 1) return -5;
  return adamw_impl(p32, m, v, g, nullptr, peer_outs, n_peers, n, grad_dtype, out_dtype, lr, beta1, beta2, eps,
                    weight_decay, step, inv_loss_scale, clip, norm_sq, skip, zero_grad, st);
}

// Fused ZeRO-1 step, first half: own[i] = Σ_r peers[r][i] for this rank's gradient shard; `peer_grads` are the dp
// peer-mapped addresses of the shard inside every rank's gradient arena, in rank order (own included).  The
// reduce-scatter is the kernel's own loads over NVLink.  Reference: all-reduce then slice
// (core/optimizer/parameter_group.py:346-394, allreduce.py).
extern "C" int sb_zero_reduce(const unsigned long long* peer_grads, int n_peers, void* own, long long n, int grad_dtype,
                              cudaStream_t st) {
  if (n <= 0) return 0;
  if (n_peers < 1 || n_peers > kMaxPeers) return -5;
  const int b = blocks((size_t)n / 4 + 1);
  if (grad_dtype == 1) {
    InPtrs<float> in{};
    in.n = n_peers;
    for (int r = 0; r < n_peers; ++r) in.p[r] = (const float*)peer_grads[r];
    peer_reduce_kernel<float><<<b, kThreads, 0, st>>>(in, (float*)own, (size_t)n);
  } else if (grad_dtype == 0) {
    InPtrs<__nv_bfloat16> in{};
    in.n = n_peers;
    for (int r = 0; r < n_peers; ++r) in.p[r] = (const __nv_bfloat16*)peer_grads[r];
    peer_reduce_kernel<__nv_bfloat16><<<b, kThreads, 0, st>>>(in, (__nv_bfloat16*)own, (size_t)n);
  } else {
    return -2;
  }
  return (int)cudaGetLastError();
}

// Data-parallel gradient all-reduce without ZeRO: `peer_chunks` are the dp peer-mapped addresses of THIS
// rank's chunk inside every rank's gradient arena, in rank order (own included); every address receives the sum.
// Reference: flatten -> fp32 copy -> NCCL all_reduce -> copy back (core/optimizer/allreduce.py:16-22).
extern "C" int sb_dp_allreduce(const unsigned long long* peer_chunks, int n_peers, long long n, int grad_dtype,
                               cudaStream_t st) {
  if (n <= 0) return 0;
  if (n_peers < 1 || n_peers > kMaxPeers) return -5;
  const int b = blocks((size_t)n / 4 + 1);
  if (grad_dtype == 1) {
    OutPtrs<float> io{};
    io.n = n_peers;
    for (int r = 0; r < n_peers; ++r) io.p[r] = (float*)peer_chunks[r];
    peer_allreduce_kernel<float><<<b, kThreads, 0, st>>>(io, (size_t)n);
  } else if (grad_dtype == 0) {
    OutPtrs<__nv_bfloat16> io{};
    io.n = n_peers;
    for (int r = 0; r < n_peers; ++r) io.p[r] = (__nv_bfloat16*)peer_chunks[r];
    peer_allreduce_kernel<__nv_bfloat16><<<b, kThreads, 0, st>>>(io, (size_t)n);
  } else {
    return -2;
  }
  return (int)cudaGetLastError();
}

extern "C" int sb_cast_from_fp32(const float* src, void* dst, long long n, int out_dtype, cudaStream_t st) {
  if (n <= 0) return 0;
  const int b = blocks(n);
  if (out_dtype == 0) cast_kernel<__nv_bfloat16><<<b, kThreads, 0, st>>>(src, (__nv_bfloat16*)dst, n);
  else if (out_dtype == 2) cast_kernel<__half><<<b, kThreads, 0, st>>>(src, (__half*)dst, n);
  else return -2;
  return (int)cudaGetLastError();
}
