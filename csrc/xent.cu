// Vocab-parallel cross-entropy for sm_100a (SURVEY §8 G8, K10/K11 replacement).
//
// The reference all-gathers full-vocab logits and runs F.cross_entropy on an fp32 copy
// (tfm/model/layers/lm_head_tied.py:41-43, tfm/model/model.py:56-76).  Here each tensor-parallel rank keeps its
// [tokens, V/tp] bf16 logits shard and
//   pass 1 (stats): per token, online (max, Σexp) over the local shard, the target logit if the target id lives in
//                   this shard, and the local arg-max (for the accuracy metric);
//   [the caller combines the 4 scalars per token across the TP group — 16 bytes/token instead of 2·V bytes];
//   pass 2 (grad):  overwrites the logits shard in place with dlogits = (softmax − onehot) · coeff[token].
// Nothing of size [tokens, V] is allocated besides the GEMM output itself.
#include "common.cuh"

namespace sb {
namespace xent {

constexpr int kThreads = 256;

struct MS {
  float m, s;
};
SB_DEVICE MS combine(MS a, MS b) {
  float m = fmaxf(a.m, b.m);
  if (m == -INFINITY) return MS{m, 0.f};
  return MS{m, a.s * __expf(a.m - m) + b.s * __expf(b.m - m)};
}

template <typename T>
SB_DEVICE void load8(const T* p, float* f);
template <>
SB_DEVICE void load8<__nv_bfloat16>(const __nv_bfloat16* p, float* f) {
  uint4 v = *reinterpret_cast<const uint4*>(p);
  float2 a = unpack_bf16(v.x), b = unpack_bf16(v.y), c = unpack_bf16(v.z), d = unpack_bf16(v.w);
  f[0] = a.x, f[1] = a.y, f[2] = b.x, f[3] = b.y, f[4] = c.x, f[5] = c.y, f[6] = d.x, f[7] = d.y;
}
template <>
SB_DEVICE void load8<float>(const float* p, float* f) {
  float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
  f[0] = a.x, f[1] = a.y, f[2] = a.z, f[3] = a.w, f[4] = b.x, f[5] = b.y, f[6] = b.z, f[7] = b.w;
}

// stats[t] = {local max, local Σexp(x - max), target logit (0 if not local), local argmax (global id, as float bits)}
template <typename T>
__global__ void __launch_bounds__(kThreads) stats_kernel(const T* __restrict__ logits, const long long* __restrict__ targets,
                                                         float4* __restrict__ stats, int vocab_local, long long ld,
                                                         long long vocab_start) {
  __shared__ float sm_m[8], sm_s[8], sm_bv[8];
  __shared__ int sm_bi[8];
  const int t = blockIdx.x;
  const T* row = logits + (size_t)t * ld;
  MS acc{-INFINITY, 0.f};
  float best = -INFINITY;
  int best_i = 0;
  for (int c = threadIdx.x * 8; c < vocab_local; c += kThreads * 8) {
    float f[8];
    if (c + 8 <= vocab_local) {
      load8<T>(row + c, f);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) f[i] = (c + i < vocab_local) ? (float)row[c + i] : -INFINITY;
    }
    float m = f[0];
    int mi = 0;
#pragma unroll
    for (int i = 1; i < 8; ++i)
      if (f[i] > m) m = f[i], mi = i;
    if (m > best) best = m, best_i = c + mi;
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += __expf(f[i] - m);
    acc = combine(acc, MS{m, s});
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MS other{__shfl_xor_sync(0xffffffffu, acc.m, o), __shfl_xor_sync(0xffffffffu, acc.s, o)};
    acc = combine(acc, other);
    float ob = __shfl_xor_sync(0xffffffffu, best, o);
    int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
    if (ob > best || (ob == best && oi < best_i)) best = ob, best_i = oi;
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) sm_m[w] = acc.m, sm_s[w] = acc.s, sm_bv[w] = best, sm_bi[w] = best_i;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < kThreads / 32; ++i) {
      acc = combine(acc, MS{sm_m[i], sm_s[i]});
      if (sm_bv[i] > best || (sm_bv[i] == best && sm_bi[i] < best_i)) best = sm_bv[i], best_i = sm_bi[i];
    }
    const long long tgt = targets[t] - vocab_start;
    float tl = 0.f;
    if (tgt >= 0 && tgt < vocab_local) tl = (float)row[tgt];
    stats[t] = make_float4(acc.m, acc.s, tl, __int_as_float((int)(best_i + vocab_start)));
  }
}

// in place: logits[t, j] <- (exp(x - lse[t]) - [j == target]) * coeff[t] * (*gscale)
template <typename T>
__global__ void __launch_bounds__(kThreads) grad_kernel(T* __restrict__ logits, const long long* __restrict__ targets,
                                                        const float* __restrict__ lse, const float* __restrict__ coeff,
                                                        const float* __restrict__ gscale, int vocab_local, long long ld,
                                                        long long vocab_start) {
  const int t = blockIdx.x;
  T* row = logits + (size_t)t * ld;
  const float l = lse[t];
  const float k = coeff[t] * (gscale ? *gscale : 1.f);
  const long long tgt = targets[t] - vocab_start;
  for (int c = threadIdx.x * 8; c < vocab_local; c += kThreads * 8) {
    if (c + 8 <= vocab_local) {
      float f[8];
      load8<T>(row + c, f);
#pragma unroll
      for (int i = 0; i < 8; ++i) f[i] = (__expf(f[i] - l) - ((c + i) == tgt ? 1.f : 0.f)) * k;
      if (sizeof(T) == 2) {
        *reinterpret_cast<uint4*>(row + c) =
            make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
      } else {
        reinterpret_cast<float4*>(row + c)[0] = make_float4(f[0], f[1], f[2], f[3]);
        reinterpret_cast<float4*>(row + c)[1] = make_float4(f[4], f[5], f[6], f[7]);
      }
    } else {
      for (int i = 0; i < 8 && c + i < vocab_local; ++i)
        row[c + i] = (T)((__expf((float)row[c + i] - l) - ((c + i) == tgt ? 1.f : 0.f)) * k);
    }
  }
}

// ---- statistics produced by the LM-head GEMM epilogue (gemm2_sm100.cu, kXentStats) -------------------------------
// combine: one thread per token merges the per-tile partials of its row, picks the target logit out of the stored
// logits and writes {max, Σexp, target logit, arg-max id} to slot `rank` of EVERY tensor-parallel rank's exchange
// buffer (NVLink peer stores; own buffer when tp == 1).  finalize (after a group barrier): merges the tp slots into
// the per-token loss, log-sum-exp and arg-max.
struct StatPeers {
  float4* p[8];
  int n;
};
__global__ void combine_tiles_kernel(const float4* __restrict__ partial, int tiles, const __nv_bfloat16* __restrict__ logits,
                                     long long ld, const long long* __restrict__ targets, int tokens, int vocab_local,
                                     long long vocab_start, StatPeers outs, int rank) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= tokens) return;
  MS acc{-INFINITY, 0.f};
  float best = -INFINITY;
  int best_i = 0;
  for (int j = 0; j < tiles; ++j) {
    const float4 q = partial[(size_t)t * tiles + j];
    acc = combine(acc, MS{q.x, q.y});
    if (q.x > best) best = q.x, best_i = __float_as_int(q.w);  // a tile's max IS its arg-max value; ties keep the lower id
  }
  const long long tgt = targets[t] - vocab_start;
  const float tl = (tgt >= 0 && tgt < vocab_local) ? __bfloat162float(logits[(size_t)t * ld + tgt]) : 0.f;
  const float4 out = make_float4(acc.m, acc.s, tl, __int_as_float((int)(best_i + vocab_start)));
  for (int r = 0; r < outs.n; ++r) outs.p[r][(size_t)rank * tokens + t] = out;
}
__global__ void finalize_kernel(const float4* __restrict__ all, int tp, int tokens, float* __restrict__ loss,
                                float* __restrict__ lse, long long* __restrict__ argmax) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= tokens) return;
  MS acc{-INFINITY, 0.f};
  float best = -INFINITY, tl = 0.f;
  int best_i = 0;
  for (int r = 0; r < tp; ++r) {
    const float4 q = all[(size_t)r * tokens + t];
    acc = combine(acc, MS{q.x, q.y});
    tl += q.z;
    if (q.x > best) best = q.x, best_i = __float_as_int(q.w);
  }
  const float l = acc.m + __logf(acc.s);
  lse[t] = l;
  loss[t] = l - tl;
  argmax[t] = best_i;
}

}  // namespace xent
}  // namespace sb

// out_ptrs: tp peer-mapped addresses of a [tp, tokens] float4 exchange buffer (own included; for tp == 1 any local
// buffer).  Follow with a group barrier and sb_xent_finalize on the local buffer.
extern "C" int sb_xent_combine(const float* partial, int tiles, const void* logits, long long ld, const long long* targets,
                               int tokens, int vocab_local, long long vocab_start, const unsigned long long* out_ptrs,
                               int tp, int rank, cudaStream_t st) {
  using namespace sb::xent;
  if (tokens <= 0) return 0;
  if (tp < 1 || tp > 8 || rank < 0 || rank >= tp) return -5;
  StatPeers o{};
  o.n = tp;
  for (int r = 0; r < tp; ++r) o.p[r] = reinterpret_cast<float4*>(out_ptrs[r]);
  combine_tiles_kernel<<<(tokens + 127) / 128, 128, 0, st>>>((const float4*)partial, tiles, (const __nv_bfloat16*)logits, ld,
                                                            targets, tokens, vocab_local, vocab_start, o, rank);
  return (int)cudaGetLastError();
}
extern "C" int sb_xent_finalize(const float* all, int tp, int tokens, float* loss, float* lse, long long* argmax,
                                cudaStream_t st) {
  using namespace sb::xent;
  if (tokens <= 0) return 0;
  finalize_kernel<<<(tokens + 127) / 128, 128, 0, st>>>((const float4*)all, tp, tokens, loss, lse, argmax);
  return (int)cudaGetLastError();
}

// dtype: 0 bf16, 1 fp32.  ld = row pitch in elements (multiple of 8).
extern "C" int sb_xent_stats(const void* logits, const long long* targets, float* stats, int tokens, int vocab_local,
                             long long ld, long long vocab_start, int dtype, cudaStream_t st) {
  using namespace sb::xent;
  if (tokens <= 0) return 0;
  if (ld % 8) return -1;
  if (dtype == 0)
    stats_kernel<__nv_bfloat16><<<tokens, kThreads, 0, st>>>((const __nv_bfloat16*)logits, targets, (float4*)stats,
                                                            vocab_local, ld, vocab_start);
  else
    stats_kernel<float><<<tokens, kThreads, 0, st>>>((const float*)logits, targets, (float4*)stats, vocab_local, ld,
                                                    vocab_start);
  return (int)cudaGetLastError();
}
extern "C" int sb_xent_grad(void* logits, const long long* targets, const float* lse, const float* coeff,
                            const float* gscale, int tokens, int vocab_local, long long ld, long long vocab_start,
                            int dtype, cudaStream_t st) {
  using namespace sb::xent;
  if (tokens <= 0) return 0;
  if (ld % 8) return -1;
  if (dtype == 0)
    grad_kernel<__nv_bfloat16><<<tokens, kThreads, 0, st>>>((__nv_bfloat16*)logits, targets, lse, coeff, gscale,
                                                           vocab_local, ld, vocab_start);
  else
    grad_kernel<float><<<tokens, kThreads, 0, st>>>((float*)logits, targets, lse, coeff, gscale, vocab_local, ld,
                                                   vocab_start);
  return (int)cudaGetLastError();
}
