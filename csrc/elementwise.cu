// Fused element-wise kernels for sm_100a (SURVEY §8 G7 backward, A1 RoPE, D1, E1).
//
//  * SwiGLU forward/backward on the (gate | up) pre-activation layout produced by the dual-GEMM epilogue
//    (reference: silu(dense_in(x)) * siglu_weight(x), core/nn/mlp.py:157-160 — 2 ATen kernels fwd, ≈6 bwd).
//  * Rotary embedding, both flavours, in place on packed q/k (reference: core/nn/rotary.py:68-139, 6–10 ATen kernels);
//    the backward pass is the same kernel with the rotation direction flipped.
//  * bias + dropout + residual (reference: tfm/model/layers/layer.py:211-216) with a counter-based Philox stream, so
//    the mask is identical on every tensor-parallel rank and under activation recomputation without RNG state forking.
//  * Vocab-parallel embedding gather (masked) and its scatter-add backward into the main-grad buffer
//    (reference: core/nn/linear/vocab_parallel_embedding.py:119-147).
#include "common.cuh"

namespace sb {
namespace ew {

// ------------------------------------------------------------------ SwiGLU
// pre: [rows, 2F] (gate cols [0,F), up cols [F,2F)), h: [rows, F]
__global__ void swiglu_fwd_kernel(const __nv_bfloat16* __restrict__ pre, __nv_bfloat16* __restrict__ h, int rows, int F) {
  const size_t n8 = (size_t)rows * F / 8;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
    const size_t e = i * 8;
    const size_t r = e / F, c = e % F;
    uint4 g = *reinterpret_cast<const uint4*>(pre + r * 2 * F + c);
    uint4 u = *reinterpret_cast<const uint4*>(pre + r * 2 * F + F + c);
    const uint32_t* gp = &g.x;
    const uint32_t* up = &u.x;
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float2 a = unpack_bf16(gp[j]), b = unpack_bf16(up[j]);
      o[j] = pack_bf16(silu_f(a.x) * b.x, silu_f(a.y) * b.y);
    }
    *reinterpret_cast<uint4*>(h + e) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// dpre = [dh * u * silu'(g) | dh * silu(g)]; optionally recomputes h (needed by the down-projection wgrad when the
// forward did not keep it)
__global__ void swiglu_bwd_kernel(const __nv_bfloat16* __restrict__ pre, const __nv_bfloat16* __restrict__ dh,
                                  __nv_bfloat16* __restrict__ dpre, __nv_bfloat16* __restrict__ h_out, int rows, int F) {
  const size_t n8 = (size_t)rows * F / 8;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
    const size_t e = i * 8;
    const size_t r = e / F, c = e % F;
    uint4 g = *reinterpret_cast<const uint4*>(pre + r * 2 * F + c);
    uint4 u = *reinterpret_cast<const uint4*>(pre + r * 2 * F + F + c);
    uint4 d = *reinterpret_cast<const uint4*>(dh + e);
    const uint32_t *gp = &g.x, *up = &u.x, *dp = &d.x;
    uint32_t og[4], ou[4], oh[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float2 a = unpack_bf16(gp[j]), b = unpack_bf16(up[j]), dd = unpack_bf16(dp[j]);
      float s0 = 1.f / (1.f + __expf(-a.x)), s1 = 1.f / (1.f + __expf(-a.y));
      float act0 = a.x * s0, act1 = a.y * s1;
      float dact0 = s0 * (1.f + a.x * (1.f - s0)), dact1 = s1 * (1.f + a.y * (1.f - s1));
      og[j] = pack_bf16(dd.x * b.x * dact0, dd.y * b.y * dact1);
      ou[j] = pack_bf16(dd.x * act0, dd.y * act1);
      oh[j] = pack_bf16(act0 * b.x, act1 * b.y);
    }
    *reinterpret_cast<uint4*>(dpre + r * 2 * F + c) = make_uint4(og[0], og[1], og[2], og[3]);
    *reinterpret_cast<uint4*>(dpre + r * 2 * F + F + c) = make_uint4(ou[0], ou[1], ou[2], ou[3]);
    if (h_out) *reinterpret_cast<uint4*>(h_out + e) = make_uint4(oh[0], oh[1], oh[2], oh[3]);
  }
}

// ------------------------------------------------------------------ activation (gelu / silu) fwd+bwd, bf16
template <int kAct>
__global__ void act_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, size_t n) {
  for (size_t i = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) * 2; i < n; i += (size_t)gridDim.x * blockDim.x * 2) {
    float2 v = unpack_bf16(*reinterpret_cast<const uint32_t*>(x + i));
    float a = kAct ? silu_f(v.x) : gelu_f(v.x), b = kAct ? silu_f(v.y) : gelu_f(v.y);
    *reinterpret_cast<uint32_t*>(y + i) = pack_bf16(a, b);
  }
}
SB_DEVICE float dgelu(float x) {
  const float k = 0.70710678118654752f;
  return 0.5f * (1.f + erff(x * k)) + x * 0.3989422804014327f * __expf(-0.5f * x * x);
}
SB_DEVICE float dsilu(float x) {
  float s = 1.f / (1.f + __expf(-x));
  return s * (1.f + x * (1.f - s));
}
template <int kAct>
__global__ void act_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                               __nv_bfloat16* __restrict__ dx, size_t n) {
  for (size_t i = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) * 2; i < n; i += (size_t)gridDim.x * blockDim.x * 2) {
    float2 v = unpack_bf16(*reinterpret_cast<const uint32_t*>(x + i));
    float2 d = unpack_bf16(*reinterpret_cast<const uint32_t*>(dy + i));
    float a = kAct ? dsilu(v.x) : dgelu(v.x), b = kAct ? dsilu(v.y) : dgelu(v.y);
    *reinterpret_cast<uint32_t*>(dx + i) = pack_bf16(a * d.x, b * d.y);
  }
}

// ------------------------------------------------------------------ rotary
// x: [tokens, heads, head_dim] bf16 with token stride `ts` elements; cos/sin: fp32 [max_pos, rot_dim/2];
// pos: int32 [tokens] (may be null -> token index modulo seq).  mode 0: interleaved pairs (rotary_complex),
// mode 1: half rotation (NeoX "rotary"). dir = +1 forward, -1 backward (inverse rotation).
__global__ void rope_kernel(__nv_bfloat16* __restrict__ x, const float* __restrict__ cs, const float* __restrict__ sn,
                            const int* __restrict__ pos, int tokens, int heads, int head_dim, int rot_dim, long long ts,
                            int seq, int mode, float dir) {
  const int half = rot_dim / 2;
  const size_t total = (size_t)tokens * heads * half;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int j = i % half;
    const size_t th = i / half;
    const int hd = th % heads;
    const int t = th / heads;
    const int p = pos ? pos[t] : (t % seq);
    const float c = cs[(size_t)p * half + j], s = dir * sn[(size_t)p * half + j];
    __nv_bfloat16* base = x + (size_t)t * ts + (size_t)hd * head_dim;
    const int i0 = mode == 0 ? 2 * j : j, i1 = mode == 0 ? 2 * j + 1 : j + half;
    const float a = __bfloat162float(base[i0]), b = __bfloat162float(base[i1]);
    base[i0] = __float2bfloat16(a * c - b * s);
    base[i1] = __float2bfloat16(a * s + b * c);
  }
}

// ------------------------------------------------------------------ fused q|k|v split + rotary (one pass)
// Up to three head segments (q, k, v) are copied from their inputs to their outputs, 16 bytes per thread; segments with
// `rotate` get the rotary embedding (dir = +1) or its inverse (dir = -1) on the way.  Forward: inputs are slices of
// the fused q|k|v GEMM output, outputs one contiguous [tokens, heads, d] buffer; backward: inputs are dq, dk, dv,
// outputs slices of the fused dY buffer the dgrad / wgrad GEMMs consume — no clone, no cat, no second rope launch.
struct RopeSeg {
  const __nv_bfloat16* in;
  __nv_bfloat16* out;
  long long in_pitch, out_pitch;  // elements per token
  int heads, rotate;
};
struct RopeSegs {
  RopeSeg s[3];
  int n, total_heads;
};
SB_DEVICE void rot_pairs_interleaved(float* v, const float* c, const float* s, float dir) {
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float a = v[2 * k], b = v[2 * k + 1], sn = dir * s[k];
    v[2 * k] = a * c[k] - b * sn;
    v[2 * k + 1] = a * sn + b * c[k];
  }
}
__global__ void qkv_rope_kernel(RopeSegs segs, const float* __restrict__ cs, const float* __restrict__ sn,
                                const int* __restrict__ pos, int tokens, int head_dim, int rot_dim, int seq, int mode,
                                float dir) {
  const int chunks = head_dim / 8;  // 16-byte chunks per head
  const int half = rot_dim / 2, half_chunks = half / 8;
  const size_t total = (size_t)tokens * segs.total_heads * chunks;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % chunks);
    const size_t th = i / chunks;
    int hd = (int)(th % segs.total_heads);
    const int t = (int)(th / segs.total_heads);
    int si = 0;
    while (si < segs.n - 1 && hd >= segs.s[si].heads) hd -= segs.s[si++].heads;
    const RopeSeg& sg = segs.s[si];
    const __nv_bfloat16* src = sg.in + (size_t)t * sg.in_pitch + (size_t)hd * head_dim;
    __nv_bfloat16* dst = sg.out + (size_t)t * sg.out_pitch + (size_t)hd * head_dim;
    const int e0 = c * 8;
    if (!sg.rotate || e0 >= rot_dim) {
      *reinterpret_cast<uint4*>(dst + e0) = *reinterpret_cast<const uint4*>(src + e0);
      continue;
    }
    const int p = pos ? pos[t] : (t % seq);
    const float* cr = cs + (size_t)p * half;
    const float* sr = sn + (size_t)p * half;
    if (mode == 0) {  // interleaved pairs: elements (2j, 2j+1), frequencies e0/2 .. e0/2+3
      const uint4 x = *reinterpret_cast<const uint4*>(src + e0);
      float v[8];
      float2 f;
      f = unpack_bf16(x.x), v[0] = f.x, v[1] = f.y;
      f = unpack_bf16(x.y), v[2] = f.x, v[3] = f.y;
      f = unpack_bf16(x.z), v[4] = f.x, v[5] = f.y;
      f = unpack_bf16(x.w), v[6] = f.x, v[7] = f.y;
      const float4 c4 = *reinterpret_cast<const float4*>(cr + e0 / 2), s4 = *reinterpret_cast<const float4*>(sr + e0 / 2);
      const float cc[4] = {c4.x, c4.y, c4.z, c4.w}, ss[4] = {s4.x, s4.y, s4.z, s4.w};
      rot_pairs_interleaved(v, cc, ss, dir);
      *reinterpret_cast<uint4*>(dst + e0) =
          make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
    } else {  // half rotation: element j pairs with j + half; the thread of the first half handles both chunks
      if (c >= half_chunks) continue;
      const uint4 xa = *reinterpret_cast<const uint4*>(src + e0), xb = *reinterpret_cast<const uint4*>(src + e0 + half);
      const uint32_t wa[4] = {xa.x, xa.y, xa.z, xa.w}, wb[4] = {xb.x, xb.y, xb.z, xb.w};
      uint32_t oa[4], ob[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 a = unpack_bf16(wa[k]), b = unpack_bf16(wb[k]);
        const float c0 = cr[e0 + 2 * k], c1 = cr[e0 + 2 * k + 1];
        const float s0 = dir * sr[e0 + 2 * k], s1 = dir * sr[e0 + 2 * k + 1];
        oa[k] = pack_bf16(a.x * c0 - b.x * s0, a.y * c1 - b.y * s1);
        ob[k] = pack_bf16(a.x * s0 + b.x * c0, a.y * s1 + b.y * c1);
      }
      *reinterpret_cast<uint4*>(dst + e0) = make_uint4(oa[0], oa[1], oa[2], oa[3]);
      *reinterpret_cast<uint4*>(dst + e0 + half) = make_uint4(ob[0], ob[1], ob[2], ob[3]);
    }
  }
}

// ------------------------------------------------------------------ bias + dropout + residual
// out = residual + dropout(x + bias); mask derived from Philox(seed, element index) — nothing is stored
__global__ void bias_dropout_add_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ bias,
                                        const __nv_bfloat16* __restrict__ residual, __nv_bfloat16* __restrict__ out,
                                        size_t n, int cols, float p, unsigned long long seed,
                                        unsigned long long offset) {
  const float scale = p > 0.f ? 1.f / (1.f - p) : 1.f;
  Philox rng(seed);
  for (size_t i = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) * 4; i < n; i += (size_t)gridDim.x * blockDim.x * 4) {
    uint2 xv = *reinterpret_cast<const uint2*>(x + i);
    float2 a = unpack_bf16(xv.x), b = unpack_bf16(xv.y);
    float v[4] = {a.x, a.y, b.x, b.y};
    if (bias) {
      const int c = i % cols;
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] += __bfloat162float(bias[c + k]);
    }
    if (p > 0.f) {
      uint4 r = rng(i / 4, offset);
      const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] = (rr[k] * 2.3283064365386963e-10f >= p) ? v[k] * scale : 0.f;
    }
    if (residual) {
      uint2 rv = *reinterpret_cast<const uint2*>(residual + i);
      float2 c0 = unpack_bf16(rv.x), c1 = unpack_bf16(rv.y);
      v[0] += c0.x, v[1] += c0.y, v[2] += c1.x, v[3] += c1.y;
    }
    *reinterpret_cast<uint2*>(out + i) = make_uint2(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]));
  }
}
// dx = dout * mask * scale  (the residual branch gradient is dout itself)
__global__ void dropout_bwd_kernel(const __nv_bfloat16* __restrict__ dout, __nv_bfloat16* __restrict__ dx, size_t n,
                                   float p, unsigned long long seed, unsigned long long offset) {
  const float scale = 1.f / (1.f - p);
  Philox rng(seed);
  for (size_t i = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) * 4; i < n; i += (size_t)gridDim.x * blockDim.x * 4) {
    uint2 dv = *reinterpret_cast<const uint2*>(dout + i);
    float2 a = unpack_bf16(dv.x), b = unpack_bf16(dv.y);
    float v[4] = {a.x, a.y, b.x, b.y};
    uint4 r = rng(i / 4, offset);
    const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = (rr[k] * 2.3283064365386963e-10f >= p) ? v[k] * scale : 0.f;
    *reinterpret_cast<uint2*>(dx + i) = make_uint2(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]));
  }
}

// ------------------------------------------------------------------ vocab-parallel embedding
// out[t] = (vocab_start <= id < vocab_end) ? W[id - vocab_start] : 0
__global__ void embedding_fwd_kernel(const long long* __restrict__ ids, const __nv_bfloat16* __restrict__ W,
                                     __nv_bfloat16* __restrict__ out, int tokens, int hidden, long long vocab_start,
                                     long long vocab_end) {
  const int per_row = hidden / 8;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < (size_t)tokens * per_row;
       i += (size_t)gridDim.x * blockDim.x) {
    const int t = i / per_row, c = (i % per_row) * 8;
    const long long id = ids[t];
    uint4 v = make_uint4(0, 0, 0, 0);
    if (id >= vocab_start && id < vocab_end) v = *reinterpret_cast<const uint4*>(W + (size_t)(id - vocab_start) * hidden + c);
    *reinterpret_cast<uint4*>(out + (size_t)t * hidden + c) = v;
  }
}
// Vocab-parallel embedding forward with the tensor-parallel "sum" done by placement: every token's row
// lives on exactly one rank, so that owner stores the row straight into the output buffer of every TP rank (NVLink peer
// stores into symmetric memory) and nobody adds anything — no dense all-reduce of [b, s, h].  With seq_per_rank > 0
// (sequence parallelism) the row goes only to the rank that owns the token's sequence position, at its compact row
// b * seq_per_rank + s % seq_per_rank: the reduce-scatter along the sequence by placement.
struct EmbPeers {
  __nv_bfloat16* p[8];
  int n;
};
__global__ void embedding_fwd_peers_kernel(const long long* __restrict__ ids, const __nv_bfloat16* __restrict__ W,
                                           EmbPeers outs, int tokens, int hidden, long long vocab_start,
                                           long long vocab_end, int seq, int seq_per_rank) {
  const int per_row = hidden / 8;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < (size_t)tokens * per_row;
       i += (size_t)gridDim.x * blockDim.x) {
    const int t = i / per_row, c = (i % per_row) * 8;
    const long long id = ids[t];
    if (id < vocab_start || id >= vocab_end) continue;  // another rank owns this row and writes it
    const uint4 v = *reinterpret_cast<const uint4*>(W + (size_t)(id - vocab_start) * hidden + c);
    if (seq_per_rank > 0) {
      const int b = t / seq, pos = t % seq;
      const size_t row = (size_t)b * seq_per_rank + pos % seq_per_rank;
      *reinterpret_cast<uint4*>(outs.p[pos / seq_per_rank] + row * hidden + c) = v;
    } else {
      for (int r = 0; r < outs.n; ++r) *reinterpret_cast<uint4*>(outs.p[r] + (size_t)t * hidden + c) = v;
    }
  }
}
template <typename G>
__global__ void embedding_bwd_kernel(const long long* __restrict__ ids, const __nv_bfloat16* __restrict__ dout,
                                     G* __restrict__ dW, int tokens, int hidden, long long vocab_start,
                                     long long vocab_end) {
  const int per_row = hidden / 2;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < (size_t)tokens * per_row;
       i += (size_t)gridDim.x * blockDim.x) {
    const int t = i / per_row, c = (i % per_row) * 2;
    const long long id = ids[t];
    if (id < vocab_start || id >= vocab_end) continue;
    float2 d = unpack_bf16(*reinterpret_cast<const uint32_t*>(dout + (size_t)t * hidden + c));
    G* dst = dW + (size_t)(id - vocab_start) * hidden + c;
    if constexpr (sizeof(G) == 4) {
      atomicAdd(reinterpret_cast<float*>(dst), d.x);
      atomicAdd(reinterpret_cast<float*>(dst) + 1, d.y);
    } else {
      atomicAdd(reinterpret_cast<__nv_bfloat162*>(dst), __floats2bfloat162_rn(d.x, d.y));
    }
  }
}

// Deterministic variant: `ids` are sorted (stable, so equal ids keep their token order) and `perm[i]` is the token the
// i-th sorted id came from.  The block that sits on the first position of a run of equal ids sums the run's rows in that
// fixed order and adds the total to the row of dW — no atomics, the same bits on every run (the reference's path,
// ATen's sort + segmented embedding backward, is deterministic as well, and its resume test compares losses to 1e-10).
template <typename G>
__global__ void embedding_bwd_sorted_kernel(const long long* __restrict__ ids, const long long* __restrict__ perm,
                                            const __nv_bfloat16* __restrict__ dout, G* __restrict__ dW, int tokens,
                                            int hidden, long long vocab_start, long long vocab_end) {
  for (int i = blockIdx.x; i < tokens; i += gridDim.x) {
    const long long id = ids[i];
    if (i > 0 && ids[i - 1] == id) continue;  // not the head of its run
    if (id < vocab_start || id >= vocab_end) continue;
    G* dst = dW + (size_t)(id - vocab_start) * hidden;
    for (int c = threadIdx.x * 2; c < hidden; c += blockDim.x * 2) {
      float ax = 0.f, ay = 0.f;
      for (int j = i; j < tokens && ids[j] == id; ++j) {
        const float2 d = unpack_bf16(*reinterpret_cast<const uint32_t*>(dout + (size_t)perm[j] * hidden + c));
        ax += d.x, ay += d.y;
      }
      if constexpr (sizeof(G) == 4) {
        float2* o = reinterpret_cast<float2*>(dst + c);
        float2 v = *o;
        v.x += ax, v.y += ay;
        *o = v;
      } else {
        uint32_t* o = reinterpret_cast<uint32_t*>(dst + c);
        const float2 v = unpack_bf16(*o);
        *o = pack_bf16(v.x + ax, v.y + ay);
      }
    }
  }
}

static int blocks_for(size_t work, int threads = 256) {
  size_t b = (work + threads - 1) / threads;
  size_t cap = 148 * 16;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace ew
}  // namespace sb

using namespace sb::ew;

extern "C" int sb_swiglu_fwd(const void* pre, void* h, int rows, int F, cudaStream_t st) {
  if (rows <= 0) return 0;
  if (F % 8) return -1;
  swiglu_fwd_kernel<<<blocks_for((size_t)rows * F / 8), 256, 0, st>>>((const __nv_bfloat16*)pre, (__nv_bfloat16*)h, rows, F);
  return (int)cudaGetLastError();
}
extern "C" int sb_swiglu_bwd(const void* pre, const void* dh, void* dpre, void* h_out, int rows, int F,
                             cudaStream_t st) {
  if (rows <= 0) return 0;
  if (F % 8) return -1;
  swiglu_bwd_kernel<<<blocks_for((size_t)rows * F / 8), 256, 0, st>>>(
      (const __nv_bfloat16*)pre, (const __nv_bfloat16*)dh, (__nv_bfloat16*)dpre, (__nv_bfloat16*)h_out, rows, F);
  return (int)cudaGetLastError();
}
// act: 0 gelu, 1 silu
extern "C" int sb_act_fwd(const void* x, void* y, long long n, int act, cudaStream_t st) {
  if (n <= 0) return 0;
  if (n % 2) return -1;
  if (act)
    act_fwd_kernel<1><<<blocks_for(n / 2), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, n);
  else
    act_fwd_kernel<0><<<blocks_for(n / 2), 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, n);
  return (int)cudaGetLastError();
}
extern "C" int sb_act_bwd(const void* x, const void* dy, void* dx, long long n, int act, cudaStream_t st) {
  if (n <= 0) return 0;
  if (n % 2) return -1;
  if (act)
    act_bwd_kernel<1><<<blocks_for(n / 2), 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, n);
  else
    act_bwd_kernel<0><<<blocks_for(n / 2), 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, n);
  return (int)cudaGetLastError();
}
extern "C" int sb_rope(void* x, const float* cos_t, const float* sin_t, const int* pos, int tokens, int heads,
                       int head_dim, int rot_dim, long long token_stride, int seq, int mode, int backward,
                       cudaStream_t st) {
  if (tokens <= 0 || rot_dim <= 0) return 0;
  rope_kernel<<<blocks_for((size_t)tokens * heads * (rot_dim / 2)), 256, 0, st>>>(
      (__nv_bfloat16*)x, cos_t, sin_t, pos, tokens, heads, head_dim, rot_dim, token_stride, seq, mode,
      backward ? -1.f : 1.f);
  return (int)cudaGetLastError();
}
// Fused split + rotary over up to three head segments.  ins/outs/in_pitch/out_pitch/heads/rotate: host arrays of
// `n_seg` entries (pointers as integers); every pointer 16-byte aligned, pitches multiples of 8 elements, head_dim and
// rot_dim/2 multiples of 8.  Reference: RotaryEmbedding(Complex).forward after the q/k/v projections
// (core/nn/attention/attention.py:560-600, rotary.py:68-203).
extern "C" int sb_qkv_rope(const unsigned long long* ins, const unsigned long long* outs, const long long* in_pitch,
                           const long long* out_pitch, const int* heads, const int* rotate, int n_seg,
                           const float* cos_t, const float* sin_t, const int* pos, int tokens, int head_dim, int rot_dim,
                           int seq, int mode, int backward, cudaStream_t st) {
  if (tokens <= 0) return 0;
  if (n_seg < 1 || n_seg > 3 || head_dim % 8 || (rot_dim / 2) % 8 || rot_dim > head_dim) return -1;
  sb::ew::RopeSegs segs{};
  segs.n = n_seg;
  for (int i = 0; i < n_seg; ++i) {
    if ((ins[i] | outs[i]) & 15 || in_pitch[i] % 8 || out_pitch[i] % 8) return -4;
    segs.s[i].in = (const __nv_bfloat16*)ins[i];
    segs.s[i].out = (__nv_bfloat16*)outs[i];
    segs.s[i].in_pitch = in_pitch[i], segs.s[i].out_pitch = out_pitch[i];
    segs.s[i].heads = heads[i], segs.s[i].rotate = rotate[i];
    segs.total_heads += heads[i];
  }
  const size_t work = (size_t)tokens * segs.total_heads * (head_dim / 8);
  qkv_rope_kernel<<<blocks_for(work), 256, 0, st>>>(segs, cos_t, sin_t, pos, tokens, head_dim, rot_dim, seq, mode,
                                                    backward ? -1.f : 1.f);
  return (int)cudaGetLastError();
}
extern "C" int sb_bias_dropout_add(const void* x, const void* bias, const void* residual, void* out, long long n,
                                   int cols, float p, unsigned long long seed, unsigned long long offset,
                                   cudaStream_t st) {
  if (n <= 0) return 0;
  if (n % 4 || cols % 4) return -1;
  bias_dropout_add_kernel<<<blocks_for(n / 4), 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)bias,
                                                            (const __nv_bfloat16*)residual, (__nv_bfloat16*)out, n, cols,
                                                            p, seed, offset);
  return (int)cudaGetLastError();
}
extern "C" int sb_dropout_bwd(const void* dout, void* dx, long long n, float p, unsigned long long seed,
                              unsigned long long offset, cudaStream_t st) {
  if (n <= 0) return 0;
  if (n % 4) return -1;
  dropout_bwd_kernel<<<blocks_for(n / 4), 256, 0, st>>>((const __nv_bfloat16*)dout, (__nv_bfloat16*)dx, n, p, seed, offset);
  return (int)cudaGetLastError();
}
extern "C" int sb_embedding_fwd(const long long* ids, const void* W, void* out, int tokens, int hidden,
                                long long vocab_start, long long vocab_end, cudaStream_t st) {
  if (tokens <= 0) return 0;
  if (hidden % 8) return -1;
  embedding_fwd_kernel<<<blocks_for((size_t)tokens * hidden / 8), 256, 0, st>>>(ids, (const __nv_bfloat16*)W,
                                                                              (__nv_bfloat16*)out, tokens, hidden,
                                                                              vocab_start, vocab_end);
  return (int)cudaGetLastError();
}
// out_ptrs: the tp peer-mapped addresses of the (symmetric) output buffer, rank order, own included.  seq == 0: every
// rank receives all `tokens` rows; seq > 0: tokens are [batch, seq] and rank r receives positions
// [r * seq / tp, (r + 1) * seq / tp) of every batch item ([batch * seq / tp, hidden] compact).  The caller brackets the
// launch with a group barrier (rows arrive from the peers' launches).  Reference: masked F.embedding then all_reduce /
// reduce_scatter of the dense [b, s, h] tensor (core/nn/linear/vocab_parallel_embedding.py:122-140).
extern "C" int sb_embedding_fwd_peers(const long long* ids, const void* W, const unsigned long long* out_ptrs, int tp,
                                      int tokens, int hidden, long long vocab_start, long long vocab_end, int seq,
                                      cudaStream_t st) {
  if (tokens <= 0) return 0;
  if (hidden % 8 || tp < 1 || tp > 8) return -1;
  if (seq > 0 && (tokens % seq || seq % tp)) return -3;
  EmbPeers o{};
  o.n = tp;
  for (int r = 0; r < tp; ++r) o.p[r] = reinterpret_cast<__nv_bfloat16*>(out_ptrs[r]);
  embedding_fwd_peers_kernel<<<blocks_for((size_t)tokens * hidden / 8), 256, 0, st>>>(
      ids, (const __nv_bfloat16*)W, o, tokens, hidden, vocab_start, vocab_end, seq, seq > 0 ? seq / tp : 0);
  return (int)cudaGetLastError();
}
// ---------------------------------------------------------------- sequence-parallel gather / reduce-scatter by peer stores
// Megatron sequence parallelism moves [b, s/tp, h] <-> [b, s, h] around every column-/row-parallel block.  These kernels
// do it with NVLink peer stores into symmetric exchange buffers instead of NCCL (reference
// core/nn/linear/utils.py:279-305): the caller follows each launch with one device barrier.
//   gather_push : every rank stores its [b, s_local, h] shard at sequence offset rank * s_local of EVERY rank's
//                 [b, s, h] buffer;
//   scatter_push: every rank stores slice j of its [b, s, h] partial sums into slot `rank` of rank j's
//                 [tp, b, s_local, h] buffer; sum_slots then adds the tp slots in rank order (fp32 accumulate).
__global__ void seq_gather_push_kernel(const uint4* __restrict__ src, EmbPeers outs, int rank, int s_local, int vec_per_row,
                                       size_t total_vec) {
  const int tp = outs.n;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total_vec; i += (size_t)gridDim.x * blockDim.x) {
    const size_t row = i / vec_per_row, c = i % vec_per_row;
    const size_t b = row / s_local, sl = row % s_local;
    const size_t drow = b * ((size_t)s_local * tp) + (size_t)rank * s_local + sl;
    const uint4 v = src[i];
    for (int r = 0; r < tp; ++r) reinterpret_cast<uint4*>(outs.p[r])[drow * vec_per_row + c] = v;
  }
}
__global__ void seq_scatter_push_kernel(const uint4* __restrict__ src, EmbPeers outs, int rank, int s_local, int vec_per_row,
                                        size_t total_vec, size_t slot_vec) {
  const int tp = outs.n;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total_vec; i += (size_t)gridDim.x * blockDim.x) {
    const size_t row = i / vec_per_row, c = i % vec_per_row;
    const size_t s = (size_t)s_local * tp;
    const size_t b = row / s, pos = row % s;
    const int owner = (int)(pos / s_local);
    const size_t drow = b * s_local + pos % s_local;
    reinterpret_cast<uint4*>(outs.p[owner])[(size_t)rank * slot_vec + drow * vec_per_row + c] = src[i];
  }
}
__global__ void sum_slots_kernel(const uint4* __restrict__ slots, uint4* __restrict__ out, int tp, size_t slot_vec) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < slot_vec; i += (size_t)gridDim.x * blockDim.x) {
    float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int r = 0; r < tp; ++r) {
      const uint4 x = slots[(size_t)r * slot_vec + i];
      float2 a = sb::unpack_bf16(x.x), b = sb::unpack_bf16(x.y), c = sb::unpack_bf16(x.z), d = sb::unpack_bf16(x.w);
      f[0] += a.x, f[1] += a.y, f[2] += b.x, f[3] += b.y, f[4] += c.x, f[5] += c.y, f[6] += d.x, f[7] += d.y;
    }
    out[i] = make_uint4(sb::pack_bf16(f[0], f[1]), sb::pack_bf16(f[2], f[3]), sb::pack_bf16(f[4], f[5]), sb::pack_bf16(f[6], f[7]));
  }
}
// gather_push with arrival counters (feeds the gated GEMM): one CTA copies one chunk of kGateChunkRows rows
// of the local shard to every rank, then adds 1 (release, system scope) to the counter of the 256-row tile the chunk
// belongs to on every rank.  A complete tile therefore shows (256 / kGateChunkRows) * calls on that buffer.
constexpr int kGateChunkRows = 8;
struct GatePeers {
  unsigned int* p[8];
};
__global__ void seq_gather_push_counted_kernel(const uint4* __restrict__ src, EmbPeers outs, GatePeers gates, int rank,
                                               int s_local, int vec_per_row, int n_chunks) {
  const int tp = outs.n;
  for (int chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
    const size_t row_first = (size_t)chunk * kGateChunkRows;
    const size_t b = row_first / s_local, sl0 = row_first % s_local;
    const size_t drow0 = b * ((size_t)s_local * tp) + (size_t)rank * s_local + sl0;  // chunks never straddle a batch item
    const size_t n_vec = (size_t)kGateChunkRows * vec_per_row;
    for (size_t i = threadIdx.x; i < n_vec; i += blockDim.x) {
      const uint4 v = src[row_first * vec_per_row + i];
      for (int r = 0; r < tp; ++r) reinterpret_cast<uint4*>(outs.p[r])[drow0 * vec_per_row + i] = v;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      const unsigned int tile = (unsigned int)(drow0 / 256);
      for (int r = 0; r < tp; ++r)
        asm volatile("red.release.sys.global.add.u32 [%0], 1;\n" ::"l"(gates.p[r] + tile) : "memory");
    }
    __syncthreads();
  }
}
static int fill_peers(EmbPeers& o, const unsigned long long* ptrs, int tp) {
  if (tp < 1 || tp > 8) return -1;
  o.n = tp;
  for (int r = 0; r < tp; ++r) o.p[r] = reinterpret_cast<__nv_bfloat16*>(ptrs[r]);
  return 0;
}
// src: [batch, s_local, hidden] bf16 contiguous; out_ptrs: tp peer-mapped [batch, s_local * tp, hidden] buffers
extern "C" int sb_seq_gather_push(const void* src, const unsigned long long* out_ptrs, int tp, int rank, int batch,
                                  int s_local, int hidden, cudaStream_t st) {
  if (batch <= 0 || s_local <= 0) return 0;
  if (hidden % 8) return -1;
  EmbPeers o{};
  if (fill_peers(o, out_ptrs, tp)) return -5;
  const size_t total = (size_t)batch * s_local * (hidden / 8);
  seq_gather_push_kernel<<<blocks_for(total), 256, 0, st>>>((const uint4*)src, o, rank, s_local, hidden / 8, total);
  return (int)cudaGetLastError();
}
// As sb_seq_gather_push, plus arrival counters: gate_ptrs are the tp peer-mapped u32 counter arrays (one counter per
// 256-row tile of the [batch * s_local * tp, hidden] gather buffer).  Requires s_local % 256 == 0.  After n calls on a
// buffer every counter of a complete tile reads n * sb_seq_gather_chunks_per_tile().
extern "C" int sb_seq_gather_chunks_per_tile() { return 256 / kGateChunkRows; }
extern "C" int sb_seq_gather_push_counted(const void* src, const unsigned long long* out_ptrs,
                                          const unsigned long long* gate_ptrs, int tp, int rank, int batch, int s_local,
                                          int hidden, cudaStream_t st) {
  if (batch <= 0 || s_local <= 0) return 0;
  if (hidden % 8 || s_local % 256) return -1;
  EmbPeers o{};
  if (fill_peers(o, out_ptrs, tp)) return -5;
  GatePeers g{};
  for (int r = 0; r < tp; ++r) g.p[r] = reinterpret_cast<unsigned int*>(gate_ptrs[r]);
  const int n_chunks = batch * s_local / kGateChunkRows;
  const int grid = n_chunks < 148 * 8 ? n_chunks : 148 * 8;
  seq_gather_push_counted_kernel<<<grid, 256, 0, st>>>((const uint4*)src, o, g, rank, s_local, hidden / 8, n_chunks);
  return (int)cudaGetLastError();
}
// src: [batch, s_local * tp, hidden] bf16 partial sums; out_ptrs: tp peer-mapped [tp, batch, s_local, hidden] slot buffers
extern "C" int sb_seq_scatter_push(const void* src, const unsigned long long* out_ptrs, int tp, int rank, int batch,
                                   int s_local, int hidden, cudaStream_t st) {
  if (batch <= 0 || s_local <= 0) return 0;
  if (hidden % 8) return -1;
  EmbPeers o{};
  if (fill_peers(o, out_ptrs, tp)) return -5;
  const size_t slot = (size_t)batch * s_local * (hidden / 8);
  seq_scatter_push_kernel<<<blocks_for(slot * tp), 256, 0, st>>>((const uint4*)src, o, rank, s_local, hidden / 8, slot * tp, slot);
  return (int)cudaGetLastError();
}
// out[n] = Σ_r slots[r][n] (bf16, fp32 accumulate, rank order); n % 8 == 0
extern "C" int sb_sum_slots(const void* slots, void* out, int tp, long long n, cudaStream_t st) {
  if (n <= 0) return 0;
  if (n % 8) return -1;
  sum_slots_kernel<<<blocks_for((size_t)n / 8), 256, 0, st>>>((const uint4*)slots, (uint4*)out, tp, (size_t)n / 8);
  return (int)cudaGetLastError();
}
extern "C" int sb_embedding_bwd(const long long* ids, const void* dout, void* dW, int grad_fp32, int tokens, int hidden,
                                long long vocab_start, long long vocab_end, cudaStream_t st) {
  if (tokens <= 0) return 0;
  if (hidden % 2) return -1;
  const int b = blocks_for((size_t)tokens * hidden / 2);
  if (grad_fp32)
    embedding_bwd_kernel<float><<<b, 256, 0, st>>>(ids, (const __nv_bfloat16*)dout, (float*)dW, tokens, hidden, vocab_start, vocab_end);
  else
    embedding_bwd_kernel<__nv_bfloat16><<<b, 256, 0, st>>>(ids, (const __nv_bfloat16*)dout, (__nv_bfloat16*)dW, tokens, hidden, vocab_start, vocab_end);
  return (int)cudaGetLastError();
}
extern "C" int sb_embedding_bwd_sorted(const long long* sorted_ids, const long long* perm, const void* dout, void* dW,
                                       int grad_fp32, int tokens, int hidden, long long vocab_start, long long vocab_end,
                                       cudaStream_t st) {
  if (tokens <= 0) return 0;
  if (hidden % 2) return -1;
  const int b = tokens < 148 * 16 ? tokens : 148 * 16;
  if (grad_fp32)
    embedding_bwd_sorted_kernel<float><<<b, 128, 0, st>>>(sorted_ids, perm, (const __nv_bfloat16*)dout, (float*)dW, tokens,
                                                          hidden, vocab_start, vocab_end);
  else
    embedding_bwd_sorted_kernel<__nv_bfloat16><<<b, 128, 0, st>>>(sorted_ids, perm, (const __nv_bfloat16*)dout,
                                                                  (__nv_bfloat16*)dW, tokens, hidden, vocab_start, vocab_end);
  return (int)cudaGetLastError();
}
