// Flash attention forward for sm_100a: tcgen05 QKᵀ and PV with TMEM accumulators, TMA-fed K/V rings,
// online softmax in fp32, two 128-row query tiles per CTA in ping-pong.
//
// Replaces the flash-attn 2 wheel the reference calls (core/nn/attention/attention.py:204-259; FA2 is an sm80-style
// mma.sync kernel with no tcgen05 path).  Covers what the reference uses it for: causal or bidirectional attention
// over packed (varlen) token-major tensors, GQA/MQA, head_dim 128, bf16.
//
//   grid = (ceil(T / 256), heads).  One CTA owns 256 query rows of one head as two 128-row tiles a and b that share
//   every K/V tile (halves the L2→SM traffic per FLOP) and keeps the tensor pipe busy with one tile's MMAs while the
//   other tile's softmax runs.  Rows carry seg_start / seg_end, so packed documents are just masks.
//
//   warp 0        TMA producer: Qa, Qb once; K through a 3-stage ring, V through a 2-stage ring (K is needed one
//                 iteration earlier than V because S(j+1) is issued before PV(j) retires)
//   warp 1        MMA issuer, per KV tile j and query tile x:  PVx(j): Ox += Px(j)·V(j)  (TS: P read from TMEM, V
//                 MN-major in smem), then Sx(j+1) = Qx·K(j+1)ᵀ (SS) — so x's next scores are ready when its softmax
//                 warps come back, while the other tile's softmax is still running
//   warp 2        TMEM allocator (512 columns: Sa Sb | Oa Ob; P aliases the first 64 columns of its S)
//   warps 4..7    softmax of tile a, warps 8..11 softmax of tile b (one query row per thread): two passes over the S
//                 row in TMEM (max, then exp2 + sum + bf16 pack), lazy rescale of O in TMEM only when the row max grew
//                 by more than 2^8, final O/l and LSE
#include <cstdlib>

#include "common.cuh"

namespace sb {
namespace attn {

constexpr int BQ = 128;   // query rows per sub-tile (two per CTA)
constexpr int BKV = 128;  // keys per tile
constexpr int D = 128;    // head dim
constexpr int kThreads = 384;
constexpr uint32_t kTileBytes = 128 * 128 * 2;  // 32 KiB: a [128, 128] bf16 tile as two 64-column swizzle slabs
constexpr int kKStages = 3;
constexpr int kVStages = 2;
constexpr uint32_t kSmemBytes = kTileBytes * (2 + kKStages + kVStages) + 1024 + 256;

// TMEM columns
constexpr uint32_t kColS = 0;    // Sa at 0, Sb at 128; Px aliases Sx[0, 64)
constexpr uint32_t kColO = 256;  // Oa at 256, Ob at 384

// Attention-probability dropout (reference core/nn/attention/attention.py:245-258, dropout_p of flash_attn_varlen_func):
// a stateless per-element hash of (seed, global head, query, key) decides keep / drop, so forward, the dK/dV blocks and
// the dQ blocks regenerate the SAME mask without storing it, the mask is identical on every replay (activation
// checkpointing) and differs between the heads of different tensor-parallel ranks (head_offset).
struct Drop {
  uint32_t thr;      // drop when hash < thr (thr = p * 2^32); 0 = dropout off
  uint32_t seed;     // per-call seed (topology dropout seed mixed with the call counter)
  int head_offset;   // global index of this rank's first head
  float inv_keep;    // 1 / (1 - p)
};
SB_DEVICE bool drop_keep(const Drop& d, int head, int q, int k) {
  uint32_t x = (uint32_t)q * 0x9E3779B1u + (uint32_t)k * 0x85EBCA77u + (uint32_t)(head + d.head_offset) * 0xC2B2AE3Du + d.seed;
  x ^= x >> 16;
  x *= 0x7FEB352Du;
  x ^= x >> 15;
  x *= 0x846CA68Bu;
  x ^= x >> 16;
  return x >= d.thr;
}

static inline Drop make_drop(float p, unsigned long long seed, int head_offset) {
  Drop d;
  d.thr = p > 0.f ? (uint32_t)((double)p * 4294967296.0) : 0u;
  if (p > 0.f && d.thr == 0) d.thr = 1;
  d.seed = (uint32_t)(seed ^ (seed >> 32)) * 0x9E3779B1u + 0x7F4A7C15u;
  d.head_offset = head_offset;
  d.inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;
  return d;
}

struct Params {
  int T, heads, kv_heads;
  float scale_log2;   // softmax scale * log2(e)
  int causal;
  Drop drop;
  const int* seg_start;  // [T] first token of the row's sequence
  const int* seg_end;    // [T] one past the last token of the row's sequence
  __nv_bfloat16* out;    // [T, heads, 128]
  float* lse;            // [heads, T]
};

struct __align__(8) Barriers {
  uint64_t q_full;
  uint64_t k_full[kKStages];
  uint64_t k_empty[kKStages];
  uint64_t v_full[kVStages];
  uint64_t v_empty[kVStages];
  uint64_t s_full[2];
  uint64_t p_full[2];
  uint64_t o_done[2];
  uint32_t tmem_base;
};

SB_DEVICE void load_tile(uint8_t* dst, const CUtensorMap* map, uint64_t* bar, int col0, int row0) {
  tma_load_2d(dst, map, bar, col0, row0);
  tma_load_2d(dst + kTileBytes / 2, map, bar, col0 + 64, row0);
}


// Softmax inner loops.  The masked variants run only on boundary tiles (diagonal / document edges); keeping the branch
// outside the unrolled element loops avoids per-element divergence bookkeeping in the common unmasked case.
template <bool kMask>
SB_DEVICE float fwd_row_max(uint32_t s_tmem, int kv0, int rs, int re) {
  float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll 1
  for (int c = 0; c < BKV; c += 32) {
    uint32_t v[32];
    tmem_ld_32x32b_x32(s_tmem + c, v);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      float a = __uint_as_float(v[i]), b = __uint_as_float(v[i + 1]), c2 = __uint_as_float(v[i + 2]), d = __uint_as_float(v[i + 3]);
      if (kMask) {
        const int key = kv0 + c + i;
        if (!(key >= rs && key < re)) a = -INFINITY;
        if (!(key + 1 >= rs && key + 1 < re)) b = -INFINITY;
        if (!(key + 2 >= rs && key + 2 < re)) c2 = -INFINITY;
        if (!(key + 3 >= rs && key + 3 < re)) d = -INFINITY;
      }
      m0 = fmaxf(m0, a), m1 = fmaxf(m1, b), m2 = fmaxf(m2, c2), m3 = fmaxf(m3, d);
    }
  }
  return fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
}

template <bool kMask, bool kDrop, bool kHalf>
SB_DEVICE void fwd_exp_pack(uint32_t s_tmem, int kv0, int rs, int re, float scale_log2, float m_use, float& l0, float& l1,
                            const Drop& drop, int head, int row) {
#pragma unroll 1
  for (int c = 0; c < BKV; c += 32) {
    uint32_t v[32];
    tmem_ld_32x32b_x32(s_tmem + c, v);
    tmem_ld_wait();
    uint32_t packed[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float a = exp2f(__uint_as_float(v[2 * i]) * scale_log2 - m_use);
      float b = exp2f(__uint_as_float(v[2 * i + 1]) * scale_log2 - m_use);
      if (kMask) {
        const int key = kv0 + c + 2 * i;
        if (!(key >= rs && key < re)) a = 0.f;
        if (!(key + 1 >= rs && key + 1 < re)) b = 0.f;
      }
      l0 += a;
      l1 += b;
      if (kDrop) {  // the row sum (softmax normaliser) is taken before dropping; P·V sees the dropped, rescaled values
        const int key = kv0 + c + 2 * i;
        a = drop_keep(drop, head, row, key) ? a * drop.inv_keep : 0.f;
        b = drop_keep(drop, head, row, key + 1) ? b * drop.inv_keep : 0.f;
      }
      packed[i] = pack16<kHalf>(a, b);
    }
    tmem_st_32x32b_x16(s_tmem + c / 2, packed);
  }
}

// kHalf: q, k, v, out (and the probabilities fed back to the tensor core) are fp16 instead of bf16.
template <bool kHalf>
__global__ void __launch_bounds__(kThreads, 1)
attention_fwd_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                     const __grid_constant__ CUtensorMap map_v, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                                      // two tiles
  uint8_t* sK = smem + kTileBytes * 2;                     // kKStages tiles
  uint8_t* sV = smem + kTileBytes * (2 + kKStages);        // kVStages tiles
  Barriers* bars = reinterpret_cast<Barriers*>(smem + kTileBytes * (2 + kKStages + kVStages));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // heaviest (latest) query tiles first, over ALL heads, then the next lighter tile of all heads, … : under a causal
  // mask the KV loop length grows with the tile index, and longest-first keeps the last wave short
  const int n_q_tiles = (p.T + 2 * BQ - 1) / (2 * BQ);
  const int q0 = (n_q_tiles - 1 - (int)blockIdx.x / p.heads) * (2 * BQ);
  const int head = (int)blockIdx.x % p.heads;
  const int kv_head = head / (p.heads / p.kv_heads);
  // KV range: from the start of the first row's sequence to the diagonal (causal) or to the end of the last row's
  // sequence (bidirectional); seg_start/seg_end are non-decreasing along the packed dimension
  const int kv_begin = (p.seg_start[q0] / BKV) * BKV;
  int n_x[2];
#pragma unroll
  for (int x = 0; x < 2; ++x) {
    const int first = q0 + x * BQ;
    if (first >= p.T) {
      n_x[x] = 0;
    } else {
      const int last = min(first + BQ, p.T) - 1;
      const int end = p.causal ? last + 1 : p.seg_end[last];
      n_x[x] = (end - kv_begin + BKV - 1) / BKV;
    }
  }
  const int n_tiles = max(n_x[0], n_x[1]);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_k);
    tma_prefetch_desc(&map_v);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(&bars->q_full, 1);
    for (int i = 0; i < kKStages; ++i) {
      mbar_init(&bars->k_full[i], 1);
      mbar_init(&bars->k_empty[i], 1);
    }
    for (int i = 0; i < kVStages; ++i) {
      mbar_init(&bars->v_full[i], 1);
      mbar_init(&bars->v_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->s_full[i], 1);
      mbar_init(&bars->p_full[i], 4);
      mbar_init(&bars->o_done[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(&bars->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_base;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(&bars->q_full, 2 * kTileBytes);
      load_tile(sQ, &map_q, &bars->q_full, head * D, q0);
      load_tile(sQ + kTileBytes, &map_q, &bars->q_full, head * D, q0 + BQ);
      // K runs one tile ahead of V; interleave the two rings in the order the MMA warp consumes them
      for (int j = 0; j < n_tiles + 1; ++j) {
        if (j < n_tiles) {
          const int st = j % kKStages;
          mbar_wait(&bars->k_empty[st], ((j / kKStages) & 1) ^ 1);
          mbar_arrive_expect_tx(&bars->k_full[st], kTileBytes);
          load_tile(sK + st * kTileBytes, &map_k, &bars->k_full[st], kv_head * D, kv_begin + j * BKV);
        }
        if (j >= 1) {
          const int jv = j - 1, st = jv % kVStages;
          mbar_wait(&bars->v_empty[st], ((jv / kVStages) & 1) ^ 1);
          mbar_arrive_expect_tx(&bars->v_full[st], kTileBytes);
          load_tile(sV + st * kTileBytes, &map_v, &bars->v_full[st], kv_head * D, kv_begin + jv * BKV);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_16(BQ, BKV, 0, 0, kHalf);  // Q, K both K-major
      constexpr uint32_t idesc_o = make_idesc_16(BQ, D, 0, 1, kHalf);    // P K-major (TMEM), V MN-major
      auto issue_s = [&](int x, int j) {
        const uint32_t q_addr = smem_u32(sQ + x * kTileBytes);
        const uint32_t k_addr = smem_u32(sK + (j % kKStages) * kTileBytes);
        const uint32_t d_tmem = tmem + kColS + x * 128;
#pragma unroll
        for (int k = 0; k < D / 16; ++k) {
          const uint32_t off = (k >> 2) * (kTileBytes / 2) + (k & 3) * 32;
          umma_f16_ss(d_tmem, make_smem_desc_sw128(q_addr + off, 16, 1024), make_smem_desc_sw128(k_addr + off, 16, 1024),
                      idesc_s, k != 0);
        }
        umma_commit(&bars->s_full[x]);
      };
      auto wait_k = [&](int j) {
        mbar_wait(&bars->k_full[j % kKStages], (j / kKStages) & 1);
        tc_fence_after();
      };
      mbar_wait(&bars->q_full, 0);
      if (n_tiles > 0) {
        wait_k(0);
        for (int x = 0; x < 2; ++x)
          if (n_x[x] > 0) issue_s(x, 0);
        umma_commit(&bars->k_empty[0]);
      }
      for (int j = 0; j < n_tiles; ++j) {
        const int vst = j % kVStages;
        mbar_wait(&bars->v_full[vst], (j / kVStages) & 1);
        if (j + 1 < n_tiles) wait_k(j + 1);
        tc_fence_after();
        const uint32_t v_addr = smem_u32(sV + vst * kTileBytes);
        for (int x = 0; x < 2; ++x) {
          if (j < n_x[x]) {
            mbar_wait(&bars->p_full[x], j & 1);
            tc_fence_after();
            const uint32_t p_tmem = tmem + kColS + x * 128;
#pragma unroll
            for (int k = 0; k < BKV / 16; ++k) {
              // V is [kv, d] in smem: N (= d) contiguous -> MN-major; 16 keys = two 8-row groups = 2 KiB per step,
              // the second 64-wide d slab lives kTileBytes/2 further (LBO)
              umma_f16_ts(tmem + kColO + x * 128, p_tmem + k * 8,
                          make_smem_desc_sw128(v_addr + k * 2048, kTileBytes / 2, 1024), idesc_o, (j | k) != 0);
            }
            umma_commit(&bars->o_done[x]);
          }
          if (j + 1 < n_x[x]) issue_s(x, j + 1);
        }
        umma_commit(&bars->v_empty[vst]);
        if (j + 1 < n_tiles) umma_commit(&bars->k_empty[(j + 1) % kKStages]);
      }
    }
  } else if (warp >= 4) {
    const int x = (warp - 4) >> 2;  // query sub-tile of this warpgroup
    const int ew = warp & 3;
    const int row = q0 + x * BQ + ew * 32 + lane;
    const bool row_ok = row < p.T;
    const int rs = row_ok ? p.seg_start[row] : 0;
    const int re = row_ok ? (p.causal ? row + 1 : p.seg_end[row]) : 0;  // row attends to keys in [rs, re)
    const uint32_t lane_base = (uint32_t)(ew * 32) << 16;
    const uint32_t s_tmem = tmem + lane_base + kColS + x * 128;
    const uint32_t o_tmem = tmem + lane_base + kColO + x * 128;
    const int n_mine = n_x[x];
    float m_ref = -INFINITY;  // reference max (log2 units) the running sums are expressed against
    float l = 0.f;
    for (int j = 0; j < n_mine; ++j) {
      const int kv0 = kv_begin + j * BKV;
      mbar_wait(&bars->s_full[x], j & 1);
      tc_fence_after();
      // warp-uniform so that the warp-collective TMEM loads inside stay convergent
      const bool need_mask = __any_sync(0xffffffffu, (kv0 < rs) || (kv0 + BKV > re));
      // pass 1: row maximum of the visible scores
      const float mx = need_mask ? fwd_row_max<true>(s_tmem, kv0, rs, re) : fwd_row_max<false>(s_tmem, kv0, rs, re);
      const float m_new = fmaxf(m_ref, mx * p.scale_log2);
      // lazy rescale: keep the old reference unless the max grew by more than 8 (p stays <= 2^8, exact in the end)
      const bool grow = m_new > m_ref + 8.f;
      float alpha = 1.f;
      if (grow) {
        alpha = (m_ref == -INFINITY) ? 0.f : exp2f(m_ref - m_new);
        m_ref = m_new;
        l *= alpha;
      }
      const float m_use = (m_ref == -INFINITY) ? 0.f : m_ref;
      // pass 2: p = exp2(s·scale − m_ref), row sum, bf16 pairs written over the already consumed S columns
      float l0 = 0.f, l1 = 0.f;
      if (p.drop.thr) {
        if (need_mask)
          fwd_exp_pack<true, true, kHalf>(s_tmem, kv0, rs, re, p.scale_log2, m_use, l0, l1, p.drop, head, row);
        else
          fwd_exp_pack<false, true, kHalf>(s_tmem, kv0, rs, re, p.scale_log2, m_use, l0, l1, p.drop, head, row);
      } else if (need_mask) {
        fwd_exp_pack<true, false, kHalf>(s_tmem, kv0, rs, re, p.scale_log2, m_use, l0, l1, p.drop, head, row);
      } else {
        fwd_exp_pack<false, false, kHalf>(s_tmem, kv0, rs, re, p.scale_log2, m_use, l0, l1, p.drop, head, row);
      }
      l += l0 + l1;
      tmem_st_wait();
      if (j > 0) {
        // O may only be touched once PV(j-1) has landed
        mbar_wait(&bars->o_done[x], (j - 1) & 1);
        tc_fence_after();
        if (__any_sync(0xffffffffu, grow)) {
#pragma unroll 1
          for (int c = 0; c < D; c += 32) {
            uint32_t v[32];
            tmem_ld_32x32b_x32(o_tmem + c, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
            tmem_st_32x32b_x16(o_tmem + c, v);
            tmem_st_32x32b_x16(o_tmem + c + 16, v + 16);
          }
          tmem_st_wait();
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->p_full[x]);
    }
    // epilogue: O / l -> bf16, LSE in natural log units
    float inv = 0.f;
    if (n_mine > 0) {
      mbar_wait(&bars->o_done[x], (n_mine - 1) & 1);
      tc_fence_after();
      inv = l > 0.f ? 1.f / l : 0.f;
    }
    __nv_bfloat16* orow = p.out + ((size_t)row * p.heads + head) * D;
#pragma unroll 1
    for (int c = 0; c < D; c += 32) {
      uint32_t v[32];
      if (n_mine > 0) {
        tmem_ld_32x32b_x32(o_tmem + c, v);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0;
      }
      if (row_ok) {
        uint4* dst = reinterpret_cast<uint4*>(orow + c);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          dst[i] = make_uint4(pack16<kHalf>(__uint_as_float(v[8 * i]) * inv, __uint_as_float(v[8 * i + 1]) * inv),
                              pack16<kHalf>(__uint_as_float(v[8 * i + 2]) * inv, __uint_as_float(v[8 * i + 3]) * inv),
                              pack16<kHalf>(__uint_as_float(v[8 * i + 4]) * inv, __uint_as_float(v[8 * i + 5]) * inv),
                              pack16<kHalf>(__uint_as_float(v[8 * i + 6]) * inv, __uint_as_float(v[8 * i + 7]) * inv));
        }
      }
    }
    if (row_ok) p.lse[(size_t)head * p.T + row] = l > 0.f ? (m_ref + log2f(l)) * 0.6931471805599453f : -INFINITY;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}
// [rows, width] bf16 with row pitch `pitch` elements; box 64 x 128, 128B swizzle
int make_map(CUtensorMap* m, const void* ptr, uint64_t width, uint64_t rows, uint64_t pitch) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return -1;
  cuuint64_t dims[2] = {width, rows};
  cuuint64_t strides[1] = {pitch * 2};
  cuuint32_t box[2] = {64, 128};
  cuuint32_t es[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS
             ? 0
             : -2;
}

}  // namespace attn
}  // namespace sb

// q: [T, heads, 128] with token pitch q_pitch (elements); k, v: [T, kv_heads, 128] with pitches k_pitch / v_pitch.
// seg_start / seg_end: int32 [T] sequence bounds of every token (packed batches).  out: [T, heads, 128] contiguous,
// lse: fp32 [heads, T].  scale: softmax scale (e.g. 1/sqrt(128)).
//
// sb_attention_fwd_ex generalises the key side: k, v hold Tk rows (Tk != T for KV-cache decoding, where the new queries
// attend to cached keys), and with causal == 0 the two arrays are simply "row t attends to keys [seg_start[t],
// seg_end[t])" in KEY index space — any mask whose bounds are non-decreasing along t: causal (seg_end[t] = t + 1),
// right-aligned causal for decoding, sliding windows (seg_start[t] = max(sequence start, t - window)), documents.
extern "C" int sb_attention_fwd_ex(const void* q, const void* k, const void* v, void* out, float* lse, const int* seg_start,
                                   const int* seg_end, int T, int Tk, int heads, int kv_heads, long long q_pitch,
                                   long long k_pitch, long long v_pitch, float scale, int causal, float dropout_p,
                                   unsigned long long dropout_seed, int head_offset, cudaStream_t stream);
extern "C" int sb_attention_fwd(const void* q, const void* k, const void* v, void* out, float* lse, const int* seg_start,
                                const int* seg_end, int T, int heads, int kv_heads, long long q_pitch, long long k_pitch,
                                long long v_pitch, float scale, int causal, cudaStream_t stream) {
  return sb_attention_fwd_ex(q, k, v, out, lse, seg_start, seg_end, T, T, heads, kv_heads, q_pitch, k_pitch, v_pitch, scale,
                             causal, 0.f, 0ull, 0, stream);
}
template <bool kHalf>
static int attention_fwd_impl(const void* q, const void* k, const void* v, void* out, float* lse, const int* seg_start,
                              const int* seg_end, int T, int Tk, int heads, int kv_heads, long long q_pitch,
                              long long k_pitch, long long v_pitch, float scale, int causal, float dropout_p,
                              unsigned long long dropout_seed, int head_offset, cudaStream_t stream) {
  using namespace sb::attn;
  if (T <= 0) return 0;
  if (heads % kv_heads || Tk <= 0 || (causal && Tk != T) || dropout_p < 0.f || dropout_p >= 1.f) return -1;
  CUtensorMap mq, mk, mv;
  if (make_map(&mq, q, (uint64_t)heads * D, T, q_pitch)) return -2;
  if (make_map(&mk, k, (uint64_t)kv_heads * D, Tk, k_pitch)) return -2;
  if (make_map(&mv, v, (uint64_t)kv_heads * D, Tk, v_pitch)) return -2;
  Params p;
  p.T = T, p.heads = heads, p.kv_heads = kv_heads, p.scale_log2 = scale * 1.4426950408889634f, p.causal = causal;
  p.seg_start = seg_start, p.seg_end = seg_end, p.out = (__nv_bfloat16*)out, p.lse = lse;
  p.drop = make_drop(dropout_p, dropout_seed, head_offset);
  static sb::PerDeviceOnce once;
  if (once.first())
    SB_CUDA_CHECK(cudaFuncSetAttribute(attention_fwd_kernel<kHalf>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kSmemBytes));
  const int blocks = (T + 2 * BQ - 1) / (2 * BQ) * heads;
  attention_fwd_kernel<kHalf><<<blocks, kThreads, kSmemBytes, stream>>>(mq, mk, mv, p);
  return (int)cudaGetLastError();
}
extern "C" int sb_attention_fwd_ex(const void* q, const void* k, const void* v, void* out, float* lse, const int* seg_start,
                                   const int* seg_end, int T, int Tk, int heads, int kv_heads, long long q_pitch,
                                   long long k_pitch, long long v_pitch, float scale, int causal, float dropout_p,
                                   unsigned long long dropout_seed, int head_offset, cudaStream_t stream) {
  return attention_fwd_impl<false>(q, k, v, out, lse, seg_start, seg_end, T, Tk, heads, kv_heads, q_pitch, k_pitch, v_pitch,
                                   scale, causal, dropout_p, dropout_seed, head_offset, stream);
}
// Same contract with fp16 tensors (the reference's fp16 + loss-scaler grid, tests/core/test_training/test_training.py).
extern "C" int sb_attention_fwd_f16(const void* q, const void* k, const void* v, void* out, float* lse, const int* seg_start,
                                    const int* seg_end, int T, int Tk, int heads, int kv_heads, long long q_pitch,
                                    long long k_pitch, long long v_pitch, float scale, int causal, float dropout_p,
                                    unsigned long long dropout_seed, int head_offset, cudaStream_t stream) {
  return attention_fwd_impl<true>(q, k, v, out, lse, seg_start, seg_end, T, Tk, heads, kv_heads, q_pitch, k_pitch, v_pitch,
                                  scale, causal, dropout_p, dropout_seed, head_offset, stream);
}

// =================================================================================================================
// Backward.  Three launches, deterministic (no atomics):
//   1. delta[h, t] = Σ_d dO·O                                   (bwd_delta_kernel)
//   2. dK, dV: one CTA per (128-key tile, kv head), streaming 64-query sub-tiles of every query head of the GQA group.
//      Scores are formed transposed, Sᵀ = K·Qᵀ and dPᵀ = V·dOᵀ (keys on TMEM lanes), so Pᵀ and dSᵀ are written back
//      to TMEM as bf16 and feed dV += Pᵀ·dO and dK += dSᵀ·Q directly as TMEM A-operands — no shared-memory round trip.
//      The group's query heads accumulate into the same TMEM dK/dV, which is the GQA reduction.
//   3. dQ: one CTA per (128-query tile, head), streaming 64-key sub-tiles: S = Q·Kᵀ, dP = dO·Vᵀ, dS (bf16, TMEM)
//      -> dQ += dS·K.
// In both kernels S/dP are double buffered in TMEM (2 x 128 columns) and two softmax warpgroups alternate over the
// stream, so the tensor pipe computes sub-tile i+1 while sub-tile i is in the exp2 / mask / pack stage.
// =================================================================================================================
namespace sb {
namespace attn {

constexpr int BS = 64;                                   // streamed sub-tile rows
constexpr uint32_t kSubBytes = BS * 128 * 2;             // 16 KiB: [64, 128] bf16 as two 64-column slabs of 8 KiB
constexpr int kRing = 3;
constexpr int kMaxSubs = 2048;                           // per-sub mask flags (T up to 131072 per packed batch)
constexpr uint32_t kStatBytes = kRing * 2 * BS * 4;      // staged lse·log2e and delta of the streamed queries
constexpr uint32_t kBwdSmemBytes = 2 * kTileBytes + kRing * 2 * kSubBytes + kStatBytes + kMaxSubs + 1024 + 256;
constexpr float kLog2e = 1.4426950408889634f;

struct BwdParams {
  int T, heads, kv_heads;
  float scale, scale_log2;
  int causal;
  const int* seg_start;
  const int* seg_end;
  const int* key_qend;  // optional [T]: one past the last query that sees key t (tighter than seg_end under a window)
  Drop drop;
  const float* lse2;    // [heads, Tp] lse · log2(e)   (Tp = padded pitch, multiple of 64 and > T)
  const float* delta;   // [heads, Tp] Σ_d dO·O
  int Tp;
  __nv_bfloat16* dq;    // [T, heads, 128]
  __nv_bfloat16* dk;    // [T, kv_heads, 128]
  __nv_bfloat16* dv;
};

struct __align__(8) BwdBarriers {
  uint64_t tile_full;
  uint64_t ring_full[kRing];
  uint64_t ring_empty[kRing];
  uint64_t s_full[2];
  uint64_t p_full[2];
  uint64_t done;
  uint32_t tmem_base;
};

SB_DEVICE void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

SB_DEVICE void load_sub(uint8_t* dst, const CUtensorMap* map, uint64_t* bar, int col0, int row0) {
  tma_load_2d(dst, map, bar, col0, row0);
  tma_load_2d(dst + kSubBytes / 2, map, bar, col0 + 64, row0);
}

template <bool kHalf>
__global__ void bwd_delta_kernel(const __nv_bfloat16* __restrict__ dout, const __nv_bfloat16* __restrict__ out,
                                 const float* __restrict__ lse, float* __restrict__ lse2, float* __restrict__ delta, int T,
                                 int heads, int Tp) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= T * heads) return;
  const int t = w / heads, h = w % heads;
  const uint2 a = *reinterpret_cast<const uint2*>(dout + (size_t)w * 128 + lane * 4);
  const uint2 b = *reinterpret_cast<const uint2*>(out + (size_t)w * 128 + lane * 4);
  float2 a0 = unpack16<kHalf>(a.x), a1 = unpack16<kHalf>(a.y), b0 = unpack16<kHalf>(b.x), b1 = unpack16<kHalf>(b.y);
  float s = warp_sum(a0.x * b0.x + a0.y * b0.y + a1.x * b1.x + a1.y * b1.y);
  if (lane == 0) {
    delta[(size_t)h * Tp + t] = s;
    lse2[(size_t)h * Tp + t] = lse[(size_t)h * T + t] * kLog2e;
  }
}

template <bool kHalf>
SB_DEVICE void store_row_16(__nv_bfloat16* dst, uint32_t tmem_addr, bool ok) {
#pragma unroll 1
  for (int c = 0; c < 128; c += 32) {
    uint32_t v[32];
    tmem_ld_32x32b_x32(tmem_addr + c, v);
    tmem_ld_wait();
    if (ok) {
      uint4* d = reinterpret_cast<uint4*>(dst + c);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        d[i] = make_uint4(pack16<kHalf>(__uint_as_float(v[8 * i]), __uint_as_float(v[8 * i + 1])),
                          pack16<kHalf>(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3])),
                          pack16<kHalf>(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5])),
                          pack16<kHalf>(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7])));
    }
  }
}

// ------------------------------------------------------------------------------------------------- dK / dV

template <bool kMask, bool kDrop, bool kHalf>
SB_DEVICE void dkdv_softmax_chunks(uint32_t t_s, uint32_t t_p, const float* st_lse2, const float* st_del, int qs0, int key,
                                   bool key_ok, const BwdParams& p, int head) {
#pragma unroll 1
  for (int c = 0; c < BS; c += 32) {
    uint32_t s[32], dp[32];
    tmem_ld_32x32b_x32(t_s + c, s);
    tmem_ld_32x32b_x32(t_p + c, dp);
    tmem_ld_wait();
    uint32_t pp[16], pd[16];
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      // four consecutive queries: their lse·log2e and delta come from shared memory as one 16-byte broadcast each
      const float4 l4 = *reinterpret_cast<const float4*>(st_lse2 + c + 2 * i);
      const float4 d4 = *reinterpret_cast<const float4*>(st_del + c + 2 * i);
      const float lv[4] = {l4.x, l4.y, l4.z, l4.w}, dl[4] = {d4.x, d4.y, d4.z, d4.w};
      float pv[4], ds[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float pr = exp2f(__uint_as_float(s[2 * i + e]) * p.scale_log2 - lv[e]);
        if (kMask) {
          const int qq = qs0 + c + 2 * i + e;
          const int qc = min(qq, p.T - 1);
          const bool vis = key_ok && qq < p.T && key >= __ldg(p.seg_start + qc) &&
                           key < (p.causal ? qq + 1 : __ldg(p.seg_end + qc));
          if (!vis) pr = 0.f;
        }
        float dpe = __uint_as_float(dp[2 * i + e]);
        pv[e] = pr;
        if (kDrop) {  // dV sees the dropped, rescaled probabilities; dS = P ∘ (D ∘ dP / (1 − p) − delta)
          const bool keep = drop_keep(p.drop, head, qs0 + c + 2 * i + e, key);
          pv[e] = keep ? pr * p.drop.inv_keep : 0.f;
          dpe = keep ? dpe * p.drop.inv_keep : 0.f;
        }
        ds[e] = pr * (dpe - dl[e]) * p.scale;
      }
      pp[i] = pack16<kHalf>(pv[0], pv[1]);
      pp[i + 1] = pack16<kHalf>(pv[2], pv[3]);
      pd[i] = pack16<kHalf>(ds[0], ds[1]);
      pd[i + 1] = pack16<kHalf>(ds[2], ds[3]);
    }
    tmem_st_32x32b_x16(t_s + c / 2, pp);  // P^T over the consumed S^T columns
    tmem_st_32x32b_x16(t_p + c / 2, pd);  // dS^T over the consumed dP^T columns
  }
}
template <bool kHalf>
SB_DEVICE void bwd_dkdv_body(const CUtensorMap& map_q, const CUtensorMap& map_do, const CUtensorMap& map_k,
                             const CUtensorMap& map_v, const BwdParams& p, const int kv_tile, const int kv_head) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = smem + kTileBytes;
  uint8_t* ring = smem + 2 * kTileBytes;  // stage: [Q_sub | dO_sub]
  float* stats = reinterpret_cast<float*>(ring + kRing * 2 * kSubBytes);  // stage: [lse2 x 64 | delta x 64]
  uint8_t* sub_mask = reinterpret_cast<uint8_t*>(stats) + kStatBytes;
  BwdBarriers* bars = reinterpret_cast<BwdBarriers*>(sub_mask + kMaxSubs);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kv0 = kv_tile * 128;
  const int kvh = kv_head;
  const int group = p.heads / p.kv_heads;
  const int kv_last = min(kv0 + 128, p.T) - 1;
  // queries that can see a key of this tile: same sequence(s), and not before the key when causal
  const int q_begin = ((p.causal ? kv0 : p.seg_start[kv0]) / BS) * BS;
  const int q_end = (p.key_qend ? p.key_qend : p.seg_end)[kv_last];
  const int n_sub = (q_end - q_begin + BS - 1) / BS;
  const int n_it = n_sub * group;
  // which query sub-tiles need element-wise masking against this key tile (diagonal, document edges, tails)
  for (int sub = threadIdx.x; sub < n_sub; sub += blockDim.x) {
    const int qs0 = q_begin + sub * BS;
    const int q_hi = min(qs0 + BS, p.T) - 1;
    sub_mask[sub] = (p.causal && qs0 < kv0 + 127) || p.seg_start[q_hi] > kv0 || p.seg_end[min(qs0, p.T - 1)] <= kv_last ||
                    qs0 + BS > p.T || kv0 + 128 > p.T;
  }

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_do);
    tma_prefetch_desc(&map_k);
    tma_prefetch_desc(&map_v);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(&bars->tile_full, 1);
    for (int i = 0; i < kRing; ++i) {
      mbar_init(&bars->ring_full[i], 1);
      mbar_init(&bars->ring_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->s_full[i], 1);
      mbar_init(&bars->p_full[i], 4);
    }
    mbar_init(&bars->done, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(&bars->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_base;
  constexpr uint32_t kColdV = 256, kColdK = 384;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(&bars->tile_full, 2 * kTileBytes);
      load_tile(sK, &map_k, &bars->tile_full, kvh * D, kv0);
      load_tile(sV, &map_v, &bars->tile_full, kvh * D, kv0);
      for (int it = 0; it < n_it; ++it) {
        const int st = it % kRing;
        const int head = kvh * group + it / n_sub;
        const int qs0 = q_begin + (it % n_sub) * BS;
        mbar_wait(&bars->ring_empty[st], ((it / kRing) & 1) ^ 1);
        mbar_arrive_expect_tx(&bars->ring_full[st], 2 * kSubBytes + 2 * BS * 4);
        load_sub(ring + st * 2 * kSubBytes, &map_q, &bars->ring_full[st], head * D, qs0);
        load_sub(ring + st * 2 * kSubBytes + kSubBytes, &map_do, &bars->ring_full[st], head * D, qs0);
        bulk_load(stats + st * 2 * BS, p.lse2 + (size_t)head * p.Tp + qs0, BS * 4, &bars->ring_full[st]);
        bulk_load(stats + st * 2 * BS + BS, p.delta + (size_t)head * p.Tp + qs0, BS * 4, &bars->ring_full[st]);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_t = make_idesc_16(128, BS, 0, 0, kHalf);   // Sᵀ / dPᵀ: [128 keys, 64 queries]
      constexpr uint32_t idesc_g = make_idesc_16(128, 128, 0, 1, kHalf);  // dV / dK : A from TMEM, B MN-major
      const uint32_t k_addr = smem_u32(sK), v_addr = smem_u32(sV);
      auto issue_scores = [&](int it) {
        const int st = it % kRing;
        mbar_wait(&bars->ring_full[st], (it / kRing) & 1);
        tc_fence_after();
        const uint32_t q_addr = smem_u32(ring + st * 2 * kSubBytes), do_addr = q_addr + kSubBytes;
        const uint32_t d_s = tmem + (it & 1) * 128, d_p = d_s + 64;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t a_off = (k >> 2) * (kTileBytes / 2) + (k & 3) * 32, b_off = (k >> 2) * (kSubBytes / 2) + (k & 3) * 32;
          umma_f16_ss(d_s, make_smem_desc_sw128(k_addr + a_off, 16, 1024), make_smem_desc_sw128(q_addr + b_off, 16, 1024),
                      idesc_t, k != 0);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t a_off = (k >> 2) * (kTileBytes / 2) + (k & 3) * 32, b_off = (k >> 2) * (kSubBytes / 2) + (k & 3) * 32;
          umma_f16_ss(d_p, make_smem_desc_sw128(v_addr + a_off, 16, 1024), make_smem_desc_sw128(do_addr + b_off, 16, 1024),
                      idesc_t, k != 0);
        }
        umma_commit(&bars->s_full[it & 1]);
      };
      mbar_wait(&bars->tile_full, 0);
      if (n_it > 0) issue_scores(0);
      for (int it = 0; it < n_it; ++it) {
        if (it + 1 < n_it) issue_scores(it + 1);
        const int st = it % kRing;
        mbar_wait(&bars->p_full[it & 1], (it >> 1) & 1);
        tc_fence_after();
        const uint32_t q_addr = smem_u32(ring + st * 2 * kSubBytes), do_addr = q_addr + kSubBytes;
        const uint32_t a_p = tmem + (it & 1) * 128, a_ds = a_p + 64;
#pragma unroll
        for (int k = 0; k < BS / 16; ++k)  // dV += Pᵀ · dO      (B: [64 queries, 128 d], d contiguous -> MN-major)
          umma_f16_ts(tmem + kColdV, a_p + k * 8, make_smem_desc_sw128(do_addr + k * 2048, kSubBytes / 2, 1024), idesc_g,
                      (it | k) != 0);
#pragma unroll
        for (int k = 0; k < BS / 16; ++k)  // dK += dSᵀ · Q
          umma_f16_ts(tmem + kColdK, a_ds + k * 8, make_smem_desc_sw128(q_addr + k * 2048, kSubBytes / 2, 1024), idesc_g,
                      (it | k) != 0);
        umma_commit(&bars->ring_empty[st]);
      }
      umma_commit(&bars->done);
    }
  } else if (warp >= 4) {
    const int wg = (warp - 4) >> 2;
    const int ew = warp & 3;
    const int key = kv0 + ew * 32 + lane;
    const bool key_ok = key < p.T;
    const uint32_t lane_base = (uint32_t)(ew * 32) << 16;
    for (int it = wg; it < n_it; it += 2) {
      const int qs0 = q_begin + (it % n_sub) * BS;
      mbar_wait(&bars->ring_full[it % kRing], (it / kRing) & 1);  // makes the staged lse2 / delta visible here
      mbar_wait(&bars->s_full[it & 1], (it >> 1) & 1);
      tc_fence_after();
      const float* st_lse2 = stats + (it % kRing) * 2 * BS;  // staged by the producer together with Q_sub / dO_sub
      const float* st_del = st_lse2 + BS;
      const uint32_t t_s = tmem + lane_base + (it & 1) * 128, t_p = t_s + 64;
      const int hq = kvh * group + it / n_sub;  // query head of this sub-tile (for the dropout mask)
      if (p.drop.thr) {
        if (sub_mask[it % n_sub])
          dkdv_softmax_chunks<true, true, kHalf>(t_s, t_p, st_lse2, st_del, qs0, key, key_ok, p, hq);
        else
          dkdv_softmax_chunks<false, true, kHalf>(t_s, t_p, st_lse2, st_del, qs0, key, key_ok, p, hq);
      } else if (sub_mask[it % n_sub]) {
        dkdv_softmax_chunks<true, false, kHalf>(t_s, t_p, st_lse2, st_del, qs0, key, key_ok, p, hq);
      } else {
        dkdv_softmax_chunks<false, false, kHalf>(t_s, t_p, st_lse2, st_del, qs0, key, key_ok, p, hq);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->p_full[it & 1]);
    }
    // epilogue: warpgroup 0 writes dV, warpgroup 1 writes dK
    mbar_wait(&bars->done, 0);
    tc_fence_after();
    __nv_bfloat16* dst = (wg == 0 ? p.dv : p.dk) + ((size_t)key * p.kv_heads + kvh) * D;
    if (n_it > 0) {
      store_row_16<kHalf>(dst, tmem + lane_base + (wg == 0 ? kColdV : kColdK), key_ok);
    } else if (key_ok) {
      for (int c = 0; c < D; c += 8) *reinterpret_cast<uint4*>(dst + c) = make_uint4(0, 0, 0, 0);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

// ------------------------------------------------------------------------------------------------- dQ

template <bool kMask, bool kDrop, bool kHalf>
SB_DEVICE void dq_softmax_chunks(uint32_t t_s, uint32_t t_p, int ks0, int rs, int re, float scale_log2, float lse2,
                                 float del, float scale, const Drop& drop, int head, int row) {
#pragma unroll 1
  for (int c = 0; c < BS; c += 32) {
    uint32_t s[32], dp[32];
    tmem_ld_32x32b_x32(t_s + c, s);
    tmem_ld_32x32b_x32(t_p + c, dp);
    tmem_ld_wait();
    uint32_t pd[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float d2[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        float pr = exp2f(__uint_as_float(s[2 * i + e]) * scale_log2 - lse2);
        if (kMask) {
          const int kk = ks0 + c + 2 * i + e;
          if (!(kk >= rs && kk < re)) pr = 0.f;
        }
        float dpe = __uint_as_float(dp[2 * i + e]);
        if (kDrop) dpe = drop_keep(drop, head, row, ks0 + c + 2 * i + e) ? dpe * drop.inv_keep : 0.f;
        d2[e] = pr * (dpe - del) * scale;
      }
      pd[i] = pack16<kHalf>(d2[0], d2[1]);
    }
    tmem_st_32x32b_x16(t_p + c / 2, pd);  // dS over the consumed dP columns
  }
}
template <bool kHalf>
SB_DEVICE void bwd_dq_body(const CUtensorMap& map_q, const CUtensorMap& map_do, const CUtensorMap& map_k,
                           const CUtensorMap& map_v, const BwdParams& p, const int q_tile, const int q_head) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sdO = smem + kTileBytes;
  uint8_t* ring = smem + 2 * kTileBytes;  // stage: [K_sub | V_sub]
  BwdBarriers* bars = reinterpret_cast<BwdBarriers*>(ring + kRing * 2 * kSubBytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = q_tile * 128;
  const int head = q_head;
  const int kvh = head / (p.heads / p.kv_heads);
  const int q_last = min(q0 + 128, p.T) - 1;
  const int kv_begin = (p.seg_start[q0] / BS) * BS;
  const int kv_end = p.causal ? q_last + 1 : p.seg_end[q_last];
  const int n_it = (kv_end - kv_begin + BS - 1) / BS;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_do);
    tma_prefetch_desc(&map_k);
    tma_prefetch_desc(&map_v);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(&bars->tile_full, 1);
    for (int i = 0; i < kRing; ++i) {
      mbar_init(&bars->ring_full[i], 1);
      mbar_init(&bars->ring_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->s_full[i], 1);
      mbar_init(&bars->p_full[i], 4);
    }
    mbar_init(&bars->done, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(&bars->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_base;
  constexpr uint32_t kColdQ = 256;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(&bars->tile_full, 2 * kTileBytes);
      load_tile(sQ, &map_q, &bars->tile_full, head * D, q0);
      load_tile(sdO, &map_do, &bars->tile_full, head * D, q0);
      for (int it = 0; it < n_it; ++it) {
        const int st = it % kRing;
        const int ks0 = kv_begin + it * BS;
        mbar_wait(&bars->ring_empty[st], ((it / kRing) & 1) ^ 1);
        mbar_arrive_expect_tx(&bars->ring_full[st], 2 * kSubBytes);
        load_sub(ring + st * 2 * kSubBytes, &map_k, &bars->ring_full[st], kvh * D, ks0);
        load_sub(ring + st * 2 * kSubBytes + kSubBytes, &map_v, &bars->ring_full[st], kvh * D, ks0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_t = make_idesc_16(128, BS, 0, 0, kHalf);   // S / dP: [128 queries, 64 keys]
      constexpr uint32_t idesc_g = make_idesc_16(128, 128, 0, 1, kHalf);  // dQ: A (dS) from TMEM, B (K_sub) MN-major
      const uint32_t q_addr = smem_u32(sQ), do_addr = smem_u32(sdO);
      auto issue_scores = [&](int it) {
        const int st = it % kRing;
        mbar_wait(&bars->ring_full[st], (it / kRing) & 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(ring + st * 2 * kSubBytes), v_addr = k_addr + kSubBytes;
        const uint32_t d_s = tmem + (it & 1) * 128, d_p = d_s + 64;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t a_off = (k >> 2) * (kTileBytes / 2) + (k & 3) * 32, b_off = (k >> 2) * (kSubBytes / 2) + (k & 3) * 32;
          umma_f16_ss(d_s, make_smem_desc_sw128(q_addr + a_off, 16, 1024), make_smem_desc_sw128(k_addr + b_off, 16, 1024),
                      idesc_t, k != 0);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t a_off = (k >> 2) * (kTileBytes / 2) + (k & 3) * 32, b_off = (k >> 2) * (kSubBytes / 2) + (k & 3) * 32;
          umma_f16_ss(d_p, make_smem_desc_sw128(do_addr + a_off, 16, 1024), make_smem_desc_sw128(v_addr + b_off, 16, 1024),
                      idesc_t, k != 0);
        }
        umma_commit(&bars->s_full[it & 1]);
      };
      mbar_wait(&bars->tile_full, 0);
      if (n_it > 0) issue_scores(0);
      for (int it = 0; it < n_it; ++it) {
        if (it + 1 < n_it) issue_scores(it + 1);
        const int st = it % kRing;
        mbar_wait(&bars->p_full[it & 1], (it >> 1) & 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(ring + st * 2 * kSubBytes);
        const uint32_t a_ds = tmem + (it & 1) * 128 + 64;
#pragma unroll
        for (int k = 0; k < BS / 16; ++k)  // dQ += dS · K_sub   (B: [64 keys, 128 d], d contiguous -> MN-major)
          umma_f16_ts(tmem + kColdQ, a_ds + k * 8, make_smem_desc_sw128(k_addr + k * 2048, kSubBytes / 2, 1024), idesc_g,
                      (it | k) != 0);
        umma_commit(&bars->ring_empty[st]);
      }
      umma_commit(&bars->done);
    }
  } else if (warp >= 4) {
    const int wg = (warp - 4) >> 2;
    const int ew = warp & 3;
    const int row = q0 + ew * 32 + lane;
    const bool row_ok = row < p.T;
    const int rc = min(row, p.T - 1);
    const int rs = row_ok ? p.seg_start[rc] : 0;
    const int re = row_ok ? (p.causal ? row + 1 : p.seg_end[rc]) : 0;
    const float lse2 = p.lse2[(size_t)head * p.Tp + rc];
    const float del = p.delta[(size_t)head * p.Tp + rc];
    const uint32_t lane_base = (uint32_t)(ew * 32) << 16;
    for (int it = wg; it < n_it; it += 2) {
      const int ks0 = kv_begin + it * BS;
      mbar_wait(&bars->s_full[it & 1], (it >> 1) & 1);
      tc_fence_after();
      const bool need_mask = __any_sync(0xffffffffu, (ks0 < rs) || (ks0 + BS > re));
      const uint32_t t_s = tmem + lane_base + (it & 1) * 128, t_p = t_s + 64;
      if (p.drop.thr) {
        if (need_mask)
          dq_softmax_chunks<true, true, kHalf>(t_s, t_p, ks0, rs, re, p.scale_log2, lse2, del, p.scale, p.drop, head, row);
        else
          dq_softmax_chunks<false, true, kHalf>(t_s, t_p, ks0, rs, re, p.scale_log2, lse2, del, p.scale, p.drop, head, row);
      } else if (need_mask) {
        dq_softmax_chunks<true, false, kHalf>(t_s, t_p, ks0, rs, re, p.scale_log2, lse2, del, p.scale, p.drop, head, row);
      } else {
        dq_softmax_chunks<false, false, kHalf>(t_s, t_p, ks0, rs, re, p.scale_log2, lse2, del, p.scale, p.drop, head, row);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->p_full[it & 1]);
    }
    mbar_wait(&bars->done, 0);
    tc_fence_after();
    if (wg == 0) {
      __nv_bfloat16* dst = p.dq + ((size_t)row * p.heads + head) * D;
      if (n_it > 0) {
        store_row_16<kHalf>(dst, tmem + lane_base + kColdQ, row_ok);
      } else if (row_ok) {
        for (int c = 0; c < D; c += 8) *reinterpret_cast<uint4*>(dst + c) = make_uint4(0, 0, 0, 0);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

static int make_map_box(CUtensorMap* m, const void* ptr, uint64_t width, uint64_t rows, uint64_t pitch, uint32_t box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return -1;
  cuuint64_t dims[2] = {width, rows};
  cuuint64_t strides[1] = {pitch * 2};
  cuuint32_t box[2] = {64, box_rows};
  cuuint32_t es[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS
             ? 0
             : -2;
}

// One launch for the whole backward: blocks [0, n_dkdv) compute dK/dV of one (key tile, kv head), blocks
// [n_dkdv, n_dkdv + n_dq) compute dQ of one (query tile, head).  Both sets are ordered heaviest first (causal: early key
// tiles / late query tiles see the most partners), so the block scheduler packs the light dQ tiles into the holes the
// long dK/dV tiles leave — with few heads per GPU (tensor parallel) the dK/dV grid alone is a single, badly balanced wave.
template <bool kHalf>
__global__ void __launch_bounds__(384, 1)
attention_bwd_kernel(const __grid_constant__ CUtensorMap kv_map_q, const __grid_constant__ CUtensorMap kv_map_do,
                     const __grid_constant__ CUtensorMap kv_map_k, const __grid_constant__ CUtensorMap kv_map_v,
                     const __grid_constant__ CUtensorMap q_map_q, const __grid_constant__ CUtensorMap q_map_do,
                     const __grid_constant__ CUtensorMap q_map_k, const __grid_constant__ CUtensorMap q_map_v,
                     const BwdParams p, const int first_block) {
  const int b = (int)blockIdx.x + first_block;
  const int tiles = (p.T + 127) / 128;
  const int n_dkdv = tiles * p.kv_heads;
  if (b < n_dkdv) {
    bwd_dkdv_body<kHalf>(kv_map_q, kv_map_do, kv_map_k, kv_map_v, p, b / p.kv_heads, b % p.kv_heads);
  } else {
    const int c = b - n_dkdv;
    bwd_dq_body<kHalf>(q_map_q, q_map_do, q_map_k, q_map_v, p, tiles - 1 - c / p.heads, c % p.heads);
  }
}

}  // namespace attn
}  // namespace sb

// row pitch (floats) of the two per-(head, token) scratch planes: a multiple of 64 with at least 64 slack, so 64-float
// bulk copies that start at any sub-tile stay in bounds and 16-byte aligned
extern "C" int sb_attention_bwd_pitch(int T) { return (T + 63) / 64 * 64 + 64; }
extern "C" long long sb_attention_bwd_scratch_floats(int T, int heads) {
  return 2ll * heads * sb_attention_bwd_pitch(T);
}

// dout, out: [T, heads, 128] contiguous; q/k/v as in sb_attention_fwd; lse: [heads, T] from the forward pass;
// scratch: fp32 buffer of sb_attention_bwd_scratch_floats(T, heads) owned by the caller (must start zero-filled or at
// least finite: the padding is read but masked); dq: [T, heads, 128], dk/dv: [T, kv_heads, 128] contiguous.
//
// sb_attention_bwd_ex: as sb_attention_bwd, plus key_qend (int32 [T], may be NULL): one past the last query row that can
// see key t, non-decreasing — lets the dK/dV blocks stop early under a sliding window (the masks themselves come from
// seg_start / seg_end, which the caller narrows per row).
extern "C" int sb_attention_bwd_ex(const void* dout, const void* q, const void* k, const void* v, const void* out,
                                   const float* lse, float* scratch, void* dq, void* dk, void* dv, const int* seg_start,
                                   const int* seg_end, const int* key_qend, int T, int heads, int kv_heads,
                                   long long q_pitch, long long k_pitch, long long v_pitch, float scale, int causal,
                                   float dropout_p, unsigned long long dropout_seed, int head_offset, cudaStream_t stream);
extern "C" int sb_attention_bwd(const void* dout, const void* q, const void* k, const void* v, const void* out,
                                const float* lse, float* scratch, void* dq, void* dk, void* dv, const int* seg_start,
                                const int* seg_end, int T, int heads, int kv_heads, long long q_pitch, long long k_pitch,
                                long long v_pitch, float scale, int causal, cudaStream_t stream) {
  return sb_attention_bwd_ex(dout, q, k, v, out, lse, scratch, dq, dk, dv, seg_start, seg_end, nullptr, T, heads, kv_heads,
                             q_pitch, k_pitch, v_pitch, scale, causal, 0.f, 0ull, 0, stream);
}
template <bool kHalf>
static int attention_bwd_impl(const void* dout, const void* q, const void* k, const void* v, const void* out,
                              const float* lse, float* scratch, void* dq, void* dk, void* dv, const int* seg_start,
                              const int* seg_end, const int* key_qend, int T, int heads, int kv_heads,
                              long long q_pitch, long long k_pitch, long long v_pitch, float scale, int causal,
                              float dropout_p, unsigned long long dropout_seed, int head_offset, cudaStream_t stream) {
  using namespace sb::attn;
  if (T <= 0) return 0;
  if (heads % kv_heads) return -1;
  if ((T + BS - 1) / BS > kMaxSubs) return -3;
  const int Tp = sb_attention_bwd_pitch(T);
  float* lse2 = scratch;
  float* delta = scratch + (size_t)heads * Tp;
  bwd_delta_kernel<kHalf><<<(T * heads * 32 + 255) / 256, 256, 0, stream>>>(
      (const __nv_bfloat16*)dout, (const __nv_bfloat16*)out, lse, lse2, delta, T, heads, Tp);
  BwdParams p;
  p.T = T, p.heads = heads, p.kv_heads = kv_heads, p.scale = scale, p.scale_log2 = scale * kLog2e, p.causal = causal;
  p.seg_start = seg_start, p.seg_end = seg_end, p.key_qend = key_qend, p.lse2 = lse2, p.delta = delta, p.Tp = Tp;
  p.drop = make_drop(dropout_p, dropout_seed, head_offset);
  p.dq = (__nv_bfloat16*)dq, p.dk = (__nv_bfloat16*)dk, p.dv = (__nv_bfloat16*)dv;
  static sb::PerDeviceOnce once;
  static bool separate = false;  // SCALING_B200_ATTN_BWD_SEPARATE=1: two launches (dK/dV, then dQ), for A/B timing
  if (once.first()) {
    SB_CUDA_CHECK(cudaFuncSetAttribute(attention_bwd_kernel<kHalf>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kBwdSmemBytes));
    const char* e = getenv("SCALING_B200_ATTN_BWD_SEPARATE");
    separate = e && e[0] == '1';
  }
  // dK/dV blocks: K/V tiles stationary (128-row boxes), Q/dO streamed (64-row boxes); dQ blocks: the other way round
  CUtensorMap kq, kdo, kk, kv_, qq, qdo, qk, qv;
  if (make_map_box(&kq, q, (uint64_t)heads * D, T, q_pitch, BS)) return -2;
  if (make_map_box(&kdo, dout, (uint64_t)heads * D, T, (uint64_t)heads * D, BS)) return -2;
  if (make_map_box(&kk, k, (uint64_t)kv_heads * D, T, k_pitch, 128)) return -2;
  if (make_map_box(&kv_, v, (uint64_t)kv_heads * D, T, v_pitch, 128)) return -2;
  if (make_map_box(&qq, q, (uint64_t)heads * D, T, q_pitch, 128)) return -2;
  if (make_map_box(&qdo, dout, (uint64_t)heads * D, T, (uint64_t)heads * D, 128)) return -2;
  if (make_map_box(&qk, k, (uint64_t)kv_heads * D, T, k_pitch, BS)) return -2;
  if (make_map_box(&qv, v, (uint64_t)kv_heads * D, T, v_pitch, BS)) return -2;
  const int tiles = (T + 127) / 128;
  const int n_dkdv = tiles * kv_heads, n_dq = tiles * heads;
  if (separate) {
    attention_bwd_kernel<kHalf><<<n_dkdv, 384, kBwdSmemBytes, stream>>>(kq, kdo, kk, kv_, qq, qdo, qk, qv, p, 0);
    attention_bwd_kernel<kHalf><<<n_dq, 384, kBwdSmemBytes, stream>>>(kq, kdo, kk, kv_, qq, qdo, qk, qv, p, n_dkdv);
  } else {
    attention_bwd_kernel<kHalf><<<n_dkdv + n_dq, 384, kBwdSmemBytes, stream>>>(kq, kdo, kk, kv_, qq, qdo, qk, qv, p, 0);
  }
  return (int)cudaGetLastError();
}
extern "C" int sb_attention_bwd_ex(const void* dout, const void* q, const void* k, const void* v, const void* out,
                                   const float* lse, float* scratch, void* dq, void* dk, void* dv, const int* seg_start,
                                   const int* seg_end, const int* key_qend, int T, int heads, int kv_heads,
                                   long long q_pitch, long long k_pitch, long long v_pitch, float scale, int causal,
                                   float dropout_p, unsigned long long dropout_seed, int head_offset, cudaStream_t stream) {
  return attention_bwd_impl<false>(dout, q, k, v, out, lse, scratch, dq, dk, dv, seg_start, seg_end, key_qend, T, heads,
                                   kv_heads, q_pitch, k_pitch, v_pitch, scale, causal, dropout_p, dropout_seed, head_offset,
                                   stream);
}
extern "C" int sb_attention_bwd_f16(const void* dout, const void* q, const void* k, const void* v, const void* out,
                                    const float* lse, float* scratch, void* dq, void* dk, void* dv, const int* seg_start,
                                    const int* seg_end, const int* key_qend, int T, int heads, int kv_heads,
                                    long long q_pitch, long long k_pitch, long long v_pitch, float scale, int causal,
                                    float dropout_p, unsigned long long dropout_seed, int head_offset, cudaStream_t stream) {
  return attention_bwd_impl<true>(dout, q, k, v, out, lse, scratch, dq, dk, dv, seg_start, seg_end, key_qend, T, heads,
                                  kv_heads, q_pitch, k_pitch, v_pitch, scale, causal, dropout_p, dropout_seed, head_offset,
                                  stream);
}
