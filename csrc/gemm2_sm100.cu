// 2-CTA (cta_group::2) variant of the tcgen05 GEMM family: a CTA pair on one TPC computes a 256x256 output tile with
// UMMA 256x256x16.  Each CTA stages its own 128 rows of A and only half (128 rows) of the B tile, so the per-SM
// shared-memory traffic per FLOP is two thirds of the single-CTA kernel's — which is what bounds that kernel
// (A 4 KiB + B 8 KiB per 128-cycle MMA step plus the same again written by TMA exceeds 128 B/clk).
//
//   both CTAs : warp 0 TMA producer (own A half + own B half; complete_tx lands on the leader's mbarrier),
//               warps 4..7 epilogue over the CTA's own 128 TMEM lanes, warp 2 TMEM allocation (cta_group::2)
//   leader    : warp 1 issues tcgen05.mma.cta_group::2 and multicast-commits to both CTAs' barriers
//
// Same operand layouts and epilogues as gemm_sm100.cu (reference call sites: core/nn/linear/*.py, core/nn/mlp.py).
#include <cstdio>
#include <cstdlib>

#include "common.cuh"

namespace sb {
namespace gemm2 {

constexpr int BM = 256;       // per cluster; 128 per CTA
constexpr int BN = 256;       // per cluster; each CTA stages 128 B rows
constexpr int BK = 64;
constexpr int UK = 16;
constexpr int kStages = 6;
constexpr int kAccStages = 2;
constexpr int kThreads = 256;
constexpr uint32_t kABytes = 128 * BK * 2;  // 16 KiB per CTA
constexpr uint32_t kBBytes = 128 * BK * 2;  // 16 KiB per CTA
constexpr uint32_t kStageBytes = kABytes + kBBytes;
constexpr uint32_t kSmemBytes = kStages * kStageBytes + 1024 + 256;
constexpr uint32_t kPeerMask = 0xFEFFFFFF;  // clears the CTA-rank bit of a shared::cluster address -> leader's copy

enum Epilogue : int { kStore = 0, kAccumulate = 1, kSwiGLU = 2, kAllReduce = 3, kXentStats = 4, kAct = 5, kRope = 6 };

// ---- fused all-reduce (kAllReduce) ----------------------------------------------------------------------------
// One launch computes this rank's partial product AND the tensor-parallel sum of it.  Every rank owns a symmetric
// workspace (peer-mapped over NVLink): per sender a slot of bf16 partial tiles plus one u32 flag per (CTA tile,
// epilogue warp).  The epilogue warps convert a 32x64 sub-tile at a time, stage it in shared memory and push it with
// bulk async copies into the slot `tp_rank` of EVERY rank's workspace (its own included), then publish the flag with a
// system-scope release store once the copies of that tile have completed.  Four extra "reducer" warps per CTA wait for
// the flags of all senders for the tiles this CTA computed, add the partials in rank order in fp32 (so every rank
// produces identical bits) and write the bf16 result (+ bias) to D.  The NVLink transfer of tile i overlaps the MMA of
// tiles i+1.., the reduction of tile i overlaps both.
constexpr int kMaxTP = 8;
constexpr int kArThreads = 512;                          // 4 control + 4 epilogue + 8 reducer warps
constexpr int kArStages = 6;                             // same TMA ring depth as the plain kernel
constexpr uint32_t kChunkBytes = 32 * 64 * 2;            // one sub-tile: 32 rows x 64 bf16
constexpr uint32_t kSlabBytes = 4 * kChunkBytes;         // one epilogue warp's 32 x 256 slab
constexpr uint32_t kCtaTileBytes = 4 * kSlabBytes;       // 128 x 256 bf16
constexpr uint32_t kHalfSlabBytes = 2 * kChunkBytes;     // 32 rows x 128 columns: what one epilogue warp stages at a time
constexpr uint32_t kStagingBytes = 4 * kHalfSlabBytes;   // 32 KiB per CTA (a full slab per warp would cost a TMA stage)
constexpr uint32_t kArSmemBytes = kArStages * (128 * BK * 2 * 2) + 256 + kStagingBytes + 1024;

struct Params {
  int M, N, K;
  void* D;
  int ldd;
  const void* bias;
  void* aux;
  int ldaux;
  int swiglu_up_row_offset;
  int tiles_m, tiles_n;
  int kb_extra;  // k-blocks of the low-rank slab behind the main K range (0 = none)
  // all-gather -> GEMM: rows of A arrive from tensor-parallel peers (NVLink stores into this GPU's gather
  // buffer); gate[tm] counts the arrived row chunks of 256-row tile tm.  The TMA producer starts a tile only once
  // gate[tm] has reached gate_target, so tiles of the local shard start at once and remote tiles as they land.
  const uint32_t* gate;
  uint32_t gate_target;
  // kRope: rotary embedding (interleaved pairs) on the first rope_cols output columns, in the accumulator registers
  const float* rope_cos;  // [positions, rope_rot / 2]
  const float* rope_sin;
  const int* rope_pos;    // [M] position of every row, or null: row % rope_seq
  int rope_seq, rope_cols, rope_head_dim, rope_rot;
  // split-K tail: the last, partial wave of tiles is cut along K into `split` parts so that it fills the machine
  int full_tiles;   // tiles [0, full_tiles) run whole; the remaining `num_tiles - full_tiles` are split
  int split;        // parts per tail tile (1 = no splitting)
  float* split_ws;  // parked partial accumulators, 128x256 fp32 per (slot, cta half); slot: see part_slot()
  unsigned int* split_counters;  // [counter index][cta half] arrivals; the last arriver reduces and resets
  // stream-K tail (stream_k = 1): tiles [0, sk_first_tile) run whole, round-robin over the clusters in waves (neighbouring
  // tiles in flight together: the A / B panels stay in L2); the flattened (tile-major) k-block space of the remaining
  // one-to-two waves of tiles is cut into one contiguous range per cluster, so every cluster does the same amount of MMA
  // work and there is no partial last wave.  Tiles that straddle a range boundary are computed in parts that meet in
  // split_ws exactly like the split tail.  (Cutting ALL tiles that way was measured 1.5x slower: 74 clusters streaming 74
  // distant regions of the tile space lose the L2 reuse of the operand panels.)
  int stream_k;
  int sk_first_tile;
  // kAllReduce only
  int tp_rank, tp_size;
  uint32_t epoch;
  // reduce-scatter mode (rs_seq > 0): rows are [batch, sequence]; rank r owns sequence positions
  // [r·rs_seq_per_rank, (r+1)·rs_seq_per_rank) of every batch item (Megatron sequence parallelism).  A 256-row tile lies
  // inside one rank's range; it is pushed to and reduced by that owner only, and D is the owner's compact [b·s/tp, N].
  int rs_seq, rs_seq_per_rank;
  // two-shot all-reduce (tp > 2): tile t is reduced by rank t % tp only — partials travel to that owner, the owner's
  // reducers store the final bf16 tile into slot `owner` of every other rank's workspace (unused there: the owner never
  // sends a partial of its own tiles) and the other ranks copy it out.  (tp - 1) / tp of the one-shot bytes per phase.
  int two_shot;
  unsigned long long* prof;       // optional [grid][8] globaltimer stamps of the phases (sb_gemm2_set_profile_buffer)
  uint32_t flag_slot;             // u32 flags per sender slot
  unsigned long long slot_bytes;  // bytes per sender slot
  uint8_t* ar_data[kMaxTP];       // rank r's workspace (this call's parity) as mapped into this process
  uint32_t* ar_flags[kMaxTP];
};

struct __align__(8) Barriers {
  uint64_t full[kStages];
  uint64_t empty[kStages];
  uint64_t acc_full[kAccStages];
  uint64_t acc_empty[kAccStages];
  uint32_t tmem_base;
  uint32_t split_last;  // split-K: 1 when this CTA is the last part of its tile to arrive
  uint32_t pushed;  // kAllReduce: epilogue-warp tile pushes completed (4 per CTA tile), read by the signal warp
};

SB_DEVICE uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
SB_DEVICE void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
SB_DEVICE void tma_load_2d_2sm(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], "
      "[%2];\n" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerMask), "r"(c0), "r"(c1)
      : "memory");
}
SB_DEVICE void umma_f16_ss_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
SB_DEVICE void umma_commit_2sm(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n" ::"r"(
          smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}
SB_DEVICE void mbar_arrive_remote(uint64_t* bar, uint32_t cta) {
  asm volatile(
      "{\n\t"
      ".reg .b32 rem;\n\t"
      "mapa.shared::cluster.u32 rem, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [rem];\n\t"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(cta)
      : "memory");
}
template <uint32_t kCols>
SB_DEVICE void tmem_alloc_2sm(uint32_t* smem_result) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(smem_result)), "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory");
}
template <uint32_t kCols>
SB_DEVICE void tmem_dealloc_2sm(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "n"(kCols) : "memory");
}

SB_DEVICE void bulk_store_s2g(void* gdst, const void* ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes)
               : "memory");
}
SB_DEVICE void bulk_commit() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
template <int kN>
SB_DEVICE void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;\n" ::"n"(kN) : "memory");
}
template <int kN>
SB_DEVICE void bulk_wait() {
  asm volatile("cp.async.bulk.wait_group %0;\n" ::"n"(kN) : "memory");
}
SB_DEVICE void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
SB_DEVICE void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
SB_DEVICE void red_release_cta_smem_inc(uint32_t* p) {
  asm volatile("red.release.cta.shared::cta.add.u32 [%0], 1;\n" ::"r"(smem_u32(p)) : "memory");
}
SB_DEVICE uint32_t ld_acquire_cta_smem(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];\n" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}
SB_DEVICE void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;\n" ::: "memory"); }
SB_DEVICE uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
SB_DEVICE unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
  return t;
}
// Phase stamp for tools/tpbench.py --profile: slot 0 kernel entry, 1 last accumulator ready, 2 last push complete,
// 3 last flag published, 4 first / 5 last tile's flags seen by the reducer, 6 reducer done.
SB_DEVICE void prof_stamp(unsigned long long* prof, int slot) {
  if (prof) prof[(size_t)blockIdx.x * 8 + slot] = global_timer_ns();
}
// Spin until *flag == epoch.  A peer that never arrives (crashed rank) traps after 30 s instead of hanging the GPU.
SB_DEVICE void wait_flag(const uint32_t* flag, uint32_t epoch) {
  if (ld_acquire_sys(flag) == epoch) return;
  const unsigned long long t0 = global_timer_ns();
  int spins = 0;
  while (ld_acquire_sys(flag) != epoch) {
    if (++spins > 64) {
      __nanosleep(100);
      if ((spins & 1023) == 0 && global_timer_ns() - t0 > 30000000000ull) __trap();
    }
  }
}
SB_DEVICE uint4 ld_cg_v4(const void* p) {
  uint4 v;
  asm volatile("ld.global.cg.v4.u32 {%0, %1, %2, %3}, [%4];\n" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
SB_DEVICE void acc_bf16x8(float* f, const uint4& x) {
  float2 a = unpack_bf16(x.x), b = unpack_bf16(x.y), c = unpack_bf16(x.z), d = unpack_bf16(x.w);
  f[0] += a.x, f[1] += a.y, f[2] += b.x, f[3] += b.y, f[4] += c.x, f[5] += c.y, f[6] += d.x, f[7] += d.y;
}

SB_DEVICE void tile_coords(int t, int tiles_m, int tiles_n, int& tm, int& tn) {
  constexpr int G = 4;
  int per_group = G * tiles_n;
  int g = t / per_group;
  int first_m = g * G;
  int gsize = min(G, tiles_m - first_m);
  int r = t - g * per_group;
  tm = first_m + r % gsize;
  tn = r / gsize;
}

// Reduce-scatter mode: which rank owns the tile whose first row is `row0`, and where that row lands in the owner's D.
SB_DEVICE int rs_owner(const Params& p, int row0) { return (row0 % p.rs_seq) / p.rs_seq_per_rank; }
SB_DEVICE int rs_out_row(const Params& p, int row) {
  const int b = row / p.rs_seq, s = row % p.rs_seq;
  return b * p.rs_seq_per_rank + s % p.rs_seq_per_rank;
}

// Which rank reduces tile t (first row tm * BM): the sequence owner in reduce-scatter mode, t % tp in two-shot mode,
// -1 = every rank (one-shot all-reduce).
SB_DEVICE int tile_owner(const Params& p, int t, int tm) {
  if (p.rs_seq > 0) return rs_owner(p, tm * BM);
  if (p.two_shot) return t % p.tp_size;
  return -1;
}

// Work unit of the persistent schedule: (tile, k-block range, part index or -1 when the tile is not split, parts of
// the tile).
struct Unit {
  int tile, kb0, kb1, part, nparts;
};
SB_DEVICE Unit get_unit(int u, const Params& p, int k_blocks) {
  Unit w;
  if (u < p.full_tiles) {
    w.tile = u, w.kb0 = 0, w.kb1 = k_blocks, w.part = -1, w.nparts = 1;
  } else {
    const int r = u - p.full_tiles;
    w.tile = p.full_tiles + r / p.split;
    w.part = r % p.split;
    w.nparts = p.split;
    w.kb0 = (int)((long long)w.part * k_blocks / p.split);
    w.kb1 = (int)((long long)(w.part + 1) * k_blocks / p.split);
  }
  return w;
}
// stream-K bookkeeping over the flattened k-block index x in [0, total): cluster c owns [sk_lo(c), sk_lo(c + 1))
SB_DEVICE long long sk_lo(int c, long long total, int nc) { return (long long)c * total / nc; }
SB_DEVICE int sk_cluster_of(long long x, long long total, int nc) { return (int)(((x + 1) * nc + total - 1) / total) - 1; }
// The units of one cluster, for all three roles (producer, MMA issuer, epilogue) to iterate identically.
struct Sched {
  int stream_k, cluster_id, nc, k_blocks, n;
  int n_dp;      // stream-K: whole tiles of this cluster before its stream-K range
  int f_tiles;   // stream-K: first tile of the stream-K region
  long long total, lo, hi;  // flattened k-block space of the stream-K region and this cluster's range in it
  int t_first;   // first tile (relative to f_tiles) of this cluster's range
};
SB_DEVICE Sched make_sched(const Params& p, int cluster_id, int nc, int k_blocks, int num_units) {
  Sched s;
  s.stream_k = p.stream_k, s.cluster_id = cluster_id, s.nc = nc, s.k_blocks = k_blocks;
  s.n_dp = 0, s.f_tiles = 0;
  if (p.stream_k) {
    s.f_tiles = p.sk_first_tile;
    s.n_dp = s.f_tiles / nc;  // sk_first_tile is a multiple of the cluster count
    s.total = (long long)(p.tiles_m * p.tiles_n - s.f_tiles) * k_blocks;
    s.lo = sk_lo(cluster_id, s.total, nc), s.hi = sk_lo(cluster_id + 1, s.total, nc);
    s.t_first = (int)(s.lo / k_blocks);
    s.n = s.n_dp + (s.hi > s.lo ? (int)((s.hi - 1) / k_blocks) - s.t_first + 1 : 0);
  } else {
    s.total = 0, s.lo = s.hi = 0, s.t_first = 0;
    s.n = cluster_id < num_units ? (num_units - cluster_id + nc - 1) / nc : 0;
  }
  return s;
}
SB_DEVICE Unit sched_unit(const Sched& s, const Params& p, int i) {
  if (!s.stream_k) return get_unit(s.cluster_id + i * s.nc, p, s.k_blocks);
  Unit w;
  if (i < s.n_dp) {
    w.tile = s.cluster_id + i * s.nc, w.kb0 = 0, w.kb1 = s.k_blocks, w.part = -1, w.nparts = 1;
    return w;
  }
  const int rel = s.t_first + (i - s.n_dp);  // tile index inside the stream-K region
  w.tile = s.f_tiles + rel;
  const long long t0 = (long long)rel * s.k_blocks, t1 = t0 + s.k_blocks;
  const long long g0 = s.lo > t0 ? s.lo : t0, g1 = s.hi < t1 ? s.hi : t1;
  w.kb0 = (int)(g0 - t0), w.kb1 = (int)(g1 - t0);
  if (w.kb0 == 0 && w.kb1 == s.k_blocks) {
    w.part = -1, w.nparts = 1;
  } else {
    const int c_first = sk_cluster_of(t0, s.total, s.nc), c_last = sk_cluster_of(t1 - 1, s.total, s.nc);
    w.part = s.cluster_id - c_first, w.nparts = c_last - c_first + 1;
  }
  return w;
}
// Where part q of a split tile parks its accumulator, and which arrival counter the tile uses.
SB_DEVICE int part_slot(const Sched& s, const Params& p, int tile, int q) {
  if (!s.stream_k) return (tile - p.full_tiles) * p.split + q;
  // stream-K: a cluster parks at most two partials per launch — the unit that begins its range (slot 0) and the one
  // that ends it (slot 1)
  const long long t0 = (long long)(tile - s.f_tiles) * s.k_blocks;
  const int c = sk_cluster_of(t0, s.total, s.nc) + q;
  const long long lo = sk_lo(c, s.total, s.nc);
  return c * 2 + ((lo > t0 ? lo : t0) == lo ? 0 : 1);
}
SB_DEVICE int part_counter(const Sched& s, const Params& p, int tile) { return s.stream_k ? tile : tile - p.full_tiles; }
constexpr int kCtaAccFloats = 128 * 256;
// Partial accumulators are stored "float4-column-major": [group of 4 columns][row][4] so that a warp (lane = row) writes
// and later reads 512 contiguous bytes per instruction.
SB_DEVICE float4* split_slot(const Params& p, int slot, uint32_t half) {
  return reinterpret_cast<float4*>(p.split_ws + ((size_t)slot * 2 + half) * kCtaAccFloats);
}
// v[0..n) += the other parts' partials of columns [c, c+n) of this thread's row, in part order (own registers take the
// place of part `own`), so the sum is the same whichever part arrives last.
template <int kN>
SB_DEVICE void split_reduce_chunk(uint32_t* v, const Params& p, const Sched& sc, const Unit& w, uint32_t half, int row_in_cta,
                                   int c) {
  float acc[kN];
  bool first = true;
  const int own = w.part;
  for (int q = 0; q < w.nparts; ++q) {
    if (q == own) {
      if (first) {
#pragma unroll
        for (int i = 0; i < kN; ++i) acc[i] = __uint_as_float(v[i]);
      } else {
#pragma unroll
        for (int i = 0; i < kN; ++i) acc[i] += __uint_as_float(v[i]);
      }
    } else {
      const float4* src = split_slot(p, part_slot(sc, p, w.tile, q), half) + (size_t)(c >> 2) * 128 + row_in_cta;
#pragma unroll
      for (int g = 0; g < kN / 4; ++g) {
        const float4 x = __ldcg(src + (size_t)g * 128);
        if (first) {
          acc[4 * g] = x.x, acc[4 * g + 1] = x.y, acc[4 * g + 2] = x.z, acc[4 * g + 3] = x.w;
        } else {
          acc[4 * g] += x.x, acc[4 * g + 1] += x.y, acc[4 * g + 2] += x.z, acc[4 * g + 3] += x.w;
        }
      }
    }
    first = false;
  }
#pragma unroll
  for (int i = 0; i < kN; ++i) v[i] = __float_as_uint(acc[i]);
}

// kHalf: the 16-bit operands (and a 16-bit OutT) are fp16 instead of bf16 — plain store / accumulate epilogues only.
template <int kAMajorMN, int kBMajorMN, int kEpi, typename OutT, bool kHalf = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kEpi == kAllReduce ? kArThreads : kThreads, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
             const __grid_constant__ CUtensorMap tma_a2, const __grid_constant__ CUtensorMap tma_b2, const Params p) {
  constexpr int kNumStages = kEpi == kAllReduce ? kArStages : kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  Barriers* bars = reinterpret_cast<Barriers*>(smem + kNumStages * kStageBytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int num_tiles = p.tiles_m * p.tiles_n;
  // k-blocks [0, kb_main) stream the main operands, [kb_main, k_blocks) the optional low-rank slab (A2 [M, K2] K-major,
  // B2 in the layout of B): D = A·Bᵀ + A2·B2ᵀ in one accumulator — LoRA as extra K
  const int kb_main = (p.K + BK - 1) / BK;
  const int k_blocks = kb_main + p.kb_extra;
  const int num_units = p.full_tiles + (num_tiles - p.full_tiles) * p.split;
  const Sched sched = make_sched(p, blockIdx.x >> 1, gridDim.x >> 1, k_blocks, num_units);
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    if (p.kb_extra > 0) {
      tma_prefetch_desc(&tma_a2);
      tma_prefetch_desc(&tma_b2);
    }
    if (kEpi == kAllReduce) prof_stamp(p.prof, 0);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kNumStages; ++i) {
      mbar_init(&bars->full[i], 1);
      mbar_init(&bars->empty[i], 1);
    }
    for (int i = 0; i < kAccStages; ++i) {
      mbar_init(&bars->acc_full[i], 1);
      mbar_init(&bars->acc_empty[i], 8);  // 4 epilogue warps in each of the two CTAs arrive on the leader's barrier
    }
    bars->pushed = 0;
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_2sm<512>(&bars->tmem_base);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int ui = 0; ui < sched.n; ++ui) {
        const Unit w = sched_unit(sched, p, ui);
        int tm, tn;
        tile_coords(w.tile, p.tiles_m, p.tiles_n, tm, tn);
        if (p.gate) {
          // acquire the peers' (generic-proxy) stores of this row tile, then order them before the async-proxy TMA reads
          const uint32_t* g = p.gate + tm;
          const unsigned long long t0 = global_timer_ns();
          int spins = 0;
          while ((int)(ld_acquire_sys(g) - p.gate_target) < 0) {
            if (++spins > 16) {
              __nanosleep(64);
              if ((spins & 1023) == 0 && global_timer_ns() - t0 > 30000000000ull) __trap();
            }
          }
          asm volatile("fence.proxy.async.global;\n" ::: "memory");
        }
        const int m0 = tm * BM + rank * 128;
        const int nb = (kEpi == kSwiGLU) ? tn * (BN / 2) + (rank ? p.swiglu_up_row_offset : 0) : tn * BN + rank * 128;
        for (int kb = w.kb0; kb < w.kb1; ++kb) {
          mbar_wait(&bars->empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          uint8_t* sb_ = sa + kABytes;
          if (leader) mbar_arrive_expect_tx(&bars->full[stage], 2 * kStageBytes);  // both CTAs' bytes land here
          const bool extra = kb >= kb_main;
          const CUtensorMap* ma = extra ? &tma_a2 : &tma_a;
          const CUtensorMap* mb = extra ? &tma_b2 : &tma_b;
          const int k0 = (extra ? kb - kb_main : kb) * BK;
          if (kAMajorMN) {
            tma_load_2d_2sm(sa, ma, &bars->full[stage], m0, k0);
            tma_load_2d_2sm(sa + 8192, ma, &bars->full[stage], m0 + 64, k0);
          } else {
            tma_load_2d_2sm(sa, ma, &bars->full[stage], k0, m0);
          }
          if (kBMajorMN) {
            tma_load_2d_2sm(sb_, mb, &bars->full[stage], nb, k0);
            tma_load_2d_2sm(sb_ + 8192, mb, &bars->full[stage], nb + 64, k0);
          } else {
            tma_load_2d_2sm(sb_, mb, &bars->full[stage], k0, nb);
          }
          if (++stage == kNumStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && lane == 0) {
      static_assert(!kHalf || kEpi == kStore || kEpi == kAccumulate, "fp16 exists for the plain epilogues");
      constexpr uint32_t idesc = make_idesc_16(BM, BN, kAMajorMN, kBMajorMN, kHalf);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int ui = 0; ui < sched.n; ++ui) {
        const Unit w = sched_unit(sched, p, ui);
        mbar_wait(&bars->acc_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = w.kb0; kb < w.kb1; ++kb) {
          mbar_wait(&bars->full[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * kStageBytes);
          const uint32_t sb_ = sa + kABytes;
#pragma unroll
          for (int k = 0; k < BK / UK; ++k) {
            const uint64_t da = kAMajorMN ? make_smem_desc_sw128(sa + k * 2048, 8192, 1024)
                                          : make_smem_desc_sw128(sa + k * 32, 16, 1024);
            const uint64_t db = kBMajorMN ? make_smem_desc_sw128(sb_ + k * 2048, 8192, 1024)
                                          : make_smem_desc_sw128(sb_ + k * 32, 16, 1024);
            umma_f16_ss_2sm(tmem_d, da, db, idesc, ((kb - w.kb0) | k) != 0);
          }
          umma_commit_2sm(&bars->empty[stage]);
          if (++stage == kNumStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_2sm(&bars->acc_full[acc]);
        if (++acc == kAccStages) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (kEpi == kAllReduce && warp == 3) {
    // ===================== signal warp (kAllReduce): publish "tile pushed" flags off the epilogue's critical path ======
    if (lane == 0) {
      uint32_t done = 0;
      for (int t = cluster_id; t < num_tiles; t += num_clusters) {
        const uint32_t ct = (uint32_t)t * 2 + rank;
        done += 4;
        while (ld_acquire_cta_smem(&bars->pushed) < done) __nanosleep(40);
        fence_acq_rel_sys();  // the four warps' bulk copies (completed, observed through `pushed`) before the flags
        int tm_, tn_;
        tile_coords(t, p.tiles_m, p.tiles_n, tm_, tn_);
        const int owner = tile_owner(p, t, tm_);
        if (owner >= 0) {
          st_relaxed_sys(p.ar_flags[owner] + (size_t)p.tp_rank * p.flag_slot + ct, p.epoch);
        } else {
          for (int r = 0; r < p.tp_size; ++r)
            st_relaxed_sys(p.ar_flags[r] + (size_t)p.tp_rank * p.flag_slot + ct, p.epoch);
        }
      }
      prof_stamp(p.prof, 3);
    }
  } else if (kEpi == kAllReduce && warp >= 8) {
    // ===================== reducer warps (kAllReduce): sum the senders' partials of this CTA's tiles =====================
    // eight warps: warp pair (2q, 2q+1) shares epilogue warp q's 32 x 256 slab, 128 columns each
    const int rw = warp - 8;
    const int slab = rw >> 1, half = rw & 1;
    const uint8_t* mine = p.ar_data[p.tp_rank];
    const uint32_t* myflags = p.ar_flags[p.tp_rank];
    const __nv_bfloat16* bias = reinterpret_cast<const __nv_bfloat16*>(p.bias);
    for (int t = cluster_id; t < num_tiles; t += num_clusters) {
      int tm, tn;
      tile_coords(t, p.tiles_m, p.tiles_n, tm, tn);
      const uint32_t ct = (uint32_t)t * 2 + rank;
      const int owner = tile_owner(p, t, tm);
      if (p.rs_seq > 0 && owner != p.tp_rank) continue;  // another rank reduces this tile and keeps the result
      // two-shot, somebody else's tile: wait for the owner's FINAL tile in slot `owner` and copy it out
      const bool copy_only = p.two_shot && owner != p.tp_rank;
      const bool spread = p.two_shot && owner == p.tp_rank;  // owner: also store the result into the peers' slots
      const int src0 = copy_only ? owner : 0, nsrc = copy_only ? 1 : p.tp_size;
      if (lane == 0) {
        for (int s = 0; s < nsrc; ++s) wait_flag(myflags + (size_t)(src0 + s) * p.flag_slot + ct, p.epoch);
        if (rw == 0) {
          if (t == cluster_id) prof_stamp(p.prof, 4);
          if (t + num_clusters >= num_tiles) prof_stamp(p.prof, 5);
        }
      }
      __syncwarp();
      const size_t slab_off = ((size_t)ct * 4 + slab) * kSlabBytes;
      const uint8_t* base = mine + (size_t)src0 * p.slot_bytes + slab_off;
      const int row0 = tm * BM + rank * 128 + slab * 32;
      const int n0 = tn * BN;
#pragma unroll 1
      for (int c4 = 2 * half; c4 < 2 * half + 2; ++c4) {
        uint4 xa[8], xb[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const size_t off = (size_t)c4 * kChunkBytes + (size_t)(lane + 32 * it) * 16;
          xa[it] = ld_cg_v4(base + off);
          xb[it] = nsrc > 1 ? ld_cg_v4(base + p.slot_bytes + off) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          float f[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
          acc_bf16x8(f, xa[it]);
          acc_bf16x8(f, xb[it]);
          const size_t off = (size_t)c4 * kChunkBytes + (size_t)(lane + 32 * it) * 16;
          for (int s = 2; s < nsrc; ++s) acc_bf16x8(f, ld_cg_v4(base + (size_t)s * p.slot_bytes + off));
          const int r_in = (lane >> 3) + 4 * it;
          const int j = (lane & 7) ^ (r_in & 7);
          const int row = row0 + r_in;
          const int col = n0 + c4 * 64 + j * 8;
          if (row >= p.M || col >= p.N) continue;
          const int drow = p.rs_seq > 0 ? rs_out_row(p, row) : row;
          __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.D) + (size_t)drow * p.ldd + col;
          const bool add_bias = bias && !copy_only;  // a copied tile already carries the bias
          if (col + 8 <= p.N) {
            if (add_bias) {
              const uint4 bv = *reinterpret_cast<const uint4*>(bias + col);
              acc_bf16x8(f, bv);
            }
            const uint4 o =
                make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
            *reinterpret_cast<uint4*>(dst) = o;
            if (spread) {
              for (int r = 0; r < p.tp_size; ++r)
                if (r != p.tp_rank)
                  *reinterpret_cast<uint4*>(p.ar_data[r] + (size_t)p.tp_rank * p.slot_bytes + slab_off + off) = o;
            }
          } else {
            for (int i = 0; i < 8 && col + i < p.N; ++i)
              dst[i] = __float2bfloat16(f[i] + (add_bias ? __bfloat162float(bias[col + i]) : 0.f));
          }
        }
      }
      if (spread) {
        // all eight reducer warps have stored the final tile into the peers' slots: publish "final ready"
        asm volatile("bar.sync 2, 256;\n" ::: "memory");
        if (rw == 0 && lane == 0) {
          fence_acq_rel_sys();
          for (int r = 0; r < p.tp_size; ++r)
            if (r != p.tp_rank) st_relaxed_sys(p.ar_flags[r] + (size_t)p.tp_rank * p.flag_slot + ct, p.epoch);
        }
      }
    }
    if (rw == 0 && lane == 0) prof_stamp(p.prof, 6);
  } else if (kEpi == kAllReduce && warp >= 4) {
    // ===================== epilogue (kAllReduce): push bf16 partial tiles into every rank's workspace =====================
    const int ew = warp & 3;
    uint8_t* half_buf = smem + kNumStages * kStageBytes + 256 + (uint32_t)ew * kHalfSlabBytes;
    int acc = 0;
    uint32_t acc_phase = 0;
    int pending_ct = -1;
    for (int t = cluster_id; t < num_tiles; t += num_clusters) {
      const uint32_t ct = (uint32_t)t * 2 + rank;
      mbar_wait(&bars->acc_full[acc], acc_phase);
      tc_fence_after();
      if (ew == 0 && lane == 0 && t + num_clusters >= num_tiles) prof_stamp(p.prof, 1);
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(ew * 32) << 16);
      const size_t slab_off = (size_t)p.tp_rank * p.slot_bytes + ((size_t)ct * 4 + ew) * kSlabBytes;
      int push_owner;
      {
        int tm_, tn_;
        tile_coords(t, p.tiles_m, p.tiles_n, tm_, tn_);
        push_owner = tile_owner(p, t, tm_);
      }
      // the slab travels as two 8 KiB halves (two 64-column chunks each) through one staging buffer per warp
#pragma unroll 1
      for (int hs = 0; hs < 2; ++hs) {
        if (lane == 0) bulk_wait_read<0>();  // the previous half's copies have drained the staging buffer
        __syncwarp();
#pragma unroll 1
        for (int cc = 0; cc < 2; ++cc) {
          const int c4 = 2 * hs + cc;
          uint8_t* buf = half_buf + cc * kChunkBytes;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            uint32_t v[32];
            tmem_ld_32x32b_x32(taddr + c4 * 64 + h * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint4 o = make_uint4(pack_bf16(__uint_as_float(v[8 * i]), __uint_as_float(v[8 * i + 1])),
                                         pack_bf16(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3])),
                                         pack_bf16(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5])),
                                         pack_bf16(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7])));
              const int pos = (h * 4 + i) ^ (lane & 7);  // 16-byte unit swizzle: conflict-free staging stores
              *reinterpret_cast<uint4*>(buf + lane * 128 + pos * 16) = o;
            }
          }
        }
        if (hs == 1) {  // accumulator fully read: hand the TMEM stage back before the last pushes
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_remote(&bars->acc_empty[acc], 0);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          const size_t off = slab_off + (size_t)hs * kHalfSlabBytes;
          if (push_owner >= 0) {
            bulk_store_s2g(p.ar_data[push_owner] + off, half_buf, kHalfSlabBytes);
          } else {
            for (int r = 0; r < p.tp_size; ++r) bulk_store_s2g(p.ar_data[r] + off, half_buf, kHalfSlabBytes);
          }
          bulk_commit();
        }
      }
      if (lane == 0) {
        if (pending_ct >= 0) {  // the previous tile's two groups have completed: hand it to the signal warp
          bulk_wait<2>();
          red_release_cta_smem_inc(&bars->pushed);
        }
        pending_ct = (int)ct;
      }
      if (++acc == kAccStages) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (lane == 0 && pending_ct >= 0) {
      bulk_wait<0>();
      red_release_cta_smem_inc(&bars->pushed);
      if (ew == 0) prof_stamp(p.prof, 2);
    }
  } else if (kEpi != kAllReduce && warp >= 4) {
    // ===================== epilogue (both CTAs, own 128 rows) =====================
    const int ew = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int ui = 0; ui < sched.n; ++ui) {
      const Unit w = sched_unit(sched, p, ui);
      int tm, tn;
      tile_coords(w.tile, p.tiles_m, p.tiles_n, tm, tn);
      const int row_in_cta = ew * 32 + lane;
      const int row = tm * BM + rank * 128 + row_in_cta;
      mbar_wait(&bars->acc_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(ew * 32) << 16);
      const bool row_ok = row < p.M;
      const bool split = w.part >= 0;
      if (split) {
        // park this part's accumulator, then find out whether the other parts are already in
        float4* mine = split_slot(p, part_slot(sched, p, w.tile, w.part), rank) + row_in_cta;
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(taddr + c, v);
          tmem_ld_wait();
#pragma unroll
          for (int g = 0; g < 8; ++g)
            __stcg(mine + (size_t)((c >> 2) + g) * 128,
                   make_float4(__uint_as_float(v[4 * g]), __uint_as_float(v[4 * g + 1]), __uint_as_float(v[4 * g + 2]),
                               __uint_as_float(v[4 * g + 3])));
        }
        __threadfence();
        asm volatile("bar.sync 1, 128;\n" ::: "memory");
        if (ew == 0 && lane == 0) {
          unsigned int* ctr = p.split_counters + part_counter(sched, p, w.tile) * 2 + rank;
          const unsigned int old = atomicAdd(ctr, 1u);
          const bool last = old == (unsigned int)(w.nparts - 1);
          if (last) *ctr = 0;  // every part has arrived: leave the counter clean for the next launch
          bars->split_last = last ? 1u : 0u;
        }
        asm volatile("bar.sync 1, 128;\n" ::: "memory");
        const bool last = bars->split_last != 0;
        asm volatile("bar.sync 1, 128;\n" ::: "memory");  // flag consumed before the next tile may overwrite it
        if (!last) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_remote(&bars->acc_empty[acc], 0);
          if (++acc == kAccStages) {
            acc = 0;
            acc_phase ^= 1;
          }
          continue;
        }
        __threadfence();
      }

      if (kEpi == kXentStats) {
        // LM head: store the bf16 logits AND the cross-entropy statistics of this row over the tile's 256 columns
        // (max, Σexp(x - max), arg-max) while the accumulator is in registers — the forward never re-reads the logits
        // (SURVEY §8 G8).  Statistics are taken from the bf16-rounded values, i.e. from exactly what is stored.
        const int n0 = tn * BN;
        __nv_bfloat16* drow = reinterpret_cast<__nv_bfloat16*>(p.D) + (size_t)row * p.ldd;
        float m_run = -INFINITY, s_run = 0.f, best = -INFINITY;
        int best_i = 0;
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(taddr + c, v);
          tmem_ld_wait();
          const int col = n0 + c;
          if (!row_ok || col >= p.N) continue;
          const int nv = min(32, p.N - col);
          uint32_t packed[16];
          float cm = -INFINITY;
          int ci = 0;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            packed[i] = pack_bf16(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1]));
            const float2 f2 = unpack_bf16(packed[i]);
            const float f0 = (2 * i < nv) ? f2.x : -INFINITY, f1 = (2 * i + 1 < nv) ? f2.y : -INFINITY;
            v[2 * i] = __float_as_uint(f0), v[2 * i + 1] = __float_as_uint(f1);
            if (f0 > cm) cm = f0, ci = 2 * i;
            if (f1 > cm) cm = f1, ci = 2 * i + 1;
          }
          if (cm > best) best = cm, best_i = col + ci;
          const float m_new = fmaxf(m_run, cm);
          float sum = 0.f;
#pragma unroll
          for (int i = 0; i < 32; ++i) sum += __expf(__uint_as_float(v[i]) - m_new);
          s_run = s_run * __expf(m_run - m_new) + sum;
          m_run = m_new;
          if (nv == 32) {
            uint4* dst = reinterpret_cast<uint4*>(drow + col);
#pragma unroll
            for (int i = 0; i < 4; ++i) dst[i] = make_uint4(packed[4 * i], packed[4 * i + 1], packed[4 * i + 2], packed[4 * i + 3]);
          } else {
            for (int i = 0; i < nv; ++i) drow[col + i] = __float2bfloat16(__uint_as_float(v[i]));
          }
        }
        if (row_ok && n0 < p.N)
          reinterpret_cast<float4*>(p.aux)[(size_t)row * p.tiles_n + tn] =
              make_float4(m_run, s_run, 0.f, __int_as_float(best_i));
      } else if (kEpi == kRope) {
        // q|k|v projection with the rotary embedding applied to the q and k columns while the row is in registers
        // (fp32, one rounding) — the separate rotary pass over [T, q|k] disappears from the forward (SURVEY §8 A1;
        // reference core/nn/rotary.py:68-90 after core/nn/attention/attention.py:509-566).
        const int n0 = tn * BN;
        __nv_bfloat16* drow = reinterpret_cast<__nv_bfloat16*>(p.D) + (size_t)row * p.ldd;
        const int half = p.rope_rot / 2;
        const int rpos = row_ok ? (p.rope_pos ? __ldg(p.rope_pos + row) : row % p.rope_seq) : 0;
        const float* cr = p.rope_cos + (size_t)rpos * half;
        const float* sr = p.rope_sin + (size_t)rpos * half;
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(taddr + c, v);
          tmem_ld_wait();
          if (split) split_reduce_chunk<32>(v, p, sched, w, rank, row_in_cta, c);
          const int col = n0 + c;
          if (!row_ok || col >= p.N) continue;
          const int d0 = col % p.rope_head_dim;  // head_dim % 32 == 0: a chunk never straddles heads
          if (col < p.rope_cols && d0 < p.rope_rot) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {  // 4 groups of 4 pairs; rope_rot % 8 == 0 keeps a group inside or outside
              if (d0 + 8 * g >= p.rope_rot) break;
              const float4 c4 = __ldg(reinterpret_cast<const float4*>(cr + d0 / 2 + 4 * g));
              const float4 s4 = __ldg(reinterpret_cast<const float4*>(sr + d0 / 2 + 4 * g));
              const float cc[4] = {c4.x, c4.y, c4.z, c4.w}, ss[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const float a = __uint_as_float(v[8 * g + 2 * k]), b = __uint_as_float(v[8 * g + 2 * k + 1]);
                v[8 * g + 2 * k] = __float_as_uint(a * cc[k] - b * ss[k]);
                v[8 * g + 2 * k + 1] = __float_as_uint(a * ss[k] + b * cc[k]);
              }
            }
          }
          if (col + 32 <= p.N) {
            uint4* dst = reinterpret_cast<uint4*>(drow + col);
#pragma unroll
            for (int i = 0; i < 4; ++i)
              dst[i] = make_uint4(pack_bf16(__uint_as_float(v[8 * i]), __uint_as_float(v[8 * i + 1])),
                                  pack_bf16(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3])),
                                  pack_bf16(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5])),
                                  pack_bf16(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7])));
          } else {
            for (int i = 0; i < 32 && col + i < p.N; ++i) drow[col + i] = __float2bfloat16(__uint_as_float(v[i]));
          }
        }
      } else if (kEpi == kAct) {
        // adapter / GELU-MLP up-projection: D = act(A·Bᵀ + bias) with the activation on the accumulator
        // registers; the bf16 pre-activation goes to aux for the backward pass.  swiglu_up_row_offset carries the
        // activation kind (0 GELU, 1 SiLU) for this epilogue.
        const int n0 = tn * BN;
        __nv_bfloat16* drow = reinterpret_cast<__nv_bfloat16*>(p.D) + (size_t)row * p.ldd;
        __nv_bfloat16* arow = p.aux ? reinterpret_cast<__nv_bfloat16*>(p.aux) + (size_t)row * p.ldaux : nullptr;
        const __nv_bfloat16* bias = reinterpret_cast<const __nv_bfloat16*>(p.bias);
        const bool silu = p.swiglu_up_row_offset != 0;
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(taddr + c, v);
          tmem_ld_wait();
          if (split) split_reduce_chunk<32>(v, p, sched, w, rank, row_in_cta, c);
          const int col = n0 + c;
          if (!row_ok || col >= p.N) continue;
          const int nv = min(32, p.N - col);
          uint32_t pre[16], act[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float x0 = __uint_as_float(v[2 * i]), x1 = __uint_as_float(v[2 * i + 1]);
            if (bias) {
              if (2 * i < nv) x0 += __bfloat162float(bias[col + 2 * i]);
              if (2 * i + 1 < nv) x1 += __bfloat162float(bias[col + 2 * i + 1]);
            }
            pre[i] = pack_bf16(x0, x1);
            const float2 r = unpack_bf16(pre[i]);  // the activation sees what the backward pass will see
            act[i] = silu ? pack_bf16(silu_f(r.x), silu_f(r.y)) : pack_bf16(gelu_f(r.x), gelu_f(r.y));
          }
          if (nv == 32) {
            uint4* dst = reinterpret_cast<uint4*>(drow + col);
#pragma unroll
            for (int i = 0; i < 4; ++i) dst[i] = make_uint4(act[4 * i], act[4 * i + 1], act[4 * i + 2], act[4 * i + 3]);
            if (arow) {
              uint4* ap = reinterpret_cast<uint4*>(arow + col);
#pragma unroll
              for (int i = 0; i < 4; ++i) ap[i] = make_uint4(pre[4 * i], pre[4 * i + 1], pre[4 * i + 2], pre[4 * i + 3]);
            }
          } else {
            for (int i = 0; i < nv; ++i) {
              const float2 a2 = unpack_bf16(act[i >> 1]), p2 = unpack_bf16(pre[i >> 1]);
              drow[col + i] = __float2bfloat16((i & 1) ? a2.y : a2.x);
              if (arow) arow[col + i] = __float2bfloat16((i & 1) ? p2.y : p2.x);
            }
          }
        }
      } else if (kEpi == kSwiGLU) {
        const int n0 = tn * (BN / 2);
        __nv_bfloat16* drow = reinterpret_cast<__nv_bfloat16*>(p.D) + (size_t)row * p.ldd;
        __nv_bfloat16* arow = p.aux ? reinterpret_cast<__nv_bfloat16*>(p.aux) + (size_t)row * p.ldaux : nullptr;
#pragma unroll 1
        for (int c = 0; c < BN / 2; c += 16) {
          uint32_t g[16], u[16];
          tmem_ld_32x32b_x16(taddr + c, g);
          tmem_ld_32x32b_x16(taddr + BN / 2 + c, u);
          tmem_ld_wait();
          if (split) {
            split_reduce_chunk<16>(g, p, sched, w, rank, row_in_cta, c);
            split_reduce_chunk<16>(u, p, sched, w, rank, row_in_cta, BN / 2 + c);
          }
          if (row_ok && n0 + c < p.N) {
            uint32_t o[8], pg[8], pu[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              pg[i] = pack_bf16(__uint_as_float(g[2 * i]), __uint_as_float(g[2 * i + 1]));
              pu[i] = pack_bf16(__uint_as_float(u[2 * i]), __uint_as_float(u[2 * i + 1]));
              float2 gr = unpack_bf16(pg[i]), ur = unpack_bf16(pu[i]);
              o[i] = pack_bf16(silu_f(gr.x) * ur.x, silu_f(gr.y) * ur.y);
            }
            if (n0 + c + 16 <= p.N) {
              uint4* dst = reinterpret_cast<uint4*>(drow + n0 + c);
              dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
              dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
              if (arow) {
                uint4* ag = reinterpret_cast<uint4*>(arow + n0 + c);
                uint4* au = reinterpret_cast<uint4*>(arow + p.swiglu_up_row_offset + n0 + c);
                ag[0] = make_uint4(pg[0], pg[1], pg[2], pg[3]);
                ag[1] = make_uint4(pg[4], pg[5], pg[6], pg[7]);
                au[0] = make_uint4(pu[0], pu[1], pu[2], pu[3]);
                au[1] = make_uint4(pu[4], pu[5], pu[6], pu[7]);
              }
            } else {
              for (int i = 0; i < 16 && n0 + c + i < p.N; ++i) {
                float gg = __bfloat162float(__float2bfloat16(__uint_as_float(g[i])));
                float uu = __bfloat162float(__float2bfloat16(__uint_as_float(u[i])));
                drow[n0 + c + i] = __float2bfloat16(silu_f(gg) * uu);
                if (arow) {
                  arow[n0 + c + i] = __float2bfloat16(gg);
                  arow[p.swiglu_up_row_offset + n0 + c + i] = __float2bfloat16(uu);
                }
              }
            }
          }
        }
      } else {
        const int n0 = tn * BN;
        OutT* drow = reinterpret_cast<OutT*>(p.D) + (size_t)row * p.ldd;
        const void* bias = p.bias;
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(taddr + c, v);
          tmem_ld_wait();
          if (split) split_reduce_chunk<32>(v, p, sched, w, rank, row_in_cta, c);
          const int col = n0 + c;
          if (!row_ok || col >= p.N) continue;
          if (col + 32 <= p.N) {
            if (bias) {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) + load16<kHalf>(bias, col + i));
            }
            if (sizeof(OutT) == 4) {
              float4* dst = reinterpret_cast<float4*>(drow + col);
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                float4 x = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                       __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
                if (kEpi == kAccumulate) {
                  float4 o = dst[i];
                  x.x += o.x, x.y += o.y, x.z += o.z, x.w += o.w;
                }
                dst[i] = x;
              }
            } else {
              uint4* dst = reinterpret_cast<uint4*>(drow + col);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                float f[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(v[8 * i + j]);
                if (kEpi == kAccumulate) {
                  uint4 o = dst[i];
                  float2 a = unpack16<kHalf>(o.x), b = unpack16<kHalf>(o.y), c2 = unpack16<kHalf>(o.z),
                         d = unpack16<kHalf>(o.w);
                  f[0] += a.x, f[1] += a.y, f[2] += b.x, f[3] += b.y, f[4] += c2.x, f[5] += c2.y, f[6] += d.x,
                      f[7] += d.y;
                }
                dst[i] = make_uint4(pack16<kHalf>(f[0], f[1]), pack16<kHalf>(f[2], f[3]), pack16<kHalf>(f[4], f[5]),
                                    pack16<kHalf>(f[6], f[7]));
              }
            }
          } else {
            for (int i = 0; i < 32 && col + i < p.N; ++i) {
              float x = __uint_as_float(v[i]);
              if (bias) x += load16<kHalf>(bias, col + i);
              if (kEpi == kAccumulate) x += (float)drow[col + i];
              drow[col + i] = (OutT)x;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(&bars->acc_empty[acc], 0);
      if (++acc == kAccStages) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2sm<512>(tmem_base);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_fn() {
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
static int make_map_2d(CUtensorMap* m, const void* ptr, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                       uint32_t box_outer) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return -1;
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    fprintf(stderr, "[scaling_b200] cuTensorMapEncodeTiled failed: CUresult=%d ptr=%p inner=%llu outer=%llu ld=%llu box=%ux%u\n",
            (int)r, ptr, (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)ld, box_inner, box_outer);
  return r == CUDA_SUCCESS ? 0 : -2;
}
static int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  }
  return n;
}

// Workspace of the split-K tail: caller-owned (the library allocates nothing), registered once per device with
// sb_gemm2_set_split_workspace.  Without a registered workspace the tail is simply not split.  The library serialises
// its use on the launching stream — GEMMs issued concurrently on several streams of one device must disable splitting.
struct SplitWorkspace {
  float* ws = nullptr;
  unsigned int* counters = nullptr;
};
static SplitWorkspace g_split_ws[64];
// partials: two slots per cluster and CTA half (stream-K) >= the split tail's one slot per tail part; counters: one per
// tile and CTA half
constexpr int kMaxStreamKTiles = 32768;
static size_t split_partial_bytes() { return (size_t)2 * num_sms() * kCtaAccFloats * sizeof(float); }
static size_t split_counter_bytes() { return (size_t)2 * kMaxStreamKTiles * sizeof(unsigned int); }
static bool stream_k_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("SCALING_B200_GEMM_STREAMK");
    on = (e && e[0] == '1') ? 1 : 0;  // opt-in until the measured A/B says otherwise
  }
  return on == 1;
}
static SplitWorkspace* split_workspace() {
  int dev = 0;
  cudaGetDevice(&dev);
  SplitWorkspace& w = g_split_ws[dev & 63];
  return w.ws ? &w : nullptr;
}
static bool split_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("SCALING_B200_GEMM_SPLITK");
    on = (e && e[0] == '0') ? 0 : 1;
  }
  return on == 1;
}

// Tensor maps of one launch: main operands plus the optional low-rank slab (copies of the main maps when there is none).
struct Maps {
  CUtensorMap a, b, a2, b2;
};
struct LowRank {
  const void* A2;  // [M, K2] row-major (K-major)
  const void* B2;  // [N, K2] row-major when b_layout == 0, [K2, N] when b_layout == 1
  int K2, lda2, ldb2;
};
static int make_maps(Maps& m, const void* A, const void* B, int M, int N, int K, int lda, int ldb, int a_layout,
                     int b_layout, const LowRank* lr, Params& p) {
  int rc;
  if (a_layout == 0)
    rc = make_map_2d(&m.a, A, K, M, lda, BK, 128);
  else
    rc = make_map_2d(&m.a, A, M, K, lda, 64, BK);
  if (rc) return 1000 + rc;
  if (b_layout == 0)
    rc = make_map_2d(&m.b, B, K, N, ldb, BK, 128);
  else
    rc = make_map_2d(&m.b, B, N, K, ldb, 64, BK);
  if (rc) return 2000 + rc;
  m.a2 = m.a, m.b2 = m.b;
  p.kb_extra = 0;
  if (lr && lr->K2 > 0) {
    if (a_layout != 0) return -11;  // the slab's A operand is K-major like the main one
    if (make_map_2d(&m.a2, lr->A2, lr->K2, M, lr->lda2, BK, 128)) return 3000;
    if (b_layout == 0)
      rc = make_map_2d(&m.b2, lr->B2, lr->K2, N, lr->ldb2, BK, 128);
    else
      rc = make_map_2d(&m.b2, lr->B2, N, lr->K2, lr->ldb2, 64, BK);
    if (rc) return 4000 + rc;
    p.kb_extra = (lr->K2 + BK - 1) / BK;
  }
  return 0;
}

// Arrival gate of the NEXT 2-CTA GEMM launched on the current device (consumed by that launch): see Params::gate.
struct PendingGate {
  const uint32_t* gate = nullptr;
  uint32_t target = 0;
};
static PendingGate g_pending_gate[64];

template <int AM, int BMj, int EPI, typename OutT, bool kHalf = false>
static int launch(const Maps& m, Params& p, cudaStream_t stream, int max_ctas) {
  auto kern = gemm2_kernel<AM, BMj, EPI, OutT, kHalf>;
  {
    int dev = 0;
    cudaGetDevice(&dev);
    PendingGate& pg = g_pending_gate[dev & 63];
    p.gate = pg.gate, p.gate_target = pg.target;
    pg.gate = nullptr;
  }
  static PerDeviceOnce once;  // the attribute is per device
  if (once.first())
    SB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)(EPI == kAllReduce ? kArSmemBytes : kSmemBytes)));
  const int tiles = p.tiles_m * p.tiles_n;
  int width = num_sms() / 2;  // clusters the machine runs at once
  if (max_ctas > 0 && width > max_ctas / 2) width = max_ctas / 2 > 0 ? max_ctas / 2 : 1;
  p.full_tiles = tiles, p.split = 1;
  // Wave quantisation: 256 tiles on 74 clusters are 3.46 waves, i.e. the machine idles 13 % of the kernel.  Cut the
  // tiles of the last, partial wave along K so that they occupy (almost) every cluster; the parts meet in an fp32
  // workspace and the last one to arrive runs the epilogue on the sum.
  if (EPI != kAllReduce && EPI != kXentStats && split_enabled()) {
    const int rem = tiles % width;
    const int k_blocks = (p.K + BK - 1) / BK + p.kb_extra;
    int parts = rem > 0 ? width / rem : 1;
    // Measured on B200 (profiles/kbench_r1_splitk.json): parking and re-reading the 128 KiB partials costs 10-25 us per
    // tail, and the un-split tail already runs at boosted clocks (half the SMs idle under the power cap), so splitting
    // pays only when each part still holds >= 112 k-blocks (K >= 14336 for two parts).
    if (parts > 4) parts = 4;
    if (parts > k_blocks / 112) parts = k_blocks / 112;
    if (parts >= 2) {
      if (SplitWorkspace* w = split_workspace()) {
        p.split = parts, p.full_tiles = tiles - rem;
        p.split_ws = w->ws, p.split_counters = w->counters;
      }
    }
  }
  int units = p.full_tiles + (tiles - p.full_tiles) * p.split;
  int clusters = units < width ? units : width;
  // stream-K tail (SCALING_B200_GEMM_STREAMK=1): whenever the tiles do not fill whole waves, the last full
  // wave and the partial one are balanced by k-blocks instead of whole tiles (see Params::stream_k).  256 tiles on 74
  // clusters then take 3.46 tile-times instead of 4, and a 16-tile GEMM with a long K occupies all 74 clusters instead of
  // 16.  Supersedes the split tail above.
  p.stream_k = 0;
  if (EPI != kAllReduce && EPI != kXentStats && split_enabled() && stream_k_enabled() && tiles % width != 0 &&
      tiles <= kMaxStreamKTiles) {
    // whole waves except the last full one run as plain tiles; the last full wave + the partial one are balanced
    const int first = tiles > width ? (tiles / width - 1) * width : 0;
    const long long total = (long long)(tiles - first) * ((p.K + BK - 1) / BK + p.kb_extra);
    if (total / width >= 8) {
      if (SplitWorkspace* w = split_workspace()) {
        p.stream_k = 1, p.sk_first_tile = first, p.full_tiles = tiles, p.split = 1;
        p.split_ws = w->ws, p.split_counters = w->counters;
        clusters = width;
      }
    }
  }
  if (EPI == kAllReduce)
    kern<<<clusters * 2, kArThreads, kArSmemBytes, stream>>>(m.a, m.b, m.a2, m.b2, p);
  else
    kern<<<clusters * 2, kThreads, kSmemBytes, stream>>>(m.a, m.b, m.a2, m.b2, p);
  return (int)cudaGetLastError();
}

}  // namespace gemm2
}  // namespace sb

using namespace sb::gemm2;

// Split-K tail workspace (see SplitWorkspace): bytes needed on the current device, and registration of a caller-owned,
// 16-byte aligned buffer of at least that size for the current device.  The words behind the fp32 partials are arrival
// counters and must be zero at registration (the kernels leave them zero).  Pass NULL to unregister.
extern "C" long long sb_gemm2_split_workspace_bytes() { return (long long)(split_partial_bytes() + split_counter_bytes()); }
extern "C" int sb_gemm2_set_split_workspace(void* base, unsigned long long bytes) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  SplitWorkspace& w = g_split_ws[dev & 63];
  if (base == nullptr) {
    w.ws = nullptr, w.counters = nullptr;
    return 0;
  }
  if (bytes < split_partial_bytes() + split_counter_bytes() || (reinterpret_cast<uintptr_t>(base) & 15)) return -2;
  w.ws = static_cast<float*>(base);
  w.counters = reinterpret_cast<unsigned int*>(static_cast<char*>(base) + split_partial_bytes());
  return 0;
}

static int gemm2_impl(const void* A, const void* B, void* D, const void* bias, int M, int N, int K, int lda, int ldb,
                      int ldd, int a_layout, int b_layout, int out_dtype, int epilogue, const LowRank* lr, int max_ctas,
                      cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  Params p{};
  Maps m;
  if (int rc = make_maps(m, A, B, M, N, K, lda, ldb, a_layout, b_layout, lr, p)) return rc;
  p.M = M, p.N = N, p.K = K, p.D = D, p.ldd = ldd, p.bias = bias;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + BN - 1) / BN;
#define SB_DISPATCH(AM, BMj)                                                                                  \
  if (a_layout == AM && b_layout == BMj) {                                                                    \
    if (epilogue == 0 && out_dtype == 0) return launch<AM, BMj, kStore, __nv_bfloat16>(m, p, stream, max_ctas);      \
    if (epilogue == 0 && out_dtype == 1) return launch<AM, BMj, kStore, float>(m, p, stream, max_ctas);              \
    if (epilogue == 1 && out_dtype == 0) return launch<AM, BMj, kAccumulate, __nv_bfloat16>(m, p, stream, max_ctas); \
    if (epilogue == 1 && out_dtype == 1) return launch<AM, BMj, kAccumulate, float>(m, p, stream, max_ctas);         \
  }
  SB_DISPATCH(0, 0)
  SB_DISPATCH(0, 1)
  SB_DISPATCH(1, 1)
  SB_DISPATCH(1, 0)
#undef SB_DISPATCH
  return -10;
}

// all-gather -> GEMM gate for the NEXT sb_gemm2_* launch on the current device (any epilogue): `gate` is a device array
// with one u32 arrival counter per 256-row tile of A, `target` the value a complete tile shows (see
// sb_seq_gather_push_counted in elementwise.cu).  One-shot: the launch clears it.  A must be row-major ([M, K]).
extern "C" int sb_gemm2_set_gate(const unsigned int* gate, unsigned int target) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  g_pending_gate[dev & 63].gate = gate;
  g_pending_gate[dev & 63].target = target;
  return 0;
}

// Same contract as sb_gemm_bf16 (gemm_sm100.cu), computed by CTA pairs (cta_group::2).
extern "C" int sb_gemm2_bf16(const void* A, const void* B, void* D, const void* bias, int M, int N, int K, int lda, int ldb,
                             int ldd, int a_layout, int b_layout, int out_dtype, int epilogue, int max_ctas,
                             cudaStream_t stream) {
  return gemm2_impl(A, B, D, bias, M, N, K, lda, ldb, ldd, a_layout, b_layout, out_dtype, epilogue, nullptr, max_ctas,
                    stream);
}

// fp16 operands on the same kernel (reference acceptance grid: fp16 + dynamic loss scaler,
// tests/core/test_training/test_training.py:149-235): fprop (a_layout 0, b_layout 0), dgrad (0, 1) and wgrad (1, 1) of a
// linear layer.  out_dtype 0 = fp16, 1 = fp32; epilogue 1 accumulates into D (fp32 main-grad arenas); bias is fp16.
// The tensor maps are shared with the bf16 path: their element type only sizes the 2-byte elements.
extern "C" int sb_gemm2_f16(const void* A, const void* B, void* D, const void* bias, int M, int N, int K, int lda, int ldb,
                            int ldd, int a_layout, int b_layout, int out_dtype, int epilogue, int max_ctas,
                            cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  Params p{};
  Maps m;
  if (int rc = make_maps(m, A, B, M, N, K, lda, ldb, a_layout, b_layout, nullptr, p)) return rc;
  p.M = M, p.N = N, p.K = K, p.D = D, p.ldd = ldd, p.bias = bias;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + BN - 1) / BN;
  if (a_layout == 0 && b_layout == 0 && epilogue == 0 && out_dtype == 0)
    return launch<0, 0, kStore, __half, true>(m, p, stream, max_ctas);
  if (a_layout == 0 && b_layout == 1 && epilogue == 0 && out_dtype == 0)
    return launch<0, 1, kStore, __half, true>(m, p, stream, max_ctas);
  if (a_layout == 1 && b_layout == 1 && epilogue == 0 && out_dtype == 0)
    return launch<1, 1, kStore, __half, true>(m, p, stream, max_ctas);
  if (a_layout == 1 && b_layout == 1 && epilogue == 0 && out_dtype == 1)
    return launch<1, 1, kStore, float, true>(m, p, stream, max_ctas);
  if (a_layout == 1 && b_layout == 1 && epilogue == 1 && out_dtype == 1)
    return launch<1, 1, kAccumulate, float, true>(m, p, stream, max_ctas);
  return -10;
}

// LoRA-fused linear: D[M,N] (+)= A·Bᵀ + A2·B2ᵀ in ONE accumulator — the low-rank pair rides as extra
// k-blocks behind the main K range, so Y = X·Wᵀ + (s·X·Aᵀ)·B_loraᵀ costs one launch and no extra pass over Y.
// A [M,K] and A2 [M,K2] are row-major; B/B2 are [N,K]/[N,K2] (b_layout 0) or [K,N]/[K2,N] (b_layout 1); K2 % 8 == 0.
// Reference: separate F.linear calls and an add (core/nn/lora.py:103-112, core/nn/attention/attention.py:709-744).
extern "C" int sb_gemm2_lowrank_bf16(const void* A, const void* B, const void* A2, const void* B2, void* D,
                                     const void* bias, int M, int N, int K, int K2, int lda, int ldb, int lda2, int ldb2,
                                     int ldd, int b_layout, int out_dtype, int epilogue, int max_ctas,
                                     cudaStream_t stream) {
  LowRank lr{A2, B2, K2, lda2, ldb2};
  return gemm2_impl(A, B, D, bias, M, N, K, lda, ldb, ldd, 0, b_layout, out_dtype, epilogue, &lr, max_ctas, stream);
}

// LM-head GEMM with the cross-entropy statistics in its epilogue: logits[M,N] = X·Wᵀ (bf16) and
// stats[M, ceil(N/256)] float4 = per row and 256-column tile (max, Σexp(x - max), unused, local arg-max column as int
// bits), taken from the stored bf16 values.  sb_xent_combine (xent.cu) merges the tiles, picks the target logit and
// exchanges the per-token result across the tensor-parallel group.  Replaces F.linear + all-gather + fp32 copy +
// cross_entropy (transformer/model/layers/lm_head_tied.py:41-43, transformer/model/model.py:56-76).
extern "C" long long sb_gemm2_xent_tiles(int N) { return (N + BN - 1) / BN; }
extern "C" int sb_gemm2_xent_stats_bf16(const void* X, const void* W, void* logits, float* stats, int M, int N, int K,
                                        int ldx, int ldw, int ldl, int max_ctas, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  Params p{};
  Maps m;
  if (int rc = make_maps(m, X, W, M, N, K, ldx, ldw, 0, 0, nullptr, p)) return rc;
  p.M = M, p.N = N, p.K = K, p.D = logits, p.ldd = ldl, p.aux = stats;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + BN - 1) / BN;
  return launch<0, 0, kXentStats, __nv_bfloat16>(m, p, stream, max_ctas);
}

// q|k|v projection + rotary in one launch: Y[M,N] = X·Wᵀ with the rotary embedding (interleaved pairs, "rotary_complex")
// applied to columns [0, rope_cols) — heads of head_dim columns, the first rot_dim of each rotated — using
// cos/sin[pos][rot_dim/2] fp32 tables and pos[M] (or row % seq when pos is NULL).  head_dim % 32 == 0, rot_dim % 8 == 0.
extern "C" int sb_gemm2_rope_bf16(const void* X, const void* W, void* Y, int M, int N, int K, int ldx, int ldw, int ldy,
                                  const float* cos_t, const float* sin_t, const int* pos, int seq, int rope_cols,
                                  int head_dim, int rot_dim, int max_ctas, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if (head_dim % 32 || rot_dim % 8 || rot_dim > head_dim || rope_cols % head_dim || rope_cols > N || seq <= 0) return -1;
  Params p{};
  Maps m;
  if (int rc = make_maps(m, X, W, M, N, K, ldx, ldw, 0, 0, nullptr, p)) return rc;
  p.M = M, p.N = N, p.K = K, p.D = Y, p.ldd = ldy;
  p.rope_cos = cos_t, p.rope_sin = sin_t, p.rope_pos = pos, p.rope_seq = seq, p.rope_cols = rope_cols;
  p.rope_head_dim = head_dim, p.rope_rot = rot_dim;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + BN - 1) / BN;
  return launch<0, 0, kRope, __nv_bfloat16>(m, p, stream, max_ctas);
}

// Linear + activation in one launch (adapter bottlenecks, GELU / SiLU MLP up-projections):
// H[M,N] = act(X·Wᵀ + bias), pre[M,N] (optional) = the bf16 pre-activation for the backward pass.  act: 0 GELU (erf),
// 1 SiLU.  Reference: F.linear followed by torch.nn.functional.gelu / silu (core/nn/mlp.py:77-89,
// transformer/model/layers/layer.py:143-186).
extern "C" int sb_gemm2_act_bf16(const void* X, const void* W, const void* bias, void* H, void* pre, int M, int N, int K,
                                 int ldx, int ldw, int ldh, int ldpre, int act, int max_ctas, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  Params p{};
  Maps m;
  if (int rc = make_maps(m, X, W, M, N, K, ldx, ldw, 0, 0, nullptr, p)) return rc;
  p.M = M, p.N = N, p.K = K, p.D = H, p.ldd = ldh, p.bias = bias, p.aux = pre, p.ldaux = ldpre;
  p.swiglu_up_row_offset = act;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + BN - 1) / BN;
  return launch<0, 0, kAct, __nv_bfloat16>(m, p, stream, max_ctas);
}

// Same contract as sb_gemm_swiglu_bf16: CTA 0 of the pair stages the gate rows, CTA 1 the up rows of the tile.
extern "C" int sb_gemm2_swiglu_bf16(const void* X, const void* W, void* H, void* pre, int M, int F, int K, int ldx, int ldw,
                                    int ldh, int ldpre, int max_ctas, cudaStream_t stream) {
  if (M <= 0 || F <= 0 || K <= 0) return 0;
  Maps m;
  if (int rc = make_map_2d(&m.a, X, K, M, ldx, BK, 128)) return 1000 + rc;
  if (int rc = make_map_2d(&m.b, W, K, 2 * (uint64_t)F, ldw, BK, 128)) return 2000 + rc;
  m.a2 = m.a, m.b2 = m.b;
  Params p{};
  p.M = M, p.N = F, p.K = K, p.D = H, p.ldd = ldh, p.aux = pre, p.ldaux = ldpre, p.swiglu_up_row_offset = F;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (F + BN / 2 - 1) / (BN / 2);
  return launch<0, 0, kSwiGLU, __nv_bfloat16>(m, p, stream, max_ctas);
}

// Optional phase profile of the fused collective kernels: a caller-owned buffer of grid x 8 u64 (148 x 8 is enough) for
// the current device; NULL switches it off.  The kernels stamp %globaltimer at their phase boundaries (see prof_stamp).
static unsigned long long* g_prof_buf[64] = {nullptr};
extern "C" int sb_gemm2_set_profile_buffer(unsigned long long* buf) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  g_prof_buf[dev & 63] = buf;
  return 0;
}

// Fused tensor-parallel GEMM + all-reduce: D = sum over ranks of op(A)·op(B) (+ bias), bf16, identical bits on every
// rank.  `data_ptrs[r]` / `flag_ptrs[r]` are rank r's symmetric workspace regions (for this call's parity) as mapped
// into this process; `epoch` must be unique per call on a workspace parity and equal on all ranks.  Replaces the
// reference's GEMM followed by torch.distributed.all_reduce (core/nn/linear/row_parallel_linear.py:150-160, and the
// backward all-reduce of core/nn/linear/utils.py copy_to_tensor_model_parallel_region).
static int fused_collective_impl(const void* A, const void* B, void* D, const void* bias, int M, int N, int K, int lda,
                                 int ldb, int ldd, int a_layout, int b_layout, const unsigned long long* data_ptrs,
                                 const unsigned long long* flag_ptrs, int tp_rank, int tp_size, unsigned int epoch,
                                 unsigned long long slot_bytes, unsigned int flag_slot, int rs_seq, int max_ctas,
                                 cudaStream_t stream, const LowRank* lr = nullptr) {
  // one-shot (every rank reduces every tile) is byte-optimal only for two ranks; beyond that tiles have owners
  static int force_two_shot = -1;
  if (force_two_shot < 0) {
    const char* e = getenv("SCALING_B200_AR_TWO_SHOT");
    force_two_shot = e ? atoi(e) : 2;  // 0 never, 1 always, 2 (default) when tp > 2
  }
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if (tp_size < 2 || tp_size > kMaxTP || tp_rank < 0 || tp_rank >= tp_size) return -20;
  Params p{};
  Maps m;
  if (int rc = make_maps(m, A, B, M, N, K, lda, ldb, a_layout, b_layout, lr, p)) return rc;
  p.M = M, p.N = N, p.K = K, p.D = D, p.ldd = ldd, p.bias = bias;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + BN - 1) / BN;
  const unsigned long long cta_tiles = 2ull * p.tiles_m * p.tiles_n;
  if (cta_tiles * kCtaTileBytes > slot_bytes || cta_tiles > flag_slot) return -21;
  {
    int dev = 0;
    cudaGetDevice(&dev);
    p.prof = g_prof_buf[dev & 63];
  }
  p.rs_seq = rs_seq, p.rs_seq_per_rank = rs_seq > 0 ? rs_seq / tp_size : 0;
  p.two_shot = (rs_seq == 0 && (force_two_shot == 1 || (force_two_shot == 2 && tp_size > 2))) ? 1 : 0;
  p.tp_rank = tp_rank, p.tp_size = tp_size, p.epoch = epoch, p.flag_slot = flag_slot, p.slot_bytes = slot_bytes;
  for (int r = 0; r < tp_size; ++r) {
    p.ar_data[r] = reinterpret_cast<uint8_t*>(data_ptrs[r]);
    p.ar_flags[r] = reinterpret_cast<uint32_t*>(flag_ptrs[r]);
  }
  if (a_layout == 0 && b_layout == 0) return launch<0, 0, kAllReduce, __nv_bfloat16>(m, p, stream, max_ctas);
  if (a_layout == 0 && b_layout == 1) return launch<0, 1, kAllReduce, __nv_bfloat16>(m, p, stream, max_ctas);
  return -10;
}

// Fused tensor-parallel GEMM + reduce-scatter along the sequence (Megatron sequence parallelism): rows of the product
// are [batch, seq]; rank r receives the sum over ranks of its sequence slice, D = [batch · seq/tp, N] compact.  Same
// workspace / epoch contract as sb_gemm2_allreduce_bf16; requires (seq / tp_size) % 256 == 0 and M % seq == 0.
// Reference: F.linear then reduce_scatter (core/nn/linear/utils.py:279-300, attention.py:704-705, mlp.py:164-165).
extern "C" int sb_gemm2_reduce_scatter_bf16(const void* A, const void* B, void* D, const void* bias, int M, int N, int K,
                                            int lda, int ldb, int ldd, int a_layout, int b_layout,
                                            const unsigned long long* data_ptrs, const unsigned long long* flag_ptrs,
                                            int tp_rank, int tp_size, unsigned int epoch, unsigned long long slot_bytes,
                                            unsigned int flag_slot, int seq, int max_ctas, cudaStream_t stream) {
  if (seq <= 0 || M % seq || seq % tp_size || (seq / tp_size) % BM) return -22;
  return fused_collective_impl(A, B, D, bias, M, N, K, lda, ldb, ldd, a_layout, b_layout, data_ptrs, flag_ptrs, tp_rank,
                               tp_size, epoch, slot_bytes, flag_slot, seq, max_ctas, stream);
}

extern "C" int sb_gemm2_allreduce_bf16(const void* A, const void* B, void* D, const void* bias, int M, int N, int K, int lda,
                                       int ldb, int ldd, int a_layout, int b_layout, const unsigned long long* data_ptrs,
                                       const unsigned long long* flag_ptrs, int tp_rank, int tp_size, unsigned int epoch,
                                       unsigned long long slot_bytes, unsigned int flag_slot, int max_ctas,
                                       cudaStream_t stream) {
  return fused_collective_impl(A, B, D, bias, M, N, K, lda, ldb, ldd, a_layout, b_layout, data_ptrs, flag_ptrs, tp_rank,
                               tp_size, epoch, slot_bytes, flag_slot, 0, max_ctas, stream);
}

// Fused all-reduce with a low-rank slab (see sb_gemm2_lowrank_bf16): D = Σ_ranks (A·Bᵀ + A2·B2ᵀ) (+ bias).
extern "C" int sb_gemm2_allreduce_lowrank_bf16(const void* A, const void* B, const void* A2, const void* B2, void* D,
                                               const void* bias, int M, int N, int K, int K2, int lda, int ldb, int lda2,
                                               int ldb2, int ldd, int b_layout, const unsigned long long* data_ptrs,
                                               const unsigned long long* flag_ptrs, int tp_rank, int tp_size,
                                               unsigned int epoch, unsigned long long slot_bytes, unsigned int flag_slot,
                                               int max_ctas, cudaStream_t stream) {
  LowRank lr{A2, B2, K2, lda2, ldb2};
  return fused_collective_impl(A, B, D, bias, M, N, K, lda, ldb, ldd, 0, b_layout, data_ptrs, flag_ptrs, tp_rank,
                               tp_size, epoch, slot_bytes, flag_slot, 0, max_ctas, stream, &lr);
}

// Bytes of workspace one sender slot needs for an [M, N] fused all-reduce, and flags per slot.
extern "C" long long sb_gemm2_allreduce_slot_bytes(int M, int N) {
  return 2ll * ((M + BM - 1) / BM) * ((N + BN - 1) / BN) * (long long)kCtaTileBytes;
}
