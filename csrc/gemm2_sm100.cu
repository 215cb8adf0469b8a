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

enum Epilogue : int { kStore = 0, kAccumulate = 1, kSwiGLU = 2 };

struct Params {
  int M, N, K;
  void* D;
  int ldd;
  const void* bias;
  void* aux;
  int ldaux;
  int swiglu_up_row_offset;
  int tiles_m, tiles_n;
};

struct __align__(8) Barriers {
  uint64_t full[kStages];
  uint64_t empty[kStages];
  uint64_t acc_full[kAccStages];
  uint64_t acc_empty[kAccStages];
  uint32_t tmem_base;
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

template <int kAMajorMN, int kBMajorMN, int kEpi, typename OutT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  Barriers* bars = reinterpret_cast<Barriers*>(smem + kStages * kStageBytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int num_tiles = p.tiles_m * p.tiles_n;
  const int k_blocks = (p.K + BK - 1) / BK;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&bars->full[i], 1);
      mbar_init(&bars->empty[i], 1);
    }
    for (int i = 0; i < kAccStages; ++i) {
      mbar_init(&bars->acc_full[i], 1);
      mbar_init(&bars->acc_empty[i], 8);  // 4 epilogue warps in each of the two CTAs arrive on the leader's barrier
    }
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
      for (int t = cluster_id; t < num_tiles; t += num_clusters) {
        int tm, tn;
        tile_coords(t, p.tiles_m, p.tiles_n, tm, tn);
        const int m0 = tm * BM + rank * 128;
        const int nb = (kEpi == kSwiGLU) ? tn * (BN / 2) + (rank ? p.swiglu_up_row_offset : 0) : tn * BN + rank * 128;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&bars->empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          uint8_t* sb_ = sa + kABytes;
          if (leader) mbar_arrive_expect_tx(&bars->full[stage], 2 * kStageBytes);  // both CTAs' bytes land here
          const int k0 = kb * BK;
          if (kAMajorMN) {
            tma_load_2d_2sm(sa, &tma_a, &bars->full[stage], m0, k0);
            tma_load_2d_2sm(sa + 8192, &tma_a, &bars->full[stage], m0 + 64, k0);
          } else {
            tma_load_2d_2sm(sa, &tma_a, &bars->full[stage], k0, m0);
          }
          if (kBMajorMN) {
            tma_load_2d_2sm(sb_, &tma_b, &bars->full[stage], nb, k0);
            tma_load_2d_2sm(sb_ + 8192, &tma_b, &bars->full[stage], nb + 64, k0);
          } else {
            tma_load_2d_2sm(sb_, &tma_b, &bars->full[stage], k0, nb);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (leader CTA only) =====================
    if (leader && lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, kAMajorMN, kBMajorMN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = cluster_id; t < num_tiles; t += num_clusters) {
        mbar_wait(&bars->acc_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < k_blocks; ++kb) {
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
            umma_f16_ss_2sm(tmem_d, da, db, idesc, (kb | k) != 0);
          }
          umma_commit_2sm(&bars->empty[stage]);
          if (++stage == kStages) {
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
  } else if (warp >= 4) {
    // ===================== epilogue (both CTAs, own 128 rows) =====================
    const int ew = warp & 3;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = cluster_id; t < num_tiles; t += num_clusters) {
      int tm, tn;
      tile_coords(t, p.tiles_m, p.tiles_n, tm, tn);
      const int row = tm * BM + rank * 128 + ew * 32 + lane;
      mbar_wait(&bars->acc_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + ((uint32_t)(ew * 32) << 16);
      const bool row_ok = row < p.M;

      if (kEpi == kSwiGLU) {
        const int n0 = tn * (BN / 2);
        __nv_bfloat16* drow = reinterpret_cast<__nv_bfloat16*>(p.D) + (size_t)row * p.ldd;
        __nv_bfloat16* arow = p.aux ? reinterpret_cast<__nv_bfloat16*>(p.aux) + (size_t)row * p.ldaux : nullptr;
#pragma unroll 1
        for (int c = 0; c < BN / 2; c += 16) {
          uint32_t g[16], u[16];
          tmem_ld_32x32b_x16(taddr + c, g);
          tmem_ld_32x32b_x16(taddr + BN / 2 + c, u);
          tmem_ld_wait();
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
        const __nv_bfloat16* bias = reinterpret_cast<const __nv_bfloat16*>(p.bias);
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(taddr + c, v);
          tmem_ld_wait();
          const int col = n0 + c;
          if (!row_ok || col >= p.N) continue;
          if (col + 32 <= p.N) {
            if (bias) {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) + __bfloat162float(bias[col + i]));
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
                  float2 a = unpack_bf16(o.x), b = unpack_bf16(o.y), c2 = unpack_bf16(o.z), d = unpack_bf16(o.w);
                  f[0] += a.x, f[1] += a.y, f[2] += b.x, f[3] += b.y, f[4] += c2.x, f[5] += c2.y, f[6] += d.x,
                      f[7] += d.y;
                }
                dst[i] = make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]),
                                    pack_bf16(f[6], f[7]));
              }
            }
          } else {
            for (int i = 0; i < 32 && col + i < p.N; ++i) {
              float x = __uint_as_float(v[i]);
              if (bias) x += __bfloat162float(bias[col + i]);
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

template <int AM, int BMj, int EPI, typename OutT>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, cudaStream_t stream, int max_ctas) {
  auto kern = gemm2_kernel<AM, BMj, EPI, OutT>;
  static bool configured = false;
  if (!configured) {
    SB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    configured = true;
  }
  int tiles = p.tiles_m * p.tiles_n;
  int clusters = num_sms() / 2;
  if (tiles < clusters) clusters = tiles;
  if (max_ctas > 0 && clusters > max_ctas / 2) clusters = max_ctas / 2 > 0 ? max_ctas / 2 : 1;
  kern<<<clusters * 2, kThreads, kSmemBytes, stream>>>(ta, tb, p);
  return (int)cudaGetLastError();
}

}  // namespace gemm2
}  // namespace sb

using namespace sb::gemm2;

// Same contract as sb_gemm_bf16 (gemm_sm100.cu), computed by CTA pairs (cta_group::2).
extern "C" int sb_gemm2_bf16(const void* A, const void* B, void* D, const void* bias, int M, int N, int K, int lda, int ldb,
                             int ldd, int a_layout, int b_layout, int out_dtype, int epilogue, int max_ctas,
                             cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  CUtensorMap ta, tb;
  int rc;
  if (a_layout == 0)
    rc = make_map_2d(&ta, A, K, M, lda, BK, 128);
  else
    rc = make_map_2d(&ta, A, M, K, lda, 64, BK);
  if (rc) return 1000 + rc;
  if (b_layout == 0)
    rc = make_map_2d(&tb, B, K, N, ldb, BK, 128);
  else
    rc = make_map_2d(&tb, B, N, K, ldb, 64, BK);
  if (rc) return 2000 + rc;
  Params p{};
  p.M = M, p.N = N, p.K = K, p.D = D, p.ldd = ldd, p.bias = bias;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + BN - 1) / BN;
#define SB_DISPATCH(AM, BMj)                                                                                          \
  if (a_layout == AM && b_layout == BMj) {                                                                            \
    if (epilogue == 0 && out_dtype == 0) return launch<AM, BMj, kStore, __nv_bfloat16>(ta, tb, p, stream, max_ctas);      \
    if (epilogue == 0 && out_dtype == 1) return launch<AM, BMj, kStore, float>(ta, tb, p, stream, max_ctas);              \
    if (epilogue == 1 && out_dtype == 0) return launch<AM, BMj, kAccumulate, __nv_bfloat16>(ta, tb, p, stream, max_ctas); \
    if (epilogue == 1 && out_dtype == 1) return launch<AM, BMj, kAccumulate, float>(ta, tb, p, stream, max_ctas);         \
  }
  SB_DISPATCH(0, 0)
  SB_DISPATCH(0, 1)
  SB_DISPATCH(1, 1)
  SB_DISPATCH(1, 0)
#undef SB_DISPATCH
  return -10;
}

// Same contract as sb_gemm_swiglu_bf16: CTA 0 of the pair stages the gate rows, CTA 1 the up rows of the tile.
extern "C" int sb_gemm2_swiglu_bf16(const void* X, const void* W, void* H, void* pre, int M, int F, int K, int ldx, int ldw,
                                    int ldh, int ldpre, int max_ctas, cudaStream_t stream) {
  if (M <= 0 || F <= 0 || K <= 0) return 0;
  CUtensorMap ta, tb;
  if (int rc = make_map_2d(&ta, X, K, M, ldx, BK, 128)) return 1000 + rc;
  if (int rc = make_map_2d(&tb, W, K, 2 * (uint64_t)F, ldw, BK, 128)) return 2000 + rc;
  Params p{};
  p.M = M, p.N = F, p.K = K, p.D = H, p.ldd = ldh, p.aux = pre, p.ldaux = ldpre, p.swiglu_up_row_offset = F;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (F + BN / 2 - 1) / (BN / 2);
  return launch<0, 0, kSwiGLU, __nv_bfloat16>(ta, tb, p, stream, max_ctas);
}
