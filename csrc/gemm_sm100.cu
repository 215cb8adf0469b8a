// tcgen05 GEMM family for sm_100a: D[M,N] (+)= A[M,K] * B[N,K]^T with fp32 accumulation in TMEM.
//
// Replaces the reference's cuBLAS calls behind F.linear (reference: core/nn/linear/column_parallel_linear.py:151,
// row_parallel_linear.py:158, mlp.py:157-161) with one persistent, warp-specialised kernel:
//   warp 0      : TMA producer (cp.async.bulk.tensor, 128B swizzle) into a kStages-deep smem ring
//   warp 1      : MMA issuer, one elected thread, tcgen05.mma cta_group::1 kind::f16, UMMA 128x256x16
//   warp 2      : TMEM allocator (512 columns = two 128x256 fp32 accumulators, double buffered)
//   warps 4..7  : epilogue, tcgen05.ld 32x32b -> registers -> fused epilogue -> global
// Operands may be K-major or MN-major (transposed) in global memory, so fprop (X·Wᵀ), dgrad (dY·W) and wgrad
// (dYᵀ·X) all run from the row-major training tensors without a transpose pass.
// Epilogues: store bf16/fp32, accumulate into an fp32/bf16 main-grad buffer (D += acc), + bias, and the
// dual-GEMM SwiGLU epilogue h = silu(X·W1ᵀ) ⊙ (X·W2ᵀ) with W1/W2 tiles loaded side by side (gate|up).
#include "common.cuh"

namespace sb {
namespace gemm {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;  // 64 bf16 = 128 bytes = one swizzle row
constexpr int UK = 16;  // UMMA K for 16-bit inputs
constexpr int kStages = 4;
constexpr int kAccStages = 2;
constexpr int kThreads = 256;
constexpr uint32_t kABytes = BM * BK * 2;  // 16 KiB
constexpr uint32_t kBBytes = BN * BK * 2;  // 32 KiB
constexpr uint32_t kStageBytes = kABytes + kBBytes;
constexpr uint32_t kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;

enum Epilogue : int { kStore = 0, kAccumulate = 1, kSwiGLU = 2 };

struct Params {
  int M, N, K;
  void* D;            // output, row-major [M, ldd]
  int ldd;
  const void* bias;   // optional bf16 [N]
  void* aux;          // SwiGLU: optional pre-activation output [M, 2*ldaux halves] (gate | up), bf16
  int ldaux;
  int swiglu_up_row_offset;  // row of B where the "up" weight starts (F for concatenated [W1; W2])
  int tiles_m, tiles_n;
};

struct __align__(8) Barriers {
  uint64_t full[kStages];
  uint64_t empty[kStages];
  uint64_t acc_full[kAccStages];
  uint64_t acc_empty[kAccStages];
  uint32_t tmem_base;
};

// tile order: sweep N inside groups of 8 M-tiles so that concurrently resident CTAs share A and B tiles in L2
SB_DEVICE void tile_coords(int t, int tiles_m, int tiles_n, int& tm, int& tn) {
  constexpr int G = 8;
  int per_group = G * tiles_n;
  int g = t / per_group;
  int first_m = g * G;
  int gsize = min(G, tiles_m - first_m);
  int r = t - g * per_group;
  tm = first_m + r % gsize;
  tn = r / gsize;
}

template <int kAMajorMN, int kBMajorMN, int kEpi, typename OutT>
__global__ void __launch_bounds__(kThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  Barriers* bars = reinterpret_cast<Barriers*>(smem + kStages * kStageBytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n;
  const int k_blocks = (p.K + BK - 1) / BK;

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
      mbar_init(&bars->acc_empty[i], 4);  // one arrival per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(&bars->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        int tm, tn;
        tile_coords(t, p.tiles_m, p.tiles_n, tm, tn);
        const int m0 = tm * BM;
        const int n0 = (kEpi == kSwiGLU) ? tn * (BN / 2) : tn * BN;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&bars->empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * kStageBytes;
          uint8_t* sb_ = sa + kABytes;
          mbar_arrive_expect_tx(&bars->full[stage], kStageBytes);
          const int k0 = kb * BK;
          if (kAMajorMN) {
            // A stored [K, M] row-major: boxes of 64 (M, inner) x 64 (K)
            tma_load_2d(sa, &tma_a, &bars->full[stage], m0, k0);
            tma_load_2d(sa + 8192, &tma_a, &bars->full[stage], m0 + 64, k0);
          } else {
            tma_load_2d(sa, &tma_a, &bars->full[stage], k0, m0);  // box 64 (K) x 128 (M)
          }
          if (kBMajorMN) {
#pragma unroll
            for (int j = 0; j < 4; ++j) tma_load_2d(sb_ + j * 8192, &tma_b, &bars->full[stage], n0 + j * 64, k0);
          } else if (kEpi == kSwiGLU) {
            tma_load_2d(sb_, &tma_b, &bars->full[stage], k0, n0);                                   // gate rows
            tma_load_2d(sb_ + 16384, &tma_b, &bars->full[stage], k0, p.swiglu_up_row_offset + n0);  // up rows
          } else {
            tma_load_2d(sb_, &tma_b, &bars->full[stage], k0, n0);  // box 64 (K) x 256 (N)
            // (two 128-row boxes: boxDim[1] is limited to 256 and 128 keeps the map reusable for SwiGLU)
            tma_load_2d(sb_ + 16384, &tma_b, &bars->full[stage], k0, n0 + 128);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, kAMajorMN, kBMajorMN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
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
            // K-major: advance 32 B inside the 128 B swizzle row; MN-major: advance two 8-row groups (2 KiB)
            const uint64_t da = kAMajorMN ? make_smem_desc_sw128(sa + k * 2048, 8192, 1024)
                                          : make_smem_desc_sw128(sa + k * 32, 16, 1024);
            const uint64_t db = kBMajorMN ? make_smem_desc_sw128(sb_ + k * 2048, 8192, 1024)
                                          : make_smem_desc_sw128(sb_ + k * 32, 16, 1024);
            umma_f16_ss(tmem_d, da, db, idesc, (kb | k) != 0);
          }
          umma_commit(&bars->empty[stage]);  // frees the smem slot once these MMAs have read it
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&bars->acc_full[acc]);  // accumulator complete -> epilogue
        if (++acc == kAccStages) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue (4 warps, 128 TMEM lanes) =====================
    const int ew = warp & 3;  // TMEM lane quarter this warp may access
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      int tm, tn;
      tile_coords(t, p.tiles_m, p.tiles_n, tm, tn);
      const int row = tm * BM + ew * 32 + lane;
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
              float g0 = __uint_as_float(g[2 * i]), g1 = __uint_as_float(g[2 * i + 1]);
              float u0 = __uint_as_float(u[2 * i]), u1 = __uint_as_float(u[2 * i + 1]);
              // round the pre-activations to bf16 first so that forward == recompute-from-saved in backward
              pg[i] = pack_bf16(g0, g1);
              pu[i] = pack_bf16(u0, u1);
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
      if (lane == 0) mbar_arrive(&bars->acc_empty[acc]);
      if (++acc == kAccStages) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// -------------------------------------------------------------------------------------------------------
// host side
// -------------------------------------------------------------------------------------------------------
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

// 2-D bf16 tensor map: `inner` contiguous elements per row, `outer` rows, row pitch ld (elements), 128B swizzle
static int make_map_2d(CUtensorMap* m, const void* ptr, uint64_t inner, uint64_t outer, uint64_t ld,
                       uint32_t box_inner, uint32_t box_outer) {
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
  auto kern = gemm_kernel<AM, BMj, EPI, OutT>;
  static PerDeviceOnce once;
  if (once.first())
    SB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
  int tiles = p.tiles_m * p.tiles_n;
  int grid = tiles < num_sms() ? tiles : num_sms();
  if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
  kern<<<grid, kThreads, kSmemBytes, stream>>>(ta, tb, p);
  return (int)cudaGetLastError();
}

}  // namespace gemm
}  // namespace sb

using namespace sb::gemm;

// C ABI ------------------------------------------------------------------------------------------------
// a_layout / b_layout: 0 = K-major (A is [M,K] row-major, B is [N,K] row-major), 1 = MN-major (A is [K,M], B is [K,N])
// out_dtype: 0 = bf16, 1 = fp32.   epilogue: 0 store, 1 accumulate (D += A·Bᵀ)
extern "C" int sb_gemm_bf16(const void* A, const void* B, void* D, const void* bias, int M, int N, int K, int lda,
                            int ldb, int ldd, int a_layout, int b_layout, int out_dtype, int epilogue, int max_ctas,
                            cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  CUtensorMap ta, tb;
  int rc;
  if (a_layout == 0)
    rc = make_map_2d(&ta, A, K, M, lda, BK, BM);
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
#define SB_DISPATCH(AM, BMj)                                                                               \
  if (a_layout == AM && b_layout == BMj) {                                                                 \
    if (epilogue == 0 && out_dtype == 0) return launch<AM, BMj, kStore, __nv_bfloat16>(ta, tb, p, stream, max_ctas); \
    if (epilogue == 0 && out_dtype == 1) return launch<AM, BMj, kStore, float>(ta, tb, p, stream, max_ctas);        \
    if (epilogue == 1 && out_dtype == 0) return launch<AM, BMj, kAccumulate, __nv_bfloat16>(ta, tb, p, stream, max_ctas); \
    if (epilogue == 1 && out_dtype == 1) return launch<AM, BMj, kAccumulate, float>(ta, tb, p, stream, max_ctas);   \
  }
  SB_DISPATCH(0, 0)
  SB_DISPATCH(0, 1)
  SB_DISPATCH(1, 1)
  SB_DISPATCH(1, 0)
#undef SB_DISPATCH
  return -10;
}

// h[M,F] = silu(X·W1ᵀ) ⊙ (X·W2ᵀ) where W = [W1; W2] is one [2F, K] row-major buffer (gate rows then up rows).
// pre (optional, may be null): [M, 2F] bf16 pre-activations (gate | up) kept for the backward pass.
extern "C" int sb_gemm_swiglu_bf16(const void* X, const void* W, void* H, void* pre, int M, int F, int K, int ldx,
                                   int ldw, int ldh, int ldpre, int max_ctas, cudaStream_t stream) {
  if (M <= 0 || F <= 0 || K <= 0) return 0;
  CUtensorMap ta, tb;
  if (int rc = make_map_2d(&ta, X, K, M, ldx, BK, BM)) return 1000 + rc;
  if (int rc = make_map_2d(&tb, W, K, 2 * (uint64_t)F, ldw, BK, 128)) return 2000 + rc;
  Params p{};
  p.M = M, p.N = F, p.K = K, p.D = H, p.ldd = ldh, p.aux = pre, p.ldaux = ldpre, p.swiglu_up_row_offset = F;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (F + BN / 2 - 1) / (BN / 2);
  return launch<0, 0, kSwiGLU, __nv_bfloat16>(ta, tb, p, stream, max_ctas);
}
