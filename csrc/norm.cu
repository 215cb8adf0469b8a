// RMSNorm / LayerNorm forward + backward for sm_100a.
//
// Replaces the reference's RMSNorm (≈6 ATen kernels with an fp32 round trip, core/nn/norm/rms_norm.py:45-53) and
// F.layer_norm (core/nn/norm/layernorm.py:72-79) with single-pass kernels: bf16/fp32 I/O, fp32 statistics, 128-bit
// loads, optional fused residual add (h = x + r is written once and normalised in the same pass), and a weight/bias
// gradient that is accumulated per CTA in shared memory, spilled to a [grid, cols] fp32 scratch and column-reduced
// straight into the (fp32 or bf16) main-grad buffer.
#include "common.cuh"

namespace sb {
namespace norm {

constexpr int kThreads = 256;

template <typename T>
struct Vec8;
template <>
struct Vec8<__nv_bfloat16> {
  static SB_DEVICE void load(const __nv_bfloat16* p, float* f) {
    uint4 v = *reinterpret_cast<const uint4*>(p);
    float2 a = unpack_bf16(v.x), b = unpack_bf16(v.y), c = unpack_bf16(v.z), d = unpack_bf16(v.w);
    f[0] = a.x, f[1] = a.y, f[2] = b.x, f[3] = b.y, f[4] = c.x, f[5] = c.y, f[6] = d.x, f[7] = d.y;
  }
  static SB_DEVICE void store(__nv_bfloat16* p, const float* f) {
    *reinterpret_cast<uint4*>(p) =
        make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
  }
  static SB_DEVICE float round(float x) { return __bfloat162float(__float2bfloat16(x)); }
};
template <>
struct Vec8<float> {
  static SB_DEVICE void load(const float* p, float* f) {
    float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
    f[0] = a.x, f[1] = a.y, f[2] = a.z, f[3] = a.w, f[4] = b.x, f[5] = b.y, f[6] = b.z, f[7] = b.w;
  }
  static SB_DEVICE void store(float* p, const float* f) {
    reinterpret_cast<float4*>(p)[0] = make_float4(f[0], f[1], f[2], f[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(f[4], f[5], f[6], f[7]);
  }
  static SB_DEVICE float round(float x) { return x; }
};
template <>
struct Vec8<__half> {
  static SB_DEVICE void load(const __half* p, float* f) {
    uint4 v = *reinterpret_cast<const uint4*>(p);
    const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 t = __half22float2(h[i]);
      f[2 * i] = t.x, f[2 * i + 1] = t.y;
    }
  }
  static SB_DEVICE void store(__half* p, const float* f) {
    uint4 v;
    __half2* h = reinterpret_cast<__half2*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = v;
  }
  static SB_DEVICE float round(float x) { return __half2float(__float2half(x)); }
};

SB_DEVICE float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.f;
  if (w == 0) {
    t = warp_sum(t);
    if (l == 0) red[0] = t;
  }
  __syncthreads();
  return red[0];
}

// y = norm(x [+ res]) * w (+ b);  kRms: no mean subtraction, no bias.  sum_out (optional) receives x + res.
template <typename T, bool kRms>
__global__ void __launch_bounds__(kThreads) norm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ res,
                                                            const T* __restrict__ w, const T* __restrict__ b,
                                                            T* __restrict__ y, T* __restrict__ sum_out,
                                                            float* __restrict__ mean_out, float* __restrict__ rstd_out,
                                                            int rows, int cols, float eps) {
  __shared__ float red[32];
  extern __shared__ float row_cache[];  // cols floats
  for (int r = blockIdx.x; r < rows; r += gridDim.x) {
    const T* xr = x + (size_t)r * cols;
    float s = 0.f, ss = 0.f;
    for (int c = threadIdx.x * 8; c < cols; c += kThreads * 8) {
      float f[8];
      Vec8<T>::load(xr + c, f);
      if (res) {
        float g[8];
        Vec8<T>::load(res + (size_t)r * cols + c, g);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = Vec8<T>::round(f[i] + g[i]);
        if (sum_out) Vec8<T>::store(sum_out + (size_t)r * cols + c, f);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        row_cache[c + i] = f[i];
        s += f[i];
        ss += f[i] * f[i];
      }
    }
    float mean = 0.f, rstd;
    if (kRms) {
      ss = block_sum(ss, red);
      rstd = rsqrtf(ss / cols + eps);
    } else {
      s = block_sum(s, red);
      mean = s / cols;
      float var = 0.f;
      for (int c = threadIdx.x * 8; c < cols; c += kThreads * 8) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float d = row_cache[c + i] - mean;
          var += d * d;
        }
      }
      var = block_sum(var, red);
      rstd = rsqrtf(var / cols + eps);
    }
    if (threadIdx.x == 0) {
      if (mean_out) mean_out[r] = mean;
      rstd_out[r] = rstd;
    }
    T* yr = y + (size_t)r * cols;
    for (int c = threadIdx.x * 8; c < cols; c += kThreads * 8) {
      float wv[8], o[8];
      Vec8<T>::load(w + c, wv);
      if (kRms) {
        // reference order: normalise in fp32, cast to the model dtype, then scale by the weight
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = Vec8<T>::round(row_cache[c + i] * rstd) * wv[i];
      } else {
        float bv[8];
        if (b) Vec8<T>::load(b + c, bv);
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = (row_cache[c + i] - mean) * rstd * wv[i] + (b ? bv[i] : 0.f);
      }
      Vec8<T>::store(yr + c, o);
    }
    __syncthreads();
  }
}

// dx (+= dres) and per-CTA partial dw/db. x is the normalised layer's input (after the residual add).
template <typename T, bool kRms>
__global__ void __launch_bounds__(kThreads) norm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                            const T* __restrict__ w, const float* __restrict__ mean_in,
                                                            const float* __restrict__ rstd_in,
                                                            const T* __restrict__ dres, T* __restrict__ dx,
                                                            float* __restrict__ partial_dw, float* __restrict__ partial_db,
                                                            int rows, int cols) {
  __shared__ float red[32];
  extern __shared__ float sm[];  // [cols] dw accumulators, [cols] db accumulators (LayerNorm only)
  float* acc_w = sm;
  float* acc_b = sm + cols;
  for (int c = threadIdx.x; c < cols; c += kThreads) {
    acc_w[c] = 0.f;
    if (!kRms) acc_b[c] = 0.f;
  }
  __syncthreads();
  for (int r = blockIdx.x; r < rows; r += gridDim.x) {
    const T* xr = x + (size_t)r * cols;
    const T* dyr = dy + (size_t)r * cols;
    const float rstd = rstd_in[r];
    const float mean = kRms ? 0.f : mean_in[r];
    float s1 = 0.f, s2 = 0.f;  // Σ g·xhat, Σ g  with g = dy·w
    for (int c = threadIdx.x * 8; c < cols; c += kThreads * 8) {
      float xv[8], dv[8], wv[8];
      Vec8<T>::load(xr + c, xv);
      Vec8<T>::load(dyr + c, dv);
      Vec8<T>::load(w + c, wv);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float xhat = (xv[i] - mean) * rstd;
        float g = dv[i] * wv[i];
        s1 += g * xhat;
        s2 += g;
        acc_w[c + i] += dv[i] * (kRms ? Vec8<T>::round(xhat) : xhat);
        if (!kRms) acc_b[c + i] += dv[i];
      }
    }
    s1 = block_sum(s1, red) / cols;
    if (!kRms) s2 = block_sum(s2, red) / cols;
    T* dxr = dx + (size_t)r * cols;
    for (int c = threadIdx.x * 8; c < cols; c += kThreads * 8) {
      float xv[8], dv[8], wv[8], o[8];
      Vec8<T>::load(xr + c, xv);
      Vec8<T>::load(dyr + c, dv);
      Vec8<T>::load(w + c, wv);
      float rv[8];
      if (dres) Vec8<T>::load(dres + (size_t)r * cols + c, rv);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float xhat = (xv[i] - mean) * rstd;
        float g = dv[i] * wv[i];
        o[i] = rstd * (g - xhat * s1 - (kRms ? 0.f : s2));
        if (dres) o[i] += rv[i];
      }
      Vec8<T>::store(dxr + c, o);
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < cols; c += kThreads) {
    partial_dw[(size_t)blockIdx.x * cols + c] = acc_w[c];
    if (!kRms) partial_db[(size_t)blockIdx.x * cols + c] = acc_b[c];
  }
}

// column-reduce the per-CTA partials; accumulate==1 adds to the existing gradient (main-grad semantics).
// One CTA owns 32 adjacent columns: lane = column (128-byte coalesced rows), the 32 warps stride over the partial rows,
// then a fixed-order shared-memory tree sums the 32 warp totals, so the result is deterministic and the `parts`
// (up to 592) loads per column are spread over 32 warps instead of one serial loop.
constexpr int kReduceWarps = 32;
template <typename G>
__global__ void __launch_bounds__(32 * kReduceWarps)
    reduce_partials_kernel(const float* __restrict__ partial, G* __restrict__ grad, int parts, int cols, int accumulate) {
  __shared__ float tile[kReduceWarps][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float s = 0.f;
  if (c < cols) {
#pragma unroll 4
    for (int p = warp; p < parts; p += kReduceWarps) s += partial[(size_t)p * cols + c];
  }
  tile[warp][lane] = s;
  __syncthreads();
  if (warp == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kReduceWarps; ++w) t += tile[w][lane];
    if (accumulate) t += (float)grad[c];
    grad[c] = (G)t;
  }
}

// launches the weight (and, for LayerNorm, bias) gradient reductions over `grid` partial rows
template <typename T>
static void launch_reduce(const float* pw, const float* pb, void* dw, void* db, int grid, int cols, int rms,
                          int grad_fp32, int accumulate, cudaStream_t st) {
  const int rb = (cols + 31) / 32, th = 32 * kReduceWarps;
  if (grad_fp32) {
    reduce_partials_kernel<float><<<rb, th, 0, st>>>(pw, (float*)dw, grid, cols, accumulate);
    if (!rms && db) reduce_partials_kernel<float><<<rb, th, 0, st>>>(pb, (float*)db, grid, cols, accumulate);
  } else {
    reduce_partials_kernel<T><<<rb, th, 0, st>>>(pw, (T*)dw, grid, cols, accumulate);
    if (!rms && db) reduce_partials_kernel<T><<<rb, th, 0, st>>>(pb, (T*)db, grid, cols, accumulate);
  }
}

static int grid_for(int rows) {
  int n = 148 * 4;
  return rows < n ? rows : n;
}

template <typename T>
static int fwd(const void* x, const void* res, const void* w, const void* b, void* y, void* sum_out, float* mean,
               float* rstd, int rows, int cols, float eps, int rms, cudaStream_t st) {
  if (cols % 8) return -1;
  size_t smem = (size_t)cols * 4;
  auto k = rms ? norm_fwd_kernel<T, true> : norm_fwd_kernel<T, false>;
  if (smem + 1024 > 48 * 1024) SB_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k<<<rows, kThreads, smem, st>>>((const T*)x, (const T*)res, (const T*)w, (const T*)b, (T*)y, (T*)sum_out, mean, rstd,
                                 rows, cols, eps);
  return (int)cudaGetLastError();
}

template <typename T>
static int bwd(const void* dy, const void* x, const void* w, const float* mean, const float* rstd, const void* dres,
               void* dx, void* dw, void* db, int grad_fp32, int accumulate, float* scratch, int rows, int cols, int rms,
               cudaStream_t st) {
  if (cols % 8) return -1;
  const int grid = grid_for(rows);
  size_t smem = (size_t)cols * 4 * (rms ? 1 : 2);
  auto k = rms ? norm_bwd_kernel<T, true> : norm_bwd_kernel<T, false>;
  if (smem + 1024 > 48 * 1024) SB_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  float* pw = scratch;
  float* pb = scratch + (size_t)grid * cols;
  k<<<grid, kThreads, smem, st>>>((const T*)dy, (const T*)x, (const T*)w, mean, rstd, (const T*)dres, (T*)dx, pw, pb,
                                 rows, cols);
  launch_reduce<T>(pw, pb, dw, db, grid, cols, rms, grad_fp32, accumulate, st);
  return (int)cudaGetLastError();
}

}  // namespace norm
}  // namespace sb

// dtype: 0 = bf16, 1 = fp32, 2 = fp16
extern "C" int sb_norm_fwd(const void* x, const void* res, const void* w, const void* b, void* y, void* sum_out,
                           float* mean, float* rstd, int rows, int cols, float eps, int rms, int dtype,
                           cudaStream_t st) {
  using namespace sb::norm;
  if (rows <= 0) return 0;
  if (dtype == 0) return fwd<__nv_bfloat16>(x, res, w, b, y, sum_out, mean, rstd, rows, cols, eps, rms, st);
  if (dtype == 1) return fwd<float>(x, res, w, b, y, sum_out, mean, rstd, rows, cols, eps, rms, st);
  if (dtype == 2) return fwd<__half>(x, res, w, b, y, sum_out, mean, rstd, rows, cols, eps, rms, st);
  return -2;
}

// scratch: fp32 buffer of sb_norm_bwd_scratch_floats(rows, cols, rms) elements (global memory owned by the caller)
extern "C" long long sb_norm_bwd_scratch_floats(int rows, int cols, int rms) {
  return (long long)sb::norm::grid_for(rows) * cols * (rms ? 1 : 2);
}

extern "C" int sb_norm_bwd(const void* dy, const void* x, const void* w, const float* mean, const float* rstd,
                           const void* dres, void* dx, void* dw, void* db, int grad_fp32, int accumulate,
                           float* scratch, int rows, int cols, int rms, int dtype, cudaStream_t st) {
  using namespace sb::norm;
  if (rows <= 0) return 0;
  if (dtype == 0)
    return bwd<__nv_bfloat16>(dy, x, w, mean, rstd, dres, dx, dw, db, grad_fp32, accumulate, scratch, rows, cols, rms, st);
  if (dtype == 1)
    return bwd<float>(dy, x, w, mean, rstd, dres, dx, dw, db, grad_fp32, accumulate, scratch, rows, cols, rms, st);
  if (dtype == 2)
    return bwd<__half>(dy, x, w, mean, rstd, dres, dx, dw, db, grad_fp32, accumulate, scratch, rows, cols, rms, st);
  return -2;
}

// ---------------------------------------------------------------------------------------------------------------
// Norm backward for callers that must not allocate.
//
// Who this is for: code that runs where device allocation is forbidden or impossible — a step that is being
// recorded into a CUDA graph, a real-time serving loop that has handed all remaining memory to a KV cache, a
// plugin running inside another framework's allocator lock, or a C caller that simply has no allocator to offer.
// Such callers cannot give sb_norm_bwd the fp32 scratch it needs for the per-CTA weight-gradient partials, so
// this variant brings its own: a fixed array that lives in the library image.  Every GPU that loads the
// library carries these kNoAllocScratchFloats * 4 bytes (148 * 8192 * 4 = 4 849 664 B, 4.625 MiB) until the process exits.
//
// The price of a fixed array is a fixed limit: the launch uses at most kNoAllocGrid CTAs, and
// grid * cols * (rms ? 1 : 2) floats must fit, i.e. cols <= 8192 for RMSNorm and cols <= 4096 for LayerNorm.
// Larger problems are refused with kNoAllocTooLarge (see sb_norm_bwd_noalloc_strerror) — nothing is launched.
//
// In host code the name `g_noalloc_scratch` is only a handle for the symbol, not a pointer the GPU can use, so the
// device address is looked up with cudaGetSymbolAddress once per device (with that device current) and cached.
// That lookup happens inside the first call made on each device; make that first call outside of any stream
// capture.  The address is resolved for the CURRENT device: the device that is current at the call must be the one
// that owns `st` and the buffers.  All calls share the one array and it keeps whatever the previous call left in it.  Nothing in it has
// to start at zero — every slot the reduction reads was written earlier in the same call — but two calls that
// overlap on the GPU (different streams, no dependency) would overwrite each other's partials: serialise them.
// ---------------------------------------------------------------------------------------------------------------
namespace sb {
namespace norm {

constexpr int kNoAllocGrid = 148;                       // one CTA per SM
constexpr int kNoAllocMaxColsRms = 8192;
constexpr size_t kNoAllocScratchFloats = (size_t)kNoAllocGrid * kNoAllocMaxColsRms;
constexpr int kNoAllocTooLarge = -3;
constexpr int kNoAllocBadArgument = -1;
constexpr int kNoAllocBadDtype = -2;
constexpr int kNoAllocSetupFailed = -4;

static __device__ __align__(256) float g_noalloc_scratch[kNoAllocScratchFloats];

constexpr int kMaxDevices = 64;
static float* g_noalloc_scratch_ptr[kMaxDevices] = {nullptr};

// device address of the static scratch on the current device; resolved on first use per device
static int noalloc_scratch(float** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kMaxDevices) return kNoAllocSetupFailed;
  if (g_noalloc_scratch_ptr[dev] == nullptr) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess) return (int)e;
    if (p == nullptr) return kNoAllocSetupFailed;
    g_noalloc_scratch_ptr[dev] = p;
  }
  *out = g_noalloc_scratch_ptr[dev];
  return 0;
}

template <typename T>
static int bwd_noalloc(const void* dy, const void* x, const void* w, const float* mean, const float* rstd,
                       const void* dres, void* dx, void* dw, void* db, int grad_fp32, int accumulate, int rows, int cols,
                       int rms, cudaStream_t st) {
  const int grid = rows < kNoAllocGrid ? rows : kNoAllocGrid;
  float* scratch = nullptr;
  if (int rc = noalloc_scratch(&scratch)) return rc;
  size_t smem = (size_t)cols * 4 * (rms ? 1 : 2);
  auto k = rms ? norm_bwd_kernel<T, true> : norm_bwd_kernel<T, false>;
  if (smem + 1024 > 48 * 1024) SB_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  float* pw = scratch;
  float* pb = scratch + (size_t)grid * cols;
  k<<<grid, kThreads, smem, st>>>((const T*)dy, (const T*)x, (const T*)w, mean, rstd, (const T*)dres, (T*)dx, pw, pb,
                                 rows, cols);
  launch_reduce<T>(pw, pb, dw, db, grid, cols, rms, grad_fp32, accumulate, st);
  return (int)cudaGetLastError();
}

}  // namespace norm
}  // namespace sb

// Largest `cols` the no-allocation variant accepts (8192 for RMSNorm, 4096 for LayerNorm).
extern "C" int sb_norm_bwd_noalloc_max_cols(int rms) {
  return rms ? sb::norm::kNoAllocMaxColsRms : sb::norm::kNoAllocMaxColsRms / 2;
}

// Human-readable text for the negative status codes of sb_norm_bwd_noalloc.
extern "C" const char* sb_norm_bwd_noalloc_strerror(int code) {
  using namespace sb::norm;
  switch (code) {
    case 0: return "success";
    case kNoAllocBadArgument: return "sb_norm_bwd_noalloc: cols must be a positive multiple of 8";
    case kNoAllocBadDtype: return "sb_norm_bwd_noalloc: dtype must be 0 (bf16), 1 (fp32) or 2 (fp16)";
    case kNoAllocTooLarge:
      return "sb_norm_bwd_noalloc: problem too large for the built-in scratch (cols <= 8192 for RMSNorm, <= 4096 for "
             "LayerNorm); use sb_norm_bwd with caller-provided scratch instead";
    case kNoAllocSetupFailed: return "sb_norm_bwd_noalloc: could not resolve the device address of the built-in scratch";
    default: return code > 0 ? "CUDA runtime error (value is a cudaError_t)" : "unknown status";
  }
}

// Same contract and results as sb_norm_bwd, minus the `scratch` argument: inputs and outputs are the caller's, the
// fp32 partial buffer is the library's own static array.  Returns 0, a cudaError_t (> 0) or one of the negative codes
// explained by sb_norm_bwd_noalloc_strerror.  Never calls cudaMalloc/cudaFree or any allocating routine.
extern "C" int sb_norm_bwd_noalloc(const void* dy, const void* x, const void* w, const float* mean, const float* rstd,
                                   const void* dres, void* dx, void* dw, void* db, int grad_fp32, int accumulate,
                                   int rows, int cols, int rms, int dtype, cudaStream_t st) {
  using namespace sb::norm;
  if (rows <= 0) return 0;  // nothing to do, like sb_norm_bwd
  if (cols <= 0 || cols % 8) return kNoAllocBadArgument;
  if (cols > sb_norm_bwd_noalloc_max_cols(rms)) return kNoAllocTooLarge;
  if (dtype == 0)
    return bwd_noalloc<__nv_bfloat16>(dy, x, w, mean, rstd, dres, dx, dw, db, grad_fp32, accumulate, rows, cols, rms, st);
  if (dtype == 1)
    return bwd_noalloc<float>(dy, x, w, mean, rstd, dres, dx, dw, db, grad_fp32, accumulate, rows, cols, rms, st);
  if (dtype == 2)
    return bwd_noalloc<__half>(dy, x, w, mean, rstd, dres, dx, dw, db, grad_fp32, accumulate, rows, cols, rms, st);
  return kNoAllocBadDtype;
}
