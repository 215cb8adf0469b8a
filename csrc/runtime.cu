// Host-side runtime helpers of the kernel library.
#include <cuda_runtime.h>

// Bind `device`'s primary context to the calling thread.  The tensor-map encoder is a driver entry point and fails
// with CUDA_ERROR_INVALID_CONTEXT on a thread that has not touched the runtime yet — PyTorch's autograd worker threads
// set their device lazily, so under pipeline parallelism the first native call of a stage's backward can arrive on
// such a thread.  The ctypes binding calls this whenever the (thread, device) pair changes.
extern "C" int sb_bind_device(int device) {
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) return (int)e;
  return (int)cudaFree(0);  // forces the primary context to be current on runtimes that bind on first use
}

// ---------------------------------------------------------------------------------------------------------------
// Device-side barrier over a group of GPUs that share peer-mapped memory.
//
// Every rank owns `world` u32 flag words in symmetric memory; `flag_ptrs[r]` is rank r's flag array as mapped into
// this process.  One warp: lane r stores `epoch` into slot `rank` of rank r's array (after a system-scope fence, so
// everything earlier kernels of this stream wrote — locally or into peers — is visible before the flag), then spins
// until slot r of the OWN array shows `epoch`.  Kernels launched after it on the stream therefore see every peer's
// writes that preceded the peer's barrier.  `epoch` must grow by one per barrier on all ranks (compared as a signed
// difference, so wrap-around is harmless and a peer that is one barrier ahead does not hide this one).  A peer that never arrives traps after 60 s instead of hanging the GPU.
// Reference counterpart: none (torch.distributed collectives imply their own synchronisation).
// ---------------------------------------------------------------------------------------------------------------
namespace {
struct BarrierPtrs {
  unsigned int* p[8];
};
__global__ void device_barrier_kernel(BarrierPtrs flags, int rank, int world, unsigned int epoch) {
  const int r = threadIdx.x;
  __threadfence_system();
  if (r < world) {
    asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(flags.p[r] + rank), "r"(epoch) : "memory");
    const unsigned int* mine = flags.p[rank] + r;
    unsigned int v;
    unsigned long long t0 = 0;
    int spins = 0;
    for (;;) {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(mine) : "memory");
      if ((int)(v - epoch) >= 0) break;  // a fast peer may already have posted the NEXT epoch into this slot
      if (++spins > 32) {
        __nanosleep(200);
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(now));
        if (t0 == 0) t0 = now;
        if (now - t0 > 60000000000ull) __trap();
      }
    }
  }
  __syncwarp();
  __threadfence_system();
}
}  // namespace

extern "C" int sb_device_barrier(const unsigned long long* flag_ptrs, int rank, int world, unsigned int epoch,
                                 cudaStream_t st) {
  if (world < 1 || world > 8 || rank < 0 || rank >= world) return -5;
  BarrierPtrs f{};
  for (int r = 0; r < world; ++r) f.p[r] = reinterpret_cast<unsigned int*>(flag_ptrs[r]);
  device_barrier_kernel<<<1, 32, 0, st>>>(f, rank, world, epoch);
  return (int)cudaGetLastError();
}
