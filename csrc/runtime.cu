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
