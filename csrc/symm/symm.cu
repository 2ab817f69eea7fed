// Symmetric (peer-addressable) device memory for tensor-parallel groups on one NVSwitch domain.
//
// Every rank allocates the same-sized buffer with cudaMalloc, exports a CUDA IPC handle, and maps the
// handles of its peers; kernels then read and write peer buffers with ordinary loads / stores that travel
// over NVLink, synchronised by flag words that live in the same kind of memory.  The reference has no
// equivalent - its tensor-parallel traffic is NCCL collectives issued from Python (reference
// smdistributed/modelparallel/torch/nn/utils.py:701-821); its only peer-memory path is the pipeline "D2D"
// transport inside the absent smplib.so (receiver-owned pool written over CUDA IPC).
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>

#define SMPC_API extern "C" __attribute__((visibility("default")))

namespace {
thread_local char g_err[512] = {0};
int fail(const char* what, cudaError_t e) {
  snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
  return 1;
}
}  // namespace

SMPC_API const char* smpc_symm_last_error() { return g_err; }

// Allocate `bytes` of zeroed device memory on the current device and export its IPC handle (64 bytes).
SMPC_API int smpc_symm_alloc(long long bytes, void** dev_ptr, unsigned char* handle_out) {
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, (size_t)bytes);
  if (e != cudaSuccess) return fail("cudaMalloc", e);
  e = cudaMemset(p, 0, (size_t)bytes);
  if (e != cudaSuccess) return fail("cudaMemset", e);
  cudaIpcMemHandle_t h;
  e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) return fail("cudaIpcGetMemHandle", e);
  static_assert(sizeof(h) == 64, "IPC handle size");
  memcpy(handle_out, &h, sizeof(h));
  *dev_ptr = p;
  return 0;
}

// Map a peer's buffer into this process (enables peer access on first use).
SMPC_API int smpc_symm_open(const unsigned char* handle, void** peer_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return fail("cudaIpcOpenMemHandle", e);
  *peer_ptr = p;
  return 0;
}

SMPC_API int smpc_symm_close(void* peer_ptr) {
  cudaError_t e = cudaIpcCloseMemHandle(peer_ptr);
  return e == cudaSuccess ? 0 : fail("cudaIpcCloseMemHandle", e);
}

SMPC_API int smpc_symm_free(void* dev_ptr) {
  cudaError_t e = cudaFree(dev_ptr);
  return e == cudaSuccess ? 0 : fail("cudaFree", e);
}

// 1 when `peer_device` can be addressed from the current device.
SMPC_API int smpc_symm_can_access_peer(int peer_device) {
  int dev = 0, ok = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (dev == peer_device) return 1;
  if (cudaDeviceCanAccessPeer(&ok, dev, peer_device) != cudaSuccess) return 0;
  return ok;
}
