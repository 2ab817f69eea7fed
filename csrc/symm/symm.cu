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

// In-library self-test of the symmetric-memory path on the current device (the counterpart of the reference's
// native test_d2d()): allocate two buffers, check both export an IPC handle, move a pattern
// host -> buffer A -> buffer B -> host and compare.  0 on success; otherwise non-zero and smpc_symm_last_error().
SMPC_API int smpc_symm_selftest(long long bytes) {
  if (bytes < 64) bytes = 64;
  bytes &= ~63LL;
  void *a = nullptr, *b = nullptr;
  unsigned char ha[64], hb[64];
  if (smpc_symm_alloc(bytes, &a, ha)) return 1;
  if (smpc_symm_alloc(bytes, &b, hb)) { cudaFree(a); return 2; }
  int rc = 0;
  unsigned char* host = new unsigned char[(size_t)bytes * 2];
  for (long long i = 0; i < bytes; ++i) host[i] = (unsigned char)((i * 131 + 7) & 0xff);
  cudaError_t e = cudaMemcpy(a, host, (size_t)bytes, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(b, a, (size_t)bytes, cudaMemcpyDeviceToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(host + bytes, b, (size_t)bytes, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) rc = 2 + fail("self-test copy", e);
  if (!rc && memcmp(host, host + bytes, (size_t)bytes) != 0) {
    snprintf(g_err, sizeof(g_err), "self-test: pattern mismatch after device round trip");
    rc = 4;
  }
  if (!rc && memcmp(ha, hb, 64) == 0) {
    snprintf(g_err, sizeof(g_err), "self-test: two allocations exported the same IPC handle");
    rc = 5;
  }
  delete[] host;
  if (smpc_symm_free(a) && !rc) rc = 6;
  if (smpc_symm_free(b) && !rc) rc = 7;
  return rc;
}
