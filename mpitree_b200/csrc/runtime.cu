// Error reporting, device queries and the peer-memory (CUDA IPC) bootstrap helpers.
// The reference's bootstrap is MPI.COMM_WORLD at import time
// (mpitree/tree/decision_tree.py:313-317); here torch.distributed exchanges the IPC
// handles produced below and kernels then address peers directly.
#include <stdio.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void mtb_set_error(const char* what, const char* why, const char* file, int line) {
  snprintf(g_err, sizeof(g_err), "%s: %s (%s:%d)", what, why, file, line);
}

extern "C" const char* mtb_last_error(void) { return g_err; }
extern "C" int mtb_version(void) { return 100; }

extern "C" int mtb_sm_count(int device) {
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return -1;
  return n;
}

extern "C" int mtb_set_device(int device) {
  MTB_CHECK(cudaSetDevice(device));
  return 0;
}

extern "C" int mtb_ipc_handle_size(void) { return (int)sizeof(cudaIpcMemHandle_t); }

extern "C" int mtb_ipc_get_handle(void* dev_ptr, void* handle_out) {
  cudaIpcMemHandle_t h;
  MTB_CHECK(cudaIpcGetMemHandle(&h, dev_ptr));
  memcpy(handle_out, &h, sizeof(h));
  return 0;
}

extern "C" int mtb_ipc_open_handle(const void* handle, void** dev_ptr_out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  MTB_CHECK(cudaIpcOpenMemHandle(dev_ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

extern "C" int mtb_ipc_close(void* dev_ptr) {
  MTB_CHECK(cudaIpcCloseMemHandle(dev_ptr));
  return 0;
}

extern "C" int mtb_enable_peer_access(int peer_device) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  MTB_CHECK(e);
  return 0;
}

// ---- raw device allocations for symmetric (peer-mapped) regions -------------------------
// cudaIpcGetMemHandle needs the base of a cudaMalloc allocation, which a sub-allocating pool
// (torch's caching allocator) does not give; these two calls are the only allocating entry
// points of the library and are used once per builder for the peer-visible region.
extern "C" int mtb_device_malloc(int64_t bytes, void** out) {
  MTB_CHECK(cudaMalloc(out, (size_t)bytes));
  return 0;
}

extern "C" int mtb_device_free(void* ptr) {
  MTB_CHECK(cudaFree(ptr));
  return 0;
}

extern "C" int mtb_memset_async(void* ptr, int value, int64_t bytes, void* stream) {
  MTB_CHECK(cudaMemsetAsync(ptr, value, (size_t)bytes, (cudaStream_t)stream));
  return 0;
}

// Device-to-device copy between any two mapped addresses (own memory or a peer's region opened with
// mtb_ipc_open_handle); the driver routes peer addresses over NVLink.
extern "C" int mtb_memcpy_async(void* dst, const void* src, int64_t bytes, void* stream) {
  if (bytes <= 0) return 0;
  MTB_CHECK(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, (cudaStream_t)stream));
  return 0;
}

extern "C" int mtb_can_access_peer(int device, int peer) {
  int ok = 0;
  if (cudaDeviceCanAccessPeer(&ok, device, peer) != cudaSuccess) return -1;
  return ok;
}
