// Symmetric-heap segments: device memory that every rank of the node maps.
// A segment is a plain cudaMalloc region exported through CUDA IPC; the
// 64-byte handle travels over the bootstrap process group (gloo) and each peer
// opens it once.  Peer access over NVLink is implied by cudaIpcOpenMemHandle
// with cudaIpcMemLazyEnablePeerAccess.
//
// Stands in for what ProcessGroupNCCL sets up behind init_process_group in the
// reference (/root/reference/src/distributed_trainer.py:62).
#include <cuda_runtime.h>
#include <cstring>

#include "../../include/dtb200.h"
#include "../common/host.h"

extern "C" int dtb_symm_alloc(size_t bytes, void** ptr_out, void* handle_out) {
  DTB_REQUIRE(bytes > 0 && ptr_out && handle_out, "dtb_symm_alloc: bad arguments");
  void* p = nullptr;
  DTB_CUDA_OK(cudaMalloc(&p, bytes));
  DTB_CUDA_OK(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    dtb::set_error("cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    return 1;
  }
  static_assert(sizeof(h) == 64, "IPC handle is 64 bytes");
  memcpy(handle_out, &h, sizeof(h));
  *ptr_out = p;
  DTB_CUDA_OK(cudaDeviceSynchronize());
  return 0;
}

extern "C" int dtb_symm_open(const void* handle, void** ptr_out) {
  DTB_REQUIRE(handle && ptr_out, "dtb_symm_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  DTB_CUDA_OK(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

extern "C" int dtb_symm_close(void* ptr) {
  if (ptr) DTB_CUDA_OK(cudaIpcCloseMemHandle(ptr));
  return 0;
}

extern "C" int dtb_symm_free(void* ptr) {
  if (ptr) DTB_CUDA_OK(cudaFree(ptr));
  return 0;
}
