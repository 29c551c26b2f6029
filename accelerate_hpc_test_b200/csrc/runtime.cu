// Library runtime: launch accounting, device queries, CUDA-IPC symmetric buffers (the data plane
// of the peer-memory collectives) and small utilities.  The reference has no counterpart: its
// data plane is whatever torch.distributed/NCCL allocates.
#include "common.cuh"
#include <string.h>

unsigned long long g_b200_launch_count = 0;

B200_API unsigned long long b200_launch_count() { return g_b200_launch_count; }
B200_API void b200_reset_launch_count() { g_b200_launch_count = 0; }

B200_API int b200_version() { return 100; }

B200_API int b200_device_info(int* sm_count, int* cc_major, int* cc_minor, long long* total_mem) {
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  B200_CUDA(cudaGetDeviceProperties(&prop, dev));
  if (sm_count) *sm_count = prop.multiProcessorCount;
  if (cc_major) *cc_major = prop.major;
  if (cc_minor) *cc_minor = prop.minor;
  if (total_mem) *total_mem = (long long)prop.totalGlobalMem;
  return 0;
}

B200_API const char* b200_error_string(int code) {
  if (code == 0) return "success";
  if (code == B200_ERR_BAD_ARG) return "b200: bad argument (shape/alignment not supported by this kernel)";
  if (code == B200_ERR_UNSUPPORTED) return "b200: unsupported configuration";
  if (code == B200_ERR_TOO_LARGE) return "b200: size exceeds the built-in limit of this entry point";
  if (code == B200_ERR_DRIVER) return "b200: driver entry point / tensor-map encode failed";
  if (code > 0) return cudaGetErrorString((cudaError_t)code);
  return "b200: unknown error";
}

// ------------------------------------------------------------------ symmetric (IPC) buffers
// Each rank cudaMalloc's one buffer, exports a 64-byte IPC handle, and opens every peer's handle.
// The kernels in comm.cu then address peers with plain ld/st over NVLink.
B200_API int b200_symm_alloc(void** ptr, long long bytes) {
  B200_CUDA(cudaMalloc(ptr, (size_t)bytes));
  B200_CUDA(cudaMemset(*ptr, 0, (size_t)bytes));
  return 0;
}
B200_API int b200_symm_free(void* ptr) {
  B200_CUDA(cudaFree(ptr));
  return 0;
}
B200_API int b200_ipc_get_handle(void* ptr, void* handle64) {
  cudaIpcMemHandle_t h;
  B200_CUDA(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle64, &h, sizeof(h));
  return 0;
}
B200_API int b200_ipc_open_handle(const void* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  B200_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
B200_API int b200_ipc_close_handle(void* ptr) {
  B200_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}

B200_API int b200_memset_async(void* ptr, int value, long long bytes, void* stream) {
  B200_CUDA(cudaMemsetAsync(ptr, value, (size_t)bytes, (cudaStream_t)stream));
  return 0;
}

B200_API int b200_memcpy_async(void* dst, const void* src, long long bytes, void* stream) {
  B200_CUDA(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, (cudaStream_t)stream));
  return 0;
}
