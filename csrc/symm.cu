// symm.cu — peer-visible device buffers for in-kernel NVLink traffic.
//
// The reference moves work and results between ranks with host-memory MPI
// collectives rooted at rank 0 (MPI_Scatter ref:636, MPI_Gather ref:889,
// MPI_Barrier ref:615).  Here each rank cudaMalloc's its exchange buffers,
// publishes a CUDA IPC handle, and maps every peer's buffer; kernels then
// load/store peer addresses directly (NVLink 5 through NVSwitch) and signal
// with system-scope release/acquire flags.  The launcher's process group is
// only used to pass the 64-byte handles around at init.
#include <cstring>

#include "errors.hpp"

using namespace b200gru;

extern "C" {

int b200gru_symm_alloc(int device, size_t bytes, void **ptr,
                       unsigned char *handle64) {
  B200_REQUIRE(ptr && handle64 && bytes > 0, "bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  B200_CUDA(cudaSetDevice(device));
  void *p = nullptr;
  B200_CUDA(cudaMalloc(&p, bytes));
  B200_CUDA(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return fail(B200GRU_ERR_CUDA, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  }
  memcpy(handle64, &h, 64);
  B200_CUDA(cudaDeviceSynchronize());
  *ptr = p;
  return 0;
}

int b200gru_symm_open(int device, const unsigned char *handle64, void **ptr) {
  B200_REQUIRE(ptr && handle64, "bad arguments");
  B200_CUDA(cudaSetDevice(device));
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void *p = nullptr;
  B200_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr = p;
  return 0;
}

int b200gru_symm_close(int device, void *ptr) {
  if (!ptr) return 0;
  B200_CUDA(cudaSetDevice(device));
  B200_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}

int b200gru_symm_free(int device, void *ptr) {
  if (!ptr) return 0;
  B200_CUDA(cudaSetDevice(device));
  B200_CUDA(cudaFree(ptr));
  return 0;
}

// kind: 1 = host->device, 2 = device->host, 3 = device->device
int b200gru_memcpy_async(void *dst, const void *src, size_t bytes, int kind,
                         void *stream) {
  cudaMemcpyKind k = kind == 1   ? cudaMemcpyHostToDevice
                     : kind == 2 ? cudaMemcpyDeviceToHost
                                 : cudaMemcpyDeviceToDevice;
  B200_CUDA(cudaMemcpyAsync(dst, src, bytes, k, static_cast<cudaStream_t>(stream)));
  return 0;
}

int b200gru_memset_async(void *dst, int value, size_t bytes, void *stream) {
  B200_CUDA(cudaMemsetAsync(dst, value, bytes, static_cast<cudaStream_t>(stream)));
  return 0;
}

__global__ void signal_flag_kernel(uint32_t *flag, uint32_t value) {
  __threadfence_system();
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
}

// Release-store `value` to a (possibly peer) flag once prior work on `stream`
// is done: the root uses it to publish "random floats are in place".
int b200gru_signal_flag(uint32_t *flag, uint32_t value, void *stream) {
  B200_REQUIRE(flag, "null flag");
  signal_flag_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(flag, value);
  B200_CUDA(cudaGetLastError());
  return 0;
}

int b200gru_can_access_peer(int device, int peer) {
  int ok = 0;
  B200_CUDA(cudaDeviceCanAccessPeer(&ok, device, peer));
  return ok;
}

}  // extern "C"
