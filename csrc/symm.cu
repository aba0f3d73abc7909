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

// ------------------------------------------------------------ diagnostics
// Store `rows` rows of `row_bytes` bytes (row pitch `pitch`) to up to 8
// destination buffers (typically peer-mapped) from `src`, with 16-byte
// stores, every warp covering contiguous bytes of a row.  Used by
// tools/p2p_bench.py to measure NVLink write bandwidth for the access pattern
// the tensor-parallel epilogue produces.
struct P2PDst { unsigned char *p[8]; };
__global__ void p2p_write_kernel(P2PDst dst, int ndst, const unsigned char *src, long long rows,
                                 int row_bytes, long long pitch) {
  const int chunks_per_row = row_bytes / 16;
  const long long total = rows * chunks_per_row;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / chunks_per_row;
    const int c = (int)(i % chunks_per_row);
    const uint4 v = *reinterpret_cast<const uint4 *>(src + r * pitch + c * 16);
    for (int d = 0; d < ndst; ++d) *reinterpret_cast<uint4 *>(dst.p[d] + r * pitch + c * 16) = v;
  }
}

extern "C" int b200gru_p2p_write_bench(void *const *dsts, int ndst, const void *src, long long rows,
                                       int row_bytes, long long pitch, int blocks, void *stream) {
  B200_REQUIRE(dsts && src && ndst >= 1 && ndst <= 8 && row_bytes % 16 == 0, "bad arguments");
  P2PDst d{};
  for (int i = 0; i < ndst; ++i) d.p[i] = static_cast<unsigned char *>(dsts[i]);
  p2p_write_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      d, ndst, static_cast<const unsigned char *>(src), rows, row_bytes, pitch);
  B200_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int b200gru_stream_synchronize(void *stream) {
  B200_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int b200gru_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}
