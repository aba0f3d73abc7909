// gru_cell_noalloc.cu — b200gru_gru_cell_fp32_noalloc: the fp32 GRU layer step
// for callers that must not allocate device memory when they call it.
//
// Who that is: a process whose GPU memory is already fully reserved by another
// owner (a serving framework that takes all of HBM for its own pool leaves
// nothing for a cudaMalloc inside a library call), and a latency-bound loop
// that cannot take the device-wide synchronisation cudaMalloc / cudaFree
// imply.  b200gru_gru_cell_fp32 does both on every call (it allocates
// 2*batch*3*hidden floats for the two gate GEMM outputs and frees them).
//
// This variant keeps that scratch in a fixed array that is part of the loaded
// library image on every GPU, so the call itself allocates nothing.  The price
// is a hard size limit and one shared buffer:
//   * 2*batch*3*hidden floats must fit in B200GRU_NOALLOC_SCRATCH_BYTES
//     (4 MiB, e.g. batch 128 at hidden 1024); larger calls are refused.
//   * every call on a GPU uses the same array, so two calls that can overlap
//     on that GPU (different streams, different threads) race.  Serialise them.
//   * the array keeps whatever the previous call left in it.  Nothing in this
//     op needs zeroed scratch — both GEMMs overwrite every element they later
//     read — so no clearing is done.
//   * the first call on each device looks the array's device address up
//     (cudaGetSymbolAddress; with CUDA's lazy module loading this is also the
//     moment the 4 MiB array appears on that GPU).  Make that first call during
//     set-up, before any stream capture begins.
//   * the scratch used is the CURRENT device's copy, and the kernels run on
//     `stream`: the current device must be the one that owns `stream` and the
//     seven operand buffers.  Device pointers that belong to another GPU are
//     refused with B200GRU_ERR_ARG (checked with cudaPointerGetAttributes,
//     which allocates nothing) instead of failing later with a launch error.
//   * operands only need the natural 4-byte alignment of float; 16-byte
//     aligned rows (pointer and in_dim/hidden multiples of 4) take the faster
//     vector loads.
// No counterpart in the reference.
#include <mutex>

#include "model.hpp"

namespace {

constexpr size_t kScratchBytes = B200GRU_NOALLOC_SCRATCH_BYTES;
constexpr size_t kScratchFloats = kScratchBytes / sizeof(float);
constexpr int kMaxDevices = 64;

// One copy of this array exists on every GPU that loads the library and stays
// until the process exits.  In host code the name is only a handle for the
// runtime; the address kernels can use comes from cudaGetSymbolAddress.
static __device__ __align__(256) float g_noalloc_scratch[kScratchFloats];

std::mutex g_mu;
float *g_dev_ptr[kMaxDevices] = {nullptr};

// Per-device set-up, with that device current: resolve the array's address.
int scratch_for_current_device(float **out) {
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices)
    return b200gru::fail(B200GRU_ERR_ARG, "device index %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_dev_ptr[dev]) {
    float *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return b200gru::fail(B200GRU_ERR_CUDA,
                           "cudaGetSymbolAddress(scratch) on device %d: %s", dev,
                           cudaGetErrorString(e));
    g_dev_ptr[dev] = p;
  }
  *out = g_dev_ptr[dev];
  return 0;
}

}  // namespace

extern "C" int b200gru_gru_cell_fp32_noalloc(
    const float *x, const float *h_in, const float *w_ih, const float *w_hh,
    const float *b_ih, const float *b_hh, float *h_out, int batch, int in_dim,
    int hidden, void *stream) {
  using namespace b200gru;
  B200_REQUIRE(x && h_in && w_ih && w_hh && b_ih && b_hh && h_out, "null argument");
  B200_REQUIRE(batch > 0 && in_dim > 0 && hidden > 0, "bad dims");
  const size_t need = 2 * (size_t)batch * 3 * (size_t)hidden;
  if (need > kScratchFloats)
    return fail(B200GRU_ERR_ARG,
                "gru_cell_fp32_noalloc: batch=%d hidden=%d needs %zu bytes of "
                "scratch, the built-in buffer holds %zu; use a smaller batch or "
                "b200gru_gru_cell_fp32",
                batch, hidden, need * sizeof(float), kScratchBytes);
  // every operand must live on the device whose scratch (and stream) is used
  {
    int cur = -1;
    B200_CUDA(cudaGetDevice(&cur));
    const void *ops[7] = {x, h_in, w_ih, w_hh, b_ih, b_hh, h_out};
    static const char *const names[7] = {"x", "h_in", "w_ih", "w_hh", "b_ih", "b_hh", "h_out"};
    for (int i = 0; i < 7; ++i) {
      if (reinterpret_cast<uintptr_t>(ops[i]) % alignof(float))
        return fail(B200GRU_ERR_ARG, "gru_cell_fp32_noalloc: %s is not 4-byte aligned", names[i]);
      cudaPointerAttributes at{};
      if (cudaPointerGetAttributes(&at, ops[i]) != cudaSuccess) {
        cudaGetLastError();
        continue;  // not known to the runtime (e.g. a foreign allocator): cannot be checked here
      }
      if (at.type == cudaMemoryTypeDevice && at.device != cur)
        return fail(B200GRU_ERR_ARG,
                    "gru_cell_fp32_noalloc: %s lives on device %d but device %d is current; make the "
                    "operands' device current (the built-in scratch is per device)",
                    names[i], at.device, cur);
    }
  }
  float *scratch = nullptr;
  if (int rc = scratch_for_current_device(&scratch)) return rc;
  return gru_cell_fp32_run(x, h_in, w_ih, w_hh, b_ih, b_hh, h_out, batch, in_dim,
                           hidden, scratch, static_cast<cudaStream_t>(stream));
}
