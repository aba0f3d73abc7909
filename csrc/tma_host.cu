// tma_host.cu — builds CUtensorMap descriptors without linking libcuda: the
// driver entry point is fetched through the runtime, so the library still
// loads (for its CPU-side functions) on a machine with no driver.
#include <cudaTypedefs.h>

#include "model.hpp"

namespace b200gru {

static PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault,
                                &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  }
  return fn;
}

// Row-major bf16 matrix [rows, cols]; box = box_rows x box_cols elements with
// the 128-byte swizzle the UMMA descriptors in sm100_ptx.cuh expect
// (box_cols * 2 bytes must be 128).  `ld` = row stride in elements (0: dense).
int make_tmap_bf16_2d(CUtensorMap *map, const void *base, uint64_t rows,
                      uint64_t cols, uint32_t box_rows, uint32_t box_cols,
                      uint64_t ld) {
  if (ld == 0) ld = cols;
  auto fn = encode_fn();
  if (!fn) return fail(B200GRU_ERR_CUDA, "cuTensorMapEncodeTiled unavailable");
  B200_REQUIRE(box_cols * 2 == 128, "box inner extent must be 128 bytes");
  B200_REQUIRE(box_rows >= 1 && box_rows <= 256, "box rows %u", box_rows);
  B200_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0 &&
                   (ld * 2) % 16 == 0 && ld >= cols,
               "tensor map base/stride must be 16-byte aligned");
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  const_cast<void *>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(B200GRU_ERR_CUDA, "cuTensorMapEncodeTiled failed: %d", (int)r);
  return 0;
}

}  // namespace b200gru
