// model.hpp — device-side model state (the reference's C5/C8/C9: raw float*
// globals at ref:348-357 filled by ref:471-612).  Here it is one heap object
// per model so several models / devices can coexist, every buffer is allocated
// once in b200gru_model_create (the reference re-mallocs per call, ref:632),
// and everything is released in b200gru_model_destroy.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <utility>
#include <vector>

#include "errors.hpp"
#include "layout.hpp"

namespace b200gru {

constexpr int kTileM = 128;     // names per tensor-core tile
constexpr int kUnitBlock = 64;  // hidden units per tile (x3 / x4 gate columns)
constexpr int kExactChunk = 8192;

struct TcLayer {
  __nv_bfloat16 *w_ih = nullptr;  // packed [H/64][n|r|z x 64][H]  (layers >= 1)
  __nv_bfloat16 *w_hh = nullptr;  // packed [H/64][r|z|n x 64][H]
  float *bias = nullptr;          // [H/64][4][64]: b_in, b_r(+hr), b_z(+hz), b_hn
  CUtensorMap map_w_ih, map_w_hh;
  CUtensorMap map_w_ih_p, map_w_hh_p;  // 96-row boxes for the CTA-pair kernel
  CUtensorMap map_h[2];           // A-operand maps over the two h buffers
  __nv_bfloat16 *h[2] = {nullptr, nullptr};  // [bpad, H] ping-pong
};

}  // namespace b200gru

struct b200gru_model {
  b200gru_config cfg{};
  int device = 0;
  int max_batch = 0;
  int bpad = 0;  // max_batch rounded up to kTileM
  bool tc_ok = false;
  std::vector<b200gru::TensorSpec> lay;
  std::vector<void *> allocs;
  // B200GRU_DEBUG builds: every allocation sits between two 4 KB guard bands
  // filled with a pattern; (user pointer, bytes) of each is kept for the checks
  std::vector<std::pair<void *, size_t>> guarded;
  int64_t device_bytes = 0;
  int last_launches = 0;

  float *params = nullptr;  // whole weight file, fp32
  // --- exact (fp32 SIMT) path
  int chunk = 0;
  float *t0_raw = nullptr;  // [C, 3H]   W_ih0 . emb[c]   (no bias)
  float *g_ih = nullptr;    // [chunk, 3H]
  float *g_hh = nullptr;    // [chunk, 3H]
  float *h32 = nullptr;     // [L, chunk, H]
  // --- shared
  float *logits = nullptr;  // [bpad, C]
  int32_t *chars = nullptr; // [bpad]
  uint8_t *alive = nullptr; // [bpad]
  // --- tensor-core path
  int ldh = 0;                       // row stride (elements) of hidden buffers and K=H weights
                                     // (H + B200GRU_LD_PAD; padding measured: no effect, default 0)
  int pw_log2 = 0;                   // hidden buffers: [H/pw panels][bpad][pw], pw = 1 << pw_log2
  // compaction of finished names (multi-launch path)
  bool compact = true;               // B200GRU_COMPACT=0 / b200gru_model_set_compaction turn it off
  int compact_min_n = 8192;          // calls with fewer names never compact (B200GRU_COMPACT_MIN)
  int32_t *slot_row = nullptr, *slot2 = nullptr;   // [bpad] original name of every row (+ scratch)
  int32_t *chars2 = nullptr, *new_idx = nullptr;   // [bpad] scratch
  int *cp_state = nullptr;           // [0] rows in use, [1] rows after the planned compaction, [2] carry it out
  __nv_bfloat16 *t0q = nullptr;      // [H/64][C][2][2][n|r|z][16] folded layer-0 input table (epilogue gather)
  bool use_pdl = true;               // programmatic dependent launch between the step kernels
  bool use_pairs = true;             // cta_group::2 layer kernel (B200GRU_CTA_PAIRS=0 disables)
  bool debug_logits = false;         // FC kernel also stores pre-softmax logits
  b200gru::TcLayer tc[B200GRU_MAX_LAYERS];
  __nv_bfloat16 *w_fc = nullptr;  // [C, H] bf16
  CUtensorMap map_w_fc;
  const float *b_fc = nullptr;    // into params

  // --- persistent weight-stationary kernel (gru_persist_sm100.cu): one launch
  // per generate for small batches
  bool ps_ok = false;
  int ps_path = 0;                   // 0 auto (by batch size), 1 multi-launch only, 2 persistent only
  int ps_max_n = 512;                // auto: use the persistent kernel up to this many names
  int ps_smem = 0, ps_stages = 0, ps_pairs = 0, ps_fc_max = 0;
  __nv_bfloat16 *ps_w_ih[B200GRU_MAX_LAYERS] = {};   // packed [H/16][n|r|z x 16][H]
  __nv_bfloat16 *ps_w_hh[B200GRU_MAX_LAYERS] = {};   // packed [H/16][r|z|n x 16][H]
  CUtensorMap ps_map_w_ih[B200GRU_MAX_LAYERS], ps_map_w_hh[B200GRU_MAX_LAYERS];
  CUtensorMap ps_map_wfc_half;       // W_fc with a box of C/2 rows (FC CTA pairs split the alphabet)
  __nv_bfloat16 *ps_t0p = nullptr;   // [H/16][C][n|r|z][16] folded layer-0 input table
  uint32_t *ps_sync = nullptr;       // per-tile counters + live counts, zeroed per call
  unsigned long long *ps_timeline = nullptr;  // in-kernel globaltimer stamps (tracing only)

  // --- tensor parallelism over the hidden dimension (tp_world == 1: off)
  int tp_world = 1, tp_rank = 0;
  bool tp_bound = false;
  bool tp_fake = false;   // B200GRU_TP_FAKE=1: single-GPU profiling of one rank's shapes
  uint32_t tp_epoch = 0;                                   // completed generate calls
  uint32_t tp_rounds[B200GRU_MAX_LAYERS][2] = {};          // write rounds per hidden buffer
  __nv_bfloat16 *tp_h[8][B200GRU_MAX_LAYERS][2] = {};      // [rank][layer][pp] hidden buffers
  uint32_t *tp_cnt[8][B200GRU_MAX_LAYERS][2] = {};         // [rank][layer][pp] -> [src][m_tile]
  uint32_t *tp_sync[8] = {};                               // [rank] -> [src] call-fence words
  uint8_t *tp_mc_base = nullptr;                           // NVSwitch multicast view of the shared region (optional)
  __nv_bfloat16 *tp_h_mc[B200GRU_MAX_LAYERS][2] = {};
  uint32_t *tp_cnt_mc[B200GRU_MAX_LAYERS][2] = {};

  // --- CUDA graphs: the 31 launches of a tensor-core generate are captured once
  // per (n, io) and replayed, which removes the per-launch CPU cost that
  // dominates small batches (B200GRU_GRAPH=0 disables)
  struct GraphEntry {
    int n = 0;
    const void *rfloats = nullptr;
    void *output = nullptr;
    bool logits = false;
    cudaGraphExec_t exec = nullptr;
    cudaGraphExec_t exec_tail = nullptr;  // second half when the call is split at `ready_event`
    // data-parallel calls: the flag addresses are part of the key, the epoch
    // VALUES change per call and are patched into these two kernel nodes
    uint32_t *done_flag = nullptr;
    const uint32_t *ready_flag = nullptr;
    cudaGraphNode_t done_node = nullptr, ready_node = nullptr;
    cudaGraph_t graph = nullptr;  // kept alive while the node handles above are in use
    int launches = 0;
    uint64_t last_use = 0;
  };
  std::vector<GraphEntry> graphs;
  uint64_t graph_clock = 0;
  bool use_graph = true;
  cudaStream_t cap_stream = nullptr;

  // --- tracing: CUDA events between launches (the reference has no timers at all)
  bool trace_on = false;
  std::vector<cudaEvent_t> trace_ev;
  std::vector<const char *> trace_names;
  int trace_used = 0;
  void trace_begin(cudaStream_t s) {
    trace_used = 0;
    trace_names.clear();
    if (trace_on) trace_mark(nullptr, s);
  }
  // records an event; `name` labels the span that ENDS here
  void trace_mark(const char *name, cudaStream_t s) {
    if (!trace_on) return;
    if (trace_used == (int)trace_ev.size()) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      trace_ev.push_back(e);
    }
    cudaEventRecord(trace_ev[trace_used++], s);
    if (name) trace_names.push_back(name);
  }

  const float *P(int k) const { return params + lay[k].offset; }
};

namespace b200gru {
// Device allocation owned by the model (engine.cu).  In B200GRU_DEBUG builds the
// block is surrounded by guard bands that check_guards() verifies.
int model_alloc_bytes(b200gru_model *m, void **out, size_t bytes);
template <typename T>
inline int model_alloc(b200gru_model *m, T **out, size_t count) {
  void *p = nullptr;
  if (int rc = model_alloc_bytes(m, &p, count * sizeof(T))) return rc;
  *out = static_cast<T *>(p);
  return 0;
}
// Debug builds: 0 if every guard band of the model still holds its pattern, else
// an error naming the damaged block (synchronises the device).  Release: 0.
int check_guards(b200gru_model *m, const char *when);
// exact path (gru_exact.cu)
int exact_sgemm_tn(const float *a, const float *w, float *c, int m, int n,
                   int k, cudaStream_t s);
int exact_generate(b200gru_model *m, int n, const b200gru_io *io,
                   cudaStream_t s);
int gru_cell_fp32_run(const float *x, const float *h_in, const float *w_ih,
                      const float *w_hh, const float *b_ih, const float *b_hh,
                      float *h_out, int batch, int in_dim, int hidden,
                      float *scratch, cudaStream_t s);
// tensor-core path (gru_tc_sm100.cu)
int tc_prepare(b200gru_model *m);
int tc_generate(b200gru_model *m, int n, const b200gru_io *io, cudaStream_t s, int part = 0);
__global__ void tc_signal_done_kernel(uint32_t *flag, uint32_t value);
// persistent path (gru_persist_sm100.cu)
int persist_prepare(b200gru_model *m);
int persist_generate(b200gru_model *m, int n, const b200gru_io *io, cudaStream_t s);
size_t tp_shared_bytes(const b200gru_model *m);
int tp_bind(b200gru_model *m, void *local_base, void *const *peer_bases, void *mc_base);
// TMA descriptor helper (tma_host.cu)
int make_tmap_bf16_2d(CUtensorMap *map, const void *base, uint64_t rows,
                      uint64_t cols, uint32_t box_rows, uint32_t box_cols,
                      uint64_t ld = 0);
}  // namespace b200gru
