// engine.cu — model lifetime and the generate entry point.
//
//   b200gru_model_create   <- namegen_initialize (ref:359-618)
//   b200gru_generate       <- namegen, one rank's slice (ref:627-890)
//   b200gru_model_destroy  <- namegen_finalize (ref:897-1033)
//
// Differences from the reference that are deliberate: the device is chosen
// explicitly (the reference never calls cudaSetDevice); every buffer is
// allocated once here, checked, tracked and freed (the reference leaks the
// file buffer, double-allocates gb_fc ref:571-577, and cudaFree()s a host
// pointer ref:1017); nothing is allocated in generate.
#include <cstdlib>

#include <nvtx3/nvToolsExt.h>

#include "model.hpp"
#include "sm100_ptx.cuh"

using namespace b200gru;

namespace {

template <typename T>
int dev_alloc(b200gru_model *m, T **out, size_t count) {
  return model_alloc(m, out, count ? count : 1);
}

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    int cur = -1;
    cudaGetDevice(&cur);
    if (prev >= 0 && cur != prev) cudaSetDevice(prev);
  }
};

__global__ void wait_flags_kernel(const uint32_t *flags, int n, uint32_t value) {
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    // epochs only grow: '>=' (wrap-safe) tolerates a peer that is already a call ahead
#ifdef B200GRU_DEBUG
    sm100::SpinGuard g;
#endif
    uint32_t v;
    while ((int32_t)((v = sm100::ld_acquire_sys(flags + i)) - value) < 0) {
      __nanosleep(200);
#ifdef B200GRU_DEBUG
      g.tick("rank flag", flags + i, value, v);
#endif
    }
  }
  __syncthreads();
  __threadfence_system();
}

#ifdef B200GRU_DEBUG
constexpr size_t kGuardBytes = 4096;
constexpr unsigned char kGuardByte = 0xA5;
__global__ void guard_check_kernel(const unsigned char *lo, const unsigned char *hi, size_t n,
                                   unsigned int *bad) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    if (lo[i] != kGuardByte || hi[i] != kGuardByte) atomicAdd(bad, 1u);
}
#endif

int create_impl(b200gru_model *m, const float *params_host) {
  const b200gru_config &c = m->cfg;
  const int H = c.hidden, C = c.num_char, L = c.num_layers, E = c.emb_dim;
  LayoutIndex ix(L);
  const int64_t n_params = m->lay.back().offset + m->lay.back().size;

  // C8: upload.  One copy of the whole file instead of 28 cudaMemcpy calls.
  if (int rc = dev_alloc(m, &m->params, (size_t)n_params)) return rc;
  B200_CUDA(cudaMemcpy(m->params, params_host, sizeof(float) * n_params,
                       cudaMemcpyHostToDevice));
  m->b_fc = m->P(ix.b_fc());

  // C9: activation buffers, sized once.
  m->chunk = m->max_batch < kExactChunk ? m->max_batch : kExactChunk;
  if (int rc = dev_alloc(m, &m->t0_raw, (size_t)C * 3 * H)) return rc;
  if (m->tp_world == 1) {  // the fp32 exact path is single-rank-per-model only
    if (int rc = dev_alloc(m, &m->g_ih, (size_t)m->chunk * 3 * H)) return rc;
    if (int rc = dev_alloc(m, &m->g_hh, (size_t)m->chunk * 3 * H)) return rc;
    if (int rc = dev_alloc(m, &m->h32, (size_t)L * m->chunk * H)) return rc;
  }
  if (int rc = dev_alloc(m, &m->logits, (size_t)m->bpad * C)) return rc;
  if (int rc = dev_alloc(m, &m->chars, (size_t)m->bpad)) return rc;
  if (int rc = dev_alloc(m, &m->alive, (size_t)m->bpad)) return rc;
  B200_CUDA(cudaMemset(m->logits, 0, sizeof(float) * (size_t)m->bpad * C));

  // Layer-0 input projection as a table: the embedding (ref:112-118) followed
  // by the three W_i*0 matvecs (ref:676,:702,:729) only depends on the
  // character id, so it is computed once per id.  Same left-to-right fma
  // order => same bits as doing it per step.
  if (int rc = exact_sgemm_tn(m->P(ix.emb()), m->P(ix.w_i(0, 0)), m->t0_raw, C,
                              3 * H, E, 0))
    return rc;

  // Tensor-core path needs 64-aligned hidden size and a single-UMMA FC.
  m->tc_ok = (H % (kUnitBlock * m->tp_world) == 0) && (C % 16 == 0) && C >= 16 && C <= 256;
  if (m->tp_world > 1 && !m->tc_ok)
    return fail(B200GRU_ERR_ARG, "tensor parallel needs hidden %% (64*world) == 0 and num_char a "
                                 "multiple of 16 in [16,256] (H=%d, world=%d, C=%d)", H, m->tp_world, C);
  if (m->tc_ok)
    if (int rc = tc_prepare(m)) return rc;
  if (m->tc_ok)
    if (int rc = persist_prepare(m)) return rc;
  B200_CUDA(cudaDeviceSynchronize());
  return 0;
}

}  // namespace

namespace b200gru {

int model_alloc_bytes(b200gru_model *m, void **out, size_t bytes) {
  if (bytes == 0) bytes = 4;
  void *p = nullptr;
#ifdef B200GRU_DEBUG
  const size_t body = (bytes + 255) / 256 * 256;
  cudaError_t e = cudaMalloc(&p, body + 2 * kGuardBytes);
#else
  cudaError_t e = cudaMalloc(&p, bytes);
#endif
  if (e != cudaSuccess)
    return fail(B200GRU_ERR_CUDA, "cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
  m->allocs.push_back(p);
  m->device_bytes += (int64_t)bytes;
#ifdef B200GRU_DEBUG
  // [guard | body rounded to 256 | guard]: the rounding slack belongs to the
  // upper guard's side and is patterned too
  B200_CUDA(cudaMemset(p, kGuardByte, body + 2 * kGuardBytes));
  p = static_cast<unsigned char *>(p) + kGuardBytes;
  m->guarded.emplace_back(p, body);
#endif
  *out = p;
  return 0;
}

int check_guards(b200gru_model *m, const char *when) {
#ifdef B200GRU_DEBUG
  unsigned int *bad = nullptr;
  B200_CUDA(cudaDeviceSynchronize());
  B200_CUDA(cudaMalloc(&bad, sizeof(unsigned int)));
  for (size_t i = 0; i < m->guarded.size(); ++i) {
    const unsigned char *body = static_cast<const unsigned char *>(m->guarded[i].first);
    B200_CUDA(cudaMemset(bad, 0, sizeof(unsigned int)));
    guard_check_kernel<<<8, 256>>>(body - kGuardBytes, body + m->guarded[i].second, kGuardBytes, bad);
    unsigned int h = 0;
    B200_CUDA(cudaMemcpy(&h, bad, sizeof(h), cudaMemcpyDeviceToHost));
    if (h) {
      cudaFree(bad);
      return fail(B200GRU_ERR_STATE, "guard band of allocation #%zu (%zu bytes) damaged (%u bytes) %s",
                  i, m->guarded[i].second, h, when);
    }
  }
  cudaFree(bad);
#else
  (void)m; (void)when;
#endif
  return 0;
}

}  // namespace b200gru

extern "C" {

int b200gru_model_check(b200gru_model *m) {
  B200_REQUIRE(m, "null model");
  DeviceGuard guard(m->device);
  return check_guards(m, "(b200gru_model_check)");
}

int b200gru_debug_build(void) {
#ifdef B200GRU_DEBUG
  return 1;
#else
  return 0;
#endif
}

int b200gru_model_create(const b200gru_config *cfg, const float *params,
                         int device, int max_batch, b200gru_model **out) {
  return b200gru_model_create_tp(cfg, params, device, max_batch, 0, 1, out);
}

size_t b200gru_tp_shared_bytes(const b200gru_model *m) {
  return (m && m->tp_world > 1) ? tp_shared_bytes(m) : 0;
}

int b200gru_tp_bind(b200gru_model *m, void *local_base, void *const *peer_bases) {
  return b200gru_tp_bind_mc(m, local_base, peer_bases, nullptr);
}

int b200gru_tp_bind_mc(b200gru_model *m, void *local_base, void *const *peer_bases, void *mc_base) {
  B200_REQUIRE(m && local_base && peer_bases, "null argument");
  DeviceGuard guard(m->device);
  return tp_bind(m, local_base, peer_bases, mc_base);
}

int b200gru_model_create_tp(const b200gru_config *cfg, const float *params,
                            int device, int max_batch, int tp_rank, int tp_world,
                            b200gru_model **out) {
  if (int rc = validate_config(cfg)) return rc;
  B200_REQUIRE(params && out, "null argument");
  B200_REQUIRE(max_batch > 0, "max_batch must be positive");
  B200_REQUIRE(tp_world >= 1 && tp_world <= 8 && tp_rank >= 0 && tp_rank < tp_world,
               "tensor-parallel rank %d of %d", tp_rank, tp_world);
  int ndev = 0;
  B200_CUDA(cudaGetDeviceCount(&ndev));
  B200_REQUIRE(device >= 0 && device < ndev, "device %d of %d", device, ndev);
  DeviceGuard guard(device);
  auto *m = new b200gru_model();
  m->cfg = *cfg;
  m->device = device;
  m->max_batch = max_batch;
  m->tp_rank = tp_rank;
  m->tp_world = tp_world;
  {
    const char *e = getenv("B200GRU_GRAPH");
    m->use_graph = !(e && e[0] == '0');
  }
  {
    // kernel path for tensor-core mode: auto picks the persistent kernel for small batches
    if (const char *e = getenv("B200GRU_PATH")) m->ps_path = atoi(e);
    if (const char *e = getenv("B200GRU_PERSIST_MAX")) m->ps_max_n = atoi(e);
    if (const char *e = getenv("B200GRU_COMPACT")) m->compact = e[0] != '0';
    if (const char *e = getenv("B200GRU_COMPACT_MIN")) m->compact_min_n = atoi(e);
  }
  {
    const char *e = getenv("B200GRU_TP_FAKE");
    m->tp_fake = tp_world > 1 && e && e[0] == '1';
  }
  m->bpad = (max_batch + 2 * kTileM - 1) / (2 * kTileM) * (2 * kTileM);  // whole CTA-pair tiles
  m->lay = build_layout(*cfg);
  {
    // pad rows of K=H matrices by 128 bytes: with H a power of two, the 128
    // rows of a TMA box would otherwise sit exactly 2^k bytes apart
    const char *e = getenv("B200GRU_LD_PAD");
    int pad = e ? atoi(e) : 0;
    if (pad < 0 || pad % 8) pad = 64;
    m->ldh = cfg->hidden + pad;
  }
  int rc = create_impl(m, params);
  if (rc) {
    for (void *p : m->allocs) cudaFree(p);
    delete m;
    return rc;
  }
  *out = m;
  return 0;
}

int b200gru_model_destroy(b200gru_model *m) {
  if (!m) return 0;
  DeviceGuard guard(m->device);
  cudaError_t first = cudaDeviceSynchronize();
  const int guards_rc = check_guards(m, "at destroy");
  for (cudaEvent_t e : m->trace_ev) cudaEventDestroy(e);
  for (auto &g : m->graphs) {
    cudaGraphExecDestroy(g.exec);
    if (g.exec_tail) cudaGraphExecDestroy(g.exec_tail);
    if (g.graph) cudaGraphDestroy(g.graph);
  }
  if (m->cap_stream) cudaStreamDestroy(m->cap_stream);
  for (void *p : m->allocs) {
    cudaError_t e = cudaFree(p);
    if (first == cudaSuccess) first = e;
  }
  delete m;
  if (first != cudaSuccess)
    return fail(B200GRU_ERR_CUDA, "destroy: %s", cudaGetErrorString(first));
  return guards_rc;
}

int64_t b200gru_model_device_bytes(const b200gru_model *m) {
  return m ? m->device_bytes : 0;
}

static int generate_impl(b200gru_model *m, int mode, int n, const b200gru_io *io, void *stream);

int b200gru_generate(b200gru_model *m, int mode, int n, const b200gru_io *io,
                     void *stream) {
  int rc = generate_impl(m, mode, n, io, stream);
#ifdef B200GRU_DEBUG
  // debug builds give up asynchrony: wait for the call and verify every guard band
  if (rc == 0 && m && !(io && (io->done_flag || io->ready_flag)) && m->tp_world == 1) {
    DeviceGuard guard(m->device);
    rc = check_guards(m, "after generate");
  }
#endif
  return rc;
}

static int generate_impl(b200gru_model *m, int mode, int n, const b200gru_io *io, void *stream) {
  B200_REQUIRE(m && io, "null argument");
  B200_REQUIRE(n >= 0 && n <= m->max_batch, "n=%d exceeds max_batch=%d", n,
               m->max_batch);
  B200_REQUIRE(io->rfloats && io->output, "io pointers are null");
  DeviceGuard guard(m->device);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  m->last_launches = 0;
  // NVTX range around the host side of the call (header-only NVTX3: a no-op
  // unless a profiler is attached); the reference has no instrumentation at all
  struct NvtxScope {
    explicit NvtxScope(const char *name) { nvtxRangePushA(name); }
    ~NvtxScope() { nvtxRangePop(); }
  } nvtx(mode == B200GRU_MODE_TC ? "b200gru_generate[tc]" : "b200gru_generate[exact]");
  if (n == 0) {
    // a rank with no names (N < ranks, ref:628 gives JPP = 0) still has to
    // report completion so the root's gather wait can finish
    if (io->done_flag && m->tp_world == 1) {
      if (int rc = b200gru_signal_flag(io->done_flag, io->done_value, stream)) return rc;
      m->last_launches = 1;
    }
    return 0;
  }
  if (mode == B200GRU_MODE_EXACT) {
    if (m->tp_world > 1)
      return fail(B200GRU_ERR_ARG, "the fp32 exact path does not run tensor parallel");
    return exact_generate(m, n, io, s);
  }
  if (mode == B200GRU_MODE_TC) {
    if (!m->tc_ok)
      return fail(B200GRU_ERR_ARG,
                  "tensor-core mode needs hidden %% 64 == 0 and num_char a "
                  "multiple of 16 in [16,256] (got H=%d, C=%d)",
                  m->cfg.hidden, m->cfg.num_char);
    // Small batches: the whole call is one persistent kernel (weights resident
    // in shared memory, flags handled in-kernel).  Large batches: one launch per
    // layer per character, replayed as a CUDA graph.
    if (m->ps_path == 2 && !m->ps_ok)
      return fail(B200GRU_ERR_ARG, "the persistent kernel cannot run this model (H=%d, L=%d) on this GPU",
                  m->cfg.hidden, m->cfg.num_layers);
    if (m->ps_ok && (m->ps_path == 2 || (m->ps_path == 0 && n <= m->ps_max_n)))
      return persist_generate(m, n, io, s);
    // Replay a captured graph when nothing about the call varies: no
    // completion flag (its value changes per call), no tensor-parallel
    // counters, no tracing.
    // Replayable: no tensor-parallel counters, no tracing.  Data-parallel flags
    // are fine: their addresses are part of the cache key and the per-call epoch
    // values are patched into the two kernel nodes that carry them.
    const bool graphable = m->use_graph && m->tp_world == 1 && !m->trace_on &&
                           !(io->ready_event && (io->done_flag || io->ready_flag));
    if (!graphable) return tc_generate(m, n, io, s);
    const bool split = io->ready_event != nullptr;  // two graphs with an event wait in between
    auto launch = [&](b200gru_model::GraphEntry &g) -> int {
      g.last_use = ++m->graph_clock;
      m->last_launches = g.launches;
      if (g.ready_node) {
        const uint32_t *flags = io->ready_flag;
        int cnt = 1;
        uint32_t value = io->ready_value;
        void *args[3] = {&flags, &cnt, &value};
        cudaKernelNodeParams p{};
        p.func = reinterpret_cast<void *>(wait_flags_kernel);
        p.gridDim = dim3(1); p.blockDim = dim3(32); p.kernelParams = args;
        B200_CUDA(cudaGraphExecKernelNodeSetParams(g.exec, g.ready_node, &p));
      }
      if (g.done_node) {
        uint32_t *flag = io->done_flag;
        uint32_t value = io->done_value;
        void *args[2] = {&flag, &value};
        cudaKernelNodeParams p{};
        p.func = reinterpret_cast<void *>(tc_signal_done_kernel);
        p.gridDim = dim3(1); p.blockDim = dim3(1); p.kernelParams = args;
        B200_CUDA(cudaGraphExecKernelNodeSetParams(g.exec, g.done_node, &p));
      }
      B200_CUDA(cudaGraphLaunch(g.exec, s));
      if (g.exec_tail) {
        B200_CUDA(cudaStreamWaitEvent(s, static_cast<cudaEvent_t>(io->ready_event), 0));
        B200_CUDA(cudaGraphLaunch(g.exec_tail, s));
      }
      return 0;
    };
    for (auto &g : m->graphs)
      if (g.n == n && g.rfloats == io->rfloats && g.output == io->output &&
          g.logits == m->debug_logits && (g.exec_tail != nullptr) == split &&
          g.done_flag == io->done_flag && g.ready_flag == io->ready_flag)
        return launch(g);
    // Capture on a private stream (the caller's may be the legacy default
    // stream, which cannot be captured); the graph is launched on the caller's.
    if (!m->cap_stream)
      B200_CUDA(cudaStreamCreateWithFlags(&m->cap_stream, cudaStreamNonBlocking));
    b200gru_model::GraphEntry e;
    for (int half = 0; half < (split ? 2 : 1); ++half) {
      cudaGraph_t graph = nullptr;
      B200_CUDA(cudaStreamBeginCapture(m->cap_stream, cudaStreamCaptureModeThreadLocal));
      int rc = tc_generate(m, n, io, m->cap_stream, split ? half + 1 : 0);
      cudaError_t ce = cudaStreamEndCapture(m->cap_stream, &graph);
      if (!rc && ce == cudaSuccess && graph && (io->done_flag || io->ready_flag)) {
        // find the two nodes whose arguments carry the per-call epoch
        size_t nn = 0;
        cudaGraphGetNodes(graph, nullptr, &nn);
        std::vector<cudaGraphNode_t> nodes(nn);
        cudaGraphGetNodes(graph, nodes.data(), &nn);
        for (cudaGraphNode_t nd : nodes) {
          cudaGraphNodeType ty;
          if (cudaGraphNodeGetType(nd, &ty) != cudaSuccess || ty != cudaGraphNodeTypeKernel) continue;
          cudaKernelNodeParams kp{};
          if (cudaGraphKernelNodeGetParams(nd, &kp) != cudaSuccess) continue;
          if (kp.func == reinterpret_cast<void *>(wait_flags_kernel)) e.ready_node = nd;
          if (kp.func == reinterpret_cast<void *>(tc_signal_done_kernel)) e.done_node = nd;
        }
        if ((io->done_flag && !e.done_node) || (io->ready_flag && !e.ready_node)) ce = cudaErrorUnknown;
      }
      cudaGraphExec_t exec = nullptr;
      if (!rc && ce == cudaSuccess && graph) ce = cudaGraphInstantiate(&exec, graph, 0);
      if (graph && exec && (e.done_node || e.ready_node)) {
        e.graph = graph;  // node handles stay valid only as long as their graph does
        graph = nullptr;
      }
      if (graph) cudaGraphDestroy(graph);
      if (rc || ce != cudaSuccess || !exec) {
        if (e.exec) cudaGraphExecDestroy(e.exec);
        if (e.graph) cudaGraphDestroy(e.graph);
        if (rc) return rc;
        cudaGetLastError();  // capture not possible here: run eagerly from now on
        m->use_graph = false;
        return tc_generate(m, n, io, s);
      }
      (half == 0 ? e.exec : e.exec_tail) = exec;
      e.launches += m->last_launches;
    }
    e.n = n; e.rfloats = io->rfloats; e.output = io->output; e.logits = m->debug_logits;
    e.done_flag = io->done_flag; e.ready_flag = io->ready_flag;
    if (m->graphs.size() >= 8) {  // evict the least recently used
      size_t victim = 0;
      for (size_t i = 1; i < m->graphs.size(); ++i)
        if (m->graphs[i].last_use < m->graphs[victim].last_use) victim = i;
      cudaGraphExecDestroy(m->graphs[victim].exec);
      if (m->graphs[victim].exec_tail) cudaGraphExecDestroy(m->graphs[victim].exec_tail);
      if (m->graphs[victim].graph) cudaGraphDestroy(m->graphs[victim].graph);
      m->graphs[victim] = e;
      return launch(m->graphs[victim]);
    }
    m->graphs.push_back(e);
    return launch(m->graphs.back());
  }
  return fail(B200GRU_ERR_ARG, "unknown mode %d", mode);
}

int b200gru_last_launch_count(const b200gru_model *m) {
  return m ? m->last_launches : 0;
}

int b200gru_model_set_debug(b200gru_model *m, int logits_tap) {
  B200_REQUIRE(m, "null model");
  m->debug_logits = logits_tap != 0;
  return 0;
}

int b200gru_model_set_path(b200gru_model *m, int path, int persist_max_names) {
  B200_REQUIRE(m, "null model");
  B200_REQUIRE(path >= 0 && path <= 2, "path must be 0 (auto), 1 (multi-launch) or 2 (persistent)");
  m->ps_path = path;
  if (persist_max_names >= 0) m->ps_max_n = persist_max_names;
  return m->ps_ok ? 1 : 0;
}

int b200gru_model_set_compaction(b200gru_model *m, int on) {
  B200_REQUIRE(m, "null model");
  if (m->compact != (on != 0)) {
    // captured graphs have the choice baked in
    for (auto &g : m->graphs) {
      cudaGraphExecDestroy(g.exec);
      if (g.exec_tail) cudaGraphExecDestroy(g.exec_tail);
      if (g.graph) cudaGraphDestroy(g.graph);
    }
    m->graphs.clear();
  }
  m->compact = on != 0;
  return 0;
}

int b200gru_model_set_trace(b200gru_model *m, int on) {
  B200_REQUIRE(m, "null model");
  m->trace_on = on != 0;
  return 0;
}

// Spans of the last traced generate: fills ms[i] / names[i], returns the count.
int b200gru_model_get_trace(b200gru_model *m, float *ms, const char **names, int max_spans) {
  B200_REQUIRE(m && ms && names, "null argument");
  DeviceGuard guard(m->device);
  const int spans = m->trace_used > 0 ? m->trace_used - 1 : 0;
  if (spans > 0) B200_CUDA(cudaEventSynchronize(m->trace_ev[m->trace_used - 1]));
  int n = spans < max_spans ? spans : max_spans;
  for (int i = 0; i < n; ++i) {
    B200_CUDA(cudaEventElapsedTime(&ms[i], m->trace_ev[i], m->trace_ev[i + 1]));
    names[i] = i < (int)m->trace_names.size() ? m->trace_names[i] : "?";
  }
  return n;
}

int b200gru_debug_logits(b200gru_model *m, float *dst_device, int n,
                         void *stream) {
  B200_REQUIRE(m && dst_device && n >= 0 && n <= m->bpad, "bad arguments");
  DeviceGuard guard(m->device);
  B200_CUDA(cudaMemcpyAsync(dst_device, m->logits,
                            sizeof(float) * (size_t)n * m->cfg.num_char,
                            cudaMemcpyDeviceToDevice,
                            static_cast<cudaStream_t>(stream)));
  return 0;
}

int b200gru_wait_flags(const uint32_t *flags, int n, uint32_t value,
                       void *stream) {
  B200_REQUIRE(flags && n > 0, "bad arguments");
  wait_flags_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(flags, n,
                                                                    value);
  B200_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
