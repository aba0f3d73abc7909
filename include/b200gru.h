/*
 * b200gru.h — C ABI of the native library behind gru_mpi_cuda_b200.
 *
 * One shared object (gru_mpi_cuda_b200/lib/libb200gru.so) is built in-tree with
 * nvcc for sm_100a and loaded from Python with ctypes.  Every function returns
 * 0 on success and a negative b200gru_status on failure; b200gru_last_error()
 * gives the message for the calling thread.  Device pointers are plain
 * `void*` obtained from `torch.Tensor.data_ptr()`; streams are `cudaStream_t`
 * passed as `void*` (0 = legacy default stream).
 *
 * Citations of the form `ref:NNN` point into /root/reference/namegensf.cu.
 */
#ifndef B200GRU_H_
#define B200GRU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200GRU_MAX_LAYERS 4

typedef enum b200gru_status {
  B200GRU_OK = 0,
  B200GRU_ERR_ARG = -1,     /* bad argument / unsupported shape */
  B200GRU_ERR_IO = -2,      /* file could not be read / written */
  B200GRU_ERR_CUDA = -3,    /* a CUDA runtime / driver call failed */
  B200GRU_ERR_STATE = -4    /* call sequence error (e.g. generate before init) */
} b200gru_status;

/* Model shape.  The reference fixes these as compile-time macros in the
 * unshipped namegen.h (ref:375-407 uses them); here they are runtime values. */
typedef struct b200gru_config {
  int32_t num_char;    /* NUM_CHAR      (default 256)  */
  int32_t emb_dim;     /* EMBEDDING_DIM (default 512)  */
  int32_t hidden;      /* HIDDEN_DIM    (default 1024) */
  int32_t num_layers;  /* stacked GRU layers (reference: 2) */
  int32_t max_len;     /* MAX_LEN       (default 10)   */
  int32_t sos;         /* SOS char id   (default 0)    */
  int32_t eos;         /* EOS char id   (default 1)    */
} b200gru_config;

const char *b200gru_last_error(void);
const char *b200gru_version(void);

/* ---------------------------------------------------------------- weights */
/* Parameter-file layout (ref:375-407): headerless little-endian fp32, tensors
 * in OFFSETk order: embedding, W_i{r,z,n} per layer, W_h{r,z,n} per layer,
 * b_i{r,z,n} per layer, b_h{r,z,n} per layer, W_fc, b_fc.
 * n_tensors = 3 + 12*num_layers (27 for the reference's two layers). */
int b200gru_num_tensors(const b200gru_config *cfg);
int64_t b200gru_param_count(const b200gru_config *cfg);
/* offsets/sizes in floats, rows/cols per tensor (cols = 1 for biases); arrays
 * must hold b200gru_num_tensors(cfg) entries. */
int b200gru_param_layout(const b200gru_config *cfg, int64_t *offsets,
                         int64_t *sizes, int32_t *rows, int32_t *cols);
/* Name of tensor k ("character_embedding", "W_ir0", ...), static storage. */
const char *b200gru_tensor_name(const b200gru_config *cfg, int k);
/* read_binary() equivalent (ref:370-372): whole file into a malloc'd buffer. */
int b200gru_read_binary(const char *path, void **out_buf, size_t *out_bytes);
int b200gru_write_binary(const char *path, const void *buf, size_t bytes);
void b200gru_free_host(void *p);
/* Deterministic random-init weights, uniform(-1/sqrt(H), 1/sqrt(H)) like
 * torch.nn.GRU's default init (splitmix64 stream; same bytes on every host). */
int b200gru_random_init(const b200gru_config *cfg, uint64_t seed, float *out,
                        int64_t n_floats);
/* N*max_len uniform floats in [0,1) — stands in for the unshipped main.cpp's
 * generator (ref:620-626). */
int b200gru_random_floats(uint64_t seed, float *out, int64_t n);

/* -------------------------------------------------------------- partition */
/* Static block split of the reference (ref:628-630): jpp = N / size, rank owns
 * [jpp*rank, jpp*rank + jpp); the N % size tail is dropped. */
void b200gru_partition(int n_names, int world_size, int rank, int *jpp,
                       int *start, int *end);

/* ----------------------------------------------------------- CPU oracle A */
/* flags */
#define B200GRU_ORACLE_LIBM 1u   /* use libm expf/tanhf instead of the shared
                                    bit-reproducible math (exact_math.h)     */
/* Inverse-CDF sampler, ref:322-333 (strict '>' and n-1 fallback). */
int b200gru_random_select(const float *probs, int n, float r);
/* Full generation on the CPU in the reference's association order
 * (ref:649-885).  `params` is the whole weight file, `rfloats` is
 * [n_names*max_len], `output` is [n_names*(max_len+1)] and is zero-filled
 * first (ref:643).  `logits_out` (optional) receives [n_names*max_len*num_char]
 * pre-softmax FC outputs (zeros after EOS). */
int b200gru_oracle_generate(const b200gru_config *cfg, const float *params,
                            int n_names, const float *rfloats, char *output,
                            float *logits_out, uint32_t flags);
/* Individual ops of the reference (ref:112-313), for unit tests. */
int b200gru_oracle_matvec(const float *w, const float *x, float *y, int n,
                          int k);
int b200gru_oracle_softmax(const float *x, float *y, int n, uint32_t flags);
int b200gru_oracle_gru_cell(const float *w_ih, const float *w_hh,
                            const float *b_ih, const float *b_hh,
                            const float *x, float *h, int in_dim, int hidden,
                            uint32_t flags);

/* ----------------------------------------------------------- device model */
typedef struct b200gru_model b200gru_model; /* opaque */

/* precision / kernel family */
#define B200GRU_MODE_EXACT 0 /* fp32 SIMT, bit-identical to oracle A           */
#define B200GRU_MODE_TC 1    /* bf16 tcgen05/TMEM/TMA, fp32 accumulate         */

/* namegen_initialize() equivalent (ref:359-618): upload + repack on `device`.
 * `params` is the host weight file contents.  `max_batch` bounds the names a
 * single generate() call may run per rank (activation buffers are sized once,
 * ref:409-468, 583-612). */
int b200gru_model_create(const b200gru_config *cfg, const float *params,
                         int device, int max_batch, b200gru_model **out);
/* Tensor-parallel variant (not in the reference, which replicates all weights
 * on every rank, ref:370-407): the hidden dimension of every layer is split
 * over `tp_world` GPUs; this rank packs and keeps only the weight rows of its
 * own hidden units.  Hidden-state buffers are full width on every rank and
 * live in a peer-visible region the caller provides:
 *   1. create on every rank, 2. allocate b200gru_tp_shared_bytes() with
 *   b200gru_symm_alloc and exchange handles, 3. b200gru_tp_bind().
 * Each layer kernel then stores its slice of h' into every rank's buffer
 * (NVLink P2P) and the next kernel's TMA producer consumes remote K-chunks as
 * their counters arrive: the all-gather is fused into the GEMM's K loop. */
int b200gru_model_create_tp(const b200gru_config *cfg, const float *params,
                            int device, int max_batch, int tp_rank,
                            int tp_world, b200gru_model **out);
size_t b200gru_tp_shared_bytes(const b200gru_model *m);
/* peer_bases[r] = mapped address of rank r's region (entry tp_rank ignored). */
int b200gru_tp_bind(b200gru_model *m, void *local_base, void *const *peer_bases);
/* Same, plus `mc_base`: the address of an NVSwitch multicast mapping
 * (cuMulticastCreate / BindMem, e.g. torch symmetric memory's multicast_ptr) of
 * the SAME region on all ranks.  With it the all-gather of a layer's new hidden
 * slice is one multimem.st per 16 bytes (the switch replicates it into every
 * rank's buffer) and one multimem.red per tile counter, instead of one store /
 * reduction per peer: NVLink egress per call drops by (tp_world - 1)x.
 * mc_base == NULL is b200gru_tp_bind.  The reference has no counterpart (every
 * rank holds all weights, ref:370-407). */
int b200gru_tp_bind_mc(b200gru_model *m, void *local_base, void *const *peer_bases,
                       void *mc_base);
/* namegen_finalize() equivalent (ref:897-1033). */
int b200gru_model_destroy(b200gru_model *m);
int64_t b200gru_model_device_bytes(const b200gru_model *m);

/* Where a generate call gets its random floats and puts its characters.  All
 * pointers are device addresses and may be peer (NVLink-mapped) addresses:
 * that is how the reference's MPI_Scatter (ref:636) and MPI_Gather (ref:889)
 * are folded into the kernels.  `rfloats` points at this rank's first float
 * (root buffer + start*max_len), `output` at this rank's first row
 * (root buffer + start*(max_len+1)).  If `done_flag` is non-null the last
 * kernel release-stores `done_value` to it (system scope) after its output
 * rows are written.  If `ready_flag` is non-null, the random floats are only
 * read once `*ready_flag == ready_value` (acquire, system scope): the wait sits
 * right before the first sampling kernel, so the GRU kernels of the first
 * character already run while the root's host-to-device copy of the floats is
 * still in flight. */
typedef struct b200gru_io {
  const float *rfloats;
  char *output;
  uint32_t *done_flag;
  uint32_t done_value;
  const uint32_t *ready_flag;
  uint32_t ready_value;
  void *ready_event; /* same idea within one process: a cudaEvent_t recorded
                        after the upload of the floats on another stream */
} b200gru_io;

/* namegen() equivalent for one rank's slice (ref:627-890, minus the MPI
 * calls): generate `n` names.  Asynchronous on `stream`. */
int b200gru_generate(b200gru_model *m, int mode, int n, const b200gru_io *io,
                     void *stream);
/* Number of kernel launches the last b200gru_generate() issued. */
int b200gru_last_launch_count(const b200gru_model *m);
/* Debug taps.  set_debug(m, 1) makes the tensor-core FC kernel also store its
 * pre-softmax logits (costly: off by default); debug_logits copies the last
 * step's logits [n, num_char] (fp32).  The exact path always keeps them. */
int b200gru_model_set_debug(b200gru_model *m, int logits_tap);
int b200gru_debug_logits(b200gru_model *m, float *dst_device, int n,
                         void *stream);

/* Kernel path of tensor-core mode.  The reference's name loop, step loop and 51
 * launches per character (ref:649, :661, :664-873) exist here in two shapes:
 *   multi-launch (1): one launch per GRU layer per character + FC/sample, weight
 *     tiles streamed from L2 — fastest for large batches;
 *   persistent (2): ONE cooperative launch per call; every CTA pair keeps its
 *     16-unit slice of all recurrent weights in shared memory for the whole call,
 *     layers / FC / sampling are chained by per-tile counters — fastest for
 *     small batches.  Needs all slices to fit (e.g. H <= 1024 with 2 layers).
 * path 0 (default) picks persistent when it is available and
 * n <= persist_max_names (default 512; < 0 leaves the threshold unchanged).
 * Returns 1 if the persistent kernel is available for this model, 0 if not.
 * Environment: B200GRU_PATH, B200GRU_PERSIST_MAX set the defaults at create. */
int b200gru_model_set_path(b200gru_model *m, int path, int persist_max_names);

/* Multi-launch tensor-core path: pack the names that are still running to the
 * front of the batch after steps 1, 3, 5, ... so that finished names stop
 * occupying GEMM rows (the reference stops a name at EOS, ref:881-882).  Each
 * point is three small launches that do nothing unless a whole 256-name tile
 * can be dropped; rows of the output keep their positions (a row map goes with
 * the names).  On by default (B200GRU_COMPACT=0 at create turns it off); off
 * while the logits tap is on and for tensor-parallel models.  The persistent
 * path skips finished 256-name tiles instead. */
int b200gru_model_set_compaction(b200gru_model *m, int on);

/* Debug build (libb200gru_debug.so, compiled with -DB200GRU_DEBUG; selected by
 * B200GRU_DEBUG=1 in the Python package).  It stands in for compute-sanitizer,
 * which this GPU pool does not allow: device-side assertions on computed global
 * offsets, shared-memory ring / TMEM stage indices and sampled ids; every model
 * allocation between two 4 KB guard bands that are verified after each
 * single-rank generate (which therefore synchronises), by b200gru_model_check and
 * at destroy; every spin loop (mbarrier, rank flags, tensor-parallel counters)
 * bounded by a 20 s deadline that traps with a message.  The reference's own bugs
 * of this kind: ref:294-300 (race), ref:1017 (cudaFree of a host pointer).
 * b200gru_debug_build() is 1 in that library, 0 in the release one;
 * b200gru_model_check() verifies the guard bands now (always 0 in release). */
int b200gru_debug_build(void);
int b200gru_model_check(b200gru_model *m);

/* Tracing (absent in the reference): with set_trace(m, 1) the tensor-core
 * generate records a CUDA event after every launch; get_trace returns the
 * per-launch device times of the last call (synchronises on its last event). */
int b200gru_model_set_trace(b200gru_model *m, int on);
int b200gru_model_get_trace(b200gru_model *m, float *ms, const char **names,
                            int max_spans);

/* Persistent path only: in-kernel timeline of the last generate run with
 * set_trace(m, 1).  `out` receives 2 x 3 x 256 words, [cta: first weight CTA,
 * first FC CTA][role: TMA producer, MMA issuer, epilogue][event]; a word is
 * (globaltimer_ns << 16) | (step << 8) | (layer << 4) | event, 0 = unused.
 * Events: 1 operand counter reached, 2 loads issued, 3 first operand landed,
 * 4 MMAs issued, 5 accumulator ready, 6 result published, 7 characters of the
 * previous step visible, 8 role started, 9 resident weights landed.
 * Returns the number of words written (0: nothing recorded). */
int b200gru_model_persist_timeline(b200gru_model *m, unsigned long long *out, int max_words);

/* Kernel that spins (ld.acquire.sys) until flags[i] >= value (wrap-safe) for all i<n;
 * rank 0 runs it after its own generate to complete the gather. */
int b200gru_wait_flags(const uint32_t *flags, int n, uint32_t value,
                       void *stream);

/* ------------------------------------------- peer-visible buffers (NVLink) */
/* cudaMalloc + zero + CUDA IPC handle (64 bytes).  Stands in for the host
 * staging buffers of MPI_Scatter / MPI_Gather (ref:632-636, :889): peers map
 * the buffer and kernels load/store it directly. */
int b200gru_symm_alloc(int device, size_t bytes, void **ptr,
                       unsigned char *handle64);
int b200gru_symm_open(int device, const unsigned char *handle64, void **ptr);
int b200gru_symm_close(int device, void *ptr);
int b200gru_symm_free(int device, void *ptr);
/* kind: 1 = H2D, 2 = D2H, 3 = D2D */
int b200gru_memcpy_async(void *dst, const void *src, size_t bytes, int kind,
                         void *stream);
int b200gru_memset_async(void *dst, int value, size_t bytes, void *stream);
/* st.release.sys of `value` to `flag` after prior work on `stream`. */
int b200gru_signal_flag(uint32_t *flag, uint32_t value, void *stream);
int b200gru_can_access_peer(int device, int peer);
/* For C callers that do not link the CUDA runtime themselves. */
int b200gru_stream_synchronize(void *stream);
int b200gru_device_count(void); /* 0 when there is no usable GPU / driver */

/* Diagnostic: write `rows` x `row_bytes` (row pitch `pitch`) from `src` to
 * `ndst` destination buffers with coalesced 16-byte stores; tools/p2p_bench.py
 * times it to measure NVLink write bandwidth per access pattern. */
int b200gru_p2p_write_bench(void *const *dsts, int ndst, const void *src,
                            long long rows, int row_bytes, long long pitch,
                            int blocks, void *stream);

/* ------------------------------------------------- stand-alone device ops */
/* D[M,N] (fp32) = A[M,K] (bf16, row-major) x B[N,K]^T (bf16, row-major) on
 * tcgen05; used by tests to validate descriptors / TMA / TMEM plumbing. */
int b200gru_gemm_bf16_tn(const void *a, const void *b, float *d, int m, int n,
                         int k, void *stream);

/* One GRU layer step for a batch in fp32, the reference's association order
 * (ref:671-763): h_out[b] = GRUCell(x[b], h_in[b]).  x [batch,in_dim],
 * h_* [batch,hidden], w_ih [3*hidden,in_dim], w_hh [3*hidden,hidden] with the
 * r,z,n blocks stacked, b_* [3*hidden]; all device pointers.  Allocates (and
 * frees) 2*batch*3*hidden floats of scratch for the two GEMM outputs, and
 * synchronises `stream` before returning. */
int b200gru_gru_cell_fp32(const float *x, const float *h_in, const float *w_ih,
                          const float *w_hh, const float *b_ih,
                          const float *b_hh, float *h_out, int batch,
                          int in_dim, int hidden, void *stream);

/* Same op for callers that cannot allocate device memory at call time (all of
 * HBM already belongs to another pool, or cudaMalloc's device-wide sync is not
 * acceptable in the loop).  No counterpart in the reference.  The GEMM scratch
 * lives in a fixed B200GRU_NOALLOC_SCRATCH_BYTES array inside the library
 * image (resident on every GPU that loads it, until exit), so:
 *   - 2*batch*3*hidden*4 bytes must fit, else B200GRU_ERR_ARG with a message;
 *   - all calls on one GPU share the array: calls that may overlap there race;
 *   - the first call on a device resolves the array's address; make it during
 *     set-up, before any stream capture;
 *   - the current device must be the one that owns `stream` and all seven
 *     buffers (the scratch is the current device's copy): device pointers of
 *     another GPU are refused with B200GRU_ERR_ARG.
 * Operands need only 4-byte alignment (16-byte aligned rows run faster).
 * Asynchronous on `stream`; inputs and outputs are the caller's, as above. */
#define B200GRU_NOALLOC_SCRATCH_BYTES (4u << 20)
int b200gru_gru_cell_fp32_noalloc(const float *x, const float *h_in,
                                  const float *w_ih, const float *w_hh,
                                  const float *b_ih, const float *b_hh,
                                  float *h_out, int batch, int in_dim,
                                  int hidden, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200GRU_H_ */
