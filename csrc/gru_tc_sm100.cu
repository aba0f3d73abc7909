// gru_tc_sm100.cu — the Blackwell hot path: batched GRU generation on
// tcgen05 tensor cores with TMA-fed operands and TMEM accumulators.
//
// What the reference does per character of ONE name — 13 matvec launches,
// 34 elementwise launches, 2 softmax launches, a D2H copy and a CPU sampler
// (ref:664-878) — becomes, for ALL of a rank's names at once, one launch per
// GRU layer plus one for FC+softmax+sample.
//
// Every GRU layer is the same fused GEMM + gate kernel (gru_layer_kernel):
//
//   acc[128 names, 256 cols] per 64-unit block, columns [ n_i | r | z | n_h ]
//     x-part : [n_i|r|z] (cols   0..191)  = x . Wx^T   one UMMA N=192 per k16
//     h-part : [r|z|n_h] (cols  64..255) += h . Wh^T   one UMMA N=192 per k16
//   r and z receive both halves in the same TMEM columns; n_i and n_h stay
//   apart because r multiplies only the recurrent half (ref:739-744).  The
//   n_h columns are zeroed (tcgen05.st) by the epilogue when it hands a TMEM
//   stage back, so every recurrent MMA can simply accumulate.
//   epilogue: r = s(acc_r + b_r), z = s(acc_z + b_z),
//             n = tanh(acc_ni + b_in + r * (acc_nh + b_hn)), h' = n + z (h - n)
//
//   layer 0: there is no x-part GEMM.  The embedding lookup and the three input
//            matvecs (ref:669, :676, :702, :729) only depend on the character
//            id, so they are the table T0[c] = W_i*0 . emb[c] + biases computed
//            once at init; the epilogue thread of (name, 32 units) fetches its
//            192 bytes of the name's row (bf16, six 32-byte loads) and adds them
//            to the recurrent accumulators.  (An earlier version fed a one-hot
//            row through the tensor core as K = NUM_CHAR extra columns: 20% of
//            layer 0's MMA time spent multiplying by zeros.)
//   layer l: x is the layer below's fresh hidden state (ref:773, :797, :821).
//
// FC (fc_sample_kernel): acc[128 names, C] = h_top . W_fc^T, then per name
// +b_fc, exp, full sum (no max subtraction, ref:302-313), inverse-CDF walk in
// index order with strict '>' and C-1 fallback (ref:322-333), output byte,
// next-step one-hot row, EOS bookkeeping (ref:876-882).
//
// Kernel anatomy (one CTA per SM, persistent over tiles):
//   warp 0    TMA producer   cp.async.bulk.tensor -> 128B-swizzled smem ring
//   warp 1    MMA issuer     tcgen05.mma kind::f16, M=128
//   warp 2    TMEM allocator 512 columns = 2 accumulator stages of 256
//   warp 4..  epilogue       tcgen05.ld -> gate math in registers -> global
// The accumulator is double-buffered in TMEM so the epilogue of tile i
// overlaps the MMAs of tile i+1.
#include <cstdlib>

#include "model.hpp"
#include "sm100_ptx.cuh"

namespace b200gru {
using namespace sm100;

constexpr int TC_BLOCK_K = 64;               // 128 bytes of bf16 per row
constexpr int TC_A_BYTES = kTileM * 128;     // 16 KB
constexpr int TC_TMEM_COLS = 512;
constexpr uint32_t kEpiSleepNs = 200;        // back-off of the epilogue's accumulator wait
constexpr uint32_t kProdSleepNs = 50;        // back-off of the producer's free-slot wait
constexpr int TC_ACC_STRIDE = 256;           // TMEM columns per accumulator

// ------------------------------------------------------------------ layers
constexpr int L_EPI_WARPS = 8;
constexpr int L_THREADS = (4 + L_EPI_WARPS) * 32;
constexpr int L_B_BYTES = 192 * 128;         // 24 KB
constexpr int L_STAGE_BYTES = TC_A_BYTES + L_B_BYTES;
constexpr int L_STAGES = 4;
constexpr int OUT_TILE_BYTES = kTileM * 128;  // 128 names x 64 units bf16, 128B-swizzled
// [align slack 1024][stages][barrier block 1024][2 staging tiles]
constexpr int OUT_TILES = 3;                  // staging tiles: the store of tile i-1 may still be reading while i+1 is written
constexpr int L_SMEM_BYTES = 1024 + L_STAGES * L_STAGE_BYTES + 1024 + OUT_TILES * OUT_TILE_BYTES;
constexpr int EPI_BAR_ID = 1;                 // named barrier of the epilogue warps


struct LayerArgs {
  int n;            // valid names (rows)
  // Names that have hit EOS are compacted away between steps (ref:881-882 stops
  // a name there): when non-null, the number of rows still in use is read from
  // device memory at kernel start and replaces n / m_tiles.
  const int *n_dev;
  int m_tiles;      // ceil(n / 128)
  int n_blocks;     // H / 64
  int kb_x;         // k-blocks (of 64) of the x operand
  int kb_h;         // k-blocks of the recurrent operand (0: h == 0, step 0)
  // Hidden buffers are stored as column panels: [H / pw panels][bpad rows][pw]
  // bf16 (pw = H is plain row-major).  Narrow panels make the 128-byte row
  // segments of a tile sit close together, which is what NVLink P2P writes
  // need (measured: 128 B segments at 16 KB pitch 180 GB/s, at <= 1 KB pitch
  // 450-600 GB/s); wide panels spread a TMA box over more L2 slices.
  int pw_log2;      // log2(panel width in elements)
  int bpad;         // rows per panel
  int h_full;       // full hidden width H (element extent of a hidden buffer = bpad * H; debug checks)
  // layer 0 only: the folded input table, [H/64][C][2 halves][2 chunks][n|r|z][16] bf16,
  // and the current character of every name
  const __nv_bfloat16 *t0q;
  const int32_t *chars;
  int num_char;
  const float *bias;            // [H/64, 4, 64]: b_in, b_r, b_z, b_hn
  const __nv_bfloat16 *h_prev;  // [bpad, H]
  __nv_bfloat16 *h_out;         // [bpad, H]
  // ---- tensor parallel over the hidden dimension (tp_world == 1: off).
  // This rank owns hidden columns [col0, col0 + 64*n_blocks).  Operand buffers
  // are full width on every rank; each rank's epilogue stores its slice into
  // every rank's copy (NVLink P2P) and bumps a per-(source rank, m-tile)
  // counter there; the TMA producer of the consuming kernel waits on the
  // counter of a source rank right before it loads that rank's K-chunk.
  int tp_world, tp_rank;
  int kb_rank;        // k-blocks per source rank in a full-width operand
  int x_full;         // x operand is a full-width hidden buffer (layers >= 1)
  int col0;
  int flag_stride;    // counters per source rank
  uint32_t expect_x, expect_h;
  const uint32_t *wait_x, *wait_h;  // local counters [src][m_tile]
  __nv_bfloat16 *h_out_peer[8];     // h_out on every rank ([tp_rank] == h_out)
  uint32_t *signal_peer[8];         // this rank's counter row [m_tile] on every rank
  // NVSwitch multicast views of the same two things (null: unicast to each peer).
  // One store / one reduction leaves this GPU and the switch replicates it into
  // every rank's copy: egress per tile drops from (world-1) copies to one.
  __nv_bfloat16 *h_out_mc;
  uint32_t *signal_mc;
};

// element offset of (row, col) in a panelled hidden buffer
__device__ __forceinline__ size_t h_offset(int row, int col, int pw_log2, int bpad) {
  const int panel = col >> pw_log2;
  return (((size_t)panel * bpad + row) << pw_log2) + (col & ((1 << pw_log2) - 1));
}
// TMA coordinates of the box starting at (row, col)
__device__ __forceinline__ void h_coords(int row, int col, int pw_log2, int bpad, int *ci, int *co) {
  const int panel = col >> pw_log2;
  *ci = col & ((1 << pw_log2) - 1);
  *co = panel * bpad + row;
}

// All lanes poll (one coalesced request); every lane has then acquired.
__device__ __forceinline__ void tp_wait_counter(const uint32_t *p, uint32_t expect) {
#ifdef B200GRU_DEBUG
  SpinGuard g;
  uint32_t v;
  while ((int32_t)((v = ld_acquire_sys(p)) - expect) < 0) {
    __nanosleep(64);
    g.tick("tensor-parallel tile counter", p, expect, v);
  }
#else
  while ((int32_t)(ld_acquire_sys(p) - expect) < 0) __nanosleep(64);
#endif
}

// 32-byte read-only global load (one request instead of two 16-byte ones: the
// table rows below are scattered per lane, so L1TEX wavefronts are what they cost)
__device__ __forceinline__ void ldg256(const void *p, uint32_t (&v)[8]) {
  asm volatile("ld.global.nc.L2::128B.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "l"(p));
}

// Gate math for one thread's (name, 32 units) slice of a finished tile:
// t addresses this warp's TMEM lanes at the slice's first n_i column.  The bf16
// result goes to the 128B-swizzled staging tile (row `r`, 16-byte chunks
// half*4 .. half*4+3), from where one thread TMA-stores the whole tile.
//   LAYER0: the x-part is not in TMEM; it is this name's row of the folded input
//           table (t0 -> [2 chunks][n|r|z][16] bf16, biases included), and the
//           accumulator only exists if the step has a recurrent half (has_acc).
//           Nothing needs re-zeroing: the tile's first MMA overwrites.
struct T0Rows { uint32_t v[2][3][8]; };  // [chunk][n|r|z][8 bf16 pairs] of one (name, 32 units)
__device__ __forceinline__ void load_t0_rows(const __nv_bfloat16 *t0, T0Rows &q) {
#pragma unroll
  for (int ch = 0; ch < 2; ++ch)
#pragma unroll
    for (int g = 0; g < 3; ++g) ldg256(t0 + ch * 48 + g * 16, q.v[ch][g]);
}

template <bool LAYER0>
__device__ __forceinline__ void gate_epilogue(uint32_t t, const uint4 (&hq)[4],
                                              const float4 *bias4, uint8_t *stage_tile,
                                              int r, int half, const T0Rows &tq, bool has_acc) {
#pragma unroll
  for (int ch = 0; ch < 2; ++ch) {
    uint32_t vn[16], vr[16], vz[16], vh[16];
    const uint32_t (&tn)[8] = tq.v[ch][0];
    const uint32_t (&tr)[8] = tq.v[ch][1];
    const uint32_t (&tz)[8] = tq.v[ch][2];
    if (!LAYER0 || has_acc) {
      if (!LAYER0) tmem_ld16(t + ch * 16, vn);
      tmem_ld16(t + 64 + ch * 16, vr);
      tmem_ld16(t + 128 + ch * 16, vz);
      tmem_ld16(t + 192 + ch * 16, vh);
      tmem_ld_wait();
    } else {
#pragma unroll
      for (int i = 0; i < 16; ++i) vr[i] = vz[i] = vh[i] = 0u;
    }
    const uint32_t hw[8] = {hq[2 * ch].x,     hq[2 * ch].y,     hq[2 * ch].z,
                            hq[2 * ch].w,     hq[2 * ch + 1].x, hq[2 * ch + 1].y,
                            hq[2 * ch + 1].z, hq[2 * ch + 1].w};
    uint32_t packed[8];
#pragma unroll
    for (int q = 0; q < 4; ++q) {  // 4 units per float4 of bias
      float bi[4] = {0.f, 0.f, 0.f, 0.f}, br[4] = {0.f, 0.f, 0.f, 0.f}, bz[4] = {0.f, 0.f, 0.f, 0.f};
      if (!LAYER0) {
        const float4 b_in = __ldg(bias4 + ch * 4 + q);
        const float4 b_r = __ldg(bias4 + 16 + ch * 4 + q);
        const float4 b_z = __ldg(bias4 + 32 + ch * 4 + q);
        bi[0] = b_in.x; bi[1] = b_in.y; bi[2] = b_in.z; bi[3] = b_in.w;
        br[0] = b_r.x; br[1] = b_r.y; br[2] = b_r.z; br[3] = b_r.w;
        bz[0] = b_z.x; bz[1] = b_z.y; bz[2] = b_z.z; bz[3] = b_z.w;
      }
      const float4 b_hn = __ldg(bias4 + 48 + ch * 4 + q);
      const float bh[4] = {b_hn.x, b_hn.y, b_hn.z, b_hn.w};
      float hn[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int i = q * 4 + e;
        const uint32_t w = hw[i >> 1];
        const float hp = __uint_as_float((i & 1) ? (w & 0xFFFF0000u) : (w << 16));
        float xn, xr, xz;
        if (LAYER0) {   // table row: bf16 pairs, biases already folded in
          const uint32_t wn = tn[i >> 1], wr = tr[i >> 1], wz = tz[i >> 1];
          xn = __uint_as_float((i & 1) ? (wn & 0xFFFF0000u) : (wn << 16));
          xr = __uint_as_float((i & 1) ? (wr & 0xFFFF0000u) : (wr << 16));
          xz = __uint_as_float((i & 1) ? (wz & 0xFFFF0000u) : (wz << 16));
        } else {        // x-part sits in the same accumulator columns as the h-part (r, z) / next to it (n)
          xn = __uint_as_float(vn[i]) + bi[e];
          xr = br[e];
          xz = bz[e];
        }
        const float rr = sigmoid_approx(__uint_as_float(vr[i]) + xr);
        const float z = sigmoid_approx(__uint_as_float(vz[i]) + xz);
        const float nn = tanh_approx(fmaf(rr, __uint_as_float(vh[i]) + bh[e], xn));
        // h' = (1-z)*n + z*h  ==  n + z*(h - n)     (ref:752-763)
        hn[e] = fmaf(z, hp - nn, nn);
      }
      __nv_bfloat162 p0 = __floats2bfloat162_rn(hn[0], hn[1]);
      __nv_bfloat162 p1 = __floats2bfloat162_rn(hn[2], hn[3]);
      packed[q * 2] = *reinterpret_cast<uint32_t *>(&p0);
      packed[q * 2 + 1] = *reinterpret_cast<uint32_t *>(&p1);
    }
    // chunks c0, c0+1 of this row, XOR-swizzled like CU_TENSOR_MAP_SWIZZLE_128B
    const int c0 = half * 4 + ch * 2;
    uint8_t *rowp = stage_tile + r * 128;
    *reinterpret_cast<uint4 *>(rowp + (((c0) ^ (r & 7)) << 4)) =
        make_uint4(packed[0], packed[1], packed[2], packed[3]);
    *reinterpret_cast<uint4 *>(rowp + (((c0 + 1) ^ (r & 7)) << 4)) =
        make_uint4(packed[4], packed[5], packed[6], packed[7]);
  }
  if (!LAYER0) {
    // re-zero this warp's n_h columns before the stage goes back to the MMAs
    uint32_t zero[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) zero[i] = 0u;
    tmem_st16(t + 192, zero);
    tmem_st16(t + 192 + 16, zero);
    tmem_st_wait();
  }
}

// Tail of a tile's epilogue, run by every epilogue thread (tid = 0..255).
//  * local copy: one thread TMA-stores the staged tile into this rank's h_out;
//  * tensor parallel: all 256 threads copy the staged tile to every peer's
//    h_out with coalesced 16-byte stores (a warp covers four full 128-byte
//    rows) through the LSU, which keeps far more NVLink writes in flight than
//    the TMA store path does (measured: ~1 GB/s per SM for remote TMA stores);
//    each warp then hands the tile to the publisher warp through an mbarrier.
struct PubRing {
  uint64_t *full;    // [2] count 8: one arrive per epilogue warp, after its peer stores
  uint64_t *empty;   // [2] count 1: publisher has read the slot
  int *m_tile;       // [2]
};
__device__ __forceinline__ void epilogue_store_tile(const CUtensorMap *map_out, const LayerArgs &args,
                                                    uint8_t *stage_tile, int col, int m_tile,
                                                    bool issuer, int tid, const PubRing &pub, int it) {
  // (every thread has already fenced its staging writes towards the async proxy)
  named_bar_sync(EPI_BAR_ID, L_EPI_WARPS * 32);   // the staged tile is complete
  if (issuer) {
    int ci, co;
    h_coords(m_tile * kTileM, col, args.pw_log2, args.bpad, &ci, &co);
    tma_store_2d(map_out, stage_tile, ci, co);
    tma_store_commit();
    // With three staging tiles this is the only reuse guard needed: the store of
    // the PREVIOUS tile has finished reading before this thread reaches the next
    // tile's barrier, i.e. before anyone writes the tile after that (same buffer).
    tma_store_wait_read<1>();
  }
  if (args.tp_world > 1) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int idx = tid + j * (L_EPI_WARPS * 32);  // 16-byte chunk of the 16 KB tile
      const int r = idx >> 3, c = idx & 7;
      const uint4 v = *reinterpret_cast<const uint4 *>(stage_tile + r * 128 + ((c ^ (r & 7)) << 4));
      const size_t off = h_offset(m_tile * kTileM + r, col + c * 8, args.pw_log2, args.bpad);
      B200_DASSERT(off + 8 <= (size_t)args.bpad * args.h_full, "peer store offset %llu beyond hidden buffer (%d x %d)",
                   (unsigned long long)off, args.bpad, args.h_full);
      if (args.h_out_mc) {
        multimem_st_v4(args.h_out_mc + off, v);   // one NVLink store, replicated by the switch
      } else {
        for (int d = 0; d < args.tp_world; ++d)
          if (d != args.tp_rank) *reinterpret_cast<uint4 *>(args.h_out_peer[d] + off) = v;
      }
    }
    // hand over to the publisher: slot (it & 1), reused every second tile
    const int slot = it & 1;
    const uint32_t ph = (uint32_t)(it >> 1) & 1;
    __syncwarp();
    if ((tid & 31) == 0) {
      mbar_wait(&pub.empty[slot], ph ^ 1);
      if (tid == 0) pub.m_tile[slot] = m_tile;
      mbar_arrive(&pub.full[slot]);  // release (CTA scope): orders this warp's peer stores
    }
  }
}
// Publisher (warp 3, one lane): for every tile of this CTA, once all epilogue
// warps have issued their peer stores, fence system-wide — this is where the
// NVLink write acknowledgements are waited for, off the epilogue's critical
// path — and bump the tile's counter on every rank.
__device__ __forceinline__ void publish_tile(const LayerArgs &args, const PubRing &pub, int it) {
  const int slot = it & 1;
  const uint32_t ph = (uint32_t)(it >> 1) & 1;
  mbar_wait(&pub.full[slot], ph);
  const int m_tile = pub.m_tile[slot];
  mbar_arrive(&pub.empty[slot]);
  __threadfence_system();
  if (args.signal_mc) {
    multimem_red_release_sys_add(args.signal_mc + m_tile, 1u);  // bumps every rank's counter at once
  } else {
    for (int d = 0; d < args.tp_world; ++d)
      red_release_sys_add(args.signal_peer[d] + m_tile, 1u);
  }
}

// ------------------------------------------------------------ TMA producer
// Streams one operand pair (A rows of `m_tile`, weight rows from `b_row`) for
// `nkb` k-blocks into the smem ring.  Run by a whole warp; one elected lane
// issues.  Everything that varies per k-block is kept as running coordinates
// so the loop body is a wait, two TMA issues and three adds.
//   full  : A is a full-width hidden buffer (panel layout; under tensor
//           parallelism it arrives in per-rank chunks that are waited for).
//   PAIR  : cta_group::2 — bytes are credited to the leader CTA's barrier.
struct Ring { int stage; uint32_t phase; };
template <bool PAIR, int STAGES, int STAGE_BYTES>
__device__ __forceinline__ void produce_part(const CUtensorMap *map_a, const CUtensorMap *map_w,
                                             int nkb, bool full, const uint32_t *wait,
                                             uint32_t expect, int m_tile, int b_row,
                                             const LayerArgs &args, uint8_t *smem,
                                             uint64_t *full_bar, uint64_t *empty_bar, bool leader,
                                             Ring &ring) {
  if (nkb == 0) return;  // step 0: h == 0, no recurrent operand
  const bool tp = full && args.tp_world > 1;
  const int nchunks = tp ? args.tp_world : 1;
  const int chunk_len = tp ? args.kb_rank : nkb;
  const int pw = full ? (1 << args.pw_log2) : 0x7fffffff;
  for (int c = 0; c < nchunks; ++c) {
    // start with the chunk this rank produced itself, then walk the ring
    int src = args.tp_rank + c;
    if (src >= args.tp_world) src -= args.tp_world;
    const int kb0 = tp ? src * args.kb_rank : 0;
    bool fence = false;
    if (tp && src != args.tp_rank) {
      tp_wait_counter(wait + src * args.flag_stride + m_tile, expect);
      fence = true;  // remote generic-proxy writes -> this chunk's TMA reads
    }
    int kw = kb0 * TC_BLOCK_K;
    int ci = kw, co = m_tile * kTileM;
    if (full) {
      ci = kw & (pw - 1);
      co += (kw >> args.pw_log2) * args.bpad;
    }
    for (int jj = 0; jj < chunk_len; ++jj) {
      B200_DASSERT(ring.stage >= 0 && ring.stage < STAGES, "smem ring stage %d of %d", ring.stage, STAGES);
      mbar_wait_relaxed(&empty_bar[ring.stage], ring.phase ^ 1, kProdSleepNs);
      if (elect_one()) {
        uint8_t *sa = smem + ring.stage * STAGE_BYTES;
        uint8_t *sb = sa + TC_A_BYTES;
        if (fence) fence_proxy_async();
        if (PAIR) {
          if (leader) mbar_expect_tx(&full_bar[ring.stage], 2 * STAGE_BYTES);
          const uint32_t bar = mapa_u32(&full_bar[ring.stage], 0);
          tma_load_2d_2sm(sa, map_a, bar, ci, co);
          tma_load_2d_2sm(sb, map_w, bar, kw, b_row);
        } else {
          mbar_expect_tx(&full_bar[ring.stage], STAGE_BYTES);
          tma_load_2d(sa, map_a, &full_bar[ring.stage], ci, co);
          tma_load_2d(sb, map_w, &full_bar[ring.stage], kw, b_row);
        }
      }
      __syncwarp();
      fence = false;
      if (++ring.stage == STAGES) { ring.stage = 0; ring.phase ^= 1; }
      kw += TC_BLOCK_K;
      ci += TC_BLOCK_K;
      if (ci == pw) { ci = 0; co += args.bpad; }
    }
  }
}

template <bool LAYER0>
__global__ void __launch_bounds__(L_THREADS, 1)
gru_layer_kernel(const __grid_constant__ CUtensorMap map_x,
                 const __grid_constant__ CUtensorMap map_h,
                 const __grid_constant__ CUtensorMap map_wx,
                 const __grid_constant__ CUtensorMap map_wh,
                 const __grid_constant__ CUtensorMap map_out, const LayerArgs args) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + L_STAGES * L_STAGE_BYTES);
  uint64_t *full_bar = bars;                       // [STAGES]
  uint64_t *empty_bar = bars + L_STAGES;           // [STAGES]
  uint64_t *tmem_full = bars + 2 * L_STAGES;       // [2]
  uint64_t *tmem_empty = bars + 2 * L_STAGES + 2;  // [2]
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * L_STAGES + 4);
  const PubRing pub{bars + 2 * L_STAGES + 5, bars + 2 * L_STAGES + 7,
                    reinterpret_cast<int *>(bars + 2 * L_STAGES + 9)};

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int kb_total = args.kb_x + args.kb_h;

  if (warp == 0 && lane == 0) {
    if (args.kb_x) { tma_prefetch_desc(&map_x); tma_prefetch_desc(&map_wx); }
    if (args.kb_h) { tma_prefetch_desc(&map_h); tma_prefetch_desc(&map_wh); }
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < L_STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], L_EPI_WARPS);  // one arrive per epilogue warp
      mbar_init(&pub.full[i], L_EPI_WARPS);
      mbar_init(&pub.empty[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<TC_TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // The n_h columns of both accumulator stages start at zero (the recurrent
  // MMAs only ever accumulate into them); later the epilogue re-zeroes them
  // each time it hands a stage back.
  if (warp >= 4) {
    const uint32_t lane_base0 = (uint32_t)((warp & 3) * 32) << 16;
    const int half0 = (warp - 4) >> 2;
    uint32_t zero[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) zero[i] = 0u;
#pragma unroll
    for (int st = 0; st < 2; ++st)
#pragma unroll
      for (int c = 0; c < 2; ++c)
        tmem_st16(tmem_base + lane_base0 + st * TC_ACC_STRIDE + 192 + half0 * 32 + c * 16, zero);
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // Everything above touched only this CTA's smem / TMEM; it may overlap the
  // tail of the previous kernel (programmatic dependent launch).  From here on
  // global data written by earlier kernels is read.
  pdl_launch_dependents();
  pdl_wait();
  // rows in use: fixed for the call, or what the last compaction left
  const int n_rows = args.n_dev ? __ldcg(args.n_dev) : args.n;
  const int total_tiles = ((n_rows + kTileM - 1) / kTileM) * args.n_blocks;

  if (warp == 0) {
    // ===================================================== TMA producer
    Ring ring{0, 0};
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int m_tile = tile / args.n_blocks;
      const int n_blk = tile % args.n_blocks;
      produce_part<false, L_STAGES, L_STAGE_BYTES>(&map_x, &map_wx, args.kb_x, args.x_full != 0,
                                                   args.wait_x, args.expect_x, m_tile, n_blk * 192,
                                                   args, smem, full_bar, empty_bar, true, ring);
      produce_part<false, L_STAGES, L_STAGE_BYTES>(&map_h, &map_wh, args.kb_h, true, args.wait_h,
                                                   args.expect_h, m_tile, n_blk * 192, args, smem,
                                                   full_bar, empty_bar, true, ring);
    }
  } else if (warp == 1) {
    // ======================================================= MMA issuer
    int stage = 0;
    uint32_t phase = 0;
    int acc_stage = 0;
    uint32_t acc_phase = 0;
    const uint32_t idesc = umma_idesc_bf16(kTileM, 192);
    const uint64_t desc0 = umma_desc_sw128(smem_u32(smem));
    // layer 0, step 0 has no GEMM at all (no x-part, h == 0): nothing to issue
    for (int tile = blockIdx.x; tile < total_tiles && kb_total > 0; tile += gridDim.x) {
      mbar_wait(&tmem_empty[acc_stage], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t d = tmem_base + acc_stage * TC_ACC_STRIDE;
      B200_DASSERT((d & 0xffffu) + 256 <= (tmem_base & 0xffffu) + TC_TMEM_COLS && acc_stage < 2 && stage < L_STAGES,
                   "accumulator stage %d / smem stage %d out of range", acc_stage, stage);
      for (int kb = 0; kb < kb_total; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t da = desc0 + (uint64_t)((stage * L_STAGE_BYTES) >> 4);
          const uint64_t db = da + (uint64_t)(TC_A_BYTES >> 4);
          if (kb < args.kb_x) {  // [n_i | r | z]: the tile's first MMA overwrites
            umma_bf16(d, da, db, idesc, kb != 0 ? 1u : 0u);
            umma_bf16(d, da + 2, db + 2, idesc, 1u);
            umma_bf16(d, da + 4, db + 4, idesc, 1u);
            umma_bf16(d, da + 6, db + 6, idesc, 1u);
          } else {               // [r | z | n_h]: accumulates (n_h pre-zeroed); layer 0 has no
                                 // x-part, there the first MMA overwrites all three
            umma_bf16(d + 64, da, db, idesc, (args.kb_x == 0 && kb == 0) ? 0u : 1u);
            umma_bf16(d + 64, da + 2, db + 2, idesc, 1u);
            umma_bf16(d + 64, da + 4, db + 4, idesc, 1u);
            umma_bf16(d + 64, da + 6, db + 6, idesc, 1u);
          }
          umma_commit(&empty_bar[stage]);  // smem slot free when MMAs retire
          if (kb == kb_total - 1) umma_commit(&tmem_full[acc_stage]);  // accumulator ready
        }
        __syncwarp();
        if (++stage == L_STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
    }
  } else if (warp == 3) {
    // ======================================= publisher (tensor parallel only)
    if (args.tp_world > 1 && lane == 0) {
      int it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it)
        publish_tile(args, pub, it);
    }
  } else if (warp >= 4) {
    // ========================================================= epilogue
    // 8 warps: warp%4 picks the TMEM lane quarter (hardware rule), warp/4
    // picks which 32 of the tile's 64 units this warp handles.
    const int quarter = warp & 3;
    const int half = (warp - 4) >> 2;
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    int acc_stage = 0;
    uint32_t acc_phase = 0;
    uint8_t *stage_out = smem + L_STAGES * L_STAGE_BYTES + 1024;  // after the barrier block
    const bool issuer = (warp == 4 && lane == 0);
    // Operands the gate math needs besides the accumulator — the previous state
    // of this thread's 32 units and, for layer 0, its 192 bytes of the name's row
    // of the folded input table — are requested ONE TILE AHEAD, right after the
    // gate math of the tile before: when the epilogue is what paces the kernel
    // the accumulator is always ready, so loads issued at the top of a tile would
    // have nothing to hide behind.  The character itself is fetched two tiles ahead.
    const int r = quarter * 32 + lane;
    const bool has_h = args.kb_h > 0;
    uint4 hq[4];
    T0Rows tq;
    int c_next = 0;
    auto char_of = [&](int tile_q) -> int {
      if (!LAYER0 || tile_q >= total_tiles) return 0;
      const int row_q = (tile_q / args.n_blocks) * kTileM + r;
      return row_q < n_rows ? args.chars[row_q] : 0;
    };
    auto request = [&](int tile_q, int c) {
      if (tile_q >= total_tiles) return;
      const int mq = tile_q / args.n_blocks, nq = tile_q % args.n_blocks;
      if (has_h) {
        B200_DASSERT(h_offset(mq * kTileM + r, args.col0 + nq * 64 + half * 32, args.pw_log2, args.bpad) + 32 <=
                         (size_t)args.bpad * args.h_full,
                     "h_prev offset beyond hidden buffer (tile %d)", tile_q);
        const uint4 *src = reinterpret_cast<const uint4 *>(
            args.h_prev + h_offset(mq * kTileM + r, args.col0 + nq * 64 + half * 32, args.pw_log2, args.bpad));
#pragma unroll
        for (int i = 0; i < 4; ++i) hq[i] = src[i];
      }
      B200_DASSERT(!LAYER0 || (c >= 0 && c < args.num_char && nq < args.n_blocks),
                   "layer-0 table row out of range: char %d of %d, unit block %d of %d", c, args.num_char, nq,
                   args.n_blocks);
      if (LAYER0) load_t0_rows(args.t0q + (((size_t)nq * args.num_char + c) * 2 + half) * 96, tq);
    };
#pragma unroll
    for (int i = 0; i < 4; ++i) hq[i] = make_uint4(0, 0, 0, 0);
    request((int)blockIdx.x, char_of((int)blockIdx.x));
    c_next = char_of((int)blockIdx.x + (int)gridDim.x);
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int m_tile = tile / args.n_blocks;
      const int n_blk = tile % args.n_blocks;
      const int col = args.col0 + n_blk * 64;
      const float4 *bias4 =
          reinterpret_cast<const float4 *>(args.bias + (size_t)n_blk * 4 * 64 + half * 32);
      uint8_t *stage_tile = stage_out + (it % OUT_TILES) * OUT_TILE_BYTES;

      const bool has_acc = kb_total > 0;
      if (has_acc) {
        mbar_wait_relaxed(&tmem_full[acc_stage], acc_phase, kEpiSleepNs);
        tc_fence_after();
      }
      const uint32_t t = tmem_base + lane_base + acc_stage * TC_ACC_STRIDE + half * 32;
      gate_epilogue<LAYER0>(t, hq, bias4, stage_tile, r, half, tq, has_acc);
      // staging writes -> async proxy, NOW: issued after the next tile's global
      // loads below, this fence would wait for them (an L2 round trip per tile)
      fence_proxy_async_smem();
      if (has_acc) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[acc_stage]);  // TMEM stage back to the MMAs
      }
      {  // next tile's operands, and the character of the tile after it
        const int c = c_next;
        c_next = char_of(tile + 2 * (int)gridDim.x);
        request(tile + (int)gridDim.x, c);
      }
      epilogue_store_tile(&map_out, args, stage_tile, col, m_tile, issuer, threadIdx.x - 128, pub, it);
      if (has_acc && ++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
    }
    if (issuer) tma_store_wait<0>();  // smem must outlive the store's reads
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<TC_TMEM_COLS>(tmem_base);
  }
}

// ------------------------------------------------- layers, CTA pairs (2-SM MMA)
// Same math, but two CTAs of one cluster (one TPC) work on a 256-name tile with
// tcgen05.mma.cta_group::2: each CTA stages its own 128 rows of A and only HALF
// of the weight block (96 of the 192 rows), the tensor cores of both SMs read
// both halves.  Shared-memory traffic per MMA flop drops by 30%, which is what
// bounds the single-CTA kernel (TMA fill + UMMA operand reads share 128 B/clk).
//   leader (cluster rank 0): issues every MMA; its `full` barrier collects the
//     TMA bytes of BOTH CTAs; its tmem_empty barrier collects the epilogue
//     arrivals of BOTH CTAs.
//   tcgen05.commit multicasts to the `empty` / tmem_full barriers of both CTAs.
constexpr int P_B_BYTES = 96 * 128;          // 12 KB: this CTA's half of B
constexpr int P_STAGE_BYTES = TC_A_BYTES + P_B_BYTES;
constexpr int P_STAGES = 6;
constexpr int P_SMEM_BYTES = 1024 + P_STAGES * P_STAGE_BYTES + 1024 + OUT_TILES * OUT_TILE_BYTES;

template <bool LAYER0>
__global__ void __launch_bounds__(L_THREADS, 1)
gru_layer_pair_kernel(const __grid_constant__ CUtensorMap map_x,
                      const __grid_constant__ CUtensorMap map_h,
                      const __grid_constant__ CUtensorMap map_wx,
                      const __grid_constant__ CUtensorMap map_wh,
                      const __grid_constant__ CUtensorMap map_out, const LayerArgs args) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + P_STAGES * P_STAGE_BYTES);
  uint64_t *full_bar = bars;                       // [STAGES] (leader's is used)
  uint64_t *empty_bar = bars + P_STAGES;           // [STAGES] (each CTA its own)
  uint64_t *tmem_full = bars + 2 * P_STAGES;       // [2]      (each CTA its own)
  uint64_t *tmem_empty = bars + 2 * P_STAGES + 2;  // [2]      (leader's is used)
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * P_STAGES + 4);
  const PubRing pub{bars + 2 * P_STAGES + 5, bars + 2 * P_STAGES + 7,
                    reinterpret_cast<int *>(bars + 2 * P_STAGES + 9)};

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;
  const int kb_total = args.kb_x + args.kb_h;

  if (warp == 0 && lane == 0) {
    if (args.kb_x) { tma_prefetch_desc(&map_x); tma_prefetch_desc(&map_wx); }
    if (args.kb_h) { tma_prefetch_desc(&map_h); tma_prefetch_desc(&map_wh); }
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < P_STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 2 * L_EPI_WARPS);  // epilogue warps of both CTAs
      mbar_init(&pub.full[i], L_EPI_WARPS);
      mbar_init(&pub.empty[i], 1);
    }
    fence_mbar_init();
  }
  cluster_sync_all();  // barrier inits visible pair-wide before any remote use
  if (warp == 2) tmem_alloc_2sm<TC_TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp >= 4) {  // n_h columns of both stages start at zero
    const uint32_t lane_base0 = (uint32_t)((warp & 3) * 32) << 16;
    const int half0 = (warp - 4) >> 2;
    uint32_t zero[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) zero[i] = 0u;
#pragma unroll
    for (int st = 0; st < 2; ++st)
#pragma unroll
      for (int c = 0; c < 2; ++c)
        tmem_st16(tmem_base + lane_base0 + st * TC_ACC_STRIDE + 192 + half0 * 32 + c * 16, zero);
    tmem_st_wait();
  }
  tc_fence_before();
  cluster_sync_all();  // both CTAs allocated + zeroed before the leader's MMAs
  tc_fence_after();
  pdl_launch_dependents();
  pdl_wait();  // see gru_layer_kernel
  const int n_rows = args.n_dev ? __ldcg(args.n_dev) : args.n;
  const int m2_tiles = (n_rows + 2 * kTileM - 1) / (2 * kTileM);    // 256-name tiles
  const int total_tiles = m2_tiles * args.n_blocks;

  if (warp == 0) {
    // ============================ TMA producer (both CTAs, own smem)
    Ring ring{0, 0};
    for (int tile = pair; tile < total_tiles; tile += num_pairs) {
      const int m2 = tile / args.n_blocks;
      const int n_blk = tile % args.n_blocks;
      const int m_tile = m2 * 2 + (int)rank;
      const int b_row = n_blk * 192 + (int)rank * 96;
      produce_part<true, P_STAGES, P_STAGE_BYTES>(&map_x, &map_wx, args.kb_x, args.x_full != 0,
                                                  args.wait_x, args.expect_x, m_tile, b_row, args,
                                                  smem, full_bar, empty_bar, leader, ring);
      produce_part<true, P_STAGES, P_STAGE_BYTES>(&map_h, &map_wh, args.kb_h, true, args.wait_h,
                                                  args.expect_h, m_tile, b_row, args, smem,
                                                  full_bar, empty_bar, leader, ring);
    }
  } else if (warp == 1) {
    // ========================================== MMA issuer (leader only)
    if (leader) {
      int stage = 0;
      uint32_t phase = 0;
      int acc_stage = 0;
      uint32_t acc_phase = 0;
      const uint32_t idesc = umma_idesc_bf16(2 * kTileM, 192);
      const uint64_t desc0 = umma_desc_sw128(smem_u32(smem));
      for (int tile = pair; tile < total_tiles && kb_total > 0; tile += num_pairs) {
        mbar_wait(&tmem_empty[acc_stage], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d = tmem_base + acc_stage * TC_ACC_STRIDE;
        B200_DASSERT((d & 0xffffu) + 256 <= (tmem_base & 0xffffu) + TC_TMEM_COLS && acc_stage < 2 && stage < P_STAGES,
                     "accumulator stage %d / smem stage %d out of range", acc_stage, stage);
        for (int kb = 0; kb < kb_total; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t da = desc0 + (uint64_t)((stage * P_STAGE_BYTES) >> 4);
            const uint64_t db = da + (uint64_t)(TC_A_BYTES >> 4);
            if (kb < args.kb_x) {
              umma_bf16_2sm(d, da, db, idesc, kb != 0 ? 1u : 0u);
              umma_bf16_2sm(d, da + 2, db + 2, idesc, 1u);
              umma_bf16_2sm(d, da + 4, db + 4, idesc, 1u);
              umma_bf16_2sm(d, da + 6, db + 6, idesc, 1u);
            } else {   // layer 0 (no x-part): the first MMA overwrites
              umma_bf16_2sm(d + 64, da, db, idesc, (args.kb_x == 0 && kb == 0) ? 0u : 1u);
              umma_bf16_2sm(d + 64, da + 2, db + 2, idesc, 1u);
              umma_bf16_2sm(d + 64, da + 4, db + 4, idesc, 1u);
              umma_bf16_2sm(d + 64, da + 6, db + 6, idesc, 1u);
            }
            umma_commit_2sm(&empty_bar[stage], 0b11);  // frees the slot in both CTAs
            if (kb == kb_total - 1) umma_commit_2sm(&tmem_full[acc_stage], 0b11);  // wakes both epilogues
          }
          __syncwarp();
          if (++stage == P_STAGES) { stage = 0; phase ^= 1; }
        }
        if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp == 3) {
    // ======================================= publisher (tensor parallel only)
    if (args.tp_world > 1 && lane == 0) {
      int it = 0;
      for (int tile = pair; tile < total_tiles; tile += num_pairs, ++it)
        publish_tile(args, pub, it);
    }
  } else if (warp >= 4) {
    // ======================================== epilogue (both CTAs, own rows)
    const int quarter = warp & 3;
    const int half = (warp - 4) >> 2;
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    int acc_stage = 0;
    uint32_t acc_phase = 0;
    uint8_t *stage_out = smem + P_STAGES * P_STAGE_BYTES + 1024;  // after the barrier block
    const bool issuer = (warp == 4 && lane == 0);
    // next tile's h_prev / table rows are requested one tile ahead, its character
    // two tiles ahead (see gru_layer_kernel)
    const int r = quarter * 32 + lane;
    const bool has_h = args.kb_h > 0;
    uint4 hq[4];
    T0Rows tq;
    int c_next = 0;
    auto char_of = [&](int tile_q) -> int {
      if (!LAYER0 || tile_q >= total_tiles) return 0;
      const int row_q = ((tile_q / args.n_blocks) * 2 + (int)rank) * kTileM + r;
      return row_q < n_rows ? args.chars[row_q] : 0;
    };
    auto request = [&](int tile_q, int c) {
      if (tile_q >= total_tiles) return;
      const int mq = (tile_q / args.n_blocks) * 2 + (int)rank, nq = tile_q % args.n_blocks;
      if (has_h) {
        B200_DASSERT(h_offset(mq * kTileM + r, args.col0 + nq * 64 + half * 32, args.pw_log2, args.bpad) + 32 <=
                         (size_t)args.bpad * args.h_full,
                     "h_prev offset beyond hidden buffer (tile %d)", tile_q);
        const uint4 *src = reinterpret_cast<const uint4 *>(
            args.h_prev + h_offset(mq * kTileM + r, args.col0 + nq * 64 + half * 32, args.pw_log2, args.bpad));
#pragma unroll
        for (int i = 0; i < 4; ++i) hq[i] = src[i];
      }
      B200_DASSERT(!LAYER0 || (c >= 0 && c < args.num_char && nq < args.n_blocks),
                   "layer-0 table row out of range: char %d of %d, unit block %d of %d", c, args.num_char, nq,
                   args.n_blocks);
      if (LAYER0) load_t0_rows(args.t0q + (((size_t)nq * args.num_char + c) * 2 + half) * 96, tq);
    };
#pragma unroll
    for (int i = 0; i < 4; ++i) hq[i] = make_uint4(0, 0, 0, 0);
    request(pair, char_of(pair));
    c_next = char_of(pair + num_pairs);
    int it = 0;
    for (int tile = pair; tile < total_tiles; tile += num_pairs, ++it) {
      const int m2 = tile / args.n_blocks;
      const int n_blk = tile % args.n_blocks;
      const int m_tile = m2 * 2 + (int)rank;
      const int col = args.col0 + n_blk * 64;
      const float4 *bias4 =
          reinterpret_cast<const float4 *>(args.bias + (size_t)n_blk * 4 * 64 + half * 32);
      uint8_t *stage_tile = stage_out + (it % OUT_TILES) * OUT_TILE_BYTES;

      const bool has_acc = kb_total > 0;
      if (has_acc) {
        mbar_wait_relaxed(&tmem_full[acc_stage], acc_phase, kEpiSleepNs);
        tc_fence_after();
      }
      const uint32_t t = tmem_base + lane_base + acc_stage * TC_ACC_STRIDE + half * 32;
      gate_epilogue<LAYER0>(t, hq, bias4, stage_tile, r, half, tq, has_acc);
      // staging writes -> async proxy, NOW: issued after the next tile's global
      // loads below, this fence would wait for them (an L2 round trip per tile)
      fence_proxy_async_smem();
      if (has_acc) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(mapa_u32(&tmem_empty[acc_stage], 0));
      }
      {  // next tile's operands, and the character of the tile after it
        const int c = c_next;
        c_next = char_of(tile + 2 * num_pairs);
        request(tile + num_pairs, c);
      }
      epilogue_store_tile(&map_out, args, stage_tile, col, m_tile, issuer, threadIdx.x - 128, pub, it);
      if (has_acc && ++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
    }
    if (issuer) tma_store_wait<0>();  // smem must outlive the store's reads
  }

  tc_fence_before();
  cluster_sync_all();  // peer smem / barriers stay valid until both are done
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2sm<TC_TMEM_COLS>(tmem_base);
  }
}

// ---------------------------------------------------------------------- FC
constexpr int F_EPI_WARPS = 8;               // two threads per name: each walks half of the columns
constexpr int F_THREADS = (4 + F_EPI_WARPS) * 32;
constexpr int F_B_BYTES_MAX = 256 * 128;     // 32 KB
constexpr int F_STAGE_BYTES = TC_A_BYTES + F_B_BYTES_MAX;
constexpr int F_STAGES = 4;
// [align slack][stages][512 B barriers | 2 KB epilogue exchange]
constexpr int F_SMEM_BYTES = 1024 + F_STAGES * F_STAGE_BYTES + 4096;

struct FcArgs {
  int n, m_tiles, kb, H, C, T, step, eos;
  const int *n_dev;         // rows in use after compaction (null: n)
  const int32_t *slot_row;  // original name of every row (null: identity); rfloats / output are indexed by it
  int *gone;                // optional: += names that hit EOS in this step (lets the compaction planner exit early)
  const float *b_fc;
  float *logits;            // optional [bpad, C] tap (tests / debugging)
  const float *rfloats;     // [n, T]   (may be a peer address)
  char *output;             // [n, T+1] (may be a peer address)
  int32_t *chars;
  uint8_t *alive;
  // tensor parallel: the top layer's hidden buffer arrives in per-rank chunks
  int tp_world, tp_rank, kb_rank, flag_stride;
  uint32_t expect_h;
  const uint32_t *wait_h;
  int pw_log2, bpad;        // panel layout of the hidden buffer
};

__global__ void __launch_bounds__(F_THREADS, 1)
fc_sample_kernel(const __grid_constant__ CUtensorMap map_h,
                 const __grid_constant__ CUtensorMap map_w, const FcArgs args) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + F_STAGES * F_STAGE_BYTES);
  uint64_t *full_bar = bars, *empty_bar = bars + F_STAGES;
  uint64_t *tmem_full = bars + 2 * F_STAGES, *tmem_empty = bars + 2 * F_STAGES + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * F_STAGES + 4);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = args.C;
  const uint32_t b_bytes = (uint32_t)C * 128u;

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&map_h); tma_prefetch_desc(&map_w); }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < F_STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], F_EPI_WARPS); }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<TC_TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();
  pdl_wait();  // prologue above may overlap the previous kernel's tail
  const int n_rows = args.n_dev ? __ldcg(args.n_dev) : args.n;
  const int m_tiles = (n_rows + kTileM - 1) / kTileM;

  if (warp == 0) {
    // TMA producer: warp-converged, one elected lane issues (see produce_part)
    int stage = 0;
    uint32_t phase = 0;
    const bool tp = args.tp_world > 1;
    const int nchunks = tp ? args.tp_world : 1;
    const int chunk_len = tp ? args.kb_rank : args.kb;
    const int pw = 1 << args.pw_log2;
    for (int tile = blockIdx.x; tile < m_tiles; tile += gridDim.x)
      for (int c = 0; c < nchunks; ++c) {
        int src = args.tp_rank + c;
        if (src >= args.tp_world) src -= args.tp_world;
        bool fence = false;
        if (tp && src != args.tp_rank) {
          tp_wait_counter(args.wait_h + src * args.flag_stride + tile, args.expect_h);
          fence = true;
        }
        int kw = tp ? src * args.kb_rank * TC_BLOCK_K : 0;
        int ci = kw & (pw - 1);
        int co = tile * kTileM + (kw >> args.pw_log2) * args.bpad;
        for (int jj = 0; jj < chunk_len; ++jj) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (elect_one()) {
            uint8_t *sa = smem + stage * F_STAGE_BYTES;
            if (fence) fence_proxy_async();
            mbar_expect_tx(&full_bar[stage], TC_A_BYTES + b_bytes);
            tma_load_2d(sa, &map_h, &full_bar[stage], ci, co);
            tma_load_2d(sa + TC_A_BYTES, &map_w, &full_bar[stage], kw, 0);
          }
          __syncwarp();
          fence = false;
          if (++stage == F_STAGES) { stage = 0; phase ^= 1; }
          kw += TC_BLOCK_K;
          ci += TC_BLOCK_K;
          if (ci == pw) { ci = 0; co += args.bpad; }
        }
      }
  } else if (warp == 1) {
    // MMA issuer: warp-converged, elected lane issues four UMMAs per k-block
    int stage = 0, acc_stage = 0;
    uint32_t phase = 0, acc_phase = 0;
    const uint32_t idesc = umma_idesc_bf16(kTileM, (uint32_t)C);
    const uint64_t desc0 = umma_desc_sw128(smem_u32(smem));
    for (int tile = blockIdx.x; tile < m_tiles; tile += gridDim.x) {
      mbar_wait(&tmem_empty[acc_stage], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t d = tmem_base + acc_stage * TC_ACC_STRIDE;
      for (int kb = 0; kb < args.kb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t da = desc0 + (uint64_t)((stage * F_STAGE_BYTES) >> 4);
          const uint64_t db = da + (uint64_t)(TC_A_BYTES >> 4);
          umma_bf16(d, da, db, idesc, kb != 0 ? 1u : 0u);
          umma_bf16(d, da + 2, db + 2, idesc, 1u);
          umma_bf16(d, da + 4, db + 4, idesc, 1u);
          umma_bf16(d, da + 6, db + 6, idesc, 1u);
          umma_commit(&empty_bar[stage]);
          if (kb == args.kb - 1) umma_commit(&tmem_full[acc_stage]);
        }
        __syncwarp();
        if (++stage == F_STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4) {
    // Two threads per name (same TMEM lane): half 0 (warps 4-7) owns columns
    // [0, split), half 1 (warps 8-11) owns [split, C).  Pass 1 turns the logits
    // into exponentials (kept in TMEM in place) and keeps one sum per 16 columns
    // in registers; the two halves swap their totals through shared memory.
    // With both totals every thread knows where its half starts in the CDF and
    // therefore which half — and which 16-column block of it — contains the
    // sampled point; only that thread walks, 16 columns at most, and does the
    // name's bookkeeping.
    const int quarter = warp & 3;
    const int half = (warp - 4) >> 2;
    const int rl = quarter * 32 + lane;           // row within the tile
    const int split = ((C / 16 + 1) / 2) * 16;
    const int cbeg = half == 0 ? 0 : split, cend = half == 0 ? split : C;
    float *part_sum = reinterpret_cast<float *>(smem + F_STAGES * F_STAGE_BYTES + 512);  // [2 parity][2][128]
    int acc_stage = 0;
    uint32_t acc_phase = 0;
    const float kLog2e = 1.4426950408889634f;
    int it = 0;
    for (int tile = blockIdx.x; tile < m_tiles; tile += gridDim.x, ++it) {
      const int row = tile * kTileM + rl;
      const bool valid = row < n_rows;
      const bool live = valid && args.alive[row] != 0;
      // the name this row holds (rows are compacted as names finish)
      const int name = (valid && args.slot_row) ? args.slot_row[row] : row;
      // this step's random float: fetched (possibly over NVLink from the root
      // rank's buffer — the MPI_Scatter of ref:636) before the accumulator wait
      const float r = live ? args.rfloats[(size_t)name * args.T + args.step] : 2.0f;
      mbar_wait_relaxed(&tmem_full[acc_stage], acc_phase, kEpiSleepNs);
      tc_fence_after();
      const uint32_t t = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc_stage * TC_ACC_STRIDE;
      // pass 1: e_i = exp(f_i + b_i), kept in TMEM in place; one sum per 16 columns
      float cs[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        cs[k] = 0.0f;
        const int c0 = cbeg + 16 * k;
        if (c0 < cend) {
          uint32_t v[16];
          tmem_ld16(t + c0, v);
          tmem_ld_wait();
          float e[16];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 b = __ldg(reinterpret_cast<const float4 *>(args.b_fc + c0) + q);
            const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float x = __uint_as_float(v[q * 4 + j]) + bb[j];
              if (args.logits && valid) args.logits[(size_t)row * C + c0 + q * 4 + j] = live ? x : 0.0f;
              e[q * 4 + j] = ex2_approx(x * kLog2e);
              v[q * 4 + j] = __float_as_uint(e[q * 4 + j]);
            }
          }
          tmem_st16(t + c0, v);
          cs[k] = (((e[0] + e[1]) + (e[2] + e[3])) + ((e[4] + e[5]) + (e[6] + e[7]))) +
                  (((e[8] + e[9]) + (e[10] + e[11])) + ((e[12] + e[13]) + (e[14] + e[15])));
        }
      }
      tmem_st_wait();
      float my_sum = 0.0f;
#pragma unroll
      for (int k = 0; k < 8; ++k) my_sum += cs[k];
      float *ps = part_sum + (it & 1) * 2 * kTileM;   // double-buffered: one barrier per tile is enough
      ps[half * kTileM + rl] = my_sum;
      named_bar_sync(EPI_BAR_ID, F_EPI_WARPS * 32);
      const float sum_lo = ps[rl], tot = sum_lo + ps[kTileM + rl];
      // first i with sum_{j<=i} e_j > r * sum (ref:322-333 on unnormalised mass,
      // no max subtraction as in ref:302-313); if the mass never exceeds it the
      // last character wins (ref:332)
      const float rs = r * tot;
      const bool fallback = !(tot > rs);
      const int owner = fallback ? 1 : (sum_lo > rs ? 0 : 1);
      const bool mine = live && owner == half;
      int sel = (mine && fallback) ? C - 1 : -1;
      float p = half == 0 ? 0.0f : sum_lo;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int c0 = cbeg + 16 * k;
        if (c0 < cend) {
          const bool need = mine && sel < 0 && (p + cs[k] > rs);
          if (__any_sync(0xffffffffu, need)) {      // tcgen05.ld is warp-collective
            uint32_t v[16];
            tmem_ld16(t + c0, v);
            tmem_ld_wait();
            if (need) {
              float q = p;
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                q += __uint_as_float(v[i]);
                if (sel < 0 && q > rs) sel = c0 + i;
              }
              if (sel < 0) sel = c0 + 15;           // block sum crossed, running sum did not (rounding)
            }
          }
          p += cs[k];
        }
      }
      if (mine && sel < 0) sel = cend - 1;
      // accumulator drained: hand the TMEM stage back before the bookkeeping
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc_stage]);
      if (valid) {
        char *orow = args.output + (size_t)name * (args.T + 1);
        if (args.step == 0 && half == 0)
          for (int l = 1; l <= args.T; ++l) orow[l] = 0;  // NUL pad, ref:643
        B200_DASSERT(!mine || (sel >= 0 && sel < C && name >= 0 && name < args.n && args.step < args.T),
                     "sampled character %d of %d, name %d of %d, step %d of %d", sel, C, name, args.n, args.step, args.T);
        if (mine) {
          orow[args.step] = (char)sel;  // EOS id is written too, ref:877
          args.chars[row] = sel;  // next step's layer-0 input: the row of the folded table it picks
          if (sel == args.eos) args.alive[row] = 0;  // ref:881
        }
      }
      if (args.gone) {
        const unsigned ended = __ballot_sync(0xffffffffu, mine && sel == args.eos);
        if (lane == 0 && ended) atomicAdd(args.gone, __popc(ended));
      }
      if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<TC_TMEM_COLS>(tmem_base);
  }
}

// ------------------------------------------------------------- init / packing
// A rank packs only the hidden units it owns: [unit0, unit0 + Hs) of H (Hs == H
// without tensor parallelism).
// fp32 [3][H][K] (gate-stacked r,z,n as in the file) -> bf16 [Hs/64][3][64][K]
// with the gate order of the packed block given by (g0,g1,g2).
__global__ void pack_gate_weights_kernel(const float *__restrict__ src,
                                         __nv_bfloat16 *__restrict__ dst, int H,
                                         int Hs, int unit0, int K, int ld, int g0, int g1, int g2) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t total = (size_t)3 * Hs * K;
  if (idx >= total) return;
  int k = (int)(idx % K);
  size_t prow = idx / K;  // packed row
  int jb = (int)(prow / 192), rem = (int)(prow % 192);
  int slot = rem / 64, u = rem % 64;
  int g = slot == 0 ? g0 : (slot == 1 ? g1 : g2);
  dst[prow * ld + k] = __float2bfloat16_rn(src[((size_t)g * H + unit0 + jb * 64 + u) * K + k]);
}

__global__ void to_bf16_kernel(const float *__restrict__ src,
                               __nv_bfloat16 *__restrict__ dst, size_t n, int cols, int ld) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < n) dst[(idx / cols) * ld + idx % cols] = __float2bfloat16_rn(src[idx]);
}

// Layer-0 input table for the epilogue gather:
//   t0q[jb][c][half][chunk][slot: n,r,z][16] = W_i{g}0 . emb[c] + b_i{g}0 (+ b_h{g}0 for r, z)
// for hidden unit unit0 + jb*64 + half*32 + chunk*16 + u, so that the 192 bytes one
// epilogue thread (name, 32 units) needs are contiguous.
__global__ void pack_t0_kernel(const float *__restrict__ t0_raw,
                               const float *__restrict__ b_ih,
                               const float *__restrict__ b_hh,
                               __nv_bfloat16 *__restrict__ t0q, int C, int H,
                               int Hs, int unit0) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)3 * Hs * C) return;
  const int u = (int)(idx % 16);
  const int slot = (int)((idx / 16) % 3);
  const int chunk = (int)((idx / 48) % 2);
  const int half = (int)((idx / 96) % 2);
  const int c = (int)((idx / 192) % C);
  const int jb = (int)(idx / ((size_t)192 * C));
  const int g = slot == 0 ? 2 : slot - 1;  // slots (n, r, z) <- gates (2, 0, 1)
  const int i = unit0 + jb * 64 + half * 32 + chunk * 16 + u;
  float v = t0_raw[(size_t)c * 3 * H + g * H + i] + b_ih[g * H + i];
  if (g < 2) v += b_hh[g * H + i];
  t0q[idx] = __float2bfloat16_rn(v);
}

// bias[jb][q][u]: q=0 b_in, q=1 b_ir+b_hr, q=2 b_iz+b_hz, q=3 b_hn.
// Layer 0 folds q=0..2 into the T0 table, so they are zero there.
__global__ void pack_bias_kernel(const float *__restrict__ b_ih,
                                 const float *__restrict__ b_hh,
                                 float *__restrict__ out, int H, int Hs, int unit0,
                                 int folded) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= 4 * Hs) return;
  int jb = idx / 256, q = (idx % 256) / 64, u = idx % 64;
  int i = unit0 + jb * 64 + u;
  float v;
  if (q == 0) v = folded ? 0.0f : b_ih[2 * H + i];
  else if (q == 1) v = folded ? 0.0f : b_ih[i] + b_hh[i];
  else if (q == 2) v = folded ? 0.0f : b_ih[H + i] + b_hh[H + i];
  else v = b_hh[2 * H + i];
  out[idx] = v;
}

// Per-call reset: chars = SOS (ref:652), alive flags.
__global__ void tc_init_kernel(int32_t *chars, uint8_t *alive, int32_t *slot_row, int *state, int n, int rows,
                               int sos) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < rows) {
    chars[idx] = sos;
    alive[idx] = idx < n;
    slot_row[idx] = idx;   // row i holds name i until the first compaction
  }
  if (idx == 0) { state[0] = n; state[1] = n; state[2] = 0; state[3] = 0; }
}

__global__ void tc_signal_done_kernel(uint32_t *flag, uint32_t value) {
  __threadfence_system();
  st_release_sys(flag, value);
}

// Tensor-parallel call fence.  wait: every peer has finished generate number
// `epoch` (its kernels no longer read or write our buffers).  signal: tell
// every peer that we have.
__global__ void tp_wait_epoch_kernel(const uint32_t *sync_local, int world, uint32_t epoch) {
  if ((int)threadIdx.x < world) {
#ifdef B200GRU_DEBUG
    SpinGuard g;
#endif
    uint32_t v;
    while ((int32_t)((v = ld_acquire_sys(sync_local + threadIdx.x)) - epoch) < 0) {
      __nanosleep(200);
#ifdef B200GRU_DEBUG
      g.tick("tensor-parallel call fence", sync_local + threadIdx.x, epoch, v);
#endif
    }
  }
}
struct TpPeers { uint32_t *p[8]; };
__global__ void tp_signal_epoch_kernel(TpPeers sync_slot, int world, uint32_t epoch) {
  if ((int)threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(sync_slot.p[threadIdx.x], epoch);
  }
}

// ------------------------------------------------------------- compaction
// The reference stops a name at EOS (ref:881-882).  Batched, a finished name
// would keep occupying a GEMM row for the rest of the call, so a few times per
// call the rows still running are packed to the front: three small launches
// that do nothing unless at least one 256-name tile can be dropped.
//   state[0] = rows in use (read by every later kernel), [1] = rows after the
//   compaction being planned, [2] = whether it is carried out.
__global__ void __launch_bounds__(1024, 1)
compact_plan_kernel(const uint8_t *__restrict__ alive, int32_t *__restrict__ new_idx, int *state) {
  __shared__ int warp_tot[32];
  __shared__ int running;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int n = state[0];
  {
    // Cheap exit: the FC kernels count the names that ended since the last
    // compaction (state[3]).  Packing moves every surviving row twice, so it is
    // only worth it when at least 1/16 of the 256-name tiles (and at least one)
    // can be dropped.
    const int tiles = (n + 2 * kTileM - 1) / (2 * kTileM);
    const int tiles_new = (n - state[3] + 2 * kTileM - 1) / (2 * kTileM);
    const int need = tiles / 16 > 1 ? tiles / 16 : 1;
    if (tiles - tiles_new < need) {
      if (tid == 0) state[2] = 0;
      return;
    }
  }
  if (tid == 0) running = 0;
  __syncthreads();
  for (int base = 0; base < n; base += 1024) {
    const int j = base + tid;
    const bool a = j < n && alive[j] != 0;
    const unsigned bal = __ballot_sync(0xffffffffu, a);
    const int pre = __popc(bal & ((1u << lane) - 1u));
    if (lane == 0) warp_tot[w] = __popc(bal);
    __syncthreads();
    int off = 0, tot = 0;
    for (int i = 0; i < 32; ++i) {   // 32 warp totals: trivial serial scan per thread
      if (i < w) off += warp_tot[i];
      tot += warp_tot[i];
    }
    if (j < n) new_idx[j] = a ? running + off + pre : -1;
    __syncthreads();
    if (tid == 0) running += tot;
    __syncthreads();
  }
  if (tid == 0) {
    const int n_new = running;
    state[1] = n_new;
    state[2] = 1;
  }
}

struct CompactBufs {
  __nv_bfloat16 *cur[B200GRU_MAX_LAYERS];   // hidden state after this step
  __nv_bfloat16 *tmp[B200GRU_MAX_LAYERS];   // the other ping-pong buffer (free until the next step writes it)
  int layers, H, pw_log2, bpad;
};

// one warp per row: move the row's hidden state of every layer to its new place (in tmp)
__global__ void compact_move_kernel(const int *state, const int32_t *__restrict__ new_idx, CompactBufs b,
                                    const int32_t *__restrict__ chars, int32_t *__restrict__ chars2,
                                    const int32_t *__restrict__ slot_row, int32_t *__restrict__ slot2) {
  if (!state[2]) return;
  const int n = state[0];
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < n; row += warps) {
    const int d = new_idx[row];
    if (d < 0) continue;
    for (int l = 0; l < b.layers; ++l)
      for (int c = lane * 8; c < b.H; c += 256)
        *reinterpret_cast<uint4 *>(b.tmp[l] + h_offset(d, c, b.pw_log2, b.bpad)) =
            *reinterpret_cast<const uint4 *>(b.cur[l] + h_offset(row, c, b.pw_log2, b.bpad));
    if (lane == 0) {
      chars2[d] = chars[row];
      slot2[d] = slot_row[row];
    }
  }
}

// copy the packed rows back and publish the new row count
__global__ void compact_commit_kernel(int *state, CompactBufs b, int32_t *__restrict__ chars,
                                      const int32_t *__restrict__ chars2, int32_t *__restrict__ slot_row,
                                      const int32_t *__restrict__ slot2, uint8_t *__restrict__ alive) {
  if (!state[2]) return;
  const int n_new = state[1];
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < n_new; row += warps) {
    for (int l = 0; l < b.layers; ++l)
      for (int c = lane * 8; c < b.H; c += 256) {
        const size_t off = h_offset(row, c, b.pw_log2, b.bpad);
        *reinterpret_cast<uint4 *>(b.cur[l] + off) = *reinterpret_cast<const uint4 *>(b.tmp[l] + off);
      }
    if (lane == 0) {
      chars[row] = chars2[row];
      slot_row[row] = slot2[row];
      alive[row] = 1;
    }
  }
  // later kernels read state[0]; nothing in THIS kernel does
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    state[0] = n_new;
    state[3] = 0;   // every name that had ended is gone from the batch now
  }
}

template <typename T>
static int tc_alloc(b200gru_model *m, T **out, size_t count) {
  return model_alloc(m, out, count);
}

// Layout of the peer-visible region a tensor-parallel rank exposes:
//   h[l][pp]            L*2 buffers of [bpad, H] bf16 (full width)
//   counters[l][pp]     [world][mt_max] uint32 per buffer
//   sync                [world] uint32 (call fence)
static size_t tp_h_bytes(const b200gru_model *m) {
  return (size_t)m->bpad * m->cfg.hidden * sizeof(__nv_bfloat16);
}
static size_t tp_counter_bytes(const b200gru_model *m) {
  return (size_t)m->tp_world * (m->bpad / kTileM) * sizeof(uint32_t);
}
static size_t tp_counter_stride(const b200gru_model *m) {  // bytes between counter blocks
  return (tp_counter_bytes(m) + 255) / 256 * 256;
}
size_t tp_shared_bytes(const b200gru_model *m) {
  const int L = m->cfg.num_layers;
  size_t b = (size_t)L * 2 * tp_h_bytes(m);
  b += (size_t)L * 2 * tp_counter_stride(m);
  b += 256;
  return b;
}

static int make_h_maps(b200gru_model *m) {
  for (int l = 0; l < m->cfg.num_layers; ++l)
    for (int pp = 0; pp < 2; ++pp)
      if (int rc = make_tmap_bf16_2d(&m->tc[l].map_h[pp], m->tc[l].h[pp],
                                     (uint64_t)m->bpad * (m->cfg.hidden >> m->pw_log2),
                                     (uint64_t)1 << m->pw_log2, kTileM, 64))
        return rc;
  return 0;
}

int tp_bind(b200gru_model *m, void *local_base, void *const *peer_bases, void *mc_base) {
  B200_REQUIRE(m->tp_world > 1, "model was not created for tensor parallelism");
  const int L = m->cfg.num_layers;
  const size_t hb = tp_h_bytes(m);
  const size_t cb = tp_counter_stride(m);
  const size_t off_cnt = (size_t)L * 2 * hb;
  const size_t off_sync = off_cnt + (size_t)L * 2 * cb;
  for (int r = 0; r < m->tp_world; ++r) {
    uint8_t *base = static_cast<uint8_t *>(r == m->tp_rank ? local_base : peer_bases[r]);
    B200_REQUIRE(base, "peer base %d is null", r);
    for (int l = 0; l < L; ++l)
      for (int pp = 0; pp < 2; ++pp) {
        m->tp_h[r][l][pp] = reinterpret_cast<__nv_bfloat16 *>(base + (size_t)(l * 2 + pp) * hb);
        m->tp_cnt[r][l][pp] = reinterpret_cast<uint32_t *>(base + off_cnt + (size_t)(l * 2 + pp) * cb);
      }
    m->tp_sync[r] = reinterpret_cast<uint32_t *>(base + off_sync);
  }
  m->tp_mc_base = static_cast<uint8_t *>(mc_base);
  if (mc_base)
    for (int l = 0; l < L; ++l)
      for (int pp = 0; pp < 2; ++pp) {
        m->tp_h_mc[l][pp] = reinterpret_cast<__nv_bfloat16 *>(m->tp_mc_base + (size_t)(l * 2 + pp) * hb);
        m->tp_cnt_mc[l][pp] = reinterpret_cast<uint32_t *>(m->tp_mc_base + off_cnt + (size_t)(l * 2 + pp) * cb);
      }
  for (int l = 0; l < L; ++l)
    for (int pp = 0; pp < 2; ++pp) m->tc[l].h[pp] = m->tp_h[m->tp_rank][l][pp];
  if (int rc = make_h_maps(m)) return rc;
  m->tp_bound = true;
  return 0;
}

int tc_prepare(b200gru_model *m) {
  const b200gru_config &c = m->cfg;
  const int H = c.hidden, C = c.num_char, L = c.num_layers;
  const int Hs = H / m->tp_world, unit0 = m->tp_rank * Hs;
  LayoutIndex ix(L);
  const unsigned TPB = 256;
  auto blocks = [&](size_t n) { return (unsigned)((n + TPB - 1) / TPB); };
  const int ld = m->ldh;
  {
    // panel width of the hidden buffers: row-major unless tensor parallel, where
    // peers write 128-byte row segments over NVLink (B200GRU_PANEL overrides)
    int pw = m->tp_world > 1 ? 64 : H;
    if (const char *e = getenv("B200GRU_PANEL")) pw = atoi(e);
    if (pw < 64 || pw > H || (pw & (pw - 1)) || H % pw) pw = (H & (H - 1)) ? 64 : H;
    m->pw_log2 = 0;
    while ((1 << m->pw_log2) < pw) ++m->pw_log2;
  }

  // layer-0 input: the folded table, gathered by the epilogue
  if (int rc = tc_alloc(m, &m->t0q, (size_t)3 * Hs * C)) return rc;
  pack_t0_kernel<<<blocks((size_t)3 * Hs * C), TPB>>>(
      m->t0_raw, m->P(ix.b_i(0, 0)), m->P(ix.b_h(0, 0)), m->t0q, C, H, Hs, unit0);

  for (int l = 0; l < L; ++l) {
    TcLayer &t = m->tc[l];
    const size_t wn = (size_t)3 * Hs * ld;   // padded rows
    const size_t we = (size_t)3 * Hs * H;    // elements to convert
    if (l > 0) {
      if (int rc = tc_alloc(m, &t.w_ih, wn)) return rc;
      // x-part column order [n_i | r | z]
      B200_CUDA(cudaMemset(t.w_ih, 0, wn * sizeof(__nv_bfloat16)));
      pack_gate_weights_kernel<<<blocks(we), TPB>>>(m->P(ix.w_i(l, 0)), t.w_ih, H, Hs, unit0, H, ld,
                                                    2, 0, 1);
      if (int rc = make_tmap_bf16_2d(&t.map_w_ih, t.w_ih, 3 * (uint64_t)Hs, H, 192, 64, ld)) return rc;
      if (int rc = make_tmap_bf16_2d(&t.map_w_ih_p, t.w_ih, 3 * (uint64_t)Hs, H, 96, 64, ld)) return rc;
    }
    if (int rc = tc_alloc(m, &t.w_hh, wn)) return rc;
    // h-part column order [r | z | n_h]
    B200_CUDA(cudaMemset(t.w_hh, 0, wn * sizeof(__nv_bfloat16)));
    pack_gate_weights_kernel<<<blocks(we), TPB>>>(m->P(ix.w_h(l, 0)), t.w_hh, H, Hs, unit0, H, ld,
                                                  0, 1, 2);
    if (int rc = make_tmap_bf16_2d(&t.map_w_hh, t.w_hh, 3 * (uint64_t)Hs, H, 192, 64, ld)) return rc;
    if (int rc = make_tmap_bf16_2d(&t.map_w_hh_p, t.w_hh, 3 * (uint64_t)Hs, H, 96, 64, ld)) return rc;
    if (int rc = tc_alloc(m, &t.bias, (size_t)4 * Hs)) return rc;
    pack_bias_kernel<<<blocks((size_t)4 * Hs), TPB>>>(m->P(ix.b_i(l, 0)), m->P(ix.b_h(l, 0)),
                                                      t.bias, H, Hs, unit0, l == 0 ? 1 : 0);
    if (m->tp_world == 1 || m->tp_fake)
      for (int pp = 0; pp < 2; ++pp) {
        if (int rc = tc_alloc(m, &t.h[pp], (size_t)m->bpad * H)) return rc;
        B200_CUDA(cudaMemset(t.h[pp], 0, sizeof(__nv_bfloat16) * (size_t)m->bpad * H));
      }
  }
  if (m->tp_world == 1 || m->tp_fake)
    if (int rc = make_h_maps(m)) return rc;
  if (int rc = tc_alloc(m, &m->slot_row, (size_t)m->bpad)) return rc;
  if (int rc = tc_alloc(m, &m->slot2, (size_t)m->bpad)) return rc;
  if (int rc = tc_alloc(m, &m->chars2, (size_t)m->bpad)) return rc;
  if (int rc = tc_alloc(m, &m->new_idx, (size_t)m->bpad)) return rc;
  if (int rc = tc_alloc(m, &m->cp_state, 4)) return rc;
  if (int rc = tc_alloc(m, &m->w_fc, (size_t)C * ld)) return rc;
  B200_CUDA(cudaMemset(m->w_fc, 0, sizeof(__nv_bfloat16) * (size_t)C * ld));
  to_bf16_kernel<<<blocks((size_t)C * H), TPB>>>(m->P(ix.w_fc()), m->w_fc, (size_t)C * H, H, ld);
  if (int rc = make_tmap_bf16_2d(&m->map_w_fc, m->w_fc, (uint64_t)C, H, C, 64, ld)) return rc;
  B200_CUDA(cudaGetLastError());

  B200_CUDA(cudaFuncSetAttribute(gru_layer_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 L_SMEM_BYTES));
  B200_CUDA(cudaFuncSetAttribute(gru_layer_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 L_SMEM_BYTES));
  B200_CUDA(cudaFuncSetAttribute(gru_layer_pair_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 P_SMEM_BYTES));
  B200_CUDA(cudaFuncSetAttribute(gru_layer_pair_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 P_SMEM_BYTES));
  {
    const char *e = getenv("B200GRU_CTA_PAIRS");
    m->use_pairs = !(e && e[0] == '0');
    e = getenv("B200GRU_PDL");
    m->use_pdl = !(e && e[0] == '0');
  }
  B200_CUDA(cudaFuncSetAttribute(fc_sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 F_SMEM_BYTES));
  return 0;
}

static int num_sms(int device) {
  static int cached[64] = {0};
  if (device < 64 && cached[device]) return cached[device];
  int n = 148;
  cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
  if (device < 64) cached[device] = n;
  return n;
}

// part: 0 = the whole call; 1 = only what runs before the random floats are
// needed (reset + the first character's layer kernels); 2 = the rest.  The
// engine captures 1 and 2 as separate graphs when the caller supplies a
// `ready_event`, and waits for the event between them.
int tc_generate(b200gru_model *m, int n, const b200gru_io *io, cudaStream_t s, int part) {
  const bool head = part != 2, tail = part != 1;
  const b200gru_config &c = m->cfg;
  const int H = c.hidden, C = c.num_char, T = c.max_len, L = c.num_layers;
  // tp_fake: profile ONE rank's share of a tensor-parallel model on a single GPU
  // (sliced shapes, no peers, no waits; the numbers it samples are meaningless)
  const int W = m->tp_fake ? 1 : m->tp_world, R = m->tp_rank;
  const int Hs = H / m->tp_world;
  const int sms = num_sms(m->device);
  int launches = 0;
  const int m_tiles = (n + kTileM - 1) / kTileM;
  const int rows = m_tiles * kTileM;
  const int mt_max = m->bpad / kTileM;
  const uint32_t per_round = (uint32_t)(Hs / kUnitBlock);  // counter bumps per (source, m-tile)
  // finished names are packed away after steps 1, 3, 5, ... (not with tensor
  // parallelism, whose counters are per tile, nor with the logits tap, whose
  // rows are per name)
  // ... nor for small calls, where the twelve extra launches cost more than dropped tiles can save
  const bool compact = m->compact && W == 1 && m->tp_world == 1 && !m->debug_logits && n >= m->compact_min_n;
  if (head) m->trace_begin(s);
  if (head && W > 1) {
    if (!m->tp_bound) return fail(B200GRU_ERR_STATE, "tensor-parallel model: call b200gru_tp_bind first");
    // nobody may touch our buffers for this call until every peer is done with the last one
    tp_wait_epoch_kernel<<<1, 32, 0, s>>>(m->tp_sync[R], W, m->tp_epoch);
    ++launches;
    m->trace_mark("tp_wait_epoch", s);
    // Counters are per call: they were zeroed at the end of the previous call
    // (see below), so the expected values only depend on this call's own
    // launches — not on how many m-tiles earlier calls happened to touch.
    for (int l = 0; l < L; ++l) m->tp_rounds[l][0] = m->tp_rounds[l][1] = 0;
  }
  if (head) {
    tc_init_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(m->chars, m->alive, m->slot_row, m->cp_state,
                                                                  n, rows, c.sos);
    ++launches;
    m->trace_mark("init", s);
  }
  for (int step = 0; step < T; ++step) {
    const int rd = step & 1, wr = rd ^ 1;  // ping-pong hidden buffers
    for (int l = 0; l < L; ++l) {
      if (!(step == 0 ? head : tail)) continue;
      TcLayer &t = m->tc[l];
      LayerArgs a{};
      a.n = n; a.m_tiles = m_tiles; a.n_blocks = Hs / kUnitBlock;
      a.n_dev = compact ? m->cp_state : nullptr;
      a.kb_x = l == 0 ? 0 : H / TC_BLOCK_K;  // layer 0: table gather in the epilogue, no x-part GEMM
      if (l == 0) { a.t0q = m->t0q; a.chars = m->chars; a.num_char = C; }
      a.kb_h = (step == 0) ? 0 : H / TC_BLOCK_K;  // h == 0 at step 0 (ref:653-654)
      a.pw_log2 = m->pw_log2; a.bpad = m->bpad; a.h_full = H; a.bias = t.bias; a.h_prev = t.h[rd]; a.h_out = t.h[wr];
      a.tp_world = W; a.tp_rank = R; a.kb_rank = Hs / TC_BLOCK_K; a.x_full = l > 0;
      a.col0 = R * Hs; a.flag_stride = mt_max;
      a.h_out_peer[0] = t.h[wr];
      if (W > 1) {
        m->tp_rounds[l][wr] += 1;  // this launch is one more write round of buffer (l, wr)
        for (int r = 0; r < W; ++r) {
          a.h_out_peer[r] = m->tp_h[r][l][wr];
          a.signal_peer[r] = m->tp_cnt[r][l][wr] + (size_t)R * mt_max;
        }
        if (m->tp_mc_base) {
          a.h_out_mc = m->tp_h_mc[l][wr];
          a.signal_mc = m->tp_cnt_mc[l][wr] + (size_t)R * mt_max;
        }
        a.wait_h = m->tp_cnt[R][l][rd];
        a.expect_h = m->tp_rounds[l][rd] * per_round;
        if (l > 0) {
          a.wait_x = m->tp_cnt[R][l - 1][wr];
          a.expect_x = m->tp_rounds[l - 1][wr] * per_round;
        }
      }
      const int tiles = a.m_tiles * a.n_blocks;
      {
        const bool pairs = m->use_pairs && m_tiles >= 2;
        cudaLaunchConfig_t cfg{};
        cudaLaunchAttribute attr[2];
        int na = 0;
        if (m->use_pdl) {  // prologue overlaps the previous kernel's tail
          attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
          attr[na].val.programmaticStreamSerializationAllowed = 1;
          ++na;
        }
        cfg.blockDim = dim3(L_THREADS);
        cfg.stream = s;
        if (pairs) {
          // 256-name tiles on CTA pairs (clusters of 2, one pair per TPC)
          const int tiles2 = ((m_tiles + 1) / 2) * a.n_blocks;
          const int npairs = tiles2 < sms / 2 ? tiles2 : sms / 2;
          cfg.gridDim = dim3(2 * npairs);
          cfg.dynamicSmemBytes = P_SMEM_BYTES;
          attr[na].id = cudaLaunchAttributeClusterDimension;
          attr[na].val.clusterDim.x = 2;
          attr[na].val.clusterDim.y = 1;
          attr[na].val.clusterDim.z = 1;
          ++na;
          cfg.attrs = attr;
          cfg.numAttrs = na;
          B200_CUDA(cudaLaunchKernelEx(&cfg, l == 0 ? gru_layer_pair_kernel<true> : gru_layer_pair_kernel<false>,
                                       l == 0 ? t.map_h[rd] : m->tc[l - 1].map_h[wr], t.map_h[rd],
                                       l == 0 ? t.map_w_hh_p : t.map_w_ih_p, t.map_w_hh_p,
                                       t.map_h[wr], a));
        } else {
          cfg.gridDim = dim3(tiles < sms ? tiles : sms);
          cfg.dynamicSmemBytes = L_SMEM_BYTES;
          cfg.attrs = attr;
          cfg.numAttrs = na;
          B200_CUDA(cudaLaunchKernelEx(&cfg, l == 0 ? gru_layer_kernel<true> : gru_layer_kernel<false>,
                                       l == 0 ? t.map_h[rd] : m->tc[l - 1].map_h[wr], t.map_h[rd],
                                       l == 0 ? t.map_w_hh : t.map_w_ih, t.map_w_hh, t.map_h[wr], a));
        }
      }
      ++launches;
      m->trace_mark(l == 0 ? "layer0" : "layerN", s);
    }
    if (!tail) break;  // part 1 ends before the first sampling kernel
    if (step == 0 && part == 0 && io->ready_event)
      B200_CUDA(cudaStreamWaitEvent(s, static_cast<cudaEvent_t>(io->ready_event), 0));
    if (step == 0 && io->ready_flag) {  // the floats are first needed here
      if (int rc = b200gru_wait_flags(io->ready_flag, 1, io->ready_value, s)) return rc;
      ++launches;
      m->trace_mark("wait_floats", s);
    }
    FcArgs a{};
    a.n = n; a.m_tiles = m_tiles; a.kb = H / TC_BLOCK_K; a.H = H; a.C = C; a.T = T;
    a.step = step; a.eos = c.eos;
    a.b_fc = m->b_fc; a.logits = m->debug_logits ? m->logits : nullptr;
    a.rfloats = io->rfloats; a.output = io->output;
    a.chars = m->chars; a.alive = m->alive;
    a.n_dev = compact ? m->cp_state : nullptr;
    a.slot_row = compact ? m->slot_row : nullptr;
    a.gone = compact ? m->cp_state + 3 : nullptr;
    a.tp_world = W; a.tp_rank = R; a.kb_rank = Hs / TC_BLOCK_K; a.flag_stride = mt_max;
    a.pw_log2 = m->pw_log2; a.bpad = m->bpad;
    if (W > 1) {
      a.wait_h = m->tp_cnt[R][L - 1][wr];
      a.expect_h = m->tp_rounds[L - 1][wr] * per_round;
    }
    {
      cudaLaunchConfig_t cfg{};
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.gridDim = dim3(m_tiles < sms ? m_tiles : sms);
      cfg.blockDim = dim3(F_THREADS);
      cfg.dynamicSmemBytes = F_SMEM_BYTES;
      cfg.stream = s;
      cfg.attrs = attr;
      cfg.numAttrs = m->use_pdl ? 1 : 0;
      B200_CUDA(cudaLaunchKernelEx(&cfg, fc_sample_kernel, m->tc[L - 1].map_h[wr], m->map_w_fc, a));
    }
    ++launches;
    m->trace_mark("fc_sample", s);
    if (compact && (step & 1) && step + 1 < T) {
      CompactBufs cb{};
      cb.layers = L; cb.H = H; cb.pw_log2 = m->pw_log2; cb.bpad = m->bpad;
      for (int l = 0; l < L; ++l) { cb.cur[l] = m->tc[l].h[wr]; cb.tmp[l] = m->tc[l].h[rd]; }
      // 8 warps (rows) per block, grid-stride; two blocks per SM keep the early
      // exit of an unneeded compaction at ~2 us per launch
      const unsigned mv_blocks = (unsigned)(rows < 8 * 296 ? (rows + 7) / 8 : 296);
      compact_plan_kernel<<<1, 1024, 0, s>>>(m->alive, m->new_idx, m->cp_state);
      compact_move_kernel<<<mv_blocks, 256, 0, s>>>(m->cp_state, m->new_idx, cb, m->chars, m->chars2, m->slot_row,
                                                    m->slot2);
      compact_commit_kernel<<<mv_blocks, 256, 0, s>>>(m->cp_state, cb, m->chars, m->chars2, m->slot_row, m->slot2,
                                                      m->alive);
      launches += 3;
      m->trace_mark("compact", s);
    }
  }
  if (tail && W > 1) {
    // Zero this rank's counters before announcing the end of the call.  Every
    // bump a peer sends here is waited for by one of our kernels in the same
    // step (the layer above, or FC for the top layer), so once our last FC
    // kernel is done no bump is in flight; peers start bumping for the next
    // call only after they have seen the epoch signal that follows.
    B200_CUDA(cudaMemsetAsync(m->tp_cnt[R][0][0], 0, (size_t)L * 2 * tp_counter_stride(m), s));
    m->tp_epoch += 1;
    TpPeers peers{};
    for (int r = 0; r < W; ++r) peers.p[r] = m->tp_sync[r] + R;
    tp_signal_epoch_kernel<<<1, 32, 0, s>>>(peers, W, m->tp_epoch);
    ++launches;
  }
  if (tail && io->done_flag) {
    tc_signal_done_kernel<<<1, 1, 0, s>>>(io->done_flag, io->done_value);
    ++launches;
  }
  B200_CUDA(cudaGetLastError());
  m->last_launches = launches;
  return 0;
}

}  // namespace b200gru

// Plain tcgen05 GEMM through the same TMA / UMMA / TMEM plumbing with a store
// epilogue — used by tests to validate descriptors in isolation.
namespace b200gru {
__global__ void __launch_bounds__(256, 1)
gemm_probe_kernel(const __grid_constant__ CUtensorMap map_a,
                  const __grid_constant__ CUtensorMap map_b, float *d, int M,
                  int N, int K, int n_tile) {
  extern __shared__ uint8_t smem_raw[];
  constexpr int STAGES = 4;
  constexpr int STAGE_BYTES = TC_A_BYTES + 256 * 128;
  uint8_t *smem = reinterpret_cast<uint8_t *>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + STAGES * STAGE_BYTES);
  uint64_t *full_bar = bars, *empty_bar = bars + STAGES;
  uint64_t *tmem_full = bars + 2 * STAGES;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * STAGES + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tile = blockIdx.y, n_blk = blockIdx.x;
  const int kb_total = K / TC_BLOCK_K;
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(tmem_full, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<256>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (warp == 0 && lane == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < kb_total; ++kb) {
      mbar_wait(&empty_bar[stage], phase ^ 1);
      uint8_t *sa = smem + stage * STAGE_BYTES;
      mbar_expect_tx(&full_bar[stage], TC_A_BYTES + n_tile * 128);
      tma_load_2d(sa, &map_a, &full_bar[stage], kb * TC_BLOCK_K, m_tile * kTileM);
      tma_load_2d(sa + TC_A_BYTES, &map_b, &full_bar[stage], kb * TC_BLOCK_K,
                  n_blk * n_tile);
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1 && lane == 0) {
    int stage = 0;
    uint32_t phase = 0;
    const uint32_t idesc = umma_idesc_bf16(kTileM, (uint32_t)n_tile);
    for (int kb = 0; kb < kb_total; ++kb) {
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
      const uint64_t da = umma_desc_sw128(sa);
      const uint64_t db = umma_desc_sw128(sa + TC_A_BYTES);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        umma_bf16(tmem_base, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
      umma_commit(&empty_bar[stage]);
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
    umma_commit(tmem_full);
  } else if (warp >= 4) {
    const int quarter = warp & 3;
    const int row = m_tile * kTileM + quarter * 32 + lane;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const uint32_t t = tmem_base + ((uint32_t)(quarter * 32) << 16);
    for (int c0 = 0; c0 < n_tile; c0 += 16) {
      uint32_t v[16];
      tmem_ld16(t + c0, v);
      tmem_ld_wait();
      if (row < M)
        for (int i = 0; i < 16; ++i) {
          int col = n_blk * n_tile + c0 + i;
          if (col < N) d[(size_t)row * N + col] = __uint_as_float(v[i]);
        }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
  }
}
}  // namespace b200gru

extern "C" int b200gru_gemm_bf16_tn(const void *a, const void *b, float *d,
                                    int m, int n, int k, void *stream) {
  using namespace b200gru;
  B200_REQUIRE(a && b && d, "null argument");
  B200_REQUIRE(m > 0 && n > 0 && k > 0 && k % 64 == 0, "k must be a multiple of 64");
  int n_tile = n >= 256 ? 256 : ((n + 15) / 16) * 16;
  B200_REQUIRE(n % n_tile == 0, "n must be <= 256 (multiple of 16) or a multiple of 256");
  CUtensorMap ma, mb;
  if (int rc = make_tmap_bf16_2d(&ma, a, (uint64_t)m, (uint64_t)k, kTileM, 64)) return rc;
  if (int rc = make_tmap_bf16_2d(&mb, b, (uint64_t)n, (uint64_t)k, (uint32_t)n_tile, 64)) return rc;
  const int smem = 4 * (TC_A_BYTES + 256 * 128) + 1024 + 256;
  B200_CUDA(cudaFuncSetAttribute(gemm_probe_kernel,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  dim3 grid(n / n_tile, (m + kTileM - 1) / kTileM);
  gemm_probe_kernel<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(
      ma, mb, d, m, n, k, n_tile);
  B200_CUDA(cudaGetLastError());
  return 0;
}
