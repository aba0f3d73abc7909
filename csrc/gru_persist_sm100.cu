// gru_persist_sm100.cu — the whole generate() as ONE persistent kernel with the
// recurrent weights resident in shared memory.
//
// The multi-launch path (gru_tc_sm100.cu) streams every weight tile from L2
// for every tile of every step and pays a kernel boundary per layer per
// character (31 launches per call).  That is the right trade for big batches —
// per SM it moves the fewest bytes through shared memory — but a small batch
// spends its time on launch gaps, prologues and idle SMs.  This kernel is the
// other end of the trade: the reference's name loop, step loop and its 51
// launches per character (ref:649, :661, :664-873) become one cooperative
// launch in which
//
//   * every CTA pair (cluster of 2, one TPC) permanently OWNS 16 hidden units of
//     every layer: its slice of W_hh (all layers) and W_ih (layers >= 1), bf16,
//     gate rows of a unit next to each other, is loaded ONCE by TMA into shared
//     memory (147 KB per CTA at H = 1024, L = 2) and stays there for all
//     MAX_LEN steps;
//   * per layer-step the pair runs tcgen05.mma (cta_group::2, M = 256 names,
//     N = 48 gate columns) with only the activations streamed from L2, gate math
//     in the TMEM epilogue, and stores its 16-unit slice of h';
//   * the layer-0 input projection is a row gather of the folded table
//     T0[c] = W_i*0 . emb[c] + biases in the epilogue (no one-hot GEMM);
//   * FC + softmax + sampling (ref:858-878) run in the same kernel on the CTAs
//     that hold no weights, one 128-name tile each;
//   * there are no grid-wide barriers: every dependency is a per-128-name-tile
//     counter in global memory (release by the producer's epilogue, acquire by
//     the consumer's TMA warp), so the recurrent half of layer 0's next step
//     already runs while FC and sampling of the current step are in flight;
//   * the DP pieces stay inside: random floats are loaded from the root rank's
//     buffer (ref:636), characters are stored into the root's rows (ref:889),
//     the wait for the floats and the completion flag are in-kernel;
//   * name tiles whose names have all hit EOS are skipped (ref:881-882).
//
// Roles per weight-holding CTA: warp 0 TMA producer, warp 1 MMA issuer (leader
// CTA only), warp 2 TMEM allocator, warps 4-11 epilogue.  FC CTAs use the same
// warp roles with cta_group::1.
#include <cstdio>
#include <cstdlib>

#include "model.hpp"
#include "sm100_ptx.cuh"

namespace b200gru {
using namespace sm100;

constexpr int PS_UNITS = 16;                 // hidden units owned by a CTA pair
constexpr int PS_ROWS = 24;                  // weight rows a CTA holds per matrix: 3 gates x 16 units / 2
constexpr int PS_N = 48;                     // UMMA N over the pair
constexpr int PS_KB_BYTES = PS_ROWS * 128;   // one 64-wide k-block of a CTA's slice (3 swizzle atoms)
constexpr int PS_A_BYTES = kTileM * 128;     // 16 KB: 128 names x 64 k
constexpr int PS_ACC_COLS = 64;              // [ n_i | r | z | n_h ] x 16 units
constexpr int PS_ACC_STAGES = 8;             // accumulator stages: the MMA warp may run 8 tiles ahead of the epilogue
constexpr int PS_TMEM_COLS = PS_ACC_STAGES * PS_ACC_COLS;  // all 512 columns
constexpr int PS_THREADS = 384;
constexpr int PS_EPI_WARPS = 8;
constexpr int PS_MAX_STAGES = 8;
constexpr int PS_FC_STAGES = 6;
constexpr int PS_EPI_BAR = 1;
constexpr uint32_t kSpinNs = 0;   // waits on the per-step critical path poll back to back

struct PersistMaps {
  CUtensorMap h[B200GRU_MAX_LAYERS][2];   // hidden buffers, box {64, 128}
  CUtensorMap whh[B200GRU_MAX_LAYERS];    // packed W_hh, box {64, 24}
  CUtensorMap wih[B200GRU_MAX_LAYERS];    // packed W_ih (layers >= 1)
  CUtensorMap wfc;                        // W_fc, box {64, C/2}: each CTA of an FC pair takes half
};

struct PersistArgs {
  int n, m_tiles, m2_tiles, mt_max;
  int T, L, H, C, kbH;
  int P;                 // CTA pairs that hold weights (H / 16)
  int n_fc;              // FC CTAs
  int stages;            // A-operand ring depth of a weight CTA
  int sos, eos;
  uint32_t w_bytes;      // resident weight bytes per CTA
  uint32_t off_ih[B200GRU_MAX_LAYERS], off_hh[B200GRU_MAX_LAYERS];  // byte offsets in the resident block
  const float *bias[B200GRU_MAX_LAYERS];     // [H/64][4][64]: b_in, b_r, b_z, b_hn (layer 0: folded into T0)
  __nv_bfloat16 *h[B200GRU_MAX_LAYERS][2];   // [bpad, H] ping-pong
  const __nv_bfloat16 *t0p;                  // [P][C][3: n,r,z][16]
  uint32_t *cnt_h;       // [L][2][mt_max]  CTAs that have stored their slice of (layer, buffer, 128-name tile)
  uint32_t *cnt_c;       // [mt_max]        2 x steps whose characters are sampled for the tile (one bump per FC CTA)
  int32_t *live;         // [T][mt_max]     names of the tile still running after step t
  uint32_t *done_cnt;    // FC CTAs that have finished the call
  const float *b_fc;
  float *logits;         // optional tap
  const float *rfloats;  // may be a peer address
  char *output;          // may be a peer address
  int32_t *chars;
  uint8_t *alive;
  uint32_t *done_flag;
  uint32_t done_value;
  const uint32_t *ready_flag;
  uint32_t ready_value;
  // optional in-kernel timeline (b200gru_model_persist_timeline): globaltimer
  // stamps of the first weight CTA and the first FC CTA, per role
  unsigned long long *timeline;
  int timeline_fc_rank;  // which CTA of the first FC pair is instrumented (B200GRU_TL_FC_RANK)
};

// ---- timeline: [cta: 0 = first weight CTA, 1 = first FC CTA][role: 0 producer, 1 MMA, 2 epilogue][PS_TL_EVENTS]
constexpr int PS_TL_EVENTS = 256;
struct Timeline {
  unsigned long long *p;
  int i;
  __device__ __forceinline__ void stamp(uint32_t code) {
    if (p && i < PS_TL_EVENTS) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      p[i++] = (t << 16) | (code & 0xffffu);
    }
  }
};
__device__ __forceinline__ Timeline timeline_for(const unsigned long long *base, int cta, int role, bool on) {
  Timeline t;
  t.p = (base && on) ? const_cast<unsigned long long *>(base) + ((size_t)cta * 3 + role) * PS_TL_EVENTS : nullptr;
  t.i = 0;
  return t;
}
// event codes: (step << 8) | (layer << 4) | what
enum { TL_WAIT_DONE = 1, TL_LOADS_ISSUED = 2, TL_MMA_START = 3, TL_MMA_ISSUED = 4, TL_ACC_READY = 5,
       TL_PUBLISHED = 6, TL_CHARS_READY = 7, TL_START = 8, TL_WEIGHTS = 9 };

// All lanes poll one address (one request); every lane has then acquired.
// A cooperative kernel whose CTAs wait on each other must never hang the GPU on
// a lost signal: after ~10 s of polling the wait traps (the launch then fails
// with an error instead of spinning forever).
constexpr uint32_t kWatchdogPolls = 1u << 24;
__device__ __noinline__ void wait_timed_out(const uint32_t *p, uint32_t expect, uint32_t seen) {
  printf("b200gru persistent kernel: block %d warp %d waited for counter %p >= %u, last saw %u\n",
         (int)blockIdx.x, (int)(threadIdx.x >> 5), (const void *)p, expect, seen);
  __trap();
}
__device__ __forceinline__ void wait_counter_gpu(const uint32_t *p, uint32_t expect) {
  uint32_t polls = 0, v;
  while ((int32_t)((v = ld_acquire_gpu(p)) - expect) < 0) {
    if (kSpinNs) __nanosleep(kSpinNs);
    if (++polls == kWatchdogPolls) wait_timed_out(p, expect, v);
  }
}

// A 256-name pair tile is skipped once both of its 128-name halves had no
// running name left two steps ago.  "Two steps ago" is the newest information
// that every role of every CTA is certain to agree on without waiting for the
// sampling of the previous step, which the recurrent MMAs of layer 0 deliberately
// do not wait for.
__device__ __forceinline__ bool pair_tile_dead(const PersistArgs &a, int step, int m2) {
  if (step < 2) return false;
  bool dead = true;
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int mt = m2 * 2 + r;
    if (mt >= a.m_tiles) continue;  // padding rows only
    wait_counter_gpu(a.cnt_c + mt, 2u * (uint32_t)(step - 1));  // already true; orders the read below
    dead = dead && (__ldcg(a.live + (size_t)(step - 2) * a.mt_max + mt) == 0);
  }
  return dead;
}

__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }

// ------------------------------------------------------------ weight CTAs
__device__ __forceinline__ void persist_main(const PersistMaps &maps, const PersistArgs &a,
                                             uint8_t *smem) {
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1;
  const int S = a.stages;

  uint8_t *wsm = smem;                               // resident weights
  uint8_t *ring = smem + a.w_bytes;                  // A-operand ring
  uint64_t *bars = reinterpret_cast<uint64_t *>(ring + (size_t)S * PS_A_BYTES);
  uint64_t *full_bar = bars;                         // [S]  (leader's is used)
  uint64_t *empty_bar = bars + PS_MAX_STAGES;        // [S]  (each CTA its own)
  uint64_t *tmem_full = bars + 2 * PS_MAX_STAGES;    // [PS_ACC_STAGES]  (each CTA its own)
  uint64_t *tmem_empty = tmem_full + PS_ACC_STAGES;  // [PS_ACC_STAGES]  (leader's is used)
  uint64_t *wbar = tmem_empty + PS_ACC_STAGES;       // resident weights landed (leader's is used)
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(wbar + 1);
  const bool tl_on = blockIdx.x == 0;

  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < PS_ACC_STAGES; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 2 * PS_EPI_WARPS);  // epilogue warps of both CTAs
    }
    mbar_init(wbar, 1);
    fence_mbar_init();
  }
  cluster_sync_all();
  if (warp == 2) tmem_alloc_2sm<PS_TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // Every column of a stage is zero whenever the stage is in the MMA warp's
  // hands, so every MMA simply accumulates and the two halves of a tile (own
  // previous state first, the layer below's fresh state last) can come in the
  // order their operands become available.
  if (warp >= 4) {
    const uint32_t lane_base0 = (uint32_t)((warp & 3) * 32) << 16;
    const int half0 = (warp - 4) >> 2;
    uint32_t zero[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) zero[i] = 0u;
#pragma unroll 1
    for (int st = 0; st < PS_ACC_STAGES; ++st)
#pragma unroll
      for (int g = 0; g < 4; ++g)
        tmem_st8(tmem_base + lane_base0 + st * PS_ACC_COLS + g * 16 + half0 * 8, zero);
    tmem_st_wait();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();

  const int kbH = a.kbH;
  if (warp == 0) {
    // =========================================================== TMA producer
    // resident weights first: every k-block of every matrix, once
    if (elect_one()) {
      if (leader) mbar_expect_tx(wbar, 2 * a.w_bytes);
      const uint32_t bar = mapa_u32(wbar, 0);
      const int row = pair * PS_N + (int)rank * PS_ROWS;
      for (int l = 0; l < a.L; ++l) {
        if (l > 0)
          for (int kb = 0; kb < kbH; ++kb)
            tma_load_2d_2sm(wsm + a.off_ih[l] + kb * PS_KB_BYTES, &maps.wih[l], bar, kb * 64, row);
        for (int kb = 0; kb < kbH; ++kb)
          tma_load_2d_2sm(wsm + a.off_hh[l] + kb * PS_KB_BYTES, &maps.whh[l], bar, kb * 64, row);
      }
    }
    __syncwarp();
    Timeline tl = timeline_for(a.timeline, 0, 0, tl_on && lane == 0);
    tl.stamp(TL_START);
    int stage = 0;
    uint32_t phase = 0;
    for (int step = 0; step < a.T; ++step) {
      const int rd = step & 1, wr = rd ^ 1;
      for (int l = 0; l < a.L; ++l)
        for (int m2 = 0; m2 < a.m2_tiles; ++m2) {
          if (pair_tile_dead(a, step, m2)) continue;
          const int mt = m2 * 2 + (int)rank;
          for (int part = 1; part >= 0; --part) {
            // part 1: own previous state (ready since the last step) first;
            // part 0: x = the layer below's fresh state, the operand on the critical path
            if (part == 0 ? (l == 0) : (step == 0)) continue;
            const int src_l = part == 0 ? l - 1 : l;
            const int src_b = part == 0 ? wr : rd;
            const uint32_t rounds = part == 0 ? (uint32_t)(step / 2 + 1) : (uint32_t)((step - 1) / 2 + 1);
            wait_counter_gpu(a.cnt_h + ((size_t)src_l * 2 + src_b) * a.mt_max + mt, rounds * (uint32_t)a.P);
            tl.stamp((step << 8) | (l << 4) | TL_WAIT_DONE);
            const CUtensorMap *map = &maps.h[src_l][src_b];
            for (int kb = 0; kb < kbH; ++kb) {
              mbar_wait_relaxed(&empty_bar[stage], phase ^ 1, 20);
              if (elect_one()) {
                if (kb == 0) fence_proxy_async();  // other SMs' generic stores -> our TMA reads
                if (leader) mbar_expect_tx(&full_bar[stage], 2 * PS_A_BYTES);
                tma_load_2d_2sm(ring + (size_t)stage * PS_A_BYTES, map, mapa_u32(&full_bar[stage], 0),
                                kb * 64, mt * kTileM);
              }
              __syncwarp();
              if (++stage == S) { stage = 0; phase ^= 1; }
            }
            tl.stamp((step << 8) | (l << 4) | TL_LOADS_ISSUED);
          }
        }
    }
  } else if (warp == 1) {
    // ==================================================== MMA issuer (leader)
    if (leader) {
      int stage = 0, acc_stage = 0;
      uint32_t phase = 0, acc_phase = 0;
      const uint32_t idesc = umma_idesc_bf16(2 * kTileM, PS_N);
      const uint64_t desc_ring = umma_desc_sw128(smem_u32(ring));
      const uint64_t desc_w = umma_desc_sw128(smem_u32(wsm));
      Timeline tl = timeline_for(a.timeline, 0, 1, tl_on && lane == 0);
      tl.stamp(TL_START);
      mbar_wait(wbar, 0);
      tc_fence_after();
      tl.stamp(TL_WEIGHTS);
      for (int step = 0; step < a.T; ++step)
        for (int l = 0; l < a.L; ++l)
          for (int m2 = 0; m2 < a.m2_tiles; ++m2) {
            if (pair_tile_dead(a, step, m2)) continue;
            const int kb_h = step > 0 ? kbH : 0;      // own previous state comes first
            const int kb_total = kb_h + (l > 0 ? kbH : 0);
            if (kb_total == 0) continue;  // layer 0, step 0: the table row is everything
            mbar_wait(&tmem_empty[acc_stage], acc_phase ^ 1);
            tc_fence_after();
            const uint32_t d = tmem_base + acc_stage * PS_ACC_COLS;
            for (int kb = 0; kb < kb_total; ++kb) {
              mbar_wait(&full_bar[stage], phase);
              tc_fence_after();
              if (kb == 0) tl.stamp((step << 8) | (l << 4) | TL_MMA_START);
              if (elect_one()) {
                const uint64_t da = desc_ring + (uint64_t)((stage * PS_A_BYTES) >> 4);
                B200_DASSERT(stage < S && acc_stage < PS_ACC_STAGES &&
                                 (kb < kb_h ? a.off_hh[l] + (kb + 1) * PS_KB_BYTES
                                            : a.off_ih[l] + (kb - kb_h + 1) * PS_KB_BYTES) <= a.w_bytes,
                             "resident weight block / ring stage out of range (layer %d k-block %d)", l, kb);
                if (kb < kb_h) {  // [r | z | n_h] += h . W_hh^T
                  const uint64_t db = desc_w + (uint64_t)((a.off_hh[l] + kb * PS_KB_BYTES) >> 4);
                  umma_bf16_2sm(d + 16, da, db, idesc, 1u);
                  umma_bf16_2sm(d + 16, da + 2, db + 2, idesc, 1u);
                  umma_bf16_2sm(d + 16, da + 4, db + 4, idesc, 1u);
                  umma_bf16_2sm(d + 16, da + 6, db + 6, idesc, 1u);
                } else {          // [n_i | r | z] += x . W_ih^T
                  const uint64_t db = desc_w + (uint64_t)((a.off_ih[l] + (kb - kb_h) * PS_KB_BYTES) >> 4);
                  umma_bf16_2sm(d, da, db, idesc, 1u);
                  umma_bf16_2sm(d, da + 2, db + 2, idesc, 1u);
                  umma_bf16_2sm(d, da + 4, db + 4, idesc, 1u);
                  umma_bf16_2sm(d, da + 6, db + 6, idesc, 1u);
                }
                umma_commit_2sm(&empty_bar[stage], 0b11);
                if (kb == kb_total - 1) umma_commit_2sm(&tmem_full[acc_stage], 0b11);
              }
              __syncwarp();
              if (++stage == S) { stage = 0; phase ^= 1; }
            }
            tl.stamp((step << 8) | (l << 4) | TL_MMA_ISSUED);
            if (++acc_stage == PS_ACC_STAGES) { acc_stage = 0; acc_phase ^= 1; }
          }
    }
  } else if (warp >= 4) {
    // ============================================================== epilogue
    // thread = (name, 8 of the pair's 16 units): warp % 4 picks the TMEM lane
    // quarter, warp / 4 the unit half.
    const int quarter = warp & 3;
    const int half = (warp - 4) >> 2;
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    const int unit0 = pair * PS_UNITS + half * 8;     // first of this thread's hidden units
    const int etid = threadIdx.x - 128;
    Timeline tl = timeline_for(a.timeline, 0, 2, tl_on && etid == 0);
    tl.stamp(TL_START);
    int acc_stage = 0;
    uint32_t acc_phase = 0;
    for (int step = 0; step < a.T; ++step) {
      const int rd = step & 1, wr = rd ^ 1;
      for (int l = 0; l < a.L; ++l)
        for (int m2 = 0; m2 < a.m2_tiles; ++m2) {
          if (pair_tile_dead(a, step, m2)) continue;
          const int mt = m2 * 2 + (int)rank;
          const int row = mt * kTileM + quarter * 32 + lane;
          const bool has_mma = (l > 0) || (step > 0);
          // previous state of this thread's units (written by this very thread)
          uint4 hq = make_uint4(0, 0, 0, 0);
          if (step > 0) hq = __ldcg(reinterpret_cast<const uint4 *>(a.h[l][rd] + (size_t)row * a.H + unit0));
          // biases: [H/64][4][64]
          const float *bl = a.bias[l] + (size_t)(unit0 >> 6) * 256 + (unit0 & 63);
          float b_in[8], b_r[8], b_z[8], b_hn[8];
#pragma unroll
          for (int v = 0; v < 2; ++v) {
            const float4 t0 = __ldg(reinterpret_cast<const float4 *>(bl) + v);
            const float4 t1 = __ldg(reinterpret_cast<const float4 *>(bl + 64) + v);
            const float4 t2 = __ldg(reinterpret_cast<const float4 *>(bl + 128) + v);
            const float4 t3 = __ldg(reinterpret_cast<const float4 *>(bl + 192) + v);
            b_in[v * 4] = t0.x; b_in[v * 4 + 1] = t0.y; b_in[v * 4 + 2] = t0.z; b_in[v * 4 + 3] = t0.w;
            b_r[v * 4] = t1.x; b_r[v * 4 + 1] = t1.y; b_r[v * 4 + 2] = t1.z; b_r[v * 4 + 3] = t1.w;
            b_z[v * 4] = t2.x; b_z[v * 4 + 1] = t2.y; b_z[v * 4 + 2] = t2.z; b_z[v * 4 + 3] = t2.w;
            b_hn[v * 4] = t3.x; b_hn[v * 4 + 1] = t3.y; b_hn[v * 4 + 2] = t3.z; b_hn[v * 4 + 3] = t3.w;
          }
          float xn[8], xr[8], xz[8];
          if (l == 0) {
            // layer-0 input projection = row of the folded table for the name's
            // current character (embedding + 3 matvecs + biases, ref:669-733)
            int c = a.sos;  // ref:652
            if (step > 0 && mt < a.m_tiles) {
              wait_counter_gpu(a.cnt_c + mt, 2u * (uint32_t)step);  // characters of step-1 are sampled (both FC CTAs)
              tl.stamp((step << 8) | (l << 4) | TL_CHARS_READY);
              if (row < a.n) c = __ldcg(a.chars + row);
            }
            B200_DASSERT(c >= 0 && c < a.C, "layer-0 table row %d of %d", c, a.C);
            const uint4 *trow = reinterpret_cast<const uint4 *>(
                a.t0p + ((size_t)pair * a.C + c) * (3 * PS_UNITS) + half * 8);
            const uint4 tn = __ldg(trow), tr = __ldg(trow + 2), tz = __ldg(trow + 4);
            const uint32_t wn[4] = {tn.x, tn.y, tn.z, tn.w}, wr4[4] = {tr.x, tr.y, tr.z, tr.w},
                           wz[4] = {tz.x, tz.y, tz.z, tz.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              xn[2 * i] = bf_lo(wn[i]); xn[2 * i + 1] = bf_hi(wn[i]);
              xr[2 * i] = bf_lo(wr4[i]); xr[2 * i + 1] = bf_hi(wr4[i]);
              xz[2 * i] = bf_lo(wz[i]); xz[2 * i + 1] = bf_hi(wz[i]);
            }
          }
          uint32_t vn[8], vr[8], vz[8], vh[8];
          if (has_mma) {
            mbar_wait_relaxed(&tmem_full[acc_stage], acc_phase, 20);
            tc_fence_after();
            tl.stamp((step << 8) | (l << 4) | TL_ACC_READY);
            const uint32_t t = tmem_base + lane_base + acc_stage * PS_ACC_COLS + half * 8;
            if (l > 0) tmem_ld8(t, vn);
            tmem_ld8(t + 16, vr);
            tmem_ld8(t + 32, vz);
            tmem_ld8(t + 48, vh);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) vr[i] = vz[i] = vh[i] = 0u;
          }
          if (l > 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              xn[i] = __uint_as_float(vn[i]);
              xr[i] = 0.0f;
              xz[i] = 0.0f;
            }
          }
          const uint32_t hw[4] = {hq.x, hq.y, hq.z, hq.w};
          uint32_t packed[4];
#pragma unroll
          for (int i = 0; i < 8; i += 2) {
            float hn2[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int j = i + e;
              const float hp = e ? bf_hi(hw[i >> 1]) : bf_lo(hw[i >> 1]);
              const float rr = sigmoid_approx(__uint_as_float(vr[j]) + xr[j] + b_r[j]);
              const float z = sigmoid_approx(__uint_as_float(vz[j]) + xz[j] + b_z[j]);
              const float nn = tanh_approx(fmaf(rr, __uint_as_float(vh[j]) + b_hn[j], xn[j] + b_in[j]));
              hn2[e] = fmaf(z, hp - nn, nn);  // (1-z)*n + z*h, ref:752-763
            }
            __nv_bfloat162 p = __floats2bfloat162_rn(hn2[0], hn2[1]);
            packed[i >> 1] = *reinterpret_cast<uint32_t *>(&p);
          }
          B200_DASSERT(row < a.mt_max * kTileM && unit0 + 8 <= a.H && acc_stage < PS_ACC_STAGES,
                       "h' store row %d unit %d / accumulator stage %d out of range", row, unit0, acc_stage);
          *reinterpret_cast<uint4 *>(a.h[l][wr] + (size_t)row * a.H + unit0) =
              make_uint4(packed[0], packed[1], packed[2], packed[3]);
          // publish: this CTA's slice of (layer, buffer, tile) is in place
          named_bar_sync(PS_EPI_BAR, PS_EPI_WARPS * 32);
          if (etid == 0) {
            // the barrier above orders every epilogue thread's stores before this
            // release (cumulativity), which publishes them at GPU scope
            red_release_gpu_add(a.cnt_h + ((size_t)l * 2 + wr) * a.mt_max + mt, 1u);
            tl.stamp((step << 8) | (l << 4) | TL_PUBLISHED);
          }
          if (has_mma) {
            // only now (off the critical path: the MMA warp has seven more
            // stages) hand the accumulator stage back, zeroed
            const uint32_t t = tmem_base + lane_base + acc_stage * PS_ACC_COLS + half * 8;
            uint32_t zero[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) zero[i] = 0u;
            if (l > 0) tmem_st8(t, zero);
            tmem_st8(t + 16, zero);
            tmem_st8(t + 32, zero);
            tmem_st8(t + 48, zero);
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa_u32(&tmem_empty[acc_stage], 0));
            if (++acc_stage == PS_ACC_STAGES) { acc_stage = 0; acc_phase ^= 1; }
          }
        }
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2sm<PS_TMEM_COLS>(tmem_base);
  }
}

// ----------------------------------------------------------------- FC CTAs
// FC + softmax + sampling (ref:858-878) for one 128-name tile per CTA PAIR.
// The two CTAs split the alphabet: CTA c computes logits of characters
// [c*C/2, (c+1)*C/2) (UMMA M = 128, N = C/2), so the GEMM on the critical path of
// every character takes half as long as on one CTA.  Inside a CTA two threads
// share a name, so a name's probabilities live in four segments.  Softmax needs
// one exchange: every thread publishes the sum of its segment's exponentials to
// both CTAs (distributed shared memory + an mbarrier that collects the warps of
// both CTAs); with the four sums every thread knows the total and where its own
// segment starts in the CDF, hence which segment contains the sampled point —
// only that segment's thread walks, 16 columns at most, and does the name's
// bookkeeping.
constexpr int PS_FC_STAGE_BYTES = PS_A_BYTES + 128 * 128;   // A tile + up to 128 rows of W_fc
constexpr int PS_FC_TMEM_COLS = 256;                          // two accumulator stages of <= 128 columns

__device__ __forceinline__ void persist_fc(const PersistMaps &maps, const PersistArgs &a, uint8_t *smem) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = a.C, L = a.L;
  const uint32_t rank = cluster_ctarank();                 // which half of the alphabet
  const int fp = ((int)blockIdx.x - 2 * a.P) >> 1;         // FC pair index
  const int n_fcp = a.n_fc >> 1;
  const int Nh = C >> 1;
  const uint32_t b_bytes = (uint32_t)Nh * 128u;
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + PS_FC_STAGES * PS_FC_STAGE_BYTES);
  uint64_t *full_bar = bars, *empty_bar = bars + PS_FC_STAGES;
  uint64_t *tmem_full = bars + 2 * PS_FC_STAGES, *tmem_empty = tmem_full + 2;
  uint64_t *xbar = tmem_empty + 2;                         // segment sums of both CTAs are in place
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(xbar + 1);
  float *seg_sum = reinterpret_cast<float *>(smem + PS_FC_STAGES * PS_FC_STAGE_BYTES + 512);  // [2][4][128]

  if (warp == 1 && lane == 0) {
    for (int i = 0; i < PS_FC_STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], PS_EPI_WARPS); }
    mbar_init(xbar, 2 * PS_EPI_WARPS);                     // epilogue warps of both CTAs
    fence_mbar_init();
  }
  cluster_sync_all();                                       // the partner's barriers exist before we signal them
  if (warp == 2) tmem_alloc<PS_FC_TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // A tile is skipped (no GEMM, no sampling) once it reported no running names;
  // every role of both CTAs reads that from the same global word.  cnt_c counts
  // two publications per step (one per CTA of the pair).
  auto tile_dead = [&](int step, int mt) -> bool {
    if (step == 0) return false;
    wait_counter_gpu(a.cnt_c + mt, 2u * (uint32_t)step);
    return __ldcg(a.live + (size_t)(step - 1) * a.mt_max + mt) == 0;
  };
  const bool tl_on = fp == 0 && (int)rank == a.timeline_fc_rank;

  if (warp == 0) {
    Timeline tl = timeline_for(a.timeline, 1, 0, tl_on && lane == 0);
    tl.stamp(TL_START);
    int stage = 0;
    uint32_t phase = 0;
    for (int step = 0; step < a.T; ++step) {
      const int wr = (step & 1) ^ 1;
      for (int mt = fp; mt < a.m_tiles; mt += n_fcp) {
        if (tile_dead(step, mt)) continue;
        wait_counter_gpu(a.cnt_h + ((size_t)(L - 1) * 2 + wr) * a.mt_max + mt,
                         (uint32_t)(step / 2 + 1) * (uint32_t)a.P);
        tl.stamp((step << 8) | TL_WAIT_DONE);
        for (int kb = 0; kb < a.kbH; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (elect_one()) {
            uint8_t *sa = smem + stage * PS_FC_STAGE_BYTES;
            if (kb == 0) fence_proxy_async();
            mbar_expect_tx(&full_bar[stage], PS_A_BYTES + b_bytes);
            tma_load_2d(sa, &maps.h[L - 1][wr], &full_bar[stage], kb * 64, mt * kTileM);
            tma_load_2d(sa + PS_A_BYTES, &maps.wfc, &full_bar[stage], kb * 64, (int)rank * Nh);
          }
          __syncwarp();
          if (++stage == PS_FC_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    int stage = 0, acc_stage = 0;
    uint32_t phase = 0, acc_phase = 0;
    const uint32_t idesc = umma_idesc_bf16(kTileM, (uint32_t)Nh);
    const uint64_t desc0 = umma_desc_sw128(smem_u32(smem));
    Timeline tl = timeline_for(a.timeline, 1, 1, tl_on && lane == 0);
    tl.stamp(TL_START);
    for (int step = 0; step < a.T; ++step)
      for (int mt = fp; mt < a.m_tiles; mt += n_fcp) {
        if (tile_dead(step, mt)) continue;
        mbar_wait(&tmem_empty[acc_stage], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d = tmem_base + acc_stage * 128;
        for (int kb = 0; kb < a.kbH; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (kb == 0) tl.stamp((step << 8) | TL_MMA_START);
          if (elect_one()) {
            const uint64_t da = desc0 + (uint64_t)((stage * PS_FC_STAGE_BYTES) >> 4);
            const uint64_t db = da + (uint64_t)(PS_A_BYTES >> 4);
            umma_bf16(d, da, db, idesc, kb != 0 ? 1u : 0u);
            umma_bf16(d, da + 2, db + 2, idesc, 1u);
            umma_bf16(d, da + 4, db + 4, idesc, 1u);
            umma_bf16(d, da + 6, db + 6, idesc, 1u);
            umma_commit(&empty_bar[stage]);
            if (kb == a.kbH - 1) umma_commit(&tmem_full[acc_stage]);
          }
          __syncwarp();
          if (++stage == PS_FC_STAGES) { stage = 0; phase ^= 1; }
        }
        tl.stamp((step << 8) | TL_MMA_ISSUED);
        if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
      }
  } else if (warp >= 4) {
    const int quarter = warp & 3;
    const int half = (warp - 4) >> 2;
    const int rl = quarter * 32 + lane;                    // row within the tile
    const int etid = threadIdx.x - 128;
    const int seg = 2 * (int)rank + half;                  // this thread's segment of the alphabet (in CDF order)
    const int split = ((Nh / 16 + 1) / 2) * 16;
    const int cbeg = half == 0 ? 0 : split, cend = half == 0 ? split : Nh;   // columns of this CTA's accumulator
    const int col0 = (int)rank * Nh;                       // character id of accumulator column 0
    const uint32_t seg_remote = mapa_u32(seg_sum, rank ^ 1u);
    const uint32_t xbar_remote = mapa_u32(xbar, rank ^ 1u);
    const uint32_t xbar_local = mapa_u32(xbar, rank);
    int acc_stage = 0;
    uint32_t acc_phase = 0;
    const float kLog2e = 1.4426950408889634f;
    Timeline tl = timeline_for(a.timeline, 1, 2, tl_on && etid == 0);
    tl.stamp(TL_START);
    if (a.ready_flag) {  // the root's upload of the random floats has landed (ref:636)
      uint32_t polls = 0, v;
      while ((int32_t)((v = ld_acquire_sys(a.ready_flag)) - a.ready_value) < 0) {
        __nanosleep(200);
        if (++polls == 4 * kWatchdogPolls) wait_timed_out(a.ready_flag, a.ready_value, v);
      }
    }
    int it = 0;
    for (int step = 0; step < a.T; ++step)
      for (int mt = fp; mt < a.m_tiles; mt += n_fcp) {
        if (tile_dead(step, mt)) {
          if (etid == 0) red_release_gpu_add(a.cnt_c + mt, 1u);  // live[step][mt] stays 0
          continue;
        }
        const int row = mt * kTileM + rl;
        const bool valid = row < a.n;
        const bool live = valid && (step == 0 || __ldcg(a.alive + row) != 0);
        // this step's random float: fetched (over NVLink from the root rank's
        // buffer when data parallel, ref:636) long before it is needed
        const float r = live ? a.rfloats[(size_t)row * a.T + step] : 2.0f;
        mbar_wait_relaxed(&tmem_full[acc_stage], acc_phase, 20);
        tc_fence_after();
        tl.stamp((step << 8) | TL_ACC_READY);
        const uint32_t t = tmem_base + ((uint32_t)(quarter * 32) << 16) + acc_stage * 128;
        // pass 1: e_i = exp(f_i + b_i) kept in TMEM in place; one sum per 16 columns
        float cs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int c0 = cbeg + 16 * k;
          if (c0 < cend) {
            uint32_t v[16];
            tmem_ld16(t + c0, v);
            tmem_ld_wait();
            float e[16];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float4 b = __ldg(reinterpret_cast<const float4 *>(a.b_fc + col0 + c0) + q);
              const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float x = __uint_as_float(v[q * 4 + j]) + bb[j];
                if (a.logits && valid) a.logits[(size_t)row * C + col0 + c0 + q * 4 + j] = live ? x : 0.0f;
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
        tl.stamp((step << 8) | 10);
        // exchange: this segment's mass to both CTAs
        float *mine_sums = seg_sum + (size_t)(it & 1) * 4 * kTileM;
        const float my_sum = ((cs[0] + cs[1]) + cs[2]) + cs[3];
        mine_sums[seg * kTileM + rl] = my_sum;
        st_shared_cluster_f32(seg_remote + (uint32_t)(((it & 1) * 4 + seg) * kTileM + rl) * 4u, my_sum);
        __syncwarp();
        if (lane == 0) {
          mbar_arrive_release_cluster(xbar_local);
          mbar_arrive_release_cluster(xbar_remote);
        }
        mbar_wait_cluster(xbar, (uint32_t)it & 1u);
        tl.stamp((step << 8) | 11);
        const float S0 = mine_sums[rl], S1 = mine_sums[kTileM + rl], S2 = mine_sums[2 * kTileM + rl],
                    S3 = mine_sums[3 * kTileM + rl];
        const float P1 = S0, P2 = P1 + S1, P3 = P2 + S2, tot = P3 + S3;
        // first i with sum_{j<=i} e_j > r * sum (ref:322-333 on unnormalised mass);
        // if the mass never exceeds it the last character wins (ref:332)
        const float rs = r * tot;
        const bool fallback = !(tot > rs);
        const int owner = fallback ? 3 : (P1 > rs ? 0 : (P2 > rs ? 1 : (P3 > rs ? 2 : 3)));
        const bool mine = live && owner == seg;
        int sel = (mine && fallback) ? C - 1 : -1;
        float p = seg == 0 ? 0.0f : (seg == 1 ? P1 : (seg == 2 ? P2 : P3));
#pragma unroll
        for (int k = 0; k < 4; ++k) {
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
                  if (sel < 0 && q > rs) sel = col0 + c0 + i;
                }
                if (sel < 0) sel = col0 + c0 + 15;      // tree sum crossed, running sum did not (rounding)
              }
            }
            p += cs[k];
          }
        }
        if (mine && sel < 0) sel = col0 + cend - 1;
        tl.stamp((step << 8) | 12);
        // accumulator drained: hand the TMEM stage back before the bookkeeping
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[acc_stage]);
        bool still = false;
        char *orow = a.output + (size_t)row * (a.T + 1);
        B200_DASSERT(!mine || (sel >= 0 && sel < C && row < a.n), "sampled character %d of %d, row %d of %d", sel, C,
                     row, a.n);
        if (mine) {
          a.chars[row] = sel;
          still = sel != a.eos;                     // ref:881
          a.alive[row] = still ? 1 : 0;
          if (step == 0) {                          // first byte + NUL padding, ref:643
            orow[0] = (char)sel;
            for (int l = 1; l <= a.T; ++l) orow[l] = 0;
          }
        }
        const unsigned running = __ballot_sync(0xffffffffu, still);
        if (lane == 0 && running) atomicAdd(a.live + (size_t)step * a.mt_max + mt, __popc(running));
        named_bar_sync(PS_EPI_BAR, PS_EPI_WARPS * 32);
        tl.stamp((step << 8) | 13);
        if (etid == 0) {
          red_release_gpu_add(a.cnt_c + mt, 1u);   // this CTA's share of the tile's characters is in place
          tl.stamp((step << 8) | TL_PUBLISHED);
        }
        // the output byte (possibly a store over NVLink into the root's row,
        // ref:877, :889) is not needed by the next step: it goes out after the
        // publication so the release above does not wait for it
        if (mine && step > 0) orow[step] = (char)sel;
        if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
        ++it;
      }
    // completion: rows of every FC CTA are written before the flag is (ref:889)
    named_bar_sync(PS_EPI_BAR, PS_EPI_WARPS * 32);
    if (etid == 0 && a.done_flag) {
      __threadfence_system();
      if (atom_add_acq_rel_gpu(a.done_cnt, 1u) == (uint32_t)a.n_fc - 1u) {
        fence_acq_rel_sys();
        st_release_sys(a.done_flag, a.done_value);
      }
    }
  }
  tc_fence_before();
  cluster_sync_all();   // the partner may still be writing our seg_sum / signalling our xbar
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<PS_FC_TMEM_COLS>(tmem_base);
  }
}

__global__ void __launch_bounds__(PS_THREADS, 1)
gru_persistent_kernel(const __grid_constant__ PersistMaps maps, const PersistArgs args) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  if ((int)blockIdx.x < 2 * args.P)
    persist_main(maps, args, smem);
  else
    persist_fc(maps, args, smem);
}

// ------------------------------------------------------------- init / host
// fp32 [3][H][K] (gates r,z,n stacked) -> bf16 [H/16][3][16][K], gate order (g0,g1,g2)
__global__ void pack_gate_weights16_kernel(const float *__restrict__ src, __nv_bfloat16 *__restrict__ dst,
                                           int H, int K, int g0, int g1, int g2) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)3 * H * K) return;
  const int k = (int)(idx % K);
  const size_t prow = idx / K;
  const int blk = (int)(prow / PS_N), rem = (int)(prow % PS_N);
  const int slot = rem / PS_UNITS, u = rem % PS_UNITS;
  const int g = slot == 0 ? g0 : (slot == 1 ? g1 : g2);
  dst[idx] = __float2bfloat16_rn(src[((size_t)g * H + blk * PS_UNITS + u) * K + k]);
}

// t0p[pair][c][slot: n,r,z][16] = W_i{g}0 . emb[c] + b_i{g}0 (+ b_h{g}0 for r, z)
__global__ void pack_t0p_kernel(const float *__restrict__ t0_raw, const float *__restrict__ b_ih,
                                const float *__restrict__ b_hh, __nv_bfloat16 *__restrict__ t0p, int C, int H) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)3 * H * C) return;
  const int u = (int)(idx % PS_UNITS);
  const int slot = (int)((idx / PS_UNITS) % 3);
  const int c = (int)((idx / (3 * PS_UNITS)) % C);
  const int pair = (int)(idx / ((size_t)3 * PS_UNITS * C));
  const int g = slot == 0 ? 2 : slot - 1;
  const int i = pair * PS_UNITS + u;
  float v = t0_raw[(size_t)c * 3 * H + g * H + i] + b_ih[g * H + i];
  if (g < 2) v += b_hh[g * H + i];
  t0p[idx] = __float2bfloat16_rn(v);
}

static size_t ps_sync_words(const b200gru_model *m) {
  const int mt_max = m->bpad / kTileM;
  return (size_t)(m->cfg.num_layers * 2 + 1) * mt_max + 64 + (size_t)m->cfg.max_len * mt_max;
}

template <typename T>
static int ps_alloc(b200gru_model *m, T **out, size_t count) {
  return model_alloc(m, out, count);
}

// Decides whether the persistent kernel can run this model on this GPU and, if
// so, packs its weights.  Called from tc_prepare's caller after the
// tensor-core path is ready (it reuses the hidden buffers, their tensor maps,
// the bias blocks and W_fc).
int persist_prepare(b200gru_model *m) {
  const b200gru_config &c = m->cfg;
  const int H = c.hidden, C = c.num_char, L = c.num_layers;
  m->ps_ok = false;
  // needs: row-major hidden buffers (H a power of two), an alphabet the FC pair can halve
  if (!m->tc_ok || m->tp_world != 1 || H % 64 != 0 || (1 << m->pw_log2) != H || C % 32 != 0) return 0;
  int sms = 148, coop = 0, max_smem = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, m->device);
  cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, m->device);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, m->device);
  const int P = H / PS_UNITS;
  const int kbH = H / 64;
  const size_t w_bytes = (size_t)(2 * L - 1) * kbH * PS_KB_BYTES;
  const size_t fc_bytes = 1024 + (size_t)PS_FC_STAGES * PS_FC_STAGE_BYTES + 512 + 4096;
  int stages = (int)(((size_t)max_smem - 1024 - 1024 - w_bytes) / PS_A_BYTES);
  if ((size_t)max_smem < 1024 + 1024 + w_bytes + 2 * PS_A_BYTES) stages = 0;
  if (stages > PS_MAX_STAGES) stages = PS_MAX_STAGES;
  if (!coop || stages < 2 || 2 * P + 2 > sms || fc_bytes > (size_t)max_smem) return 0;
  const size_t main_bytes = 1024 + w_bytes + (size_t)stages * PS_A_BYTES + 1024;
  m->ps_smem = (int)(main_bytes > fc_bytes ? main_bytes : fc_bytes);
  m->ps_stages = stages;
  m->ps_pairs = P;
  m->ps_fc_max = (sms - 2 * P) & ~1;

  LayoutIndex ix(L);
  const unsigned TPB = 256;
  auto blocks = [&](size_t n) { return (unsigned)((n + TPB - 1) / TPB); };
  const size_t we = (size_t)3 * H * H;
  for (int l = 0; l < L; ++l) {
    if (l > 0) {
      if (int rc = ps_alloc(m, &m->ps_w_ih[l], we)) return rc;
      pack_gate_weights16_kernel<<<blocks(we), TPB>>>(m->P(ix.w_i(l, 0)), m->ps_w_ih[l], H, H, 2, 0, 1);
      if (int rc = make_tmap_bf16_2d(&m->ps_map_w_ih[l], m->ps_w_ih[l], 3 * (uint64_t)H, H, PS_ROWS, 64)) return rc;
    }
    if (int rc = ps_alloc(m, &m->ps_w_hh[l], we)) return rc;
    pack_gate_weights16_kernel<<<blocks(we), TPB>>>(m->P(ix.w_h(l, 0)), m->ps_w_hh[l], H, H, 0, 1, 2);
    if (int rc = make_tmap_bf16_2d(&m->ps_map_w_hh[l], m->ps_w_hh[l], 3 * (uint64_t)H, H, PS_ROWS, 64)) return rc;
  }
  if (int rc = ps_alloc(m, &m->ps_t0p, (size_t)3 * H * C)) return rc;
  pack_t0p_kernel<<<blocks((size_t)3 * H * C), TPB>>>(m->t0_raw, m->P(ix.b_i(0, 0)), m->P(ix.b_h(0, 0)),
                                                      m->ps_t0p, C, H);
  if (int rc = make_tmap_bf16_2d(&m->ps_map_wfc_half, m->w_fc, (uint64_t)C, H, C / 2, 64, m->ldh)) return rc;
  if (int rc = ps_alloc(m, &m->ps_sync, ps_sync_words(m))) return rc;
  B200_CUDA(cudaMemset(m->ps_sync, 0, ps_sync_words(m) * sizeof(uint32_t)));
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaFuncSetAttribute(gru_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 m->ps_smem));
  // all CTAs must be co-resident: ask the occupancy calculator about clusters of 2
  {
    cudaLaunchConfig_t cfg{};
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.gridDim = dim3(2 * P + 2);
    cfg.blockDim = dim3(PS_THREADS);
    cfg.dynamicSmemBytes = m->ps_smem;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int clusters = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&clusters, gru_persistent_kernel, &cfg);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return 0;
    }
    if (clusters < P + 1) return 0;
    if (2 * clusters < 2 * P + m->ps_fc_max) m->ps_fc_max = 2 * (clusters - P);
  }
  m->ps_ok = true;
  return 0;
}

int persist_generate(b200gru_model *m, int n, const b200gru_io *io, cudaStream_t s) {
  const b200gru_config &c = m->cfg;
  const int H = c.hidden, L = c.num_layers;
  if (!m->ps_ok) return fail(B200GRU_ERR_STATE, "persistent kernel is not available for this model");
  PersistMaps maps;
  PersistArgs a{};
  a.n = n;
  a.m_tiles = (n + kTileM - 1) / kTileM;
  a.m2_tiles = (a.m_tiles + 1) / 2;
  a.mt_max = m->bpad / kTileM;
  a.T = c.max_len; a.L = L; a.H = H; a.C = c.num_char; a.kbH = H / 64;
  a.P = m->ps_pairs;
  a.n_fc = 2 * a.m_tiles < m->ps_fc_max ? 2 * a.m_tiles : m->ps_fc_max;  // one CTA pair per 128-name tile
  a.stages = m->ps_stages;
  a.sos = c.sos; a.eos = c.eos;
  uint32_t off = 0;
  for (int l = 0; l < L; ++l) {
    if (l > 0) { a.off_ih[l] = off; off += (uint32_t)a.kbH * PS_KB_BYTES; }
    a.off_hh[l] = off; off += (uint32_t)a.kbH * PS_KB_BYTES;
    a.bias[l] = m->tc[l].bias;
    for (int pp = 0; pp < 2; ++pp) {
      a.h[l][pp] = m->tc[l].h[pp];
      maps.h[l][pp] = m->tc[l].map_h[pp];
    }
    maps.whh[l] = m->ps_map_w_hh[l];
    if (l > 0) maps.wih[l] = m->ps_map_w_ih[l];
  }
  a.w_bytes = off;
  maps.wfc = m->ps_map_wfc_half;
  a.t0p = m->ps_t0p;
  a.cnt_h = m->ps_sync;
  a.cnt_c = a.cnt_h + (size_t)L * 2 * a.mt_max;
  a.done_cnt = a.cnt_c + a.mt_max;
  a.live = reinterpret_cast<int32_t *>(a.done_cnt + 64);
  a.b_fc = m->b_fc;
  a.logits = m->debug_logits ? m->logits : nullptr;
  a.rfloats = io->rfloats;
  a.output = io->output;
  a.chars = m->chars;
  a.alive = m->alive;
  a.done_flag = io->done_flag;
  a.done_value = io->done_value;
  a.ready_flag = io->ready_flag;
  a.ready_value = io->ready_value;
  a.timeline = nullptr;
  a.timeline_fc_rank = 0;
  if (m->trace_on) {
    if (const char *e = getenv("B200GRU_TL_FC_RANK")) a.timeline_fc_rank = atoi(e) & 1;  // tracing also records the in-kernel timeline of two CTAs
    if (!m->ps_timeline)
      if (int rc = ps_alloc(m, &m->ps_timeline, (size_t)2 * 3 * PS_TL_EVENTS)) return rc;
    B200_CUDA(cudaMemsetAsync(m->ps_timeline, 0, sizeof(unsigned long long) * 2 * 3 * PS_TL_EVENTS, s));
    a.timeline = m->ps_timeline;
  }

  m->trace_begin(s);
  if (io->ready_event) B200_CUDA(cudaStreamWaitEvent(s, static_cast<cudaEvent_t>(io->ready_event), 0));
  B200_CUDA(cudaMemsetAsync(m->ps_sync, 0, ps_sync_words(m) * sizeof(uint32_t), s));
  cudaLaunchConfig_t cfg{};
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeCooperative;
  attr[1].val.cooperative = 1;
  cfg.gridDim = dim3(2 * a.P + a.n_fc);
  cfg.blockDim = dim3(PS_THREADS);
  cfg.dynamicSmemBytes = m->ps_smem;
  cfg.stream = s;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  B200_CUDA(cudaLaunchKernelEx(&cfg, gru_persistent_kernel, maps, a));
  m->trace_mark("persistent", s);
  m->last_launches = 1;
  return 0;
}

}  // namespace b200gru

// In-kernel timeline of the last traced persistent generate: `out` receives
// 2 x 3 x 256 words, [cta: first weight CTA, first FC CTA][role: TMA producer,
// MMA issuer, epilogue][event]; each word is (globaltimer_ns << 16) | code with
// code = (step << 8) | (layer << 4) | event, 0 = unused slot.
extern "C" int b200gru_model_persist_timeline(b200gru_model *m, unsigned long long *out, int max_words) {
  using namespace b200gru;
  B200_REQUIRE(m && out, "null argument");
  const int words = 2 * 3 * PS_TL_EVENTS;
  B200_REQUIRE(max_words >= words, "buffer must hold %d words", words);
  if (!m->ps_timeline) return 0;
  B200_CUDA(cudaMemcpy(out, m->ps_timeline, sizeof(unsigned long long) * words, cudaMemcpyDeviceToHost));
  return words;
}
