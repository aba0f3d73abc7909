// gru_tc_sm100.cu — the Blackwell hot path: batched GRU generation on
// tcgen05 tensor cores with TMA-fed operands and TMEM accumulators.
//
// What the reference does per character of ONE name — 13 matvec launches,
// 34 elementwise launches, 2 softmax launches, a D2H copy and a CPU sampler
// (ref:664-878) — becomes, for ALL of a rank's names at once, one launch per
// GRU layer plus one for FC+softmax+sample:
//
//   layer 0 :  acc[128 names, (r|z|n) x 64 units] = h0 . W_hh0^T      (UMMA)
//              + gather of T0[char] (= W_ih0.emb[char] + biases, folded at
//              init; replaces ref:669,:676,:702,:729) -> gates -> h0'
//   layer l :  acc[.., r|z|n_i] = x . W_ih^T ; acc[.., r|z] += h . W_hh^T ;
//              acc[.., n_h] = h . W_hn^T  -> gates -> h'  (ref:770-854)
//   FC      :  acc[128 names, C] = h_top . W_fc^T -> +b_fc -> softmax (full sum,
//              no max subtraction, ref:302-313) -> inverse-CDF walk in index
//              order with strict '>' (ref:322-333) -> output byte + next char.
//
// Kernel anatomy (one CTA per SM, persistent over tiles, 256 threads):
//   warp 0   TMA producer   cp.async.bulk.tensor -> 128B-swizzled smem ring
//   warp 1   MMA issuer     tcgen05.mma kind::f16, M=128, N in {192,128,64,C}
//   warp 2   TMEM allocator 512 columns = 2 accumulator stages of 256
//   warp 4-7 epilogue       tcgen05.ld -> gate math in registers -> global
// The accumulator is double-buffered in TMEM so the epilogue of tile i
// overlaps the MMAs of tile i+1.
//
// Weights are repacked once at init so that a 64-unit block's r, z and n rows
// are contiguous: one TMA box / one UMMA covers all three gates and the gate
// math for a (name, unit) pair is local to one thread.
#include "model.hpp"
#include "sm100_ptx.cuh"

namespace b200gru {
using namespace sm100;

enum : int { MODE_L0 = 0, MODE_LN = 1, MODE_FC = 2 };

constexpr int TC_THREADS = 256;
constexpr int TC_BLOCK_K = 64;               // 128 bytes of bf16 per row
constexpr int TC_A_BYTES = kTileM * 128;     // 16 KB
constexpr int TC_TMEM_COLS = 512;
constexpr int TC_ACC_STRIDE = 256;           // TMEM columns per accumulator

struct TcArgs {
  int n;            // valid names (rows)
  int m_tiles;      // ceil(n / 128)
  int n_blocks;     // H / 64 (1 for FC)
  int kb_x;         // k-blocks (of 64) of the x operand  (0: none)
  int kb_h;         // k-blocks of the recurrent operand  (0: h == 0, step 0)
  int H, C, T, step, eos;
  const float *t0;          // [C, H/64, 3, 64]           (L0)
  const float *bias;        // [H/64, 4, 64]              (L0: only q=3, LN: all)
  const int32_t *chars;     // [bpad]                     (L0 gather index)
  const __nv_bfloat16 *h_prev;  // [bpad, H]
  __nv_bfloat16 *h_out;         // [bpad, H]
  // FC
  const float *b_fc;
  float *logits;            // optional [bpad, C] tap
  const float *rfloats;     // [n, T]   (may be a peer address)
  char *output;             // [n, T+1] (may be a peer address)
  int32_t *chars_out;
  uint8_t *alive;
};

template <int MODE>
struct TcCfg {
  static constexpr int B_ROWS_MAX = (MODE == MODE_FC) ? 256 : 192;
  static constexpr int B_BYTES_MAX = B_ROWS_MAX * 128;
  static constexpr int STAGE_BYTES = TC_A_BYTES + B_BYTES_MAX;
  static constexpr int STAGES = (MODE == MODE_FC) ? 4 : 5;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256;
};

template <int MODE>
__global__ void __launch_bounds__(TC_THREADS, 1)
gru_tc_kernel(const __grid_constant__ CUtensorMap map_x,
              const __grid_constant__ CUtensorMap map_h,
              const __grid_constant__ CUtensorMap map_wx,
              const __grid_constant__ CUtensorMap map_wh, const TcArgs args) {
  using Cfg = TcCfg<MODE>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t *full_bar = bars;                  // [STAGES]
  uint64_t *empty_bar = bars + STAGES;        // [STAGES]
  uint64_t *tmem_full = bars + 2 * STAGES;    // [2]
  uint64_t *tmem_empty = bars + 2 * STAGES + 2;  // [2]
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_tiles = args.m_tiles * args.n_blocks;
  const int b_rows = (MODE == MODE_FC) ? args.C : 192;
  const uint32_t b_bytes = (uint32_t)b_rows * 128u;
  const int kb_total = args.kb_x + args.kb_h;

  if (warp == 0 && lane == 0) {
    if (args.kb_x) { tma_prefetch_desc(&map_x); tma_prefetch_desc(&map_wx); }
    if (args.kb_h) { tma_prefetch_desc(&map_h); tma_prefetch_desc(&map_wh); }
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 4);  // one arrive per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<TC_TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================================================== TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int m_tile = tile / args.n_blocks;
        const int n_blk = tile % args.n_blocks;
        for (int kb = 0; kb < kb_total; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t *sa = smem + stage * Cfg::STAGE_BYTES;
          uint8_t *sb = sa + TC_A_BYTES;
          mbar_expect_tx(&full_bar[stage], TC_A_BYTES + b_bytes);
          const bool is_x = kb < args.kb_x;
          const int k = (is_x ? kb : kb - args.kb_x) * TC_BLOCK_K;
          tma_load_2d(sa, is_x ? &map_x : &map_h, &full_bar[stage], k,
                      m_tile * kTileM);
          tma_load_2d(sb, is_x ? &map_wx : &map_wh, &full_bar[stage], k,
                      n_blk * b_rows);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================================= MMA issuer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int acc_stage = 0;
      uint32_t acc_phase = 0;
      const uint32_t idesc_full = umma_idesc_bf16(kTileM, (uint32_t)b_rows);
      const uint32_t idesc_rz = umma_idesc_bf16(kTileM, 128);
      const uint32_t idesc_n = umma_idesc_bf16(kTileM, 64);
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty[acc_stage], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d = tmem_base + acc_stage * TC_ACC_STRIDE;
        for (int kb = 0; kb < kb_total; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * Cfg::STAGE_BYTES);
          const uint32_t sb = sa + TC_A_BYTES;
          const uint64_t da = umma_desc_sw128(sa);
          const uint64_t db = umma_desc_sw128(sb);
          const bool is_x = kb < args.kb_x;
          const int kh = kb - args.kb_x;
#pragma unroll
          for (int k = 0; k < TC_BLOCK_K / 16; ++k) {
            const uint64_t koff = (uint64_t)(k * 32 >> 4);  // 32 B per UMMA_K
            if (MODE != MODE_LN || is_x) {
              // one MMA over every gate column of this unit block
              umma_bf16(d, da + koff, db + koff, idesc_full,
                        (kb | k) != 0 ? 1u : 0u);
            } else {
              // recurrent half of a stacked layer: r,z accumulate onto the
              // x-projection, n_h goes to its own columns (r multiplies it)
              umma_bf16(d, da + koff, db + koff, idesc_rz,
                        (args.kb_x | kh | k) != 0 ? 1u : 0u);
              umma_bf16(d + 192, da + koff, db + koff + (128 * 128 >> 4),
                        idesc_n, (kh | k) != 0 ? 1u : 0u);
            }
          }
          umma_commit(&empty_bar[stage]);  // smem slot free when MMAs retire
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tmem_full[acc_stage]);  // accumulator ready
        if (++acc_stage == 2) { acc_stage = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ========================================================= epilogue
    const int quarter = warp & 3;  // TMEM lane quarter this warp may touch
    int acc_stage = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int m_tile = tile / args.n_blocks;
      const int n_blk = tile % args.n_blocks;
      const int row = m_tile * kTileM + quarter * 32 + lane;
      const bool valid = row < args.n;
      mbar_wait(&tmem_full[acc_stage], acc_phase);
      tc_fence_after();
      const uint32_t t =
          tmem_base + ((uint32_t)(quarter * 32) << 16) + acc_stage * TC_ACC_STRIDE;

      if constexpr (MODE == MODE_L0 || MODE == MODE_LN) {
        const bool has_acc = kb_total > 0;
        const bool has_h = args.kb_h > 0;
        const float *bias = args.bias + (size_t)n_blk * 4 * 64;
        const float *t0row = nullptr;
        if (MODE == MODE_L0)
          t0row = args.t0 +
                  ((size_t)args.chars[row] * args.n_blocks + n_blk) * 3 * 64;
        const size_t hoff = (size_t)row * args.H + (size_t)n_blk * 64;
#pragma unroll 1
        for (int ch = 0; ch < 4; ++ch) {
          uint32_t vr[16], vz[16], vn[16], vh[16];
          if (has_acc) {
            tmem_ld16(t + ch * 16, vr);
            tmem_ld16(t + 64 + ch * 16, vz);
            tmem_ld16(t + 128 + ch * 16, vn);
            if (MODE == MODE_LN && has_h) tmem_ld16(t + 192 + ch * 16, vh);
            tmem_ld_wait();
          }
          if (!has_acc) {
#pragma unroll
            for (int i = 0; i < 16; ++i) vr[i] = vz[i] = vn[i] = 0u;
          }
          if (!(MODE == MODE_LN && has_h)) {
#pragma unroll
            for (int i = 0; i < 16; ++i) vh[i] = 0u;
          }
          // previous hidden, 16 bf16 = 32 bytes
          float hp[16];
          if (has_h) {
            const uint4 *src =
                reinterpret_cast<const uint4 *>(args.h_prev + hoff + ch * 16);
            uint4 q0 = src[0], q1 = src[1];
            const uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              hp[2 * i] = __uint_as_float(w[i] << 16);
              hp[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) hp[i] = 0.0f;
          }
          uint32_t packed[8];
#pragma unroll
          for (int i = 0; i < 16; i += 2) {
            float hn[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int u = ch * 16 + i + e;
              float ar = __uint_as_float(vr[i + e]);
              float az = __uint_as_float(vz[i + e]);
              float an = __uint_as_float(vn[i + e]);
              float r, z, nn;
              if (MODE == MODE_L0) {
                // T0 already holds W_i.x + b_i (+ b_hr / b_hz)
                r = sigmoid_approx(ar + __ldg(t0row + u));
                z = sigmoid_approx(az + __ldg(t0row + 64 + u));
                nn = tanh_approx(
                    fmaf(r, an + __ldg(bias + 3 * 64 + u), __ldg(t0row + 128 + u)));
              } else {
                float ah = __uint_as_float(vh[i + e]);
                r = sigmoid_approx(ar + __ldg(bias + u));
                z = sigmoid_approx(az + __ldg(bias + 64 + u));
                nn = tanh_approx(fmaf(r, ah + __ldg(bias + 3 * 64 + u),
                                      an + __ldg(bias + 2 * 64 + u)));
              }
              // h' = (1-z)*n + z*h  ==  n + z*(h - n)     (ref:752-763)
              hn[e] = fmaf(z, hp[i + e] - nn, nn);
            }
            __nv_bfloat162 p2 = __floats2bfloat162_rn(hn[0], hn[1]);
            packed[i / 2] = *reinterpret_cast<uint32_t *>(&p2);
          }
          if (valid) {
            uint4 *dst = reinterpret_cast<uint4 *>(args.h_out + hoff + ch * 16);
            dst[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
            dst[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
          }
        }
      } else {
        // ---------------------------------- FC + softmax + sample (per name)
        const int C = args.C;
        const bool live = valid && args.alive[row] != 0;
        const float kLog2e = 1.4426950408889634f;
        float sum = 0.0f;
#pragma unroll 1
        for (int c0 = 0; c0 < C; c0 += 16) {
          uint32_t v[16];
          tmem_ld16(t + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float x = __uint_as_float(v[i]) + __ldg(args.b_fc + c0 + i);
            sum += ex2_approx(x * kLog2e);
            if (args.logits && valid)
              args.logits[(size_t)row * C + c0 + i] = live ? x : 0.0f;
          }
        }
        const float inv = 1.0f / sum;
        const float r =
            live ? args.rfloats[(size_t)row * args.T + args.step] : 2.0f;
        float psum = 0.0f;
        int sel = C - 1;
        bool found = false;
#pragma unroll 1
        for (int c0 = 0; c0 < C; c0 += 16) {
          // warp-uniform early exit once every lane has its character
          if (__all_sync(0xffffffffu, found || !live)) break;
          uint32_t v[16];
          tmem_ld16(t + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float x = __uint_as_float(v[i]) + __ldg(args.b_fc + c0 + i);
            psum += ex2_approx(x * kLog2e) * inv;
            if (!found && psum > r) {
              sel = c0 + i;
              found = true;
            }
          }
        }
        if (valid) {
          char *orow = args.output + (size_t)row * (args.T + 1);
          if (args.step == 0)
            for (int l = 1; l <= args.T; ++l) orow[l] = 0;  // NUL pad, ref:643
          if (live) {
            orow[args.step] = (char)sel;  // EOS id is written too, ref:877
            args.chars_out[row] = sel;
            if (sel == args.eos) args.alive[row] = 0;  // ref:881
          }
        }
      }
      // accumulator drained: hand the TMEM stage back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc_stage]);
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
// fp32 [3][H][K] (gate-stacked, as in the file) -> bf16 [H/64][3][64][K]
__global__ void pack_gate_weights_kernel(const float *__restrict__ src,
                                         __nv_bfloat16 *__restrict__ dst, int H,
                                         int K) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t total = (size_t)3 * H * K;
  if (idx >= total) return;
  int k = (int)(idx % K);
  size_t prow = idx / K;  // packed row
  int jb = (int)(prow / 192), rem = (int)(prow % 192);
  int g = rem / 64, u = rem % 64;
  dst[idx] = __float2bfloat16_rn(src[((size_t)g * H + jb * 64 + u) * K + k]);
}

__global__ void to_bf16_kernel(const float *__restrict__ src,
                               __nv_bfloat16 *__restrict__ dst, size_t n) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < n) dst[idx] = __float2bfloat16_rn(src[idx]);
}

// t0_tc[c][jb][g][u] = W_i{g}0 . emb[c] + b_i{g}0 (+ b_h{g}0 for r,z)
__global__ void pack_t0_kernel(const float *__restrict__ t0_raw,
                               const float *__restrict__ b_ih,
                               const float *__restrict__ b_hh,
                               float *__restrict__ t0_tc, int C, int H) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)C * 3 * H) return;
  int c = (int)(idx / (3 * H));
  int rem = (int)(idx % (3 * H));
  int jb = rem / 192, g = (rem % 192) / 64, u = rem % 64;
  int i = jb * 64 + u;
  float v = t0_raw[(size_t)c * 3 * H + g * H + i] + b_ih[g * H + i];
  if (g < 2) v += b_hh[g * H + i];
  t0_tc[idx] = v;
}

// bias[jb][q][u]: q=0 b_ir+b_hr, q=1 b_iz+b_hz, q=2 b_in, q=3 b_hn
__global__ void pack_bias_kernel(const float *__restrict__ b_ih,
                                 const float *__restrict__ b_hh,
                                 float *__restrict__ out, int H) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= 4 * H) return;
  int jb = idx / 256, q = (idx % 256) / 64, u = idx % 64;
  int i = jb * 64 + u;
  float v;
  if (q == 0) v = b_ih[i] + b_hh[i];
  else if (q == 1) v = b_ih[H + i] + b_hh[H + i];
  else if (q == 2) v = b_ih[2 * H + i];
  else v = b_hh[2 * H + i];
  out[idx] = v;
}

__global__ void tc_init_kernel(int32_t *chars, uint8_t *alive, int n, int bpad,
                               int sos) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < bpad) {
    chars[i] = sos;       // ref:652
    alive[i] = i < n;
  }
}

__global__ void tc_signal_done_kernel(uint32_t *flag, uint32_t value) {
  __threadfence_system();
  st_release_sys(flag, value);
}

template <typename T>
static int tc_alloc(b200gru_model *m, T **out, size_t count) {
  void *p = nullptr;
  size_t bytes = count * sizeof(T);
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess)
    return fail(B200GRU_ERR_CUDA, "cudaMalloc(%zu bytes) failed: %s", bytes,
                cudaGetErrorString(e));
  m->allocs.push_back(p);
  m->device_bytes += (int64_t)bytes;
  *out = static_cast<T *>(p);
  return 0;
}

int tc_prepare(b200gru_model *m) {
  const b200gru_config &c = m->cfg;
  const int H = c.hidden, C = c.num_char, L = c.num_layers;
  LayoutIndex ix(L);
  const unsigned TPB = 256;
  auto blocks = [&](size_t n) { return (unsigned)((n + TPB - 1) / TPB); };

  if (int rc = tc_alloc(m, &m->t0_tc, (size_t)C * 3 * H)) return rc;
  pack_t0_kernel<<<blocks((size_t)C * 3 * H), TPB>>>(
      m->t0_raw, m->P(ix.b_i(0, 0)), m->P(ix.b_h(0, 0)), m->t0_tc, C, H);

  for (int l = 0; l < L; ++l) {
    TcLayer &t = m->tc[l];
    const size_t wn = (size_t)3 * H * H;
    if (l > 0) {
      if (int rc = tc_alloc(m, &t.w_ih, wn)) return rc;
      pack_gate_weights_kernel<<<blocks(wn), TPB>>>(m->P(ix.w_i(l, 0)), t.w_ih,
                                                    H, H);
      if (int rc = make_tmap_bf16_2d(&t.map_w_ih, t.w_ih, 3 * (uint64_t)H, H,
                                     192, 64))
        return rc;
    }
    if (int rc = tc_alloc(m, &t.w_hh, wn)) return rc;
    pack_gate_weights_kernel<<<blocks(wn), TPB>>>(m->P(ix.w_h(l, 0)), t.w_hh, H,
                                                  H);
    if (int rc =
            make_tmap_bf16_2d(&t.map_w_hh, t.w_hh, 3 * (uint64_t)H, H, 192, 64))
      return rc;
    if (int rc = tc_alloc(m, &t.bias, (size_t)4 * H)) return rc;
    pack_bias_kernel<<<blocks((size_t)4 * H), TPB>>>(
        m->P(ix.b_i(l, 0)), m->P(ix.b_h(l, 0)), t.bias, H);
    for (int pp = 0; pp < 2; ++pp) {
      if (int rc = tc_alloc(m, &t.h[pp], (size_t)m->bpad * H)) return rc;
      B200_CUDA(cudaMemset(t.h[pp], 0, sizeof(__nv_bfloat16) * (size_t)m->bpad * H));
      if (int rc = make_tmap_bf16_2d(&t.map_h[pp], t.h[pp], (uint64_t)m->bpad, H,
                                     kTileM, 64))
        return rc;
    }
  }
  if (int rc = tc_alloc(m, &m->w_fc, (size_t)C * H)) return rc;
  to_bf16_kernel<<<blocks((size_t)C * H), TPB>>>(m->P(ix.w_fc()), m->w_fc,
                                                 (size_t)C * H);
  if (int rc = make_tmap_bf16_2d(&m->map_w_fc, m->w_fc, (uint64_t)C, H, C, 64))
    return rc;
  B200_CUDA(cudaGetLastError());

  B200_CUDA(cudaFuncSetAttribute(gru_tc_kernel<MODE_L0>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 TcCfg<MODE_L0>::SMEM_BYTES));
  B200_CUDA(cudaFuncSetAttribute(gru_tc_kernel<MODE_LN>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 TcCfg<MODE_LN>::SMEM_BYTES));
  B200_CUDA(cudaFuncSetAttribute(gru_tc_kernel<MODE_FC>,
                                 cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 TcCfg<MODE_FC>::SMEM_BYTES));
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

int tc_generate(b200gru_model *m, int n, const b200gru_io *io, cudaStream_t s) {
  const b200gru_config &c = m->cfg;
  const int H = c.hidden, C = c.num_char, T = c.max_len, L = c.num_layers;
  const int sms = num_sms(m->device);
  int launches = 0;
  const int m_tiles = (n + kTileM - 1) / kTileM;
  const int rows = m_tiles * kTileM;
  tc_init_kernel<<<(rows + 255) / 256, 256, 0, s>>>(m->chars, m->alive, n, rows,
                                                    c.sos);
  ++launches;
  for (int step = 0; step < T; ++step) {
    const int rd = step & 1, wr = rd ^ 1;  // ping-pong hidden buffers
    for (int l = 0; l < L; ++l) {
      TcLayer &t = m->tc[l];
      TcArgs a{};
      a.n = n; a.m_tiles = m_tiles; a.n_blocks = H / kUnitBlock;
      a.kb_x = (l == 0) ? 0 : H / TC_BLOCK_K;
      a.kb_h = (step == 0) ? 0 : H / TC_BLOCK_K;  // h == 0 at step 0 (ref:653-654)
      a.H = H; a.C = C; a.T = T; a.step = step; a.eos = c.eos;
      a.t0 = m->t0_tc; a.bias = t.bias; a.chars = m->chars;
      a.h_prev = t.h[rd]; a.h_out = t.h[wr];
      const int tiles = a.m_tiles * a.n_blocks;
      const int grid = tiles < sms ? tiles : sms;
      if (l == 0) {
        gru_tc_kernel<MODE_L0><<<grid, TC_THREADS, TcCfg<MODE_L0>::SMEM_BYTES, s>>>(
            t.map_h[rd], t.map_h[rd], t.map_w_hh, t.map_w_hh, a);
      } else {
        gru_tc_kernel<MODE_LN><<<grid, TC_THREADS, TcCfg<MODE_LN>::SMEM_BYTES, s>>>(
            m->tc[l - 1].map_h[wr], t.map_h[rd], t.map_w_ih, t.map_w_hh, a);
      }
      ++launches;
    }
    TcArgs a{};
    a.n = n; a.m_tiles = m_tiles; a.n_blocks = 1;
    a.kb_x = H / TC_BLOCK_K; a.kb_h = 0;
    a.H = H; a.C = C; a.T = T; a.step = step; a.eos = c.eos;
    a.b_fc = m->b_fc; a.logits = m->logits;
    a.rfloats = io->rfloats; a.output = io->output;
    a.chars_out = m->chars; a.alive = m->alive;
    const int grid = m_tiles < sms ? m_tiles : sms;
    const CUtensorMap &mx = m->tc[L - 1].map_h[wr];
    gru_tc_kernel<MODE_FC><<<grid, TC_THREADS, TcCfg<MODE_FC>::SMEM_BYTES, s>>>(
        mx, mx, m->map_w_fc, m->map_w_fc, a);
    ++launches;
  }
  if (io->done_flag) {
    tc_signal_done_kernel<<<1, 1, 0, s>>>(io->done_flag, io->done_value);
    ++launches;
  }
  B200_CUDA(cudaGetLastError());
  m->last_launches = launches;
  return 0;
}

}  // namespace b200gru

// Plain tcgen05 GEMM through the same pipeline (FC mode with the sampling
// epilogue replaced by a store) — used by tests to validate the TMA / UMMA
// descriptor plumbing in isolation.
namespace b200gru {
__global__ void __launch_bounds__(TC_THREADS, 1)
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
  gemm_probe_kernel<<<grid, TC_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(
      ma, mb, d, m, n, k, n_tile);
  B200_CUDA(cudaGetLastError());
  return 0;
}
