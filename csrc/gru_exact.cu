// gru_exact.cu — fp32 SIMT batched generator: the correctness anchor.
//
// Replaces the reference's per-character sequence of 51 launches on one name
// (ref:664-873) by batched kernels over all of a rank's names, but keeps every
// output element's arithmetic *identical* to the reference's association
// order, so sampled names match Oracle A (csrc/cpu_oracle.cpp) bit for bit:
//   * each GEMM output is one accumulator updated left-to-right over K with
//     fused multiply-add, exactly like `matvec` (ref:233-242);
//   * gate math uses explicitly rounded adds/multiplies in the order of the
//     elementwise launches (ref:686-763);
//   * softmax is full-sum-then-divide in index order without max subtraction
//     (the commented CPU body, ref:302-313 — not the racy kernel ref:294-300);
//   * sampling is the inverse-CDF walk of random_select (ref:322-333), on the
//     device, so there is no per-character D2H copy (ref:873).
#include "exact_math.h"
#include "model.hpp"
#include "sm100_ptx.cuh"

namespace b200gru {

// ---------------------------------------------------------------- SGEMM (TN)
// C[m,n] = sum_k A[m,k] * W[n,k]; both operands K-contiguous.  Classic
// 64x64x16 smem tiling, 4x4 register micro-tile, one accumulator per output,
// k strictly ascending => same rounding sequence as the oracle's dot_lr().
constexpr int SG_BM = 64, SG_BN = 64, SG_BK = 16, SG_T = 256;

__global__ void __launch_bounds__(SG_T)
sgemm_tn_kernel(const float *__restrict__ A, const float *__restrict__ W,
                float *__restrict__ C, int M, int N, int K) {
  __shared__ float As[SG_BK][SG_BM + 4];
  __shared__ float Ws[SG_BK][SG_BN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * SG_BM, n0 = blockIdx.x * SG_BN;
  const int tx = tid % 16, ty = tid / 16;  // 16x16 threads, 4x4 outputs each
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

  const int lrow = tid / 4;        // 0..63
  const int lk = (tid % 4) * 4;    // 0,4,8,12
  // float4 loads need 16-byte aligned rows: K % 4 == 0 AND aligned base pointers
  // (operands are often `params + offset`, and offsets such as num_char*emb_dim
  // need not be multiples of 4 floats)
  const bool vec_ok = (K % 4 == 0) &&
                      (((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(W)) & 15) == 0);
  for (int k0 = 0; k0 < K; k0 += SG_BK) {
    // load A tile (64 x 16) and W tile (64 x 16), transposed into smem
    {
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      int r = m0 + lrow;
      if (r < M) {
        const float *src = A + (size_t)r * K + k0 + lk;
        if (vec_ok && k0 + lk + 3 < K) {
          float4 t = *reinterpret_cast<const float4 *>(src);
          v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
        } else {
          for (int q = 0; q < 4; ++q)
            if (k0 + lk + q < K) v[q] = src[q];
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) As[lk + q][lrow] = v[q];
    }
    {
      float v[4] = {0.f, 0.f, 0.f, 0.f};
      int r = n0 + lrow;
      if (r < N) {
        const float *src = W + (size_t)r * K + k0 + lk;
        if (vec_ok && k0 + lk + 3 < K) {
          float4 t = *reinterpret_cast<const float4 *>(src);
          v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
        } else {
          for (int q = 0; q < 4; ++q)
            if (k0 + lk + q < K) v[q] = src[q];
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) Ws[lk + q][lrow] = v[q];
    }
    __syncthreads();
    const int kend = min(SG_BK, K - k0);  // no padded FMAs: keeps -0.0 intact
    for (int kk = 0; kk < kend; ++kk) {
      float a[4], w[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) w[j] = Ws[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
          acc[i][j] = __fmaf_rn(w[j], a[i], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int r = m0 + ty * 4 + i;
    if (r >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int c = n0 + tx * 4 + j;
      if (c < N) C[(size_t)r * N + c] = acc[i][j];
    }
  }
}

int exact_sgemm_tn(const float *a, const float *w, float *c, int m, int n,
                   int k, cudaStream_t s) {
  if (m <= 0 || n <= 0) return 0;
  dim3 grid((n + SG_BN - 1) / SG_BN, (m + SG_BM - 1) / SG_BM);
  sgemm_tn_kernel<<<grid, SG_T, 0, s>>>(a, w, c, m, n, k);
  B200_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------ kernels
// Reset per-name state (ref:652-659 does this with two H2D copies per name).
__global__ void exact_init_kernel(float *h32, size_t h_elems, int32_t *chars,
                                  uint8_t *alive, int n, int sos) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t j = i; j < h_elems; j += stride) h32[j] = 0.0f;
  for (size_t j = i; j < (size_t)n; j += stride) {
    chars[j] = sos;
    alive[j] = 1;
  }
}

// Gate math of one layer for all names (ref:686-763 / 782-854).  gi is either
// the layer-0 table row of the name's current character or the W_ih GEMM
// output; gh is the W_hh GEMM output.
__global__ void exact_gates_kernel(const float *__restrict__ gi_table,
                                   const int32_t *__restrict__ chars,
                                   const float *__restrict__ gi_gemm,
                                   const float *__restrict__ gh,
                                   const float *__restrict__ b_ih,
                                   const float *__restrict__ b_hh,
                                   float *__restrict__ h, int n, int H) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)n * H) return;
  int b = (int)(idx / H), i = (int)(idx % H);
  B200_DASSERT(!gi_table || chars[b] >= 0, "negative character id %d in row %d", gi_table ? chars[b] : 0, b);
  const float *gi = gi_table ? gi_table + (size_t)chars[b] * 3 * H
                             : gi_gemm + (size_t)b * 3 * H;
  const float *ghb = gh + (size_t)b * 3 * H;
  float r = b200_sigmoidf(__fadd_rn(
      __fadd_rn(__fadd_rn(gi[i], b_ih[i]), ghb[i]), b_hh[i]));
  float z = b200_sigmoidf(__fadd_rn(
      __fadd_rn(__fadd_rn(gi[H + i], b_ih[H + i]), ghb[H + i]), b_hh[H + i]));
  float nn = b200_tanhf(__fadd_rn(
      __fadd_rn(gi[2 * H + i], b_ih[2 * H + i]),
      __fmul_rn(r, __fadd_rn(ghb[2 * H + i], b_hh[2 * H + i]))));
  float hp = h[idx];
  float a = __fadd_rn(1.0f, -z);
  h[idx] = __fadd_rn(__fmul_rn(a, nn), __fmul_rn(z, hp));
}

// FC bias + softmax + inverse-CDF sample + output write, one thread per name
// (the index-order sum and CDF walk are inherently sequential per name).
// rfloats/output may be peer addresses (MPI_Scatter/Gather fold, ref:636,:889).
__global__ void exact_sample_kernel(float *__restrict__ logits,
                                    const float *__restrict__ b_fc,
                                    const float *__restrict__ rfloats,
                                    char *__restrict__ output,
                                    int32_t *__restrict__ chars,
                                    uint8_t *__restrict__ alive, int n, int C,
                                    int T, int step, int eos, int name0) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n) return;
  float *f = logits + (size_t)b * C;
  char *row = output + (size_t)(name0 + b) * (T + 1);
  if (step == 0)
    for (int l = 1; l <= T; ++l) row[l] = 0;  // NUL padding (ref:643)
  if (!alive[b]) {
    for (int i = 0; i < C; ++i) f[i] = 0.0f;  // oracle logits are 0 after EOS
    return;
  }
  float sum = 0.0f;
  for (int i = 0; i < C; ++i) {
    float x = __fadd_rn(f[i], b_fc[i]);
    f[i] = x;
    sum = __fadd_rn(sum, b200_expf(x));
  }
  float r = rfloats[(size_t)(name0 + b) * T + step];
  float psum = 0.0f;
  int sel = C - 1;
  for (int i = 0; i < C; ++i) {
    psum = __fadd_rn(psum, __fdiv_rn(b200_expf(f[i]), sum));
    if (psum > r) {
      sel = i;
      break;
    }
  }
  B200_DASSERT(sel >= 0 && sel < C && step < T, "sampled character %d of %d at step %d of %d", sel, C, step, T);
  row[step] = (char)sel;
  chars[b] = sel;
  if (sel == eos) alive[b] = 0;
}

// One batched GRU layer step with caller-chosen scratch: the two gate GEMMs
// (ref:676-681 etc.) into scratch, then the elementwise gate sequence
// (ref:686-763).  scratch holds 2 * batch * 3 * hidden floats.
int gru_cell_fp32_run(const float *x, const float *h_in, const float *w_ih,
                      const float *w_hh, const float *b_ih, const float *b_hh,
                      float *h_out, int batch, int in_dim, int hidden,
                      float *scratch, cudaStream_t s) {
  float *g_ih = scratch;
  float *g_hh = scratch + (size_t)batch * 3 * hidden;
  if (int rc = exact_sgemm_tn(x, w_ih, g_ih, batch, 3 * hidden, in_dim, s))
    return rc;
  if (int rc = exact_sgemm_tn(h_in, w_hh, g_hh, batch, 3 * hidden, hidden, s))
    return rc;
  if (h_out != h_in)
    B200_CUDA(cudaMemcpyAsync(h_out, h_in, sizeof(float) * (size_t)batch * hidden,
                              cudaMemcpyDeviceToDevice, s));
  size_t tot = (size_t)batch * hidden;
  exact_gates_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(
      nullptr, nullptr, g_ih, g_hh, b_ih, b_hh, h_out, batch, hidden);
  B200_CUDA(cudaGetLastError());
  return 0;
}

__global__ void signal_done_kernel(uint32_t *flag, uint32_t value) {
  __threadfence_system();
  sm100::st_release_sys(flag, value);
}

int exact_generate(b200gru_model *m, int n, const b200gru_io *io,
                   cudaStream_t s) {
  const b200gru_config &c = m->cfg;
  const int H = c.hidden, C = c.num_char, T = c.max_len, L = c.num_layers;
  LayoutIndex ix(L);
  int launches = 0;
  for (int base = 0; base < n; base += m->chunk) {
    const int nb = min(m->chunk, n - base);
    exact_init_kernel<<<296, 256, 0, s>>>(m->h32, (size_t)L * m->chunk * H,
                                         m->chars, m->alive, nb, c.sos);
    ++launches;
    for (int step = 0; step < T; ++step) {
      const float *x = nullptr;
      for (int l = 0; l < L; ++l) {
        float *hl = m->h32 + (size_t)l * m->chunk * H;
        if (l > 0) {
          if (int rc = exact_sgemm_tn(x, m->P(ix.w_i(l, 0)), m->g_ih, nb, 3 * H,
                                      H, s))
            return rc;
          ++launches;
        }
        if (int rc = exact_sgemm_tn(hl, m->P(ix.w_h(l, 0)), m->g_hh, nb, 3 * H,
                                    H, s))
          return rc;
        ++launches;
        size_t tot = (size_t)nb * H;
        exact_gates_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(
            l == 0 ? m->t0_raw : nullptr, m->chars, m->g_ih, m->g_hh,
            m->P(ix.b_i(l, 0)), m->P(ix.b_h(l, 0)), hl, nb, H);
        ++launches;
        x = hl;
      }
      if (int rc = exact_sgemm_tn(x, m->P(ix.w_fc()), m->logits, nb, C, H, s))
        return rc;
      ++launches;
      if (base == 0 && step == 0 && io->ready_event)
        B200_CUDA(cudaStreamWaitEvent(s, static_cast<cudaEvent_t>(io->ready_event), 0));
      if (base == 0 && step == 0 && io->ready_flag) {  // the floats are first needed here
        if (int rc = b200gru_wait_flags(io->ready_flag, 1, io->ready_value, s)) return rc;
        ++launches;
      }
      exact_sample_kernel<<<(nb + 127) / 128, 128, 0, s>>>(
          m->logits, m->b_fc, io->rfloats, io->output, m->chars, m->alive, nb,
          C, T, step, c.eos, base);
      ++launches;
    }
  }
  if (io->done_flag) {
    signal_done_kernel<<<1, 1, 0, s>>>(io->done_flag, io->done_value);
    ++launches;
  }
  B200_CUDA(cudaGetLastError());
  m->last_launches = launches;
  return 0;
}

}  // namespace b200gru

// Stand-alone op: one GRU layer step for a batch, fp32, reference association
// order (bit-identical to b200gru_oracle_gru_cell per row).  The two GEMM
// outputs need 2*batch*3*hidden floats of scratch, allocated here per call.
extern "C" int b200gru_gru_cell_fp32(const float *x, const float *h_in,
                                     const float *w_ih, const float *w_hh,
                                     const float *b_ih, const float *b_hh,
                                     float *h_out, int batch, int in_dim,
                                     int hidden, void *stream) {
  using namespace b200gru;
  B200_REQUIRE(x && h_in && w_ih && w_hh && b_ih && b_hh && h_out, "null argument");
  B200_REQUIRE(batch > 0 && in_dim > 0 && hidden > 0, "bad dims");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  float *scratch = nullptr;
  B200_CUDA(cudaMalloc(&scratch, sizeof(float) * 2 * (size_t)batch * 3 * hidden));
  int rc = gru_cell_fp32_run(x, h_in, w_ih, w_hh, b_ih, b_hh, h_out, batch,
                             in_dim, hidden, scratch, s);
  cudaError_t e = cudaStreamSynchronize(s);
  cudaFree(scratch);
  if (rc) return rc;
  if (e != cudaSuccess)
    return fail(B200GRU_ERR_CUDA, "gru_cell_fp32: %s", cudaGetErrorString(e));
  return 0;
}
