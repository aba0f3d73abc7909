/*
 * namegen_demo.c — the library from plain C, no CUDA headers needed.
 *
 * Plays the role of the reference's missing main.cpp (namegensf.cu:21, :625):
 * make weights and random floats, initialize, generate N names, finalize.
 * Falls back to the CPU oracle when no GPU is usable.
 *
 *   gcc -O2 -I include examples/namegen_demo.c -L gru_mpi_cuda_b200/lib \
 *       -lb200gru -Wl,-rpath,$PWD/gru_mpi_cuda_b200/lib -o namegen_demo
 *   ./namegen_demo [n_names] [exact|tc] [out.bin]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "b200gru.h"

#define CHECK(call)                                                        \
  do {                                                                     \
    int rc__ = (call);                                                     \
    if (rc__ < 0) {                                                        \
      fprintf(stderr, "%s failed (%d): %s\n", #call, rc__, b200gru_last_error()); \
      return 1;                                                            \
    }                                                                      \
  } while (0)

int main(int argc, char **argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 8;
  const int mode = (argc > 2 && strcmp(argv[2], "exact") == 0) ? B200GRU_MODE_EXACT : B200GRU_MODE_TC;
  b200gru_config cfg = {64, 64, 128, 2, 6, 0, 1}; /* small shape; see GRUConfig.tiny() */
  const int T = cfg.max_len;

  const int64_t n_params = b200gru_param_count(&cfg);
  float *params = (float *)malloc(sizeof(float) * (size_t)n_params);
  float *rf = (float *)malloc(sizeof(float) * (size_t)n * T);
  char *out = (char *)calloc((size_t)n, (size_t)T + 1);
  if (!params || !rf || !out) return 1;
  CHECK(b200gru_random_init(&cfg, 11, params, n_params));
  CHECK(b200gru_random_floats(5, rf, (int64_t)n * T));

  if (b200gru_device_count() == 0) {
    printf("no GPU: CPU oracle\n");
    CHECK(b200gru_oracle_generate(&cfg, params, n, rf, out, NULL, 0));
  } else {
    b200gru_model *m = NULL;
    void *d_rf = NULL, *d_out = NULL;
    unsigned char handle[64];
    CHECK(b200gru_model_create(&cfg, params, 0, n, &m));          /* namegen_initialize */
    CHECK(b200gru_symm_alloc(0, sizeof(float) * (size_t)n * T, &d_rf, handle));
    CHECK(b200gru_symm_alloc(0, (size_t)n * (T + 1), &d_out, handle));
    CHECK(b200gru_memcpy_async(d_rf, rf, sizeof(float) * (size_t)n * T, 1, NULL));
    b200gru_io io;
    memset(&io, 0, sizeof(io));
    io.rfloats = (const float *)d_rf;
    io.output = (char *)d_out;
    CHECK(b200gru_generate(m, mode, n, &io, NULL));               /* namegen */
    CHECK(b200gru_memcpy_async(out, d_out, (size_t)n * (T + 1), 2, NULL));
    CHECK(b200gru_stream_synchronize(NULL));
    printf("GPU, %s path, %d launches\n", mode == B200GRU_MODE_TC ? "tcgen05" : "fp32 exact",
           b200gru_last_launch_count(m));
    CHECK(b200gru_symm_free(0, d_rf));
    CHECK(b200gru_symm_free(0, d_out));
    CHECK(b200gru_model_destroy(m));                              /* namegen_finalize */
  }
  for (int i = 0; i < n && i < 8; ++i) {
    printf("%4d:", i);
    for (int l = 0; l <= T; ++l) printf(" %3d", (unsigned char)out[i * (T + 1) + l]);
    printf("\n");
  }
  if (argc > 3) {
    FILE *f = fopen(argv[3], "wb");
    if (f) {
      fwrite(out, 1, (size_t)n * (T + 1), f);
      fclose(f);
    }
  }
  free(params); free(rf); free(out);
  return 0;
}
