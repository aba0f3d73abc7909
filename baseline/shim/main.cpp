// main.cpp — stand-in for the driver the reference refers to but does not ship
// ("Defined in main.cpp", /root/reference/namegensf.cu:21; "allocated at
// main.cpp", :625).  It owns mpi_rank / mpi_size, makes the random floats,
// allocates the output, calls the reference's three functions and times
// namegen() the only way a host-driven program can be timed on the device:
// wall clock between cudaDeviceSynchronize() fences, max over ranks.
//
// usage: namegen_ref <weights.bin> <N> <seed> <steps> <warmup> [out.bin]
#include <cuda_runtime.h>
#include <mpi.h>

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "namegen.h"
#include "util.h"

int mpi_rank = 0, mpi_size = 1;

void *read_binary(const char *filename, size_t *size) {
  FILE *f = fopen(filename, "rb");
  if (!f) {
    fprintf(stderr, "cannot open %s\n", filename);
    exit(1);
  }
  fseek(f, 0, SEEK_END);
  size_t sz = (size_t)ftell(f);
  rewind(f);
  void *buf = malloc(sz);
  if (fread(buf, 1, sz, f) != sz) {
    fprintf(stderr, "short read on %s\n", filename);
    exit(1);
  }
  fclose(f);
  *size = sz;
  return buf;
}

// same generator as b200gru_random_floats (csrc/weights_io.cpp) so both arms
// see identical inputs
static void fill_random(uint64_t seed, float *out, int64_t n) {
  uint64_t s = seed ^ 0xD1B54A32D192ED03ull;
  for (int64_t i = 0; i < n; ++i) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    out[i] = (float)(z >> 40) * (1.0f / 16777216.0f);
  }
}

int main(int argc, char **argv) {
  if (argc < 6) {
    fprintf(stderr, "usage: %s weights N seed steps warmup [out.bin]\n", argv[0]);
    return 2;
  }
  MPI_Init(&argc, &argv);
  MPI_Comm_rank(MPI_COMM_WORLD, &mpi_rank);
  MPI_Comm_size(MPI_COMM_WORLD, &mpi_size);
  const int N = atoi(argv[2]);
  const int seed = atoi(argv[3]);
  const int steps = atoi(argv[4]);
  const int warmup = atoi(argv[5]);

  std::vector<float> rf;
  std::vector<char> out;
  if (mpi_rank == 0) {
    rf.resize((size_t)N * MAX_LEN);
    fill_random((uint64_t)seed, rf.data(), (int64_t)rf.size());
    out.assign((size_t)N * (MAX_LEN + 1), 0);
  }
  namegen_initialize(N, seed, argv[1]);
  double total_ms = 0.0;
  for (int it = 0; it < warmup + steps; ++it) {
    cudaDeviceSynchronize();
    MPI_Barrier(MPI_COMM_WORLD);
    auto t0 = std::chrono::steady_clock::now();
    namegen(N, rf.data(), out.data());
    cudaDeviceSynchronize();
    auto t1 = std::chrono::steady_clock::now();
    double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    ms = shim_allreduce_max(ms);
    if (it >= warmup) total_ms += ms;
    if (mpi_rank == 0)
      fprintf(stderr, "[ref] iter %d %s %.3f ms\n", it, it < warmup ? "warmup" : "timed", ms);
  }
  if (mpi_rank == 0) {
    int generated = (N / mpi_size) * mpi_size;
    printf("{\"impl\": \"reference\", \"n_names\": %d, \"generated\": %d, \"ranks\": %d, "
           "\"steps\": %d, \"warmup\": %d, \"ms_per_step\": %.6f, \"names_per_sec\": %.3f}\n",
           N, generated, mpi_size, steps, warmup, total_ms / (steps > 0 ? steps : 1),
           steps > 0 ? generated / (total_ms / steps / 1e3) : 0.0);
    if (argc > 6) {
      FILE *f = fopen(argv[6], "wb");
      if (f) {
        fwrite(out.data(), 1, out.size(), f);
        fclose(f);
      }
    }
  }
  namegen_finalize();
  MPI_Finalize();
  return 0;
}
