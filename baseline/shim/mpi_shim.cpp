// mpi_shim.cpp — shared-memory implementation of baseline/shim/mpi.h.
#include "mpi.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace {
constexpr size_t kDataBytes = size_t(512) << 20;  // staging area
struct Header {
  std::atomic<int> arrived;
  std::atomic<int> generation;
  std::atomic<int> attached;
  double slots[64];
};
int g_rank = 0, g_size = 1;
Header *g_hdr = nullptr;
char *g_data = nullptr;
char g_name[128];

void barrier() {
  if (g_size == 1) return;
  int gen = g_hdr->generation.load(std::memory_order_acquire);
  if (g_hdr->arrived.fetch_add(1, std::memory_order_acq_rel) == g_size - 1) {
    g_hdr->arrived.store(0, std::memory_order_relaxed);
    g_hdr->generation.fetch_add(1, std::memory_order_release);
  } else {
    while (g_hdr->generation.load(std::memory_order_acquire) == gen) usleep(20);
  }
}
}  // namespace

int MPI_Init(int *, char ***) {
  const char *r = getenv("RANK"), *s = getenv("WORLD_SIZE");
  g_rank = r ? atoi(r) : 0;
  g_size = s ? atoi(s) : 1;
  if (g_size == 1) return 0;
  const char *tag = getenv("B200GRU_SHM_TAG");
  snprintf(g_name, sizeof(g_name), "/b200gru_ref_%s", tag ? tag : "default");
  size_t total = sizeof(Header) + kDataBytes;
  int fd = -1;
  if (g_rank == 0) {
    shm_unlink(g_name);
    fd = shm_open(g_name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0 || ftruncate(fd, (off_t)total) != 0) {
      perror("shm_open/ftruncate");
      exit(1);
    }
  } else {
    for (int tries = 0; tries < 30000 && fd < 0; ++tries) {
      fd = shm_open(g_name, O_RDWR, 0600);
      struct stat st;
      if (fd >= 0 && (fstat(fd, &st) != 0 || (size_t)st.st_size < total)) {
        close(fd);
        fd = -1;
      }
      if (fd < 0) usleep(1000);
    }
    if (fd < 0) {
      fprintf(stderr, "rank %d: could not attach %s\n", g_rank, g_name);
      exit(1);
    }
  }
  void *p = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) {
    perror("mmap");
    exit(1);
  }
  g_hdr = static_cast<Header *>(p);
  g_data = static_cast<char *>(p) + sizeof(Header);
  g_hdr->attached.fetch_add(1);
  while (g_hdr->attached.load() < g_size) usleep(100);
  barrier();
  return 0;
}

int MPI_Finalize(void) {
  if (g_size > 1) {
    barrier();
    if (g_rank == 0) shm_unlink(g_name);
  }
  return 0;
}

int MPI_Get_processor_name(char *name, int *len) {
  if (gethostname(name, MPI_MAX_PROCESSOR_NAME) != 0) strcpy(name, "localhost");
  name[MPI_MAX_PROCESSOR_NAME - 1] = 0;
  *len = (int)strlen(name);
  return 0;
}
int MPI_Comm_rank(MPI_Comm, int *rank) { *rank = g_rank; return 0; }
int MPI_Comm_size(MPI_Comm, int *size) { *size = g_size; return 0; }
int MPI_Barrier(MPI_Comm) { barrier(); return 0; }

int MPI_Scatter(const void *sendbuf, int sendcount, MPI_Datatype sendtype,
                void *recvbuf, int recvcount, MPI_Datatype recvtype, int root,
                MPI_Comm) {
  size_t bytes = (size_t)recvcount * recvtype;
  if (g_size == 1) {
    memcpy(recvbuf, sendbuf, bytes);
    return 0;
  }
  if (bytes * g_size > kDataBytes) {
    fprintf(stderr, "shim: scatter too large\n");
    exit(1);
  }
  if (g_rank == root) memcpy(g_data, sendbuf, (size_t)sendcount * sendtype * g_size);
  barrier();
  memcpy(recvbuf, g_data + bytes * g_rank, bytes);
  barrier();
  return 0;
}

int MPI_Gather(const void *sendbuf, int sendcount, MPI_Datatype sendtype,
               void *recvbuf, int recvcount, MPI_Datatype recvtype, int root,
               MPI_Comm) {
  size_t bytes = (size_t)sendcount * sendtype;
  if (g_size == 1) {
    memcpy(recvbuf, sendbuf, bytes);
    return 0;
  }
  if (bytes * g_size > kDataBytes) {
    fprintf(stderr, "shim: gather too large\n");
    exit(1);
  }
  memcpy(g_data + bytes * g_rank, sendbuf, bytes);
  barrier();
  if (g_rank == root) memcpy(recvbuf, g_data, (size_t)recvcount * recvtype * g_size);
  barrier();
  return 0;
}

double shim_allreduce_max(double v) {
  if (g_size == 1) return v;
  g_hdr->slots[g_rank] = v;
  barrier();
  double m = g_hdr->slots[0];
  for (int i = 1; i < g_size; ++i) m = g_hdr->slots[i] > m ? g_hdr->slots[i] : m;
  barrier();
  return m;
}
