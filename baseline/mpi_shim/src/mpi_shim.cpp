// Single-node MPI stand-in for building and timing the reference (see ../include/mpi.h and
// baseline/README.md).  Baseline-only code: never linked into tenzing_b200.
//
// Ranks are processes of one host.  They find each other through a POSIX shared-memory segment
// named after TZMPI_JOB; rank and size come from TZMPI_RANK/TZMPI_SIZE (or RANK/WORLD_SIZE, so a
// torchrun-style launcher works too).
//
//   point to point   per (source, destination) a ring of 128-byte descriptors in shared memory.
//                    A send posts a descriptor; the receiver matches by tag in posting order.
//     device buffer  rendezvous, receiver pulls: the descriptor carries the CUDA IPC handle of
//                    the sender's allocation and the offset inside it; the receiver opens the
//                    handle once per allocation (cached), issues cudaMemcpyAsync on one of
//                    kStreams internal non-blocking streams and completes on a CUDA event, then
//                    marks the descriptor done, which completes the sender's request.
//     host buffer    eager through a named shared-memory object (set-up traffic only).
//   collectives      barrier (sense counter), bcast and allreduce through per-rank slots.
//   progress         every blocking call progresses all pending receives of the process, so
//                    "post everything, then wait in any order" cannot deadlock.
#include "mpi.h"

#include <cuda_runtime.h>
#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <set>
#include <string>
#include <vector>

namespace {

constexpr int kMaxRanks = 16;
constexpr uint64_t kQueue = 1024;         // descriptors per (src, dst) ring
constexpr size_t kCollBytes = 256 * 1024; // per-rank collective slot
constexpr int kStreams = 4;
constexpr uint64_t kMagic = 0x545a4d5049534d31ull; // "TZMPISM1"

enum Kind : int32_t { kZero = 0, kHost = 1, kDevice = 2, kSameProcess = 3 };

struct alignas(128) Desc {
  std::atomic<uint32_t> state; // 0 empty, 1 posted, 2 matched, 3 done
  int32_t tag;
  uint64_t seq;
  uint64_t bytes;
  int32_t kind;
  int32_t pad;
  uint64_t offset; // device: offset inside the allocation; same process: the source pointer
  unsigned char handle[64];
};
static_assert(sizeof(Desc) == 128, "descriptor layout");

struct PairQueue {
  alignas(64) std::atomic<uint64_t> head; // descriptors ever posted
  Desc ring[kQueue];
};

struct Shared {
  std::atomic<uint64_t> magic;
  int32_t size;
  std::atomic<int32_t> attached;
  std::atomic<int32_t> aborted;
  alignas(64) std::atomic<uint32_t> barCount;
  alignas(64) std::atomic<uint32_t> barGen;
  alignas(64) unsigned char coll[kMaxRanks][kCollBytes];
  PairQueue q[1]; // size * size, index src * size + dst
};

struct Request {
  enum Type { kFree, kSend, kRecv, kMulti } type = kFree;
  bool done = false;
  // send: complete once the receiver marked the descriptor done (or the slot moved on)
  Desc *desc = nullptr;
  uint64_t seq = 0;
  // recv
  int src = -1, tag = 0;
  void *buf = nullptr;
  uint64_t capacity = 0, received = 0;
  bool matched = false;
  cudaEvent_t ev = nullptr;
  // multi (Ialltoallv)
  std::vector<int> children;
};

struct State {
  bool initialized = false, finalized = false;
  int rank = 0, size = 1;
  std::string job;
  Shared *sh = nullptr;
  size_t shBytes = 0;
  std::vector<Request> reqs;
  std::vector<int> freeReqs;
  std::vector<int> pendingRecvs;
  uint64_t sendSeq[kMaxRanks] = {};
  uint64_t tail[kMaxRanks] = {};
  std::set<uint64_t> matchedAhead[kMaxRanks];
  cudaStream_t streams[kStreams] = {};
  int nextStream = 0;
  bool cudaReady = false;
  std::vector<cudaEvent_t> eventPool;
  // sender side: allocation base -> IPC handle; receiver side: (src rank, handle) -> mapped base
  std::map<uintptr_t, cudaIpcMemHandle_t> exported;
  std::map<std::pair<int, std::string>, void *> imported;
} g;

const int kTypeSize[] = {0, 1, 1, 4, 4, 8, 4, 8, 8, 8, 8, 2, 1, 8, 8, 4, 4, 1};

[[noreturn]] void die(const char *what) {
  std::fprintf(stderr, "[tz-mpi rank %d] fatal: %s\n", g.rank, what);
  if (g.sh) g.sh->aborted.store(1);
  _exit(70);
}
#define CU(x)                                                                                      \
  do {                                                                                             \
    cudaError_t e_ = (x);                                                                          \
    if (e_ != cudaSuccess) {                                                                       \
      std::fprintf(stderr, "[tz-mpi rank %d] %s:%d %s\n", g.rank, __FILE__, __LINE__,             \
                   cudaGetErrorString(e_));                                                        \
      die("CUDA error");                                                                           \
    }                                                                                              \
  } while (0)

void check_abort() {
  if (g.sh && g.sh->aborted.load(std::memory_order_relaxed)) {
    std::fprintf(stderr, "[tz-mpi rank %d] another rank aborted\n", g.rank);
    _exit(71);
  }
}

inline void pause_a_little(unsigned &spins) {
  if (++spins < 2000) {
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  } else {
    sched_yield();
    if ((spins & 0xfff) == 0) check_abort();
  }
}

const char *env_first(std::initializer_list<const char *> names) {
  for (const char *n : names) {
    const char *v = std::getenv(n);
    if (v && *v) return v;
  }
  return nullptr;
}

std::string shm_name() { return "/tzmpi_" + std::to_string(getuid()) + "_" + g.job; }
std::string msg_name(int src, int dst, uint64_t seq) {
  return "/tzmpi_" + std::to_string(getuid()) + "_" + g.job + "_m" + std::to_string(src) + "_" +
         std::to_string(dst) + "_" + std::to_string(seq);
}

PairQueue &queue(int src, int dst) { return g.sh->q[size_t(src) * size_t(g.size) + size_t(dst)]; }

void ensure_cuda() {
  if (g.cudaReady) return;
  for (auto &s : g.streams) CU(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  g.cudaReady = true;
}

cudaEvent_t get_event() {
  if (!g.eventPool.empty()) {
    cudaEvent_t e = g.eventPool.back();
    g.eventPool.pop_back();
    return e;
  }
  cudaEvent_t e;
  CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  return e;
}

bool is_device_pointer(const void *p) {
  cudaPointerAttributes a;
  cudaError_t e = cudaPointerGetAttributes(&a, p);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice;
}

// base address and size of the allocation that contains p (driver API through the runtime's
// entry-point query, so the shim does not link libcuda)
void allocation_range(const void *p, uintptr_t &base, size_t &size) {
  typedef int (*GetRange)(unsigned long long *, size_t *, unsigned long long);
  static GetRange fn = [] {
    void *f = nullptr;
    cudaDriverEntryPointQueryResult r;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &f, cudaEnableDefault, &r) != cudaSuccess || !f)
      die("cuMemGetAddressRange not available");
    return reinterpret_cast<GetRange>(f);
  }();
  unsigned long long b = 0;
  size_t s = 0;
  if (fn(&b, &s, (unsigned long long)(uintptr_t)p) != 0) die("cuMemGetAddressRange failed");
  base = uintptr_t(b);
  size = s;
}

int new_request() {
  int i;
  if (!g.freeReqs.empty()) {
    i = g.freeReqs.back();
    g.freeReqs.pop_back();
  } else {
    i = int(g.reqs.size());
    g.reqs.emplace_back();
  }
  g.reqs[size_t(i)] = Request();
  return i;
}

void release_request(int i) {
  g.reqs[size_t(i)] = Request();
  g.freeReqs.push_back(i);
}

// ---- receiving ---------------------------------------------------------------------------
// first unmatched descriptor from src with this tag (posting order), or nullptr
Desc *find_match(int src, int tag, uint64_t *seqOut) {
  PairQueue &pq = queue(src, g.rank);
  const uint64_t head = pq.head.load(std::memory_order_acquire);
  for (uint64_t s = g.tail[src]; s < head; ++s) {
    if (g.matchedAhead[src].count(s)) continue;
    Desc &d = pq.ring[s % kQueue];
    if (d.state.load(std::memory_order_acquire) != 1 || d.seq != s) break; // not visible yet
    if (tag == MPI_ANY_TAG || d.tag == tag) {
      *seqOut = s;
      return &d;
    }
  }
  return nullptr;
}

void mark_matched(int src, uint64_t s) {
  if (s == g.tail[src]) {
    ++g.tail[src];
    auto &ahead = g.matchedAhead[src];
    while (!ahead.empty() && *ahead.begin() == g.tail[src]) {
      ahead.erase(ahead.begin());
      ++g.tail[src];
    }
  } else {
    g.matchedAhead[src].insert(s);
  }
}

void *import_allocation(int src, const unsigned char *handle) {
  const std::pair<int, std::string> key(src, std::string(reinterpret_cast<const char *>(handle), 64));
  auto it = g.imported.find(key);
  if (it != g.imported.end()) return it->second;
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, sizeof(h));
  void *base = nullptr;
  CU(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
  g.imported[key] = base;
  return base;
}

// one step of a receive request; returns true when it is complete
bool progress_recv(Request &r) {
  if (r.done) return true;
  if (!r.matched) {
    uint64_t seq = 0;
    Desc *d = find_match(r.src, r.tag, &seq);
    if (!d) return false;
    if (d->bytes > r.capacity) die("message longer than the receive buffer");
    mark_matched(r.src, seq);
    d->state.store(2, std::memory_order_relaxed);
    r.matched = true;
    r.desc = d;
    r.received = d->bytes;
    r.tag = d->tag;
    if (d->kind == kZero || d->bytes == 0) {
      d->state.store(3, std::memory_order_release);
      r.done = true;
      return true;
    }
    if (d->kind == kHost) {
      const std::string name = msg_name(r.src, g.rank, seq);
      int fd = shm_open(name.c_str(), O_RDONLY, 0600);
      if (fd < 0) die("cannot open a message segment");
      void *m = mmap(nullptr, d->bytes, PROT_READ, MAP_SHARED, fd, 0);
      close(fd);
      if (m == MAP_FAILED) die("cannot map a message segment");
      if (is_device_pointer(r.buf)) CU(cudaMemcpy(r.buf, m, d->bytes, cudaMemcpyHostToDevice));
      else std::memcpy(r.buf, m, d->bytes);
      munmap(m, d->bytes);
      shm_unlink(name.c_str());
      d->state.store(3, std::memory_order_release);
      r.done = true;
      return true;
    }
    ensure_cuda();
    const void *srcPtr = d->kind == kSameProcess
                             ? reinterpret_cast<const void *>(uintptr_t(d->offset))
                             : static_cast<const char *>(import_allocation(r.src, d->handle)) + d->offset;
    cudaStream_t s = g.streams[g.nextStream];
    g.nextStream = (g.nextStream + 1) % kStreams;
    CU(cudaMemcpyAsync(r.buf, srcPtr, d->bytes, cudaMemcpyDefault, s));
    r.ev = get_event();
    CU(cudaEventRecord(r.ev, s));
  }
  const cudaError_t q = cudaEventQuery(r.ev);
  if (q == cudaErrorNotReady) return false;
  CU(q);
  g.eventPool.push_back(r.ev);
  r.ev = nullptr;
  r.desc->state.store(3, std::memory_order_release);
  r.done = true;
  return true;
}

void progress_all() {
  for (size_t i = 0; i < g.pendingRecvs.size();) {
    Request &r = g.reqs[size_t(g.pendingRecvs[i])];
    if (r.type != Request::kRecv || progress_recv(r)) {
      g.pendingRecvs[i] = g.pendingRecvs.back();
      g.pendingRecvs.pop_back();
    } else {
      ++i;
    }
  }
}

bool test_request(int idx) {
  Request &r = g.reqs[size_t(idx)];
  switch (r.type) {
  case Request::kFree: return true;
  case Request::kSend:
    if (!r.done && (r.desc->seq != r.seq || r.desc->state.load(std::memory_order_acquire) == 3)) r.done = true;
    return r.done;
  case Request::kRecv: return r.done || progress_recv(r);
  case Request::kMulti: {
    bool all = true;
    for (int c : r.children) all = test_request(c) && all;
    return all;
  }
  }
  return true;
}

void finish_request(int idx, MPI_Status *status) {
  Request &r = g.reqs[size_t(idx)];
  if (status && r.type == Request::kRecv) {
    status->MPI_SOURCE = r.src;
    status->MPI_TAG = r.tag;
    status->MPI_ERROR = MPI_SUCCESS;
    status->tz_bytes = (long long)r.received;
  }
  if (r.type == Request::kMulti)
    for (int c : r.children) release_request(c);
  release_request(idx);
}

void wait_request(MPI_Request *request, MPI_Status *status) {
  if (!request || *request == MPI_REQUEST_NULL) return;
  const int idx = *request;
  if (idx < 0 || size_t(idx) >= g.reqs.size()) die("bad request handle");
  unsigned spins = 0;
  while (!test_request(idx)) {
    progress_all();
    pause_a_little(spins);
  }
  finish_request(idx, status);
  *request = MPI_REQUEST_NULL;
}

int post_send(const void *buf, uint64_t bytes, int dest, int tag) {
  if (dest < 0 || dest >= g.size) die("send to a rank outside the world");
  PairQueue &pq = queue(g.rank, dest);
  const uint64_t seq = g.sendSeq[dest]++;
  Desc &d = pq.ring[seq % kQueue];
  // the slot's previous message (seq - kQueue) must have been received
  unsigned spins = 0;
  while (seq >= kQueue && d.state.load(std::memory_order_acquire) != 3) {
    progress_all();
    pause_a_little(spins);
  }
  const int idx = new_request();
  Request &r = g.reqs[size_t(idx)];
  r.type = Request::kSend;
  r.desc = &d;
  r.seq = seq;
  d.state.store(0, std::memory_order_relaxed);
  d.tag = tag;
  d.seq = seq;
  d.bytes = bytes;
  d.offset = 0;
  if (bytes == 0) {
    d.kind = kZero;
    r.done = true;
  } else if (is_device_pointer(buf)) {
    if (dest == g.rank) {
      d.kind = kSameProcess;
      d.offset = uint64_t(uintptr_t(buf));
    } else {
      uintptr_t base = 0;
      size_t size = 0;
      allocation_range(buf, base, size);
      auto it = g.exported.find(base);
      if (it == g.exported.end()) {
        cudaIpcMemHandle_t h;
        CU(cudaIpcGetMemHandle(&h, reinterpret_cast<void *>(base)));
        it = g.exported.emplace(base, h).first;
      }
      d.kind = kDevice;
      d.offset = uint64_t(uintptr_t(buf) - base);
      std::memcpy(d.handle, &it->second, sizeof(cudaIpcMemHandle_t));
    }
  } else {
    // eager: the payload goes into a named segment the receiver unlinks
    const std::string name = msg_name(g.rank, dest, seq);
    shm_unlink(name.c_str());
    int fd = shm_open(name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0 || ftruncate(fd, off_t(bytes)) != 0) die("cannot create a message segment");
    void *m = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (m == MAP_FAILED) die("cannot map a message segment");
    std::memcpy(m, buf, bytes);
    munmap(m, bytes);
    d.kind = kHost;
    r.done = true; // buffered: the user buffer may be reused
  }
  d.state.store(1, std::memory_order_release);
  pq.head.store(seq + 1, std::memory_order_release);
  return idx;
}

int post_recv(void *buf, uint64_t bytes, int source, int tag) {
  if (source < 0 || source >= g.size) die("receive from a rank outside the world (wildcards are not supported)");
  const int idx = new_request();
  Request &r = g.reqs[size_t(idx)];
  r.type = Request::kRecv;
  r.src = source;
  r.tag = tag;
  r.buf = buf;
  r.capacity = bytes;
  g.pendingRecvs.push_back(idx);
  progress_recv(g.reqs[size_t(idx)]);
  return idx;
}

void barrier() {
  if (g.size == 1) return;
  const uint32_t gen = g.sh->barGen.load(std::memory_order_acquire);
  if (g.sh->barCount.fetch_add(1, std::memory_order_acq_rel) + 1 == uint32_t(g.size)) {
    g.sh->barCount.store(0, std::memory_order_relaxed);
    g.sh->barGen.store(gen + 1, std::memory_order_release);
  } else {
    unsigned spins = 0;
    while (g.sh->barGen.load(std::memory_order_acquire) == gen) {
      progress_all();
      pause_a_little(spins);
    }
  }
}

template <typename T> void reduce_into(T *acc, const T *in, size_t n, MPI_Op op) {
  for (size_t i = 0; i < n; ++i) {
    if (op == MPI_MAX) acc[i] = in[i] > acc[i] ? in[i] : acc[i];
    else if (op == MPI_MIN) acc[i] = in[i] < acc[i] ? in[i] : acc[i];
    else acc[i] = acc[i] + in[i];
  }
}

void reduce_bytes(void *acc, const void *in, size_t bytes, MPI_Datatype t, MPI_Op op) {
  switch (t) {
  case MPI_DOUBLE: reduce_into(static_cast<double *>(acc), static_cast<const double *>(in), bytes / 8, op); break;
  case MPI_FLOAT: reduce_into(static_cast<float *>(acc), static_cast<const float *>(in), bytes / 4, op); break;
  case MPI_INT:
  case MPI_INT32_T: reduce_into(static_cast<int *>(acc), static_cast<const int *>(in), bytes / 4, op); break;
  case MPI_UNSIGNED:
  case MPI_UINT32_T: reduce_into(static_cast<unsigned *>(acc), static_cast<const unsigned *>(in), bytes / 4, op); break;
  case MPI_LONG:
  case MPI_LONG_LONG:
  case MPI_INT64_T: reduce_into(static_cast<long long *>(acc), static_cast<const long long *>(in), bytes / 8, op); break;
  case MPI_UNSIGNED_LONG:
  case MPI_UNSIGNED_LONG_LONG:
  case MPI_UINT64_T:
    reduce_into(static_cast<unsigned long long *>(acc), static_cast<const unsigned long long *>(in), bytes / 8, op);
    break;
  default: die("allreduce: unsupported datatype");
  }
}

int type_size(MPI_Datatype t) {
  if (t <= 0 || t >= int(sizeof(kTypeSize) / sizeof(kTypeSize[0]))) die("unknown datatype");
  return kTypeSize[t];
}

} // namespace

extern "C" {

int MPI_Init(int *, char ***) {
  if (g.initialized) return MPI_SUCCESS;
  const char *r = env_first({"TZMPI_RANK", "RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK"});
  const char *s = env_first({"TZMPI_SIZE", "WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE"});
  const char *j = env_first({"TZMPI_JOB", "MASTER_PORT"});
  g.rank = r ? std::atoi(r) : 0;
  g.size = s ? std::atoi(s) : 1;
  g.job = j ? j : "0";
  if (g.size < 1 || g.size > kMaxRanks || g.rank < 0 || g.rank >= g.size) die("bad rank / size in the environment");
  g.shBytes = sizeof(Shared) + sizeof(PairQueue) * (size_t(g.size) * size_t(g.size));
  const std::string name = shm_name();
  int fd = -1;
  if (g.rank == 0) {
    shm_unlink(name.c_str()); // a stale segment of a crashed job with the same id
    fd = shm_open(name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0 || ftruncate(fd, off_t(g.shBytes)) != 0) die("cannot create the job segment");
  } else {
    // rank 0 creates the segment; wait for it (and for its magic word)
    for (int tries = 0; fd < 0; ++tries) {
      fd = shm_open(name.c_str(), O_RDWR, 0600);
      if (fd >= 0) {
        struct stat st;
        if (fstat(fd, &st) == 0 && size_t(st.st_size) >= g.shBytes) break;
        close(fd);
        fd = -1;
      }
      if (tries > 60000) die("the job segment never appeared (is rank 0 running?)");
      usleep(1000);
    }
  }
  void *m = mmap(nullptr, g.shBytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (m == MAP_FAILED) die("cannot map the job segment");
  g.sh = static_cast<Shared *>(m);
  if (g.rank == 0) {
    // a fresh POSIX segment is zero-filled: all counters and descriptor states start at 0
    g.sh->size = g.size;
    g.sh->magic.store(kMagic, std::memory_order_release);
  } else {
    unsigned spins = 0;
    while (g.sh->magic.load(std::memory_order_acquire) != kMagic) pause_a_little(spins);
    if (g.sh->size != g.size) die("ranks disagree about the world size");
  }
  g.sh->attached.fetch_add(1);
  g.initialized = true;
  barrier();
  // everybody is attached: the name is no longer needed (the memory lives until the last unmap)
  if (g.rank == 0) shm_unlink(name.c_str());
  return MPI_SUCCESS;
}

int MPI_Init_thread(int *argc, char ***argv, int required, int *provided) {
  if (provided) *provided = required < MPI_THREAD_FUNNELED ? required : MPI_THREAD_FUNNELED;
  return MPI_Init(argc, argv);
}

int MPI_Finalize(void) {
  if (!g.initialized || g.finalized) return MPI_SUCCESS;
  barrier();
  g.finalized = true;
  return MPI_SUCCESS;
}

int MPI_Initialized(int *flag) {
  *flag = g.initialized ? 1 : 0;
  return MPI_SUCCESS;
}
int MPI_Finalized(int *flag) {
  *flag = g.finalized ? 1 : 0;
  return MPI_SUCCESS;
}
int MPI_Abort(MPI_Comm, int errorcode) {
  if (g.sh) g.sh->aborted.store(1);
  _exit(errorcode ? errorcode : 1);
}
int MPI_Comm_rank(MPI_Comm comm, int *rank) {
  *rank = comm == MPI_COMM_SELF ? 0 : g.rank;
  return MPI_SUCCESS;
}
int MPI_Comm_size(MPI_Comm comm, int *size) {
  *size = comm == MPI_COMM_SELF ? 1 : g.size;
  return MPI_SUCCESS;
}
int MPI_Get_processor_name(char *name, int *resultlen) {
  if (gethostname(name, MPI_MAX_PROCESSOR_NAME - 1) != 0) std::strcpy(name, "localhost");
  name[MPI_MAX_PROCESSOR_NAME - 1] = 0;
  *resultlen = int(std::strlen(name));
  return MPI_SUCCESS;
}
double MPI_Wtime(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return double(ts.tv_sec) + double(ts.tv_nsec) * 1e-9;
}
double MPI_Wtick(void) { return 1e-9; }

int MPI_Barrier(MPI_Comm comm) {
  if (comm != MPI_COMM_SELF) barrier();
  return MPI_SUCCESS;
}

int MPI_Bcast(void *buffer, int count, MPI_Datatype datatype, int root, MPI_Comm comm) {
  if (comm == MPI_COMM_SELF || g.size == 1) return MPI_SUCCESS;
  const size_t total = size_t(count) * size_t(type_size(datatype));
  char *p = static_cast<char *>(buffer);
  for (size_t off = 0; off < total; off += kCollBytes) {
    const size_t n = total - off < kCollBytes ? total - off : kCollBytes;
    if (g.rank == root) std::memcpy(g.sh->coll[root], p + off, n);
    barrier();
    if (g.rank != root) std::memcpy(p + off, g.sh->coll[root], n);
    barrier();
  }
  return MPI_SUCCESS;
}

int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype datatype, MPI_Op op,
                  MPI_Comm comm) {
  const size_t total = size_t(count) * size_t(type_size(datatype));
  if (sendbuf != MPI_IN_PLACE && sendbuf != recvbuf) std::memcpy(recvbuf, sendbuf, total);
  if (comm == MPI_COMM_SELF || g.size == 1) return MPI_SUCCESS;
  char *p = static_cast<char *>(recvbuf);
  for (size_t off = 0; off < total; off += kCollBytes) {
    const size_t n = total - off < kCollBytes ? total - off : kCollBytes;
    std::memcpy(g.sh->coll[g.rank], p + off, n);
    barrier();
    for (int r = 0; r < g.size; ++r)
      if (r != g.rank) reduce_bytes(p + off, g.sh->coll[r], n, datatype, op);
    barrier();
  }
  return MPI_SUCCESS;
}

int MPI_Isend(const void *buf, int count, MPI_Datatype datatype, int dest, int tag, MPI_Comm, MPI_Request *request) {
  *request = post_send(buf, uint64_t(count) * uint64_t(type_size(datatype)), dest, tag);
  return MPI_SUCCESS;
}

int MPI_Irecv(void *buf, int count, MPI_Datatype datatype, int source, int tag, MPI_Comm, MPI_Request *request) {
  *request = post_recv(buf, uint64_t(count) * uint64_t(type_size(datatype)), source, tag);
  return MPI_SUCCESS;
}

int MPI_Send(const void *buf, int count, MPI_Datatype datatype, int dest, int tag, MPI_Comm comm) {
  MPI_Request r;
  MPI_Isend(buf, count, datatype, dest, tag, comm, &r);
  wait_request(&r, nullptr);
  return MPI_SUCCESS;
}

int MPI_Recv(void *buf, int count, MPI_Datatype datatype, int source, int tag, MPI_Comm comm, MPI_Status *status) {
  MPI_Request r;
  MPI_Irecv(buf, count, datatype, source, tag, comm, &r);
  wait_request(&r, status);
  return MPI_SUCCESS;
}

int MPI_Wait(MPI_Request *request, MPI_Status *status) {
  wait_request(request, status);
  return MPI_SUCCESS;
}

int MPI_Waitall(int count, MPI_Request reqs[], MPI_Status statuses[]) {
  for (int i = 0; i < count; ++i) wait_request(&reqs[i], statuses ? &statuses[i] : nullptr);
  return MPI_SUCCESS;
}

int MPI_Test(MPI_Request *request, int *flag, MPI_Status *status) {
  *flag = 1;
  if (!request || *request == MPI_REQUEST_NULL) return MPI_SUCCESS;
  progress_all();
  if (test_request(*request)) {
    finish_request(*request, status);
    *request = MPI_REQUEST_NULL;
  } else {
    *flag = 0;
  }
  return MPI_SUCCESS;
}

int MPI_Iprobe(int source, int tag, MPI_Comm, int *flag, MPI_Status *status) {
  if (source < 0 || source >= g.size) die("probe: wildcards are not supported");
  progress_all();
  uint64_t seq = 0;
  Desc *d = find_match(source, tag, &seq);
  *flag = d ? 1 : 0;
  if (d && status) {
    status->MPI_SOURCE = source;
    status->MPI_TAG = d->tag;
    status->MPI_ERROR = MPI_SUCCESS;
    status->tz_bytes = (long long)d->bytes;
  }
  return MPI_SUCCESS;
}

int MPI_Probe(int source, int tag, MPI_Comm comm, MPI_Status *status) {
  int flag = 0;
  unsigned spins = 0;
  for (;;) {
    MPI_Iprobe(source, tag, comm, &flag, status);
    if (flag) return MPI_SUCCESS;
    pause_a_little(spins);
  }
}

int MPI_Get_count(const MPI_Status *status, MPI_Datatype datatype, int *count) {
  const int ts = type_size(datatype);
  *count = status->tz_bytes % ts ? MPI_UNDEFINED : int(status->tz_bytes / ts);
  return MPI_SUCCESS;
}

int MPI_Ialltoallv(const void *sendbuf, const int sendcounts[], const int sdispls[], MPI_Datatype sendtype,
                   void *recvbuf, const int recvcounts[], const int rdispls[], MPI_Datatype recvtype, MPI_Comm,
                   MPI_Request *request) {
  const int ss = type_size(sendtype), rs = type_size(recvtype);
  const int idx = new_request();
  std::vector<int> children;
  constexpr int kTag = 0x7a11;
  for (int r = 0; r < g.size; ++r)
    children.push_back(post_recv(static_cast<char *>(recvbuf) + size_t(rdispls[r]) * size_t(rs),
                                 uint64_t(recvcounts[r]) * uint64_t(rs), r, kTag));
  for (int r = 0; r < g.size; ++r)
    children.push_back(post_send(static_cast<const char *>(sendbuf) + size_t(sdispls[r]) * size_t(ss),
                                 uint64_t(sendcounts[r]) * uint64_t(ss), r, kTag));
  Request &m = g.reqs[size_t(idx)];
  m.type = Request::kMulti;
  m.children = children;
  *request = idx;
  return MPI_SUCCESS;
}

int MPI_Type_size(MPI_Datatype datatype, int *size) {
  *size = type_size(datatype);
  return MPI_SUCCESS;
}

int MPI_Error_string(int errorcode, char *string, int *resultlen) {
  std::snprintf(string, MPI_MAX_ERROR_STRING, "tz-mpi error %d", errorcode);
  *resultlen = int(std::strlen(string));
  return MPI_SUCCESS;
}

} // extern "C"
