// CUDA-stream semantics on CPU threads, for host execution (tenzing/host_exec.hpp).
//
// With Platform::enable_host_streams() every platform stream becomes a worker thread with a
// FIFO of tasks.  The host thread that runs a sequence only ISSUES: a BoundGpuOp becomes a task
// on its stream's thread, CudaEventRecord / CudaStreamWaitEvent become marker tasks,
// CudaEventSync / StreamSync block the host - exactly the ordering guarantees CUDA gives and no
// others:
//   * issue order   host -> stream   (what the host wrote before the launch is visible to the op)
//   * stream order  op -> next op of the same stream
//   * event order   record -> every wait on that event issued after the record
//   * host syncs    event / stream completion -> host
// Kernels on different streams therefore really run concurrently unless the schedule orders
// them, and ThreadSanitizer sees precisely the happens-before edges the schedule provides: a
// missing CudaEventRecord/CudaStreamWaitEvent pair, or a missing edge in a workload graph, shows
// up as a reported data race on the memory the two ops touch.  This is a dynamic race detector
// for schedules (the reference has none: SURVEY 5.2).
//
// The implementation avoids every synchronisation that is not one of the four above (no shared
// mutex between streams, no condition variables): queues are single-producer/single-consumer
// linked lists published with release stores, events and completion counters are atomics.
#pragma once

#include <atomic>
#include <cstdint>
#include <deque>
#include <exception>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <unordered_map>
#include <vector>

namespace tenzing {

class HostStreams {
public:
  explicit HostStreams(int nStreams);
  ~HostStreams();
  HostStreams(const HostStreams &) = delete;
  HostStreams &operator=(const HostStreams &) = delete;

  int num_streams() const { return int(workers_.size()); }
  // all of these are called by the one host thread that runs sequences
  void enqueue(int stream, std::function<void()> task);
  void record(int event, int stream);      // cudaEventRecord
  void stream_wait(int stream, int event); // cudaStreamWaitEvent (no-op if never recorded)
  void event_sync(int event);              // cudaEventSynchronize
  void stream_sync(int stream);            // cudaStreamSynchronize
  void sync_all();                         // every stream idle
  // tasks executed so far (all streams) - for tests
  uint64_t tasks_done() const;

private:
  struct Node {
    std::function<void()> task;
    bool marker = false; // event record / wait: still executed after a task of the stream failed
    std::atomic<Node *> next{nullptr};
  };
  struct Worker {
    std::thread thread;
    Node *tail = nullptr;                 // producer side (host thread only)
    std::atomic<uint64_t> issued{0};      // host thread writes, host thread reads
    std::atomic<uint64_t> done{0};        // worker publishes (release) after every task
    std::atomic<bool> stop{false};
    std::deque<std::unique_ptr<Node>> nodes; // owned by the host thread, trimmed at syncs
    std::exception_ptr error;             // first exception of a task (read after a sync)
    std::atomic<bool> failed{false};
  };
  struct EventState {
    uint64_t issued = 0;                  // host thread only: records issued so far
    std::atomic<uint64_t> done{0};        // generation of the latest completed record
  };
  std::vector<std::unique_ptr<Worker>> workers_;
  std::unordered_map<int, std::unique_ptr<EventState>> events_; // host thread only (the map)

  EventState &event(int id);
  void push(int stream, std::function<void()> task, bool marker);
  void wait_idle(Worker &w);
  void worker_loop(Worker *w, Node *first);
  void rethrow_if_failed();
  static void backoff(unsigned &spins);
};

} // namespace tenzing
