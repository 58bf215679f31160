#include "tenzing/host_streams.hpp"

#include <chrono>

#include "tenzing/macro_at.hpp"

namespace tenzing {

void HostStreams::backoff(unsigned &spins) {
  // busy for a moment, then progressively lazier: an idle stream thread must not eat a core
  if (++spins < 256) {
    std::this_thread::yield();
  } else if (spins < 4096) {
    std::this_thread::sleep_for(std::chrono::microseconds(20));
  } else {
    std::this_thread::sleep_for(std::chrono::microseconds(500));
  }
}

HostStreams::HostStreams(int nStreams) {
  for (int i = 0; i < nStreams; ++i) {
    auto w = std::make_unique<Worker>();
    w->nodes.push_back(std::make_unique<Node>()); // dummy head: the consumer's starting point
    Node *first = w->nodes.back().get();
    w->tail = first;
    Worker *raw = w.get();
    w->thread = std::thread([this, raw, first] { worker_loop(raw, first); });
    workers_.push_back(std::move(w));
  }
}

HostStreams::~HostStreams() {
  for (auto &w : workers_) w->stop.store(true, std::memory_order_release);
  for (auto &w : workers_)
    if (w->thread.joinable()) w->thread.join();
}

void HostStreams::worker_loop(Worker *w, Node *cur) {
  unsigned spins = 0;
  for (;;) {
    Node *next = cur->next.load(std::memory_order_acquire);
    if (!next) {
      if (w->stop.load(std::memory_order_acquire)) return;
      backoff(spins);
      continue;
    }
    spins = 0;
    if (next->marker || !w->failed.load(std::memory_order_relaxed)) {
      try {
        next->task();
      } catch (...) {
        if (!w->failed.load(std::memory_order_relaxed)) {
          w->error = std::current_exception();
          w->failed.store(true, std::memory_order_relaxed);
        }
      }
    }
    cur = next;
    w->done.store(w->done.load(std::memory_order_relaxed) + 1, std::memory_order_release);
  }
}

void HostStreams::push(int stream, std::function<void()> task, bool marker) {
  if (stream < 0 || stream >= int(workers_.size()))
    THROW_RUNTIME("host streams: stream " << stream << " does not exist");
  Worker &w = *workers_[size_t(stream)];
  auto node = std::make_unique<Node>();
  node->task = std::move(task);
  node->marker = marker;
  Node *raw = node.get();
  w.nodes.push_back(std::move(node));
  w.issued.store(w.issued.load(std::memory_order_relaxed) + 1, std::memory_order_relaxed);
  w.tail->next.store(raw, std::memory_order_release); // issue order: host -> stream
  w.tail = raw;
}

void HostStreams::enqueue(int stream, std::function<void()> task) {
  push(stream, std::move(task), false);
}

HostStreams::EventState &HostStreams::event(int id) {
  auto it = events_.find(id);
  if (it == events_.end()) it = events_.emplace(id, std::make_unique<EventState>()).first;
  return *it->second;
}

void HostStreams::record(int eventId, int stream) {
  EventState &es = event(eventId);
  const uint64_t gen = ++es.issued;
  EventState *p = &es;
  // completes when everything issued to `stream` before it has completed (FIFO)
  push(stream, [p, gen] {
    uint64_t cur = p->done.load(std::memory_order_relaxed);
    while (cur < gen && !p->done.compare_exchange_weak(cur, gen, std::memory_order_release,
                                                       std::memory_order_relaxed)) {
    }
  }, true);
}

void HostStreams::stream_wait(int stream, int eventId) {
  EventState &es = event(eventId);
  const uint64_t gen = es.issued; // the most recent record at the time of the call
  if (gen == 0) return;           // never recorded: cudaStreamWaitEvent is a no-op
  EventState *p = &es;
  push(stream, [p, gen] {
    unsigned spins = 0;
    while (p->done.load(std::memory_order_acquire) < gen) backoff(spins);
  }, true);
}

void HostStreams::event_sync(int eventId) {
  EventState &es = event(eventId);
  unsigned spins = 0;
  while (es.done.load(std::memory_order_acquire) < es.issued) backoff(spins);
  rethrow_if_failed();
}

void HostStreams::wait_idle(Worker &w) {
  const uint64_t want = w.issued.load(std::memory_order_relaxed);
  unsigned spins = 0;
  while (w.done.load(std::memory_order_acquire) < want) backoff(spins);
  // everything but the node the consumer stands on can go
  while (w.nodes.size() > 1) w.nodes.pop_front();
}

void HostStreams::stream_sync(int stream) {
  if (stream < 0 || stream >= int(workers_.size()))
    THROW_RUNTIME("host streams: stream " << stream << " does not exist");
  wait_idle(*workers_[size_t(stream)]);
  rethrow_if_failed();
}

void HostStreams::sync_all() {
  for (auto &w : workers_) wait_idle(*w);
  rethrow_if_failed();
}

uint64_t HostStreams::tasks_done() const {
  uint64_t n = 0;
  for (const auto &w : workers_) n += w->done.load(std::memory_order_acquire);
  return n;
}

void HostStreams::rethrow_if_failed() {
  for (auto &w : workers_) {
    if (!w->failed.load(std::memory_order_relaxed)) continue;
    // only safe to read the exception once the stream is idle
    wait_idle(*w);
    std::exception_ptr e = w->error;
    w->error = nullptr;
    w->failed.store(false, std::memory_order_relaxed);
    if (e) std::rethrow_exception(e);
  }
}

} // namespace tenzing
