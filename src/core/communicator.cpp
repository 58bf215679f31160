#include "tenzing/communicator.hpp"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <exception>
#include <thread>

#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <poll.h>
#include <unistd.h>

#include "tenzing/macro_at.hpp"

namespace tenzing {

// ---- conveniences ----
std::string Communicator::bcast_string(const std::string &s, int root) {
  int64_t n = static_cast<int64_t>(s.size());
  bcast(&n, sizeof(n), root);
  std::string out = (rank() == root) ? s : std::string(size_t(n), '\0');
  if (n > 0) bcast(&out[0], size_t(n), root);
  return out;
}

int64_t Communicator::bcast_int(int64_t v, int root) {
  bcast(&v, sizeof(v), root);
  return v;
}

std::vector<std::string> Communicator::allgather_string(const std::string &s) {
  const int n = size();
  std::vector<int64_t> lens(n);
  int64_t mine = static_cast<int64_t>(s.size());
  allgather(&mine, lens.data(), sizeof(int64_t));
  const size_t maxLen = size_t(*std::max_element(lens.begin(), lens.end()));
  std::vector<std::string> out(n);
  if (maxLen == 0) return out;
  std::string padded = s;
  padded.resize(maxLen, '\0');
  std::string all(maxLen * n, '\0');
  allgather(padded.data(), &all[0], maxLen);
  for (int r = 0; r < n; ++r) out[r] = all.substr(size_t(r) * maxLen, size_t(lens[r]));
  return out;
}

// ---- loopback ----
void LoopbackWorld::run(int size,
                        const std::function<void(int, std::shared_ptr<Communicator>)> &body) {
  auto world = std::make_shared<LoopbackWorld>(size);
  std::vector<std::thread> threads;
  std::vector<std::exception_ptr> errors(size);
  for (int r = 0; r < size; ++r) {
    threads.emplace_back([&, r] {
      try {
        body(r, std::make_shared<LoopbackComm>(world, r));
      } catch (...) {
        errors[r] = std::current_exception();
        world->abort();
      }
    });
  }
  for (auto &t : threads) t.join();
  // report the root cause, not the "world aborted" that it caused on the other ranks
  std::exception_ptr first;
  for (auto &e : errors) {
    if (!e) continue;
    bool secondary = false;
    try {
      std::rethrow_exception(e);
    } catch (const LoopbackAborted &) {
      secondary = true;
    } catch (...) {
    }
    if (!secondary) std::rethrow_exception(e);
    if (!first) first = e;
  }
  if (first) std::rethrow_exception(first);
}

void LoopbackWorld::abort() {
  std::lock_guard<std::mutex> lk(mu_);
  aborted_ = true;
  cv_.notify_all();
}

template <typename F>
void LoopbackComm::rendezvous(const void *src, void *dst, F &&atRendezvous) {
  LoopbackWorld &w = *world_;
  std::unique_lock<std::mutex> lk(w.mu_);
  if (w.aborted_) throw LoopbackAborted();
  w.src_[rank_] = src;
  w.dst_[rank_] = dst;
  const uint64_t gen = w.generation_;
  if (++w.arrived_ == w.size_) {
    atRendezvous(w.src_, w.dst_);
    w.arrived_ = 0;
    ++w.generation_;
    w.cv_.notify_all();
  } else {
    w.cv_.wait(lk, [&] { return w.generation_ != gen || w.aborted_; });
    if (w.generation_ == gen) throw LoopbackAborted();
  }
}

void LoopbackComm::barrier() {
  rendezvous(nullptr, nullptr, [](const std::vector<const void *> &, const std::vector<void *> &) {});
}

void LoopbackComm::bcast(void *buf, size_t bytes, int root) {
  rendezvous(buf, buf, [&](const std::vector<const void *> &src, const std::vector<void *> &dst) {
    for (size_t r = 0; r < dst.size(); ++r)
      if (int(r) != root) std::memcpy(dst[r], src[root], bytes);
  });
}

void LoopbackComm::allreduce_max(double *vals, size_t n) {
  rendezvous(vals, vals, [&](const std::vector<const void *> &, const std::vector<void *> &dst) {
    for (size_t i = 0; i < n; ++i) {
      double m = static_cast<double *>(dst[0])[i];
      for (void *p : dst) m = std::max(m, static_cast<double *>(p)[i]);
      for (void *p : dst) static_cast<double *>(p)[i] = m;
    }
  });
}

void LoopbackComm::allgather(const void *in, void *out, size_t bytesPerRank) {
  rendezvous(in, out, [&](const std::vector<const void *> &src, const std::vector<void *> &dst) {
    for (void *d : dst)
      for (size_t r = 0; r < src.size(); ++r)
        std::memcpy(static_cast<char *>(d) + r * bytesPerRank, src[r], bytesPerRank);
  });
}

// ---- sockets ----
void SocketComm::send_all(int fd, const void *buf, size_t n) {
  const char *p = static_cast<const char *>(buf);
  while (n) {
    ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
    if (k <= 0) THROW_RUNTIME("SocketComm: send failed (peer gone?)");
    p += k;
    n -= size_t(k);
  }
}

void SocketComm::recv_all(int fd, void *buf, size_t n) {
  char *p = static_cast<char *>(buf);
  while (n) {
    ssize_t k = ::recv(fd, p, n, 0);
    if (k <= 0) THROW_RUNTIME("SocketComm: recv failed (peer gone?)");
    p += k;
    n -= size_t(k);
  }
}

SocketComm::SocketComm(int rank, int size, const std::string &addr, int port, double timeoutSec)
    : rank_(rank), size_(size) {
  log::set_rank(rank);
  if (size_ == 1) return;
  sockaddr_in sa{};
  sa.sin_family = AF_INET;
  sa.sin_port = htons(static_cast<uint16_t>(port));
  if (inet_pton(AF_INET, addr.c_str(), &sa.sin_addr) != 1) sa.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
  const int one = 1;
  if (rank_ == 0) {
    listenFd_ = ::socket(AF_INET, SOCK_STREAM, 0);
    setsockopt(listenFd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    if (::bind(listenFd_, reinterpret_cast<sockaddr *>(&sa), sizeof(sa)) != 0)
      THROW_RUNTIME("SocketComm: cannot bind " << addr << ":" << port);
    ::listen(listenFd_, size_);
    fds_.assign(size_, -1);
    const auto deadline =
        std::chrono::steady_clock::now() + std::chrono::duration<double>(timeoutSec);
    for (int i = 1; i < size_; ++i) {
      // a rank that never starts must not leave rank 0 in accept() for ever
      const double left =
          std::chrono::duration<double>(deadline - std::chrono::steady_clock::now()).count();
      pollfd pfd{listenFd_, POLLIN, 0};
      if (left <= 0 || ::poll(&pfd, 1, int(left * 1000) + 1) <= 0)
        THROW_RUNTIME("SocketComm: only " << (i - 1) << " of " << (size_ - 1)
                                          << " ranks connected within " << timeoutSec << " s");
      int fd = ::accept(listenFd_, nullptr, nullptr);
      if (fd < 0) THROW_RUNTIME("SocketComm: accept failed");
      setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
      int32_t peer = -1;
      recv_all(fd, &peer, sizeof(peer));
      if (peer <= 0 || peer >= size_ || fds_[peer] != -1) THROW_RUNTIME("SocketComm: bad hello");
      fds_[peer] = fd;
    }
  } else {
    const auto deadline =
        std::chrono::steady_clock::now() + std::chrono::duration<double>(timeoutSec);
    int fd = -1;
    while (true) {
      fd = ::socket(AF_INET, SOCK_STREAM, 0);
      if (::connect(fd, reinterpret_cast<sockaddr *>(&sa), sizeof(sa)) == 0) break;
      ::close(fd);
      if (std::chrono::steady_clock::now() > deadline)
        THROW_RUNTIME("SocketComm: cannot reach rank 0 at " << addr << ":" << port);
      std::this_thread::sleep_for(std::chrono::milliseconds(20));
    }
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
    int32_t me = rank_;
    send_all(fd, &me, sizeof(me));
    fds_.assign(1, fd);
  }
}

SocketComm::~SocketComm() {
  for (int fd : fds_)
    if (fd >= 0) ::close(fd);
  if (listenFd_ >= 0) ::close(listenFd_);
}

std::shared_ptr<Communicator> SocketComm::from_env(int portOffset) {
  const char *r = std::getenv("RANK");
  const char *w = std::getenv("WORLD_SIZE");
  if (!r || !w) return nullptr;
  const char *a = std::getenv("MASTER_ADDR");
  const char *p = std::getenv("MASTER_PORT");
  const int port = (p ? std::atoi(p) : 29500) + portOffset;
  return std::make_shared<SocketComm>(std::atoi(r), std::atoi(w), a ? a : "127.0.0.1", port);
}

void SocketComm::allgather(const void *in, void *out, size_t n) {
  if (size_ == 1) {
    std::memcpy(out, in, n);
    return;
  }
  if (rank_ == 0) {
    char *o = static_cast<char *>(out);
    std::memcpy(o, in, n);
    for (int r = 1; r < size_; ++r) recv_all(fds_[r], o + size_t(r) * n, n);
    for (int r = 1; r < size_; ++r) send_all(fds_[r], o, n * size_t(size_));
  } else {
    send_all(fds_[0], in, n);
    recv_all(fds_[0], out, n * size_t(size_));
  }
}

void SocketComm::barrier() {
  char token = 0;
  std::vector<char> all(size_);
  allgather(&token, all.data(), 1);
}

void SocketComm::bcast(void *buf, size_t bytes, int root) {
  if (size_ == 1 || bytes == 0) return;
  if (rank_ == 0) {
    if (root != 0) recv_all(fds_[root], buf, bytes);
    for (int r = 1; r < size_; ++r)
      if (r != root) send_all(fds_[r], buf, bytes);
  } else if (rank_ == root) {
    send_all(fds_[0], buf, bytes);
  } else {
    recv_all(fds_[0], buf, bytes);
  }
}

void SocketComm::allreduce_max(double *vals, size_t n) {
  if (size_ == 1 || n == 0) return;
  std::vector<double> all(n * size_t(size_));
  allgather(vals, all.data(), n * sizeof(double));
  for (size_t i = 0; i < n; ++i) {
    double m = all[i];
    for (int r = 1; r < size_; ++r) m = std::max(m, all[size_t(r) * n + i]);
    vals[i] = m;
  }
}

} // namespace tenzing
