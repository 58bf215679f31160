#include "tenzing/ops_comm.hpp"

#include "tenzing/host_exec.hpp"

namespace tenzing {

P2PTransport::P2PTransport(bool nullDevice) : null_(nullDevice) {
  if (null_) return;
  for (auto &s : send_) CUDA_RUNTIME(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  CUDA_RUNTIME(cudaStreamCreateWithFlags(&recv_, cudaStreamNonBlocking));
}

P2PTransport::~P2PTransport() {
  for (auto s : send_)
    if (s) cudaStreamDestroy(s);
  if (recv_) cudaStreamDestroy(recv_);
}

Request::~Request() {
  if (sent) cudaEventDestroy(sent);
}

void Request::wait() {
  if (host) {
    // a host-executed send was complete when it was posted; a receive polls its flag
    if (kind == Kind::Recv) host::flags_wait(flag, 1, epoch);
    kind = Kind::None;
    return;
  }
  if (!transport || transport->is_null()) return;
  if (kind == Kind::Send) {
    if (sent) CUDA_RUNTIME(cudaEventSynchronize(sent));
  } else if (kind == Kind::Recv) {
    CUDA_RUNTIME(cudaError_t(tzb_flags_wait(flag, 1, epoch, transport->recv_stream())));
    CUDA_RUNTIME(cudaStreamSynchronize(transport->recv_stream()));
  }
  kind = Kind::None;
}

static void mark_sent(Request &r, const std::shared_ptr<P2PTransport> &tr, uint64_t epoch,
                      cudaStream_t stream) {
  r.kind = Request::Kind::Send;
  r.transport = tr;
  r.epoch = epoch;
  if (!r.sent) CUDA_RUNTIME(cudaEventCreateWithFlags(&r.sent, cudaEventDisableTiming));
  CUDA_RUNTIME(cudaEventRecord(r.sent, stream));
}

void Isend::run(Platform &plat) {
  if (plat.host_execution()) {
    if (args_.bytes) host::copy_bytes(args_.peerBuf, args_.buf, args_.bytes);
    if (args_.peerFlag) host::flag_signal(args_.peerFlag, plat.epoch());
    if (args_.request) {
      args_.request->kind = Request::Kind::Send;
      args_.request->host = true;
      args_.request->epoch = plat.epoch();
    }
    return;
  }
  if (tr_->is_null() || plat.is_null()) return;
  cudaStream_t st = tr_->next_send_stream();
  if (args_.bytes)
    CUDA_RUNTIME(cudaMemcpyAsync(args_.peerBuf, args_.buf, args_.bytes, cudaMemcpyDeviceToDevice, st));
  TzbFlagList l;
  l.n = 1;
  l.flag[0] = args_.peerFlag;
  CUDA_RUNTIME(cudaError_t(tzb_flags_signal(&l, plat.epoch(), st)));
  if (args_.request) mark_sent(*args_.request, tr_, plat.epoch(), st);
}

void Irecv::run(Platform &plat) {
  if (!args_.request) return;
  args_.request->kind = Request::Kind::Recv;
  args_.request->transport = tr_;
  args_.request->flag = args_.flag;
  args_.request->epoch = plat.epoch();
  args_.request->host = plat.host_execution();
}

void Wait::run(Platform &plat) {
  if ((plat.is_null() && !plat.host_execution()) || !request_) return;
  request_->wait();
}

void OwningWaitall::run(Platform &plat) {
  if (plat.is_null() && !plat.host_execution()) return;
  for (auto &r : reqs_) r->wait();
}

void MultiWait::run(Platform &plat) {
  if (plat.is_null() && !plat.host_execution()) return;
  for (auto &r : reqs_) r->wait();
}

void Ialltoallv::run(Platform &plat) {
  const size_t n = args_.peerBufs.size();
  if (plat.host_execution()) {
    for (size_t p = 0; p < n; ++p)
      if (args_.sendBytes[p])
        host::copy_bytes(args_.peerBufs[p],
                         static_cast<const char *>(args_.sendbuf) + args_.sendOffsets[p],
                         args_.sendBytes[p]);
    for (size_t p = 0; p < n; ++p)
      if (args_.peerFlags[p]) host::flag_signal(args_.peerFlags[p], plat.epoch());
    if (args_.request) {
      args_.request->kind = Request::Kind::Send;
      args_.request->host = true;
      args_.request->epoch = plat.epoch();
    }
    return;
  }
  if (tr_->is_null() || plat.is_null()) return;
  // one collective = one stream: its single request completes when the last copy and flag are out
  cudaStream_t st = tr_->next_send_stream();
  for (size_t p = 0; p < n; ++p) {
    if (args_.sendBytes[p])
      CUDA_RUNTIME(cudaMemcpyAsync(args_.peerBufs[p],
                                   static_cast<const char *>(args_.sendbuf) + args_.sendOffsets[p],
                                   args_.sendBytes[p], cudaMemcpyDeviceToDevice, st));
  }
  for (size_t p0 = 0; p0 < n; p0 += 32) {
    TzbFlagList l;
    l.n = 0;
    for (size_t p = p0; p < n && p < p0 + 32; ++p)
      if (args_.peerFlags[p]) l.flag[l.n++] = args_.peerFlags[p];
    CUDA_RUNTIME(cudaError_t(tzb_flags_signal(&l, plat.epoch(), st)));
  }
  if (args_.request) mark_sent(*args_.request, tr_, plat.epoch(), st);
}

} // namespace tenzing
