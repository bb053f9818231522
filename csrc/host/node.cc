#include "node.h"

#include <arpa/inet.h>
#include <fcntl.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <sys/eventfd.h>
#include <sys/socket.h>
#include <unistd.h>

#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <stdexcept>

namespace aar {

using Clock = std::chrono::steady_clock;

// ------------------------------------------------------------------ byte helpers
namespace {

struct Writer {
  std::vector<uint8_t> b;
  template <class T>
  void put(T v) {
    const auto* p = reinterpret_cast<const uint8_t*>(&v);
    b.insert(b.end(), p, p + sizeof(T));
  }
  void str(const std::string& s) {
    put<uint32_t>(static_cast<uint32_t>(s.size()));
    b.insert(b.end(), s.begin(), s.end());
  }
  void floats(const std::vector<float>& v) {
    put<uint32_t>(static_cast<uint32_t>(v.size()));
    const auto* p = reinterpret_cast<const uint8_t*>(v.data());
    b.insert(b.end(), p, p + v.size() * sizeof(float));
  }
};

// Largest frame a node accepts: a chunk of 2^28 floats plus its header.  (Akka remoting, which the
// reference runs on, has the same kind of limit: maximum-frame-size, 128 kB unless configured.)
constexpr uint32_t kMaxFrameBytes = (1u << 30) + 64;

struct Reader {
  const uint8_t* p;
  size_t n;
  template <class T>
  T get() {
    if (n < sizeof(T)) throw std::runtime_error("short frame");
    T v;
    std::memcpy(&v, p, sizeof(T));
    p += sizeof(T);
    n -= sizeof(T);
    return v;
  }
  std::string str() {
    const uint32_t len = get<uint32_t>();
    if (n < len) throw std::runtime_error("short frame (string)");
    std::string s(reinterpret_cast<const char*>(p), len);
    p += len;
    n -= len;
    return s;
  }
  std::vector<float> floats() {
    const uint32_t len = get<uint32_t>();
    if (n < len * sizeof(float)) throw std::runtime_error("short frame (floats)");
    std::vector<float> v(len);
    std::memcpy(v.data(), p, len * sizeof(float));
    p += len * sizeof(float);
    n -= len * sizeof(float);
    return v;
  }
};

std::vector<uint8_t> encode_scatter(const ScatterBlock& m) {
  Writer w;
  w.put<int32_t>(m.src_id);
  w.put<int32_t>(m.dest_id);
  w.put<int32_t>(m.chunk_id);
  w.put<int32_t>(m.round);
  w.floats(m.value);
  return std::move(w.b);
}
ScatterBlock decode_scatter(Reader r) {
  ScatterBlock m;
  m.src_id = r.get<int32_t>();
  m.dest_id = r.get<int32_t>();
  m.chunk_id = r.get<int32_t>();
  m.round = r.get<int32_t>();
  m.value = r.floats();
  return m;
}
std::vector<uint8_t> encode_reduce(const ReduceBlock& m) {
  Writer w;
  w.put<int32_t>(m.src_id);
  w.put<int32_t>(m.dest_id);
  w.put<int32_t>(m.chunk_id);
  w.put<int32_t>(m.round);
  w.put<int32_t>(m.count);
  w.floats(m.value);
  return std::move(w.b);
}
ReduceBlock decode_reduce(Reader r) {
  ReduceBlock m;
  m.src_id = r.get<int32_t>();
  m.dest_id = r.get<int32_t>();
  m.chunk_id = r.get<int32_t>();
  m.round = r.get<int32_t>();
  m.count = r.get<int32_t>();
  m.value = r.floats();
  return m;
}

void set_nonblocking(int fd) {
  const int fl = fcntl(fd, F_GETFL, 0);
  fcntl(fd, F_SETFL, fl | O_NONBLOCK);
  const int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
}

constexpr Endpoint kSelfEp = 1;
constexpr Endpoint kMasterEp = 2;
constexpr Endpoint kPeerBase = 100;
constexpr Endpoint kConnBase = 1000;

}  // namespace

// ------------------------------------------------------------------ TcpNode
TcpNode::TcpNode() { wake_fd_ = eventfd(0, EFD_NONBLOCK | EFD_CLOEXEC); }

TcpNode::~TcpNode() {
  stop();
  if (wake_fd_ >= 0) close(wake_fd_);
}

int TcpNode::listen_on(const std::string& host, int port) {
  listen_host_ = host;
  listen_fd_ = socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
  if (listen_fd_ < 0) throw std::runtime_error("socket() failed");
  const int one = 1;
  setsockopt(listen_fd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
  sockaddr_in sa{};
  sa.sin_family = AF_INET;
  sa.sin_port = htons(static_cast<uint16_t>(port));
  if (inet_pton(AF_INET, host.c_str(), &sa.sin_addr) != 1) throw std::runtime_error("bad host " + host);
  if (bind(listen_fd_, reinterpret_cast<sockaddr*>(&sa), sizeof(sa)) != 0)
    throw std::runtime_error("bind " + host + ":" + std::to_string(port) + " failed: " + std::strerror(errno));
  if (::listen(listen_fd_, 128) != 0) throw std::runtime_error("listen failed");
  socklen_t len = sizeof(sa);
  getsockname(listen_fd_, reinterpret_cast<sockaddr*>(&sa), &len);
  listen_port_ = ntohs(sa.sin_port);
  fcntl(listen_fd_, F_SETFL, fcntl(listen_fd_, F_GETFL, 0) | O_NONBLOCK);
  return listen_port_;
}

void TcpNode::start() {
  if (running_.exchange(true)) return;
  stop_flag_.store(false);
  thread_ = std::thread([this] { io_loop(); });
}

void TcpNode::stop() {
  if (running_.load()) {
    stop_flag_.store(true);
    wake();
    if (thread_.joinable() && std::this_thread::get_id() != thread_.get_id()) thread_.join();
    running_.store(false);
  }
  // Orderly close.  close() on a socket that still has UNREAD input makes the kernel send a reset instead of a FIN,
  // and a reset lets the peer's stack throw away what it has received from us but not yet read — e.g. the last
  // completion reports of a worker that leaves while StartAllreduce frames it never looked at are still queued.
  // So: half-close (our data, then FIN), read and discard until the peer closes too or a short grace period ends.
  for (auto& [id, c] : conns_)
    if (c.fd >= 0) shutdown(c.fd, SHUT_WR);
  const auto grace_end = Clock::now() + std::chrono::milliseconds(300);
  for (auto& [id, c] : conns_) {
    if (c.fd < 0) continue;
    const int fl = fcntl(c.fd, F_GETFL, 0);
    fcntl(c.fd, F_SETFL, fl | O_NONBLOCK);
    uint8_t sink[4096];
    for (;;) {
      const ssize_t got = ::read(c.fd, sink, sizeof(sink));
      if (got > 0) continue;
      if (got == 0 || (errno != EAGAIN && errno != EWOULDBLOCK && errno != EINTR)) break;  // peer closed / error
      const auto left = std::chrono::duration_cast<std::chrono::milliseconds>(grace_end - Clock::now()).count();
      if (left <= 0) break;
      pollfd pf{c.fd, POLLIN, 0};
      poll(&pf, 1, static_cast<int>(left));
    }
    close(c.fd);
  }
  conns_.clear();
  if (listen_fd_ >= 0) {
    close(listen_fd_);
    listen_fd_ = -1;
  }
}

void TcpNode::wake() {
  const uint64_t one = 1;
  if (wake_fd_ >= 0) (void)!write(wake_fd_, &one, sizeof(one));
}

TcpNode::Conn* TcpNode::conn(int id) {
  auto it = conns_.find(id);
  return it == conns_.end() ? nullptr : &it->second;
}

int TcpNode::connect_to(const PeerAddress& a, int retries) {
  const int fd = socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
  if (fd < 0) throw std::runtime_error("socket() failed");
  sockaddr_in sa{};
  sa.sin_family = AF_INET;
  sa.sin_port = htons(static_cast<uint16_t>(a.port));
  inet_pton(AF_INET, a.host.c_str(), &sa.sin_addr);
  int rc = -1;
  for (int attempt = 0; attempt <= retries; ++attempt) {  // 50 ms apart
    rc = ::connect(fd, reinterpret_cast<sockaddr*>(&sa), sizeof(sa));
    if (rc == 0 || (errno != ECONNREFUSED && errno != EINTR) || attempt == retries) break;
    std::this_thread::sleep_for(std::chrono::milliseconds(50));
  }
  if (rc != 0) {
    close(fd);
    throw std::runtime_error("connect " + a.host + ":" + std::to_string(a.port) + " failed: " +
                             std::strerror(errno));
  }
  set_nonblocking(fd);
  const int id = next_conn_++;
  Conn& c = conns_[id];
  c.fd = fd;
  c.last_seen = Clock::now();
  return id;
}

void TcpNode::queue_frame(int conn_id, Frame kind, const std::vector<uint8_t>& payload) {
  std::vector<uint8_t> f(5 + payload.size());
  const uint32_t n = static_cast<uint32_t>(payload.size());
  std::memcpy(f.data(), &n, 4);
  f[4] = static_cast<uint8_t>(kind);
  if (n) std::memcpy(f.data() + 5, payload.data(), n);
  {
    std::lock_guard<std::mutex> g(cmd_mu_);
    cmd_frames_.emplace_back(conn_id, std::move(f));
  }
  // The IO thread drains the queue itself before it polls again: only other threads need to wake it.
  if (std::this_thread::get_id() != io_thread_id_) wake();
}

void TcpNode::queue_local(Message m) {
  {
    std::lock_guard<std::mutex> g(cmd_mu_);
    cmd_local_.push_back(std::move(m));
  }
  if (std::this_thread::get_id() != io_thread_id_) wake();
}

void TcpNode::close_conn(int conn_id) {
  if (Conn* c = conn(conn_id)) c->closing = true;
}

void TcpNode::drain_commands() {
  std::deque<std::pair<int, std::vector<uint8_t>>> frames;
  {
    std::lock_guard<std::mutex> g(cmd_mu_);
    frames.swap(cmd_frames_);
  }
  for (auto& [id, bytes] : frames)
    if (Conn* c = conn(id)) c->out.push_back(std::move(bytes));
}

bool TcpNode::pump_reads(int id, Conn& c) {
  uint8_t buf[1 << 16];
  for (;;) {
    const ssize_t got = ::read(c.fd, buf, sizeof(buf));
    if (got > 0) {
      c.in.insert(c.in.end(), buf, buf + got);
      c.last_seen = Clock::now();
      if (got < static_cast<ssize_t>(sizeof(buf))) break;
    } else if (got == 0) {
      return false;  // peer closed
    } else {
      if (errno == EAGAIN || errno == EWOULDBLOCK) break;
      if (errno == EINTR) continue;
      return false;
    }
  }
  size_t off = 0;
  while (c.in.size() - off >= 5) {
    uint32_t n;
    std::memcpy(&n, c.in.data() + off, 4);
    if (n > kMaxFrameBytes) {  // not one of ours (or corrupt): do not buffer gigabytes waiting for it
      std::fprintf(stderr, "[aar node] frame of %u bytes announced on connection %d: closing it\n", n, id);
      return false;
    }
    if (c.in.size() - off < 5 + static_cast<size_t>(n)) break;
    const Frame kind = static_cast<Frame>(c.in[off + 4]);
    try {
      on_frame(id, kind, c.in.data() + off + 5, n);
    } catch (const std::exception& e) {
      std::fprintf(stderr, "[aar node] frame kind %d dropped: %s\n", static_cast<int>(kind), e.what());
    }
    off += 5 + n;
  }
  if (off) c.in.erase(c.in.begin(), c.in.begin() + off);
  return true;
}

void TcpNode::pump_writes(Conn& c) {
  while (!c.out.empty()) {
    auto& front = c.out.front();
    const ssize_t put = ::send(c.fd, front.data() + c.out_off, front.size() - c.out_off, MSG_NOSIGNAL);
    if (put > 0) {
      c.out_off += static_cast<size_t>(put);
      if (c.out_off == front.size()) {
        c.out.pop_front();
        c.out_off = 0;
      }
    } else {
      if (errno == EAGAIN || errno == EWOULDBLOCK) return;
      if (errno == EINTR) continue;
      c.closing = true;
      return;
    }
  }
}

void TcpNode::io_loop() {
  io_thread_id_ = std::this_thread::get_id();
  auto last_tick = Clock::now();
  std::vector<pollfd> fds;
  std::vector<int> ids;
  while (!stop_flag_.load()) {
    drain_commands();
    // Local (self-addressed) messages are handled between socket events, which is the
    // back-of-the-mailbox position `self ! msg` has in the reference.
    for (;;) {
      Message m;
      {
        std::lock_guard<std::mutex> g(cmd_mu_);
        if (cmd_local_.empty()) break;
        m = std::move(cmd_local_.front());
        cmd_local_.pop_front();
      }
      on_local(m);
    }
    drain_commands();
    for (auto& [id, c] : conns_) pump_writes(c);

    fds.clear();
    ids.clear();
    fds.push_back({wake_fd_, POLLIN, 0});
    ids.push_back(-1);
    if (listen_fd_ >= 0) {
      fds.push_back({listen_fd_, POLLIN, 0});
      ids.push_back(-2);
    }
    for (auto& [id, c] : conns_) {
      short ev = POLLIN;
      if (!c.out.empty()) ev |= POLLOUT;
      fds.push_back({c.fd, ev, 0});
      ids.push_back(id);
    }
    bool local_pending;
    {
      std::lock_guard<std::mutex> g(cmd_mu_);
      local_pending = !cmd_local_.empty() || !cmd_frames_.empty();
    }
    poll(fds.data(), fds.size(), local_pending ? 0 : 50);

    for (size_t i = 0; i < fds.size(); ++i) {
      if (!fds[i].revents) continue;
      if (ids[i] == -1) {
        uint64_t junk;
        while (read(wake_fd_, &junk, sizeof(junk)) > 0) {}
      } else if (ids[i] == -2) {
        for (;;) {
          const int fd = accept4(listen_fd_, nullptr, nullptr, SOCK_CLOEXEC);
          if (fd < 0) break;
          set_nonblocking(fd);
          const int id = next_conn_++;
          Conn& c = conns_[id];
          c.fd = fd;
          c.last_seen = Clock::now();
        }
      } else {
        Conn* c = conn(ids[i]);
        if (!c) continue;
        if (fds[i].revents & (POLLIN | POLLHUP | POLLERR))
          if (!pump_reads(ids[i], *c)) c->closing = true;
        if (fds[i].revents & POLLOUT) pump_writes(*c);
      }
    }
    for (auto it = conns_.begin(); it != conns_.end();) {
      if (it->second.closing) {
        const int id = it->first;
        close(it->second.fd);
        it = conns_.erase(it);
        on_closed(id);
      } else {
        ++it;
      }
    }
    const auto now = Clock::now();
    if (now - last_tick > std::chrono::milliseconds(100)) {
      last_tick = now;
      on_tick();
    }
  }
  // Flush what is still queued (e.g. Shutdown frames) before leaving.
  drain_commands();
  for (auto& [id, c] : conns_) {
    const int fl = fcntl(c.fd, F_GETFL, 0);
    fcntl(c.fd, F_SETFL, fl & ~O_NONBLOCK);
    pump_writes(c);
  }
}

// ------------------------------------------------------------------ MasterNode
MasterNode::MasterNode(MasterConfig cfg, MasterNodeOptions opts)
    : engine_(kSelfEp, cfg, this, opts.verbose), opts_(opts) {}

void MasterNode::send(Endpoint to, Message msg) {
  const int conn_id = static_cast<int>(to - kConnBase);
  switch (kind_of(msg)) {
    case MsgKind::Init: {
      const auto& x = std::get<InitWorkers>(msg);
      Writer w;
      w.put<int32_t>(x.worker_num);
      w.put<int32_t>(x.dest_id);
      w.put<float>(x.th_reduce);
      w.put<float>(x.th_complete);
      w.put<int32_t>(x.max_lag);
      w.put<int64_t>(x.data_size);
      w.put<int64_t>(x.max_chunk_size);
      w.put<int32_t>(x.start_round);
      w.put<int32_t>(x.max_round);
      w.put<uint32_t>(static_cast<uint32_t>(x.workers.size()));
      for (const auto& [id, ep] : x.workers) {
        const PeerAddress& a = addr_of_conn_[static_cast<int>(ep - kConnBase)];
        w.put<int32_t>(id);
        w.put<int32_t>(a.port);
        w.str(a.host);
      }
      queue_frame(conn_id, Frame::Init, w.b);
      break;
    }
    case MsgKind::Start: {
      Writer w;
      w.put<int32_t>(std::get<StartAllreduce>(msg).round);
      queue_frame(conn_id, Frame::Start, w.b);
      break;
    }
    default:
      break;  // the master originates nothing else
  }
}

void MasterNode::note_progress() {
  std::lock_guard<std::mutex> g(state_mu_);
  round_snapshot_ = engine_.round();
  alive_ = static_cast<int>(engine_.workers().size());
  if (engine_.finished() && !finished_) {
    finished_ = true;
    if (opts_.shutdown_workers_on_finish)
      for (const auto& [id, ep] : engine_.workers())
        queue_frame(static_cast<int>(ep - kConnBase), Frame::Shutdown, {});
  }
  state_cv_.notify_all();
}

void MasterNode::on_frame(int conn_id, Frame kind, const uint8_t* p, size_t n) {
  Reader r{p, n};
  switch (kind) {
    case Frame::Register: {
      PeerAddress a;
      a.port = r.get<int32_t>();
      a.host = r.str();
      addr_of_conn_[conn_id] = a;
      if (Conn* c = conn(conn_id)) c->tag = kConnBase + conn_id;
      {
        std::lock_guard<std::mutex> g(state_mu_);
        registered_ += 1;
      }
      const int id = engine_.member_up(kConnBase + conn_id);
      if (opts_.verbose)
        std::printf("----Detect member %s:%d up, id %d\n", a.host.c_str(), a.port, id);
      note_progress();
      break;
    }
    case Frame::Complete: {
      CompleteAllreduce c;
      c.src_id = r.get<int32_t>();
      c.round = r.get<int32_t>();
      // a completion report counts towards the thAllreduce gate: only the registered worker holding that id may
      // send it (a stray connection must not be able to open rounds)
      const auto owner = engine_.workers().find(c.src_id);
      if (owner == engine_.workers().end() || owner->second != kConnBase + conn_id)
        throw std::runtime_error("completion report for worker " + std::to_string(c.src_id) +
                                 " from a connection that does not hold that id");
      engine_.handle(c);
      note_progress();
      break;
    }
    case Frame::Heartbeat:
    default:
      break;
  }
}

void MasterNode::on_closed(int conn_id) {
  if (addr_of_conn_.erase(conn_id)) {
    if (opts_.verbose) std::printf("----worker on connection %d is terminated, removing it from the set\n", conn_id);
    engine_.terminated(kConnBase + conn_id);
    note_progress();
  }
}

void MasterNode::on_tick() {
  const auto now = Clock::now();
  for (const auto& [conn_id, addr] : addr_of_conn_) {
    Conn* c = conn(conn_id);
    if (!c) continue;
    const double quiet = std::chrono::duration<double>(now - c->last_seen).count();
    if (quiet > opts_.down_after_s) close_conn(conn_id);  // unreachable for too long: down it
  }
}

bool MasterNode::wait_finished(double timeout_s) {
  std::unique_lock<std::mutex> g(state_mu_);
  return state_cv_.wait_for(g, std::chrono::duration<double>(timeout_s), [&] { return finished_; });
}
int MasterNode::round() {
  std::lock_guard<std::mutex> g(state_mu_);
  return round_snapshot_;
}
int MasterNode::registered() {
  std::lock_guard<std::mutex> g(state_mu_);
  return registered_;
}
int MasterNode::alive() {
  std::lock_guard<std::mutex> g(state_mu_);
  return alive_;
}

// ------------------------------------------------------------------ WorkerNode
WorkerNode::WorkerNode(PeerAddress master, DataSource source, DataSink sink, WorkerNodeOptions opts)
    : master_addr_(std::move(master)), opts_(opts) {
  if (source && sink) {
    DataSink counted = [this, sink](const std::vector<float>& d, const std::vector<int>& c, int it) {
      sink(d, c, it);
      rounds_done_.fetch_add(1);
    };
    engine_ = std::make_unique<WorkerEngine>(kSelfEp, std::move(source), std::move(counted), this, opts.engine);
  }
}

void WorkerNode::begin() {
  if (listen_port_ == 0) listen_on("127.0.0.1", 0);
  master_conn_ = connect_to(master_addr_, 400);  // the master may still be starting: retry ~20 s
  Writer w;
  w.put<int32_t>(listen_port_);
  w.str(listen_host_);
  queue_frame(master_conn_, Frame::Register, w.b);
  last_beat_ = Clock::now();
  start();
}

int WorkerNode::conn_for_peer(int peer_id) {
  auto it = peer_conn_.find(peer_id);
  if (it != peer_conn_.end() && conn(it->second)) return it->second;
  if (dead_peers_.count(peer_id)) throw std::runtime_error("peer is down");
  int id;
  try {
    // Peers listen before they register, so a refused connection means the peer is gone.
    id = connect_to(peer_addr_.at(peer_id), 0);
  } catch (const std::exception&) {
    dead_peers_.insert(peer_id);
    throw;
  }
  peer_conn_[peer_id] = id;
  Writer w;
  w.put<int32_t>(worker_id_.load());
  queue_frame(id, Frame::PeerHello, w.b);
  return id;
}

void WorkerNode::send(Endpoint to, Message msg) {
  if (to == kSelfEp) {
    queue_local(std::move(msg));
    return;
  }
  if (to == kMasterEp) {
    if (const auto* c = std::get_if<CompleteAllreduce>(&msg)) {
      Writer w;
      w.put<int32_t>(c->src_id);
      w.put<int32_t>(c->round);
      queue_frame(master_conn_, Frame::Complete, w.b);
    }
    return;
  }
  const int peer = static_cast<int>(to - kPeerBase);
  int cid;
  try {
    cid = conn_for_peer(peer);
  } catch (const std::exception&) {
    return;  // unreachable peer: at-most-once delivery, the thresholds absorb it
  }
  if (const auto* s = std::get_if<ScatterBlock>(&msg)) queue_frame(cid, Frame::Scatter, encode_scatter(*s));
  else if (const auto* r = std::get_if<ReduceBlock>(&msg)) queue_frame(cid, Frame::Reduce, encode_reduce(*r));
}

void WorkerNode::report_complete(int src_id, int round) {
  Writer w;
  w.put<int32_t>(src_id);
  w.put<int32_t>(round);
  queue_frame(master_conn_, Frame::Complete, w.b);
  rounds_done_.fetch_add(1);
}

void WorkerNode::deliver(const Message& m) {
  if (handler_) handler_(m);
  else if (engine_) engine_->handle(m);
}

void WorkerNode::on_local(const Message& m) { deliver(m); }

void WorkerNode::on_frame(int conn_id, Frame kind, const uint8_t* p, size_t n) {
  Reader r{p, n};
  // Control frames are the coordinator's: a connection that is not the one this worker opened to it cannot
  // re-initialise the worker, start rounds or shut it down.
  if ((kind == Frame::Init || kind == Frame::Start || kind == Frame::Shutdown) && conn_id != master_conn_)
    throw std::runtime_error("control frame from a connection that is not the coordinator's");
  switch (kind) {
    case Frame::Init: {
      InitWorkers x;
      x.worker_num = r.get<int32_t>();
      x.dest_id = r.get<int32_t>();
      x.th_reduce = r.get<float>();
      x.th_complete = r.get<float>();
      x.max_lag = r.get<int32_t>();
      x.data_size = r.get<int64_t>();
      x.max_chunk_size = r.get<int64_t>();
      x.start_round = r.get<int32_t>();
      x.max_round = r.get<int32_t>();
      x.master = kMasterEp;
      const uint32_t peers = r.get<uint32_t>();
      const PeerAddress me{listen_host_, listen_port_};
      for (uint32_t i = 0; i < peers; ++i) {
        const int id = r.get<int32_t>();
        PeerAddress a;
        a.port = r.get<int32_t>();
        a.host = r.str();
        auto known = peer_addr_.find(id);
        if (known == peer_addr_.end() || !(known->second == a)) {  // new or replaced peer: forget the old route
          dead_peers_.erase(id);
          auto pc = peer_conn_.find(id);
          if (pc != peer_conn_.end()) {
            close_conn(pc->second);
            peer_conn_.erase(pc);
          }
        }
        peer_addr_[id] = a;
        x.workers[id] = (a == me) ? kSelfEp : kPeerBase + id;
      }
      if (worker_id_.load() == -1) worker_id_.store(x.dest_id);
      deliver(x);
      break;
    }
    case Frame::Start:
      deliver(StartAllreduce{r.get<int32_t>()});
      break;
    case Frame::Scatter:
      deliver(decode_scatter(r));
      break;
    case Frame::Reduce:
      deliver(decode_reduce(r));
      break;
    case Frame::Shutdown: {
      std::lock_guard<std::mutex> g(state_mu_);
      shutdown_ = true;
      state_cv_.notify_all();
      break;
    }
    case Frame::PeerHello:
    default:
      break;
  }
}

void WorkerNode::on_closed(int conn_id) {
  if (conn_id == master_conn_) {  // coordinator gone: nothing left to gate rounds
    std::lock_guard<std::mutex> g(state_mu_);
    shutdown_ = true;
    state_cv_.notify_all();
  }
  for (auto it = peer_conn_.begin(); it != peer_conn_.end();) {
    if (it->second == conn_id) {
      dead_peers_.insert(it->first);  // messages to it become dead letters from now on
      it = peer_conn_.erase(it);
    } else {
      ++it;
    }
  }
}

void WorkerNode::on_tick() {
  const auto now = Clock::now();
  if (std::chrono::duration<double>(now - last_beat_).count() >= opts_.heartbeat_s) {
    last_beat_ = now;
    queue_frame(master_conn_, Frame::Heartbeat, {});
  }
}

bool WorkerNode::wait_shutdown(double timeout_s) {
  std::unique_lock<std::mutex> g(state_mu_);
  return state_cv_.wait_for(g, std::chrono::duration<double>(timeout_s), [&] { return shutdown_; });
}

// ------------------------------------------------------------------ built-in source/sink
DataSource make_ramp_source(int64_t data_size) {
  auto ramp = std::make_shared<std::vector<float>>(static_cast<size_t>(data_size));
  for (int64_t i = 0; i < data_size; ++i) (*ramp)[static_cast<size_t>(i)] = static_cast<float>(i);
  return [ramp](int) { return *ramp; };
}

DataSink make_throughput_sink(int64_t data_size, int checkpoint, int assert_multiple,
                              std::shared_ptr<ThroughputStats> stats) {
  auto tic = std::make_shared<Clock::time_point>(Clock::now());
  return [=](const std::vector<float>& data, const std::vector<int>& count, int iteration) {
    if (checkpoint <= 0 || iteration % checkpoint != 0 || iteration == 0) return;
    const double elapsed = std::chrono::duration<double>(Clock::now() - *tic).count();
    const double mbytes = static_cast<double>(data.size()) * 4.0 * checkpoint / 1.0e6;
    std::printf("----Data output at #%d - %g s\n", iteration, elapsed);
    std::printf("%2.1f Mbytes in %2.1f seconds at %4.3f MBytes/sec\n", mbytes, elapsed, mbytes / elapsed);
    std::fflush(stdout);
    if (stats) {
      stats->checkpoints.fetch_add(1);
      stats->last_mbytes_per_s.store(mbytes / elapsed);
    }
    if (assert_multiple > 0) {
      bool ok = static_cast<int64_t>(data.size()) == data_size;
      for (size_t i = 0; ok && i < data.size(); ++i)
        ok = data[i] == static_cast<float>(i) * assert_multiple && count[i] == assert_multiple;
      if (!ok) {
        if (stats) stats->failures.fetch_add(1);
        std::fprintf(stderr,
                     "Expected output should be multiple of input, and counts equal to the multiple. "
                     "Check if all thresholds are 1 (iteration %d)\n",
                     iteration);
      }
    }
    *tic = Clock::now();
  };
}

}  // namespace aar
