// TCP control plane and CPU data plane: the coordinator ("master node") and the worker
// processes that talk to it and to each other.
//
// This is the stand-in for the reference's L0 layer — Akka remoting over Netty TCP plus
// cluster membership (src/main/resources/application.conf:1-34, AllreduceMaster.scala:
// 30-52,66-74): workers register with the coordinator on a configured port (MemberUp),
// get ids in arrival order, receive one InitWorkers carrying every peer's address, and
// then exchange ScatterBlock/ReduceBlock frames directly, all-to-all, with per-pair FIFO
// order.  A closed connection or `down_after` seconds of missed heartbeats removes the
// worker (Terminated / auto-down-unreachable-after).
//
// Every node runs one poll()-driven IO thread; engines are only ever touched from that
// thread, which gives them the single-threaded mailbox discipline of an actor.
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "master_engine.h"
#include "messages.h"
#include "worker_engine.h"

namespace pybind11 {
class module_;
}

namespace aar {

struct PeerAddress {
  std::string host;
  int port = 0;
  bool operator==(const PeerAddress& o) const { return port == o.port && host == o.host; }
};

// Frame kinds on the wire: 0..4 are MsgKind, the rest is membership plumbing.
enum class Frame : uint8_t {
  Init = 0, Start = 1, Scatter = 2, Reduce = 3, Complete = 4,
  Register = 16, Heartbeat = 17, PeerHello = 18, Shutdown = 19,
};

class TcpNode : public Mailer {
 public:
  TcpNode();
  ~TcpNode() override;
  TcpNode(const TcpNode&) = delete;

  // Binds 127.0.0.1:`port` (0 = ephemeral) and returns the bound port.
  int listen_on(const std::string& host, int port);
  void start();  // spawns the IO thread
  void stop();   // joins it, closes sockets
  bool running() const { return running_.load(); }
  int port() const { return listen_port_; }

 protected:
  struct Conn {
    int fd = -1;
    std::vector<uint8_t> in;
    std::deque<std::vector<uint8_t>> out;
    size_t out_off = 0;
    std::chrono::steady_clock::time_point last_seen;
    Endpoint tag = kNoEndpoint;  // set by subclasses once the peer has identified itself
    bool closing = false;
  };

  // --- called on the IO thread ---
  virtual void on_frame(int conn_id, Frame kind, const uint8_t* p, size_t n) = 0;
  virtual void on_closed(int /*conn_id*/) {}
  virtual void on_tick() {}
  virtual void on_local(const Message& /*m*/) {}  // message an engine sent to itself

  int connect_to(const PeerAddress& a, int retries = 0);  // returns conn id (blocking connect)
  void queue_frame(int conn_id, Frame kind, const std::vector<uint8_t>& payload);  // any thread
  void queue_local(Message m);                                                     // any thread
  void close_conn(int conn_id);
  Conn* conn(int id);
  void request_stop() { stop_flag_.store(true); wake(); }

  std::string listen_host_ = "127.0.0.1";
  int listen_port_ = 0;

 private:
  void io_loop();
  void wake();
  void drain_commands();
  bool pump_reads(int id, Conn& c);
  void pump_writes(Conn& c);

  int listen_fd_ = -1;
  int wake_fd_ = -1;
  std::thread thread_;
  std::thread::id io_thread_id_{};
  std::atomic<bool> running_{false}, stop_flag_{false};
  std::map<int, Conn> conns_;
  int next_conn_ = 1;
  std::mutex cmd_mu_;
  std::deque<std::pair<int, std::vector<uint8_t>>> cmd_frames_;
  std::deque<Message> cmd_local_;
};

struct MasterNodeOptions {
  double down_after_s = 10.0;  // application.conf: auto-down-unreachable-after = 10s
  bool shutdown_workers_on_finish = true;
  bool verbose = false;
};

// Rank-0 coordinator process: MasterEngine behind a TCP listener.
class MasterNode : public TcpNode {
 public:
  MasterNode(MasterConfig cfg, MasterNodeOptions opts = {});
  ~MasterNode() override { stop(); }
  void send(Endpoint to, Message msg) override;
  bool wait_finished(double timeout_s);
  int round();
  int registered();
  int alive();

 protected:
  void on_frame(int conn_id, Frame kind, const uint8_t* p, size_t n) override;
  void on_closed(int conn_id) override;
  void on_tick() override;

 private:
  void note_progress();
  MasterEngine engine_;
  MasterNodeOptions opts_;
  std::map<int, PeerAddress> addr_of_conn_;  // registered workers' listen addresses
  std::mutex state_mu_;
  std::condition_variable state_cv_;
  bool finished_ = false;
  int round_snapshot_ = -1, registered_ = 0, alive_ = 0;
};

struct WorkerNodeOptions {
  double heartbeat_s = 1.0;
  WorkerOptions engine;
};

// A worker process.  Either hosts a WorkerEngine (CPU backend: the whole protocol runs
// over TCP) or, with `set_handler`, forwards control messages to a caller-supplied
// handler (GPU backend: the data plane lives in symmetric memory and only
// Init/Start/Complete travel through the coordinator).
class WorkerNode : public TcpNode {
 public:
  WorkerNode(PeerAddress master, DataSource source, DataSink sink, WorkerNodeOptions opts = {});
  ~WorkerNode() override { stop(); }
  using Handler = std::function<void(const Message&)>;
  void set_handler(Handler h) { handler_ = std::move(h); }

  void send(Endpoint to, Message msg) override;
  void report_complete(int src_id, int round);  // any thread (GPU backend)
  bool wait_shutdown(double timeout_s);
  WorkerEngine* engine() { return engine_.get(); }
  int worker_id() const { return worker_id_.load(); }
  int rounds_completed() const { return rounds_done_.load(); }

  void begin();  // listen + start + register with the master

 protected:
  void on_frame(int conn_id, Frame kind, const uint8_t* p, size_t n) override;
  void on_closed(int conn_id) override;
  void on_tick() override;
  void on_local(const Message& m) override;

 private:
  void deliver(const Message& m);
  int conn_for_peer(int peer_id);

  PeerAddress master_addr_;
  WorkerNodeOptions opts_;
  std::unique_ptr<WorkerEngine> engine_;
  Handler handler_;
  int master_conn_ = -1;
  std::map<int, PeerAddress> peer_addr_;
  std::map<int, int> peer_conn_;
  std::set<int> dead_peers_;
  std::atomic<int> worker_id_{-1};
  std::atomic<int> rounds_done_{0};
  std::chrono::steady_clock::time_point last_beat_;
  std::mutex state_mu_;
  std::condition_variable state_cv_;
  bool shutdown_ = false;
};

// Ramp source / throughput sink of the reference's worker launcher
// (AllreduceWorker.scala:324-343): source [0..n-1]; every `checkpoint` iterations the
// sink prints elapsed seconds and MBytes/sec, and with assert_multiple > 0 checks
// output == input*k and count == k.
DataSource make_ramp_source(int64_t data_size);
struct ThroughputStats {
  std::atomic<int> checkpoints{0};
  std::atomic<int> failures{0};
  std::atomic<double> last_mbytes_per_s{0.0};
};
DataSink make_throughput_sink(int64_t data_size, int checkpoint, int assert_multiple,
                              std::shared_ptr<ThroughputStats> stats);

void bind_nodes(pybind11::module_& m);

}  // namespace aar
