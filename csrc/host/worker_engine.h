// Host protocol engine of one allreduce worker.
//
// A message-for-message restatement of the reference worker's behaviour
// (AllreduceWorker.scala:7-301): init, round start, catch-up, fetch + scatter,
// threshold-triggered reduce and broadcast, threshold-triggered completion, flush to
// the sink, report to the master, ring advance, out-of-order completion, stale-message
// drop and future-round self-start.  It is the CPU backend, and the oracle the GPU
// path is validated against.  It knows nothing about transports: outgoing messages go
// through a `Mailer`, and delivery to an endpoint equal to `self` happens inline, which
// is what the reference does when a peer's ActorRef is the worker itself
// (AllreduceWorker.scala:228-229,260-261).
#pragma once
#include <deque>
#include <functional>
#include <memory>
#include <set>
#include <string>
#include <vector>

#include "aar/protocol_math.h"
#include "buffers.h"
#include "messages.h"

namespace aar {

// AllReduceInputRequest(iteration) => AllReduceInput(data)   (DataWrapper.scala:3-5)
using DataSource = std::function<std::vector<float>(int iteration)>;
// AllReduceOutput(data, count, iteration) => Unit            (DataWrapper.scala:7)
using DataSink = std::function<void(const std::vector<float>& data, const std::vector<int>& count,
                                    int iteration)>;

struct WorkerOptions {
  // Reference behaviour: any exception inside a handler is printed and swallowed
  // (AllreduceWorker.scala:287-299).  With `rethrow` the error propagates instead.
  bool rethrow = false;
  // Reference behaviour: a partition with fewer than P block starts crashes init
  // (n=9, P=4).  When false (default) trailing owners get empty blocks.
  bool strict_partition = false;
  bool verbose = false;
  // A message may name a round this far beyond the newest one started here, no further.  The catch-up path
  // force-completes every round in between (AllreduceWorker.scala:100-106), so a corrupt or hostile round number
  // would otherwise keep the handler busy for billions of iterations; a worker that is legitimately this far
  // behind has long been declared dead by the coordinator.  Messages beyond the bound are recorded as errors.
  int max_round_jump = 1 << 16;
};

class WorkerEngine {
 public:
  WorkerEngine(Endpoint self, DataSource source, DataSink sink, Mailer* mailer,
               WorkerOptions opts = {});

  void handle(const Message& m);

  // Introspection (tests, launchers, the device engine reuses the same bookkeeping).
  int id() const { return id_; }
  int round() const { return round_; }
  int max_round() const { return max_round_; }
  int max_scattered() const { return max_scattered_; }
  int peer_count() const { return static_cast<int>(peers_.size()); }
  const std::set<int>& completed() const { return completed_; }
  const std::vector<std::string>& errors() const { return errors_; }
  int scatter_required() const { return scatter_ ? scatter_->required() : 0; }
  int complete_required() const { return reduce_ ? reduce_->required() : 0; }
  const Partition& partition() const { return part_; }

 private:
  void on_init(const InitWorkers& m);
  void check_round(int round) const;  // throws for a round absurdly far ahead (WorkerOptions::max_round_jump)
  void on_start(const StartAllreduce& m);
  void on_scatter(const ScatterBlock& m);
  void on_reduce(const ReduceBlock& m);

  void fetch(int iteration);
  void scatter();
  void broadcast(const std::vector<float>& data, int chunk, int bcast_round, int count);
  void complete(int done_round, int row);
  void post(int peer_id, Message msg, bool scatter_kind);
  void guarded(const char* where, const std::function<void()>& body);

  Endpoint self_;
  DataSource source_;
  DataSink sink_;
  Mailer* mailer_;
  WorkerOptions opts_;

  int id_ = -1;
  Endpoint master_ = kNoEndpoint;
  std::map<int, Endpoint> peers_;
  int peer_num_ = 0;
  float th_reduce_ = 1.f, th_complete_ = 1.f;
  int max_lag_ = 0;
  int round_ = -1;
  int max_round_ = -1;
  int max_scattered_ = -1;
  std::set<int> completed_;

  int64_t data_size_ = 0;
  int64_t max_chunk_ = 1024;
  std::vector<float> data_;
  Partition part_;
  std::unique_ptr<ScatterBuffer> scatter_;
  std::unique_ptr<ReduceBuffer> reduce_;

  std::deque<Message> before_init_;  // held until InitWorkers arrives, then replayed in order
  std::vector<std::string> errors_;
};

}  // namespace aar
