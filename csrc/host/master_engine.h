// Round-gating master (control plane), transport-agnostic.
//
// Behaviour of the reference's AllreduceMaster actor (AllreduceMaster.scala:12-90):
// waits for `total_workers` registrations, hands out ids in arrival order, sends one
// InitWorkers per worker, starts round 0, and starts round r+1 once
// numComplete >= totalWorkers*thAllreduce (fp32) reports for round r are in, up to and
// including `max_round`.  Terminated workers leave the map but the gate keeps using the
// configured total (AllreduceMaster.scala:46-52,58).
#pragma once
#include <map>

#include "messages.h"

namespace aar {

struct MasterConfig {
  int total_workers = 2;
  float th_allreduce = 1.f;
  float th_reduce = 1.f;
  float th_complete = 0.8f;
  int max_lag = 1;
  int64_t data_size = 10;
  int max_round = 100;
  int64_t max_chunk_size = 2;
  // Extension: when a worker is removed from a RUNNING job, send the survivors a refreshed InitWorkers (they
  // treat a second InitWorkers as a peer-map refresh, AllreduceWorker.scala:87-89).  The reference leaves the
  // stale ActorRef in every worker's map and lets the sends turn into dead letters; a GPU worker has no dead
  // letters — it would push into, and wait on, the memory of a process that is gone — so the CUDA backend
  // needs to be told.  Off by default: the engine then behaves message for message like the reference.
  bool announce_removals = false;
};

class MasterEngine {
 public:
  MasterEngine(Endpoint self, MasterConfig cfg, Mailer* mailer, bool verbose = false);

  // MemberUp with role "worker" (AllreduceMaster.scala:36-44,66-74). Returns the new id.
  int member_up(Endpoint worker);
  // Deathwatch notification (AllreduceMaster.scala:46-52).
  void terminated(Endpoint worker);
  void handle(const Message& m);  // CompleteAllreduce

  int round() const { return round_; }
  int num_complete() const { return num_complete_; }
  const std::map<int, Endpoint>& workers() const { return workers_; }
  const MasterConfig& config() const { return cfg_; }
  bool finished() const { return round_ >= cfg_.max_round && done_last_; }

 private:
  void init_workers(int start_round);
  void start_allreduce();

  Endpoint self_;
  MasterConfig cfg_;
  Mailer* mailer_;
  bool verbose_;
  std::map<int, Endpoint> workers_;
  int next_id_ = 0;
  int round_ = -1;
  int num_complete_ = 0;
  bool done_last_ = false;
  std::map<int, int> early_;  // completion reports for rounds the master has not started yet
};

}  // namespace aar
