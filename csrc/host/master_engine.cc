#include "master_engine.h"

#include <cstdio>

#include "aar/protocol_math.h"

namespace aar {

MasterEngine::MasterEngine(Endpoint self, MasterConfig cfg, Mailer* mailer, bool verbose)
    : self_(self), cfg_(cfg), mailer_(mailer), verbose_(verbose) {}

int MasterEngine::member_up(Endpoint worker) {
  // Ids follow arrival order.  The reference uses workers.size, which collides with a live id after a
  // removal; here a newcomer takes the lowest id that is free (so a replacement inherits the block of the
  // worker it replaces), which is the same thing until the first removal.
  int id = 0;
  while (workers_.count(id)) ++id;
  if (id >= cfg_.total_workers && round_ != -1) id = next_id_ > id ? next_id_ : id;  // surplus worker: unused id
  next_id_ = id + 1 > next_id_ ? id + 1 : next_id_;
  workers_[id] = worker;
  if (round_ == -1) {
    if (static_cast<int>(workers_.size()) >= cfg_.total_workers) {
      if (verbose_)
        std::printf("----%zu (out of %d) workers are up\n", workers_.size(), cfg_.total_workers);
      init_workers(0);
      round_ = 0;
      start_allreduce();
    }
  } else if (id < cfg_.total_workers) {
    // The job is already running (the reference never initialises such a node): everybody
    // gets the refreshed peer map — existing workers only update their peers (AllreduceWorker.scala:87-89)
    // — and the newcomer is initialised at the current round and told to start it.
    if (verbose_) std::printf("----worker joined a running job as id %d at round %d\n", id, round_);
    init_workers(round_);
    mailer_->send(worker, StartAllreduce{round_});
  }
  return id;
}

void MasterEngine::terminated(Endpoint worker) {
  bool removed = false;
  for (auto it = workers_.begin(); it != workers_.end();) {
    if (it->second == worker) {
      it = workers_.erase(it);
      removed = true;
    } else {
      ++it;
    }
  }
  if (removed && cfg_.announce_removals && round_ != -1 && !finished()) init_workers(round_);
}

void MasterEngine::handle(const Message& m) {
  const auto* c = std::get_if<CompleteAllreduce>(&m);
  if (!c) return;
  if (c->round > round_) {
    // Cannot happen with host-gated workers (they learn of a round from the master).  Device-gated
    // workers free-run and report asynchronously, so a report may arrive before the master's own
    // bookkeeping has reached that round: keep it instead of dropping it.
    early_[c->round] += 1;
    return;
  }
  if (c->round != round_) return;  // only the current round counts; late reports are ignored
  num_complete_ += 1;
  while (master_gate(num_complete_, cfg_.total_workers, cfg_.th_allreduce)) {
    if (round_ >= cfg_.max_round) {
      done_last_ = true;  // rounds 0..max_round inclusive have been run
      break;
    }
    if (verbose_)
      std::printf("----%d (out of %d) workers complete round %d\n", num_complete_, cfg_.total_workers, round_);
    round_ += 1;
    start_allreduce();
    auto it = early_.find(round_);
    if (it == early_.end()) break;
    num_complete_ = it->second;
    early_.erase(it);
  }
}

void MasterEngine::init_workers(int start_round) {
  for (const auto& [id, ep] : workers_) {
    InitWorkers init;
    init.workers = workers_;
    init.worker_num = cfg_.total_workers;
    init.master = self_;
    init.dest_id = id;
    init.th_reduce = cfg_.th_reduce;
    init.th_complete = cfg_.th_complete;
    init.max_lag = cfg_.max_lag;
    init.data_size = cfg_.data_size;
    init.max_chunk_size = cfg_.max_chunk_size;
    init.start_round = start_round;
    init.max_round = cfg_.max_round;
    mailer_->send(ep, std::move(init));
  }
}

void MasterEngine::start_allreduce() {
  if (verbose_) std::printf("\n----Start allreduce round %d\n", round_);
  num_complete_ = 0;
  for (const auto& [id, ep] : workers_) mailer_->send(ep, StartAllreduce{round_});
}

}  // namespace aar
