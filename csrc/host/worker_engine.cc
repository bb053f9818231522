#include "worker_engine.h"

#include <algorithm>
#include <cstdio>
#include <stdexcept>

namespace aar {

WorkerEngine::WorkerEngine(Endpoint self, DataSource source, DataSink sink, Mailer* mailer,
                           WorkerOptions opts)
    : self_(self),
      source_(std::move(source)),
      sink_(std::move(sink)),
      mailer_(mailer),
      opts_(opts) {}

void WorkerEngine::guarded(const char* where, const std::function<void()>& body) {
  if (opts_.rethrow) {
    body();
    return;
  }
  try {
    body();
  } catch (const std::exception& e) {
    errors_.push_back(std::string(where) + ": " + e.what());
    std::fprintf(stderr, "[aar worker %d] error in %s: %s\n", id_, where, e.what());
  }
}

void WorkerEngine::handle(const Message& m) {
  if (id_ == -1 && !std::holds_alternative<InitWorkers>(m)) {
    // The reference re-queues to itself until InitWorkers shows up
    // (AllreduceWorker.scala:95-97,120-122,132-134); parking the message and replaying
    // it after init gives the same order without spinning the mailbox.
    before_init_.push_back(m);
    return;
  }
  switch (kind_of(m)) {
    case MsgKind::Init:
      guarded("init worker", [&] { on_init(std::get<InitWorkers>(m)); });
      break;
    case MsgKind::Start:
      guarded("start all reduce", [&] { on_start(std::get<StartAllreduce>(m)); });
      break;
    case MsgKind::Scatter:
      guarded("scatter block", [&] { on_scatter(std::get<ScatterBlock>(m)); });
      break;
    case MsgKind::Reduce:
      guarded("reduce block", [&] { on_reduce(std::get<ReduceBlock>(m)); });
      break;
    case MsgKind::Complete:
      break;  // a worker never receives completion reports
  }
}

void WorkerEngine::on_init(const InitWorkers& m) {
  if (id_ != -1) {  // later InitWorkers only refresh the peer map (AllreduceWorker.scala:87-89)
    peers_ = m.workers;
    return;
  }
  Partition part(m.data_size, m.worker_num);
  if (opts_.strict_partition && !part.reference_defined())
    throw std::out_of_range("partition of " + std::to_string(m.data_size) + " elements over " +
                            std::to_string(m.worker_num) + " workers has fewer block starts than workers");
  id_ = m.dest_id;
  master_ = m.master;
  peer_num_ = m.worker_num;
  peers_ = m.workers;
  th_reduce_ = m.th_reduce;
  th_complete_ = m.th_complete;
  max_lag_ = m.max_lag;
  // A replacement that joins a running job starts at the master's current round instead of replaying
  // (and force-completing) the whole history; start_round is 0 in the reference's flow.
  round_ = m.start_round;
  max_round_ = m.start_round - 1;
  max_scattered_ = m.start_round - 1;
  completed_.clear();
  data_size_ = m.data_size;
  max_chunk_ = m.max_chunk_size;
  data_.assign(static_cast<size_t>(data_size_), 0.f);
  part_ = part;

  const int slots = max_lag_ + 1;
  scatter_ = std::make_unique<ScatterBuffer>(part_.size(id_), peer_num_, slots, th_reduce_, max_chunk_);
  reduce_ = std::make_unique<ReduceBuffer>(part_.max_block(), total_reduced_chunks(part_, max_chunk_),
                                           data_size_, peer_num_, slots, th_complete_, max_chunk_, 0);
  if (opts_.verbose) {
    std::printf("----Actor id = %d\n----Number of peers / total peers = %zu / %d\n", id_, peers_.size(),
                peer_num_);
    std::printf("----Thresholds: thReduce = %g, thComplete = %g, maxLag = %d\n", th_reduce_,
                th_complete_, max_lag_);
    std::printf("----Size of scatter buffer: %d x %d x %lld. threshold : %d\n", slots, peer_num_,
                static_cast<long long>(scatter_->width()), scatter_->required());
    std::printf("----Size of reduce buffer: %d x %d x %lld. threshold: %d\n", slots, peer_num_,
                static_cast<long long>(reduce_->width()), reduce_->required());
  }
  std::deque<Message> parked;
  parked.swap(before_init_);
  for (auto& held : parked) handle(held);
}

void WorkerEngine::check_round(int round) const {
  if (static_cast<int64_t>(round) - max_round_ > opts_.max_round_jump)
    throw std::out_of_range("round " + std::to_string(round) + " is more than " +
                            std::to_string(opts_.max_round_jump) + " rounds beyond round " +
                            std::to_string(max_round_) + ", the newest started here: dropped");
}

void WorkerEngine::on_start(const StartAllreduce& m) {
  check_round(m.round);
  max_round_ = std::max(max_round_, m.round);
  // Fallen more than maxLag behind: force every chunk of the oldest round out with
  // whatever has arrived, then complete it (AllreduceWorker.scala:100-106).  `round_`
  // is read live: an inline self-delivery may complete the round mid-loop.
  while (round_ < max_round_ - max_lag_) {
    const int chunks = scatter_->chunks();
    for (int k = 0; k < chunks; ++k) {
      auto [sum, cnt] = scatter_->reduce(0, k);
      broadcast(sum, k, round_, cnt);
    }
    complete(round_, 0);
  }
  while (max_scattered_ < max_round_) {
    fetch(max_scattered_ + 1);
    scatter();
    max_scattered_ += 1;
  }
  for (auto it = completed_.begin(); it != completed_.end();)
    it = (*it < round_) ? completed_.erase(it) : std::next(it);
}

void WorkerEngine::on_scatter(const ScatterBlock& m) {
  if (m.dest_id != id_)
    throw std::logic_error("scatter block for worker " + std::to_string(m.dest_id) +
                           " delivered to worker " + std::to_string(id_));
  if (m.round < round_ || completed_.count(m.round)) return;  // outdated
  check_round(m.round);
  if (m.round <= max_round_) {
    const int row = m.round - round_;
    scatter_->store(m.value.data(), m.value.size(), row, m.src_id, m.chunk_id);
    if (scatter_->at_threshold(row, m.chunk_id)) {
      auto [sum, cnt] = scatter_->reduce(row, m.chunk_id);
      broadcast(sum, m.chunk_id, m.round, cnt);
    }
  } else {  // first contact from a future round: start it ourselves, then retry
    mailer_->send(self_, StartAllreduce{m.round});
    mailer_->send(self_, m);
  }
}

void WorkerEngine::on_reduce(const ReduceBlock& m) {
  if (static_cast<int64_t>(m.value.size()) > max_chunk_)
    throw std::runtime_error("Reduced block of size " + std::to_string(m.value.size()) +
                             " is larger than expected.. Max msg size is " + std::to_string(max_chunk_));
  if (m.dest_id != id_)
    throw std::runtime_error("Message with destination " + std::to_string(m.dest_id) +
                             " was incorrectly routed to node " + std::to_string(id_));
  if (m.round < round_ || completed_.count(m.round)) return;  // outdated
  check_round(m.round);
  if (m.round <= max_round_) {
    const int row = m.round - round_;
    reduce_->store(m.value.data(), m.value.size(), row, m.src_id, m.chunk_id, m.count);
    if (reduce_->at_threshold(row)) complete(m.round, row);
  } else {
    mailer_->send(self_, StartAllreduce{m.round});
    mailer_->send(self_, m);
  }
}

void WorkerEngine::fetch(int iteration) {
  std::vector<float> in = source_(iteration);
  if (static_cast<int64_t>(in.size()) != data_size_)
    throw std::invalid_argument("Input data size " + std::to_string(in.size()) +
                                " is different from initialization time " + std::to_string(data_size_) + "!");
  data_ = std::move(in);
}

void WorkerEngine::post(int peer_id, Message msg, bool scatter_kind) {
  const Endpoint to = peers_.at(peer_id);
  if (to == self_) {  // straight function call, ahead of the remaining sends
    if (scatter_kind) on_scatter(std::get<ScatterBlock>(msg));
    else on_reduce(std::get<ReduceBlock>(msg));
  } else {
    mailer_->send(to, std::move(msg));
  }
}

void WorkerEngine::scatter() {
  const int known = static_cast<int>(peers_.size());  // loop length is the KNOWN peer count
  for (int i = 0; i < known; ++i) {
    const int dest = (i + id_) % peer_num_;  // rotation starts at self
    if (!peers_.count(dest)) continue;
    const int64_t base = part_.start(dest);
    const int64_t bsize = part_.size(dest);
    const int chunks = num_chunks(bsize, max_chunk_);
    for (int c = 0; c < chunks; ++c) {
      const int64_t lo = chunk_begin(c, max_chunk_, bsize), hi = chunk_end(c, max_chunk_, bsize);
      ScatterBlock blk;
      blk.value.assign(data_.begin() + base + lo, data_.begin() + base + hi);
      blk.src_id = id_;
      blk.dest_id = dest;
      blk.chunk_id = c;
      blk.round = max_scattered_ + 1;
      post(dest, std::move(blk), true);
    }
  }
}

void WorkerEngine::broadcast(const std::vector<float>& data, int chunk, int bcast_round, int count) {
  const int known = static_cast<int>(peers_.size());
  for (int i = 0; i < known; ++i) {
    const int dest = (i + id_) % peer_num_;
    if (!peers_.count(dest)) continue;
    ReduceBlock blk;
    blk.value = data;
    blk.src_id = id_;
    blk.dest_id = dest;
    blk.chunk_id = chunk;
    blk.round = bcast_round;
    blk.count = count;
    post(dest, std::move(blk), false);
  }
}

void WorkerEngine::complete(int done_round, int row) {
  auto [out, counts] = reduce_->assemble(row);
  sink_(out, counts, done_round);
  data_.clear();
  if (master_ != kNoEndpoint) mailer_->send(master_, CompleteAllreduce{id_, done_round});
  completed_.insert(done_round);
  if (round_ == done_round) {
    do {
      round_ += 1;
      scatter_->advance();
      reduce_->advance();
    } while (completed_.count(round_));
  }
}

}  // namespace aar
