#include "local_bus.h"

#include <sstream>

namespace aar {

Endpoint LocalBus::add_endpoint(Handler h) {
  const Endpoint ep = next_++;
  handlers_[ep] = std::move(h);
  return ep;
}

Endpoint LocalBus::add_probe() {
  const Endpoint ep = next_++;
  inboxes_[ep];
  return ep;
}

void LocalBus::send(Endpoint to, Message msg) { queue_.emplace_back(to, std::move(msg)); }

double LocalBus::next_uniform() {  // xorshift64*: small, seedable, good enough for fault injection
  rng_state_ ^= rng_state_ >> 12;
  rng_state_ ^= rng_state_ << 25;
  rng_state_ ^= rng_state_ >> 27;
  return static_cast<double>((rng_state_ * 0x2545F4914F6CDD1Dull) >> 11) / 9007199254740992.0;
}

size_t LocalBus::run(size_t max_steps) {
  size_t steps = 0;
  while (!queue_.empty() && steps < max_steps) {
    auto [to, msg] = std::move(queue_.front());
    queue_.pop_front();
    ++steps;
    if (rng_state_ != 0 && (kind_of(msg) == MsgKind::Scatter || kind_of(msg) == MsgKind::Reduce)) {
      const double u = next_uniform();
      if (u < drop_) {
        ++chaos_dropped_;
        continue;
      }
      if (u < drop_ + delay_ && !queue_.empty()) {
        queue_.emplace_back(to, std::move(msg));
        continue;
      }
    }
    if (auto box = inboxes_.find(to); box != inboxes_.end()) {
      box->second.push_back(std::move(msg));
    } else if (auto h = handlers_.find(to); h != handlers_.end()) {
      h->second(msg);
    } else {
      ++dropped_;  // dead letter
    }
  }
  return steps;
}

static void join(std::ostringstream& os, const std::vector<float>& v) {
  os << "[";
  for (size_t i = 0; i < v.size(); ++i) os << (i ? ", " : "") << v[i];
  os << "]";
}

std::string describe(const Message& m) {
  std::ostringstream os;
  switch (kind_of(m)) {
    case MsgKind::Init: {
      const auto& x = std::get<InitWorkers>(m);
      os << "InitWorkers(peers=" << x.workers.size() << "/" << x.worker_num << ", destId=" << x.dest_id
         << ", thReduce=" << x.th_reduce << ", thComplete=" << x.th_complete << ", maxLag=" << x.max_lag
         << ", dataSize=" << x.data_size << ", maxChunkSize=" << x.max_chunk_size << ")";
      break;
    }
    case MsgKind::Start:
      os << "StartAllreduce(" << std::get<StartAllreduce>(m).round << ")";
      break;
    case MsgKind::Scatter: {
      const auto& x = std::get<ScatterBlock>(m);
      os << "ScatterBlock(";
      join(os, x.value);
      os << ", src=" << x.src_id << ", dest=" << x.dest_id << ", chunk=" << x.chunk_id
         << ", round=" << x.round << ")";
      break;
    }
    case MsgKind::Reduce: {
      const auto& x = std::get<ReduceBlock>(m);
      os << "ReduceBlock(";
      join(os, x.value);
      os << ", src=" << x.src_id << ", dest=" << x.dest_id << ", chunk=" << x.chunk_id
         << ", round=" << x.round << ", count=" << x.count << ")";
      break;
    }
    case MsgKind::Complete: {
      const auto& x = std::get<CompleteAllreduce>(m);
      os << "CompleteAllreduce(src=" << x.src_id << ", round=" << x.round << ")";
      break;
    }
  }
  return os.str();
}

}  // namespace aar
