// Deterministic in-process message bus.
//
// Plays the part Akka's mailboxes play in the reference's single-actor tests
// (src/test/scala/AllreduceSpec.scala:812-818): every endpoint is either a real engine
// or a probe that records what it receives.  One global FIFO keeps delivery order
// reproducible; a message an engine sends to itself goes to the back of the queue, the
// way `self ! msg` does.
#pragma once
#include <deque>
#include <functional>
#include <map>
#include <utility>
#include <vector>

#include "messages.h"

namespace aar {

class LocalBus : public Mailer {
 public:
  using Handler = std::function<void(const Message&)>;

  Endpoint add_endpoint(Handler h);  // engine
  Endpoint add_probe();              // records
  void rebind(Endpoint ep, Handler h) { handlers_[ep] = std::move(h); }

  void send(Endpoint to, Message msg) override;
  // Delivers until the queue drains (or `max_steps` deliveries); returns deliveries made.
  size_t run(size_t max_steps = 1u << 22);
  size_t pending() const { return queue_.size(); }

  // Fault injection for tests (the reference has none): every ScatterBlock / ReduceBlock in
  // transit is dropped with probability `drop` or pushed back to the end of the queue with probability
  // `delay` (reordering across senders; per-pair FIFO is NOT preserved for delayed messages).  Control
  // messages (Init/Start/Complete) always get through.  Deterministic for a given seed.
  void set_chaos(uint64_t seed, double drop, double delay) {
    rng_state_ = seed ? seed : 0x9e3779b97f4a7c15ull;
    drop_ = drop;
    delay_ = delay;
  }
  size_t chaos_dropped() const { return chaos_dropped_; }

  std::vector<Message>& inbox(Endpoint probe) { return inboxes_[probe]; }
  size_t dropped() const { return dropped_; }

 private:
  Endpoint next_ = 1;
  std::map<Endpoint, Handler> handlers_;
  std::map<Endpoint, std::vector<Message>> inboxes_;
  std::deque<std::pair<Endpoint, Message>> queue_;
  size_t dropped_ = 0;
  double next_uniform();
  uint64_t rng_state_ = 0;
  double drop_ = 0.0, delay_ = 0.0;
  size_t chaos_dropped_ = 0;
};

}  // namespace aar
