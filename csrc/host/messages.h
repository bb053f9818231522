// Wire protocol of the threshold/lag allreduce (host side).
//
// Same five messages as the reference's AllreduceMessage.scala:7-21, with an opaque
// 64-bit `Endpoint` standing in for Akka's ActorRef.  On the GPU path these messages
// collapse into round-tagged flags in symmetric memory (csrc/cuda/fused_allreduce.cu);
// the structs here are what the CPU backend, the oracle and the tests exchange.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <variant>
#include <vector>

namespace aar {

using Endpoint = int64_t;
constexpr Endpoint kNoEndpoint = -1;

struct InitWorkers {
  std::map<int, Endpoint> workers;  // id -> address of every known peer (may be partial)
  int worker_num = 0;               // configured P, not workers.size()
  Endpoint master = kNoEndpoint;
  int dest_id = -1;
  float th_reduce = 1.f;
  float th_complete = 1.f;
  int max_lag = 0;
  int64_t data_size = 0;
  int64_t max_chunk_size = 1024;
  // Extension (not in the reference): first round a worker that joins a RUNNING job takes part in.
  // 0 for the initial broadcast, so the normal start-up is message-for-message the reference's.
  int start_round = 0;
  // Extension: the last round of the job (maxRound), -1 = not told.  The reference's workers never learn it — they
  // only react to StartAllreduce — but a worker that gates rounds itself (device / shared-memory gate) has to know
  // where to stop.
  int max_round = -1;
};

struct StartAllreduce {
  int round = 0;
};

struct ScatterBlock {
  std::vector<float> value;
  int src_id = 0, dest_id = 0, chunk_id = 0, round = 0;
};

struct ReduceBlock {
  std::vector<float> value;
  int src_id = 0, dest_id = 0, chunk_id = 0, round = 0;
  int count = 0;  // how many peers contributed to `value`
};

struct CompleteAllreduce {
  int src_id = 0, round = 0;
};

using Message = std::variant<InitWorkers, StartAllreduce, ScatterBlock, ReduceBlock, CompleteAllreduce>;

enum class MsgKind : uint8_t { Init = 0, Start = 1, Scatter = 2, Reduce = 3, Complete = 4 };
inline MsgKind kind_of(const Message& m) { return static_cast<MsgKind>(m.index()); }
std::string describe(const Message& m);

// Anything that can carry a message to an endpoint: the in-process bus, the TCP
// transport, or a test probe.
class Mailer {
 public:
  virtual ~Mailer() = default;
  virtual void send(Endpoint to, Message msg) = 0;
};

}  // namespace aar
