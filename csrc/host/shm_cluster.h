// The DEVICE protocol restated for CPU threads in shared memory.
//
// csrc/cuda/fused_allreduce.cu turns the reference's five messages into stores and round-tagged flags in a
// symmetric region per rank.  This class runs the same protocol — same region layout (aar/device_layout.h),
// same partition/chunk/threshold arithmetic (aar/protocol_math.h), same rules: thresholds fire at ">= required"
// with the true count, slot = round % (maxLag+1), a sender skips a peer whose slot still holds a round more than
// maxLag behind, the thAllreduce gate reads the replicated `done` words, waits are bounded by a timeout or closed
// by a force tag — with one CPU thread per rank, host memory for the regions and acquire/release atomics for the
// flags.  Three uses:
//
//   * an oracle for the device semantics (partial sums, counts, zeros for missing chunks) that needs no GPU;
//   * a data-race check of the protocol AS IMPLEMENTED: data words are plain floats, so ThreadSanitizer sees any
//     read that is not ordered after its writer by a flag (scripts/sanitize_host.sh);
//   * a single-node CPU backend without sockets.
//
// What it leaves out of the kernel: the multi-CTA work split (one thread does a rank's scatter, reduce and
// completion in turn), bf16 slots, the one-shot and NVLS variants.
#pragma once
#include <array>
#include <atomic>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "aar/device_layout.h"

namespace aar {

struct ShmConfig {
  int world = 2;
  int64_t n = 16;
  int64_t max_chunk = 4;
  int max_lag = 1;
  float th_reduce = 1.f, th_complete = 1.f;
  float th_allreduce = 1.f;  // <= 0: no gate
  double timeout_ms = 2000.0;
  float scale = 1.f;       // source scale fused into the scatter (the kernel's `scale`)
  bool normalize = false;  // divide every reduced chunk by its contributor count (the kernel's `normalize`)
  // Extension over the kernel's protocol (candidate for it): a sender that leaves a peer out because of
  // back-pressure says so in a word of its own, so the peer does not sit out a timeout waiting for data that was
  // never sent.  See "skip notices" in shm_cluster.cc.
  bool skip_notice = true;
};

struct ShmInjection {
  uint32_t mute_mask = 0;                          // ranks that scatter nothing this round
  std::array<uint32_t, kMaxRanks> skip_dest{};     // per rank: destinations it leaves out
  std::array<uint64_t, kMaxRanks> delay_ns{};      // per rank: stall before the round starts
};

struct ShmRoundInfo {
  int reduced_full = 0;     // my chunks reduced with every rank present
  int reduced_partial = 0;  // my chunks reduced with fewer (threshold or timeout)
  int chunks_present = 0;   // (owner, chunk) results this rank assembled
  bool timed_out = false;   // some wait ended on the clock
  uint32_t dest_mask = 0;   // peers this rank pushed to
};

struct ShmFreeRunStats {
  int rounds = 0;
  double seconds = 0;
  int64_t mismatches = 0;       // elements whose sum does not name exactly `count` contributors
  int64_t partial_elements = 0; // elements completed with fewer than `world` contributors
  int64_t missing_elements = 0; // elements completed with count 0
  int timeouts = 0;
  int forced = 0;
};

class ShmCluster {
 public:
  explicit ShmCluster(const ShmConfig& cfg);  // private memory: every rank is a thread of this process
  // Named POSIX shared memory: one process per rank, like the reference's one JVM per worker.  `create` makes (and on
  // destruction unlinks) the segment; the others attach, wait for it to be initialised and check that it was made
  // for the same configuration.
  // A segment whose creator process is gone is a leftover of an earlier run with the same name: an attacher keeps
  // waiting for the new one unless `accept_orphan` (a replacement joining a job whose creator has died).
  ShmCluster(const ShmConfig& cfg, const std::string& shm_name, bool create, double attach_timeout_s = 30.0,
             bool accept_orphan = false);
  ~ShmCluster();
  ShmCluster(const ShmCluster&) = delete;
  ShmCluster& operator=(const ShmCluster&) = delete;

  // One rank's round (tag = round + 1): returns when that rank has completed it.  One caller thread per rank;
  // a rank issues its rounds in increasing order.
  ShmRoundInfo rank_round(int rank, int round, const float* src, float* out, int32_t* counts,
                          const ShmInjection* inj = nullptr);
  // The same round on every rank, one thread each.
  std::vector<ShmRoundInfo> run_round(int round, const std::vector<const float*>& srcs, const std::vector<float*>& outs,
                                      const std::vector<int32_t*>& counts, const ShmInjection* inj = nullptr);
  // Every rank runs `rounds` rounds on its own thread, as fast as the protocol lets it; with probability
  // `stall_prob` a rank sleeps `stall_us` before a round.  Rank q contributes 2^q to every element, so a sum names
  // its contributors and is checked against the count.  A rank more than maxLag rounds behind the front-runner
  // closes its waits at once (the coordinator's catch-up, AllreduceWorker.scala:100-106).
  // Rounds continue after the last one issued on this cluster.
  ShmFreeRunStats run_free(int rounds, double stall_prob, int stall_us, uint64_t seed);
  int next_round() const;  // lowest round number no rank has issued yet
  int issued(int rank) const;  // lowest round number `rank` may still issue

  void force_complete(int rank, int round);  // waits of rounds <= `round` on `rank` close immediately
  // Peers `rank` has been told are gone (bit q = rank q; the kernel's dead-peer mask, DeviceAllreduce.set_dead_peers):
  // nothing is pushed to them, nothing waits for them, and the gate asks for no more than the live ranks can give.
  void set_dead_peers(int rank, uint32_t mask);
  const SymLayout& layout() const { return sym_; }
  int req_reduce() const { return req_reduce_; }
  int req_complete() const { return req_complete_; }
  int gate_count() const { return gate_count_; }
  int total_chunks() const { return total_chunks_; }

 private:
  void init_geometry();
  void place(uint8_t* block);  // carve control words and regions out of one zero-filled block
  static size_t control_bytes();
  size_t region_bytes() const;
  uint8_t* region(int rank) const { return regions_[static_cast<size_t>(rank)]; }

  ShmConfig cfg_;
  SymLayout sym_;
  Partition part_;
  int L_ = 1, req_reduce_ = 0, req_complete_ = 0, gate_count_ = 0, total_chunks_ = 0;
  uint8_t* block_ = nullptr;     // control words + world regions, malloc'ed or mmap'ed
  size_t block_bytes_ = 0;
  std::string shm_name_;         // empty: private memory
  bool shm_owner_ = false;
  std::vector<uint8_t*> regions_;
  // control words (inside block_, so that processes attached to the same segment share them)
  std::atomic<uint32_t>* force_ = nullptr;     // [rank]
  std::atomic<int>* progress_ = nullptr;       // [rank]: round it is working on (free-run catch-up)
  std::atomic<int>* next_round_ = nullptr;     // [rank]: lowest round it may still issue (tags never repeat)
  std::atomic<uint32_t>* notice_ = nullptr;    // [dest][src]: highest tag for which src left dest out
  std::atomic<uint32_t>* dead_ = nullptr;      // [rank]: peers that rank has been told are gone
};

}  // namespace aar
