#include "shm_cluster.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <random>
#include <stdexcept>
#include <string>
#include <thread>

#include <fcntl.h>
#include <signal.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <cerrno>

namespace aar {

namespace {

using Clock = std::chrono::steady_clock;

// Flags live inside the raw region, like on the device; GCC's atomic builtins work on plain words and are what
// ThreadSanitizer models.
inline uint32_t ld_acquire(const uint32_t* p) { return __atomic_load_n(p, __ATOMIC_ACQUIRE); }
inline uint64_t ld_acquire(const uint64_t* p) { return __atomic_load_n(p, __ATOMIC_ACQUIRE); }
inline void st_release(uint32_t* p, uint32_t v) { __atomic_store_n(p, v, __ATOMIC_RELEASE); }
inline void st_release(uint64_t* p, uint64_t v) { __atomic_store_n(p, v, __ATOMIC_RELEASE); }

// Waiting: spin briefly (the flag usually flips within microseconds), then yield, then sleep — a rank that waits
// for a straggler must not burn the cores (or the container's CPU quota) the straggler needs to get there.
inline void relax(int& spins) {
  ++spins;
  if (spins < 128) {
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  } else if (spins < 192) {
    std::this_thread::yield();
  } else {
    std::this_thread::sleep_for(std::chrono::microseconds(spins < 1024 ? 20 : 200));
  }
}

struct ShmHeader {  // first words of a shared segment
  uint32_t magic, ready;
  int32_t world, max_lag;
  int64_t n, max_chunk;
  uint64_t total_bytes;
  int64_t owner_pid;
};
constexpr uint32_t kShmMagic = 0x41415253u;  // "AARS"
constexpr size_t kHeaderBytes = 256;
}  // namespace

size_t ShmCluster::control_bytes() {
  // header + force[16] + progress[16] + next_round[16] + notice[16*16] + dead[16], rounded to a page
  return static_cast<size_t>(SymLayout::up(kHeaderBytes + 4 * kMaxRanks * 4 + kMaxRanks * kMaxRanks * 4, 4096));
}
size_t ShmCluster::region_bytes() const { return static_cast<size_t>(SymLayout::up(sym_.total_bytes, 4096)); }

void ShmCluster::init_geometry() {
  const ShmConfig& cfg = cfg_;
  if (cfg.world < 1 || cfg.world > kMaxRanks) throw std::invalid_argument("ShmCluster: world must be 1..16");
  if (cfg.n <= 0 || cfg.max_chunk <= 0 || cfg.max_lag < 0) throw std::invalid_argument("ShmCluster: bad size arguments");
  L_ = cfg.max_lag + 1;
  sym_ = SymLayout::make(cfg.n, cfg.max_chunk, cfg.world, L_, 4, 0, 0);
  part_ = Partition(cfg.n, cfg.world, sym_.block);
  total_chunks_ = total_reduced_chunks(part_, cfg.max_chunk);
  req_reduce_ = threshold_count(cfg.th_reduce, cfg.world);
  req_complete_ = threshold_count(cfg.th_complete, total_chunks_);
  gate_count_ = 0;
  if (cfg.th_allreduce > 0.f)  // smallest k with master_gate(k) true (AllreduceMaster.scala:58)
    while (gate_count_ < cfg.world && !master_gate(gate_count_, cfg.world, cfg.th_allreduce)) ++gate_count_;
  block_bytes_ = control_bytes() + static_cast<size_t>(cfg.world) * region_bytes();
}

void ShmCluster::place(uint8_t* block) {
  // Zero-filled memory is a valid initial state for every word below (lock-free 32-bit atomics, value 0).
  block_ = block;
  uint8_t* p = block + kHeaderBytes;
  force_ = reinterpret_cast<std::atomic<uint32_t>*>(p);
  p += kMaxRanks * 4;
  progress_ = reinterpret_cast<std::atomic<int>*>(p);
  p += kMaxRanks * 4;
  next_round_ = reinterpret_cast<std::atomic<int>*>(p);
  p += kMaxRanks * 4;
  notice_ = reinterpret_cast<std::atomic<uint32_t>*>(p);
  p += kMaxRanks * kMaxRanks * 4;
  dead_ = reinterpret_cast<std::atomic<uint32_t>*>(p);
  regions_.clear();
  for (int r = 0; r < cfg_.world; ++r) regions_.push_back(block + control_bytes() + static_cast<size_t>(r) * region_bytes());
}

ShmCluster::ShmCluster(const ShmConfig& cfg) : cfg_(cfg) {
  static_assert(sizeof(std::atomic<uint32_t>) == 4 && sizeof(std::atomic<int>) == 4, "control words are 32-bit");
  init_geometry();
  uint8_t* p = static_cast<uint8_t*>(std::aligned_alloc(4096, block_bytes_));
  if (p == nullptr) throw std::bad_alloc();
  std::memset(p, 0, block_bytes_);
  place(p);
}

ShmCluster::ShmCluster(const ShmConfig& cfg, const std::string& shm_name, bool create, double attach_timeout_s,
                       bool accept_orphan)
    : cfg_(cfg), shm_name_(shm_name), shm_owner_(create) {
  init_geometry();
  if (shm_name.empty() || shm_name[0] != '/') throw std::invalid_argument("ShmCluster: shared memory name must start with '/'");
  int fd = -1;
  if (create) {
    shm_unlink(shm_name.c_str());  // a stale segment of a crashed run
    fd = shm_open(shm_name.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) throw std::runtime_error("shm_open(" + shm_name + ", create): " + std::strerror(errno));
    // posix_fallocate reserves the pages now: a /dev/shm that is too small (64 MB by default in containers) is an
    // error message here instead of a SIGBUS at the first store into the region
    int rc = ftruncate(fd, static_cast<off_t>(block_bytes_)) != 0 ? errno : 0;
    if (rc == 0) rc = posix_fallocate(fd, 0, static_cast<off_t>(block_bytes_));
    if (rc != 0) {
      const std::string why = std::string(std::strerror(rc)) + " (" + std::to_string(block_bytes_ >> 20) +
                              " MiB of shared memory needed: world x (maxLag+1) x 2 x message bytes)";
      close(fd);
      shm_unlink(shm_name.c_str());
      throw std::runtime_error("ftruncate(" + shm_name + "): " + why);
    }
  } else {
    const auto give_up = Clock::now() + std::chrono::duration_cast<Clock::duration>(std::chrono::duration<double>(attach_timeout_s));
    for (;;) {  // the creator may not have got there yet, or the name may still point at a dead run's segment
      fd = shm_open(shm_name.c_str(), O_RDWR, 0600);
      struct stat st {};
      if (fd >= 0 && fstat(fd, &st) == 0 && static_cast<size_t>(st.st_size) >= kHeaderBytes) {
        ShmHeader h{};
        const bool got = pread(fd, &h, sizeof(h), 0) == static_cast<ssize_t>(sizeof(h));
        const bool live = got && h.magic == kShmMagic && h.ready == 1u && h.owner_pid > 0 &&
                          (kill(static_cast<pid_t>(h.owner_pid), 0) == 0 || errno == EPERM);
        if (got && h.magic == kShmMagic && h.ready == 1u && (live || accept_orphan) &&
            static_cast<size_t>(st.st_size) >= block_bytes_)
          break;
      }
      if (fd >= 0) close(fd);
      fd = -1;
      if (Clock::now() > give_up) throw std::runtime_error("shared memory segment " + shm_name + " did not appear");
      std::this_thread::sleep_for(std::chrono::milliseconds(5));
    }
  }
  void* m = mmap(nullptr, block_bytes_, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (m == MAP_FAILED) {
    if (create) shm_unlink(shm_name.c_str());
    throw std::runtime_error("mmap(" + shm_name + "): " + std::strerror(errno));
  }
  place(static_cast<uint8_t*>(m));  // (a fresh segment is zero-filled by the kernel)
  auto* hdr = reinterpret_cast<ShmHeader*>(block_);
  uint32_t* ready = &hdr->ready;
  if (create) {
    hdr->magic = kShmMagic;
    hdr->world = cfg.world;
    hdr->max_lag = cfg.max_lag;
    hdr->n = cfg.n;
    hdr->max_chunk = cfg.max_chunk;
    hdr->total_bytes = block_bytes_;
    hdr->owner_pid = static_cast<int64_t>(getpid());
    st_release(ready, 1u);
  } else {
    const auto give_up = Clock::now() + std::chrono::duration_cast<Clock::duration>(std::chrono::duration<double>(attach_timeout_s));
    while (ld_acquire(ready) != 1u) {
      if (Clock::now() > give_up) {
        munmap(block_, block_bytes_);
        throw std::runtime_error("shared memory segment " + shm_name + " was never initialised");
      }
      std::this_thread::sleep_for(std::chrono::milliseconds(1));
    }
    if (hdr->magic != kShmMagic || hdr->world != cfg.world || hdr->max_lag != cfg.max_lag || hdr->n != cfg.n ||
        hdr->max_chunk != cfg.max_chunk || hdr->total_bytes != block_bytes_) {
      munmap(block_, block_bytes_);
      throw std::runtime_error("shared memory segment " + shm_name + " was created for a different configuration");
    }
  }
}

ShmCluster::~ShmCluster() {
  if (block_ == nullptr) return;
  if (shm_name_.empty()) {
    std::free(block_);
  } else {
    munmap(block_, block_bytes_);
    if (shm_owner_) shm_unlink(shm_name_.c_str());
  }
}

int ShmCluster::issued(int rank) const {
  if (rank < 0 || rank >= cfg_.world) throw std::invalid_argument("ShmCluster: rank out of range");
  return next_round_[static_cast<size_t>(rank)].load();
}

int ShmCluster::next_round() const {
  int m = 0;
  for (int r = 0; r < cfg_.world; ++r) m = std::max(m, next_round_[static_cast<size_t>(r)].load());
  return m;
}

void ShmCluster::set_dead_peers(int rank, uint32_t mask) {
  if (rank < 0 || rank >= cfg_.world) throw std::invalid_argument("ShmCluster::set_dead_peers: rank out of range");
  dead_[static_cast<size_t>(rank)].store(mask & ~(1u << rank), std::memory_order_release);
}

void ShmCluster::force_complete(int rank, int round) {
  if (rank < 0 || rank >= cfg_.world) throw std::invalid_argument("ShmCluster: rank out of range");
  if (round < 0) return;
  auto& f = force_[static_cast<size_t>(rank)];
  uint32_t want = static_cast<uint32_t>(round) + 1u, cur = f.load();
  while (cur < want && !f.compare_exchange_weak(cur, want)) {
  }
}

ShmRoundInfo ShmCluster::rank_round(int rank, int round, const float* src, float* out, int32_t* counts,
                                    const ShmInjection* inj) {
  const int P = cfg_.world;
  if (rank < 0 || rank >= P) throw std::invalid_argument("ShmCluster: rank out of range");
  // a round tag that was used before would find its own stale flags: rounds only go up
  if (round < next_round_[static_cast<size_t>(rank)].load())
    throw std::invalid_argument("ShmCluster: rank " + std::to_string(rank) + " already issued round " + std::to_string(round));
  next_round_[static_cast<size_t>(rank)].store(round + 1);
  const uint32_t tag = static_cast<uint32_t>(round) + 1u;
  const int slot = ring_slot(round, L_);
  uint8_t* mine = region(rank);
  ShmRoundInfo info;
  const bool muted = inj != nullptr && ((inj->mute_mask >> rank) & 1u);
  const uint32_t skip = inj != nullptr ? inj->skip_dest[static_cast<size_t>(rank)] : 0u;
  if (inj != nullptr && inj->delay_ns[static_cast<size_t>(rank)] != 0)
    std::this_thread::sleep_for(std::chrono::nanoseconds(inj->delay_ns[static_cast<size_t>(rank)]));

  const auto t0 = Clock::now();
  const auto timeout = std::chrono::duration_cast<Clock::duration>(std::chrono::duration<double, std::milli>(cfg_.timeout_ms));
  auto forced = [&] { return force_[static_cast<size_t>(rank)].load(std::memory_order_relaxed) >= tag; };
  auto expired = [&](int periods) {
    if (forced()) return true;
    if (Clock::now() - t0 > periods * timeout) {
      info.timed_out = true;
      return true;
    }
    return false;
  };
  const uint32_t* done = reinterpret_cast<const uint32_t*>(mine + sym_.off_done);

  // ---- thAllreduce gate (AllreduceMaster.scala:56-62, evaluated here from the replicated `done` words)
  if (gate_count_ > 0 && tag > 1) {
    int spins = 0;
    for (;;) {
      int have = 0;
      for (int q = 0; q < P; ++q) have += (q == rank || ld_acquire(done + q) + 1u >= tag) ? 1 : 0;
      const int alive = P - __builtin_popcount(dead_[static_cast<size_t>(rank)].load(std::memory_order_acquire) & ((1u << P) - 1u));
      if (have >= (gate_count_ < alive ? gate_count_ : alive) || expired(1)) break;
      relax(spins);
    }
  }
  const double dbg_gate = std::chrono::duration<double, std::milli>(Clock::now() - t0).count();
  // ---- back-pressure: leave out a peer whose ring slot still holds a round more than maxLag behind
  uint32_t dest_ok = 0;
  const uint32_t dead_now = dead_[static_cast<size_t>(rank)].load(std::memory_order_acquire);
  for (int q = 0; q < P; ++q) {
    const uint32_t d = q == rank ? tag : ld_acquire(done + q);
    if (d + static_cast<uint32_t>(L_) >= tag && !((dead_now >> q) & 1u)) dest_ok |= 1u << q;
  }
  info.dest_mask = dest_ok;
  // ---- skip notices.  A peer left out here will get neither my slices nor my reduced block for this round; without
  // a word from me it can only find that out by waiting for its timeout (or for the coordinator's catch-up).  The
  // notice is a per-(destination, source) word OUTSIDE the slot ring — the slot is exactly what must not be touched
  // — holding the highest tag for which the source left the destination out.  Release order makes it safe to read:
  // whoever sees notice >= t also sees every flag this rank raised for rounds before the one that wrote it.
  if (cfg_.skip_notice)
    for (int q = 0; q < P; ++q)
      if (!((dest_ok >> q) & 1u) && !((dead_now >> q) & 1u)) notice_[static_cast<size_t>(q) * P + rank].store(tag, std::memory_order_release);
  auto left_out_by = [&](int s) {  // will nothing come from rank s: it is known dead, or it told me that it skipped me
    if ((dead_[static_cast<size_t>(rank)].load(std::memory_order_acquire) >> s) & 1u) return true;
    return cfg_.skip_notice && notice_[static_cast<size_t>(rank) * P + s].load(std::memory_order_acquire) >= tag;
  };

  // ---- phase S: my slice of every owner's block into that owner's scat row [me], then the arrival flag
  if (!muted) {
    for (int i = 0; i < P; ++i) {
      const int d = (i + rank) % P;  // rotation starts at self (AllreduceWorker.scala:213-216)
      if (((skip >> d) & 1u) || !((dest_ok >> d) & 1u)) continue;
      const int64_t bsize = part_.size(d);
      const int nch = num_chunks(bsize, cfg_.max_chunk);
      float* row = reinterpret_cast<float*>(region(d) + sym_.off_scat) + sym_.scat_elem(slot, rank);
      uint32_t* arr = reinterpret_cast<uint32_t*>(region(d) + sym_.off_arr);
      for (int ch = 0; ch < nch; ++ch) {
        const int64_t cb = chunk_begin(ch, cfg_.max_chunk, bsize), ce = chunk_end(ch, cfg_.max_chunk, bsize);
        if (cfg_.scale == 1.f) {
          std::memcpy(row + cb, src + part_.start(d) + cb, static_cast<size_t>(ce - cb) * sizeof(float));
        } else {
          const float* from = src + part_.start(d);
          for (int64_t k = cb; k < ce; ++k) row[k] = from[k] * cfg_.scale;
        }
        st_release(arr + sym_.arr_index(slot, rank, ch), tag);
      }
    }
  }

  // ---- phase R: reduce my block chunk by chunk as thresholds are met, broadcast into every peer's sink row [me]
  {
    const int64_t bsize = part_.size(rank);
    const int nch = num_chunks(bsize, cfg_.max_chunk);
    const uint32_t* arr = reinterpret_cast<const uint32_t*>(mine + sym_.off_arr);
    const float* rows = reinterpret_cast<const float*>(mine + sym_.off_scat);
    std::vector<char> reduced(static_cast<size_t>(nch), 0);
    std::vector<float> acc(static_cast<size_t>(cfg_.max_chunk));
    int left = nch, spins = 0;
    while (left > 0) {
      bool progressed = false;
      const bool give_up = expired(1);
      for (int ch = 0; ch < nch; ++ch) {
        if (reduced[static_cast<size_t>(ch)]) continue;
        uint32_t have = 0;
        for (int s = 0; s < P; ++s)
          if (ld_acquire(arr + sym_.arr_index(slot, s, ch)) == tag) have |= 1u << s;
        int cnt = __builtin_popcount(have);
        if (cnt < req_reduce_ && !give_up) {
          // everyone who has not arrived has said it never will: decide now with what is here
          int absent = 0;
          for (int s = 0; s < P; ++s) absent += (!((have >> s) & 1u) && left_out_by(s)) ? 1 : 0;
          if (absent == 0 || cnt + absent < P) continue;
          // (flags are re-read: a slice that landed between the two looks is simply included)
          for (int s = 0; s < P; ++s)
            if (ld_acquire(arr + sym_.arr_index(slot, s, ch)) == tag) have |= 1u << s;
          cnt = __builtin_popcount(have);
        }
        const int64_t cb = chunk_begin(ch, cfg_.max_chunk, bsize), ce = chunk_end(ch, cfg_.max_chunk, bsize);
        const int64_t len = ce - cb;
        std::fill(acc.begin(), acc.begin() + len, 0.f);
        for (int s = 0; s < P; ++s) {  // peer order, like the reference's reduce (ScatteredDataBuffer.scala:20-30)
          if (!((have >> s) & 1u)) continue;
          const float* __restrict r = rows + sym_.scat_elem(slot, s) + cb;
          float* __restrict a = acc.data();
          for (int64_t k = 0; k < len; ++k) a[k] += r[k];
        }
        if (cfg_.normalize && cnt > 0) {
          const float post = 1.f / static_cast<float>(cnt);
          for (int64_t k = 0; k < len; ++k) acc[static_cast<size_t>(k)] *= post;
        }
        for (int i = 0; i < P; ++i) {
          const int d = (i + rank) % P;
          if (!((dest_ok >> d) & 1u)) continue;
          float* to = reinterpret_cast<float*>(region(d) + sym_.off_sink) + sym_.sink_elem(slot, rank) + cb;
          std::memcpy(to, acc.data(), static_cast<size_t>(len) * sizeof(float));
          st_release(reinterpret_cast<uint64_t*>(region(d) + sym_.off_rflag) + sym_.arr_index(slot, rank, ch),
                     (static_cast<uint64_t>(tag) << 32) | static_cast<uint32_t>(cnt));
        }
        reduced[static_cast<size_t>(ch)] = 1;
        --left;
        progressed = true;
        (cnt == P ? info.reduced_full : info.reduced_partial) += 1;
      }
      if (!progressed) relax(spins);
      else spins = 0;
    }
  }

  const double dbg_reduce = std::chrono::duration<double, std::milli>(Clock::now() - t0).count();
  // ---- phase C: wait for enough reduced chunks, then assemble output and counts from what is there
  const uint64_t* rflag = reinterpret_cast<const uint64_t*>(mine + sym_.off_rflag);
  {
    int spins = 0;
    for (;;) {
      int seen = 0, never = 0;
      for (int o = 0; o < P; ++o) {
        const int nch = num_chunks(part_.size(o), cfg_.max_chunk);
        const bool gone = left_out_by(o);  // read BEFORE the flags: a notice seen here postdates the owner's pushes
        for (int ch = 0; ch < nch; ++ch) {
          const bool here = static_cast<uint32_t>(ld_acquire(rflag + sym_.arr_index(slot, o, ch)) >> 32) == tag;
          seen += here ? 1 : 0;
          never += (!here && gone) ? 1 : 0;
        }
      }
      if (seen >= req_complete_ || seen >= total_chunks_ || expired(2)) break;
      if (never > 0 && seen + never >= total_chunks_) break;  // the rest was never sent: nothing left to wait for
      relax(spins);
    }
  }
  for (int o = 0; o < P; ++o) {
    const int64_t bsize = part_.size(o);
    const int nch = num_chunks(bsize, cfg_.max_chunk);
    const float* row = reinterpret_cast<const float*>(mine + sym_.off_sink) + sym_.sink_elem(slot, o);
    for (int ch = 0; ch < nch; ++ch) {
      const int64_t cb = chunk_begin(ch, cfg_.max_chunk, bsize), ce = chunk_end(ch, cfg_.max_chunk, bsize);
      const int64_t g = part_.start(o) + cb;
      const uint64_t f = ld_acquire(rflag + sym_.arr_index(slot, o, ch));
      if (static_cast<uint32_t>(f >> 32) == tag) {
        std::memcpy(out + g, row + cb, static_cast<size_t>(ce - cb) * sizeof(float));
        if (counts != nullptr) std::fill(counts + g, counts + g + (ce - cb), static_cast<int32_t>(static_cast<uint32_t>(f)));
        info.chunks_present += 1;
      } else {  // missing chunk: zeros and count 0 (ReducedDataBuffer.scala:29-58)
        std::fill(out + g, out + g + (ce - cb), 0.f);
        if (counts != nullptr) std::fill(counts + g, counts + g + (ce - cb), 0);
      }
    }
  }
  static const bool debug = std::getenv("AAR_SHM_DEBUG") != nullptr;  // AAR_SHM_DEBUG=1: report rounds over 3 ms
  if (debug) {
    const double ms = std::chrono::duration<double, std::milli>(Clock::now() - t0).count();
    if (ms > 3.0)
      std::fprintf(stderr, "[shm] rank %d round %d took %.1f ms (gate %.1f, reduce %.1f) dest_ok %x full %d partial %d present %d forced %d\n",
                   rank, round, ms, dbg_gate, dbg_reduce, dest_ok, info.reduced_full, info.reduced_partial, info.chunks_present, forced() ? 1 : 0);
  }
  // ---- CompleteAllreduce(rank, round) to everyone
  for (int i = 0; i < P; ++i) {
    const int q = (i + rank) % P;
    if ((dead_now >> q) & 1u) continue;
    st_release(reinterpret_cast<uint32_t*>(region(q) + sym_.off_done) + rank, tag);
  }
  return info;
}

std::vector<ShmRoundInfo> ShmCluster::run_round(int round, const std::vector<const float*>& srcs,
                                                const std::vector<float*>& outs, const std::vector<int32_t*>& counts,
                                                const ShmInjection* inj) {
  const size_t P = static_cast<size_t>(cfg_.world);
  if (srcs.size() != P || outs.size() != P || counts.size() != P)
    throw std::invalid_argument("ShmCluster::run_round: one source, output and count buffer per rank");
  for (size_t r = 0; r < P; ++r)  // refuse here, on the caller's thread, what rank_round would refuse on a worker thread
    if (round < next_round_[r].load())
      throw std::invalid_argument("ShmCluster: rank " + std::to_string(r) + " already issued round " + std::to_string(round));
  std::vector<ShmRoundInfo> infos(P);
  std::vector<std::exception_ptr> errors(P);
  std::vector<std::thread> threads;
  for (size_t r = 0; r < P; ++r)
    threads.emplace_back([&, r] {
      try {
        infos[r] = rank_round(static_cast<int>(r), round, srcs[r], outs[r], counts[r], inj);
      } catch (...) {
        errors[r] = std::current_exception();
      }
    });
  for (auto& t : threads) t.join();
  for (auto& e : errors)
    if (e) std::rethrow_exception(e);
  return infos;
}

ShmFreeRunStats ShmCluster::run_free(int rounds, double stall_prob, int stall_us, uint64_t seed) {
  const int P = cfg_.world;
  const size_t n = static_cast<size_t>(cfg_.n);
  std::vector<ShmFreeRunStats> per(static_cast<size_t>(P));
  std::vector<std::thread> threads;
  const auto t0 = Clock::now();
  const int first = next_round();
  for (int r = 0; r < P; ++r) progress_[static_cast<size_t>(r)].store(first, std::memory_order_relaxed);
  for (int r = 0; r < P; ++r) {
    threads.emplace_back([&, r] {
      std::mt19937_64 rng(seed * 1000003ull + static_cast<uint64_t>(r));
      std::uniform_real_distribution<double> u(0.0, 1.0);
      std::vector<float> src(n, static_cast<float>(1u << r)), out(n);
      std::vector<int32_t> cnt(n);
      ShmFreeRunStats& st = per[static_cast<size_t>(r)];
      for (int round = first; round < first + rounds; ++round) {
        progress_[static_cast<size_t>(r)].store(round, std::memory_order_relaxed);
        if (stall_prob > 0 && u(rng) < stall_prob) std::this_thread::sleep_for(std::chrono::microseconds(stall_us));
        int front = 0;
        for (int q = 0; q < P; ++q) front = std::max(front, progress_[static_cast<size_t>(q)].load(std::memory_order_relaxed));
        if (front - round > cfg_.max_lag) {  // more than maxLag behind: close everything up to the lag window now
          force_complete(r, front - cfg_.max_lag - 1);
          st.forced += 1;
        }
        const ShmRoundInfo info = rank_round(r, round, src.data(), out.data(), cnt.data(), nullptr);
        st.timeouts += info.timed_out ? 1 : 0;
        for (size_t i = 0; i < n; ++i) {
          const uint32_t bits = static_cast<uint32_t>(out[i]);
          if (static_cast<float>(bits) != out[i] || __builtin_popcount(bits) != cnt[i] || (bits >> P) != 0) {
            if (st.mismatches < 3 && std::getenv("AAR_SHM_DEBUG") != nullptr)
              std::fprintf(stderr, "[shm] rank %d round %d elem %zu: out %g count %d\n", r, round, i, out[i], cnt[i]);
            st.mismatches += 1;
          }
          if (cnt[i] == 0) st.missing_elements += 1;
          else if (cnt[i] < P) st.partial_elements += 1;
        }
      }
      progress_[static_cast<size_t>(r)].store(first + rounds - 1, std::memory_order_relaxed);
    });
  }
  for (auto& t : threads) t.join();
  ShmFreeRunStats total;
  total.rounds = rounds;
  total.seconds = std::chrono::duration<double>(Clock::now() - t0).count();
  for (const auto& s : per) {
    total.mismatches += s.mismatches;
    total.partial_elements += s.partial_elements;
    total.missing_elements += s.missing_elements;
    total.timeouts += s.timeouts;
    total.forced += s.forced;
  }
  return total;
}

}  // namespace aar
