// Fused scatter -> reduce -> broadcast allreduce round for sm_100a, over NVLink peer memory.
//
// One launch per rank per round does what the reference spreads over five message handlers:
//
//   phase S  scatter     source cast/scale, chunk slices pushed straight into each owner's
//                        `scat` slot + arrival flag            (AllreduceWorker.scala:212-238,
//                                                               AllReduceBuffer.scala:25-32)
//   phase R  reduce      wait for >= int(thReduce*P) arrivals of a chunk, fp32 sum over the
//                        arrived peers in peer order, contributor count, optional 1/count,
//                        push into every peer's `sink` slot + reduced flag
//                                                              (ScatteredDataBuffer.scala:9-32,
//                                                               AllreduceWorker.scala:177-181,252-268)
//   phase C  complete    wait for >= int(thComplete*totalChunks) reduced chunks, assemble the
//                        output and the per-element counts, zeros/count 0 for missing chunks,
//                        publish CompleteAllreduce             (ReducedDataBuffer.scala:13-66,
//                                                               AllreduceWorker.scala:270-285)
//
// Messages are stores into a peer's symmetric region followed by a release-store of a
// round-tagged flag; receivers acquire-load the flag.  Differences from the reference that come
// with polling instead of mailboxes: thresholds fire at "count >= required" with the true count
// reported (the reference fires at equality), a round may also be closed by a timeout or by the
// host's force flag (the reference's catch-up, AllreduceWorker.scala:100-106), and a sender
// skips a peer whose ring slot is still occupied by a round more than maxLag behind.
//
// Grid: block 0 of each rank is the control CTA (one warp: polls reduced flags, takes the
// completion decision); blocks 1..ctas move data.  In emulation mode blockIdx.y is the rank and
// one cooperative launch runs every rank of the job on one GPU.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>
#include <mutex>
#include <string>

#include "aar/device_layout.h"
#include "aar_b200.h"
#include "device_utils.cuh"

namespace aar {

constexpr int kThreads = 512;
constexpr int kUnitsPerThread = 4;

struct KParams {
  AarLaunch a;
  SymLayout sym;
  ScratchLayout scr;
  Partition part;
  int part_s, tpc_s;  // scatter/complete phases: elements per work item, items per full chunk
  int part_r, tpc_r;  // reduce phase: elements per work item, items per full chunk
  int part_c, tpc_c;  // complete phase
  int n_s, n_r, n_c;  // work CTAs per role (scatter / reduce+broadcast / complete); 0,0,0 = every CTA does all three
  int total_chunks;   // reduced chunks that exist this round (empty trailing blocks excluded)
  int vec_ok;         // 16-byte paths usable (alignment of pointers, step and chunk size)
};

struct Ctx {
  int rank, vr, P, slot;
  uint32_t tag;
  uint8_t* my_sym;
  uint8_t* scratch;
  volatile uint32_t* host;
  const uint32_t* force;  // device word: rounds <= force-1 close their waits immediately (catch-up)
  uint64_t deadline;
};

__device__ __forceinline__ void note_timeout(const Ctx& c) {
  if (c.host != nullptr) c.host[2] = c.tag;  // last round closed by a timeout (plain store: host memory)
}
// The force word lives in DEVICE memory (the host updates it with an async copy): a kernel that polls
// host memory over PCIe while it waits slows every system-scope fence on the GPU down by an order of
// magnitude (measured: profiles/fence_probe.md).  Waiting loops still look at it, and at the timer,
// only every kSlowPoll-th iteration.
__device__ __forceinline__ bool forced(const Ctx& c) { return ld_volatile(c.force) >= c.tag; }
constexpr int kSlowPoll = 16;
__device__ __forceinline__ bool give_up(const Ctx& c, int& spins, uint64_t deadline) {
  if ((++spins & (kSlowPoll - 1)) != 0) return false;
  if (forced(c)) return true;
  if (globaltimer_ns() > deadline) {
    note_timeout(c);
    return true;
  }
  return false;
}

// ---------------------------------------------------------------- tile movers
// src (TIn, local) -> dst (TW, possibly a peer), scaled.
template <typename TIn, typename TW>
__device__ __forceinline__ void push_tile(const TIn* __restrict__ src, TW* __restrict__ dst, int len, float scale,
                                          bool vec) {
  constexpr int E = 16 / (sizeof(TIn) > sizeof(TW) ? sizeof(TIn) : sizeof(TW));
  using UI = Unit<TIn, E>;
  using UW = Unit<TW, E>;
  if (vec) {
    const int units = len / E;
    const auto* s = reinterpret_cast<const typename UI::raw_t*>(src);
    auto* d = reinterpret_cast<typename UW::raw_t*>(dst);
    for (int u0 = threadIdx.x; u0 < units; u0 += kThreads * kUnitsPerThread) {
      typename UI::raw_t r[kUnitsPerThread];
#pragma unroll
      for (int j = 0; j < kUnitsPerThread; ++j) {
        const int u = u0 + j * kThreads;
        if (u < units) r[j] = ld_stream(s + u);
      }
#pragma unroll
      for (int j = 0; j < kUnitsPerThread; ++j) {
        const int u = u0 + j * kThreads;
        if (u < units) {
          float v[E];
          UI::unpack(r[j], v);
#pragma unroll
          for (int e = 0; e < E; ++e) v[e] *= scale;
          d[u] = UW::pack(v);
        }
      }
    }
    for (int i = units * E + threadIdx.x; i < len; i += kThreads) dst[i] = from_float<TW>(to_float(src[i]) * scale);
  } else {
    for (int i = threadIdx.x; i < len; i += kThreads) dst[i] = from_float<TW>(to_float(src[i]) * scale);
  }
}

// Sum of the arrived peers' rows -> every peer's sink row.  `rows[s]` is where peer s's contribution
// to this tile lives: my own scat slot (push mode) or peer s's staging slot over NVLink (pull mode).
// B units per thread per sweep and 8/B peers per load batch: eight 16-byte loads in flight per thread
// whatever the peer count; every element is still accumulated in peer order 0..P-1.
template <typename TW, int B>
__device__ __forceinline__ void reduce_units(const TW* const* __restrict__ rows, uint32_t mask, int P,
                                             TW* const* __restrict__ dst, int ndst, int units, float post) {
  constexpr int E = 16 / sizeof(TW);
  constexpr int PB = 8 / B;
  using U = Unit<TW, E>;
  using raw_t = typename U::raw_t;
  for (int u0 = threadIdx.x; u0 < units; u0 += kThreads * B) {
    float acc[B][E];
#pragma unroll
    for (int b = 0; b < B; ++b)
#pragma unroll
      for (int e = 0; e < E; ++e) acc[b][e] = 0.f;
    for (int s0 = 0; s0 < P; s0 += PB) {
      raw_t r[B][PB];
#pragma unroll
      for (int b = 0; b < B; ++b)
#pragma unroll
        for (int j = 0; j < PB; ++j) {
          const int u = u0 + b * kThreads, s = s0 + j;
          if (u < units && s < P && ((mask >> s) & 1u)) r[b][j] = ld_fresh(reinterpret_cast<const raw_t*>(rows[s]) + u);
        }
#pragma unroll
      for (int b = 0; b < B; ++b)
#pragma unroll
        for (int j = 0; j < PB; ++j) {
          const int u = u0 + b * kThreads, s = s0 + j;
          if (u < units && s < P && ((mask >> s) & 1u)) {
            float v[E];
            U::unpack(r[b][j], v);
#pragma unroll
            for (int e = 0; e < E; ++e) acc[b][e] += v[e];
          }
        }
    }
#pragma unroll
    for (int b = 0; b < B; ++b) {
      const int u = u0 + b * kThreads;
      if (u < units) {
#pragma unroll
        for (int e = 0; e < E; ++e) acc[b][e] *= post;
        const raw_t out = U::pack(acc[b]);
        for (int q = 0; q < ndst; ++q) reinterpret_cast<raw_t*>(dst[q])[u] = out;
      }
    }
  }
}

template <typename TW>
__device__ __forceinline__ void reduce_tile(const TW* const* __restrict__ rows, uint32_t mask, int P,
                                            TW* const* __restrict__ dst, int ndst, int len, float post, bool vec) {
  constexpr int E = 16 / sizeof(TW);
  int done = 0;
  if (vec) {
    const int units = len / E;
    constexpr int BMAX = E == 8 ? 2 : 4;  // 8-wide units carry twice the accumulators
    if (P <= 2) reduce_units<TW, BMAX>(rows, mask, P, dst, ndst, units, post);
    else if (P <= 4) reduce_units<TW, 2>(rows, mask, P, dst, ndst, units, post);
    else reduce_units<TW, 1>(rows, mask, P, dst, ndst, units, post);
    done = units * E;
  }
  for (int i = done + threadIdx.x; i < len; i += kThreads) {
    float acc = 0.f;
    for (int s = 0; s < P; ++s)
      if ((mask >> s) & 1u) acc += to_float(ld_fresh_scalar(rows[s] + i));
    const TW o = from_float<TW>(acc * post);
    for (int q = 0; q < ndst; ++q) dst[q][i] = o;
  }
}

// NVLS form of the same tile: one multimem.ld_reduce pulls the switch-reduced sum of every rank's
// staging slot, one multimem.st lands the result in every rank's sink slot.  Only used when all P
// contributions are present (an in-switch reduction cannot leave a peer out).
template <typename TW, int B>
__device__ __noinline__ void nvls_reduce_tile(const TW* __restrict__ mc_src, TW* __restrict__ mc_dst, int len,
                                                 float post) {
  constexpr int E = 16 / sizeof(TW);
  using U = Unit<TW, E>;
  const int units = len / E;
  const uint4* s = reinterpret_cast<const uint4*>(mc_src);
  uint4* d = reinterpret_cast<uint4*>(mc_dst);
  for (int u0 = threadIdx.x; u0 < units; u0 += kThreads * B) {
    uint4 r[B];
#pragma unroll
    for (int j = 0; j < B; ++j) {
      const int u = u0 + j * kThreads;
      if (u < units) r[j] = multimem_ld_reduce_add<TW>(s + u);
    }
#pragma unroll
    for (int j = 0; j < B; ++j) {
      const int u = u0 + j * kThreads;
      if (u < units) {
        if (post != 1.f) {
          float v[E];
          U::unpack(r[j], v);
#pragma unroll
          for (int e = 0; e < E; ++e) v[e] *= post;
          r[j] = U::pack(v);
        }
        multimem_st(d + u, r[j]);
      }
    }
  }
}

// sink row (local, written by the owner) -> user output + per-element counts.
template <typename TW, typename TOut>
__device__ __forceinline__ void flush_tile(const TW* __restrict__ sink, TOut* __restrict__ out,
                                           int32_t* __restrict__ counts, int len, bool present, int cnt, bool vec) {
  constexpr int E = 16 / (sizeof(TOut) > sizeof(TW) ? sizeof(TOut) : sizeof(TW));
  using UW = Unit<TW, E>;
  using UO = Unit<TOut, E>;
  if (vec) {
    const int units = len / E;
    for (int u0 = threadIdx.x; u0 < units; u0 += kThreads * kUnitsPerThread) {
      typename UW::raw_t r[kUnitsPerThread];
      if (present) {
#pragma unroll
        for (int j = 0; j < kUnitsPerThread; ++j) {
          const int u = u0 + j * kThreads;
          if (u < units) r[j] = ld_fresh(reinterpret_cast<const typename UW::raw_t*>(sink) + u);
        }
      }
#pragma unroll
      for (int j = 0; j < kUnitsPerThread; ++j) {
        const int u = u0 + j * kThreads;
        if (u < units) {
          float v[E];
          if (present) {
            UW::unpack(r[j], v);
          } else {
#pragma unroll
            for (int e = 0; e < E; ++e) v[e] = 0.f;
          }
          reinterpret_cast<typename UO::raw_t*>(out)[u] = UO::pack(v);
          if (counts != nullptr) {
            int4* cp = reinterpret_cast<int4*>(counts + static_cast<int64_t>(u) * E);
#pragma unroll
            for (int e = 0; e < E / 4; ++e) cp[e] = make_int4(cnt, cnt, cnt, cnt);
          }
        }
      }
    }
    for (int i = units * E + threadIdx.x; i < len; i += kThreads) {
      out[i] = present ? from_float<TOut>(to_float(ld_fresh_scalar(sink + i))) : from_float<TOut>(0.f);
      if (counts != nullptr) counts[i] = cnt;
    }
  } else {
    for (int i = threadIdx.x; i < len; i += kThreads) {
      out[i] = present ? from_float<TOut>(to_float(ld_fresh_scalar(sink + i))) : from_float<TOut>(0.f);
      if (counts != nullptr) counts[i] = cnt;
    }
  }
}

// ---------------------------------------------------------------- shared decisions
// Set of peers this launch may push to: a peer whose ring slot still holds a round more than
// maxLag behind is skipped for the whole launch (first CTA to look decides for all).
__device__ __forceinline__ uint32_t decide_dest_mask(const KParams& p, const Ctx& c) {
  uint32_t* word = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_fin_cnt) + 1;
  uint32_t m = ld_volatile(word);
  if (m & kDecided) return m & ~kDecided;
  const uint32_t* done = reinterpret_cast<const uint32_t*>(c.my_sym + p.sym.off_done);
  uint32_t ok = 0;
  for (int q = 0; q < c.P; ++q) {
    const uint32_t d = (q == c.rank) ? c.tag : ld_acquire_sys(done + q);
    if (d + static_cast<uint32_t>(p.a.L) >= c.tag) ok |= 1u << q;
  }
  const uint32_t old = atomicCAS(word, 0u, ok | kDecided);
  return (old & kDecided) ? (old & ~kDecided) : ok;
}

// thAllreduce gate: round r may start once `gate_count` peers have finished round r-1
// (AllreduceMaster.scala:56-62, evaluated on every rank from the replicated `done` flags).
__device__ __forceinline__ void wait_gate(const KParams& p, const Ctx& c) {
  if (p.a.gate_count <= 0 || c.tag <= 1) return;
  const uint32_t* done = reinterpret_cast<const uint32_t*>(c.my_sym + p.sym.off_done);
  int spins = 0;
  for (;;) {
    int have = 0;
    for (int q = 0; q < c.P; ++q) have += (q == c.rank || ld_volatile(done + q) + 1 >= c.tag) ? 1 : 0;
    if (have >= p.a.gate_count) break;
    if (give_up(c, spins, c.deadline)) return;
    __nanosleep(64);
  }
  __threadfence();  // acquire side of the peers' CompleteAllreduce
}

// All CTAs of a rank report here; the last one resets the scratch and publishes completion.
__device__ __forceinline__ void finish_cta(const KParams& p, const Ctx& c) {
  uint32_t* fin = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_fin_cnt);
  __threadfence();
  const uint32_t old = atomicAdd(fin, 1u);
  if (old != static_cast<uint32_t>(p.a.ctas) + 1u) return;  // ctas work CTAs + the two control warps
  st_volatile(fin, 0u);
  st_volatile(fin + 1, 0u);
  st_volatile(fin + 3, 0u);  // (fin[2] is the host's force tag: not ours to reset)
  uint32_t* cw = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_complete);
  const int words = (c.P * p.sym.cpb + 31) / 32;
  for (int i = 0; i < 4 + words; ++i) st_volatile(cw + i, 0u);
  __threadfence();  // nothing but flags follows: the data of this round was fenced when its flags were raised
  for (int i = 0; i < c.P; ++i) {  // CompleteAllreduce(rank, round) to everyone (AllreduceWorker.scala:276)
    const int q = (i + c.rank) % c.P;
    st_relaxed_sys(reinterpret_cast<uint32_t*>(p.a.sym[q] + p.sym.off_done) + c.rank, c.tag);
  }
  if (c.host != nullptr) c.host[1] = c.tag;
}

// ---------------------------------------------------------------- control CTA
// Publisher warp: the only place flags leave this rank.  Data CTAs bump a local counter per
// (destination, chunk) behind a gpu-scope fence — which, for stores that went to a peer, returns only
// once the peer has acknowledged them; this warp polls the counters and raises the flags of every chunk
// that has all its parts in: one scan, one fence, then all the flag stores.
//
// A system-scope fence on this machine waits for the NVLink traffic of the WHOLE GPU to drain (tens of
// microseconds while a round is streaming; measured in profiles/fence_probe.md), which is why it is paid
// here, once per batch, and never by a data CTA.  `fast_fences` (AAR_FAST_FENCES=1, experiments only)
// downgrades it to gpu scope; that was observed to let a flag overtake multimem.st data on 8 GPUs.
constexpr int kPubWords = 8;  // up to 256 (destination, chunk) + chunk entries per scan
__device__ void publisher_warp(const KParams& p, const Ctx& c) {
  const int lane = threadIdx.x & 31;
  const int P = c.P, cpb = p.sym.cpb;
  uint32_t* scat_cnt = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_scat_cnt);
  uint32_t* red_cnt = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_red_cnt);
  uint32_t* red_mask = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_red_mask);
  const uint32_t* fin = reinterpret_cast<const uint32_t*>(c.scratch + p.scr.off_fin_cnt);
  const int s_entries = P * cpb;
  const int entries = s_entries + cpb;
  const int64_t my_bsize = p.part.size(c.rank);
  bool last_pass = false;
  for (;;) {
    const bool closing = ld_volatile(fin + 3) >= static_cast<uint32_t>(p.a.ctas);  // read BEFORE the scan
    for (int base0 = 0; base0 < entries; base0 += 32 * kPubWords) {
      uint32_t mine = 0;  // bit w: entry base0 + 32*w + lane is complete and unpublished
#pragma unroll
      for (int w = 0; w < kPubWords; ++w) {
        const int e = base0 + 32 * w + lane;
        bool ready = false;
        if (e < s_entries) {  // ScatterBlock (d, ch): all parts out?
          const int d = e / cpb, ch = e % cpb;
          const int64_t bs = p.part.size(d);
          if (ch < num_chunks(bs, p.a.max_chunk)) {
            const int want = static_cast<int>(ceil_div(chunk_end(ch, p.a.max_chunk, bs) - chunk_begin(ch, p.a.max_chunk, bs), p.part_s));
            ready = ld_volatile(scat_cnt + e) == static_cast<uint32_t>(want);
          }
        } else if (e < entries) {  // ReduceBlock (me, ch): all parts reduced and broadcast?
          const int ch = e - s_entries;
          if (ch < num_chunks(my_bsize, p.a.max_chunk)) {
            const int want = static_cast<int>(ceil_div(chunk_end(ch, p.a.max_chunk, my_bsize) - chunk_begin(ch, p.a.max_chunk, my_bsize), p.part_r));
            ready = ld_volatile(red_cnt + ch) == static_cast<uint32_t>(want);
          }
        }
        if (ready) mine |= 1u << w;
      }
      if (!__any_sync(0xffffffffu, mine != 0)) continue;
      // Data that left this GPU (pushed scatter slices, broadcast results — unicast or multimem.st) must
      // have landed on the peers before its flag does: that takes a system-scope fence.  Slices that were
      // only staged locally (NVLS scatter) are ordered by a gpu-scope fence: the peers read them through
      // this GPU's L2.
      bool remote = false;
#pragma unroll
      for (int w = 0; w < kPubWords; ++w)
        if ((mine >> w) & 1u) remote = remote || !(p.a.algo == AAR_ALGO_NVLS && base0 + 32 * w + lane < s_entries);
      if (__any_sync(0xffffffffu, remote) && !p.a.fast_fences) __threadfence_system();
      else __threadfence();
#pragma unroll
      for (int w = 0; w < kPubWords; ++w) {
        if (!((mine >> w) & 1u)) continue;
        const int e = base0 + 32 * w + lane;
        if (e < s_entries) {
          const int d = e / cpb, ch = e % cpb;
          st_volatile(scat_cnt + e, 0u);
          st_relaxed_sys(reinterpret_cast<uint32_t*>(p.a.sym[d] + p.sym.off_arr) + p.sym.arr_index(c.slot, c.rank, ch), c.tag);
        } else {
          const int ch = e - s_entries;
          const uint32_t mask = ld_volatile(red_mask + ch) & ~kDecided;
          const uint32_t dest_ok = ld_volatile(fin + 1) & ~kDecided;
          st_volatile(red_cnt + ch, 0u);
          st_volatile(red_mask + ch, 0u);
          const uint64_t f = (static_cast<uint64_t>(c.tag) << 32) | static_cast<uint32_t>(__popc(mask));
          for (int i = 0; i < P; ++i) {
            const int q = (i + c.rank) % P;
            if ((dest_ok >> q) & 1u)
              st_relaxed_sys(reinterpret_cast<uint64_t*>(p.a.sym[q] + p.sym.off_rflag) + p.sym.arr_index(c.slot, c.rank, ch), f);
          }
        }
      }
    }
    if (last_pass) break;
    if (closing) last_pass = true;  // every work CTA is past phase R: one more scan picks up the rest
    else __nanosleep(20);
  }
  if (lane == 0) finish_cta(p, c);
}

__device__ void control_cta(const KParams& p, const Ctx& c) {
  if (threadIdx.x >= 64) return;
  if (threadIdx.x >= 32) {
    publisher_warp(p, c);
    return;
  }
  const int lane = threadIdx.x;
  const int cpb = p.sym.cpb;
  const int entries = c.P * cpb;
  const int words = (entries + 31) / 32;
  const uint64_t* rflag = reinterpret_cast<const uint64_t*>(c.my_sym + p.sym.off_rflag) + (int64_t)c.slot * entries;
  uint32_t* cw = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_complete);
  uint32_t* bitmap = cw + 4;
  int spins = 0;
  for (;;) {
    int seen = 0;
    for (int w = 0; w < words; ++w) {
      const int idx = w * 32 + lane;
      bool present = false;
      if (idx < entries) {
        const int o = idx / cpb, ch = idx % cpb;
        if (ch < num_chunks(p.part.size(o), p.a.max_chunk))
          present = static_cast<uint32_t>(ld_volatile64(rflag + idx) >> 32) == c.tag;
      }
      const uint32_t m = __ballot_sync(0xffffffffu, present);
      seen += __popc(m);
      if (lane == 0 && m != 0) st_volatile(bitmap + w, m);
    }
    bool close = seen >= p.a.req_complete || seen >= p.total_chunks;
    if (!close && lane == 0) close = give_up(c, spins, c.deadline + p.a.timeout_ns);  // owners get one timeout to
    close = __shfl_sync(0xffffffffu, close ? 1 : 0, 0) != 0 || close;                 // reduce, the round one more
    if (close) break;
    __nanosleep(200);
  }
  if (lane == 0) {
    __threadfence();
    st_volatile(cw, kDecided);
    finish_cta(p, c);
  }
}

// ---------------------------------------------------------------- TMA staging ring
// Pure copies (scatter of an unconverted source, flush of the sink slot) never touch registers:
// one elected thread streams 32 KB tiles global -> shared -> global with cp.async.bulk, three
// loads ahead of the store.  The destination may be a peer GPU (the store goes out over NVLink).
constexpr int kStageBytes = 32768;
constexpr int kStages = 4;
constexpr int kSmemBytes = (kStages + 1) * kStageBytes + 64;  // ring + one tile of counts + mbarriers

struct TmaRing {
  uint8_t* buf;
  uint8_t* cnt_tile;
  uint64_t* full;
  uint32_t issued;  // tiles this CTA has pushed through the ring so far (stage and parity bookkeeping)
};

__device__ __forceinline__ void mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
  const uint64_t t0 = globaltimer_ns();
  while (!mbar_try_wait(bar, parity))
    if (globaltimer_ns() - t0 > 4000000000ull) __trap();  // a bulk copy that never lands is a bug, not a wait
}

// Thread 0 only.  `bytes` and both pointers are 16-byte aligned.  With `counts_dst`, every data tile is
// followed by the matching stretch of the pre-filled counts tile (elem_bytes = size of a data element).
__device__ __forceinline__ void tma_copy(TmaRing& R, uint8_t* dst, const uint8_t* src, int64_t bytes,
                                         uint8_t* counts_dst, int elem_bytes) {
  const int nt = static_cast<int>((bytes + kStageBytes - 1) / kStageBytes);
  const uint32_t base = R.issued;
  auto tile_bytes = [&](int i) -> uint32_t {
    const int64_t left = bytes - static_cast<int64_t>(i) * kStageBytes;
    return static_cast<uint32_t>(left < kStageBytes ? left : kStageBytes);
  };
  auto issue_load = [&](int i) {
    const uint32_t st = (base + i) % kStages;
    mbar_expect_tx(&R.full[st], tile_bytes(i));
    tma_load_1d(R.buf + st * kStageBytes, src + static_cast<int64_t>(i) * kStageBytes, tile_bytes(i), &R.full[st]);
  };
  for (int i = 0; i < nt && i < kStages - 1; ++i) issue_load(i);
  for (int i = 0; i < nt; ++i) {
    const uint32_t st = (base + i) % kStages;
    mbar_wait_bounded(&R.full[st], ((base + i) / kStages) & 1u);
    const uint32_t tb = tile_bytes(i);
    tma_store_1d(dst + static_cast<int64_t>(i) * kStageBytes, R.buf + st * kStageBytes, tb);
    if (counts_dst != nullptr) {
      uint32_t cb = tb / elem_bytes * 4;  // bytes of int32 counts covering this tile
      uint8_t* cd = counts_dst + static_cast<int64_t>(i) * (kStageBytes / elem_bytes) * 4;
      while (cb > 0) {
        const uint32_t piece = cb < static_cast<uint32_t>(kStageBytes) ? cb : kStageBytes;
        tma_store_1d(cd, R.cnt_tile, piece);
        cd += piece;
        cb -= piece;
      }
    }
    tma_commit();
    if (i + kStages - 1 < nt) {
      tma_wait_read<1>();  // every store but the newest has finished reading its stage
      issue_load(i + kStages - 1);
    }
  }
  tma_wait_all<0>();  // all stores complete before any flag is raised or the ring is reused
  asm volatile("fence.proxy.async;" ::: "memory");
  R.issued = base + nt;
}

// ---------------------------------------------------------------- work CTAs
template <typename TIn, typename TW, bool kNvls>
__device__ void work_cta(const KParams& p, const Ctx& c) {
  extern __shared__ __align__(1024) uint8_t dyn_smem[];
  __shared__ uint32_t sh_word;
  __shared__ int sh_cnt;
  __shared__ int sh_tile_cnt;
  __shared__ TW* sh_dst[kMaxRanks];
  __shared__ const TW* sh_rows[kMaxRanks];
  const int P = c.P;
  const int cpb = p.sym.cpb;
  const int wid = blockIdx.x - 1;
  // Role of this CTA.  With enough CTAs the three phases run CONCURRENTLY on disjoint CTA groups and
  // the round is pipelined chunk by chunk: scatter feeds HBM/NVLink, reduce+broadcast is NVLink-bound,
  // completion is HBM-bound, and the flags between them are the only coupling.  With few CTAs (small
  // messages) every CTA walks through all three phases in order.
  const bool split = p.n_s > 0;
  const bool do_s = !split || wid < p.n_s;
  const bool do_r = !split || (wid >= p.n_s && wid < p.n_s + p.n_r);
  const bool do_c = !split || wid >= p.n_s + p.n_r;
  const int s_idx = wid, s_stride = split ? p.n_s : p.a.ctas;
  // (indices are clamped to 0 for CTAs outside the role: their item count is 0, so the loops do not run)
  const int r_idx = split ? max(wid - p.n_s, 0) : wid, r_stride = split ? p.n_r : p.a.ctas;
  const int c_idx = split ? max(wid - p.n_s - p.n_r, 0) : wid, c_stride = split ? p.n_c : p.a.ctas;
  const bool vec = p.vec_ok != 0;
  const bool tma = vec && p.a.use_tma != 0;
  const TIn* src = static_cast<const TIn*>(p.a.src[c.vr]);
  TIn* out = static_cast<TIn*>(p.a.out[c.vr]);
  int32_t* counts = p.a.counts[c.vr];
  uint32_t* scat_cnt = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_scat_cnt);
  uint32_t* red_cnt = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_red_cnt);
  uint32_t* red_mask = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_red_mask);
  uint64_t* trace = p.a.trace != nullptr
                        ? p.a.trace + (static_cast<int64_t>(c.vr) * (p.a.ctas + 1) + blockIdx.x) * 16
                        : nullptr;
  uint64_t wait_r = 0, wait_c = 0, pub_s = 0;
  if (trace != nullptr && threadIdx.x == 0) trace[0] = globaltimer_ns();
  TmaRing ring;
  ring.buf = dyn_smem;
  ring.cnt_tile = dyn_smem + kStages * kStageBytes;
  ring.full = reinterpret_cast<uint64_t*>(dyn_smem + (kStages + 1) * kStageBytes);
  ring.issued = 0;

  if (threadIdx.x == 0) {
    if (tma) {
      for (int i = 0; i < kStages; ++i) mbar_init(&ring.full[i], 1);
      fence_barrier_init();
    }
    sh_tile_cnt = -1;
    if (p.a.delay_ns[c.rank] != 0) {  // injected straggler
      const uint64_t until = globaltimer_ns() + p.a.delay_ns[c.rank];
      while (globaltimer_ns() < until) __nanosleep(256);
    }
    wait_gate(p, c);
    sh_word = decide_dest_mask(p, c);
    if (trace != nullptr) trace[1] = globaltimer_ns();
  }
  __syncthreads();
  const uint32_t dest_ok = sh_word;
  const uint32_t all_ranks = P >= 32 ? 0xffffffffu : ((1u << P) - 1u);
  const bool muted = (p.a.mute_mask >> c.rank) & 1u;
  const uint32_t skip = p.a.skip_dest_mask[c.rank];
  __syncthreads();

  // ---- phase S: scatter my message, chunk-major, destinations rotated from self ----
  // Push mode: the slice for owner d goes straight into d's scat row [me] over NVLink.
  // NVLS mode: it is staged in MY scat row [d]; the switch will pull and sum it.
  {
    const bool pure_copy = tma && sizeof(TIn) == sizeof(TW) && p.a.scale == 1.0f;
    const int tile = p.part_s, tpc = p.tpc_s;
    const int items = do_s ? cpb * P * tpc : 0;
    for (int item = s_idx; item < items; item += s_stride) {
      const int k = item % tpc;
      const int d = ((item / tpc) % P + c.rank) % P;
      const int ch = item / (tpc * P);
      const int64_t bsize = p.part.size(d);
      if (ch >= num_chunks(bsize, p.a.max_chunk)) continue;
      const int64_t cb = chunk_begin(ch, p.a.max_chunk, bsize), ce = chunk_end(ch, p.a.max_chunk, bsize);
      const int ntiles = static_cast<int>(ceil_div(ce - cb, tile));
      if (k >= ntiles) continue;
      if (muted || ((skip >> d) & 1u) || !((dest_ok >> d) & 1u)) continue;
      const int64_t off = cb + static_cast<int64_t>(k) * tile;
      const int len = static_cast<int>(min(static_cast<int64_t>(tile), ce - off));
      TW* dst = kNvls ? reinterpret_cast<TW*>(c.my_sym + p.sym.off_scat) + p.sym.scat_elem(c.slot, d) + off
                      : reinterpret_cast<TW*>(p.a.sym[d] + p.sym.off_scat) + p.sym.scat_elem(c.slot, c.rank) + off;
      const TIn* from = src + p.part.start(d) + off;
      if (reinterpret_cast<const void*>(from) == reinterpret_cast<const void*>(dst)) {
        // zero-copy source: the caller produced its message straight into this round's staging slot
      } else if (pure_copy && (static_cast<int64_t>(len) * sizeof(TW)) % 16 == 0) {
        if (threadIdx.x == 0)
          tma_copy(ring, reinterpret_cast<uint8_t*>(dst), reinterpret_cast<const uint8_t*>(from),
                   static_cast<int64_t>(len) * sizeof(TW), nullptr, sizeof(TW));
      } else {
        push_tile<TIn, TW>(from, dst, len, p.a.scale, vec);
      }
      __syncthreads();
      if (threadIdx.x == 0) {  // the publisher warp of the control CTA raises the flag once all parts are in
        const uint64_t w0 = trace != nullptr ? globaltimer_ns() : 0;
        __threadfence();
        atomicAdd(scat_cnt + d * cpb + ch, 1u);
        if (trace != nullptr) pub_s += globaltimer_ns() - w0;
      }
    }
  }
  if (trace != nullptr && threadIdx.x == 0) trace[2] = globaltimer_ns();

  // ---- phase C as a re-entrant step: assemble output + counts as reduced chunks are acknowledged ----
  // `complete_step(false)` handles this CTA's next completion item only if its chunk has already been
  // acknowledged, `complete_step(true)` waits for it.  The reduce loop below calls the non-blocking form after
  // every item, so the HBM-bound flush of chunks that peers have already delivered overlaps the NVLink-bound
  // reduce+broadcast instead of running as a tail after it.
  const uint32_t* cw = reinterpret_cast<const uint32_t*>(c.scratch + p.scr.off_complete);
  const uint32_t* bitmap = cw + 4;
  const uint64_t* rflag = reinterpret_cast<const uint64_t*>(c.my_sym + p.sym.off_rflag) + p.sym.arr_index(c.slot, 0, 0);
  const bool c_pure_copy = tma && sizeof(TIn) == sizeof(TW);
  const int c_tile = p.part_c, c_tpc = p.tpc_c;
  const int c_items = do_c ? cpb * P * c_tpc : 0;
  int c_item = c_idx;
  auto complete_step = [&](bool blocking) -> bool {
    int k = 0, o = 0, ch = 0, ntiles = 0;
    int64_t bsize = 0, cb = 0, ce = 0;
    for (; c_item < c_items; c_item += c_stride) {  // skip slots of the item grid that hold no data
      k = c_item % c_tpc;
      o = ((c_item / c_tpc) % P + c.rank) % P;
      ch = c_item / (c_tpc * P);
      bsize = p.part.size(o);
      if (ch >= num_chunks(bsize, p.a.max_chunk)) continue;
      cb = chunk_begin(ch, p.a.max_chunk, bsize);
      ce = chunk_end(ch, p.a.max_chunk, bsize);
      ntiles = static_cast<int>(ceil_div(ce - cb, c_tile));
      if (k < ntiles) break;
    }
    if (c_item >= c_items) return false;
    const int64_t off = cb + static_cast<int64_t>(k) * c_tile;
    const int len = static_cast<int>(min(static_cast<int64_t>(c_tile), ce - off));
    const int idx = o * cpb + ch;
    if (threadIdx.x == 0) {
      const uint64_t w0 = trace != nullptr ? globaltimer_ns() : 0;
      uint32_t state = 0;  // 0 not acknowledged yet, 1 present, 2 closed without it
      for (;;) {
        if ((ld_volatile(bitmap + (idx >> 5)) >> (idx & 31)) & 1u) {
          state = 1;
          break;
        }
        if (ld_volatile(cw) & kDecided) {  // closed: the snapshot is final
          state = ((ld_volatile(bitmap + (idx >> 5)) >> (idx & 31)) & 1u) ? 1 : 2;
          break;
        }
        if (!blocking) break;
        if (globaltimer_ns() > c.deadline + 2 * p.a.timeout_ns + 1000000000ull) {  // the control CTA never answered
          note_timeout(c);
          state = 2;
          break;
        }
        __nanosleep(64);
      }
      int cnt = 0;
      if (state == 1) {
        cnt = static_cast<int>(static_cast<uint32_t>(ld_volatile64(rflag + idx)));
        __threadfence();
      }
      sh_word = state;
      sh_cnt = cnt;
      if (trace != nullptr && blocking) wait_c += globaltimer_ns() - w0;
    }
    __syncthreads();
    const uint32_t state = sh_word;
    const int cnt = sh_cnt;
    __syncthreads();  // sh_word / sh_cnt are reused by the reduce loop
    if (state == 0) return false;
    const bool present = state == 1;
    const TW* sink = reinterpret_cast<const TW*>(c.my_sym + p.sym.off_sink) + p.sym.sink_elem(c.slot, o) + off;
    const int64_t g = p.part.start(o) + off;
    if (present && reinterpret_cast<const void*>(sink) == reinterpret_cast<const void*>(out + g)) {
      // zero-copy sink: `out` IS this round's sink slot; only the contributor counts are left to write
      if (counts != nullptr) {
        int4* cp = reinterpret_cast<int4*>(counts + g);
        const bool c16 = (reinterpret_cast<uintptr_t>(counts + g) & 15u) == 0;
        const int quads = c16 ? len / 4 : 0;
        for (int i = threadIdx.x; i < quads; i += kThreads) cp[i] = make_int4(cnt, cnt, cnt, cnt);
        for (int i = quads * 4 + threadIdx.x; i < len; i += kThreads) counts[g + i] = cnt;
      }
    } else if (c_pure_copy && present && (static_cast<int64_t>(len) * sizeof(TW)) % 16 == 0) {
      if (counts != nullptr && sh_tile_cnt != cnt) {  // (re)fill the shared tile of counts
        int4* ct = reinterpret_cast<int4*>(ring.cnt_tile);
        for (int i = threadIdx.x; i < kStageBytes / 16; i += kThreads) ct[i] = make_int4(cnt, cnt, cnt, cnt);
        fence_proxy_async_smem();
        __syncthreads();
        if (threadIdx.x == 0) sh_tile_cnt = cnt;
      }
      if (threadIdx.x == 0)
        tma_copy(ring, reinterpret_cast<uint8_t*>(out + g), reinterpret_cast<const uint8_t*>(sink),
                 static_cast<int64_t>(len) * sizeof(TW), counts != nullptr ? reinterpret_cast<uint8_t*>(counts + g) : nullptr,
                 sizeof(TW));
    } else {
      flush_tile<TW, TIn>(sink, out + g, counts != nullptr ? counts + g : nullptr, len, present, cnt, vec);
    }
    __syncthreads();
    c_item += c_stride;
    return true;
  };

  // ---- phase R: reduce my block chunk by chunk and broadcast ----
  {
    const int64_t bsize = p.part.size(c.rank);
    const int nch = num_chunks(bsize, p.a.max_chunk);
    const uint32_t* arr = reinterpret_cast<const uint32_t*>(c.my_sym + p.sym.off_arr) + p.sym.arr_index(c.slot, 0, 0);
    const int tile = p.part_r, tpc = p.tpc_r;
    const int items = do_r ? nch * tpc : 0;
    for (int item = r_idx; item < items; item += r_stride) {
      const int k = item % tpc;
      const int ch = item / tpc;
      const int64_t cb = chunk_begin(ch, p.a.max_chunk, bsize), ce = chunk_end(ch, p.a.max_chunk, bsize);
      const int ntiles = static_cast<int>(ceil_div(ce - cb, tile));
      if (k >= ntiles) continue;
      const int64_t off = cb + static_cast<int64_t>(k) * tile;
      const int len = static_cast<int>(min(static_cast<int64_t>(tile), ce - off));
      if (threadIdx.x == 0) {
        const uint64_t w0 = trace != nullptr ? globaltimer_ns() : 0;
        uint32_t m = ld_volatile(red_mask + ch);
        int spins = 0;
        while (!(m & kDecided)) {
          uint32_t have = 0;
          for (int s = 0; s < P; ++s)
            if (ld_volatile(arr + s * cpb + ch) == c.tag) have |= 1u << s;
          if (__popc(have) >= p.a.req_reduce || give_up(c, spins, c.deadline)) {
            // first CTA to see the threshold met (or to give up) fixes the contributor set of the chunk
            const uint32_t old = atomicCAS(red_mask + ch, 0u, have | kDecided);
            m = (old & kDecided) ? old : (have | kDecided);
            break;
          }
          __nanosleep(100);
          m = ld_volatile(red_mask + ch);
        }
        m &= ~kDecided;
        __threadfence();  // acquire side: the contributors' data is read from L2 / the switch, never from L1
        sh_word = m;
        if (trace != nullptr) wait_r += globaltimer_ns() - w0;
      }
      if (threadIdx.x < P) {
        const int q = (threadIdx.x + c.rank) % P;
        TW* mine = reinterpret_cast<TW*>(c.my_sym + p.sym.off_sink) + p.sym.sink_elem(c.slot, c.rank) + off;
        // A peer that is skipped this launch gets its copy redirected to my own sink row.
        sh_dst[threadIdx.x] = ((dest_ok >> q) & 1u)
                                  ? reinterpret_cast<TW*>(p.a.sym[q] + p.sym.off_sink) + p.sym.sink_elem(c.slot, c.rank) + off
                                  : mine;
        const int s = threadIdx.x;
        sh_rows[s] = kNvls ? reinterpret_cast<const TW*>(p.a.sym[s] + p.sym.off_scat) + p.sym.scat_elem(c.slot, c.rank) + off
                           : reinterpret_cast<const TW*>(c.my_sym + p.sym.off_scat) + p.sym.scat_elem(c.slot, s) + off;
      }
      __syncthreads();
      const uint32_t mask = sh_word;
      const int cnt = __popc(mask);
      const float post = (p.a.normalize && cnt > 0) ? 1.f / static_cast<float>(cnt) : 1.f;
      constexpr int EW = 16 / sizeof(TW);
      if (kNvls && mask == all_ranks && dest_ok == all_ranks && vec && len % EW == 0) {
        const int64_t so = p.sym.off_scat + (p.sym.scat_elem(c.slot, c.rank) + off) * static_cast<int64_t>(sizeof(TW));
        const int64_t to = p.sym.off_sink + (p.sym.sink_elem(c.slot, c.rank) + off) * static_cast<int64_t>(sizeof(TW));
        // 8 switch requests in flight per thread; nvls_unroll = 16 doubles that (spills a little: experiments)
        if (p.a.nvls_unroll == 16)
          nvls_reduce_tile<TW, 16>(reinterpret_cast<const TW*>(p.a.mc_base + so), reinterpret_cast<TW*>(p.a.mc_base + to), len, post);
        else
          nvls_reduce_tile<TW, 8>(reinterpret_cast<const TW*>(p.a.mc_base + so), reinterpret_cast<TW*>(p.a.mc_base + to), len, post);
      } else {
        reduce_tile<TW>(sh_rows, mask, P, sh_dst, P, len, post, vec);
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(red_cnt + ch, 1u);
      }
      if (p.a.overlap_flush) {
        while (complete_step(false)) {
        }
      }
    }
  }
  if (threadIdx.x == 0) {  // tells the publisher that this CTA will add nothing more to the S and R counters
    __threadfence();
    atomicAdd(reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_fin_cnt) + 3, 1u);
  }
  if (trace != nullptr && threadIdx.x == 0) trace[3] = globaltimer_ns();

  // ---- phase C, rest: whatever completion items the reduce loop did not get to, now blocking ----
  while (complete_step(true)) {
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (trace != nullptr) {
      trace[4] = globaltimer_ns();
      trace[5] = wait_r;
      trace[6] = wait_c;
      trace[7] = pub_s;
      trace[8] = split ? (do_s ? 1 : (do_r ? 2 : 3)) : 0;
    }
    finish_cta(p, c);
  }
}

template <typename TIn, typename TW, bool kNvls>
__global__ void __launch_bounds__(kThreads, 1) aar_twoshot_kernel(const __grid_constant__ KParams p) {
  Ctx c;
  c.vr = p.a.emu ? blockIdx.y : 0;
  c.rank = p.a.emu ? static_cast<int>(blockIdx.y) : p.a.rank;
  c.P = p.a.world;
  c.slot = p.a.slot;
  c.tag = p.a.tag;
  c.my_sym = p.a.sym[c.rank];
  c.scratch = p.a.scratch[c.vr];
  c.host = p.a.host_flags != nullptr ? p.a.host_flags + 4 * c.vr : nullptr;
  c.force = reinterpret_cast<const uint32_t*>(c.scratch + p.scr.off_fin_cnt) + 2;
  c.deadline = globaltimer_ns() + p.a.timeout_ns;
  if (blockIdx.x == 0) control_cta(p, c);
  else work_cta<TIn, TW, kNvls>(p, c);
}

// ---------------------------------------------------------------- one-shot (small messages)
// Latency path for thresholds = 1: every rank pushes its WHOLE message into every peer's one-shot
// slot, then each rank sums the P copies it holds in peer order.  One NVLink hop instead of two, no
// owner, no control CTA, no scratch traffic: CTA k owns slice k of the message end to end and has its
// own arrival flag per sender, so CTAs never talk to each other.  All ranks add the same values in
// the same order, so they still end up bit-identical (what the owner-reduces rule gives the two-shot
// path).  A sender that does not show up before the timeout is left out and the count says so.
//
// kPull selects the pull form of the same idea: every rank only STAGES its slice in its own one-shot row
// 0 (a local copy, so a gpu-scope fence orders it) and raises its flag on the peers; each rank then sums
// the P staged copies by reading them over NVLink — with a multicast alias as ONE multimem.ld_reduce per
// 16 bytes (the switch adds the P copies), otherwise with P unicast loads.  No system-scope fence and no
// second hop; the price is P x N bytes leaving every GPU, so it is meant for the 0.5-2 MB range.
template <typename TIn, typename TW, bool kPull>
__global__ void __launch_bounds__(kThreads, 1) aar_oneshot_kernel(const __grid_constant__ KParams p) {
  __shared__ uint32_t sh_mask;
  __shared__ const TW* sh_rows[kMaxRanks];
  const int vr = p.a.emu ? blockIdx.y : 0;
  const int rank = p.a.emu ? static_cast<int>(blockIdx.y) : p.a.rank;
  const int P = p.a.world, slot = p.a.slot;
  const uint32_t tag = p.a.tag;
  uint8_t* my_sym = p.a.sym[rank];
  uint8_t* scratch = p.a.scratch[vr];
  Ctx c;
  c.rank = rank;
  c.vr = vr;
  c.P = P;
  c.slot = slot;
  c.tag = tag;
  c.my_sym = my_sym;
  c.scratch = scratch;
  c.host = p.a.host_flags != nullptr ? p.a.host_flags + 4 * vr : nullptr;
  c.force = reinterpret_cast<const uint32_t*>(scratch + p.scr.off_fin_cnt) + 2;
  c.deadline = globaltimer_ns() + p.a.timeout_ns;
  const int nct = gridDim.x;
  const bool vec = p.vec_ok != 0;
  // slice of this CTA, in multiples of 8 elements so 16-byte accesses stay aligned
  const int64_t per = ceil_div(ceil_div(p.a.n, nct), 8) * 8;
  const int64_t lo = min(static_cast<int64_t>(blockIdx.x) * per, p.a.n);
  const int64_t hi = min(lo + per, p.a.n);
  const int len = static_cast<int>(hi - lo);
  const TIn* src = static_cast<const TIn*>(p.a.src[vr]);
  TIn* out = static_cast<TIn*>(p.a.out[vr]);
  int32_t* counts = p.a.counts[vr];

  if (threadIdx.x == 0) {
    if (p.a.delay_ns[rank] != 0) {
      const uint64_t until = globaltimer_ns() + p.a.delay_ns[rank];
      while (globaltimer_ns() < until) __nanosleep(256);
    }
    wait_gate(p, c);
  }
  __syncthreads();
  const bool muted = (p.a.mute_mask >> rank) & 1u;
  const uint32_t skip = p.a.skip_dest_mask[rank];

  // ---- push my slice to every peer (self first) — or, pull form, stage it once locally — then raise my flag
  if (len > 0 && !muted) {
    if (kPull) {
      TW* dst = reinterpret_cast<TW*>(my_sym + p.sym.off_os) + p.sym.os_elem(slot, 0) + lo;
      push_tile<TIn, TW>(src + lo, dst, len, p.a.scale, vec);
    } else {
      for (int i = 0; i < P; ++i) {
        const int q = (i + rank) % P;
        if ((skip >> q) & 1u) continue;
        TW* dst = reinterpret_cast<TW*>(p.a.sym[q] + p.sym.off_os) + p.sym.os_elem(slot, rank) + lo;
        push_tile<TIn, TW>(src + lo, dst, len, p.a.scale, vec);
      }
    }
  }
  if (threadIdx.x < P)  // where sender s's copy of my slice lives: its own staging row (pull) or my row s (push)
    sh_rows[threadIdx.x] = kPull ? reinterpret_cast<const TW*>(p.a.sym[threadIdx.x] + p.sym.off_os) + p.sym.os_elem(slot, 0) + lo
                                 : reinterpret_cast<const TW*>(my_sym + p.sym.off_os) + p.sym.os_elem(slot, threadIdx.x) + lo;
  __syncthreads();
  if (threadIdx.x == 0 && !muted) {
    if (p.a.fast_fences || kPull) __threadfence();  // staged locally: the peers read it through this GPU's L2
    else __threadfence_system();  // the slices must have landed on the peers before the flags do
    for (int i = 0; i < P; ++i) {
      const int q = (i + rank) % P;
      if ((skip >> q) & 1u) continue;
      st_relaxed_sys(reinterpret_cast<uint32_t*>(p.a.sym[q] + p.sym.off_os_arr) +
                         (static_cast<int64_t>(slot) * P + rank) * kMaxOneShotCtas + blockIdx.x,
                     tag);
    }
  }

  // ---- wait for every sender's slice, then reduce locally
  if (threadIdx.x == 0) {
    const uint32_t* arr = reinterpret_cast<const uint32_t*>(my_sym + p.sym.off_os_arr) +
                          static_cast<int64_t>(slot) * P * kMaxOneShotCtas + blockIdx.x;
    const uint32_t all = (1u << P) - 1u;
    uint32_t have = 0;
    int spins = 0;
    for (;;) {
      for (int s = 0; s < P; ++s)
        if (!((have >> s) & 1u) && ld_volatile(arr + s * kMaxOneShotCtas) == tag) have |= 1u << s;
      if (have == all || __popc(have) >= p.a.req_reduce && p.a.req_reduce < P) break;
      if (give_up(c, spins, c.deadline)) break;
    }
    __threadfence();
    sh_mask = have;
  }
  __syncthreads();
  const uint32_t mask = sh_mask;
  const int cnt = __popc(mask);
  const float post = (p.a.normalize && cnt > 0) ? 1.f / static_cast<float>(cnt) : 1.f;
  if (len > 0) {
    constexpr int E = 16 / (sizeof(TIn) > sizeof(TW) ? sizeof(TIn) : sizeof(TW));
    using UW = Unit<TW, E>;
    using UO = Unit<TIn, E>;
    int done = 0;
    const uint32_t all_ranks = (1u << P) - 1u;
    if (kPull && vec && sizeof(typename UW::raw_t) == 16 && p.a.mc_base != nullptr && p.a.oneshot_pull == 1 &&
        mask == all_ranks) {
      // in-switch reduction of the P staged copies: one request per 16 bytes, eight in flight per thread
      const int units = len / E;
      const uint4* mc = reinterpret_cast<const uint4*>(
          p.a.mc_base + p.sym.off_os + (p.sym.os_elem(slot, 0) + lo) * static_cast<int64_t>(sizeof(TW)));
      for (int u0 = threadIdx.x; u0 < units; u0 += kThreads * 8) {
        uint4 r[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int u = u0 + j * kThreads;
          if (u < units) r[j] = multimem_ld_reduce_add<TW>(mc + u);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int u = u0 + j * kThreads;
          if (u < units) {
            float v[E];
            UW::unpack(*reinterpret_cast<const typename UW::raw_t*>(&r[j]), v);
#pragma unroll
            for (int e = 0; e < E; ++e) v[e] *= post;
            reinterpret_cast<typename UO::raw_t*>(out + lo)[u] = UO::pack(v);
            if (counts != nullptr) {
              int4* cp = reinterpret_cast<int4*>(counts + lo + static_cast<int64_t>(u) * E);
#pragma unroll
              for (int e = 0; e < E / 4; ++e) cp[e] = make_int4(cnt, cnt, cnt, cnt);
            }
          }
        }
      }
      done = units * E;
    } else if (vec) {
      const int units = len / E;
      for (int u = threadIdx.x; u < units; u += kThreads) {
        float acc[E];
#pragma unroll
        for (int e = 0; e < E; ++e) acc[e] = 0.f;
        for (int s0 = 0; s0 < P; s0 += 8) {
          typename UW::raw_t r[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int s = s0 + j;
            if (s < P && ((mask >> s) & 1u)) r[j] = ld_fresh(reinterpret_cast<const typename UW::raw_t*>(sh_rows[s]) + u);
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int s = s0 + j;
            if (s < P && ((mask >> s) & 1u)) {
              float v[E];
              UW::unpack(r[j], v);
#pragma unroll
              for (int e = 0; e < E; ++e) acc[e] += v[e];
            }
          }
        }
#pragma unroll
        for (int e = 0; e < E; ++e) acc[e] *= post;
        reinterpret_cast<typename UO::raw_t*>(out + lo)[u] = UO::pack(acc);
        if (counts != nullptr) {
          int4* cp = reinterpret_cast<int4*>(counts + lo + static_cast<int64_t>(u) * E);
#pragma unroll
          for (int e = 0; e < E / 4; ++e) cp[e] = make_int4(cnt, cnt, cnt, cnt);
        }
      }
      done = units * E;
    }
    for (int i = done + threadIdx.x; i < len; i += kThreads) {
      float acc = 0.f;
      for (int s = 0; s < P; ++s)
        if ((mask >> s) & 1u) acc += to_float(ld_fresh_scalar(sh_rows[s] + i));
      out[lo + i] = from_float<TIn>(acc * post);
      if (counts != nullptr) counts[lo + i] = cnt;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {  // last CTA publishes CompleteAllreduce
    uint32_t* fin = reinterpret_cast<uint32_t*>(scratch + p.scr.off_fin_cnt);
    __threadfence();
    if (atomicAdd(fin, 1u) == static_cast<uint32_t>(nct - 1)) {
      st_volatile(fin, 0u);
      __threadfence();
      for (int i = 0; i < P; ++i) {
        const int q = (i + rank) % P;
        st_relaxed_sys(reinterpret_cast<uint32_t*>(p.a.sym[q] + p.sym.off_done) + rank, tag);
      }
      if (c.host != nullptr) c.host[1] = tag;
    }
  }
}

// ---------------------------------------------------------------- host side
constexpr int kMaxDevices = 64;
thread_local std::string g_error;
int fail(int code, const std::string& what) {
  g_error = what;
  return code;
}

template <typename TIn, typename TW, bool kNvls>
int launch_twoshot(const KParams& kp, cudaStream_t stream) {
  auto* fn = aar_twoshot_kernel<TIn, TW, kNvls>;
  // The shared-memory opt-in is a per-device attribute of the function: once per (instantiation, device).
  static std::atomic<bool> configured[kMaxDevices];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) return fail(AAR_ECUDA, "cudaGetDevice failed");
  if (dev >= kMaxDevices || !configured[dev].load(std::memory_order_acquire)) {
    cudaError_t ea = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    if (ea != cudaSuccess) return fail(AAR_ECUDA, std::string("cudaFuncSetAttribute(smem): ") + cudaGetErrorString(ea));
    if (dev < kMaxDevices) configured[dev].store(true, std::memory_order_release);
  }
  const int smem = kp.a.use_tma ? kSmemBytes : 0;
  dim3 grid(kp.a.ctas + 1, kp.a.emu ? kp.a.world : 1);
  cudaError_t e;
  if (kp.a.emu) {
    void* args[] = {const_cast<KParams*>(&kp)};
    e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(fn), grid, dim3(kThreads), args, smem, stream);
  } else {
    fn<<<grid, kThreads, smem, stream>>>(kp);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) return fail(AAR_ECUDA, std::string("two-shot launch: ") + cudaGetErrorString(e));
  return 1;
}

template <typename TIn, typename TW>
int launch_oneshot(const KParams& kp, cudaStream_t stream) {
  int nct = static_cast<int>(ceil_div(kp.a.n, 4096));
  if (nct > kp.a.ctas) nct = kp.a.ctas;
  if (nct > kMaxOneShotCtas) nct = kMaxOneShotCtas;
  if (nct < 1) nct = 1;
  dim3 grid(nct, kp.a.emu ? kp.a.world : 1);
  cudaError_t e;
  auto* fn = kp.a.oneshot_pull ? aar_oneshot_kernel<TIn, TW, true> : aar_oneshot_kernel<TIn, TW, false>;
  if (kp.a.emu) {
    void* args[] = {const_cast<KParams*>(&kp)};
    e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(fn), grid, dim3(kThreads), args, 0, stream);
  } else {
    fn<<<grid, kThreads, 0, stream>>>(kp);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) return fail(AAR_ECUDA, std::string("one-shot launch: ") + cudaGetErrorString(e));
  return 1;
}

template <typename TIn, typename TW>
int launch_algo(const KParams& kp, cudaStream_t stream) {
  if (kp.a.algo == AAR_ALGO_ONESHOT) return launch_oneshot<TIn, TW>(kp, stream);
  if (kp.a.algo == AAR_ALGO_NVLS) return launch_twoshot<TIn, TW, true>(kp, stream);
  return launch_twoshot<TIn, TW, false>(kp, stream);
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace aar

using namespace aar;

extern "C" {

const char* aar_last_error(void) { return g_error.c_str(); }

static int wire_bytes_of(int dtype) { return dtype == AAR_BF16 ? 2 : 4; }

int64_t aar_sym_bytes(int64_t n, int64_t max_chunk, int world, int L, int wire_dtype, int64_t os_cap,
                      int64_t block_step) {
  return SymLayout::make(n, max_chunk, world, L, wire_bytes_of(wire_dtype), os_cap, block_step).total_bytes;
}
int64_t aar_sym_flag_bytes(int64_t n, int64_t max_chunk, int world, int L, int wire_dtype, int64_t os_cap,
                           int64_t block_step) {
  return SymLayout::make(n, max_chunk, world, L, wire_bytes_of(wire_dtype), os_cap, block_step).flags_bytes();
}
int64_t aar_sym_scat_offset(int64_t n, int64_t max_chunk, int world, int L, int wire_dtype, int64_t os_cap,
                            int64_t block_step) {
  return SymLayout::make(n, max_chunk, world, L, wire_bytes_of(wire_dtype), os_cap, block_step).off_scat;
}
int64_t aar_sym_sink_offset(int64_t n, int64_t max_chunk, int world, int L, int wire_dtype, int64_t os_cap,
                            int64_t block_step) {
  return SymLayout::make(n, max_chunk, world, L, wire_bytes_of(wire_dtype), os_cap, block_step).off_sink;
}
int64_t aar_scratch_bytes(int64_t n, int64_t max_chunk, int world, int64_t block_step) {
  const SymLayout s = SymLayout::make(n, max_chunk, world, 1, 4, 0, block_step);
  return ScratchLayout::make(world, s.cpb).total_bytes;
}

int64_t aar_scratch_force_offset(int64_t n, int64_t max_chunk, int world, int64_t block_step) {
  const SymLayout s = SymLayout::make(n, max_chunk, world, 1, 4, 0, block_step);
  return ScratchLayout::make(world, s.cpb).off_fin_cnt + 8;
}

int aar_abi_sizeof_launch(void) { return static_cast<int>(sizeof(AarLaunch)); }

int aar_max_emu_ctas(int world, int device) {
  int sms = 0, per_sm = 0;
  cudaFuncSetAttribute(aar_twoshot_kernel<float, float, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return AAR_ECUDA;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, aar_twoshot_kernel<float, float, false>, kThreads, kSmemBytes) != cudaSuccess)
    return AAR_ECUDA;
  const int resident = sms * (per_sm > 0 ? per_sm : 1);
  return resident / world - 1;
}

int aar_allreduce_round(const AarLaunch* a, void* stream_v) {
  if (a == nullptr) return fail(AAR_EINVAL, "null launch");
  if (a->world < 1 || a->world > AAR_MAX_RANKS) return fail(AAR_EINVAL, "world must be 1..16");
  if (a->n <= 0 || a->max_chunk <= 0 || a->L < 1 || a->ctas < 1) return fail(AAR_EINVAL, "bad size/ring/cta arguments");
  if (a->tag == 0) return fail(AAR_EINVAL, "tag is round+1 and must be > 0");
  if (a->block_step != 0 && a->block_step * a->world < a->n) return fail(AAR_EINVAL, "block_step too small");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  KParams kp;
  kp.a = *a;
  const int wb = wire_bytes_of(a->wire_dtype);
  kp.sym = SymLayout::make(a->n, a->max_chunk, a->world, a->L, wb, a->os_cap, a->block_step);
  kp.scr = ScratchLayout::make(a->world, kp.sym.cpb);
  kp.part = Partition(a->n, a->world, kp.sym.block);
  kp.total_chunks = total_reduced_chunks(kp.part, a->max_chunk);
  const int in_bytes = a->in_dtype == AAR_BF16 ? 2 : 4;
  const int unit_elems = 16 / (in_bytes > wb ? in_bytes : wb);
  // Work items: a chunk is cut into equal parts (multiples of one full CTA sweep) so that each phase has
  // about two items per CTA; flags stay per chunk, a self-resetting counter tells the last part to raise it.
  const int64_t sweep = static_cast<int64_t>(kThreads) * kUnitsPerThread * unit_elems;
  const int64_t chunk_len = a->max_chunk < kp.sym.block ? a->max_chunk : (kp.sym.block > 0 ? kp.sym.block : 1);
  // Optional role split (see work_cta).
  kp.n_s = kp.n_r = kp.n_c = 0;
  // (Measured on 2 and 8 B200s: one CTA keeps at most ~160 KB in flight, so a phase needs most of the
  // SMs to reach HBM or NVLink speed and splitting the grid three ways loses more than the overlap
  // gains.  The split therefore only happens on request: role_s / role_c > 0.)
  if (a->role_s > 0 && a->role_c > 0 && a->role_s + a->role_c < a->ctas && a->algo != AAR_ALGO_ONESHOT) {
    kp.n_s = a->role_s;
    kp.n_c = a->role_c;
    kp.n_r = a->ctas - kp.n_s - kp.n_c;
  }
  // Work items: a chunk is cut into equal parts (multiples of one full CTA sweep) so that each phase has at
  // least two items per CTA, and no part is larger than `part_kb` (default 256 KB): all CTAs of a phase then
  // walk through memory inside a window of ctas x part, which keeps the pages they touch at any one time
  // within TLB reach even for GB-sized messages.
  const int64_t cap_elems = (static_cast<int64_t>(a->part_kb > 0 ? a->part_kb : 256) << 10) / wb;
  auto split = [&](int64_t items_without_split, int ctas_of_role, int& part, int& tpc) {
    int64_t want = ceil_div(2 * static_cast<int64_t>(ctas_of_role), items_without_split > 0 ? items_without_split : 1);
    const int64_t for_cap = ceil_div(chunk_len, cap_elems > sweep ? cap_elems : sweep);
    if (want < for_cap) want = for_cap;
    const int64_t most = ceil_div(chunk_len, sweep);
    if (want > most) want = most;
    if (want < 1) want = 1;
    int64_t len = ceil_div(ceil_div(chunk_len, want), sweep) * sweep;
    part = static_cast<int>(len);
    tpc = static_cast<int>(ceil_div(chunk_len, len));
  };
  split(static_cast<int64_t>(kp.sym.cpb) * a->world, kp.n_s > 0 ? kp.n_s : a->ctas, kp.part_s, kp.tpc_s);
  split(kp.sym.cpb, kp.n_r > 0 ? kp.n_r : a->ctas, kp.part_r, kp.tpc_r);
  split(static_cast<int64_t>(kp.sym.cpb) * a->world, kp.n_c > 0 ? kp.n_c : a->ctas, kp.part_c, kp.tpc_c);
  // 16-byte paths need every block start, chunk start and pointer aligned for both dtypes.
  const int64_t align_elems = 16 / (in_bytes < wb ? in_bytes : wb);
  bool vec = (kp.sym.block % align_elems == 0) && (a->max_chunk % align_elems == 0 || a->max_chunk >= kp.sym.block);
  const int ranks_here = a->emu ? a->world : 1;
  for (int r = 0; r < ranks_here; ++r) {
    if (a->src[r] == nullptr || a->out[r] == nullptr || a->scratch[r] == nullptr)
      return fail(AAR_EINVAL, "src/out/scratch pointer missing");
    vec = vec && aligned16(a->src[r]) && aligned16(a->out[r]) && (a->counts[r] == nullptr || aligned16(a->counts[r]));
  }
  for (int r = 0; r < a->world; ++r)
    if (a->sym[r] == nullptr) return fail(AAR_EINVAL, "symmetric base pointer missing");
  kp.vec_ok = vec ? 1 : 0;

  if (a->algo == AAR_ALGO_NVLS) {
    if (a->mc_base == nullptr || a->emu) return fail(AAR_EUNSUPPORTED, "NVLS needs a multicast alias of the symmetric region");
    if (a->req_reduce < a->world) return fail(AAR_EINVAL, "NVLS reduces in the switch over all ranks: thReduce must be 1");
  } else if (a->algo == AAR_ALGO_ONESHOT) {
    if (a->n > a->os_cap) return fail(AAR_ETOOBIG, "message larger than the one-shot capacity of the region");
    if (a->L < 2 && a->gate_count <= 0)
      return fail(AAR_EINVAL, "one-shot needs maxLag >= 1 or the thAllreduce gate (a peer could overwrite a slot in use)");
  } else if (a->algo != AAR_ALGO_TWOSHOT) {
    return fail(AAR_EUNSUPPORTED, "unknown algorithm");
  }
  if (a->in_dtype == AAR_F32 && a->wire_dtype == AAR_F32) return launch_algo<float, float>(kp, stream);
  if (a->in_dtype == AAR_BF16 && a->wire_dtype == AAR_BF16) return launch_algo<__nv_bfloat16, __nv_bfloat16>(kp, stream);
  if (a->in_dtype == AAR_BF16 && a->wire_dtype == AAR_F32) return launch_algo<__nv_bfloat16, float>(kp, stream);
  return fail(AAR_EUNSUPPORTED, "dtype combination not supported (f32/f32, bf16/bf16, bf16/f32)");
}

// ---------------------------------------------------------------- entry point that allocates nothing
// The scratch of aar_allreduce_round() is normally a buffer the caller allocated.  This variant keeps it
// in the library image instead: a fixed array in device global memory, 16 KB per (emulated) rank, that
// every GPU loading the library carries until the process exits.
#define AAR_STATIC_SCRATCH_PER_RANK 16384
#define AAR_STATIC_SCRATCH_MAX_DEVICES 64

}  // extern "C"

namespace aar {
static __device__ __align__(256) unsigned char g_static_scratch[AAR_MAX_RANKS * AAR_STATIC_SCRATCH_PER_RANK];
}

extern "C" {

int aar_allreduce_round_static_scratch(const AarLaunch* a, void* stream_v) {
  if (a == nullptr) return fail(AAR_EINVAL, "null launch");
  if (a->world < 1 || a->world > AAR_MAX_RANKS || a->n <= 0 || a->max_chunk <= 0)
    return fail(AAR_EINVAL, "bad world/size arguments");
  const int64_t need = aar_scratch_bytes(a->n, a->max_chunk, a->world, a->block_step);
  if (need > AAR_STATIC_SCRATCH_PER_RANK)
    return fail(AAR_ETOOBIG, "this configuration needs " + std::to_string(need) +
                                 " bytes of scratch per rank, the built-in scratch holds " +
                                 std::to_string(AAR_STATIC_SCRATCH_PER_RANK) +
                                 ": use fewer chunks per block (a larger maxChunkSize) or aar_allreduce_round() "
                                 "with a scratch buffer of your own");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(AAR_ECUDA, std::string("cudaGetDevice: ") + cudaGetErrorString(e));
  if (dev < 0 || dev >= AAR_STATIC_SCRATCH_MAX_DEVICES) return fail(AAR_EUNSUPPORTED, "device ordinal beyond 63");

  // Set-up, once per device, with that device current: in host code the array's name is only a stand-in,
  // the address a kernel can use comes from cudaGetSymbolAddress().
  static std::mutex mu;
  static unsigned char* base_of[AAR_STATIC_SCRATCH_MAX_DEVICES] = {};
  unsigned char* base = nullptr;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (base_of[dev] == nullptr) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_static_scratch);
      if (e != cudaSuccess || p == nullptr)
        return fail(AAR_ECUDA, std::string("cudaGetSymbolAddress(built-in scratch): ") + cudaGetErrorString(e));
      base_of[dev] = static_cast<unsigned char*>(p);
    }
    base = base_of[dev];
  }

  AarLaunch with_scratch = *a;
  const int ranks_here = a->emu ? a->world : 1;
  for (int r = 0; r < AAR_MAX_RANKS; ++r)
    with_scratch.scratch[r] = r < ranks_here ? base + static_cast<size_t>(r) * AAR_STATIC_SCRATCH_PER_RANK : nullptr;
  // The array still holds whatever the previous call left (counters of a round that timed out, another
  // geometry's layout, a force tag): everything the kernel expects to find zero is cleared first, on the
  // caller's stream so the order with earlier launches on that stream is kept.
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  e = cudaMemsetAsync(base, 0, static_cast<size_t>(ranks_here) * AAR_STATIC_SCRATCH_PER_RANK, stream);
  if (e != cudaSuccess) return fail(AAR_ECUDA, std::string("clearing the built-in scratch: ") + cudaGetErrorString(e));
  return aar_allreduce_round(&with_scratch, stream_v);
}

}  // extern "C"
