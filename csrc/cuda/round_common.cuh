// Protocol pieces shared by the allreduce kernels of one round: launch parameters, bounded waits, the
// launch prologue (thAllreduce gate, back-pressure mask, skip notices), the control CTA (completion poller
// + flag publisher) and the end-of-round bookkeeping.
//
// Reference behaviour these stand for (paths under /root/reference/src/main/scala/sample/cluster/allreduce):
//   gate                 AllreduceMaster.scala:56-62      (evaluated on every rank from replicated `done` words)
//   skip unknown peers   AllreduceWorker.scala:213-216,254-256   (-> destination mask, skip notices, dead mask)
//   stale-round drop     AllreduceWorker.scala:155-156,172-173   (-> round-tagged flags)
//   completion threshold buffer/ReducedDataBuffer.scala:13-17,60-66
//   CompleteAllreduce    AllreduceWorker.scala:270-285
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "aar/device_layout.h"
#include "aar_b200.h"
#include "device_utils.cuh"

namespace aar {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;

struct KParams {
  AarLaunch a;
  SymLayout sym;
  ScratchLayout scr;
  Partition part;
  int part_s, tpc_s;  // scatter phase: elements per work item, items per full chunk
  int part_r, tpc_r;  // reduce phase
  int part_c, tpc_c;  // complete phase
  int n_s, n_r, n_c;  // generic kernel only: work CTAs per role; 0,0,0 = every CTA does all three
  int total_chunks;   // reduced chunks that exist this round (empty trailing blocks excluded)
  int nch[kMaxRanks]; // chunks of every owner's block (host-computed: keeps 64-bit divisions out of the kernels)
  int vec_ok;         // 16-byte paths usable (alignment of pointers, step and chunk size)
  int zc_src;         // the source IS this round's staging slot: phase S moves nothing, only raises flags
  int zc_out;         // the output IS this round's sink slot: phase C writes only the contributor counts
  int spec_counts;    // stream kernel: counts are written as P up front and repaired if a chunk came short
  int msg_chunks;     // direct mode: chunks of the whole message (flags are per message chunk there)
  int nvls_chunks;    // NVLS: leading chunks of every block that go through the switch; the rest take the unicast push path
  int self_publish;   // stream kernel: data warps raise their own flags, the publisher warp only sends the zero-copy ones
  int warps_s, warps_r;  // stream kernel: warps per CTA that scatter / reduce+broadcast (the rest complete)
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

__device__ __forceinline__ Ctx make_ctx(const KParams& p) {
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
  return c;
}

__device__ __forceinline__ uint32_t* fin_words(const KParams& p, const Ctx& c) {
  return reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_fin_cnt);
}
__device__ __forceinline__ uint32_t all_ranks_mask(int P) { return P >= 32 ? 0xffffffffu : ((1u << P) - 1u); }
// parts of `part` elements a chunk of `len` elements is cut into (both < 2^31: checked on the host)
__device__ __forceinline__ uint32_t parts_of(int64_t len, int part) {
  return (static_cast<uint32_t>(len) + static_cast<uint32_t>(part) - 1u) / static_cast<uint32_t>(part);
}

// Direct mode keeps one u32 arrival flag per (slot, sender, message chunk); it borrows the ReduceBlock flag area,
// which that mode does not use otherwise (2 x cpb words per (slot, sender): enough for the chunks of a whole
// message on two ranks).
__device__ __forceinline__ uint32_t* direct_flag(const KParams& p, uint8_t* region, int slot, int src, int m) {
  return reinterpret_cast<uint32_t*>(region + p.sym.off_rflag) + (static_cast<int64_t>(slot) * p.sym.P + src) * (2 * p.sym.cpb) + m;
}
__device__ __forceinline__ int64_t msg_chunk_len(const KParams& p, int m) {
  const int64_t lo = static_cast<int64_t>(m) * p.a.max_chunk;
  const int64_t hi = lo + p.a.max_chunk < p.a.n ? lo + p.a.max_chunk : p.a.n;
  return hi - lo;
}

__device__ __forceinline__ void note_timeout(const Ctx& c) {
  if (c.host != nullptr) c.host[2] = c.tag;  // last round closed by a timeout (plain store: host memory)
}
// The force word and the dead-peer mask live in DEVICE memory (the host updates them with async copies): a
// kernel that polls host memory over PCIe while it waits slows every system-scope fence on the GPU down by
// an order of magnitude (measured: profiles/fence_probe.md).  Waiting loops look at them, and at the timer,
// only every kSlowPoll-th iteration.
__device__ __forceinline__ bool forced(const Ctx& c) { return ld_volatile(c.force) >= c.tag; }
constexpr int kSlowPoll = 16;
__device__ __forceinline__ bool give_up_now(const Ctx& c, uint64_t deadline) {
  if (forced(c)) return true;
  if (globaltimer_ns() > deadline) {
    note_timeout(c);
    return true;
  }
  return false;
}
__device__ __forceinline__ bool give_up(const Ctx& c, int& spins, uint64_t deadline) {
  if ((++spins & (kSlowPoll - 1)) != 0) return false;
  return give_up_now(c, deadline);
}

// Peers that will never deliver anything for this round, as far as this rank can know: the host marked
// them dead (coordinator saw Terminated, AllreduceMaster.scala:46-52), or they told us — with a skip
// notice — that they left us out of this round or a later one.  Warp-collective; lane q answers for peer q.
// A notice is read with acquire: whoever sees notice >= tag also sees every flag that peer raised before
// (the kernel that wrote the notice started after the peer's earlier rounds had finished).
__device__ __forceinline__ uint32_t peers_gone(const KParams& p, const Ctx& c) {
  const int lane = threadIdx.x & 31;
  const uint32_t dead = ld_volatile(fin_words(p, c) + 4);
  bool gone = false;
  if (lane < c.P && lane != c.rank) {
    const uint32_t* notice = reinterpret_cast<const uint32_t*>(c.my_sym + p.sym.off_notice);
    gone = ((dead >> lane) & 1u) || ld_acquire_sys(notice + lane) >= c.tag;
  }
  return __ballot_sync(0xffffffffu, gone);
}

// ---------------------------------------------------------------- launch prologue (warp-collective)
// thAllreduce gate: round r may start once `gate_count` peers have finished round r-1
// (AllreduceMaster.scala:56-62, evaluated on every rank from the replicated `done` flags).  Peers known
// dead can never finish anything: the gate then asks for no more than the live ranks can give.
__device__ __forceinline__ void wait_gate_warp(const KParams& p, const Ctx& c) {
  if (p.a.gate_count <= 0 || c.tag <= 1) return;
  const int lane = threadIdx.x & 31;
  const uint32_t* done = reinterpret_cast<const uint32_t*>(c.my_sym + p.sym.off_done);
  int spins = 0;
  for (;;) {
    const bool fin = lane < c.P && (lane == c.rank || ld_acquire_sys(done + lane) + 1 >= c.tag);
    const int have = __popc(__ballot_sync(0xffffffffu, fin));
    const uint32_t dead = ld_volatile(fin_words(p, c) + 4) & all_ranks_mask(c.P);
    const int alive = c.P - __popc(dead);
    if (have >= (p.a.gate_count < alive ? p.a.gate_count : alive)) break;
    bool stop = false;
    if ((++spins & (kSlowPoll - 1)) == 0) stop = __shfl_sync(0xffffffffu, (lane == 0 && give_up_now(c, c.deadline)) ? 1 : 0, 0) != 0;
    if (stop) break;
    __nanosleep(64);
  }
}

// Set of peers this launch may push to.  A peer whose ring slot still holds a round more than maxLag behind
// (done[q] + L < tag) or that is known dead is skipped for the whole launch; the first warp to look decides
// for all CTAs (one CAS), and tells every skipped peer so with a skip notice: that peer will get neither
// slices nor reduced blocks from this rank for this round and must not wait for them.
__device__ __forceinline__ uint32_t decide_dest_mask_warp(const KParams& p, const Ctx& c) {
  const int lane = threadIdx.x & 31;
  uint32_t* word = fin_words(p, c) + 1;
  uint32_t m = ld_volatile(word);
  m = __shfl_sync(0xffffffffu, m, 0);
  if (m & kDecided) return m & ~kDecided;
  const uint32_t* done = reinterpret_cast<const uint32_t*>(c.my_sym + p.sym.off_done);
  const uint32_t dead = ld_volatile(fin_words(p, c) + 4);
  bool ok = false;
  if (lane < c.P) {
    // acquire: what we are about to write into q's slot must come after q's last reads of it
    const uint32_t d = (lane == c.rank) ? c.tag : ld_acquire_sys(done + lane);
    ok = (d + static_cast<uint32_t>(p.a.L) >= c.tag) && !((dead >> lane) & 1u);
  }
  const uint32_t mine = __ballot_sync(0xffffffffu, ok);
  uint32_t old = 0;
  if (lane == 0) old = atomicCAS(word, 0u, mine | kDecided);
  old = __shfl_sync(0xffffffffu, old, 0);
  if (old & kDecided) return old & ~kDecided;
  if (lane < c.P && lane != c.rank && !((mine >> lane) & 1u) && !((dead >> lane) & 1u))
    st_relaxed_sys(reinterpret_cast<uint32_t*>(p.a.sym[lane] + p.sym.off_notice) + c.rank, c.tag);
  return mine;
}

// All CTAs of a rank report here; the last one resets the scratch and publishes completion.
__device__ __forceinline__ void finish_cta(const KParams& p, const Ctx& c) {
  uint32_t* fin = fin_words(p, c);
  __threadfence();
  const uint32_t old = atomicAdd(fin, 1u);
  if (old != static_cast<uint32_t>(p.a.ctas) + 1u) return;  // ctas work CTAs + the two control warps
  const uint32_t dead = ld_volatile(fin + 4);
  st_volatile(fin, 0u);
  st_volatile(fin + 1, 0u);
  st_volatile(fin + 3, 0u);  // (fin[2] force tag and fin[4] dead mask are the host's: not ours to reset)
  uint32_t* cw = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_complete);
  const int words = (c.P * p.sym.cpb + 31) / 32;
  for (int i = 0; i < 4 + words; ++i) st_volatile(cw + i, 0u);
  __threadfence();  // nothing but flags follows: the data of this round was fenced when its flags were raised
  for (int i = 0; i < c.P; ++i) {  // CompleteAllreduce(rank, round) to everyone (AllreduceWorker.scala:276)
    const int q = (i + c.rank) % c.P;
    if ((dead >> q) & 1u) continue;
    st_relaxed_sys(reinterpret_cast<uint32_t*>(p.a.sym[q] + p.sym.off_done) + c.rank, c.tag);
  }
  if (c.host != nullptr) c.host[1] = c.tag;
}

// ---------------------------------------------------------------- control CTA
// Publisher warp: the only place flags leave this rank.  Data warps bump a local counter per
// (destination, chunk) behind a gpu-scope fence; this warp polls the counters and raises the flags of every
// chunk that has all its parts in: one scan, one fence, then all the flag stores.
//
// A system-scope fence on this machine waits for the NVLink traffic of the WHOLE GPU to drain (tens of
// microseconds while a round is streaming; measured in profiles/fence_probe.md), which is why it is paid
// here, once per batch, and never by a data warp.  `fast_fences` (AAR_FAST_FENCES=1, experiments only)
// downgrades it to gpu scope; that was observed to let a flag overtake multimem.st data on 8 GPUs.
constexpr int kPubWords = 8;  // up to 256 (destination, chunk) + chunk entries per scan
__device__ void publisher_warp(const KParams& p, const Ctx& c) {
  const int lane = threadIdx.x & 31;
  const int P = c.P, cpb = p.sym.cpb;
  uint32_t* scat_cnt = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_scat_cnt);
  uint32_t* red_cnt = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_red_cnt);
  uint32_t* red_mask = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_red_mask);
  const uint32_t* fin = fin_words(p, c);
  const int s_entries = P * cpb;
  const int entries = s_entries + cpb;
  const int64_t my_bsize = p.part.size(c.rank);
  const uint32_t all = all_ranks_mask(P);
  // the NVLS alias reaches every rank's copy of a flag word with one store
  uint8_t* const mc = (!p.a.emu && !p.a.fast_fences) ? p.a.mc_base : nullptr;

  if (p.zc_src) {
    // The caller produced its message straight into this round's staging slot before the launch: there
    // is nothing to move in phase S, the ScatterBlock flags can go out at once.
    if (lane == 0 && p.a.delay_ns[c.rank] != 0) {  // injected straggler
      const uint64_t until = globaltimer_ns() + p.a.delay_ns[c.rank];
      while (globaltimer_ns() < until) __nanosleep(256);
    }
    __syncwarp();
    wait_gate_warp(p, c);
    const uint32_t dest_ok = decide_dest_mask_warp(p, c);
    const bool muted = (p.a.mute_mask >> c.rank) & 1u;
    const uint32_t skip = p.a.skip_dest_mask[c.rank];
    if (!muted) {
      __threadfence();
      if (mc != nullptr && p.a.algo == AAR_ALGO_NVLS && dest_ok == all && skip == 0) {
        for (int ch = lane; ch < p.nvls_chunks; ch += 32)
          multimem_st_u32(reinterpret_cast<uint32_t*>(mc + p.sym.off_arr) + p.sym.arr_index(c.slot, c.rank, ch), c.tag);
      } else {
        for (int e = lane; e < s_entries; e += 32) {
          const int d = e / cpb, ch = e % cpb;
          if (((skip >> d) & 1u) || !((dest_ok >> d) & 1u)) continue;
          if (ch >= p.nch[d]) continue;
          if (p.a.algo == AAR_ALGO_NVLS && ch >= p.nvls_chunks) continue;  // (a hybrid round's push chunks are flagged by whoever scatters them)
          st_relaxed_sys(reinterpret_cast<uint32_t*>(p.a.sym[d] + p.sym.off_arr) + p.sym.arr_index(c.slot, c.rank, ch), c.tag);
        }
      }
    }
  }

  if (p.self_publish) {  // nothing to scan: the data warps raise the flags of what they finish
    if (lane == 0) finish_cta(p, c);
    return;
  }
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
          if (p.a.algo == AAR_ALGO_DIRECT) {  // entry e = message chunk e, pushed to every peer
            if (e < p.msg_chunks) ready = ld_volatile(scat_cnt + e) == parts_of(msg_chunk_len(p, e), p.part_s);
          } else if (!p.zc_src || (p.a.algo == AAR_ALGO_NVLS && ch >= p.nvls_chunks)) {
            const int64_t bs = p.part.size(d);
            if (ch < p.nch[d]) {
              const uint32_t want = parts_of(chunk_end(ch, p.a.max_chunk, bs) - chunk_begin(ch, p.a.max_chunk, bs), p.part_s);
              ready = ld_volatile(scat_cnt + e) == want;
            }
          }
        } else if (e < entries) {  // ReduceBlock (me, ch): all parts reduced and broadcast?
          const int ch = e - s_entries;
          if (ch < p.nch[c.rank]) {
            const uint32_t want = parts_of(chunk_end(ch, p.a.max_chunk, my_bsize) - chunk_begin(ch, p.a.max_chunk, my_bsize), p.part_r);
            ready = ld_volatile(red_cnt + ch) == want;
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
        if ((mine >> w) & 1u) {
          const int e = base0 + 32 * w + lane;
          remote = remote || !(e < s_entries && e % cpb < p.nvls_chunks);  // (nvls_chunks is 0 without NVLS)
        }
      if (__any_sync(0xffffffffu, remote) && !p.a.fast_fences) __threadfence_system();
      else __threadfence();
#pragma unroll
      for (int w = 0; w < kPubWords; ++w) {
        if (!((mine >> w) & 1u)) continue;
        const int e = base0 + 32 * w + lane;
        if (e < s_entries && p.a.algo == AAR_ALGO_DIRECT) {
          const uint32_t dest_ok = ld_volatile(fin + 1) & ~kDecided;
          st_volatile(scat_cnt + e, 0u);
          for (int q = 0; q < P; ++q)
            if (q != c.rank && ((dest_ok >> q) & 1u)) st_relaxed_sys(direct_flag(p, p.a.sym[q], c.slot, c.rank, e), c.tag);
        } else if (e < s_entries) {
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
          if (mc != nullptr && dest_ok == all) {
            multimem_st_u64(reinterpret_cast<uint64_t*>(mc + p.sym.off_rflag) + p.sym.arr_index(c.slot, c.rank, ch), f);
          } else {
            for (int i = 0; i < P; ++i) {
              const int q = (i + c.rank) % P;
              if ((dest_ok >> q) & 1u)
                st_relaxed_sys(reinterpret_cast<uint64_t*>(p.a.sym[q] + p.sym.off_rflag) + p.sym.arr_index(c.slot, c.rank, ch), f);
            }
          }
        }
      }
    }
    if (last_pass) break;
    if (closing) last_pass = true;  // every work CTA has added all it ever will: one more scan picks up the rest
    else __nanosleep(20);
  }
  if (lane == 0) finish_cta(p, c);
}

// Completion poller (warp 0 of the control CTA): snapshots the ReduceBlock flags of this round into a bitmap
// and takes the completion decision (ReducedDataBuffer.scala:60-66, at ">= required"): enough chunks, or
// nothing more can arrive (every missing chunk belongs to an owner that is dead or left us out), or the
// clock / the host's force tag says so.
__device__ void poller_warp(const KParams& p, const Ctx& c) {
  const int lane = threadIdx.x & 31;
  const int cpb = p.sym.cpb;
  const int entries = c.P * cpb;
  const int words = (entries + 31) / 32;
  const uint64_t* rflag = reinterpret_cast<const uint64_t*>(c.my_sym + p.sym.off_rflag) + (int64_t)c.slot * entries;
  uint32_t* cw = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_complete);
  uint32_t* bitmap = cw + 4;
  int spins = 0;
  uint32_t gone = 0;  // owners that will deliver nothing more (read BEFORE the flags of the scan that uses it)
  bool full = false;
  for (;;) {
    int seen = 0, never = 0;
    bool all_p = true;
    for (int w = 0; w < words; ++w) {
      const int idx = w * 32 + lane;
      bool present = false, exists = false, short_cnt = false;
      int o = 0;
      if (idx < entries) {
        o = idx / cpb;
        const int ch = idx % cpb;
        exists = ch < p.nch[o] && (p.a.op != AAR_OP_REDUCE_SCATTER || o == c.rank);  // (a reduce-scatter ends with my own block)
        if (exists) {
          const uint64_t f = ld_volatile64(rflag + idx);
          present = static_cast<uint32_t>(f >> 32) == c.tag;
          short_cnt = present && static_cast<uint32_t>(f) != static_cast<uint32_t>(c.P);
        }
      }
      const uint32_t m = __ballot_sync(0xffffffffu, present);
      seen += __popc(m);
      never += __popc(__ballot_sync(0xffffffffu, exists && !present && ((gone >> o) & 1u)));
      all_p = all_p && !__any_sync(0xffffffffu, short_cnt);
      if (lane == 0 && m != 0) st_volatile(bitmap + w, m);
    }
    full = all_p && seen >= p.total_chunks;
    bool close = seen >= p.a.req_complete || seen >= p.total_chunks || (never > 0 && seen + never >= p.total_chunks);
    if (!close && (++spins & (kSlowPoll - 1)) == 0) {
      gone = peers_gone(p, c);
      // owners get one timeout to reduce, the round one more
      close = __shfl_sync(0xffffffffu, (lane == 0 && give_up_now(c, c.deadline + p.a.timeout_ns)) ? 1 : 0, 0) != 0;
    }
    if (close) break;
    __nanosleep(100);
  }
  if (lane == 0) {
    st_volatile(cw + 1, full ? 1u : 0u);
    __threadfence();
    st_volatile(cw, kDecided);
    finish_cta(p, c);
  }
}

__device__ void control_cta(const KParams& p, const Ctx& c) {
  if (threadIdx.x >= 64) return;
  if (threadIdx.x >= 32) publisher_warp(p, c);
  else poller_warp(p, c);
}

}  // namespace aar
