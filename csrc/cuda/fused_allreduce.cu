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
#include <map>
#include <mutex>
#include <utility>
#include <string>
#include <type_traits>

#include "aar/device_layout.h"
#include "aar_b200.h"
#include "device_utils.cuh"
#include "round_common.cuh"
#include "stream_kernel.cuh"

namespace aar {

constexpr int kUnitsPerThread = 4;

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
                                             TW* const* __restrict__ dst, int ndst, TW* mc_dst, int units, float post) {
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
        if (mc_dst != nullptr) multimem_st(reinterpret_cast<uint4*>(mc_dst) + u, out);  // one store, the switch replicates it
        else
          for (int q = 0; q < ndst; ++q) reinterpret_cast<raw_t*>(dst[q])[u] = out;
      }
    }
  }
}

template <typename TW>
__device__ __forceinline__ void reduce_tile(const TW* const* __restrict__ rows, uint32_t mask, int P,
                                            TW* const* __restrict__ dst, int ndst, TW* mc_dst, int len, float post, bool vec) {
  constexpr int E = 16 / sizeof(TW);
  int done = 0;
  if (vec) {
    const int units = len / E;
    constexpr int BMAX = E == 8 ? 2 : 4;  // 8-wide units carry twice the accumulators
    if (P <= 2) reduce_units<TW, BMAX>(rows, mask, P, dst, ndst, mc_dst, units, post);
    else if (P <= 4) reduce_units<TW, 2>(rows, mask, P, dst, ndst, mc_dst, units, post);
    else reduce_units<TW, 1>(rows, mask, P, dst, ndst, mc_dst, units, post);
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
  }
  if (threadIdx.x < 32) {  // gate, destination mask and skip notices: one lane per peer, one ballot each
    if (threadIdx.x == 0 && p.a.delay_ns[c.rank] != 0) {  // injected straggler
      const uint64_t until = globaltimer_ns() + p.a.delay_ns[c.rank];
      while (globaltimer_ns() < until) __nanosleep(256);
    }
    __syncwarp();
    wait_gate_warp(p, c);
    const uint32_t m = decide_dest_mask_warp(p, c);
    if (threadIdx.x == 0) {
      sh_word = m;
      if (trace != nullptr) trace[1] = globaltimer_ns();
    }
  }
  __syncthreads();
  const uint32_t dest_ok = sh_word;
  const uint32_t all_ranks = all_ranks_mask(P);
  const bool muted = (p.a.mute_mask >> c.rank) & 1u;
  const uint32_t skip = p.a.skip_dest_mask[c.rank];
  __syncthreads();

  // ---- phase S: scatter my message, chunk-major, destinations rotated from self ----
  // Push mode: the slice for owner d goes straight into d's scat row [me] over NVLink.
  // NVLS mode: it is staged in MY scat row [d]; the switch will pull and sum it.
  {
    const bool pure_copy = tma && sizeof(TIn) == sizeof(TW) && p.a.scale == 1.0f;
    const int tile = p.part_s, tpc = p.tpc_s;
    const int items = (do_s && !p.zc_src) ? cpb * P * tpc : 0;  // (zero-copy source: the publisher raised the flags at launch)
    for (int item = s_idx; item < items; item += s_stride) {
      const int k = item % tpc;
      const int d = ((item / tpc) % P + c.rank) % P;
      const int ch = item / (tpc * P);
      const int64_t bsize = p.part.size(d);
      if (ch >= p.nch[d]) continue;
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
      if (ch >= p.nch[o]) continue;
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
      if (state == 1) cnt = static_cast<int>(static_cast<uint32_t>(ld_acquire_sys(rflag + idx)));  // acquire: the owner's data
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
    const int nch = p.nch[c.rank];
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
      if (threadIdx.x < 32) {
        const uint64_t w0 = trace != nullptr ? globaltimer_ns() : 0;
        // first warp to see the threshold met (or nothing more to come, or to give up) fixes the contributor set
        const uint32_t m = wait_chunk_mask(p, c, ch, threadIdx.x);
        if (threadIdx.x == 0) {
          sh_word = m;
          if (trace != nullptr) wait_r += globaltimer_ns() - w0;
        }
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
        // ReduceBlock to everyone: P unicast stores per unit, or (mc_bcast, NVLS alias present, nobody skipped) one
        // multimem.st that the switch replicates — independent of how the sum was formed (AllreduceWorker.scala:252-268)
        TW* mc_dst = nullptr;
        if (p.a.mc_bcast && !p.a.emu && p.a.mc_base != nullptr && dest_ok == all_ranks && vec && len % EW == 0)
          mc_dst = reinterpret_cast<TW*>(p.a.mc_base + p.sym.off_sink) + p.sym.sink_elem(c.slot, c.rank) + off;
        reduce_tile<TW>(sh_rows, mask, P, sh_dst, P, mc_dst, len, post, vec);
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
  const Ctx c = make_ctx(p);
  const int vr = c.vr, rank = c.rank, P = c.P, slot = c.slot;
  const uint32_t tag = c.tag;
  uint8_t* my_sym = c.my_sym;
  uint8_t* scratch = c.scratch;
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

  // Gate, destination mask and skip notices, as in the two-shot kernels (one lane per peer).  The mask is the
  // device form of the reference's stale-round drop (AllreduceWorker.scala:172-173): a peer whose ring slot
  // still holds a round more than maxLag behind gets nothing from this launch — a rank that gave up twice in a
  // row must not overwrite a copy that a stalled peer is still summing — and is told so with a notice.
  if (threadIdx.x < 32) {
    if (threadIdx.x == 0 && p.a.delay_ns[rank] != 0) {
      const uint64_t until = globaltimer_ns() + p.a.delay_ns[rank];
      while (globaltimer_ns() < until) __nanosleep(256);
    }
    __syncwarp();
    wait_gate_warp(p, c);
    const uint32_t m = decide_dest_mask_warp(p, c);
    if (threadIdx.x == 0) sh_mask = m;
  }
  __syncthreads();
  const uint32_t dest_ok = sh_mask;
  const uint32_t all_ranks = all_ranks_mask(P);
  bool muted = (p.a.mute_mask >> rank) & 1u;
  if (kPull) {
    // Pull form: peers read MY staging row, so the hazard is a lagging peer still reading the copy of round
    // r - L that I am about to overwrite.  A mask cannot protect my own row; if anybody alive is that far
    // behind I sit this round out instead (contribute nothing) and tell everyone not to wait for me.
    const uint32_t dead = ld_volatile(fin_words(p, c) + 4);
    if (((dest_ok | dead) & all_ranks) != all_ranks) {
      muted = true;
      if (threadIdx.x < P && static_cast<int>(threadIdx.x) != rank && !((dead >> threadIdx.x) & 1u))
        st_relaxed_sys(reinterpret_cast<uint32_t*>(p.a.sym[threadIdx.x] + p.sym.off_notice) + rank, tag);
    }
  }
  __syncthreads();  // sh_mask is reused below
  const uint32_t skip = p.a.skip_dest_mask[rank] | (~dest_ok & all_ranks);

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
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    const uint32_t* arr = reinterpret_cast<const uint32_t*>(my_sym + p.sym.off_os_arr) +
                          static_cast<int64_t>(slot) * P * kMaxOneShotCtas + blockIdx.x;
    uint32_t have = 0;
    int spins = 0;
    for (;;) {
      have |= __ballot_sync(0xffffffffu, lane < P && ld_volatile(arr + lane * kMaxOneShotCtas) == tag);
      if (have == all_ranks || (__popc(have) >= p.a.req_reduce && p.a.req_reduce < P)) break;
      if ((++spins & (kSlowPoll - 1)) == 0) {
        const uint32_t gone = peers_gone(p, c) & ~have;  // dead, or told us they left us out: nothing more to come
        if (gone != 0 && __popc(have | gone) >= P) {
          have |= __ballot_sync(0xffffffffu, lane < P && ld_volatile(arr + lane * kMaxOneShotCtas) == tag);
          break;
        }
        if (__shfl_sync(0xffffffffu, (lane == 0 && give_up_now(c, c.deadline)) ? 1 : 0, 0) != 0) break;
      }
    }
    if (lane < P && ((have >> lane) & 1u)) (void)ld_acquire_sys(arr + lane * kMaxOneShotCtas);  // acquire: the senders' data
    __syncwarp();
    if (lane == 0) sh_mask = have;
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
      const uint32_t dead = ld_volatile(fin + 4);
      st_volatile(fin, 0u);
      st_volatile(fin + 1, 0u);  // the destination mask of this launch
      __threadfence();
      for (int i = 0; i < P; ++i) {
        const int q = (i + rank) % P;
        if ((dead >> q) & 1u) continue;
        st_relaxed_sys(reinterpret_cast<uint32_t*>(p.a.sym[q] + p.sym.off_done) + rank, tag);
      }
      if (c.host != nullptr) c.host[1] = tag;
    }
  }
}

// ---------------------------------------------------------------- host side
static int wire_bytes_of(int dtype) { return dtype == AAR_BF16 ? 2 : 4; }
thread_local std::string g_error;
int fail(int code, const std::string& what) {
  g_error = what;
  return code;
}

// CTAs of `fn` that can be resident at once on the current device (the kernels spin on their peers'
// flags and on their own control CTA: a CTA that is not resident would never be waited out).
template <typename F>
int resident_ctas(F* fn, int smem, int* out) {
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, int> cache;  // (kernel, device) -> resident CTAs
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) return fail(AAR_ECUDA, "cudaGetDevice failed");
  std::lock_guard<std::mutex> lock(mu);
  const auto key = std::make_pair(reinterpret_cast<const void*>(fn), dev);
  auto it = cache.find(key);
  if (it == cache.end()) {
    int sms = 0, per_sm = 0;
    if (smem > 0) {  // the shared-memory opt-in is a per-device attribute of the function
      cudaError_t ea = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      if (ea != cudaSuccess) return fail(AAR_ECUDA, std::string("cudaFuncSetAttribute(smem): ") + cudaGetErrorString(ea));
    }
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kThreads, smem) != cudaSuccess)
      return fail(AAR_ECUDA, "occupancy query failed");
    it = cache.emplace(key, sms * (per_sm > 0 ? per_sm : 1)).first;
  }
  *out = it->second;
  return 0;
}

template <typename F>
int launch_round(F* fn, const KParams& kp_in, int smem, cudaStream_t stream, const char* what) {
  int resident = 0;
  if (int rc = resident_ctas(fn, smem, &resident)) return rc;
  const int ranks_here = kp_in.a.emu ? kp_in.a.world : 1;
  // Every CTA of the launch must be resident (they wait on each other and on the peers): `ctas` is clamped
  // to what this device holds.  The work split only depends on kp.a.ctas, so clamping here is enough.
  const int most = resident / ranks_here - 1;
  if (most < 1) return fail(AAR_EINVAL, std::string(what) + ": this device cannot keep a control CTA and a work CTA for " +
                                            std::to_string(ranks_here) + " ranks resident at once");
  KParams kp = kp_in;
  if (kp.a.ctas > most) kp.a.ctas = most;
  dim3 grid(kp.a.ctas + 1, ranks_here);
  cudaError_t e;
  if (kp.a.emu) {
    void* args[] = {&kp};
    e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(fn), grid, dim3(kThreads), args, smem, stream);
  } else {
    fn<<<grid, kThreads, smem, stream>>>(kp);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) return fail(AAR_ECUDA, std::string(what) + " launch: " + cudaGetErrorString(e));
  return 1;
}

template <typename TIn, typename TW, bool kNvls>
int launch_twoshot(const KParams& kp, cudaStream_t stream) {
  // (the TMA ring is reserved even when this launch never touches it: the occupancy this kernel is
  // launched against must not depend on the launch)
  return launch_round(aar_twoshot_kernel<TIn, TW, kNvls>, kp, kSmemBytes, stream, "two-shot");
}

template <typename TW, bool kNvls>
int launch_stream(const KParams& kp, cudaStream_t stream) {
  return launch_round(aar_stream_kernel<TW, kNvls>, kp, 0, stream, "stream");
}

template <typename TIn, typename TW>
int launch_oneshot(const KParams& kp, cudaStream_t stream) {
  int nct = static_cast<int>(ceil_div(kp.a.n, 4096));  // one CTA per 4096 elements, as many as the flag array holds (`ctas` is for the two-shot kernels)
  if (nct > kMaxOneShotCtas) nct = kMaxOneShotCtas;
  auto* fn = kp.a.oneshot_pull ? aar_oneshot_kernel<TIn, TW, true> : aar_oneshot_kernel<TIn, TW, false>;
  int resident = 0;  // CTAs wait for their peers' CTAs: all of them (of every emulated rank) must be resident
  if (int rc = resident_ctas(fn, 0, &resident)) return rc;
  const int ranks_here = kp.a.emu ? kp.a.world : 1;
  if (nct > resident / ranks_here) nct = resident / ranks_here;
  if (nct < 1) nct = 1;
  dim3 grid(nct, ranks_here);
  cudaError_t e;
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
int launch_algo(const KParams& kp, cudaStream_t stream, bool stream_kernel) {
  if (kp.a.algo == AAR_ALGO_ONESHOT) return launch_oneshot<TIn, TW>(kp, stream);
  if constexpr (std::is_same<TIn, TW>::value) {
    if (stream_kernel)
      return kp.a.algo == AAR_ALGO_NVLS ? launch_stream<TW, true>(kp, stream) : launch_stream<TW, false>(kp, stream);
  }
  if (kp.a.algo == AAR_ALGO_NVLS) return launch_twoshot<TIn, TW, true>(kp, stream);
  return launch_twoshot<TIn, TW, false>(kp, stream);
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Everything aar_allreduce_round() derives from a launch description before it picks a kernel.
int prepare_round(const AarLaunch* a, KParams& kp, bool& use_stream) {
  if (a == nullptr) return fail(AAR_EINVAL, "null launch");
  if (a->world < 1 || a->world > AAR_MAX_RANKS) return fail(AAR_EINVAL, "world must be 1..16");
  if (a->n <= 0 || a->max_chunk <= 0 || a->L < 1 || a->ctas < 1) return fail(AAR_EINVAL, "bad size/ring/cta arguments");
  if (a->tag == 0) return fail(AAR_EINVAL, "tag is round+1 and must be > 0");
  if (a->slot < 0 || a->slot >= a->L) return fail(AAR_EINVAL, "slot must be round % L, in 0..L-1");
  if (!a->emu && (a->rank < 0 || a->rank >= a->world)) return fail(AAR_EINVAL, "rank must be in 0..world-1");
  if (a->os_cap < 0) return fail(AAR_EINVAL, "negative one-shot capacity");
  if (a->block_step != 0 && a->block_step * a->world < a->n) return fail(AAR_EINVAL, "block_step too small");
  // Capacity layout: the region and the scratch may have been laid out for a LARGER message than this round moves.
  // Offsets and strides (row pitch, flags per block) then come from the capacity, the partition from this round's n;
  // the kernels only ever use the strides, so rounds of different lengths share one region and one ring.
  const bool capped = a->n_cap > 0;
  if (capped) {
    if (a->n > a->n_cap) return fail(AAR_ETOOBIG, "message longer than the capacity the region was laid out for");
    if (a->block_step <= 0 || a->block_step_cap <= 0)
      return fail(AAR_EINVAL, "a capacity layout needs explicit block steps (block_step and block_step_cap)");
    if (a->block_step > a->block_step_cap || a->block_step_cap * a->world < a->n_cap)
      return fail(AAR_EINVAL, "block_step must not exceed block_step_cap, and block_step_cap x world must cover n_cap");
  } else if (a->n_cap < 0 || a->block_step_cap != 0) {
    return fail(AAR_EINVAL, "block_step_cap without n_cap");
  }
  kp.a = *a;
  const int wb = wire_bytes_of(a->wire_dtype);
  kp.sym = SymLayout::make(capped ? a->n_cap : a->n, a->max_chunk, a->world, a->L, wb, a->os_cap,
                           capped ? a->block_step_cap : a->block_step);
  if (capped) kp.sym.block = a->block_step;  // host arithmetic below is about THIS round (device code reads block_pad / cpb)
  kp.scr = ScratchLayout::make(a->world, kp.sym.cpb);
  kp.part = Partition(a->n, a->world, kp.sym.block);
  kp.total_chunks = total_reduced_chunks(kp.part, a->max_chunk);
  for (int r = 0; r < kMaxRanks; ++r) kp.nch[r] = r < a->world ? num_chunks(kp.part.size(r), a->max_chunk) : 0;
  kp.zc_src = kp.zc_out = kp.spec_counts = 0;
  kp.warps_s = kp.warps_r = 0;
  kp.self_publish = 0;
  kp.msg_chunks = 0;
  kp.nvls_chunks = a->algo == AAR_ALGO_NVLS ? kp.sym.cpb : 0;
  kp.n_s = kp.n_r = kp.n_c = 0;
  const int in_bytes = a->in_dtype == AAR_BF16 ? 2 : 4;
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
    // every rank sums the copies it holds on its own: only with all contributions required do the ranks agree
    if (a->req_reduce < a->world || a->req_complete < kp.total_chunks)
      return fail(AAR_EINVAL, "one-shot needs thReduce = thComplete = 1 (ranks reduce independently)");
  } else if (a->algo == AAR_ALGO_DIRECT) {
    if (a->world != 2) return fail(AAR_EINVAL, "the direct algorithm is for two ranks (it moves (P-1) x N per rank)");
    if (a->n > a->os_cap) return fail(AAR_ETOOBIG, "message larger than the one-shot capacity of the region (direct needs os_cap >= n)");
    if (a->L < 2 && a->gate_count <= 0)
      return fail(AAR_EINVAL, "direct needs maxLag >= 1 or the thAllreduce gate (a peer could overwrite a row in use)");
    if (a->req_reduce < a->world || a->req_complete < kp.total_chunks)
      return fail(AAR_EINVAL, "direct needs thReduce = thComplete = 1 (ranks reduce independently)");
    if (a->op != AAR_OP_ALLREDUCE) return fail(AAR_EUNSUPPORTED, "direct runs whole rounds only");
  } else if (a->algo != AAR_ALGO_TWOSHOT) {
    return fail(AAR_EUNSUPPORTED, "unknown algorithm");
  }
  if (a->op < AAR_OP_ALLREDUCE || a->op > AAR_OP_ALL_GATHER) return fail(AAR_EINVAL, "op must be AAR_OP_ALLREDUCE, _REDUCE_SCATTER or _ALL_GATHER");
  if (a->kernel < 0 || a->kernel > 2) return fail(AAR_EINVAL, "kernel must be 0 (auto), 1 (generic) or 2 (stream)");
  // The stream kernel moves whole 16-byte units only: besides aligned block and chunk starts it needs a message
  // length that is a multiple of the unit (else the last block would end in a partial unit — the generic kernel
  // has a scalar tail for that).
  const bool stream_ok = vec && a->n % align_elems == 0 && a->in_dtype == a->wire_dtype && a->algo != AAR_ALGO_ONESHOT;
  if (a->kernel == 2 && !stream_ok && a->algo != AAR_ALGO_ONESHOT)  // (the one-shot kernel is its own: `kernel` does not apply)
    return fail(AAR_EINVAL, "the stream kernel needs 16-byte aligned pointers, block step, chunk size and message length, and src dtype == wire dtype");
  use_stream = stream_ok && a->kernel != 1;
  if (a->algo == AAR_ALGO_DIRECT && (!use_stream || a->n % align_elems != 0))
    return fail(AAR_EUNSUPPORTED, "direct runs on the stream kernel only (aligned buffers and message length, src dtype == wire dtype)");
  if (a->op != AAR_OP_ALLREDUCE) {
    if (!use_stream) return fail(AAR_EUNSUPPORTED, "reduce-scatter / all-gather rounds run on the stream kernel only (aligned buffers, src dtype == wire dtype, not one-shot)");
    if (a->op == AAR_OP_REDUCE_SCATTER) {  // the round ends with my own block: its chunks always get reduced
      kp.total_chunks = kp.nch[a->emu ? 0 : a->rank];
      kp.a.req_complete = kp.total_chunks;
      if (a->emu) return fail(AAR_EUNSUPPORTED, "reduce-scatter rounds are not emulated (per-rank completion counts)");
    }
  }

  // Zero-copy: the caller's tensors ARE this round's staging / sink slots (real ranks only).
  if (!a->emu && a->algo != AAR_ALGO_ONESHOT && a->algo != AAR_ALGO_DIRECT && in_bytes == wb && a->op == AAR_OP_ALLREDUCE) {
    const uint8_t* src = static_cast<const uint8_t*>(a->src[0]);
    const uint8_t* out = static_cast<const uint8_t*>(a->out[0]);
    bool zs = a->scale == 1.0f, zo = true;
    for (int d = 0; d < a->world; ++d) {
      if (kp.part.size(d) == 0) continue;
      const uint8_t* stage = a->algo == AAR_ALGO_NVLS
                                 ? a->sym[a->rank] + kp.sym.off_scat + kp.sym.scat_elem(a->slot, d) * wb
                                 : a->sym[d] + kp.sym.off_scat + kp.sym.scat_elem(a->slot, a->rank) * wb;
      zs = zs && stage == src + kp.part.start(d) * wb;
      zo = zo && a->sym[a->rank] + kp.sym.off_sink + kp.sym.sink_elem(a->slot, d) * wb == out + kp.part.start(d) * wb;
    }
    kp.zc_src = zs ? 1 : 0;
    kp.zc_out = zo ? 1 : 0;
  }

  if (a->algo != AAR_ALGO_ONESHOT) {
    // Every CTA of the launch must be resident (they wait on each other and on the peers): `ctas` is clamped
    // to what this device holds, before the work is split.
    int resident = 0;
    const int rc = use_stream ? resident_ctas(aar_stream_kernel<float, false>, 0, &resident)
                              : resident_ctas(aar_twoshot_kernel<float, float, false>, kSmemBytes, &resident);
    if (rc != 0) return rc;
    const int most = resident / ranks_here - 1;
    if (most < 1) return fail(AAR_EINVAL, "this device cannot keep a control CTA and a work CTA per rank resident at once");
    if (kp.a.ctas > most) kp.a.ctas = most;
  }
  const int64_t chunk_len = a->max_chunk < kp.sym.block ? a->max_chunk : (kp.sym.block > 0 ? kp.sym.block : 1);
  if (chunk_len >= (int64_t{1} << 31)) return fail(AAR_ETOOBIG, "chunks of 2^31 elements or more: use a smaller maxChunkSize");
  if (use_stream && a->algo == AAR_ALGO_DIRECT) {
    // Two roles: push my message to the peer, sum "mine + the peer's" into the output.  Chunks are those of the
    // whole message; nothing is reduced by an owner, so there are no ReduceBlock flags and nothing to complete.
    const int64_t mlen = a->max_chunk < a->n ? a->max_chunk : a->n;
    kp.msg_chunks = static_cast<int>(ceil_div(a->n, a->max_chunk));
    if (kp.msg_chunks > 2 * kp.sym.cpb) return fail(AAR_EINVAL, "direct: more message chunks than flag words");
    for (int r = 0; r < kMaxRanks; ++r) kp.nch[r] = 0;
    kp.total_chunks = 0;
    kp.a.req_complete = 0;
    // (2 GPUs, 64 MB: 2-3 pushing warps per CTA 0.148 ms, 4: 0.150, 6: 0.162, 8: 0.170 — the link is full long before
    // the SMs are; more pushers only add contention)
    int ws = a->role_s > 0 ? a->role_s : 3;
    if (ws > kWarps - 2) ws = kWarps - 2;
    kp.warps_s = ws;
    kp.warps_r = kWarps - ws;
    const int64_t sweep = 32 * (16 / wb);
    const int64_t cap = a->part_sc_kb > 0 ? (static_cast<int64_t>(a->part_sc_kb) << 10) / wb : 16 * sweep;
    auto tile_of = [&](int role_warps, int& part, int& tpc) {
      const int64_t warps = static_cast<int64_t>(kp.a.ctas) * role_warps;
      int64_t t = ceil_div(ceil_div(a->n, 2 * warps), sweep) * sweep;
      if (t < 4 * sweep) t = 4 * sweep;
      if (t > cap) t = ceil_div(cap, sweep) * sweep;
      const int64_t whole = ceil_div(mlen, sweep) * sweep;
      if (t > whole) t = whole;
      part = static_cast<int>(t);
      tpc = static_cast<int>(ceil_div(mlen, t));
    };
    tile_of(kp.warps_s, kp.part_s, kp.tpc_s);
    tile_of(kp.warps_r, kp.part_r, kp.tpc_r);
    kp.part_c = kp.part_r;
    kp.tpc_c = 0;
    kp.self_publish = a->self_publish == 2 ? 1 : 0;  // default: the publisher warp batches the chunk flags
    return 0;
  }
  if (use_stream) {
    // Warps of every CTA are split between the three phases by what each has to move.  Scatter and completion are
    // copies (HBM, or NVLink egress for the push scatter); reduce+broadcast is the NVLink-bound part and gets the
    // rest.  With a zero-copy source there is nothing to scatter; with a zero-copy sink and all thresholds 1 the
    // completion phase only streams counts out of a shared tile (one lane per warp issues bulk stores).
    const bool spec = kp.zc_out && a->counts[0] != nullptr && a->req_reduce >= a->world && a->req_complete >= kp.total_chunks;
    // Hybrid NVLS: the last `push_chunks` chunks of every block bypass the switch's reduction (unicast push +
    // local sum), so that the links carry more than the in-switch reduce path can take on its own.
    // (Not with a zero-copy source: row [q] of the staging slot is the caller's slice for owner q in the NVLS layout
    // and the landing row of q's pushed slice in the push layout — a pushed chunk would overwrite source data the
    // peer has yet to read.)
    if (a->algo == AAR_ALGO_NVLS && a->push_chunks > 0 && !kp.zc_src)
      kp.nvls_chunks = kp.sym.cpb - a->push_chunks > 0 ? kp.sym.cpb - a->push_chunks : 0;
    const bool hybrid = a->algo == AAR_ALGO_NVLS && kp.nvls_chunks < kp.sym.cpb;
    int ws = ((kp.zc_src && !hybrid) || a->op == AAR_OP_ALL_GATHER) ? 0 : (kp.zc_src ? 3 : 5), wc = spec ? 1 : (kp.zc_out && a->counts[0] == nullptr ? 1 : 5);
    if (a->role_s > 0 && ws > 0) ws = a->role_s;   // (experiments: AAR_ROLE_S / AAR_ROLE_C warps per CTA)
    if (a->role_c > 0) wc = a->role_c;
    if (ws + wc > kWarps - 2) return fail(AAR_EINVAL, "role_s + role_c leave fewer than 2 of the 16 warps of a CTA for the reduce phase");
    kp.warps_s = ws;
    kp.warps_r = kWarps - ws - wc;
    // Measured on 8 GPUs: with 15 reduce warps per CTA a 1 GB round takes 3.35 ms, with 6 (the ordinary-tensor
    // split) 2.32 ms; 256 MB 0.69 vs 0.65 ms; 64 MB and below the other way round (0.180 vs 0.196 ms).  Thousands of
    // warps, each on its own tile, spread the in-flight multimem requests over more pages than the TLBs reach.
    if (static_cast<int64_t>(a->n) * wb >= (int64_t{128} << 20) && kp.warps_r > 6 && a->role_s <= 0 && a->role_c <= 0) kp.warps_r = 6;
    // Warp tiles: about two items per warp of the phase's role, between 4 and 64 warp sweeps (2-32 KB) each; flags
    // stay per chunk, a self-resetting counter tells the publisher when the last tile of a chunk is in.
    const int64_t sweep = 32 * (16 / wb);
    // The copies (scatter, completion) take small tiles: the first ScatterBlock flags of a round can go out only
    // when the first wave of scatter tiles is through, and the round ends one wave of completion tiles after the
    // last ReduceBlock flag — the shorter a wave, the shorter the fill and the drain of the pipeline.  Reduce tiles
    // each start with a flag wait, so they are larger.
    const int64_t cap_r = a->part_kb > 0 ? (static_cast<int64_t>(a->part_kb) << 10) / wb : 64 * sweep;
    const int64_t cap_sc = a->part_sc_kb > 0 ? (static_cast<int64_t>(a->part_sc_kb) << 10) / wb : 16 * sweep;
    auto tile_of = [&](int64_t phase_elems, int role_warps, int64_t cap, int& part, int& tpc) {
      const int64_t warps = static_cast<int64_t>(kp.a.ctas) * (role_warps > 0 ? role_warps : 1);
      int64_t t = ceil_div(ceil_div(phase_elems, 2 * warps), sweep) * sweep;
      if (t < 4 * sweep) t = 4 * sweep;
      if (t > cap) t = ceil_div(cap, sweep) * sweep;
      const int64_t whole = ceil_div(chunk_len, sweep) * sweep;
      if (t > whole) t = whole;
      part = static_cast<int>(t);
      tpc = static_cast<int>(ceil_div(chunk_len, t));
    };
    tile_of(a->n, ws, cap_sc, kp.part_s, kp.tpc_s);
    tile_of(kp.sym.block, kp.warps_r, cap_r, kp.part_r, kp.tpc_r);
    tile_of(a->n, wc, cap_sc, kp.part_c, kp.tpc_c);
    kp.spec_counts = spec ? 1 : 0;
    // who raises the flags: see AarLaunch.self_publish.  Measured on 2 GPUs (push, 64 MB): publisher 0.134 ms,
    // data warps 0.156 ms — a system-scope fence per chunk in the middle of the stream costs more than the
    // publisher's scan latency saves; NVLS stages locally (gpu-scope fences), where the data warps win.
    kp.self_publish = a->self_publish == 2 ? 1 : (a->self_publish == 1 ? 0 : (a->algo == AAR_ALGO_NVLS ? 1 : 0));
    if (a->op != AAR_OP_ALLREDUCE) kp.self_publish = 1;  // (the publisher warp only knows the full round's flag set)
    return 0;
  }
  const int unit_elems = 16 / (in_bytes > wb ? in_bytes : wb);
  const int64_t sweep = static_cast<int64_t>(kThreads) * kUnitsPerThread * unit_elems;
  // Optional role split (see work_cta).
  // (Measured on 2 and 8 B200s: one CTA keeps at most ~160 KB in flight, so a phase needs most of the
  // SMs to reach HBM or NVLink speed and splitting the grid three ways loses more than the overlap
  // gains.  The split therefore only happens on request: role_s / role_c > 0.)
  if (a->role_s > 0 && a->role_c > 0 && a->role_s + a->role_c < kp.a.ctas && a->algo != AAR_ALGO_ONESHOT) {
    kp.n_s = a->role_s;
    kp.n_c = a->role_c;
    kp.n_r = kp.a.ctas - kp.n_s - kp.n_c;
  }
  // Work items: a chunk is cut into equal parts (multiples of one full CTA sweep) so that each phase has at
  // least two items per CTA, and no part is larger than `part_kb` (default 256 KB): all CTAs of a phase then
  // walk through memory inside a window of ctas x part, which keeps the pages they touch at any one time
  // within TLB reach even for GB-sized messages; flags stay per chunk, a self-resetting counter tells the last
  // part to raise it.
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
  split(static_cast<int64_t>(kp.sym.cpb) * a->world, kp.n_s > 0 ? kp.n_s : kp.a.ctas, kp.part_s, kp.tpc_s);
  split(kp.sym.cpb, kp.n_r > 0 ? kp.n_r : kp.a.ctas, kp.part_r, kp.tpc_r);
  split(static_cast<int64_t>(kp.sym.cpb) * a->world, kp.n_c > 0 ? kp.n_c : kp.a.ctas, kp.part_c, kp.tpc_c);
  return 0;
}

}  // namespace aar

using namespace aar;

extern "C" {

const char* aar_last_error(void) { return g_error.c_str(); }

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

int64_t aar_scratch_dead_offset(int64_t n, int64_t max_chunk, int world, int64_t block_step) {
  const SymLayout s = SymLayout::make(n, max_chunk, world, 1, 4, 0, block_step);
  return ScratchLayout::make(world, s.cpb).off_fin_cnt + 16;
}

int aar_abi_sizeof_launch(void) { return static_cast<int>(sizeof(AarLaunch)); }

// Work CTAs per rank that fit next to each other (and their control CTA) on `device`: what `ctas` is
// clamped against.  kernel: 1 = generic (one CTA per SM), 2 = stream (two per SM).
int aar_max_ctas(int kernel, int world_on_device, int device) {
  int prev = 0, resident = 0;
  if (cudaGetDevice(&prev) != cudaSuccess) return fail(AAR_ECUDA, "cudaGetDevice failed");
  if (device != prev && cudaSetDevice(device) != cudaSuccess) return fail(AAR_ECUDA, "cudaSetDevice failed");
  int rc = kernel == 2 ? resident_ctas(aar_stream_kernel<float, false>, 0, &resident)
                       : resident_ctas(aar_twoshot_kernel<float, float, false>, kSmemBytes, &resident);
  if (device != prev) cudaSetDevice(prev);
  if (rc != 0) return rc;
  if (world_on_device < 1) world_on_device = 1;
  return resident / world_on_device - 1;
}

int aar_max_emu_ctas(int world, int device) { return aar_max_ctas(1, world, device); }

int aar_allreduce_round(const AarLaunch* a, void* stream_v) {
  KParams kp;
  bool use_stream = false;
  if (int rc = prepare_round(a, kp, use_stream)) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  if (a->in_dtype == AAR_F32 && a->wire_dtype == AAR_F32) return launch_algo<float, float>(kp, stream, use_stream);
  if (a->in_dtype == AAR_BF16 && a->wire_dtype == AAR_BF16) return launch_algo<__nv_bfloat16, __nv_bfloat16>(kp, stream, use_stream);
  if (a->in_dtype == AAR_BF16 && a->wire_dtype == AAR_F32) return launch_algo<__nv_bfloat16, float>(kp, stream, use_stream);
  return fail(AAR_EUNSUPPORTED, "dtype combination not supported (f32/f32, bf16/bf16, bf16/f32)");
}

// Work CTAs aar_allreduce_round() would actually launch for this description (after the residency clamp).
int aar_round_ctas(const AarLaunch* a) {
  KParams kp;
  bool use_stream = false;
  if (int rc = prepare_round(a, kp, use_stream)) return rc;
  return kp.a.ctas;
}

// Which kernel aar_allreduce_round() would run for this launch: 1 generic, 2 stream, 3 one-shot; < 0 = error.
int aar_round_kernel(const AarLaunch* a) {
  KParams kp;
  bool use_stream = false;
  if (int rc = prepare_round(a, kp, use_stream)) return rc;
  return a->algo == AAR_ALGO_ONESHOT ? 3 : (use_stream ? 2 : 1);
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
  const int64_t need = a->n_cap > 0 ? aar_scratch_bytes(a->n_cap, a->max_chunk, a->world, a->block_step_cap)
                                    : aar_scratch_bytes(a->n, a->max_chunk, a->world, a->block_step);
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
