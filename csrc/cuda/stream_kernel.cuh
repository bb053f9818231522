// Stream kernel: the same fused round as the generic kernel (fused_allreduce.cu), rebuilt for bytes in
// flight.  Same symmetric layout, same flags, same control CTA — a rank running this kernel and a rank
// running the generic one interoperate — but the data path is organised differently:
//
//   * the unit of work is a WARP tile (2-32 KB), not a CTA tile: no __syncthreads between items, a warp
//     that waits for a flag stalls nobody else, and a tile's "part done" is one fence + one atomic by
//     lane 0;
//   * no dynamic shared memory and <= 64 registers: two 512-thread CTAs per SM, all co-resident
//     (the kernel spins on its peers, so every CTA of the grid must be resident);
//   * the three phases run CONCURRENTLY on disjoint warps of every CTA (scatter / reduce+broadcast /
//     complete, split by the host in proportion to what each has to move): a scatter warp never waits, a
//     reduce warp waits only for its chunk's ScatterBlock flags, a completion warp only for its chunk's
//     ReduceBlock flag, so no wait can hold up work that is ready, the round is pipelined chunk by chunk
//     (the HBM-bound copies of one chunk overlap the NVLink-bound reduce of another) and there is no cycle
//     of waits.  Within a role, tiles are dealt chunk-major, round-robin over all CTAs — statically: handing
//     them out through one ticket counter per kind (so that warps of an exhausted role could help the others)
//     was measured on one GPU at 0.169 ms against 0.104 ms; thousands of warps on three counters serialise.  (A first version
//     interleaved the three kinds in one list per warp; a blocked reduce item then delayed the scatter
//     item behind it, and the delays chained from wave to wave: 0.205 ms for one GPU's 64 MB against
//     0.123 ms for the generic kernel.)
//   * with a multicast alias the reduce is one multimem.ld_reduce + one multimem.st per 16 bytes, eight
//     in flight per lane (NVLS); without it — or for chunks that close short of all P contributors —
//     the same tile is summed from unicast loads in peer order 0..P-1;
//   * per-chunk flags are raised either by the publisher warp of the control CTA (it scans the part counters and
//     pays one fence per batch) or by the data warp that completes the last tile of a chunk (no scan latency, one
//     fence per chunk; AarLaunch.self_publish);
//   * per-element contributor counts are pure fill traffic: one elected lane streams them out of an 8 KB
//     shared tile with cp.async.bulk (TMA) while the other lanes move data.  When the output is the sink
//     slot itself and every threshold is 1 they are written up front as P and repaired after the
//     completion decision if some chunk came short (timeout, dead or skipped peer).
//
// Requirements (the host falls back to the generic kernel otherwise): 16-byte aligned pointers, block
// step and chunk size; source dtype == wire dtype.
//
// Reference path being replaced: AllreduceWorker.scala:212-268 (scatter, broadcast),
// buffer/ScatteredDataBuffer.scala:20-32 (reduce), buffer/ReducedDataBuffer.scala:26-53 (flush with counts).
#pragma once
#include "round_common.cuh"

namespace aar {

constexpr int kCntTile = 2048;  // int32 counts in the shared fill tile: one bulk store covers 8 KB of counts

// ---------------------------------------------------------------- warp tile movers (16-byte units)
template <typename TW>
__device__ __forceinline__ uint4 scale_unit(const uint4& r, float scale) {
  constexpr int E = 16 / sizeof(TW);
  using U = Unit<TW, E>;
  float v[E];
  U::unpack(r, v);
#pragma unroll
  for (int e = 0; e < E; ++e) v[e] *= scale;
  return U::pack(v);
}

// local source -> staging slot (my own, or a peer's over NVLink), source scale fused
template <typename TW>
__device__ __forceinline__ void warp_copy_tile(const uint4* __restrict__ s, uint4* __restrict__ d, int units,
                                               float scale, int lane) {
  constexpr int U = 8;
  int u0 = lane;
  for (; u0 + 32 * (U - 1) < units; u0 += 32 * U) {  // full sweeps: U loads in flight, no predicates
    uint4 r[U];
#pragma unroll
    for (int j = 0; j < U; ++j) r[j] = ld_stream(s + u0 + 32 * j);
#pragma unroll
    for (int j = 0; j < U; ++j) d[u0 + 32 * j] = scale != 1.f ? scale_unit<TW>(r[j], scale) : r[j];
  }
  for (; u0 < units; u0 += 32) {
    const uint4 r = ld_stream(s + u0);
    d[u0] = scale != 1.f ? scale_unit<TW>(r, scale) : r;
  }
}

// NVLS: the switch sums every rank's staging slot on the way in and replicates the result on the way out.
// `bcast`: the destination is a multicast address too (allreduce); else an ordinary local one (reduce-scatter).
template <typename TW>
__device__ __forceinline__ void warp_nvls_tile(const uint4* __restrict__ mc_src, uint4* __restrict__ dst, int units,
                                               float post, bool bcast, int lane) {
  uint4* const mc_dst = dst;
  constexpr int U = 8;
  int u0 = lane;
  for (; u0 + 32 * (U - 1) < units; u0 += 32 * U) {  // full sweeps: U switch requests in flight per lane
    uint4 r[U];
#pragma unroll
    for (int j = 0; j < U; ++j) r[j] = multimem_ld_reduce_add<TW>(mc_src + u0 + 32 * j);
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const uint4 v = post != 1.f ? scale_unit<TW>(r[j], post) : r[j];
      if (bcast) multimem_st(mc_dst + u0 + 32 * j, v);
      else dst[u0 + 32 * j] = v;
    }
  }
  for (; u0 < units; u0 += 32) {
    const uint4 r = multimem_ld_reduce_add<TW>(mc_src + u0);
    const uint4 v = post != 1.f ? scale_unit<TW>(r, post) : r;
    if (bcast) multimem_st(mc_dst + u0, v);
    else dst[u0] = v;
  }
}

// Sum of the contributing peers' rows, peer order 0..P-1, fp32 accumulate -> every destination's sink row
// (P unicast stores, or one multimem.st).  UB units per lane and 4/UB peers per load batch: four 16-byte
// loads in flight per lane whatever the peer count (32 resident warps per SM keep ~64 KB in flight).
// kPull: row s is peer s's staging slot for MY block (NVLS layout, read over NVLink); else my own scat row s.
template <typename TW, int UB, bool kPull>
__device__ __forceinline__ void warp_reduce_units(const KParams& p, const Ctx& c, int64_t elem, int units, uint32_t mask,
                                                  uint32_t dest_ok, uint4* mc_dst, float post, int lane) {
  constexpr int E = 16 / sizeof(TW);
  constexpr int PB = 4 / UB;
  using U = Unit<TW, E>;
  const int P = c.P;
  auto row = [&](int s) -> const uint4* {
    const int64_t e = kPull ? p.sym.scat_elem(c.slot, c.rank) : p.sym.scat_elem(c.slot, s);
    return reinterpret_cast<const uint4*>((kPull ? p.a.sym[s] : c.my_sym) + p.sym.off_scat + (e + elem) * static_cast<int64_t>(sizeof(TW)));
  };
  const int64_t sink_off = p.sym.off_sink + (p.sym.sink_elem(c.slot, c.rank) + elem) * static_cast<int64_t>(sizeof(TW));
  // `nb` units per lane this sweep: UB on full sweeps, fewer on the tail
  auto sweep = [&](int u0, int nb) {
    float acc[UB][E];
#pragma unroll
    for (int b = 0; b < UB; ++b)
#pragma unroll
      for (int e = 0; e < E; ++e) acc[b][e] = 0.f;
    for (int s0 = 0; s0 < P; s0 += PB) {
      uint4 r[UB][PB];
#pragma unroll
      for (int j = 0; j < PB; ++j)
#pragma unroll
        for (int b = 0; b < UB; ++b) r[b][j] = make_uint4(0u, 0u, 0u, 0u);  // (bf16 / fp32 zero: adding it changes nothing)
#pragma unroll
      for (int j = 0; j < PB; ++j) {
        const int s = s0 + j;
        if (s < P && ((mask >> s) & 1u)) {
          const uint4* rp = row(s) + u0;
#pragma unroll
          for (int b = 0; b < UB; ++b)
            if (b < nb) r[b][j] = ld_fresh(rp + 32 * b);
        }
      }
#pragma unroll
      for (int j = 0; j < PB; ++j)
#pragma unroll
        for (int b = 0; b < UB; ++b) {
          float v[E];
          U::unpack(r[b][j], v);
#pragma unroll
          for (int e = 0; e < E; ++e) acc[b][e] += v[e];
        }
    }
#pragma unroll
    for (int b = 0; b < UB; ++b) {
      if (b < nb) {
        const int u = u0 + 32 * b;
#pragma unroll
        for (int e = 0; e < E; ++e) acc[b][e] *= post;
        const uint4 out = U::pack(acc[b]);
        if (mc_dst != nullptr) {
          multimem_st(mc_dst + u, out);
        } else {
          for (int i = 0; i < P; ++i) {
            const int q = (i + c.rank) % P;  // own copy first, like the reference's self-delivery
            if ((dest_ok >> q) & 1u) reinterpret_cast<uint4*>(p.a.sym[q] + sink_off)[u] = out;
          }
        }
      }
    }
  };
  int u0 = lane;
  for (; u0 + 32 * (UB - 1) < units; u0 += 32 * UB) sweep(u0, UB);
  if (UB > 1)
    for (; u0 < units; u0 += 32) sweep(u0, 1);
}

// All-gather half on its own: my block goes from the caller's tensor straight into every destination's sink row
// (ReduceBlock with a single contributor).
__device__ __forceinline__ void warp_bcast_tile(const KParams& p, const Ctx& c, const uint4* __restrict__ src, int64_t sink_off,
                                                int units, uint32_t dest_ok, uint4* mc_dst, int lane) {
  constexpr int U = 4;
  auto put = [&](int u, const uint4& v) {
    if (mc_dst != nullptr) {
      multimem_st(mc_dst + u, v);
    } else {
      for (int i = 0; i < c.P; ++i) {
        const int q = (i + c.rank) % c.P;
        if ((dest_ok >> q) & 1u) reinterpret_cast<uint4*>(p.a.sym[q] + sink_off)[u] = v;
      }
    }
  };
  int u0 = lane;
  for (; u0 + 32 * (U - 1) < units; u0 += 32 * U) {
    uint4 r[U];
#pragma unroll
    for (int j = 0; j < U; ++j) r[j] = ld_stream(src + u0 + 32 * j);
#pragma unroll
    for (int j = 0; j < U; ++j) put(u0 + 32 * j, r[j]);
  }
  for (; u0 < units; u0 += 32) put(u0, ld_stream(src + u0));
}

// Contributor counts of `n` consecutive elements (n % 4 == 0, 16-byte aligned): bulk stores out of the
// shared tile when it holds this value, LSU stores otherwise.  Returns true if lane 0 issued bulk stores.
__device__ __forceinline__ bool warp_fill_counts(int32_t* __restrict__ counts, int n, int cnt, const int32_t* tile,
                                                 int tile_value, bool tma, int lane) {
  if (tma && cnt == tile_value) {
    if (lane == 0) {
      for (int i = 0; i < n; i += kCntTile) {
        const int piece = n - i < kCntTile ? n - i : kCntTile;
        tma_store_1d(counts + i, tile, static_cast<uint32_t>(piece) * 4u);
      }
      tma_commit();
    }
    return true;
  }
  int4* cp = reinterpret_cast<int4*>(counts);
  const int4 v = make_int4(cnt, cnt, cnt, cnt);
  for (int i = lane; i < n / 4; i += 32) cp[i] = v;
  return false;
}

// sink row (written by its owner, over NVLink) -> user output
template <typename TW>
__device__ __forceinline__ void warp_flush_tile(const uint4* __restrict__ sink, uint4* __restrict__ out, int units,
                                                bool present, int lane) {
  constexpr int U = 8;
  if (!present) {
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (int u = lane; u < units; u += 32) out[u] = z;
    return;
  }
  int u0 = lane;
  for (; u0 + 32 * (U - 1) < units; u0 += 32 * U) {
    uint4 r[U];
#pragma unroll
    for (int j = 0; j < U; ++j) r[j] = ld_fresh(sink + u0 + 32 * j);
#pragma unroll
    for (int j = 0; j < U; ++j) out[u0 + 32 * j] = r[j];
  }
  for (; u0 < units; u0 += 32) out[u0] = ld_fresh(sink + u0);
}

// ---------------------------------------------------------------- waits (warp-collective)
// Contributor set of my chunk `ch`: fixed once per chunk (one CAS) so that every tile of the chunk sums the
// same peers.  Fires at ">= required" arrivals with the true set (ScatteredDataBuffer.scala:9-13 fires at
// equality), or when nothing more can arrive (every absent peer is dead or has left us out), or on the
// clock / the host's force tag.
__device__ __forceinline__ uint32_t wait_chunk_mask(const KParams& p, const Ctx& c, int ch, int lane) {
  uint32_t* red_mask = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_red_mask) + ch;
  const uint32_t* arr = reinterpret_cast<const uint32_t*>(c.my_sym + p.sym.off_arr) + p.sym.arr_index(c.slot, 0, 0) + ch;
  const int P = c.P, cpb = p.sym.cpb;
  uint32_t m;
  int spins = 0;
  for (;;) {
    m = ld_volatile(red_mask);
    m = __shfl_sync(0xffffffffu, m, 0);
    if (m & kDecided) break;
    uint32_t have = __ballot_sync(0xffffffffu, lane < P && ld_volatile(arr + lane * cpb) == c.tag);
    bool decide = __popc(have) >= p.a.req_reduce;
    if (!decide && (++spins & (kSlowPoll - 1)) == 0) {
      const uint32_t gone = peers_gone(p, c) & ~have;
      if (gone != 0 && __popc(have | gone) >= P) {
        // (flags are re-read after the notices: a slice that landed in between is simply included)
        have = __ballot_sync(0xffffffffu, lane < P && ld_volatile(arr + lane * cpb) == c.tag);
        decide = true;
      } else {
        decide = __shfl_sync(0xffffffffu, (lane == 0 && give_up_now(c, c.deadline)) ? 1 : 0, 0) != 0;
      }
    }
    if (decide) {
      uint32_t old = 0;
      if (lane == 0) old = atomicCAS(red_mask, 0u, have | kDecided);
      old = __shfl_sync(0xffffffffu, old, 0);
      m = (old & kDecided) ? old : (have | kDecided);
      break;
    }
    __nanosleep(64);
  }
  m &= ~kDecided;
  // acquire side of the contributors' ScatterBlock flags, extended to the whole warp by the barrier
  if (lane < P && ((m >> lane) & 1u)) (void)ld_acquire_sys(arr + lane * cpb);
  __syncwarp();
  return m;
}

// Has owner o's reduced chunk `ch` arrived?  Returns the contributor count (>= 0) or -1 when the round was
// closed without it.  `direct`: every threshold is 1, so the flag itself may be trusted before the poller
// has seen it (a late chunk can only ever be the right one).
__device__ __forceinline__ int wait_reduced(const KParams& p, const Ctx& c, int o, int ch, bool direct, int lane) {
  const uint32_t* cw = reinterpret_cast<const uint32_t*>(c.scratch + p.scr.off_complete);
  const int idx = o * p.sym.cpb + ch;
  const uint64_t* rf = reinterpret_cast<const uint64_t*>(c.my_sym + p.sym.off_rflag) + p.sym.arr_index(c.slot, 0, 0) + idx;
  int cnt = -1;
  if (lane == 0) {
    const uint32_t* bm = cw + 4 + (idx >> 5);
    const uint64_t stop = c.deadline + 2 * p.a.timeout_ns + 1000000000ull;
    int spins = 0;
    for (;;) {
      bool here = (ld_volatile(bm) >> (idx & 31)) & 1u;
      if (!here && direct) here = static_cast<uint32_t>(ld_volatile64(rf) >> 32) == c.tag;
      if (!here && (ld_volatile(cw) & kDecided)) {  // closed: the snapshot is final
        here = (ld_volatile(bm) >> (idx & 31)) & 1u;
        if (!here) break;
      }
      if (here) {
        cnt = static_cast<int>(static_cast<uint32_t>(ld_acquire_sys(rf)));
        break;
      }
      if ((++spins & (kSlowPoll - 1)) == 0 && globaltimer_ns() > stop) {  // the control CTA never answered
        note_timeout(c);
        break;
      }
      __nanosleep(64);
    }
  }
  cnt = __shfl_sync(0xffffffffu, cnt, 0);  // (also carries lane 0's acquire to the warp: bar.warp.sync semantics)
  return cnt;
}

// ---------------------------------------------------------------- work items
// Each item kind is its own (non-inlined) function: the register allocator then solves three small problems
// instead of one large one, and the kernel stays at 64 registers — two CTAs per SM — without spilling on the
// streaming loops.
struct WarpEnv {
  uint32_t dest_ok, skip;
  int lane;
  bool muted, full, use_tma;
  uint8_t* mc;              // multicast alias (nullptr in emulation or without NVLS)
  const int32_t* cnt_tile;  // shared tile of counts, filled with P
  uint64_t* wait_r;         // tracing only (thread 0 of a traced launch), else nullptr
  uint64_t* wait_c;
  uint64_t* first_r;
};

// S: my slice of owner d's chunk -> d's scat row [me] (push) or MY scat row [d] (NVLS staging)
template <typename TW, bool kNvls>
__device__ __noinline__ void scatter_item(const KParams& p, const Ctx& c, const WarpEnv& w, int d, int ch, int t) {
  constexpr int E = 16 / sizeof(TW);
  const int64_t bsize = p.part.size(d);
  if (ch >= p.nch[d]) return;
  const int64_t cb = chunk_begin(ch, p.a.max_chunk, bsize), ce = chunk_end(ch, p.a.max_chunk, bsize);
  const int64_t off = cb + static_cast<int64_t>(t) * p.part_s;
  if (off >= ce) return;
  if (w.muted || ((w.skip >> d) & 1u) || !((w.dest_ok >> d) & 1u)) return;
  // NVLS stages the slice in MY scat row [d] for the switch to pull; the trailing `push` chunks of a hybrid round
  // (and every chunk without NVLS) go straight into owner d's scat row [me] over NVLink.
  const bool push_ch = !kNvls || ch >= p.nvls_chunks;
  if (p.zc_src && !push_ch) return;  // staged by the caller before the launch, flagged by the publisher warp
  const int len = static_cast<int>(min(static_cast<int64_t>(p.part_s), ce - off));
  TW* dst = !push_ch ? reinterpret_cast<TW*>(c.my_sym + p.sym.off_scat) + p.sym.scat_elem(c.slot, d) + off
                     : reinterpret_cast<TW*>(p.a.sym[d] + p.sym.off_scat) + p.sym.scat_elem(c.slot, c.rank) + off;
  const TW* from = static_cast<const TW*>(p.a.src[c.vr]) + p.part.start(d) + off;
  if (reinterpret_cast<const void*>(from) != reinterpret_cast<const void*>(dst))
    warp_copy_tile<TW>(reinterpret_cast<const uint4*>(from), reinterpret_cast<uint4*>(dst), len / E, p.a.scale, w.lane);
  __syncwarp();
  if (w.lane == 0) {
    // Part done.  Whoever brings the count to the number of parts of the chunk raises its ScatterBlock flag —
    // no publisher warp to wait for.  Ordering: every part's stores precede (fence) its increment; the last
    // incrementer has observed all of them and fences again before the flag, at system scope when the slice went
    // to a peer over NVLink (a gpu-scope fence is enough for a slice staged locally: the peers read it through
    // this GPU's L2).  That fence stalls one warp out of thousands, not a CTA.
    __threadfence();
    uint32_t* cnt = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_scat_cnt) + d * p.sym.cpb + ch;
    if (!p.self_publish) {
      atomicAdd(cnt, 1u);  // the publisher warp of the control CTA watches the count
    } else if (atomicAdd(cnt, 1u) + 1u == parts_of(ce - cb, p.part_s)) {
      st_volatile(cnt, 0u);
      if (!push_ch || p.a.fast_fences) __threadfence();
      else __threadfence_system();
      st_relaxed_sys(reinterpret_cast<uint32_t*>(p.a.sym[d] + p.sym.off_arr) + p.sym.arr_index(c.slot, c.rank, ch), c.tag);
    }
  }
}

// R: reduce one tile of my chunk and broadcast it (ReduceBlock)
template <typename TW, bool kNvls>
__device__ __noinline__ void reduce_item(const KParams& p, const Ctx& c, const WarpEnv& w, int ch, int t) {
  constexpr int E = 16 / sizeof(TW);
  const int64_t bsize = p.part.size(c.rank);
  if (ch < 0 || ch >= p.nch[c.rank]) return;
  const int64_t cb = chunk_begin(ch, p.a.max_chunk, bsize), ce = chunk_end(ch, p.a.max_chunk, bsize);
  const int64_t off = cb + static_cast<int64_t>(t) * p.part_r;
  if (off >= ce) return;
  const int len = static_cast<int>(min(static_cast<int64_t>(p.part_r), ce - off));
  const uint64_t w0 = w.wait_r != nullptr ? globaltimer_ns() : 0;
  // (an all-gather has nothing to wait for: the block is the caller's, the only contributor is this rank)
  const uint32_t mask = p.a.op == AAR_OP_ALL_GATHER ? (1u << c.rank) : wait_chunk_mask(p, c, ch, w.lane);
  if (w.wait_r != nullptr) {
    const uint64_t w1 = globaltimer_ns();
    *w.wait_r += w1 - w0;
    if (*w.first_r == 0) *w.first_r = w1;
  }
  const uint32_t all = all_ranks_mask(c.P);
  const int cnt_mask_popc = __popc(mask);
  const int cnt = cnt_mask_popc;
  const float post = (p.a.normalize && cnt > 0) ? 1.f / static_cast<float>(cnt) : 1.f;
  const int64_t sink_off = p.sym.off_sink + (p.sym.sink_elem(c.slot, c.rank) + off) * static_cast<int64_t>(sizeof(TW));
  const bool push_ch = !kNvls || ch >= p.nvls_chunks;
  // who gets the reduced tile: every peer this launch may push to — or, in a reduce-scatter, nobody but me
  const uint32_t bdest = p.a.op == AAR_OP_REDUCE_SCATTER ? (1u << c.rank) : w.dest_ok;
  uint4* mc_dst = (w.mc != nullptr && bdest == all && (!push_ch || p.a.mc_bcast || p.a.op == AAR_OP_ALL_GATHER))
                      ? reinterpret_cast<uint4*>(w.mc + sink_off) : nullptr;
  if (p.a.op == AAR_OP_ALL_GATHER) {
    const uint4* from = reinterpret_cast<const uint4*>(static_cast<const TW*>(p.a.src[c.vr]) + off);
    warp_bcast_tile(p, c, from, sink_off, len / E, bdest, mc_dst, w.lane);
  } else if (!push_ch && mask == all && w.mc != nullptr && (mc_dst != nullptr || p.a.op == AAR_OP_REDUCE_SCATTER)) {
    // in-switch reduction; the result goes to every rank's sink row through the switch again, or — reduce-scatter
    // — into my own sink row only
    const int64_t so = p.sym.off_scat + (p.sym.scat_elem(c.slot, c.rank) + off) * static_cast<int64_t>(sizeof(TW));
    uint4* dst = mc_dst != nullptr ? mc_dst : reinterpret_cast<uint4*>(c.my_sym + sink_off);
    warp_nvls_tile<TW>(reinterpret_cast<const uint4*>(w.mc + so), dst, len / E, post, mc_dst != nullptr, w.lane);
  } else if (push_ch) {  // contributions were pushed into my own scat rows
    if (c.P == 1) warp_reduce_units<TW, 4, false>(p, c, off, len / E, mask, bdest, mc_dst, post, w.lane);
    else if (c.P == 2) warp_reduce_units<TW, 2, false>(p, c, off, len / E, mask, bdest, mc_dst, post, w.lane);
    else warp_reduce_units<TW, 1, false>(p, c, off, len / E, mask, bdest, mc_dst, post, w.lane);
  } else {  // NVLS chunk that closed short of all P contributors: unicast pulls from the peers' staging slots
    warp_reduce_units<TW, 1, true>(p, c, off, len / E, mask, bdest, mc_dst, post, w.lane);
  }
  __syncwarp();
  if (w.lane == 0) {
    // Part done; the last part of the chunk raises the ReduceBlock flag (tag | contributor count) on every
    // destination, behind a system-scope fence: the reduced data went out over NVLink.
    __threadfence();
    uint32_t* cnt = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_red_cnt) + ch;
    if (!p.self_publish) {
      atomicAdd(cnt, 1u);
    } else if (atomicAdd(cnt, 1u) + 1u == parts_of(ce - cb, p.part_r)) {
      uint32_t* red_mask = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_red_mask) + ch;
      st_volatile(cnt, 0u);
      st_volatile(red_mask, 0u);
      if (p.a.fast_fences) __threadfence();
      else __threadfence_system();
      const uint64_t f = (static_cast<uint64_t>(c.tag) << 32) | static_cast<uint32_t>(cnt_mask_popc);
      if (w.mc != nullptr && bdest == all && !p.a.fast_fences) {
        multimem_st_u64(reinterpret_cast<uint64_t*>(w.mc + p.sym.off_rflag) + p.sym.arr_index(c.slot, c.rank, ch), f);
      } else {
        for (int i = 0; i < c.P; ++i) {
          const int q = (i + c.rank) % c.P;
          if ((bdest >> q) & 1u)
            st_relaxed_sys(reinterpret_cast<uint64_t*>(p.a.sym[q] + p.sym.off_rflag) + p.sym.arr_index(c.slot, c.rank, ch), f);
        }
      }
    }
  }
}

// C: owner o's reduced chunk -> output + per-element counts (ReducedDataBuffer.scala:26-53).  `repair`: second
// look, after the completion decision, at a tile whose counts were written speculatively.  Returns true when
// lane 0 issued bulk stores (to be waited for before the CTA retires).
template <typename TW>
__device__ __noinline__ bool complete_item(const KParams& p, const Ctx& c, const WarpEnv& w, int o, int ch, int t, bool repair) {
  constexpr int E = 16 / sizeof(TW);
  const int64_t bsize = p.part.size(o);
  if (ch < 0 || ch >= p.nch[o]) return false;
  if (p.a.op == AAR_OP_REDUCE_SCATTER && o != c.rank) return false;  // out holds my block only
  const int64_t cb = chunk_begin(ch, p.a.max_chunk, bsize), ce = chunk_end(ch, p.a.max_chunk, bsize);
  const int64_t off = cb + static_cast<int64_t>(t) * p.part_c;
  if (off >= ce) return false;
  const int len = static_cast<int>(min(static_cast<int64_t>(p.part_c), ce - off));
  const int64_t g = (p.a.op == AAR_OP_REDUCE_SCATTER ? 0 : p.part.start(o)) + off;
  int32_t* counts = p.a.counts[c.vr];
  if (p.spec_counts && !repair) return warp_fill_counts(counts + g, len, c.P, w.cnt_tile, c.P, w.use_tma, w.lane);
  const uint64_t w0 = w.wait_c != nullptr ? globaltimer_ns() : 0;
  const int cnt = wait_reduced(p, c, o, ch, w.full, w.lane);
  if (w.wait_c != nullptr) *w.wait_c += globaltimer_ns() - w0;
  if (repair && cnt == c.P) return false;
  const bool present = cnt >= 0;
  TW* out = static_cast<TW*>(p.a.out[c.vr]);
  const TW* sink = reinterpret_cast<const TW*>(c.my_sym + p.sym.off_sink) + p.sym.sink_elem(c.slot, o) + off;
  if (!(present && reinterpret_cast<const void*>(sink) == reinterpret_cast<const void*>(out + g)))
    warp_flush_tile<TW>(reinterpret_cast<const uint4*>(sink), reinterpret_cast<uint4*>(out + g), len / E, present, w.lane);
  if (counts == nullptr) return false;
  return warp_fill_counts(counts + g, len, present ? cnt : 0, w.cnt_tile, c.P, w.use_tma && !repair, w.lane);
}

// ---------------------------------------------------------------- direct mode (two ranks, large messages)
// Push: tile t of message chunk m -> every peer's one-shot row [me], source scale fused.
template <typename TW>
__device__ __noinline__ void direct_push_item(const KParams& p, const Ctx& c, const WarpEnv& w, int m, int t) {
  constexpr int E = 16 / sizeof(TW);
  const int64_t clen = msg_chunk_len(p, m);
  const int64_t off = static_cast<int64_t>(t) * p.part_s;
  if (off >= clen || w.muted) return;
  const int len = static_cast<int>(min(static_cast<int64_t>(p.part_s), clen - off));
  const int64_t g = static_cast<int64_t>(m) * p.a.max_chunk + off;
  const TW* from = static_cast<const TW*>(p.a.src[c.vr]) + g;
  for (int i = 1; i < c.P; ++i) {
    const int q = (i + c.rank) % c.P;
    if (((w.skip >> q) & 1u) || !((w.dest_ok >> q) & 1u)) continue;
    TW* dst = reinterpret_cast<TW*>(p.a.sym[q] + p.sym.off_os) + p.sym.os_elem(c.slot, c.rank) + g;
    warp_copy_tile<TW>(reinterpret_cast<const uint4*>(from), reinterpret_cast<uint4*>(dst), len / E, p.a.scale, w.lane);
  }
  __syncwarp();
  if (w.lane == 0) {
    __threadfence();
    uint32_t* cnt = reinterpret_cast<uint32_t*>(c.scratch + p.scr.off_scat_cnt) + m;
    if (!p.self_publish) {
      atomicAdd(cnt, 1u);  // the publisher warp raises the chunk's flag
    } else if (atomicAdd(cnt, 1u) + 1u == parts_of(clen, p.part_s)) {
      st_volatile(cnt, 0u);
      if (p.a.fast_fences) __threadfence();
      else __threadfence_system();
      for (int q = 0; q < c.P; ++q)
        if (q != c.rank && !((w.skip >> q) & 1u) && ((w.dest_ok >> q) & 1u))
          st_relaxed_sys(direct_flag(p, p.a.sym[q], c.slot, c.rank, m), c.tag);
    }
  }
}

// Sum: out[tile] = sum over the ranks whose chunk m has landed (mine is the source itself), rank order, fp32.
template <typename TW>
__device__ __noinline__ bool direct_sum_item(const KParams& p, const Ctx& c, const WarpEnv& w, int m, int t) {
  constexpr int E = 16 / sizeof(TW);
  using U = Unit<TW, E>;
  const int64_t clen = msg_chunk_len(p, m);
  const int64_t off = static_cast<int64_t>(t) * p.part_r;
  if (off >= clen) return false;
  const int len = static_cast<int>(min(static_cast<int64_t>(p.part_r), clen - off));
  const int64_t g = static_cast<int64_t>(m) * p.a.max_chunk + off;
  const int P = c.P, lane = w.lane;
  // wait for every peer's chunk (or for word that it will not come)
  const uint64_t w0 = w.wait_r != nullptr ? globaltimer_ns() : 0;
  const uint32_t all = all_ranks_mask(P);
  uint32_t have = w.muted ? 0u : (1u << c.rank);
  {
    const uint32_t* f = lane < P ? direct_flag(p, c.my_sym, c.slot, lane, m) : nullptr;
    int spins = 0;
    for (;;) {
      have |= __ballot_sync(0xffffffffu, lane < P && lane != c.rank && ld_volatile(f) == c.tag);
      if ((have | (1u << c.rank)) == all) break;
      if ((++spins & (kSlowPoll - 1)) == 0) {
        const uint32_t gone = peers_gone(p, c) & ~have;
        if (gone != 0 && ((have | gone | (1u << c.rank)) == all)) {
          have |= __ballot_sync(0xffffffffu, lane < P && lane != c.rank && ld_volatile(f) == c.tag);
          break;
        }
        if (__shfl_sync(0xffffffffu, (lane == 0 && give_up_now(c, c.deadline)) ? 1 : 0, 0) != 0) break;
      }
      __nanosleep(64);
    }
    if (lane < P && lane != c.rank && ((have >> lane) & 1u)) (void)ld_acquire_sys(f);  // acquire: the sender's data
    __syncwarp();
  }
  if (w.wait_r != nullptr) {
    const uint64_t w1 = globaltimer_ns();
    *w.wait_r += w1 - w0;
    if (*w.first_r == 0) *w.first_r = w1;
  }
  const int cnt = __popc(have);
  const float post = (p.a.normalize && cnt > 0) ? 1.f / static_cast<float>(cnt) : 1.f;
  const uint4* mine = reinterpret_cast<const uint4*>(static_cast<const TW*>(p.a.src[c.vr]) + g);
  uint4* out = reinterpret_cast<uint4*>(static_cast<TW*>(p.a.out[c.vr]) + g);
  const int units = len / E;
  constexpr int UB = 4;  // units per lane and sweep: with two ranks, eight 16-byte loads in flight
  auto sweep = [&](int u0, int nb) {
    float acc[UB][E];
#pragma unroll
    for (int b = 0; b < UB; ++b)
#pragma unroll
      for (int e = 0; e < E; ++e) acc[b][e] = 0.f;
    for (int s = 0; s < P; ++s) {
      if (!((have >> s) & 1u)) continue;
      const uint4* row = s == c.rank ? mine
                                     : reinterpret_cast<const uint4*>(reinterpret_cast<const TW*>(c.my_sym + p.sym.off_os) + p.sym.os_elem(c.slot, s) + g);
      uint4 r[UB];
#pragma unroll
      for (int b = 0; b < UB; ++b) r[b] = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
      for (int b = 0; b < UB; ++b)
        if (b < nb) r[b] = s == c.rank ? ld_stream(row + u0 + 32 * b) : ld_fresh(row + u0 + 32 * b);
#pragma unroll
      for (int b = 0; b < UB; ++b) {
        float v[E];
        U::unpack(r[b], v);
        const float sc = s == c.rank ? p.a.scale : 1.f;  // (the peers' copies were scaled on the way out)
#pragma unroll
        for (int e = 0; e < E; ++e) acc[b][e] += v[e] * sc;
      }
    }
#pragma unroll
    for (int b = 0; b < UB; ++b)
      if (b < nb) {
#pragma unroll
        for (int e = 0; e < E; ++e) acc[b][e] *= post;
        out[u0 + 32 * b] = U::pack(acc[b]);
      }
  };
  int u0 = lane;
  for (; u0 + 32 * (UB - 1) < units; u0 += 32 * UB) sweep(u0, UB);
  for (; u0 < units; u0 += 32) sweep(u0, 1);
  int32_t* counts = p.a.counts[c.vr];
  if (counts == nullptr) return false;
  return warp_fill_counts(counts + g, len, cnt, w.cnt_tile, P, w.use_tma, lane);
}

// ---------------------------------------------------------------- the kernel
template <typename TW, bool kNvls>
__global__ void __launch_bounds__(kThreads, 2) aar_stream_kernel(const __grid_constant__ KParams p) {
  const Ctx c = make_ctx(p);
  if (blockIdx.x == 0) {
    control_cta(p, c);
    return;
  }
  __shared__ __align__(128) int32_t cnt_tile[kCntTile];
  __shared__ uint32_t sh_dest_ok;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int P = c.P, cpb = p.sym.cpb;
  uint64_t* trace = p.a.trace != nullptr
                        ? p.a.trace + (static_cast<int64_t>(c.vr) * (p.a.ctas + 1) + blockIdx.x) * 16
                        : nullptr;
  const bool tracer0 = trace != nullptr && threadIdx.x == 0;
  uint64_t wait_r = 0, wait_c = 0, first_r = 0;
  if (tracer0) trace[0] = globaltimer_ns();

  for (int i = threadIdx.x; i < kCntTile / 4; i += kThreads) reinterpret_cast<int4*>(cnt_tile)[i] = make_int4(P, P, P, P);
  fence_proxy_async_smem();
  if (warp == 0) {
    if (lane == 0 && p.a.delay_ns[c.rank] != 0) {  // injected straggler
      const uint64_t until = globaltimer_ns() + p.a.delay_ns[c.rank];
      while (globaltimer_ns() < until) __nanosleep(256);
    }
    __syncwarp();
    wait_gate_warp(p, c);
    const uint32_t m = decide_dest_mask_warp(p, c);
    if (lane == 0) sh_dest_ok = m;
  }
  __syncthreads();
  if (tracer0) trace[1] = globaltimer_ns();
  WarpEnv w;
  w.dest_ok = sh_dest_ok;
  w.skip = p.a.skip_dest_mask[c.rank];
  w.lane = lane;
  w.muted = (p.a.mute_mask >> c.rank) & 1u;
  w.full = p.a.req_reduce >= P && p.a.req_complete >= p.total_chunks;
  w.use_tma = p.a.use_tma != 0;
  w.mc = p.a.emu ? nullptr : p.a.mc_base;
  w.cnt_tile = cnt_tile;

  // Role of this warp, and its position among the warps of that role across the grid (consecutive tiles land
  // on different CTAs, so the grid walks through a chunk together and its flag goes out early).
  const int cta = blockIdx.x - 1;
  const int role = warp < p.warps_s ? 0 : (warp < p.warps_s + p.warps_r ? 1 : 2);
  const int role_warps = role == 0 ? p.warps_s : (role == 1 ? p.warps_r : kWarps - p.warps_s - p.warps_r);
  const int rw = (warp - (role == 0 ? 0 : (role == 1 ? p.warps_s : p.warps_s + p.warps_r))) * p.a.ctas + cta;
  const int stride = role_warps * p.a.ctas;
  // tracing: lane 0 of the first warp of each role keeps the time line of its role for this CTA
  const bool tracer = trace != nullptr && lane == 0 && rw == cta;
  w.wait_r = tracer ? &wait_r : nullptr;
  w.wait_c = tracer ? &wait_c : nullptr;
  w.first_r = tracer ? &first_r : nullptr;
  bool bulk = false;
  if (p.a.algo == AAR_ALGO_DIRECT) {
    if (role == 0) {
      const int total = p.msg_chunks * p.tpc_s;
      for (int it = rw; it < total; it += stride) direct_push_item<TW>(p, c, w, it / p.tpc_s, it % p.tpc_s);
    } else if (role == 1) {
      const int total = p.msg_chunks * p.tpc_r;
      for (int it = rw; it < total; it += stride) bulk |= direct_sum_item<TW>(p, c, w, it / p.tpc_r, it % p.tpc_r);
    }
  } else if (role == 0) {
    if ((!p.zc_src || (kNvls && p.nvls_chunks < cpb)) && p.a.op != AAR_OP_ALL_GATHER) {
      const int per_chunk = P * p.tpc_s, total = per_chunk * cpb;
      for (int it = rw; it < total; it += stride) {
        const int ch = it / per_chunk, rem = it - ch * per_chunk;
        scatter_item<TW, kNvls>(p, c, w, (rem / p.tpc_s + c.rank) % P, ch, rem % p.tpc_s);
      }
    }
  } else if (role == 1) {
    const int total = p.tpc_r * p.nch[c.rank];
    for (int it = rw; it < total; it += stride) reduce_item<TW, kNvls>(p, c, w, it / p.tpc_r, it % p.tpc_r);
  } else {
    const int per_chunk = P * p.tpc_c, total = per_chunk * cpb;
    for (int it = rw; it < total; it += stride) {
      const int ch = it / per_chunk, rem = it - ch * per_chunk;
      bulk |= complete_item<TW>(p, c, w, (rem / p.tpc_c + c.rank) % P, ch, rem % p.tpc_c, false);
    }
  }

  if (p.spec_counts && role == 2) {
    // Counts went out as P before anything was known.  Usually that is what they are; if the poller saw a
    // chunk missing or short of contributors, walk the completion items again and repair those.
    const uint32_t* cw = reinterpret_cast<const uint32_t*>(c.scratch + p.scr.off_complete);
    uint32_t ok = 0;
    if (lane == 0) {
      const uint64_t stop = c.deadline + 2 * p.a.timeout_ns + 1000000000ull;
      const uint64_t w0 = tracer ? globaltimer_ns() : 0;
      int spins = 0;
      while (!(ld_volatile(cw) & kDecided)) {
        if ((++spins & (kSlowPoll - 1)) == 0 && globaltimer_ns() > stop) break;
        __nanosleep(64);
      }
      ok = (ld_volatile(cw) & kDecided) ? ld_volatile(cw + 1) : 0u;
      if (tracer) wait_c += globaltimer_ns() - w0;
    }
    ok = __shfl_sync(0xffffffffu, ok, 0);
    if (ok == 0) {
      if (bulk && lane == 0) {  // the speculative bulk stores must have landed before ordinary stores correct them
        tma_wait_all<0>();
        asm volatile("fence.proxy.async;" ::: "memory");
      }
      __syncwarp();
      const int per_chunk = P * p.tpc_c, total = per_chunk * cpb;
      for (int it = rw; it < total; it += stride) {
        const int ch = it / per_chunk, rem = it - ch * per_chunk;
        complete_item<TW>(p, c, w, (rem / p.tpc_c + c.rank) % P, ch, rem % p.tpc_c, true);
      }
    }
  }
  if (bulk && lane == 0) tma_wait_all<0>();  // this lane's count stores have landed
  if (tracer) {  // [2] scatter role done, [3] reduce role done, [4] complete role done, [5]/[6] ns waited in R / C
    trace[2 + role] = globaltimer_ns();
    if (role == 1) {
      trace[5] = wait_r;
      trace[9] = first_r;
    }
    if (role == 2) trace[6] = wait_c;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(fin_words(p, c) + 3, 1u);  // tells the publisher that this CTA will add nothing more to its counters
    if (trace != nullptr) {
      trace[10] = globaltimer_ns();  // all warps of the CTA done
      trace[8] = 4;                  // role 4: a CTA of the stream kernel
    }
    finish_cta(p, c);
  }
}

}  // namespace aar
