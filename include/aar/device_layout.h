// Layout of one rank's symmetric region and of the kernel's local scratch.
//
// Every rank allocates the same region and maps every peer's copy (NVLink P2P, and an NVLS
// multicast alias where the driver exposes one).  The region is the device-side form of the
// reference's two ring buffers plus its mailbox:
//
//   scat  [L][P][block_pad]   contributions of every peer to MY block   (ScatteredDataBuffer)
//   sink  [L][P][block_pad]   reduced blocks pushed by their owners      (ReducedDataBuffer)
//   arr   [L][P][cpb] u32     "ScatterBlock arrived" round tags, written by the sender
//   rflag [L][P][cpb] u64     "ReduceBlock arrived": round tag << 32 | contributor count
//   done  [P] u32             CompleteAllreduce: rounds each peer has finished (tag = round+1)
//   notice[P] u32             skip notices: notice[s] = highest tag for which peer s left ME out of its
//                             scatter and broadcast (back-pressure); lets a wait give up without the clock
//   os    [L][P][os_cap]      one-shot path (small messages): every peer's WHOLE message
//   os_arr[L][P][64] u32      one-shot arrival tags, one per (sender, CTA slice)
//
// L = maxLag + 1 ring slots, slot = round % L (buffer/AllReduceBuffer.scala:34-42); P = peers;
// cpb = chunks per block.  A message is a store into the destination's region followed by a
// release-store of the flag; a round tag that does not match is a stale or future message.
#pragma once
#include <stdint.h>

#include "aar/protocol_math.h"

namespace aar {

constexpr int kMaxRanks = 16;       // one NVSwitch domain
constexpr int kAlignElems = 128;    // block padding, in elements (keeps every block 512 B aligned)
constexpr int kMaxOneShotCtas = 64; // one-shot: every CTA owns a slice and its own arrival flags

struct SymLayout {
  int P, L;
  int wire_bytes;          // 4 = fp32 slots, 2 = bf16 slots
  int64_t n, max_chunk;
  int64_t block;           // block step: ceil(n / P) (the reference's), or an aligned step >= that
  int64_t block_pad;       // block rounded up to kAlignElems
  int cpb;                 // chunks per (full) block
  int64_t os_cap;          // one-shot capacity in elements (0 = none)
  int64_t off_arr, off_rflag, off_done, off_notice, off_os_arr, off_scat, off_sink, off_os, total_bytes;

  AAR_HD static int64_t up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

  AAR_HD static SymLayout make(int64_t n, int64_t max_chunk, int P, int L, int wire_bytes, int64_t os_cap,
                               int64_t step = 0) {
    SymLayout s;
    s.P = P;
    s.L = L;
    s.wire_bytes = wire_bytes;
    s.n = n;
    s.max_chunk = max_chunk;
    s.block = step > 0 ? step : ceil_div(n, P);
    s.block_pad = up(s.block > 0 ? s.block : 1, kAlignElems);
    s.cpb = num_chunks(s.block, max_chunk) > 0 ? num_chunks(s.block, max_chunk) : 1;
    s.os_cap = up(os_cap, kAlignElems);
    int64_t off = 0;
    s.off_arr = off;
    off = up(off + (int64_t)L * P * s.cpb * 4, 256);
    s.off_rflag = off;
    off = up(off + (int64_t)L * P * s.cpb * 8, 256);
    s.off_done = off;
    s.off_notice = off + kMaxRanks * 4;  // same 256-byte line group: done[16] | notice[16]
    off = up(off + 2 * kMaxRanks * 4, 256);
    s.off_os_arr = off;
    off = up(off + (int64_t)L * P * kMaxOneShotCtas * 4, 4096);
    s.off_scat = off;
    off = up(off + (int64_t)L * P * s.block_pad * wire_bytes, 4096);
    s.off_sink = off;
    off = up(off + (int64_t)L * P * s.block_pad * wire_bytes, 4096);
    s.off_os = off;
    off = up(off + (int64_t)L * P * s.os_cap * wire_bytes, 4096);
    s.total_bytes = off;
    return s;
  }
  AAR_HD int64_t flags_bytes() const { return off_scat; }  // everything that must start zeroed

  AAR_HD int64_t arr_index(int slot, int src, int chunk) const { return ((int64_t)slot * P + src) * cpb + chunk; }
  AAR_HD int64_t scat_elem(int slot, int src) const { return ((int64_t)slot * P + src) * block_pad; }
  AAR_HD int64_t sink_elem(int slot, int owner) const { return ((int64_t)slot * P + owner) * block_pad; }
  AAR_HD int64_t os_elem(int slot, int src) const { return ((int64_t)slot * P + src) * os_cap; }
};

// Local (non-symmetric) scratch of one rank: self-resetting tile counters and the per-chunk /
// per-round decisions that all CTAs of a launch must agree on.  Zero at allocation; every
// launch leaves it zero again.
struct ScratchLayout {
  int P, cpb;
  int64_t off_scat_cnt;   // u32 [P][cpb]   tiles of (dest, chunk) already pushed
  int64_t off_red_cnt;    // u32 [cpb]      tiles of my chunk already reduced + broadcast
  int64_t off_red_mask;   // u32 [cpb]      decided contributor mask | kDecided
  int64_t off_fin_cnt;    // u32 [0] CTAs that finished; [1] decided set of scatter destinations;
                          //     [2] force tag, written by the HOST (never reset by the kernel);
                          //     [3] work CTAs that are past phase R;
                          //     [4] bit mask of peers known dead, written by the HOST (never reset by the kernel)
  int64_t off_complete;   // u32 [0] completion decision, [1] 1 = every chunk arrived with all P contributors,
                          //     [4 + words] snapshot bitmap of (owner, chunk)
  int64_t total_bytes;

  AAR_HD static ScratchLayout make(int P, int cpb) {
    ScratchLayout s;
    s.P = P;
    s.cpb = cpb;
    int64_t off = 0;
    s.off_scat_cnt = off;
    off += (int64_t)P * cpb * 4;
    s.off_red_cnt = off;
    off += (int64_t)cpb * 4;
    s.off_red_mask = off;
    off += (int64_t)cpb * 4;
    s.off_fin_cnt = off;
    off += 32;
    s.off_complete = off;
    off += 16 + ((int64_t)P * cpb + 31) / 32 * 4;
    s.total_bytes = SymLayout::up(off, 256);
    return s;
  }
};

constexpr uint32_t kDecided = 0x80000000u;

}  // namespace aar
