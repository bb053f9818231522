// Shared host/device arithmetic of the threshold/lag allreduce protocol.
//
// One header, included by the C++ host engine (csrc/host) and by the sm_100a
// kernels (csrc/cuda), so the partitioner, the chunker, the three thresholds and
// the master gate cannot drift apart between the CPU oracle and the GPU path.
//
// Reference behaviour being reproduced (paths relative to /root/reference):
//   block partition   src/main/scala/sample/cluster/allreduce/AllreduceWorker.scala:240-250
//   chunking          AllreduceWorker.scala:218-224, buffer/AllReduceBuffer.scala:44-46
//   thReduce          buffer/ScatteredDataBuffer.scala:9
//   thComplete        buffer/ReducedDataBuffer.scala:13-17
//   thAllreduce gate  AllreduceMaster.scala:56-62
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define AAR_HD __host__ __device__ __forceinline__
#else
#define AAR_HD inline
#endif

namespace aar {

AAR_HD int64_t ceil_div(int64_t a, int64_t b) { return b > 0 ? (a + b - 1) / b : 0; }

// int(float32(th) * float32(n)): the product is rounded to fp32 BEFORE truncation.
// 0.7f*10 -> 7 here, while widening 0.7f to double first would give 6.
AAR_HD int threshold_count(float th, int n) {
  float prod = th * static_cast<float>(n);
  return static_cast<int>(prod);
}

// Master gate: Int >= Int*Float is evaluated in fp32 (AllreduceMaster.scala:58).
AAR_HD bool master_gate(int num_complete, int total_workers, float th_allreduce) {
  return static_cast<float>(num_complete) >= static_cast<float>(total_workers) * th_allreduce;
}

// Block partition of `n` elements over `P` owners: step = ceil(n/P), owner i holds
// [i*step, (i+1)*step) clipped to n.  Whenever the reference's partitioner is
// defined (ceil(n/step) == P) this is the same split; where the reference indexes
// out of range (e.g. n=9, P=4) the trailing owners simply get empty blocks.
struct Partition {
  int64_t n;
  int P;
  int64_t step;

  AAR_HD Partition() : n(0), P(0), step(0) {}
  AAR_HD Partition(int64_t n_, int P_) : n(n_), P(P_), step(P_ > 0 ? ceil_div(n_, P_) : 0) {}
  // Explicit step >= ceil(n/P): the device path may round blocks up to an alignment unit
  // (trailing owners then hold short or empty blocks).
  AAR_HD Partition(int64_t n_, int P_, int64_t step_) : n(n_), P(P_), step(step_) {}

  AAR_HD int64_t start(int i) const {
    int64_t s = static_cast<int64_t>(i) * step;
    return s < n ? s : n;
  }
  AAR_HD int64_t end(int i) const { return i >= P - 1 ? n : start(i + 1); }
  AAR_HD int64_t size(int i) const { return end(i) - start(i); }
  AAR_HD int64_t max_block() const { return P > 0 ? size(0) : 0; }
  AAR_HD int64_t min_block() const { return P > 0 ? size(P - 1) : 0; }
  // True when the reference's Array.range(0, n, step) has exactly P entries.
  AAR_HD bool reference_defined() const { return P > 0 && step > 0 && ceil_div(n, step) == P; }
};

AAR_HD int num_chunks(int64_t block_size, int64_t max_chunk) {
  return static_cast<int>(ceil_div(block_size, max_chunk));
}
// Half-open element range of chunk c inside a block of size b.
AAR_HD int64_t chunk_begin(int c, int64_t max_chunk, int64_t b) {
  int64_t s = static_cast<int64_t>(c) * max_chunk;
  return s < b ? s : b;
}
AAR_HD int64_t chunk_end(int c, int64_t max_chunk, int64_t b) {
  int64_t e = static_cast<int64_t>(c + 1) * max_chunk;
  return e < b ? e : b;
}

// Number of reduced chunks a worker can receive in one round: every owner's block,
// chunked.  Equals numChunks(maxBlock)*(P-1) + numChunks(minBlock) whenever the
// reference partition is defined (ReducedDataBuffer.scala:14-15).
AAR_HD int total_reduced_chunks(const Partition& part, int64_t max_chunk) {
  int t = 0;
  for (int i = 0; i < part.P; ++i) t += num_chunks(part.size(i), max_chunk);
  return t;
}

// Reference formula, kept for the buffer-level API where max/min block are given
// directly (ReducedDataBufferSpec builds buffers that way).
AAR_HD int total_reduced_chunks_ref(int64_t max_block, int64_t min_block, int P, int64_t max_chunk) {
  return num_chunks(max_block, max_chunk) * (P - 1) + num_chunks(min_block, max_chunk);
}

// Ring slot of a round: the reference rotates `temporalOffset` once per `round += 1`
// (AllReduceBuffer.scala:34-42, AllreduceWorker.scala:279-283), so row = msgRound-round
// lands on physical slot msgRound % L.
AAR_HD int ring_slot(int64_t round, int L) { return static_cast<int>(round % L); }

}  // namespace aar
