// Lag-ring buffers of the host protocol engine.
//
// Behavioural counterparts of the reference's buffer package
// (buffer/AllReduceBuffer.scala:3-47, ScatteredDataBuffer.scala:3-41,
// ReducedDataBuffer.scala:5-73) on flat storage: one contiguous float array
// [slots][peers][width] plus per-slot arrival counters, addressed through a rotating
// head instead of re-allocating a row per round.
#pragma once
#include <cstddef>
#include <cstdint>
#include <utility>
#include <vector>

namespace aar {

// Ring of `slots` rounds in flight; every slot holds one `width`-float row per peer.
class LagRing {
 public:
  LagRing(int64_t width, int peers, int slots, int64_t max_chunk);
  virtual ~LagRing() = default;

  // Copies `len` floats to (row, src) at element offset chunk*max_chunk and counts one
  // arrival for that chunk id.  A duplicate overwrites and still counts
  // (AllReduceBuffer.scala:25-32).  Throws std::out_of_range when the data would run
  // past the row, like the JVM's ArrayIndexOutOfBoundsException.
  void store(const float* data, size_t len, int row, int src, int chunk);

  // Rotates by one round; the slot that becomes the newest row is zeroed
  // (AllReduceBuffer.scala:38-42).
  virtual void advance();

  int64_t width() const { return width_; }
  int peers() const { return peers_; }
  int slots() const { return slots_; }
  int64_t max_chunk() const { return max_chunk_; }
  int chunks() const { return chunks_; }
  int arrivals(int row, int chunk) const { return arrivals_[slot_of(row) * chunks_ + chunk]; }
  int arrivals_total(int row) const;
  const float* peer_row(int row, int src) const { return &cells_[cell_index(slot_of(row), src)]; }

 protected:
  int slot_of(int row) const { return slots_ > 0 ? (row + head_) % slots_ : 0; }
  size_t cell_index(int slot, int src) const {
    return (static_cast<size_t>(slot) * peers_ + src) * static_cast<size_t>(width_);
  }

  int64_t width_;
  int peers_, slots_;
  int64_t max_chunk_;
  int chunks_;
  int head_ = 0;
  std::vector<float> cells_;
  std::vector<int> arrivals_;
};

// Contributions of every peer to MY block; reduces a chunk once exactly
// int(thReduce*P) contributions are in (ScatteredDataBuffer.scala:9-13).
class ScatterBuffer : public LagRing {
 public:
  ScatterBuffer(int64_t block_size, int peers, int slots, float th_reduce, int64_t max_chunk);
  int required() const { return required_; }
  bool at_threshold(int row, int chunk) const { return arrivals(row, chunk) == required_; }
  // Sum over ALL peer rows of the chunk's element range (absent peers are zeros) and
  // the number of arrivals so far (ScatteredDataBuffer.scala:20-32).
  std::pair<std::vector<float>, int> reduce(int row, int chunk) const;

 private:
  int required_;
};

// Reduced chunks from every block owner; completes a round once exactly
// int(thComplete*totalChunks) of them are in (ReducedDataBuffer.scala:13-17,60-66).
class ReduceBuffer : public LagRing {
 public:
  ReduceBuffer(int64_t max_block, int64_t min_block, int64_t total_size, int peers, int slots,
               float th_complete, int64_t max_chunk);
  // General form used by the worker engine when trailing blocks may be empty.
  ReduceBuffer(int64_t max_block, int total_chunks, int64_t total_size, int peers, int slots,
               float th_complete, int64_t max_chunk, int /*tag*/);

  void store(const float* data, size_t len, int row, int src, int chunk, int count);
  void advance() override;
  int required() const { return required_; }
  bool at_threshold(int row) const { return arrivals_total(row) == required_; }
  // Concatenated blocks and per-element contributor counts; chunks never received are
  // zeros with count 0 (ReducedDataBuffer.scala:26-53).
  std::pair<std::vector<float>, std::vector<int>> assemble(int row) const;

 private:
  int64_t total_size_;
  int required_;
  std::vector<int> contributors_;  // [slots][peers*chunks]
};

}  // namespace aar
