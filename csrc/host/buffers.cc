#include "buffers.h"

#include <algorithm>
#include <stdexcept>
#include <string>

#include "aar/protocol_math.h"

namespace aar {

LagRing::LagRing(int64_t width, int peers, int slots, int64_t max_chunk)
    : width_(width),
      peers_(peers),
      slots_(slots),
      max_chunk_(max_chunk > 0 ? max_chunk : 1),
      chunks_(num_chunks(width, max_chunk > 0 ? max_chunk : 1)),
      cells_(static_cast<size_t>(slots) * peers * static_cast<size_t>(width), 0.f),
      arrivals_(static_cast<size_t>(slots) * chunks_, 0) {}

void LagRing::store(const float* data, size_t len, int row, int src, int chunk) {
  if (src < 0 || src >= peers_ || row < 0 || row >= slots_)
    throw std::out_of_range("LagRing::store: row/src outside the ring");
  const int64_t at = static_cast<int64_t>(chunk) * max_chunk_;
  if (chunk < 0 || at + static_cast<int64_t>(len) > width_)
    throw std::out_of_range("LagRing::store: chunk " + std::to_string(chunk) + " of " +
                            std::to_string(len) + " elements overruns a row of " +
                            std::to_string(width_));
  const int slot = slot_of(row);
  std::copy(data, data + len, cells_.begin() + cell_index(slot, src) + at);
  if (chunk >= chunks_) throw std::out_of_range("LagRing::store: chunk id beyond chunk count");
  arrivals_[slot * chunks_ + chunk] += 1;
}

void LagRing::advance() {
  if (slots_ == 0) return;
  const int recycled = head_;  // oldest slot becomes the newest row
  head_ = (head_ + 1) % slots_;
  auto first = cells_.begin() + cell_index(recycled, 0);
  std::fill(first, first + static_cast<size_t>(peers_) * width_, 0.f);
  std::fill(arrivals_.begin() + recycled * chunks_, arrivals_.begin() + (recycled + 1) * chunks_, 0);
}

int LagRing::arrivals_total(int row) const {
  const int slot = slot_of(row);
  int t = 0;
  for (int c = 0; c < chunks_; ++c) t += arrivals_[slot * chunks_ + c];
  return t;
}

ScatterBuffer::ScatterBuffer(int64_t block_size, int peers, int slots, float th_reduce,
                             int64_t max_chunk)
    : LagRing(block_size, peers, slots, max_chunk), required_(threshold_count(th_reduce, peers)) {}

std::pair<std::vector<float>, int> ScatterBuffer::reduce(int row, int chunk) const {
  const int64_t lo = static_cast<int64_t>(chunk) * max_chunk_;
  const int64_t hi = std::min<int64_t>(width_, static_cast<int64_t>(chunk + 1) * max_chunk_);
  std::vector<float> sum(static_cast<size_t>(std::max<int64_t>(hi - lo, 0)), 0.f);
  for (int p = 0; p < peers_; ++p) {  // fixed peer order 0..P-1, fp32 accumulate
    const float* src = peer_row(row, p) + lo;
    for (size_t j = 0; j < sum.size(); ++j) sum[j] += src[j];
  }
  return {std::move(sum), arrivals(row, chunk)};
}

ReduceBuffer::ReduceBuffer(int64_t max_block, int64_t min_block, int64_t total_size, int peers,
                           int slots, float th_complete, int64_t max_chunk)
    : ReduceBuffer(max_block,
                   total_reduced_chunks_ref(max_block, min_block, peers, max_chunk > 0 ? max_chunk : 1),
                   total_size, peers, slots, th_complete, max_chunk, 0) {}

ReduceBuffer::ReduceBuffer(int64_t max_block, int total_chunks, int64_t total_size, int peers,
                           int slots, float th_complete, int64_t max_chunk, int)
    : LagRing(max_block, peers, slots, max_chunk),
      total_size_(total_size),
      required_(threshold_count(th_complete, total_chunks)),
      contributors_(static_cast<size_t>(slots) * peers * chunks_, 0) {}

void ReduceBuffer::store(const float* data, size_t len, int row, int src, int chunk, int count) {
  LagRing::store(data, len, row, src, chunk);
  contributors_[(static_cast<size_t>(slot_of(row)) * peers_ + src) * chunks_ + chunk] = count;
}

void ReduceBuffer::advance() {
  if (slots_ == 0) return;
  const int recycled = head_;
  LagRing::advance();
  const size_t per_slot = static_cast<size_t>(peers_) * chunks_;
  std::fill(contributors_.begin() + recycled * per_slot,
            contributors_.begin() + (recycled + 1) * per_slot, 0);
}

std::pair<std::vector<float>, std::vector<int>> ReduceBuffer::assemble(int row) const {
  std::vector<float> data(static_cast<size_t>(total_size_), 0.f);
  std::vector<int> counts(static_cast<size_t>(total_size_), 0);
  const int slot = slot_of(row);
  int64_t filled = 0, counted = 0;
  for (int p = 0; p < peers_; ++p) {
    const int64_t take = std::min<int64_t>(total_size_ - filled, width_);
    if (take > 0) {
      const float* src = peer_row(row, p);
      std::copy(src, src + take, data.begin() + filled);
      filled += take;
    }
    for (int c = 0; c < chunks_; ++c) {
      int64_t span = std::min<int64_t>(max_chunk_, width_ - max_chunk_ * c);
      span = std::min<int64_t>(span, total_size_ - counted);
      if (span <= 0) continue;
      const int who = contributors_[(static_cast<size_t>(slot) * peers_ + p) * chunks_ + c];
      std::fill(counts.begin() + counted, counts.begin() + counted + span, who);
      counted += span;
    }
  }
  return {std::move(data), std::move(counts)};
}

}  // namespace aar
