// Native blend-index builder: the C++ equivalent of the Rust `blended_dataset_loop.sample`
// wheel the reference calls from core/data/blended_dataset.py:316-329.
//
// Rule: repeatedly emit one sample from the dataset whose sampled/target ratio is lowest (ties -> lowest dataset
// index), until EVERY dataset has delivered its target count.  The reference's pure-Python variant of the loop
// (tfm/data/legacy_blended_dataset.py:214-229) applies the same rule but stops when the FIRST dataset is exhausted;
// the wheel's contract is "run to exhaustion" (the blended length equals the sum of the targets,
// tests/transformer/test_blended_dataset.py:58), which is what this routine implements — the legacy output is a
// prefix of it (tests/test_core_units_cpu.py::test_blend_index_extends_the_legacy_python_loop).  Instead of an O(N·D) argmin scan this keeps a binary heap keyed on the exact rational
// sampled/target (compared by cross multiplication in 128-bit integers, so there is no floating-point tie noise):
// O(N log D).
#include <stdint.h>

#include <algorithm>
#include <queue>
#include <vector>

namespace {
struct Entry {
  int64_t sampled;
  int64_t target;
  int32_t index;
};
struct Greater {
  bool operator()(const Entry& a, const Entry& b) const {
    // a > b  <=>  a.sampled/a.target > b.sampled/b.target, ties broken by dataset index
    __int128 l = (__int128)a.sampled * b.target;
    __int128 r = (__int128)b.sampled * a.target;
    if (l != r) return l > r;
    return a.index > b.index;
  }
};
}  // namespace

// counts[n_datasets] -> out[total, 2] int64 rows (dataset_index, index_in_dataset); returns rows written or -1
extern "C" long long sb_blend_index(const long long* counts, int n_datasets, long long* out, long long capacity) {
  std::priority_queue<Entry, std::vector<Entry>, Greater> heap;
  long long total = 0;
  for (int i = 0; i < n_datasets; ++i) {
    if (counts[i] <= 0) continue;
    heap.push(Entry{0, counts[i], i});
    total += counts[i];
  }
  if (total > capacity) return -1;
  long long w = 0;
  while (!heap.empty()) {
    Entry e = heap.top();
    heap.pop();
    out[2 * w] = e.index;
    out[2 * w + 1] = e.sampled;
    ++w;
    if (++e.sampled < e.target) heap.push(e);
  }
  return w;
}

// Packed-window gather: copy `n_pieces` (doc, start, end) slices of a flat token array into one
// contiguous output window — the inner loop of TextDataset.__getitem__ (tfm/data/text_dataset.py:371-385).
extern "C" long long sb_gather_window_i32(const int32_t* tokens, const int64_t* doc_offsets, const int64_t* pieces,
                                          int n_pieces, int32_t* out, long long capacity) {
  long long w = 0;
  for (int i = 0; i < n_pieces; ++i) {
    const int64_t doc = pieces[3 * i], s = pieces[3 * i + 1], e = pieces[3 * i + 2];
    const int32_t* src = tokens + doc_offsets[doc] + s;
    const long long n = e - s;
    if (w + n > capacity) return -1;
    std::copy(src, src + n, out + w);
    w += n;
  }
  return w;
}
