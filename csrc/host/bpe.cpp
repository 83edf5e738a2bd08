// Byte-level BPE tokenizer core (train + encode), C ABI — the native counterpart of the HF
// `tokenizers` (Rust) calls in the reference: Tokenizer(models.BPE()) + BpeTrainer + encode_batch
// (LLM_Distributed_Trainning/PyTorch/transformer_basics/GPTLike_wikitext2_fixed_pe.py:52-122,
//  …/GPTLike_wikitext2.py:49-62). Rust is not available in this build environment, and tokenisation
// of a corpus is a host-side hot loop, so it is C++ with OpenMP over documents.
//
// Model: ids 0..255 are the raw bytes; merge i creates id 256+i from an ordered pair of ids.
// Pre-tokenisation splits on whitespace boundaries (a leading space stays attached to the word, as
// ByteLevel pre-tokenisers do) so merges never cross words. Training counts pairs over the unique
// words weighted by frequency and repeatedly merges the most frequent pair (ties: smallest pair).
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#define B200_HOST_API extern "C" __attribute__((visibility("default")))

namespace {

struct PairHash {
  size_t operator()(const std::pair<int32_t, int32_t>& p) const {
    return (static_cast<size_t>(static_cast<uint32_t>(p.first)) << 32) ^ static_cast<uint32_t>(p.second);
  }
};

static void split_words(const uint8_t* text, int64_t n, std::vector<std::pair<int64_t, int64_t>>& out) {
  // word = optional run of spaces/newlines attached to the following non-space run
  int64_t i = 0;
  while (i < n) {
    int64_t s = i;
    while (i < n && (text[i] == ' ' || text[i] == '\n' || text[i] == '\t' || text[i] == '\r')) ++i;
    if (i - s > 1) {  // keep at most one leading whitespace byte with the word; the rest stand alone
      out.emplace_back(s, i - 1);
      s = i - 1;
    }
    while (i < n && !(text[i] == ' ' || text[i] == '\n' || text[i] == '\t' || text[i] == '\r')) ++i;
    if (i > s) out.emplace_back(s, i);
  }
}

static void apply_merges(std::vector<int32_t>& w, const std::unordered_map<std::pair<int32_t, int32_t>, int32_t, PairHash>& rank) {
  // greedy lowest-rank-first merging, as in the classic BPE encoder
  while (w.size() >= 2) {
    int32_t best = INT32_MAX;
    size_t pos = 0;
    for (size_t i = 0; i + 1 < w.size(); ++i) {
      auto it = rank.find({w[i], w[i + 1]});
      if (it != rank.end() && it->second < best) { best = it->second; pos = i; }
    }
    if (best == INT32_MAX) break;
    w[pos] = 256 + best;
    w.erase(w.begin() + pos + 1);
  }
}

}  // namespace

// Train: merges_out receives up to max_merges pairs (2 ints each); returns the number of merges learnt.
B200_HOST_API int64_t b200_bpe_train(const uint8_t* text, int64_t n, int32_t max_merges, int32_t min_frequency,
                                     int32_t* merges_out) {
  std::vector<std::pair<int64_t, int64_t>> spans;
  split_words(text, n, spans);
  std::unordered_map<std::string, int64_t> freq;
  for (auto& sp : spans) freq[std::string(reinterpret_cast<const char*>(text + sp.first), sp.second - sp.first)]++;
  std::vector<std::vector<int32_t>> words;
  std::vector<int64_t> counts;
  words.reserve(freq.size());
  for (auto& kv : freq) {
    std::vector<int32_t> w(kv.first.size());
    for (size_t i = 0; i < kv.first.size(); ++i) w[i] = static_cast<uint8_t>(kv.first[i]);
    words.push_back(std::move(w));
    counts.push_back(kv.second);
  }
  int32_t learnt = 0;
  for (; learnt < max_merges; ++learnt) {
    std::unordered_map<std::pair<int32_t, int32_t>, int64_t, PairHash> pc;
    for (size_t wi = 0; wi < words.size(); ++wi) {
      const auto& w = words[wi];
      for (size_t i = 0; i + 1 < w.size(); ++i) pc[{w[i], w[i + 1]}] += counts[wi];
    }
    std::pair<int32_t, int32_t> best{-1, -1};
    int64_t bc = 0;
    for (auto& kv : pc)
      if (kv.second > bc || (kv.second == bc && kv.first < best)) { bc = kv.second; best = kv.first; }
    if (bc < min_frequency || bc <= 0) break;
    merges_out[2 * learnt] = best.first;
    merges_out[2 * learnt + 1] = best.second;
    const int32_t nid = 256 + learnt;
#pragma omp parallel for schedule(static)
    for (long wi = 0; wi < (long)words.size(); ++wi) {
      auto& w = words[wi];
      size_t o = 0;
      for (size_t i = 0; i < w.size(); ++i) {
        if (i + 1 < w.size() && w[i] == best.first && w[i + 1] == best.second) { w[o++] = nid; ++i; }
        else w[o++] = w[i];
      }
      w.resize(o);
    }
  }
  return learnt;
}

// Encode one text; returns the number of ids written (or -needed if out_cap is too small).
B200_HOST_API int64_t b200_bpe_encode(const uint8_t* text, int64_t n, const int32_t* merges, int32_t n_merges,
                                      int32_t* out, int64_t out_cap) {
  std::unordered_map<std::pair<int32_t, int32_t>, int32_t, PairHash> rank;
  rank.reserve(n_merges * 2);
  for (int32_t i = 0; i < n_merges; ++i) rank[{merges[2 * i], merges[2 * i + 1]}] = i;
  std::vector<std::pair<int64_t, int64_t>> spans;
  split_words(text, n, spans);
  std::vector<std::vector<int32_t>> enc(spans.size());
  std::unordered_map<std::string, std::vector<int32_t>> cache;  // filled serially below for repeated words
#pragma omp parallel for schedule(dynamic, 256)
  for (long si = 0; si < (long)spans.size(); ++si) {
    const auto& sp = spans[si];
    std::vector<int32_t> w(sp.second - sp.first);
    for (int64_t i = 0; i < sp.second - sp.first; ++i) w[i] = text[sp.first + i];
    apply_merges(w, rank);
    enc[si] = std::move(w);
  }
  int64_t total = 0;
  for (auto& e : enc) total += (int64_t)e.size();
  if (total > out_cap) return -total;
  int64_t o = 0;
  for (auto& e : enc) {
    memcpy(out + o, e.data(), e.size() * sizeof(int32_t));
    o += (int64_t)e.size();
  }
  return total;
}
