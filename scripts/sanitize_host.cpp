// Driver for scripts/sanitize_host.sh: exercises every entry point of libb200host (csrc/host/*.cpp) under
// AddressSanitizer + UndefinedBehaviorSanitizer with exact-size heap buffers, so any out-of-bounds access, misaligned
// access or signed overflow inside the BPE trainer / encoder or the host AdamW aborts the run.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

extern "C" {
int64_t b200_bpe_train(const uint8_t* text, int64_t n, int32_t max_merges, int32_t min_frequency, int32_t* merges_out);
int64_t b200_bpe_encode(const uint8_t* text, int64_t n, const int32_t* merges, int32_t n_merges, int32_t* out, int64_t out_cap);
void b200_cpu_adamw(float* p, const float* g, float* m, float* v, uint16_t* p_bf16, int64_t n, float lr, float beta1, float beta2,
                    float eps, float weight_decay, int64_t step, float grad_scale);
double b200_cpu_sumsq(const float* g, int64_t n, float scale);
}

static int fail(const char* what) {
  std::fprintf(stderr, "FAILED: %s\n", what);
  return 1;
}

int main() {
  // ---- BPE: train on a corpus with multi-byte UTF-8, runs of whitespace, a trailing partial word ----
  std::string corpus;
  const char* lines[] = {"the quick brown fox jumps over the lazy dog\n", "pack my box with five dozen liquor jugs\t\t\n",
                         "\xe9\xa9\xac\xe5\x93\xa5\xe6\x95\x99\xe8\x82\xb2 \xe4\xba\x91\xe5\x8e\x9f\xe7\x94\x9f   devops sre\n", "tail"};
  for (int r = 0; r < 200; ++r)
    for (const char* l : lines) corpus += l;
  const int32_t want = 300;
  std::vector<int32_t> merges(2 * want);
  const int64_t got = b200_bpe_train(reinterpret_cast<const uint8_t*>(corpus.data()), (int64_t)corpus.size(), want, 2, merges.data());
  if (got <= 0 || got > want) return fail("bpe_train merge count");
  for (int64_t i = 0; i < got; ++i)
    if (merges[2 * i] < 0 || merges[2 * i] >= 256 + i || merges[2 * i + 1] < 0 || merges[2 * i + 1] >= 256 + i) return fail("merge refers forward");
  // encode texts of awkward sizes into exact-size buffers (the documented capacity is one id per byte, at least one)
  const std::string texts[] = {"", " ", "a", "the quick brown fox", "\xe9\xa9\xac\xe5\x93\xa5", corpus.substr(0, 4097), std::string(1000, ' ')};
  for (const std::string& t : texts) {
    std::vector<int32_t> out(t.empty() ? 1 : t.size());
    const int64_t k = b200_bpe_encode(reinterpret_cast<const uint8_t*>(t.data()), (int64_t)t.size(), merges.data(), (int32_t)got, out.data(),
                                      (int64_t)out.size());
    if (k < 0 || k > (int64_t)out.size()) return fail("bpe_encode count");
    // decode back through the merge table: the bytes must be the input
    std::vector<std::string> piece(256 + got);
    for (int b = 0; b < 256; ++b) piece[b] = std::string(1, (char)b);
    for (int64_t i = 0; i < got; ++i) piece[256 + i] = piece[merges[2 * i]] + piece[merges[2 * i + 1]];
    std::string back;
    for (int64_t i = 0; i < k; ++i) back += piece[out[i]];
    if (back != t) return fail("bpe round trip");
  }
  // a too-small output buffer must be respected
  {
    const std::string t = "the quick brown fox jumps";
    std::vector<int32_t> out(3);
    const int64_t k = b200_bpe_encode(reinterpret_cast<const uint8_t*>(t.data()), (int64_t)t.size(), merges.data(), (int32_t)got, out.data(), 3);
    if (k > 3) return fail("bpe_encode ignored out_cap");
  }
  // ---- host AdamW: sizes around the vector width, with and without the bf16 mirror ----
  for (int64_t n : {1LL, 7LL, 8LL, 9LL, 63LL, 1000LL, 100003LL}) {
    std::vector<float> p(n), g(n), m(n, 0.f), v(n, 0.f), ref(n);
    std::vector<uint16_t> lo(n);
    for (int64_t i = 0; i < n; ++i) {
      p[i] = ref[i] = std::sin(0.1f * (float)i);
      g[i] = std::cos(0.3f * (float)i);
    }
    for (int64_t step = 1; step <= 3; ++step) {
      b200_cpu_adamw(p.data(), g.data(), m.data(), v.data(), (step & 1) ? lo.data() : nullptr, n, 1e-2f, 0.9f, 0.999f, 1e-8f, 0.01f, step, 1.0f);
    }
    double diff = 0;
    for (int64_t i = 0; i < n; ++i) diff += std::fabs(p[i] - ref[i]);
    if (!(diff > 0) || !std::isfinite(diff)) return fail("adamw did not move the parameters");
    const double s = b200_cpu_sumsq(g.data(), n, 1.0f);
    double want_s = 0;
    for (int64_t i = 0; i < n; ++i) want_s += (double)g[i] * g[i];
    if (std::fabs(s - want_s) > 1e-6 * (1 + want_s)) return fail("sumsq");
  }
  std::puts("sanitize_host: all entry points ran clean under ASan + UBSan");
  return 0;
}
