// tb_synth.cpp -- deterministic synthetic C4-shaped corpus for bench.py and the
// parity tests.  Host-only helper (g++), not part of the hot path.  Every document is generated
// from splitmix64(seed, doc index), so any slice of the corpus can be produced independently.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace {

struct Rng {
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed) {}
  uint64_t next() { uint64_t z = (s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }
  double uni() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
  uint32_t below(uint32_t n) { return (uint32_t)(((next() >> 32) * (uint64_t)n) >> 32); }
  int range(int lo, int hi) { return lo + (int)below((uint32_t)(hi - lo + 1)); }  // inclusive
  bool chance(double p) { return uni() < p; }
  double normal() { double u1 = uni(), u2 = uni(); if (u1 < 1e-300) u1 = 1e-300; return std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586 * u2); }
};

struct Vocab {
  std::vector<std::string> words;
  std::vector<double> cdf;
  explicit Vocab(uint64_t seed) {
    const int V = 50000;
    Rng r(seed ^ 0xC0FFEEull);
    const char* stops[8] = {"the", "be", "to", "of", "and", "that", "have", "with"};  // ranks 1-8
    words.reserve(V);
    for (int i = 0; i < V; i++) {
      if (i < 8) { words.emplace_back(stops[i]); continue; }
      // shifted Poisson(mean 4.7): 1 + Poisson(3.7), clipped to [1, 20]
      int k = 0; double L = std::exp(-3.7), p = 1.0;
      do { k++; p *= r.uni(); } while (p > L);
      int len = std::min(20, std::max(1, k));
      std::string w;
      for (int j = 0; j < len; j++) w.push_back("etaoinshrdlcumwfgypbvkjxqz"[std::min<uint32_t>(25, (uint32_t)(r.uni() * r.uni() * 26))]);
      words.push_back(w);
    }
    cdf.resize(V);
    double acc = 0;
    for (int i = 0; i < V; i++) { acc += 1.0 / std::pow((double)(i + 1), 1.07); cdf[i] = acc; }
    for (auto& c : cdf) c /= acc;
  }
  const std::string& sample(Rng& r) const {
    double u = r.uni();
    size_t i = std::lower_bound(cdf.begin(), cdf.end(), u) - cdf.begin();
    return words[std::min(i, words.size() - 1)];
  }
};

const char* POLICY[6] = {"terms of use", "privacy policy", "cookie policy", "uses cookies", "use of cookies", "use cookies"};

void sentence(const Vocab& v, Rng& r, std::string& out, int nwords, bool terminal) {
  for (int i = 0; i < nwords; i++) {
    const std::string& w = v.sample(r);
    size_t at = out.size();
    out += w;
    if (i == 0 && out[at] >= 'a' && out[at] <= 'z') out[at] = (char)(out[at] - 32);
    if (i + 1 < nwords) { if (r.chance(0.04)) out += ','; out += ' '; }
  }
  if (terminal) { double u = r.uni(); out += u < 0.85 ? '.' : u < 0.93 ? '?' : '!'; }
}

void make_doc(const Vocab& v, uint64_t seed, uint64_t idx, double median, double sigma, std::string& out) {
  Rng r(seed * 0x9E3779B97F4A7C15ull + idx * 0xD1B54A32D192ED03ull + 1);
  out.clear();
  double target = median * std::exp(sigma * r.normal());
  target = std::min(65536.0, std::max(64.0, target));
  const bool lorem = r.chance(0.002), curly = r.chance(0.02), boiler = r.chance(0.05), phrase = r.chance(0.05),
             longtok = r.chance(0.001);
  std::string boiler_line, phrase_text;
  if (boiler) { sentence(v, r, boiler_line, r.range(5, 12), true); }
  if (phrase) { sentence(v, r, phrase_text, r.range(6, 12), false); }
  int boiler_left = boiler ? r.range(3, 20) : 0, phrase_left = phrase ? r.range(5, 50) : 0;
  bool first_para = true;
  while ((double)out.size() < target) {
    if (!first_para) out += "\n\n";
    first_para = false;
    int nlines = r.range(1, 6);
    for (int l = 0; l < nlines && (double)out.size() < target; l++) {
      if (l) out += '\n';
      if (boiler_left && r.chance(0.3)) { out += boiler_line; boiler_left--; continue; }
      if (r.chance(0.05)) out += r.chance(0.5) ? "- " : "\xE2\x80\xA2 ";
      if (r.chance(0.10)) { sentence(v, r, out, r.range(1, 2), true); continue; }   // fewer than 3 words
      int ns = r.range(1, 5);
      bool no_term = r.chance(0.20);
      for (int s = 0; s < ns; s++) {
        if (s) out += ' ';
        if (phrase_left && r.chance(0.4)) { out += phrase_text; out += '.'; phrase_left--; continue; }
        sentence(v, r, out, r.range(8, 25), !(no_term && s + 1 == ns));
        if (r.chance(0.05)) { out += " ["; out += std::to_string(r.range(1, 99)); if (r.chance(0.3)) { out += ", "; out += std::to_string(r.range(1, 99)); } out += ']'; }
      }
      if (r.chance(0.01)) out += " javascript";
      if (r.chance(0.01)) { out += ' '; out += POLICY[r.below(6)]; out += '.'; }
      if (r.chance(0.02)) out += "...";
      if (r.chance(0.01)) out += " # #";
    }
  }
  if (lorem) out += " Lorem ipsum dolor sit amet.";
  if (curly) out += r.chance(0.5) ? " function() { return 1; }" : " }";
  if (longtok) { out += ' '; out.append((size_t)r.range(1001, 1500), 'x'); out += '.'; }
}

struct Corpus { std::vector<std::string> parts; std::vector<std::vector<int64_t>> lens; int64_t n, total; };

}  // namespace

extern "C" {

void* tbs_generate(int64_t n, uint64_t seed, int64_t first_index, double median_bytes, double sigma, int nthreads, int64_t* total) {
  static Vocab* vocab = nullptr;
  static uint64_t vocab_seed = 0;
  const uint64_t vs = 0x7EB1A57Eull;  // the vocabulary is shared by every corpus
  if (!vocab || vocab_seed != vs) { delete vocab; vocab = new Vocab(vs); vocab_seed = vs; }
  if (nthreads < 1) nthreads = 1;
  auto* c = new Corpus();
  c->n = n; c->parts.resize(nthreads); c->lens.resize(nthreads);
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; t++) {
    th.emplace_back([&, t]() {
      int64_t a = n * t / nthreads, b = n * (t + 1) / nthreads;
      std::string doc;
      c->parts[t].reserve((size_t)((b - a) * median_bytes * 1.7) + 1024);
      for (int64_t i = a; i < b; i++) {
        make_doc(*vocab, seed, (uint64_t)(first_index + i), median_bytes, sigma, doc);
        c->parts[t] += doc;
        c->lens[t].push_back((int64_t)doc.size());
      }
    });
  }
  for (auto& x : th) x.join();
  c->total = 0;
  for (auto& p : c->parts) c->total += (int64_t)p.size();
  *total = c->total;
  return c;
}
void tbs_copy(void* h, uint8_t* out, int64_t* offsets) {
  auto* c = (Corpus*)h;
  int64_t pos = 0, d = 0;
  for (size_t t = 0; t < c->parts.size(); t++) {
    memcpy(out + pos, c->parts[t].data(), c->parts[t].size());
    int64_t p = pos;
    for (int64_t l : c->lens[t]) { offsets[d++] = p; p += l; }
    pos += (int64_t)c->parts[t].size();
  }
  offsets[d] = pos;
}
void tbs_free(void* h) { delete (Corpus*)h; }

}  // extern "C"
