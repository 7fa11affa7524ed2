// tb_hostutil.h -- host-side helpers shared by the C ABI (tb_api.cu) and the host emulation
// harness (tests/host_emul): building the stop-word table the kernels probe.
#pragma once
#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/textblaster_b200.h"
#include "tb_lanes.cuh"

// Entries are inserted exactly as configured: the reference does not lower-case the list
// (gopher_quality.rs:46-58), only the word being looked up (:186).
inline std::vector<SwEntry> tb_build_stop_table(const std::vector<std::string>& words, uint32_t* mask, uint32_t* maxlen,
                                                 uint32_t* lenmask, uint32_t* firstmask, uint32_t* secondmask) {
  uint32_t T = 16;
  while (T < 2 * words.size() + 2) T <<= 1;
  std::vector<SwEntry> tab(T);
  memset(tab.data(), 0, T * sizeof(SwEntry));
  *maxlen = 0; *lenmask = 0; *firstmask = 0; *secondmask = 0;
  for (const auto& w : words) {
    LowKey k; k.init();
    for (unsigned char c : w) k.push(c);
    if (k.len == 0) continue;  // the empty string can never equal a word
    if (k.len > *maxlen) *maxlen = k.len;
    *lenmask |= 1u << (k.len < 31u ? k.len : 31u);
    *firstmask |= sw_first_bit((unsigned char)w[0]);
    *secondmask |= w.size() >= 2 ? sw_first_bit((unsigned char)w[1]) : 0xFFFFFFFFu;
    uint32_t s = (uint32_t)sw_slot_hash(k.lo, k.hi, k.ext, k.len) & (T - 1);
    bool dup = false;
    while (tab[s].used) {
      if (tab[s].len == k.len && tab[s].lo == k.lo && tab[s].hi == k.hi && tab[s].ext == k.ext) { dup = true; break; }
      s = (s + 1) & (T - 1);
    }
    if (!dup) tab[s] = SwEntry{k.lo, k.hi, k.ext, k.len, 1};
  }
  *mask = T - 1;
  return tab;
}
inline std::vector<std::string> tb_stop_words_of(const tb_gq_params& p) {
  std::vector<std::string> words;
  if (p.has & TB_GQ_HAS_STOP_WORDS) {
    for (int i = 0; i < p.n_stop_words; i++)
      words.emplace_back((const char*)p.stop_words_utf8 + p.stop_words_offsets[i],
                         (size_t)(p.stop_words_offsets[i + 1] - p.stop_words_offsets[i]));
  } else {
    words = {"the", "be", "to", "of", "and", "that", "have", "with"};  // DEFAULT_STOP_WORDS, gopher_quality.rs:10
  }
  return words;
}
inline C4Cfg tb_c4cfg_of(const tb_c4_params& q) {
  return C4Cfg{(uint32_t)q.remove_citations, (uint32_t)q.filter_no_terminal_punct, (uint32_t)q.filter_lorem_ipsum,
               (uint32_t)q.filter_javascript, (uint32_t)q.filter_curly_bracket, (uint32_t)q.filter_policy,
               q.min_num_sentences, q.min_words_per_line, q.max_word_length};
}

// units of the per-document scratch of a chunk whose content buffer spans cb bytes and holds n documents
// (tb_wunit / tb_lunit in tb_kernels.cuh: the last document ends at most at unit f(cb) + n)
inline size_t tb_word_units(size_t cb, size_t n) { return 2 * cb / 3 + n + 2; }
inline size_t tb_line_units(size_t cb, size_t n) { return cb / 2 + n + 2; }

// bytes the sentence-per-line copy of a chunk (nb bytes, n documents) can take: see tb_sent_off (tb_bodies.cuh)
inline size_t tb_grown_bytes(int64_t nb, int64_t n) { return (size_t)nb + (size_t)nb / 2 + 2 * (size_t)n + 64; }

// Sizes of the chunks a batch of B bytes is cut into (tb_api.cu run_impl; the documents are dealt to them in order,
// so a chunk ends up a little under its target).  Ramp up by doubling from 64 MiB (host input: the first copy to the
// device is exposed) or 256 MiB (text already on the device), plateau at C, ramp down by thirds to 48 MiB (the
// results of a chunk leave while the next, smaller one is computed).  The sizes sum to B.
// `floor_bytes`: no chunk smaller than this (a chunk of a few hundred long documents cannot fill the GPU: the caller
// passes the size of some thousand average documents).
inline std::vector<int64_t> tb_chunk_targets(int64_t B, int64_t C, bool device_input, int64_t floor_bytes = 0) {
  if (C < 1) C = 1;
  const int64_t lo_up = std::min<int64_t>(C, std::max<int64_t>(floor_bytes, (device_input ? 256ll : 64ll) << 20)),
                lo_down = std::min<int64_t>(C, std::max<int64_t>(floor_bytes, 48ll << 20));
  std::vector<int64_t> up, down;  // ramp sizes, smallest first
  for (int64_t v = lo_up; v < C; v *= 2) up.push_back(v);
  for (int64_t v = lo_down; v < C; v *= 3) down.push_back(v);
  // take ramp steps from both ends, smallest first, while they fit; what is left is the plateau
  std::vector<int64_t> head, tail;
  int64_t used = 0;
  for (size_t k = 0; k < std::max(up.size(), down.size()); k++) {
    if (k < up.size() && used + up[k] <= B) { head.push_back(up[k]); used += up[k]; }
    if (k < down.size() && used + down[k] <= B) { tail.push_back(down[k]); used += down[k]; }
  }
  int64_t mid = B - used;
  // a remainder smaller than its neighbours would break the ramp: it joins the largest ramp step that stays <= C
  if (mid > 0 && !(head.empty() && tail.empty())) {
    int64_t* big = nullptr;
    for (auto& v : head) if (!big || v > *big) big = &v;
    for (auto& v : tail) if (!big || v > *big) big = &v;
    if (mid < *big && *big + mid <= C) { *big += mid; mid = 0; }
  }
  std::vector<int64_t> targets = head;
  if (mid > 0) { const int64_t k = (mid + C - 1) / C; for (int64_t i = 0; i < k; i++) targets.push_back(mid / k + (i < mid % k ? 1 : 0)); }
  for (size_t k = tail.size(); k-- > 0;) targets.push_back(tail[k]);
  if (targets.empty()) targets.push_back(C);
  return targets;
}

// UcTables::wslead: the lead bytes of the non-ASCII White_Space characters, from the cinfo trie
inline void tb_compute_wslead(const uint16_t* stage1, const uint16_t* stage2, uint32_t out[8]) {
  for (int i = 0; i < 8; i++) out[i] = 0;
  for (uint32_t cp = 0x80; cp < 0x110000u; cp++) {
    uint32_t ci = stage2[(uint32_t)stage1[cp >> 8] * 256u + (cp & 255u)];
    if (!(ci & CI_WS)) continue;
    uint32_t lead = cp < 0x800u ? (0xC0u | (cp >> 6)) : cp < 0x10000u ? (0xE0u | (cp >> 12)) : (0xF0u | (cp >> 18));
    out[lead >> 5] |= 1u << (lead & 31u);
  }
}
