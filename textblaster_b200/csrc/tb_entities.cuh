// tb_entities.cuh -- HTML entity decoding of the text column, as a pre-pass in front of the filter chain
// (SURVEY 8(f) row 1).  The reference decodes in its reader, before any step sees the text:
//   html_escape::decode_html_entities(raw_text)      /root/reference src/pipeline/readers/parquet_reader.rs:178-179
// `html-escape 0.2.13` (Cargo.lock) is not part of the reference tree and the reference has no test with an
// entity in it, so this restates the crate's documented behaviour -- PARITY UNPINNED:
//   &name;    a WHATWG named character reference, with its semicolon, case-sensitive  -> its 1-2 characters
//   &#D..D;   decimal,  &#xH..H; / &#XH..H;  hexadecimal                              -> char::from_u32(value)
// Anything else stays as it is: an unknown name, a missing semicolon, no digits, a value that does not fit
// u32 or is not a scalar value (surrogates, > 0x10FFFF).  Choices the documentation does not settle, stated
// here so that they can be checked against the crate one day: a candidate ends at the first byte that cannot
// belong to it (names are ASCII letters and digits), so a later '&' starts a new candidate ("AT&T &amp; co"
// decodes the "&amp;"); U+0000 is a scalar value and is produced for "&#0;"; a candidate longer than
// TB_ENT_MAX_SPAN bytes (only reachable with ~20 leading zeros) is left alone.
#pragma once
#include "tb_common.cuh"

#define TB_ENT_MAX_SPAN 40  // '&' .. ';' inclusive
struct EntEntry { char name[32]; uint8_t nlen, rlen; uint8_t repl[6]; };

// Candidate starting at the '&' at position a of [.., re).  Returns the replacement length (0: not an entity);
// *span = bytes of the entity including '&' and ';', *repl = the replacement bytes, first byte lowest.
TB_HD uint32_t ent_parse(const uint8_t* t, int64_t a, int64_t re, const EntEntry* tab, uint32_t ntab, uint32_t* span, uint64_t* repl) {
  int64_t p = a + 1;
  const int64_t lim = a + TB_ENT_MAX_SPAN < re ? a + TB_ENT_MAX_SPAN : re;
  if (p >= lim) return 0;
  uint32_t b = tb_byte(t, p);
  if (b == '#') {
    p++;
    if (p >= lim) return 0;
    b = tb_byte(t, p);
    const bool hex = b == 'x' || b == 'X';
    if (hex) p++;
    uint64_t v = 0;
    int64_t d0 = p;
    for (; p < lim; p++) {
      b = tb_byte(t, p);
      uint32_t dv;
      if (b >= '0' && b <= '9') dv = b - '0';
      else if (hex && b >= 'a' && b <= 'f') dv = b - 'a' + 10;
      else if (hex && b >= 'A' && b <= 'F') dv = b - 'A' + 10;
      else break;
      v = v * (hex ? 16u : 10u) + dv;
      if (v > 0xFFFFFFFFull) return 0;  // u32 overflow: parse error
    }
    if (p >= lim || p == d0 || b != ';') return 0;
    if (v > 0x10FFFFull || (v >= 0xD800ull && v <= 0xDFFFull)) return 0;  // char::from_u32 is None
    const uint32_t c = (uint32_t)v;
    *span = (uint32_t)(p - a + 1);
    if (c < 0x80u) { *repl = c; return 1; }
    if (c < 0x800u) { *repl = (0xC0u | (c >> 6)) | ((uint64_t)(0x80u | (c & 63u)) << 8); return 2; }
    if (c < 0x10000u) { *repl = (0xE0u | (c >> 12)) | ((uint64_t)(0x80u | ((c >> 6) & 63u)) << 8) | ((uint64_t)(0x80u | (c & 63u)) << 16); return 3; }
    *repl = (0xF0u | (c >> 18)) | ((uint64_t)(0x80u | ((c >> 12) & 63u)) << 8) | ((uint64_t)(0x80u | ((c >> 6) & 63u)) << 16) |
            ((uint64_t)(0x80u | (c & 63u)) << 24);
    return 4;
  }
  // named: ASCII letters and digits up to ';'
  const int64_t n0 = p;
  for (; p < lim; p++) {
    b = tb_byte(t, p);
    if (!((b >= '0' && b <= '9') || ((b | 0x20u) >= 'a' && (b | 0x20u) <= 'z'))) break;
  }
  const uint32_t nlen = (uint32_t)(p - n0);
  if (p >= lim || b != ';' || nlen == 0 || nlen > 31) return 0;
  uint32_t lo = 0, hi = ntab;  // binary search over the names (bytewise order, shorter prefix first)
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    const EntEntry& e = tab[mid];
    int cmp = 0;
    const uint32_t m = nlen < e.nlen ? nlen : e.nlen;
    for (uint32_t k = 0; k < m && cmp == 0; k++) cmp = (int)tb_byte(t, n0 + k) - (int)(uint8_t)e.name[k];
    if (cmp == 0) cmp = (int)nlen - (int)e.nlen;
    if (cmp == 0) {
      uint64_t r = 0;
      for (uint32_t k = 0; k < e.rlen; k++) r |= (uint64_t)e.repl[k] << (8 * k);
      *repl = r; *span = nlen + 2;
      return e.rlen;
    }
    if (cmp < 0) hi = mid; else lo = mid + 1;
  }
  return 0;
}
