// tb_common.cuh -- shared device/host-emulation helpers: Unicode property access, UTF-8
// stepping, the 2 x Mersenne-31 polynomial hash.  Everything here is per-lane code (no warp
// collectives) so the same functions can be driven serially by the host emulation harness in
// tests/host_emul (test infrastructure; never part of the shipped library).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define TB_HD __host__ __device__ __forceinline__
#define TB_D __device__ __forceinline__
#define TB_HD_NOINLINE static __host__ __device__ __noinline__
#define TB_D_NOINLINE static __device__ __noinline__
#else
#define TB_HD inline
#define TB_D inline
#define TB_HD_NOINLINE inline
#define TB_D_NOINLINE inline
#endif

// ---- cinfo bit layout (tools/gen_unicode_tables.py) ------------------------------------------
#define CI_WB(ci) ((ci) & 31u)
#define CI_SB(ci) (((ci) >> 5) & 15u)
#define CI_WS (1u << 9)
#define CI_PUNCT (1u << 10)
#define CI_ALPHA (1u << 11)
#define CI_EXTPICT (1u << 12)
#define CI_ND (1u << 13)
#define CI_COMPLEX (1u << 14)
#define CI_LOWERS (1u << 15)

// Word_Break / Sentence_Break class numbers -- must match unicode_tables.inc (static_assert'ed in
// tb_tables.cu).
enum {
  WB_OTHER = 0, WB_CR, WB_LF, WB_NEWLINE, WB_EXTEND, WB_ZWJ, WB_RI, WB_FORMAT, WB_KATAKANA, WB_HEBREW,
  WB_ALETTER, WB_SQ, WB_DQ, WB_MIDNUMLET, WB_MIDLETTER, WB_MIDNUM, WB_NUMERIC, WB_EXTENDNUMLET, WB_WSEG,
  WB_SOT = 31  // pseudo class: nothing before
};
enum {
  SB_OTHER = 0, SB_CR, SB_LF, SB_SEP, SB_EXTEND, SB_FORMAT, SB_SP, SB_LOWER, SB_UPPER, SB_OLETTER,
  SB_NUMERIC, SB_ATERM, SB_SCONTINUE, SB_STERM, SB_CLOSE,
  SB_SOT = 15
};
#define BIT(c) (1u << (c))

// Device-resident Unicode tables (uploaded once per context).
struct UcTables {
  const uint16_t* stage1;   // [0x1100]
  const uint16_t* stage2;   // [nblocks * 256]
  const uint32_t* lower;    // [n_lower][2] sorted by code point
  const uint32_t* cased;    // [n_cased][2] ranges [a,b)
  const uint32_t* caseign;  // [n_caseign][2]
  const uint64_t* powtab;   // [TB_POWTAB_ALL] packed B^k for both moduli, k <= 1024, then B^(1024 k), k = 1 .. 1024
  int n_lower, n_cased, n_caseign;
  uint32_t wslead[8];       // bit b: some non-ASCII White_Space character starts with byte b (tb_compute_wslead)
};
#define TB_POWTAB 1025
#define TB_POWTAB_ALL (TB_POWTAB + 1024)  // + the coarse table: entry TB_POWTAB - 1 + k = B^(1024 k), k = 1 .. 1024

#if defined(__CUDA_ARCH__)
#define TB_LDG(p) __ldg(p)
#else
#define TB_LDG(p) (*(p))
#endif

// Text view of one document range; positions are absolute byte indices into `t`.
struct TextRange {
  const uint8_t* t;
  int64_t rs, re;  // [rs, re): sot at rs, eot at re
};

TB_HD uint32_t tb_byte(const uint8_t* t, int64_t i) { return TB_LDG(t + i); }
TB_HD bool tb_is_cont(uint32_t b) { return (b & 0xC0u) == 0x80u; }

// Decode the code point whose lead byte is at i (valid UTF-8 assumed, i < re). Returns length.
TB_HD int tb_decode(const uint8_t* t, int64_t i, uint32_t b0, uint32_t* cp) {
  if (b0 < 0x80u) { *cp = b0; return 1; }
  if (b0 < 0xE0u) { *cp = ((b0 & 0x1Fu) << 6) | (tb_byte(t, i + 1) & 0x3Fu); return 2; }
  if (b0 < 0xF0u) { *cp = ((b0 & 0x0Fu) << 12) | ((tb_byte(t, i + 1) & 0x3Fu) << 6) | (tb_byte(t, i + 2) & 0x3Fu); return 3; }
  *cp = ((b0 & 0x07u) << 18) | ((tb_byte(t, i + 1) & 0x3Fu) << 12) | ((tb_byte(t, i + 2) & 0x3Fu) << 6) | (tb_byte(t, i + 3) & 0x3Fu);
  return 4;
}
TB_HD int tb_u8len(uint32_t b0) { return b0 < 0x80u ? 1 : b0 < 0xE0u ? 2 : b0 < 0xF0u ? 3 : 4; }

TB_HD uint32_t tb_cinfo_trie(const UcTables& u, uint32_t cp) {
  if (cp >= 0x110000u) return 0;
  uint32_t blk = TB_LDG(u.stage1 + (cp >> 8));
  return TB_LDG(u.stage2 + blk * 256u + (cp & 255u));
}

// Classifier: ASCII from a 128-entry table (shared memory on the device), the rest via the trie.
struct Classifier {
  const uint16_t* ascii;  // [128] cinfo of the ASCII code points, followed by 128 bytes of fast-path class flags
  UcTables u;
  uint32_t key_maxlen = 0xFFFFFFFFu;  // longest stop word: longer words need no lower-cased key
  uint32_t key_lenmask = 0xFFFFFFFFu, key_firstmask = 0xFFFFFFFFu;  // cheap reject by length / first letter (SwTable)
  TB_HD const uint8_t* aflags() const { return (const uint8_t*)(ascii + 128); }
  TB_HD uint32_t of_cp(uint32_t cp) const { return cp < 128u ? ascii[cp] : tb_cinfo_trie(u, cp); }
  // classify the code point starting at i; returns cinfo, sets *len (and *cp)
  TB_HD uint32_t at(const uint8_t* t, int64_t i, int* len, uint32_t* cp) const {
    uint32_t b0 = tb_byte(t, i);
    if (b0 < 0x80u) { *len = 1; *cp = b0; return ascii[b0]; }
    *len = tb_decode(t, i, b0, cp);
    return tb_cinfo_trie(u, *cp);
  }
  // position of the lead byte of the code point that ends just before `i` (i > lo)
  TB_HD int64_t prev_lead(const uint8_t* t, int64_t i, int64_t lo) const {
    int64_t j = i - 1;
    while (j > lo && tb_is_cont(tb_byte(t, j))) j--;
    return j;
  }
};

TB_HD bool tb_in_ranges(const uint32_t* r, int n, uint32_t cp) {
  int lo = 0, hi = n;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    uint32_t a = TB_LDG(r + 2 * mid), b = TB_LDG(r + 2 * mid + 1);
    if (cp < a) hi = mid; else if (cp >= b) lo = mid + 1; else return true;
  }
  return false;
}
TB_HD uint32_t tb_simple_lower(const UcTables& u, uint32_t cp) {
  int lo = 0, hi = u.n_lower;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (TB_LDG(u.lower + 2 * mid) < cp) lo = mid + 1; else hi = mid;
  }
  if (lo < u.n_lower && TB_LDG(u.lower + 2 * lo) == cp) return TB_LDG(u.lower + 2 * lo + 1);
  return cp;
}

// ---- 2 x Mersenne-31 polynomial hash ---------------------------------------------------------
// H(b_0..b_{L-1}) = sum (b_k + 1) * B^(L-1-k)  mod p,  p = 2^31 - 1, two independent bases.
// It composes under concatenation: H(xy) = H(x) * B^|y| + H(y), which is what the concatenated
// n-gram keys of find_all_duplicate (text.rs:250) need.  Two keys are treated as equal when both
// 31-bit residues agree.  The two bases are drawn at random when a context is created (tb_ctx_create; below 2^29, which
// lets tb_hpush skip the reduction between steps) and reach the kernels as entry 1 of the power table: for two different
// strings of length <= L the two residues agree with probability <= (L / 2^28)^2 over that draw, whatever the text.
#define TB_P31 0x7FFFFFFFu
#define TB_BASE_BITS 29
// the fixed pair the tests and the host emulation use (any pair in [2^28, 2^29) does)
#define TB_B1_DEFAULT 0x1F0B4A27u
#define TB_B2_DEFAULT 0x0C4D3A11u
TB_HD uint32_t tb_mm(uint32_t a, uint32_t b) {  // a, b <= p  ->  result <= p (p itself stands for 0)
  const uint64_t x = (uint64_t)a * b;  // < 2^62: both halves of the first fold are below 2^31, their sum fits 32 bits
  const uint32_t s = ((uint32_t)x & TB_P31) + (uint32_t)(x >> 31);
  return (s & TB_P31) + (s >> 31);
}
TB_HD uint32_t tb_am(uint32_t a, uint32_t b) {  // a, b <= p -> <= p
  uint32_t x = a + b;
  return (x & TB_P31) + (x >> 31);
}
// bring the cache line of p towards L1 (no register, no wait); nothing on the host
TB_HD void tb_prefetch(const void* p) {
#if defined(__CUDA_ARCH__)
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#else
  (void)p;
#endif
}
TB_HD uint32_t tb_canon31(uint32_t x) { return x == TB_P31 ? 0u : x; }

struct Hash2 {
  uint32_t a, b;
  TB_HD void init() { a = 0; b = 0; }
  // pw1: the packed bases (entry 1 of the power table)
  TB_HD void push(uint32_t byte, uint64_t pw1) { a = tb_am(tb_mm(a, (uint32_t)(pw1 >> 32)), byte + 1u); b = tb_am(tb_mm(b, (uint32_t)pw1), byte + 1u); }
  TB_HD uint64_t pack() const { return ((uint64_t)tb_canon31(a) << 32) | tb_canon31(b); }
};
TB_HD Hash2 tb_unpack(uint64_t v) { Hash2 h; h.a = (uint32_t)(v >> 32); h.b = (uint32_t)v; return h; }
// x * B^len(y) + y  with pw = packed (B1^len, B2^len)
TB_HD Hash2 tb_hcat(Hash2 x, Hash2 y, uint64_t pw) {
  Hash2 r;
  r.a = tb_am(tb_mm(x.a, (uint32_t)(pw >> 32)), y.a);
  r.b = tb_am(tb_mm(x.b, (uint32_t)pw), y.b);
  return r;
}
// (B1^len, B2^len) packed, by square-and-multiply: the rare element longer than the table (out of line)
TB_HD_NOINLINE uint64_t tb_pow_long(const uint64_t* powtab, uint64_t len) {
  const uint64_t pw1 = TB_LDG(powtab + 1);
  uint32_t ra = 1, rb = 1, ba = (uint32_t)(pw1 >> 32), bb = (uint32_t)pw1;
  uint64_t e = len;
  while (e) {
    if (e & 1) { ra = tb_mm(ra, ba); rb = tb_mm(rb, bb); }
    ba = tb_mm(ba, ba); bb = tb_mm(bb, bb);
    e >>= 1;
  }
  return ((uint64_t)ra << 32) | rb;
}
// (lengths up to a megabyte -- long lines, paragraphs -- from two table entries and one multiplication per modulus)
TB_HD uint64_t tb_pow_pack(const uint64_t* powtab, uint64_t len) {
  if (len < TB_POWTAB) return TB_LDG(powtab + len);
  if (len < (1024ull << 10)) {
    const uint64_t hi = TB_LDG(powtab + (TB_POWTAB - 1) + (len >> 10)), lo = TB_LDG(powtab + (len & 1023u));
    return ((uint64_t)tb_mm((uint32_t)(hi >> 32), (uint32_t)(lo >> 32)) << 32) | tb_mm((uint32_t)hi, (uint32_t)lo);
  }
  return tb_pow_long(powtab, len);
}
// ---- suffix signature of a byte string: its last 16 bytes, four bits of each, the last byte in the low nibble
// (a string shorter than 16 bytes leaves the upper nibbles zero).  Equal strings have equal signatures, and so
// do their suffixes of any common length k <= 16 (mask with tb_sig_mask(k)): the repetition pass uses that as a
// cheap necessary condition for two concatenated n-gram keys to be equal (tb_rep.cuh).
TB_HD uint64_t tb_sig_push(uint64_t s, uint32_t byte) { return (s << 4) | (uint64_t)(byte & 15u); }
TB_HD uint64_t tb_sig_cat(uint64_t x, uint64_t y, uint32_t ylen) { return ylen >= 16u ? y : (y | (x << (4u * ylen))); }  // x followed by y
TB_HD uint64_t tb_sig_mask(uint32_t k) { return k >= 16u ? ~0ull : ((1ull << (4u * k)) - 1ull); }

// the power table of a pair of bases: out[k] = (B1^k mod p) << 32 | (B2^k mod p), k < TB_POWTAB; out has TB_POWTAB_ALL entries
TB_HD void tb_fill_powtab(uint64_t* out, uint32_t b1, uint32_t b2) {
  uint32_t a = 1, b = 1;
  for (int i = 0; i < TB_POWTAB; i++) { out[i] = ((uint64_t)a << 32) | b; a = tb_mm(a, b1); b = tb_mm(b, b2); }
  // the coarse table: powers of B^1024
  const uint32_t s1 = (uint32_t)(out[1024] >> 32), s2 = (uint32_t)out[1024];
  a = s1; b = s2;
  for (int k = 1; k <= 1024; k++) { out[TB_POWTAB - 1 + k] = ((uint64_t)a << 32) | b; a = tb_mm(a, s1); b = tb_mm(b, s2); }
}
TB_HD uint64_t tb_mix64(uint64_t x) {  // splitmix64 finaliser, for table slot selection only
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31;
  return x;
}
