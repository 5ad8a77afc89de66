// Raw DEFLATE (RFC 1951) decoder for whole-member inflation of .npz episode files.
//
// zlib's inflate runs at ~120 MB/s on episode data (float32 states and uint8 frames compress
// into literal-heavy streams); this one keeps a 64-bit bit buffer refilled with one unaligned
// load, resolves most codes with one lookup in an 11-bit table (entries carry base and extra-bit
// count), decodes runs of literals without refilling and copies matches eight bytes at a time.
// It only handles the case the ingestion needs — the whole stream into one buffer of known size —
// and reports anything unexpected as an error; the caller (npz_ingest.cu) verifies the CRC-32 of
// the result and falls back to zlib, so a defect here costs speed, never data.
// Host-only.  Replaces the inflate inside np.load (utils/replay_buffer.py:26-29).
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <string.h>

namespace exb {
namespace finf {

constexpr int LIT_BITS = 11, DIST_BITS = 8, PRE_BITS = 7;
constexpr int LIT_CAP = 8192, DIST_CAP = 8192;
// table entry: bits 0-4 code bits to consume (for F_SUB: the primary bits), bits 8-12 extra-bit
// count (for F_SUB: index bits of the subtable), bits 13-15 flags, bits 16-31 payload (literal,
// base length, base distance, or first index of the subtable).  0 = no such code.
constexpr uint32_t F_SUB = 1u << 13, F_EOB = 1u << 14, F_LIT = 1u << 15;

struct Tables {
  uint32_t lit[LIT_CAP];
  uint32_t dist[DIST_CAP];
};

static inline uint64_t load64(const uint8_t* p) {
  uint64_t v;
  memcpy(&v, p, 8);
  return v;   // little-endian hosts only (x86-64, aarch64)
}
static inline void store64(uint8_t* p, uint64_t v) { memcpy(p, &v, 8); }

static inline uint32_t reverse_bits(uint32_t code, int len) {
  uint32_t r = 0;
  for (int i = 0; i < len; ++i) {
    r = (r << 1) | (code & 1u);
    code >>= 1;
  }
  return r;
}

static const uint16_t LEN_BASE[29] = {3,  4,  5,  6,  7,  8,  9,  10, 11,  13,  15,  17,  19,  23, 27,
                                      31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
static const uint8_t LEN_EXTRA[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2,
                                      2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
static const uint16_t DIST_BASE[30] = {1,   2,   3,   4,   5,   7,    9,    13,   17,   25,
                                       33,  49,  65,  97,  129, 193,  257,  385,  513,  769,
                                       1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
static const uint8_t DIST_EXTRA[30] = {0, 0, 0, 0, 1, 1, 2, 2,  3,  3,  4,  4,  5,  5,  6,
                                       6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

// kind 0: literal/length alphabet, 1: distance alphabet, 2: code-length alphabet (payload = symbol)
static inline uint32_t symbol_entry(int kind, int s) {
  if (kind == 2) return (uint32_t)s << 16;
  if (kind == 1) return s < 30 ? ((uint32_t)DIST_BASE[s] << 16) | ((uint32_t)DIST_EXTRA[s] << 8) : 0u;
  if (s < 256) return ((uint32_t)s << 16) | F_LIT;
  if (s == 256) return F_EOB;
  if (s < 286) return ((uint32_t)LEN_BASE[s - 257] << 16) | ((uint32_t)LEN_EXTRA[s - 257] << 8);
  return 0u;
}

// Canonical Huffman decoding table from code lengths.  False on an over-subscribed code or a
// table that does not fit.
static inline bool build_table(const uint8_t* lens, int n, int pbits, int kind, uint32_t* table,
                               int cap) {
  int count[16] = {0};
  for (int s = 0; s < n; ++s) count[lens[s]]++;
  count[0] = 0;
  int left = 1;
  for (int l = 1; l <= 15; ++l) {
    left = (left << 1) - count[l];
    if (left < 0) return false;
  }
  uint32_t first[16];
  uint32_t code = 0;
  for (int l = 1; l <= 15; ++l) {
    code = (code + (uint32_t)count[l - 1]) << 1;
    first[l] = code;
  }
  const int psize = 1 << pbits;
  const uint32_t pmask = (uint32_t)psize - 1u;
  memset(table, 0, sizeof(uint32_t) * (size_t)psize);
  // pass 1: the longest code behind each primary slot decides its subtable's size
  uint8_t submax[1 << LIT_BITS];
  bool any_long = false;
  uint32_t next[16];
  memcpy(next, first, sizeof(next));
  memset(submax, 0, (size_t)psize);
  for (int s = 0; s < n; ++s) {
    const int l = lens[s];
    if (!l) continue;
    const uint32_t rev = reverse_bits(next[l]++, l);
    if (l > pbits) {
      any_long = true;
      uint8_t& m = submax[rev & pmask];
      if (m < l) m = (uint8_t)l;
    }
  }
  int used = psize;
  if (any_long) {
    for (int i = 0; i < psize; ++i) {
      if (!submax[i]) continue;
      const int sbits = submax[i] - pbits;
      if (used + (1 << sbits) > cap) return false;
      table[i] = ((uint32_t)used << 16) | F_SUB | ((uint32_t)sbits << 8) | (uint32_t)pbits;
      memset(table + used, 0, sizeof(uint32_t) << sbits);
      used += 1 << sbits;
    }
  }
  // pass 2: fill
  memcpy(next, first, sizeof(next));
  for (int s = 0; s < n; ++s) {
    const int l = lens[s];
    if (!l) continue;
    const uint32_t rev = reverse_bits(next[l]++, l);
    const uint32_t ent = symbol_entry(kind, s);
    if (ent == 0u && kind != 2) continue;   // symbol outside the alphabet: its slots stay 0
    if (l <= pbits) {
      for (uint32_t i = rev; i < (uint32_t)psize; i += 1u << l) table[i] = ent | (uint32_t)l;
    } else {
      const uint32_t sub = table[rev & pmask];
      const int sbits = (int)((sub >> 8) & 31u);
      const uint32_t start = sub >> 16;
      for (uint32_t i = rev >> pbits; i < (1u << sbits); i += 1u << (l - pbits))
        table[start + i] = ent | (uint32_t)(l - pbits);
    }
  }
  return true;
}

struct Bits {
  const uint8_t* p;
  const uint8_t* end;
  uint64_t buf = 0;
  int cnt = 0;   // valid bits in buf; negative after reading past the end of the input
  inline void refill() {
    if (end - p >= 8) {
      buf |= load64(p) << cnt;
      p += (63 - cnt) >> 3;
      cnt |= 56;
    } else {
      while (cnt <= 56 && p < end) {
        buf |= (uint64_t)*p++ << cnt;
        cnt += 8;
      }
    }
  }
  inline uint32_t peek(int n) const { return (uint32_t)(buf & ((1ull << n) - 1ull)); }
  inline void drop(int n) {
    buf >>= n;
    cnt -= n;
  }
  inline uint32_t take(int n) {
    const uint32_t v = peek(n);
    drop(n);
    return v;
  }
};

static inline const Tables* fixed_tables() {
  static const Tables* t = [] {
    Tables* f = new Tables;
    uint8_t l[288];
    for (int i = 0; i < 144; ++i) l[i] = 8;
    for (int i = 144; i < 256; ++i) l[i] = 9;
    for (int i = 256; i < 280; ++i) l[i] = 7;
    for (int i = 280; i < 288; ++i) l[i] = 8;
    build_table(l, 288, LIT_BITS, 0, f->lit, LIT_CAP);
    uint8_t d[32];
    for (int i = 0; i < 32; ++i) d[i] = 5;
    build_table(d, 32, DIST_BITS, 1, f->dist, DIST_CAP);
    return f;
  }();
  return t;
}

// Dynamic block header -> tables.  False on a malformed header.
static inline bool read_dynamic(Bits& b, Tables* t) {
  static const uint8_t ORDER[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
  b.refill();
  const int hlit = (int)b.take(5) + 257, hdist = (int)b.take(5) + 1, hclen = (int)b.take(4) + 4;
  if (hlit > 286 || hdist > 30) return false;
  uint8_t pl[19] = {0};
  for (int i = 0; i < hclen; ++i) {
    if ((i & 7) == 0) b.refill();
    pl[ORDER[i]] = (uint8_t)b.take(3);
  }
  uint32_t pre[1 << PRE_BITS];
  if (!build_table(pl, 19, PRE_BITS, 2, pre, 1 << PRE_BITS)) return false;
  uint8_t lens[286 + 30];
  int n = 0;
  const int total = hlit + hdist;
  while (n < total) {
    b.refill();
    const uint32_t e = pre[b.peek(PRE_BITS)];
    if ((e & 31u) == 0u) return false;
    b.drop((int)(e & 31u));
    const int sym = (int)(e >> 16);
    if (sym < 16) {
      lens[n++] = (uint8_t)sym;
    } else {
      int rep;
      uint8_t v = 0;
      if (sym == 16) {
        if (n == 0) return false;
        v = lens[n - 1];
        rep = 3 + (int)b.take(2);
      } else if (sym == 17) {
        rep = 3 + (int)b.take(3);
      } else {
        rep = 11 + (int)b.take(7);
      }
      if (n + rep > total) return false;
      while (rep--) lens[n++] = v;
    }
    if (b.cnt < 0) return false;
  }
  if (lens[256] == 0) return false;   // no end-of-block code
  if (!build_table(lens, hlit, LIT_BITS, 0, t->lit, LIT_CAP)) return false;
  return build_table(lens + hlit, hdist, DIST_BITS, 1, t->dist, DIST_CAP);
}

// Inflate the whole raw-deflate stream in[0, in_len) into out[0, out_len); out must have 8 spare
// bytes behind out_len (match copies move whole words).  Returns 0 when the stream ends exactly
// at out_len bytes, a non-zero code otherwise.
static inline int inflate_all(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len,
                              Tables* scratch) {
  Bits b;
  b.p = in;
  b.end = in + in_len;
  uint8_t* o = out;
  uint8_t* const o_end = out + out_len;
  const uint32_t lmask = (1u << LIT_BITS) - 1u, dmask = (1u << DIST_BITS) - 1u;
  for (;;) {
    b.refill();
    const uint32_t bfinal = b.take(1), type = b.take(2);
    if (b.cnt < 0) return 1;
    if (type == 0) {
      b.drop(b.cnt & 7);
      b.refill();
      if (b.cnt < 32) return 2;
      const uint32_t len = b.take(16), nlen = b.take(16);
      if (len != (~nlen & 0xffffu)) return 3;
      if ((size_t)(o_end - o) < len) return 4;
      uint32_t left = len;
      // the bit buffer holds whole bytes here; the read pointer is exactly cnt / 8 bytes ahead
      const uint8_t* src = b.p - (b.cnt >> 3);
      if ((size_t)(b.end - src) < left) return 5;
      memcpy(o, src, left);
      o += left;
      b.p = src + left;
      b.buf = 0;
      b.cnt = 0;
    } else if (type == 3) {
      return 6;
    } else {
      const Tables* t;
      if (type == 1) {
        t = fixed_tables();
      } else {
        if (!read_dynamic(b, scratch)) return 7;
        t = scratch;
      }
      const uint32_t* lt = t->lit;
      const uint32_t* dt = t->dist;
      for (;;) {
        b.refill();
        uint32_t e = lt[b.buf & lmask];
        if (e & F_SUB) {
          b.drop(LIT_BITS);
          e = lt[(e >> 16) + b.peek((int)((e >> 8) & 31u))];
        }
        if (e & F_LIT) {
          // runs of literals: up to three codes from one refill (15 + 11 + 11 bits)
          if (o_end - o < 3) {
            if (o >= o_end) return 8;
            b.drop((int)(e & 31u));
            *o++ = (uint8_t)(e >> 16);
            continue;
          }
          b.drop((int)(e & 31u));
          *o++ = (uint8_t)(e >> 16);
          e = lt[b.buf & lmask];
          if ((e & (F_LIT | F_SUB)) != F_LIT) continue;
          b.drop((int)(e & 31u));
          *o++ = (uint8_t)(e >> 16);
          e = lt[b.buf & lmask];
          if ((e & (F_LIT | F_SUB)) != F_LIT) continue;
          b.drop((int)(e & 31u));
          *o++ = (uint8_t)(e >> 16);
          continue;
        }
        if ((e & 31u) == 0u) return 9;
        b.drop((int)(e & 31u));
        if (e & F_EOB) break;
        const int eb = (int)((e >> 8) & 31u);
        size_t len = (e >> 16) + b.take(eb);
        uint32_t d = dt[b.buf & dmask];
        if (d & F_SUB) {
          b.drop(DIST_BITS);
          d = dt[(d >> 16) + b.peek((int)((d >> 8) & 31u))];
        }
        if ((d & 31u) == 0u) return 10;
        b.drop((int)(d & 31u));
        const int db = (int)((d >> 8) & 31u);
        const size_t dist = (d >> 16) + b.take(db);
        if (b.cnt < 0) return 11;
        if (dist > (size_t)(o - out) || len > (size_t)(o_end - o)) return 12;
        const uint8_t* s = o - dist;
        if (dist >= 8) {
          // whole words; may write up to 7 bytes past the match (inside the 8 spare bytes)
          uint8_t* w = o;
          o += len;
          do {
            store64(w, load64(s));
            w += 8;
            s += 8;
          } while (w < o);
        } else {
          while (len--) *o++ = *s++;
        }
      }
      if (b.cnt < 0) return 13;
    }
    if (bfinal) break;
  }
  return o == o_end ? 0 : 14;
}

}  // namespace finf
}  // namespace exb
