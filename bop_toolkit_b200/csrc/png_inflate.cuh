// RFC 1950/1951 (zlib / DEFLATE) decoder core of the depth-PNG input path.
//
// Replaces the libpng + zlib decode behind `inout.load_depth` (/root/reference
// bop_toolkit_lib/inout.py:39-46, `imageio.imread` of the 16-bit depth PNGs read by
// scripts/eval_calc_errors.py:236-243).  One DEFLATE stream is serial by construction (symbol
// boundaries are only known after decoding), so the parallelism is ACROSS images: on the device
// one warp owns one image, lane 0 runs `inflate_run` below with its Huffman tables in shared
// memory and hands long LZ77 matches to the whole warp (EV_COPY).  The same source compiles for
// the host (`bop_png_decode_selftest_host`), which is how the CPU test-suite checks the decoder
// bit for bit against zlib without a GPU.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define BOP_HD __host__ __device__ __forceinline__
#define BOP_HD_NOINLINE __host__ __device__ __noinline__
#else
#define BOP_HD inline
#define BOP_HD_NOINLINE
#endif

namespace bop_png {

constexpr int kLitBits = 10;    // primary lookup width of the literal/length code
constexpr int kDistBits = 8;    // primary lookup width of the distance code
constexpr uint32_t kCoopMin = 16;  // matches at least this long are copied by the whole warp
constexpr uint32_t kRingWords = 1024;   // device input ring per stream (4 KB of shared memory)
constexpr uint32_t kRingRefill = 512;   // words loaded per EV_INPUT
constexpr uint32_t kRingMargin = 256;   // the reader asks for input this far before the ring's end:
                                        // more than one dynamic block header (<= 160 words) needs

enum {
  EV_DONE = 0,           // final block finished
  EV_COPY = 1,           // caller copies m_len bytes from distance m_dist, then adds m_len to o
  EV_INPUT = 2,          // caller loads more of the stream into the ring and raises br.limit
  EV_BAD_HEADER = -1,    // not a zlib stream (CM != 8, preset dictionary, bad check bits)
  EV_BAD_BLOCK = -2,     // reserved block type / stored LEN != ~NLEN
  EV_BAD_CODES = -3,     // over-subscribed or otherwise invalid Huffman code description
  EV_BAD_SYMBOL = -4,    // undecodable code, length/distance symbol out of range
  EV_BAD_DISTANCE = -5,  // distance reaches before the start of the output
  EV_OVERFLOW = -6,      // more output than the image holds
  EV_TRUNCATED = -7,     // input ended inside the stream
  EV_SHORT = -8,         // stream ended before the image was complete (set by the caller)
  EV_BAD_FILTER = -9,    // PNG row filter type above 4 (set by the unfilter pass)
  EV_BAD_CHECKSUM = -10  // Adler-32 of the inflated bytes differs from the stream's trailer
};

// Per-stream decode tables: shared memory on the device (3.7 KB per warp).
struct Tables {
  uint16_t lit_fast[1 << kLitBits];    // (symbol << 4) | code length, 0 = longer than kLitBits
  uint16_t dist_fast[1 << kDistBits];
  uint16_t lit_sym[288], dist_sym[32], len_sym[20];   // symbols in canonical order
  uint16_t lit_count[16], dist_count[16], len_count[16];
  uint8_t lengths[320];
};

struct BitReader {
  // Word w of the 4-byte aligned, zero padded stream lives at in[w & mask]: the whole stream on
  // the host (mask = ~0), a shared-memory ring on the device that the warp refills whenever the
  // reader passes `limit` (EV_INPUT).
  const uint32_t* in;
  uint32_t mask, limit;
  uint32_t nwords, ipos;
  uint32_t nextw;       // word `ipos`, loaded one refill ahead so its latency is hidden
  uint64_t buf;
  int cnt;
  BOP_HD void init(const uint32_t* p, uint32_t n, uint32_t ring_mask, uint32_t lim) {
    in = p; nwords = n; mask = ring_mask; limit = lim; ipos = 0; buf = 0; cnt = 0;
    nextw = n > 0 ? p[0] : 0u;
  }
  BOP_HD void refill() {   // afterwards cnt > 32
    if (cnt <= 32) {
      buf |= (uint64_t)nextw << cnt;
      cnt += 32;
      ++ipos;
      nextw = ipos < nwords ? in[ipos & mask] : 0u;
    }
  }
  BOP_HD bool starved() const { return ipos > limit; }
  BOP_HD uint32_t peek(int n) const { return (uint32_t)buf & ((1u << n) - 1u); }
  BOP_HD void drop(int n) { buf >>= n; cnt -= n; }
  BOP_HD uint32_t bits(int n) { uint32_t v = peek(n); drop(n); return v; }
  // words past the end read as zero; two of them can be buffered before a bit of them is used
  BOP_HD bool overrun() const { return ipos > nwords + 2u; }
};

struct State {
  BitReader br;
  uint8_t* out;
  uint32_t o, cap;
  int in_block;     // 1 while inside a Huffman-coded block
  int final_block;
  uint32_t stored_left;   // bytes of a stored block still to copy
  uint32_t m_len, m_dist;
  uint32_t adler;         // the stream's Adler-32 trailer (RFC 1950), valid after EV_DONE
};

// Canonical-code decode one bit at a time (zlib's contrib/puff method): used for the code-length
// code and for literal/length or distance codes longer than the primary table.
BOP_HD int decode_slow(BitReader& br, const uint16_t* count, const uint16_t* sym, int maxlen) {
  int code = 0, first = 0, index = 0;
  uint64_t b = br.buf;
  for (int len = 1; len <= maxlen; ++len) {
    code |= (int)(b & 1u);
    b >>= 1;
    const int c = count[len];
    if (code - c < first) {
      br.buf = b;
      br.cnt -= len;
      return sym[index + (code - first)];
    }
    index += c;
    first += c;
    first <<= 1;
    code <<= 1;
  }
  return -1;
}

// Builds count[] / sym[] (canonical order) from code lengths and, if `fast`, the primary lookup
// table.  Returns <0 for an over-subscribed set, 0 for a complete one, >0 for an incomplete one.
BOP_HD_NOINLINE int construct(uint16_t* count, uint16_t* sym, const uint8_t* length, int n,
                              uint16_t* fast, int fastbits) {
  uint16_t offs[16];
  for (int l = 0; l < 16; ++l) count[l] = 0;
  for (int s = 0; s < n; ++s) count[length[s]]++;
  int left = 1;
  for (int l = 1; l < 16; ++l) {
    left <<= 1;
    left -= count[l];
    if (left < 0) return -1;
  }
  offs[1] = 0;
  for (int l = 1; l < 15; ++l) offs[l + 1] = (uint16_t)(offs[l] + count[l]);
  for (int s = 0; s < n; ++s)
    if (length[s]) sym[offs[length[s]]++] = (uint16_t)s;
  if (fast) {
    const int size = 1 << fastbits;
    for (int i = 0; i < size; ++i) fast[i] = 0;
    uint32_t code = 0;
    int idx = 0;
    for (int l = 1; l <= fastbits; ++l) {
      for (int k = 0; k < count[l]; ++k) {
        uint32_t rev = 0;
        for (int b = 0; b < l; ++b) rev |= ((code >> b) & 1u) << (l - 1 - b);
        const uint16_t e = (uint16_t)((sym[idx++] << 4) | l);
        for (int j = (int)rev; j < size; j += 1 << l) fast[j] = e;
        ++code;
      }
      code <<= 1;
    }
  }
  return left;
}

BOP_HD_NOINLINE int build_fixed(Tables& t) {
  int s = 0;
  for (; s < 144; ++s) t.lengths[s] = 8;
  for (; s < 256; ++s) t.lengths[s] = 9;
  for (; s < 280; ++s) t.lengths[s] = 7;
  for (; s < 288; ++s) t.lengths[s] = 8;
  construct(t.lit_count, t.lit_sym, t.lengths, 288, t.lit_fast, kLitBits);
  for (s = 0; s < 30; ++s) t.lengths[s] = 5;
  construct(t.dist_count, t.dist_sym, t.lengths, 30, t.dist_fast, kDistBits);
  return 0;
}

BOP_HD_NOINLINE int build_dynamic(BitReader& br, Tables& t) {
  br.refill();
  const int nlen = (int)br.bits(5) + 257;
  const int ndist = (int)br.bits(5) + 1;
  const int ncode = (int)br.bits(4) + 4;
  if (nlen > 286 || ndist > 30) return EV_BAD_CODES;
  // order of the code-length code lengths (RFC 1951 3.2.7), 5 bits each
  const uint64_t ord_a = 16ull | 17ull << 5 | 18ull << 10 | 0ull << 15 | 8ull << 20 | 7ull << 25 |
                         9ull << 30 | 6ull << 35 | 10ull << 40 | 5ull << 45 | 11ull << 50 | 4ull << 55;
  const uint64_t ord_b = 12ull | 3ull << 5 | 13ull << 10 | 2ull << 15 | 14ull << 20 | 1ull << 25 |
                         15ull << 30;
  for (int i = 0; i < 19; ++i) t.lengths[i] = 0;
  for (int i = 0; i < ncode; ++i) {
    br.refill();
    const int pos = i < 12 ? (int)((ord_a >> (5 * i)) & 31u) : (int)((ord_b >> (5 * (i - 12))) & 31u);
    t.lengths[pos] = (uint8_t)br.bits(3);
  }
  if (construct(t.len_count, t.len_sym, t.lengths, 19, nullptr, 0) < 0) return EV_BAD_CODES;
  int index = 0;
  while (index < nlen + ndist) {
    br.refill();
    if (br.overrun()) return EV_TRUNCATED;
    int s = decode_slow(br, t.len_count, t.len_sym, 7);
    if (s < 0) return EV_BAD_CODES;
    if (s < 16) {
      t.lengths[index++] = (uint8_t)s;
    } else {
      int prev = 0, rep;
      if (s == 16) {
        if (index == 0) return EV_BAD_CODES;
        prev = t.lengths[index - 1];
        rep = 3 + (int)br.bits(2);
      } else if (s == 17) {
        rep = 3 + (int)br.bits(3);
      } else {
        rep = 11 + (int)br.bits(7);
      }
      if (index + rep > nlen + ndist) return EV_BAD_CODES;
      while (rep--) t.lengths[index++] = (uint8_t)prev;
    }
  }
  if (t.lengths[256] == 0) return EV_BAD_CODES;   // no end-of-block code
  int left = construct(t.lit_count, t.lit_sym, t.lengths, nlen, t.lit_fast, kLitBits);
  if (left < 0 || (left > 0 && nlen - t.lit_count[0] != 1)) return EV_BAD_CODES;
  left = construct(t.dist_count, t.dist_sym, t.lengths + nlen, ndist, t.dist_fast, kDistBits);
  if (left < 0 || (left > 0 && ndist - t.dist_count[0] != 1)) return EV_BAD_CODES;
  return 0;
}

// zlib header (RFC 1950): CM = 8, window <= 32 KB, no preset dictionary, FCHECK.
BOP_HD int inflate_begin(State& s, const uint32_t* in, uint32_t nwords, uint32_t ring_mask,
                         uint32_t limit, uint8_t* out, uint32_t cap) {
  s.br.init(in, nwords, ring_mask, limit);
  s.stored_left = 0;
  s.out = out;
  s.o = 0;
  s.cap = cap;
  s.in_block = 0;
  s.final_block = 0;
  s.m_len = s.m_dist = 0;
  s.adler = 0;
  s.br.refill();
  const uint32_t cmf = s.br.bits(8), flg = s.br.bits(8);
  if ((cmf & 15u) != 8u || (cmf >> 4) > 7u || (flg & 32u) || ((cmf << 8) | flg) % 31u) return EV_BAD_HEADER;
  return 0;
}

// Decodes until the stream ends (EV_DONE), a long match needs the caller (EV_COPY) or an error.
// Matches shorter than kCoopMin are copied here.
BOP_HD_NOINLINE int inflate_run(State& s, Tables& t) {
  BitReader br = s.br;
  uint8_t* const out = s.out;
#ifdef __CUDA_ARCH__
  // the device caller keeps the input ring and the tables in shared memory and writes to global
  // memory: lets the compiler use LDS / STG instead of generic accesses
  __builtin_assume(__isShared(br.in));
  __builtin_assume(__isShared(&t));
  __builtin_assume(__isGlobal(out));
#endif
  uint32_t o = s.o;
  const uint32_t cap = s.cap;
  int ev = EV_DONE;
  for (;;) {
    if (s.stored_left) {
      uint32_t left = s.stored_left;
      while (left && !br.starved()) {
        br.refill();
        out[o++] = (uint8_t)br.bits(8);
        --left;
      }
      s.stored_left = left;
      if (left) { ev = EV_INPUT; break; }
      if (br.overrun()) { ev = EV_TRUNCATED; break; }
      continue;
    }
    if (!s.in_block) {
      if (s.final_block) {
        // the trailer: four bytes, big-endian, on the next byte boundary
        if (br.starved()) { ev = EV_INPUT; break; }
        br.drop(br.cnt & 7);
        uint32_t a = 0;
        for (int k = 0; k < 4; ++k) {
          br.refill();
          a = (a << 8) | br.bits(8);
        }
        s.adler = a;
        ev = br.overrun() ? (int)EV_TRUNCATED : (int)EV_DONE;
        break;
      }
      if (br.starved()) { ev = EV_INPUT; break; }
      br.refill();
      if (br.overrun()) { ev = EV_TRUNCATED; break; }
      s.final_block = (int)br.bits(1);
      const uint32_t type = br.bits(2);
      if (type == 0) {                 // stored
        br.drop(br.cnt & 7);
        br.refill();
        const uint32_t len = br.bits(16);
        br.refill();
        const uint32_t nlen = br.bits(16);
        if ((len ^ nlen) != 0xffffu) { ev = EV_BAD_BLOCK; break; }
        if (o + len > cap) { ev = EV_OVERFLOW; break; }
        s.stored_left = len;
        continue;
      } else if (type == 1) {
        build_fixed(t);
      } else if (type == 2) {
        const int rc = build_dynamic(br, t);
        if (rc) { ev = rc; break; }
      } else {
        ev = EV_BAD_BLOCK;
        break;
      }
      s.in_block = 1;
    }
    // ---- Huffman-coded symbols --------------------------------------------------------
    // One warp instruction per step of one lane: what counts here is the number of instructions
    // and taken branches per symbol, so the checks that only guard against corrupt input are
    // folded (one range test per match, input exhaustion looked at only when a word is loaded).
    for (;;) {
      // Runs of literals: one table look-up, one combined test (a literal whose code fits the
      // primary table has an entry in [1, 4095]; 0 = long code, >= 4096 = end of block / length),
      // everything else drops to the general step below.
      for (;;) {
        if (br.cnt <= 32) {
          if (br.starved()) break;
          br.refill();
        }
        const uint32_t e = t.lit_fast[br.peek(kLitBits)];
        if (e - 1u >= 4095u || o >= cap) break;
        br.drop((int)(e & 15u));
        out[o++] = (uint8_t)(e >> 4);
      }
      if (br.cnt <= 32) {
        if (br.starved()) { ev = EV_INPUT; goto done; }
        br.refill();
      }
      int sym;
      {
        const uint32_t e = t.lit_fast[br.peek(kLitBits)];
        if (e & 15u) { br.drop((int)(e & 15u)); sym = (int)(e >> 4); }
        else {
          sym = decode_slow(br, t.lit_count, t.lit_sym, 15);
          if (sym < 0) { ev = EV_BAD_SYMBOL; goto done; }
        }
      }
      if (sym < 256) {
        if (o >= cap) { ev = EV_OVERFLOW; goto done; }
        out[o++] = (uint8_t)sym;
        continue;
      }
      if (sym == 256) {
        s.in_block = 0;
        if (br.overrun()) { ev = EV_TRUNCATED; goto done; }
        break;
      }
      sym -= 257;
      if (sym >= 29) { ev = EV_BAD_SYMBOL; goto done; }
      uint32_t len = 3u + (uint32_t)sym;
      if (sym >= 8) {
        const int eb = (sym - 4) >> 2;
        len = sym == 28 ? 258u : ((4u + ((uint32_t)sym & 3u)) << eb) + 3u + br.bits(eb);
      }
      br.refill();
      uint32_t ds;
      {
        const uint32_t e = t.dist_fast[br.peek(kDistBits)];
        if (e & 15u) { br.drop((int)(e & 15u)); ds = e >> 4; }
        else ds = (uint32_t)decode_slow(br, t.dist_count, t.dist_sym, 15);   // -1 -> out of range
      }
      if (ds >= 30u) { ev = EV_BAD_SYMBOL; goto done; }
      uint32_t dist = 1u + ds;
      if (ds >= 4u) {
        const int eb = (int)(ds >> 1) - 1;
        dist = ((2u + (ds & 1u)) << eb) + 1u + br.bits(eb);
      }
      if (dist > o || o + len > cap) { ev = dist > o ? EV_BAD_DISTANCE : EV_OVERFLOW; goto done; }
      if (len >= kCoopMin) {
        s.m_len = len;
        s.m_dist = dist;
        ev = EV_COPY;
        goto done;
      }
      // 3 <= len < kCoopMin.  Without overlap every load is issued before the first store (one
      // memory round trip, not one per byte) and the two switches execute exactly len loads and
      // len stores.
      const uint8_t* src = out + o - dist;
      uint8_t* dst = out + o;
      o += len;
      if (dist >= len) {
        uint8_t b0, b1, b2, b3 = 0, b4 = 0, b5 = 0, b6 = 0, b7 = 0, b8 = 0, b9 = 0, b10 = 0, b11 = 0,
                b12 = 0, b13 = 0, b14 = 0;
        switch (len) {
          case 15: b14 = src[14];  // fallthrough
          case 14: b13 = src[13];  // fallthrough
          case 13: b12 = src[12];  // fallthrough
          case 12: b11 = src[11];  // fallthrough
          case 11: b10 = src[10];  // fallthrough
          case 10: b9 = src[9];    // fallthrough
          case 9: b8 = src[8];     // fallthrough
          case 8: b7 = src[7];     // fallthrough
          case 7: b6 = src[6];     // fallthrough
          case 6: b5 = src[5];     // fallthrough
          case 5: b4 = src[4];     // fallthrough
          case 4: b3 = src[3];     // fallthrough
          default: b2 = src[2]; b1 = src[1]; b0 = src[0];
        }
        switch (len) {
          case 15: dst[14] = b14;  // fallthrough
          case 14: dst[13] = b13;  // fallthrough
          case 13: dst[12] = b12;  // fallthrough
          case 12: dst[11] = b11;  // fallthrough
          case 11: dst[10] = b10;  // fallthrough
          case 10: dst[9] = b9;    // fallthrough
          case 9: dst[8] = b8;     // fallthrough
          case 8: dst[7] = b7;     // fallthrough
          case 7: dst[6] = b6;     // fallthrough
          case 6: dst[5] = b5;     // fallthrough
          case 5: dst[4] = b4;     // fallthrough
          case 4: dst[3] = b3;     // fallthrough
          default: dst[2] = b2; dst[1] = b1; dst[0] = b0;
        }
      } else if (dist == 1) {
        const uint8_t v = src[0];
        for (uint32_t k = 0; k < len; ++k) dst[k] = v;
      } else {
        for (uint32_t k = 0; k < len; ++k) dst[k] = src[k];
      }
    }
  }
done:
  s.br = br;
  s.o = o;
  return ev;
}

}  // namespace bop_png
