// RFC 1950/1951 (zlib / DEFLATE) decoder core of the depth-PNG input path.
//
// Replaces the libpng + zlib decode behind `inout.load_depth` (/root/reference
// bop_toolkit_lib/inout.py:39-46, `imageio.imread` of the 16-bit depth PNGs read by
// scripts/eval_calc_errors.py:236-243).  One DEFLATE stream is serial by construction (symbol
// boundaries are only known after decoding), so the parallelism is ACROSS images: on the device
// one warp owns one image.  The decode loop is WARP-REPLICATED: all 32 lanes run `inflate_run`
// below on identical state (the Huffman tables and the input ring are in shared memory, so a
// table look-up is one broadcast load for the warp and costs the issue slot it cost when a single
// lane decoded), lane 0 stores the literals, and every LZ77 match -- whatever its length -- is
// copied by all lanes at once: no hand-off between a decoding lane and a copying warp, and no
// byte-by-byte copies of the short matches that make up most of a depth PNG.  Tables are built by
// lane 0 and the bit reader is broadcast afterwards (the builders are read-modify-write code).
// The same source compiles for the host (`bop_png_decode_selftest_host`, lane 0 of 1), which is
// how the CPU test-suite checks the decoder bit for bit against zlib without a GPU.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define BOP_HD __host__ __device__ __forceinline__
#define BOP_HD_NOINLINE __host__ __device__ __noinline__
#else
#define BOP_HD inline
#define BOP_HD_NOINLINE
#endif

#ifdef __CUDA_ARCH__
#define BOP_PNG_WARP_SYNC() __syncwarp()
#else
#define BOP_PNG_WARP_SYNC() ((void)0)
#endif

namespace bop_png {

constexpr int kLitBits = 10;    // primary lookup width of the literal/length code
constexpr int kDistBits = 8;    // primary lookup width of the distance code
constexpr uint32_t kRingWords = 1024;   // device input ring per stream (4 KB of shared memory)
constexpr uint32_t kRingRefill = 512;   // words loaded per EV_INPUT
constexpr uint32_t kRingMargin = 256;   // the reader asks for input this far before the ring's end:
                                        // more than one dynamic block header (<= 160 words) needs

enum {
  EV_DONE = 0,           // final block finished
  EV_COPY = 1,           // (unused since the decode loop is warp-replicated: matches are copied inside inflate_run)
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

// Decodes until the stream ends (EV_DONE), more input is needed (EV_INPUT) or an error.
// Device: called by ALL lanes of the warp that owns the stream with identical `s` (lane = 0..31,
// nlanes = 32); every lane follows the same control flow.  Host: lane 0 of 1.
BOP_HD_NOINLINE int inflate_run(State& s, Tables& t, int lane = 0, int nlanes = 1) {
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
        const uint8_t v = (uint8_t)br.bits(8);
        if (lane == 0) out[o] = v;
        ++o;
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
      } else if (type == 1 || type == 2) {
        // lane 0 builds the tables (every lane is done with the previous block's), then hands
        // the bit reader -- advanced past a dynamic block's header -- to the other lanes
        int rc = 0;
        BOP_PNG_WARP_SYNC();
        if (lane == 0) rc = type == 1 ? build_fixed(t) : build_dynamic(br, t);
#ifdef __CUDA_ARCH__
        if (nlanes > 1) {
          rc = __shfl_sync(0xffffffffu, rc, 0);
          br.buf = __shfl_sync(0xffffffffu, (unsigned long long)br.buf, 0);
          br.cnt = __shfl_sync(0xffffffffu, br.cnt, 0);
          br.ipos = __shfl_sync(0xffffffffu, br.ipos, 0);
          br.nextw = __shfl_sync(0xffffffffu, br.nextw, 0);
        }
#endif
        BOP_PNG_WARP_SYNC();
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
      // Runs of literals: one table look-up and one combined test per literal (a literal whose code
      // fits the primary table has an entry in [1, 4095]; 0 = long code, >= 4096 = end of block /
      // length); everything else drops to the general step below.  After a refill at least 33 bits
      // are valid and such a literal takes at most kLitBits = 10 of them, so THREE are decoded per
      // refill check, and the output bound is checked once for the three.
      for (;;) {
        if (br.cnt <= 32) {
          if (br.starved()) break;
          br.refill();
        }
        if (o + 3u > cap) break;
        uint32_t e = t.lit_fast[br.peek(kLitBits)];
        if (e - 1u >= 4095u) break;
        br.drop((int)(e & 15u));
        if (lane == 0) out[o] = (uint8_t)(e >> 4);
        ++o;
        e = t.lit_fast[br.peek(kLitBits)];
        if (e - 1u >= 4095u) break;
        br.drop((int)(e & 15u));
        if (lane == 0) out[o] = (uint8_t)(e >> 4);
        ++o;
        e = t.lit_fast[br.peek(kLitBits)];
        if (e - 1u >= 4095u) break;
        br.drop((int)(e & 15u));
        if (lane == 0) out[o] = (uint8_t)(e >> 4);
        ++o;
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
        if (lane == 0) out[o] = (uint8_t)sym;
        ++o;
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
      // The match, copied by all lanes: every source byte lies in the window [o - dist, o) that was
      // complete before this match began (i % dist wraps an overlapping match onto it), so one
      // pass needs no ordering inside the match; the barrier orders it behind the stores that
      // filled the window (lane 0's literals, earlier matches).
      {
        BOP_PNG_WARP_SYNC();
        const uint8_t* src = out + o - dist;
        uint8_t* dst = out + o;
        o += len;
        if (dist >= len) {
          for (uint32_t i = (uint32_t)lane; i < len; i += (uint32_t)nlanes) dst[i] = src[i];
        } else if (dist == 1) {
          const uint8_t v = src[0];
          for (uint32_t i = (uint32_t)lane; i < len; i += (uint32_t)nlanes) dst[i] = v;
        } else {
          for (uint32_t i = (uint32_t)lane; i < len; i += (uint32_t)nlanes) dst[i] = src[i % dist];
        }
      }
    }
  }
done:
  s.br = br;
  s.o = o;
  return ev;
}

}  // namespace bop_png
