// inflate_core.cuh — RFC 1951 DEFLATE decoder for one BGZF block, written as a scalar routine that
// one GPU lane runs from start to finish (32 independent streams per warp).
//
// Replaces, for the BAM->IR path, the external `gzip -cd` process that feeds the reference's
// irfinder binary (SURVEY.md §1 "Below L1", Appendix A.1).  Format authority: RFC 1951 and SAM v1 §4.1.
//
// Design notes (see DESIGN.md "S1 inflate"):
//  * Huffman decode is a serial bit chain, so the unit of parallelism is the BGZF block: one lane per
//    block, thousands of blocks in flight.
//  * Decode tables are single-level LUTs of LIT_BITS / DIST_BITS bits held in shared memory and
//    interleaved by lane (word index * LANES + lane) so a warp's 32 random lookups never bank-conflict.
//    Codes longer than the LUT width (rare) fall back to a canonical bit-serial walk over the
//    code-length histogram, held in per-lane local memory.
//  * Output bytes go through a 4-byte accumulator so literal runs leave as aligned 32-bit stores;
//    LZ77 copies load up to 8 source bytes before storing any, so their latencies overlap.
//
// The routine is __host__ __device__ so tests/emu can run the very same code on the CPU against
// zlib; the shipped library only ever instantiates it inside kernels.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define IRF_HD __host__ __device__ __forceinline__
#else
#define IRF_HD inline
#endif

namespace irf {

enum InflateStatus : int {
  INF_OK = 0,
  INF_ERR_BTYPE = 1,       // reserved block type 3
  INF_ERR_STORED = 2,      // LEN != ~NLEN
  INF_ERR_CODELEN = 3,     // bad code-length sequence / over-subscribed set
  INF_ERR_SYMBOL = 4,      // invalid literal/length or distance code
  INF_ERR_DIST = 5,        // distance reaches before the start of the block
  INF_ERR_OVERRUN_OUT = 6, // more output than ISIZE
  INF_ERR_OVERRUN_IN = 7,  // ran past the end of CDATA
  INF_ERR_ISIZE = 8,       // stream ended with output length != ISIZE
};

// Lane-interleaved table of 16-bit entries living in shared memory (LANES = 32) or a plain array (LANES = 1).
template <int LANES>
struct LaneTable {
  uint16_t* base;  // start of the warp's table region
  int lane;
  IRF_HD uint16_t get(uint32_t idx) const { return base[(((idx >> 1) * LANES + lane) << 1) | (idx & 1)]; }
  IRF_HD void set(uint32_t idx, uint16_t v) const { base[(((idx >> 1) * LANES + lane) << 1) | (idx & 1)] = v; }
};

struct BitReader {
  const uint32_t* wp;
  const uint32_t* wend;  // first word that lies wholly beyond the input
  uint64_t bb;
  int bc;
  IRF_HD void init(const uint8_t* p, uint32_t len) {
    uintptr_t a = (uintptr_t)p; int mis = (int)(a & 3);
    wp = (const uint32_t*)(a - mis); wend = (const uint32_t*)((a + len + 3) & ~(uintptr_t)3);
    bb = (uint64_t)(*wp++) >> (8 * mis); bc = 32 - 8 * mis;
  }
  IRF_HD void refill() { if (bc <= 32) { bb |= (uint64_t)(*wp++) << bc; bc += 32; } }  // afterwards bc >= 33
  IRF_HD uint32_t peek(int n) const { return (uint32_t)bb & ((1u << n) - 1u); }
  IRF_HD void drop(int n) { bb >>= n; bc -= n; }
  IRF_HD uint32_t take(int n) { uint32_t v = peek(n); drop(n); return v; }
  IRF_HD bool overrun() const { return wp > wend + 2; }  // two words of look-ahead are legitimate
};

// Output cursor with a 4-byte accumulator.
struct OutCursor {
  uint8_t* out;      // start of this block's output
  uint32_t op;       // bytes produced so far (including those still in acc)
  uint32_t cap;      // ISIZE
  uint32_t acc; int nacc;
  IRF_HD void init(uint8_t* o, uint32_t c) { out = o; op = 0; cap = c; acc = 0; nacc = 0; }
  IRF_HD void flush_bytes() {
    uint8_t* p = out + op - nacc;
    for (int i = 0; i < nacc; i++) p[i] = (uint8_t)(acc >> (8 * i));
    acc = 0; nacc = 0;
  }
  IRF_HD void emit(uint32_t byte) {
    acc |= byte << (8 * nacc); nacc++; op++;
    if ((((uintptr_t)out + op) & 3) == 0) {
      if (nacc == 4) { *(uint32_t*)(out + op - 4) = acc; acc = 0; nacc = 0; }
      else flush_bytes();
    }
  }
};

// Per-lane Huffman set: LUT in LaneTable + canonical fallback arrays in local memory.
template <int LANES, int BITS, int MAXSYM>
struct Huff {
  LaneTable<LANES> lut;          // entry = sym << 4 | len (len 1..BITS); 0 => long code or invalid
  uint16_t count[16];            // codes per length
  uint16_t symbol[MAXSYM];       // symbols ordered by (length, value)
  bool has_long;

  // lens[0..n): code lengths.  Returns false on an over-subscribed set, or an incomplete one
  // (allowed only for a single-code set, as in zlib).
  IRF_HD bool build(const uint8_t* lens, int n) {
    for (int i = 0; i < 16; i++) count[i] = 0;
    for (int i = 0; i < n; i++) count[lens[i]]++;
    for (uint32_t i = 0; i < (1u << BITS); i++) lut.set(i, 0);
    if (count[0] == n) { has_long = false; return true; }  // no codes: any use is an error later
    int left = 1;
    for (int l = 1; l < 16; l++) { left <<= 1; left -= count[l]; if (left < 0) return false; }
    if (left > 0 && (n - count[0]) != 1) return false;
    uint16_t offs[16]; offs[1] = 0;
    for (int l = 1; l < 15; l++) offs[l + 1] = offs[l] + count[l];
    uint16_t next_code[16];
    { uint32_t code = 0; for (int l = 1; l < 16; l++) { code = (code + (l > 1 ? count[l - 1] : 0)) << 1; next_code[l] = (uint16_t)code; } }
    has_long = false;
    for (int s = 0; s < n; s++) {
      int l = lens[s]; if (!l) continue;
      symbol[offs[l]++] = (uint16_t)s;
      uint32_t code = next_code[l]++;
      if (l <= BITS) {
        uint32_t rev = 0; for (int b = 0; b < l; b++) rev |= ((code >> b) & 1u) << (l - 1 - b);
        uint16_t e = (uint16_t)(s << 4 | l);
        for (uint32_t i = rev; i < (1u << BITS); i += (1u << l)) lut.set(i, e);
      } else has_long = true;
    }
    return true;
  }
  // canonical bit-serial decode (used when the LUT entry is empty); returns -1 on an invalid code
  IRF_HD int slow(BitReader& br) const {
    int code = 0, first = 0, index = 0;
    for (int l = 1; l < 16; l++) {
      code |= (int)((br.bb >> (l - 1)) & 1u);
      int c = count[l];
      if (code - c < first) { br.drop(l); return symbol[index + (code - first)]; }
      index += c; first += c; first <<= 1; code <<= 1;
    }
    return -1;
  }
  IRF_HD int decode(BitReader& br) const {
    uint16_t e = lut.get(br.peek(BITS));
    int l = e & 15;
    if (l) { br.drop(l); return e >> 4; }
    return slow(br);
  }
};

// Decodes one raw DEFLATE stream of `clen` bytes at `cdata` into exactly `isize` bytes at `out`.
// Both buffers need 8 readable bytes of slack after their ends (the caller pads its allocations).
template <int LANES, int LIT_BITS, int DIST_BITS>
IRF_HD int inflate_block(const uint8_t* cdata, uint32_t clen, uint8_t* out, uint32_t isize,
                         uint16_t* lit_lut_base, uint16_t* dist_lut_base, int lane) {
  BitReader br; br.init(cdata, clen);
  OutCursor oc; oc.init(out, isize);
  Huff<LANES, LIT_BITS, 288> hl; hl.lut.base = lit_lut_base; hl.lut.lane = lane;
  Huff<LANES, DIST_BITS, 32> hd; hd.lut.base = dist_lut_base; hd.lut.lane = lane;
  uint8_t lens[320];
  int bfinal;
  do {
    br.refill();
    bfinal = (int)br.take(1); int btype = (int)br.take(2);
    if (btype == 0) {
      br.drop(br.bc & 7); br.refill();
      uint32_t len = br.take(16); br.refill(); uint32_t nlen = br.take(16);
      if ((len ^ nlen) != 0xffffu) return INF_ERR_STORED;
      if (oc.op + len > oc.cap) return INF_ERR_OVERRUN_OUT;
      for (uint32_t i = 0; i < len; i++) { br.refill(); oc.emit(br.take(8)); if (br.overrun()) return INF_ERR_OVERRUN_IN; }
      continue;
    }
    if (btype == 3) return INF_ERR_BTYPE;
    if (btype == 1) {
      int i = 0;
      for (; i < 144; i++) lens[i] = 8; for (; i < 256; i++) lens[i] = 9; for (; i < 280; i++) lens[i] = 7; for (; i < 288; i++) lens[i] = 8;
      if (!hl.build(lens, 288)) return INF_ERR_CODELEN;
      for (i = 0; i < 32; i++) lens[i] = 5;  // 30 and 31 complete the code but never occur
      if (!hd.build(lens, 32)) return INF_ERR_CODELEN;
    } else {
      br.refill();
      int hlit = (int)br.take(5) + 257, hdist = (int)br.take(5) + 1, hclen = (int)br.take(4) + 4;
      if (hlit > 286 || hdist > 30) return INF_ERR_CODELEN;
      const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
      for (int i = 0; i < 19; i++) lens[i] = 0;
      for (int i = 0; i < hclen; i++) { br.refill(); lens[order[i]] = (uint8_t)br.take(3); }
      // the code-length code (<= 7 bits) borrows the distance LUT when it is wide enough, else decodes bit-serially
      Huff<LANES, (DIST_BITS < 7 ? DIST_BITS : 7), 19> hc; hc.lut.base = dist_lut_base; hc.lut.lane = lane;
      if (!hc.build(lens, 19)) return INF_ERR_CODELEN;
      int idx = 0;
      while (idx < hlit + hdist) {
        br.refill();
        int sym = hc.decode(br);
        if (sym < 0) return INF_ERR_CODELEN;
        if (sym < 16) lens[idx++] = (uint8_t)sym;
        else {
          int rep, val = 0;
          if (sym == 16) { if (idx == 0) return INF_ERR_CODELEN; val = lens[idx - 1]; rep = 3 + (int)br.take(2); }
          else if (sym == 17) rep = 3 + (int)br.take(3);
          else rep = 11 + (int)br.take(7);
          if (idx + rep > hlit + hdist) return INF_ERR_CODELEN;
          while (rep--) lens[idx++] = (uint8_t)val;
        }
        if (br.overrun()) return INF_ERR_OVERRUN_IN;
      }
      if (lens[256] == 0) return INF_ERR_CODELEN;
      if (!hl.build(lens, hlit)) return INF_ERR_CODELEN;
      if (!hd.build(lens + hlit, hdist)) return INF_ERR_CODELEN;
    }
    // ---- symbol loop
    for (;;) {
      br.refill();
      int sym = hl.decode(br);
      if (sym < 256) {
        if (sym < 0) return INF_ERR_SYMBOL;
        if (oc.op >= oc.cap) return INF_ERR_OVERRUN_OUT;
        oc.emit((uint32_t)sym);
        continue;
      }
      if (sym == 256) break;
      if (sym > 285) return INF_ERR_SYMBOL;
      uint32_t len;
      if (sym < 265) len = (uint32_t)sym - 254u;
      else if (sym == 285) len = 258;
      else { int xb = (sym - 261) >> 2; len = ((4u + ((uint32_t)(sym - 261) & 3u)) << xb) + 3u + br.take(xb); }
      br.refill();
      int ds = hd.decode(br);
      if (ds < 0 || ds > 29) return INF_ERR_SYMBOL;
      uint32_t dist;
      if (ds < 4) dist = (uint32_t)ds + 1u;
      else { int xb = (ds >> 1) - 1; dist = ((2u + ((uint32_t)ds & 1u)) << xb) + 1u + br.take(xb); }
      if (dist > oc.op) return INF_ERR_DIST;
      if (oc.op + len > oc.cap) return INF_ERR_OVERRUN_OUT;
      if (br.overrun()) return INF_ERR_OVERRUN_IN;
      // ---- LZ77 copy
      if (dist >= 8) {
        while (len) {
          oc.flush_bytes();
          const uint8_t* src = oc.out + oc.op - dist;
          uint32_t k = len < 8 ? len : 8;
          uint8_t t[8];
#pragma unroll
          for (int j = 0; j < 8; j++) t[j] = (uint32_t)j < k ? src[j] : (uint8_t)0;
#pragma unroll
          for (int j = 0; j < 8; j++) if ((uint32_t)j < k) oc.emit(t[j]);
          len -= k;
        }
      } else {
        oc.flush_bytes();
        const uint8_t* src = oc.out + oc.op - dist;
        uint64_t pat = 0;
#pragma unroll
        for (int j = 0; j < 7; j++) if ((uint32_t)j < dist) pat |= (uint64_t)src[j] << (8 * j);
        uint32_t ph = 0;
        for (uint32_t i = 0; i < len; i++) { oc.emit((uint32_t)(pat >> (8 * ph)) & 255u); ph++; if (ph == dist) ph = 0; }
      }
    }
    if (br.overrun()) return INF_ERR_OVERRUN_IN;
  } while (!bfinal);
  oc.flush_bytes();
  return oc.op == isize ? INF_OK : INF_ERR_ISIZE;
}

}  // namespace irf
