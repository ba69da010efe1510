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
//  * Output: bytes are stored to global memory one by one; long-distance LZ77 matches load up to 8 source bytes in one step and
//    store them in the next, so the load latency hides behind decode work.
//  * This one-phase decoder is inflate_variant 0-11; the default kernel is the two-phase one in inflate_tok.cuh, which reuses
//    BitReader, LaneTable and Huff from this file.
//
// The routine is __host__ __device__ so tests/emu can run the very same code on the CPU against
// zlib; the shipped library only ever instantiates it inside kernels.
#pragma once
#include <stdint.h>
#include <type_traits>

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

// Lane-interleaved table of 16-bit entries living in shared memory (LANES = 32) or a plain array (LANES = 1):
// entry idx of lane l sits at base[idx * LANES + l].  Two neighbouring lanes share a 4-byte bank, so a warp's 32 random
// look-ups cost at most two wavefronts, and the address is one multiply-add.
template <int LANES, class E = uint16_t>
struct LaneTable {
  E* base;  // start of the warp's table region, already offset by the lane
  IRF_HD void bind(E* region, int lane) { base = region + lane; }
  IRF_HD E get(uint32_t idx) const { return base[idx * LANES]; }
  IRF_HD void set(uint32_t idx, E v) const { base[idx * LANES] = v; }
};

struct BitReader {
  const uint64_t* wp;    // next 8-byte word to load
  const uint8_t* end;    // one past the last input byte
  uint64_t bb;           // bit buffer, consumed from bit 0
  uint64_t cur, nxt, nx2; // the 8 bytes being fed into bb and the 16 bytes after them: every load is issued four refills
                         // (about four decode steps) before its first use, enough for a DRAM-latency miss to land
  int bc;                // valid bits in bb
  int half;              // 32-bit halves of `cur` already consumed (0..2)
  IRF_HD void init(const uint8_t* p, uint32_t len) {
    uintptr_t a = (uintptr_t)p; int mis = (int)(a & 7);
    wp = (const uint64_t*)(a - mis); end = p + len;
    uint64_t first = *wp++; cur = *wp++; nxt = *wp++; nx2 = *wp++; half = 0;
    bb = first >> (8 * mis); bc = 64 - 8 * mis;
  }
  IRF_HD void refill() {   // afterwards bc >= 33
    if (bc <= 32) {
      uint32_t w = half ? (uint32_t)(cur >> 32) : (uint32_t)cur;
      bb |= (uint64_t)w << bc; bc += 32;
      if (half) { cur = nxt; nxt = nx2; nx2 = *wp++; half = 0; } else half = 1;
    }
  }
  IRF_HD uint32_t peek(int n) const { return (uint32_t)bb & ((1u << n) - 1u); }
  IRF_HD void drop(int n) { bb >>= n; bc -= n; }
  IRF_HD uint32_t take(int n) { uint32_t v = peek(n); drop(n); return v; }
  // bytes consumed beyond the end of the input: the words held in cur/nxt/bb are look-ahead, not consumption
  IRF_HD bool overrun() const { return (const uint8_t*)wp - 16 - (half ? 4 : 8) - (bc >> 3) > end + 4; }
};

// Output cursor: bytes are stored to global memory one by one.  (Round 1 also had a mode that kept each lane's last 2^k output bytes in a
// shared-memory ring and flushed whole sectors; it lost more to occupancy than it won and its flush faulted on some output alignments, so it
// was removed in round 2.  The two-phase kernel in inflate_tok.cuh solves the same problem - partial-sector writes - by writing every output
// byte in its resolve phase.)
template <int LANES>
struct OutCursor {
  uint8_t* out;      // start of this block's output
  uint32_t op;       // bytes produced so far
  uint32_t cap;      // ISIZE
  uintptr_t pz;      // output sectors below this address have been pre-written (see prezero)
  IRF_HD void init(uint8_t* o, uint32_t c) { out = o; op = 0; cap = c; pz = ((uintptr_t)o + 31) & ~(uintptr_t)31; }
  IRF_HD void emit(uint32_t byte) { out[op] = (uint8_t)byte; op++; }
  // A 1-byte store that misses in L2 makes L2 fetch the whole 32-byte sector from DRAM first.  Writing whole sectors of zeros a
  // little ahead of the byte stores creates those sectors in L2 without a fetch, so the byte stores that follow hit.  Only sectors
  // that lie wholly inside this block's output are touched (the first and last partial sectors are shared with the neighbouring
  // blocks, i.e. with other lanes).
  IRF_HD void prezero(uint32_t lead) {
    uintptr_t want = (uintptr_t)out + op + lead, last = ((uintptr_t)out + cap) & ~(uintptr_t)31;
    if (want > last) want = last;
    while (pz + 32 <= want) {
#ifdef __CUDA_ARCH__
      uint4* d = (uint4*)pz; d[0] = make_uint4(0, 0, 0, 0); d[1] = make_uint4(0, 0, 0, 0);
#else
      for (int j = 0; j < 8; j++) ((uint32_t*)pz)[j] = 0;
#endif
      pz += 32;
    }
  }
};

// Per-lane Huffman set: a BITS-wide LUT plus, for the rare longer codes, the canonical-code limits.
// aux[l] (l = 1..15) = upper bound of the length-l codes, left-justified to 15 bits; aux[16 + l] = index of the
// first length-l symbol in `symbol` minus the first length-l code.  `symbol` (ordered by length, value) stays in
// per-lane local memory: it is touched once per long code.
// E = uint16_t: LUT entry = sym << 4 | len, 0 => long code or invalid.  E = uint8_t (alphabets of at most 31 symbols and
// BITS <= 8: the distance and code-length codes): entry = sym << 3 | (len - 1), 0xFF => long code or invalid; halves the
// table's shared memory.
// CAUX = 1: compact aux layout, only the entries a decoder with a BITS-wide LUT can ask for: [limit of length BITS+1 .. 15][offset of length
// BITS+1 .. 15], 2 * (15 - BITS) entries per lane instead of 32 (the two-phase kernel fits seven CTAs per SM that way).
template <int LANES, int BITS, int MAXSYM, class E = uint16_t, int CAUX = 0>
struct Huff {
  static constexpr int AUXN = 15 - BITS;
  static IRF_HD int ai_lim(int l) { return CAUX ? l - (BITS + 1) : l; }
  static IRF_HD int ai_off(int l) { return CAUX ? AUXN + l - (BITS + 1) : 16 + l; }
  static_assert(sizeof(E) == 2 || (MAXSYM <= 32 && BITS <= 7), "byte entries hold 5 symbol bits and 3 length bits; 0xFF is reserved");
  static IRF_HD E enc(int sym, int len) { return sizeof(E) == 2 ? (E)(sym << 4 | len) : (E)(sym << 3 | (len - 1)); }
  static IRF_HD E none() { return sizeof(E) == 2 ? (E)0 : (E)0xFF; }
  static IRF_HD int len_of(E e) { return sizeof(E) == 2 ? (int)(e & 15) : (e == (E)0xFF ? 0 : (int)(e & 7) + 1); }
  static IRF_HD int sym_of(E e) { return sizeof(E) == 2 ? (int)(e >> 4) : (int)(e >> 3); }
  LaneTable<LANES, E> lut;       // 2^BITS entries per lane
  LaneTable<LANES> aux;          // 32 entries per lane
  uint16_t* symbol;              // MAXSYM entries in the lane's local memory (kept outside this struct so its scalars stay in registers)

  // lens[0..n): code lengths.  Returns false on an over-subscribed set, or an incomplete one
  // (allowed only for a single-code set, as in zlib).
  IRF_HD bool build(const uint8_t* lens, int n) {
    uint16_t count[16];
    for (int i = 0; i < 16; i++) count[i] = 0;
    for (int i = 0; i < n; i++) count[lens[i]]++;
    for (uint32_t i = 0; i < (1u << BITS); i++) lut.set(i, none());
    for (int i = 0; i < (CAUX ? 2 * AUXN : 32); i++) aux.set(i, 0);
    if (count[0] == n) return true;  // no codes: any use is an error later
    int left = 1;
    for (int l = 1; l < 16; l++) { left <<= 1; left -= count[l]; if (left < 0) return false; }
    if (left > 0 && (n - count[0]) != 1) return false;
    uint16_t offs[16], next_code[16]; offs[1] = 0;
    for (int l = 1; l < 15; l++) offs[l + 1] = offs[l] + count[l];
    { uint32_t code = 0; for (int l = 1; l < 16; l++) { code = (code + (l > 1 ? count[l - 1] : 0)) << 1; next_code[l] = (uint16_t)code;
        if (!CAUX || l > BITS) { aux.set(ai_lim(l), (uint16_t)((code + count[l]) << (15 - l) > 0xffffu ? 0xffffu : (code + count[l]) << (15 - l))); aux.set(ai_off(l), (uint16_t)(offs[l] - code)); } } }
    for (int s = 0; s < n; s++) {
      int l = lens[s]; if (!l) continue;
      symbol[offs[l]++] = (uint16_t)s;
      uint32_t code = next_code[l]++;
      if (l <= BITS) {
#ifdef __CUDA_ARCH__
        uint32_t rev = __brev(code) >> (32 - l);
#else
        uint32_t rev = 0; for (int b = 0; b < l; b++) rev |= ((code >> b) & 1u) << (l - 1 - b);
#endif
        E e = enc(s, l);
        for (uint32_t i = rev; i < (1u << BITS); i += (1u << l)) lut.set(i, e);
      }
    }
    return true;
  }
  // index into `symbol` of the long code at the head of the bit buffer (drops its bits); -1 on an invalid code
  IRF_HD int slow_index(BitReader& br, uint32_t code15) const {
    uint32_t lim[16];
#pragma unroll
    for (int l = BITS + 1; l < 16; l++) lim[l] = aux.get(ai_lim(l));
    int len = 0;
#pragma unroll
    for (int l = 15; l > BITS; l--) if (code15 < lim[l]) len = l;
    if (!len) return -1;
    uint32_t idx = (uint32_t)(uint16_t)((code15 >> (15 - len)) + aux.get(ai_off(len)));
    if (idx >= (uint32_t)MAXSYM) return -1;
    br.drop(len);
    return (int)idx;
  }
  IRF_HD int slow15(BitReader& br, uint32_t code15) const { int idx = slow_index(br, code15); return idx < 0 ? -1 : symbol[idx]; }
  static IRF_HD uint32_t head15(const BitReader& br) {   // the next 15 bits, first bit most significant
#ifdef __CUDA_ARCH__
    return __brev((uint32_t)br.bb) >> 17;
#else
    uint32_t x = (uint32_t)br.bb & 0x7fffu, c = 0;
    for (int b = 0; b < 15; b++) c |= ((x >> b) & 1u) << (14 - b);
    return c;
#endif
  }
  IRF_HD int decode(BitReader& br) const {
    E e = lut.get(br.peek(BITS));
    int l = len_of(e);
    if (l) { br.drop(l); return sym_of(e); }
    return slow15(br, head15(br));
  }
};

// One lane's decoder as a state machine.  A GPU warp advances all 32 lanes one step() at a time under a
// warp-wide vote, so the lanes re-converge every iteration instead of drifting into 32 serialised
// instruction streams (the first version of this kernel did exactly that and ran ~50x slower).
// Per step a lane decodes one symbol and/or copies up to 8 bytes of a pending LZ77 match; block headers
// (Huffman table builds) are a third, rarer state.
enum LaneState : int { LS_HEADER = 0, LS_SYMBOL = 1, LS_COPY = 2, LS_FINISH = 3, LS_DONE = 4, LS_SLOWLIT = 5, LS_SLOWDIST = 6 };

// DIST8: the distance LUT (and the code-length LUT that borrows its storage) uses byte entries (see Huff).
template <int LANES, int LIT_BITS, int DIST_BITS, int PREZERO = 0, int DIST8 = 0>
struct InflateLane {
  using DistEntry = typename std::conditional<DIST8 != 0, uint8_t, uint16_t>::type;
  static constexpr size_t DIST_LUT_U16 = DIST8 ? (size_t)(1u << DIST_BITS) * LANES / 2 + ((size_t)(1u << DIST_BITS) * LANES & 1) : (size_t)(1u << DIST_BITS) * LANES;   // the distance LUT's size in 16-bit units
  BitReader br; OutCursor<LANES> oc; uint32_t tick;
  Huff<LANES, LIT_BITS, 288> hl; Huff<LANES, DIST_BITS, 32, DistEntry> hd;
  int state, status, bfinal; uint32_t copy_len, copy_dist, isize;
  uint32_t pend_n; uint32_t pend[8];  // bytes of a match chunk loaded in the previous step, one register each so nothing has to
                                      // wait for a load just to pack it (untouched until stored in this step)
  uint16_t slow_sym;   // long-code symbol fetched from local memory in the previous step

  // sym_store: 320 uint16 of lane-private scratch (288 literal/length symbols + 32 distance symbols)
  IRF_HD void start(const uint8_t* cdata, uint32_t clen, uint8_t* out, uint32_t isize_, uint16_t* lit_lut_base, uint16_t* dist_lut_base, int lane, uint16_t* sym_store) {
    hl.symbol = sym_store; hd.symbol = sym_store + 288; tick = 0;
    br.init(cdata, clen); oc.init(out, isize_); isize = isize_; if (PREZERO) oc.prezero(256);
    hl.lut.bind(lit_lut_base, lane); hd.lut.bind((DistEntry*)dist_lut_base, lane);
    // canonical-limit tables sit behind the two LUTs: [lit LUT][dist LUT][lit aux (32)][dist aux (32)], all lane-interleaved
    hd.aux.bind(dist_lut_base + DIST_LUT_U16, lane); hl.aux.bind(dist_lut_base + DIST_LUT_U16 + (size_t)32 * LANES, lane);
    state = LS_HEADER; status = INF_OK; bfinal = 0; copy_len = copy_dist = 0; pend_n = 0; slow_sym = 0;
  }
  IRF_HD void idle() { state = LS_DONE; status = INF_OK; pend_n = 0; tick = 0; oc.init(nullptr, 0); }   // the driver loop keeps stepping idle lanes: their cursor must be inert
  IRF_HD bool done() const { return state == LS_DONE; }
  IRF_HD void fail(int code) { status = code; state = LS_DONE; }

  IRF_HD void header() {
    uint8_t lens[320];
    br.refill();
    bfinal = (int)br.take(1); int btype = (int)br.take(2);
    if (btype == 0) {  // stored block: rare in BAM, copied here in one go
      br.drop(br.bc & 7); br.refill();
      uint32_t len = br.take(16); br.refill(); uint32_t nlen = br.take(16);
      if ((len ^ nlen) != 0xffffu) return fail(INF_ERR_STORED);
      if (oc.op + len > oc.cap) return fail(INF_ERR_OVERRUN_OUT);
      for (uint32_t i = 0; i < len; i++) { br.refill(); if (PREZERO && (i & 31u) == 0) oc.prezero(64); oc.emit(br.take(8)); if (br.overrun()) return fail(INF_ERR_OVERRUN_IN); }
      state = bfinal ? LS_FINISH : LS_HEADER; return;
    }
    if (btype == 3) return fail(INF_ERR_BTYPE);
    if (btype == 1) {
      int i = 0;
      for (; i < 144; i++) lens[i] = 8; for (; i < 256; i++) lens[i] = 9; for (; i < 280; i++) lens[i] = 7; for (; i < 288; i++) lens[i] = 8;
      if (!hl.build(lens, 288)) return fail(INF_ERR_CODELEN);
      for (i = 0; i < 32; i++) lens[i] = 5;  // 30 and 31 complete the code but never occur
      if (!hd.build(lens, 32)) return fail(INF_ERR_CODELEN);
    } else {
      br.refill();
      int hlit = (int)br.take(5) + 257, hdist = (int)br.take(5) + 1, hclen = (int)br.take(4) + 4;
      if (hlit > 286 || hdist > 30) return fail(INF_ERR_CODELEN);
      const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
      for (int i = 0; i < 19; i++) lens[i] = 0;
      for (int i = 0; i < hclen; i++) { br.refill(); lens[order[i]] = (uint8_t)br.take(3); }
      // the code-length code (<= 7 bits) borrows the distance LUT region
      Huff<LANES, (DIST_BITS < 7 ? DIST_BITS : 7), 19, DistEntry> hc; hc.symbol = hd.symbol; hc.lut.base = hd.lut.base; hc.aux.base = hd.aux.base;
      if (!hc.build(lens, 19)) return fail(INF_ERR_CODELEN);
      int idx = 0;
      while (idx < hlit + hdist) {
        br.refill();
        int sym = hc.decode(br);
        if (sym < 0) return fail(INF_ERR_CODELEN);
        if (sym < 16) lens[idx++] = (uint8_t)sym;
        else {
          int rep, val = 0;
          if (sym == 16) { if (idx == 0) return fail(INF_ERR_CODELEN); val = lens[idx - 1]; rep = 3 + (int)br.take(2); }
          else if (sym == 17) rep = 3 + (int)br.take(3);
          else rep = 11 + (int)br.take(7);
          if (idx + rep > hlit + hdist) return fail(INF_ERR_CODELEN);
          while (rep--) lens[idx++] = (uint8_t)val;
        }
        if (br.overrun()) return fail(INF_ERR_OVERRUN_IN);
      }
      if (lens[256] == 0) return fail(INF_ERR_CODELEN);
      if (!hl.build(lens, hlit)) return fail(INF_ERR_CODELEN);
      if (!hd.build(lens + hlit, hdist)) return fail(INF_ERR_CODELEN);
    }
    state = LS_SYMBOL;
  }

  // A literal/length symbol is in hand (from the LUT, or a long code resolved in the previous step).
  IRF_HD void handle_litlen(int sym) {
    if (sym < 256) {
      if (oc.op >= oc.cap) return fail(INF_ERR_OVERRUN_OUT);
      oc.emit((uint32_t)sym);
      return;
    }
    if (sym == 256) { if (br.overrun()) return fail(INF_ERR_OVERRUN_IN); state = bfinal ? LS_FINISH : LS_HEADER; return; }
    if (sym > 285) return fail(INF_ERR_SYMBOL);
    uint32_t len;
    if (sym < 265) len = (uint32_t)sym - 254u;
    else if (sym == 285) len = 258;
    else { int xb = (sym - 261) >> 2; len = ((4u + ((uint32_t)(sym - 261) & 3u)) << xb) + 3u + br.take(xb); }
    copy_len = len;
    br.refill();
    DistEntry e = hd.lut.get(br.peek(DIST_BITS)); int l = hd.len_of(e);
    if (l) { br.drop(l); return handle_dist(hd.sym_of(e)); }
    int idx = hd.slow_index(br, hd.head15(br));
    if (idx < 0) return fail(INF_ERR_SYMBOL);
    slow_sym = hd.symbol[idx]; state = LS_SLOWDIST;   // local-memory load: consumed in the next step
  }
  IRF_HD void handle_dist(int ds) {
    if (ds > 29) return fail(INF_ERR_SYMBOL);
    uint32_t dist;
    if (ds < 4) dist = (uint32_t)ds + 1u;
    else { int xb = (ds >> 1) - 1; dist = ((2u + ((uint32_t)ds & 1u)) << xb) + 1u + br.take(xb); }
    if (dist > oc.op) return fail(INF_ERR_DIST);
    if (oc.op + copy_len > oc.cap) return fail(INF_ERR_OVERRUN_OUT);
    if (br.overrun()) return fail(INF_ERR_OVERRUN_IN);
    copy_dist = dist; state = LS_COPY;
  }
  IRF_HD void symbol() {
    br.refill();  // >= 33 bits: enough for three LUT-resolved literals, or one length code + extra bits
    uint16_t e = 0; int l = 0, nlit = 0;
#pragma unroll
    for (int r = 0; r < 3; r++) {
      e = hl.lut.get(br.peek(LIT_BITS)); l = e & 15;
      if (l && (e >> 4) < 256 && oc.op < oc.cap) { br.drop(l); oc.emit((uint32_t)(e >> 4)); nlit++; } else break;
    }
    if (nlit) return;
    if (l) { br.drop(l); return handle_litlen(e >> 4); }
    int idx = hl.slow_index(br, hl.head15(br));
    if (idx < 0) return fail(INF_ERR_SYMBOL);
    slow_sym = hl.symbol[idx]; state = LS_SLOWLIT;      // local-memory load: consumed in the next step
  }

  // LZ77 copy, software-pipelined across steps: copy_load() issues the loads of up to 8 source bytes into registers,
  // copy_store() at the top of the NEXT step stores them, so the L2 latency of the loads hides behind the other lanes'
  // (and this lane's) decode work instead of stalling the warp.  For dist < 8 the source index wraps inside the period,
  // which makes overlapping (run-length) matches correct without reloading.
  IRF_HD void copy_store() {   // bytes fetched from global memory in the previous step
    uint8_t* dst = oc.out + oc.op;
#pragma unroll
    for (int j = 0; j < 8; j++) if ((uint32_t)j < pend_n) dst[j] = (uint8_t)pend[j];
    oc.op += pend_n;
    pend_n = 0;
    if (!copy_len) state = LS_SYMBOL;
  }
  IRF_HD void copy_load() {   // consumed by copy_store() in the next step
    uint32_t k = copy_len < 8 ? copy_len : 8, w = 0;
    const uint8_t* src = oc.out + oc.op - copy_dist;
#pragma unroll
    for (int j = 0; j < 8; j++) {
      if ((uint32_t)j < k) pend[j] = src[w];
      w++; if (w == copy_dist) w = 0;
    }
    pend_n = k; copy_len -= k;
  }

  // allow_header: the GPU driver loop batches header work (a table build costs thousands of instructions, and in lock-step
  // every lane pays for it), so a lane waiting in LS_HEADER only proceeds when enough lanes wait with it.
  IRF_HD void step(bool allow_header = true) {
    if (pend_n) copy_store();
    if (state == LS_SLOWLIT) { state = LS_SYMBOL; handle_litlen(slow_sym); }
    else if (state == LS_SLOWDIST) handle_dist(slow_sym);
    if (state == LS_HEADER && allow_header) header();
    if (state == LS_SYMBOL) symbol();
    if (state == LS_COPY) copy_load();
    if (state == LS_FINISH) { status = oc.op == isize ? INF_OK : INF_ERR_ISIZE; state = LS_DONE; }
    if (PREZERO && (++tick & 7u) == 0) oc.prezero(256);   // every lane, every 8th step
  }
};

// Decodes one raw DEFLATE stream of `clen` bytes at `cdata` into exactly `isize` bytes at `out` (scalar driver,
// used by the host emulation in tests/emu).  Both buffers need 8 readable bytes of slack after their ends.
template <int LANES, int LIT_BITS, int DIST_BITS, int PREZERO = 0, int DIST8 = 0>
IRF_HD int inflate_block(const uint8_t* cdata, uint32_t clen, uint8_t* out, uint32_t isize,
                         uint16_t* lit_lut_base, uint16_t* dist_lut_base, int lane) {
  InflateLane<LANES, LIT_BITS, DIST_BITS, PREZERO, DIST8> L; uint16_t sym_store[320];
  L.start(cdata, clen, out, isize, lit_lut_base, dist_lut_base, lane, sym_store);
  while (!L.done()) L.step();
  return L.status;
}

}  // namespace irf
