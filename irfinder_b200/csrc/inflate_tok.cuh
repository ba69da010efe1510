// inflate_tok.cuh — two-phase DEFLATE for BGZF blocks: a lane decodes the Huffman stream of one block into a literal stream and a
// list of LZ77 match tokens (phase A); afterwards the whole warp writes the output of each of its 32 blocks, literals and matches
// (phase B, inflate_tok.cu).  Replaces `gzip -cd` of the reference's pipeline; format authority RFC 1951.
//
// Why two phases (measured on the one-phase kernel of round 1, profiles/README.md): with the LZ77 copies inside the decode
// loop a warp's lanes sat in different states (table build / symbol / copy) and only ~11 of 32 were active per instruction; each
// copy was a dependent load from a 64 KiB-per-lane working set that missed L2 (13.6 DRAM bytes per algorithmic byte); every
// output byte was one store instruction.  Here the decode loop does ONE thing per iteration in every lane — one literal/length
// symbol (plus its distance symbol) — so lanes stay converged and, because zlib closes a DEFLATE block after a fixed number of
// symbols, reach their next table build in nearly the same iteration; matches become 4-byte tokens that the warp resolves later
// with all 32 lanes working on ONE block (64 KiB working set per warp, not per lane).
//
// Token (32 bits):  lit_len << 24 | match_len << 15 | (dist - 1)
//   lit_len   0..255  literal bytes that precede the match (taken from the literal stream in order)
//   match_len 0, 3..258; 0 = no match: the token only carries literals (a run longer than 255, or the tail of the block)
//   dist      1..32768
// Literal stream: the literal bytes of a block in order, packed four to a word.  Tokens grow from the bottom of the lane's scratch
// region, literal words from its top (word j of the stream at index TOK_CAP - 1 - j): a block never needs more than isize/3 + 2
// words for both (a literal costs a quarter of a word, a match at least three output bytes per word).
// Phase A therefore writes nothing to the output: every output byte is stored by the resolve phase, whole blocks at a time, so
// sectors reach DRAM once and complete (the first version stored literals in place during decode; the match bytes filled in later
// re-opened sectors that had long left L2: 5.1 DRAM bytes per algorithmic byte).
//
// __host__ __device__ like inflate_core.cuh: tests/emu runs this very routine on the CPU against zlib.
#pragma once
#include "inflate_core.cuh"

namespace irf {

constexpr uint32_t TOK_CAP = 65536 / 3 + 65536 / 255 + 8;   // tokens one BGZF block can need (worst case): size of a lane's scratch region
IRF_HD uint32_t tok_make(uint32_t lit, uint32_t mlen, uint32_t dist) { return lit << 24 | mlen << 15 | (dist - 1u); }
IRF_HD uint32_t tok_lit(uint32_t t) { return t >> 24; }
IRF_HD uint32_t tok_mlen(uint32_t t) { return (t >> 15) & 511u; }
IRF_HD uint32_t tok_dist(uint32_t t) { return (t & 32767u) + 1u; }

// Decode states of a lane (LS_* of inflate_core.cuh are reused): HEADER (a table build is due), SYMBOL (decoding), FINISH, DONE.
//
// Written for SIMT execution: iter() is ONE straight body that every decoding lane of a warp runs together — refill, one
// literal/length look-up, then either up to two literals or one match (length + distance + token).  The first version of this
// routine kept the round-1 layout (separate states for codes longer than the LUT, a second inlined copy of the symbol handler,
// an error branch per check) and measured 9 of 32 lanes active per instruction: every one of those branches splits the warp and
// the split paths run one after the other.  Here a long code is resolved on the spot, errors are collected in one flag that is
// looked at once per vote, and the literal counter of a token is a subtraction at match time instead of an increment per literal.
// FUSE = 1: an iteration is [up to two literals] + [one match]: a match whose code follows the literals in the LUT is taken in the same
// iteration (FUSE = 0: an iteration ends after its literals).  On BAM data written by zlib level 1, 60 % of the iterations are matches and
// most literal runs are one or two bytes long, so the lanes of a warp walk through the same literal-literal-match sequence together
// instead of spending an iteration of their own on the literals: 23 % fewer iterations on the benchmark data.
template <int LANES, int LIT_BITS, int DIST_BITS, int FUSE = (LIT_BITS <= 6 ? 1 : 0)>   // (a wider LUT leaves too few bits for the fused sequence)
struct TokLane {
  BitReader br;
  // shared-memory tables of a lane: literal/length LUT (16-bit entries), distance LUT (byte entries), compact canonical-limit tables
  using HL = Huff<LANES, LIT_BITS, 288, uint16_t, 1>; using HD = Huff<LANES, DIST_BITS, 32, uint8_t, 1>;
  static constexpr size_t DIST_LUT_U16 = ((size_t)(1u << DIST_BITS) * LANES + 1) / 2;          // the distance LUT's size in 16-bit units
  static constexpr size_t TABLE_U16 = (size_t)(1u << LIT_BITS) * LANES + DIST_LUT_U16 + (size_t)(2 * HD::AUXN + 2 * HL::AUXN) * LANES;   // all tables of LANES lanes
  HL hl; HD hd;
  uint32_t op, cap;            // output bytes accounted for so far (literals + matches), ISIZE
  uint32_t acc, sh;            // literal bytes of the current word of the literal stream; bit position of the next byte in it
  uint32_t* tok; uint32_t ntok, nlw, last_end;   // tokens written, literal words written (from the top), output offset where the previous match ended
  int state, status, bfinal; uint32_t bad;   // bad: an InflateStatus noticed inside iter(); turned into a failure by check()

  // sym_store: 320 uint16 of lane-private scratch (288 literal/length symbols + 32 distance symbols); tok_region: TOK_CAP words
  IRF_HD void start(const uint8_t* cdata, uint32_t clen, uint32_t isize, uint16_t* lit_lut_base, uint16_t* dist_lut_base, int lane, uint16_t* sym_store, uint32_t* tok_region) {
    hl.symbol = sym_store; hd.symbol = sym_store + 288;
    br.init(cdata, clen); op = 0; cap = isize; acc = 0; sh = 0;
    tok = tok_region; ntok = 0; nlw = 0; last_end = 0;
    hl.lut.bind(lit_lut_base, lane); hd.lut.bind((uint8_t*)dist_lut_base, lane);
    // canonical-limit tables sit behind the two LUTs: [lit LUT][dist LUT (bytes)][dist aux][lit aux], all lane-interleaved
    hd.aux.bind(dist_lut_base + DIST_LUT_U16, lane); hl.aux.bind(dist_lut_base + DIST_LUT_U16 + (size_t)(2 * HD::AUXN) * LANES, lane);
    state = LS_HEADER; status = INF_OK; bfinal = 0; bad = 0;
  }
  IRF_HD void idle() { state = LS_DONE; status = INF_OK; ntok = 0; nlw = 0; op = 0; cap = 0; bad = 0; }
  IRF_HD bool done() const { return state == LS_DONE; }
  IRF_HD void fail(int code) { status = code; state = LS_DONE; }

  IRF_HD void put_literal(uint32_t byte) {   // caller has checked op < cap
    acc |= byte << sh; sh += 8; op++;
    if (sh == 32) { tok[TOK_CAP - 1 - nlw] = acc; nlw++; acc = 0; sh = 0; }
  }
  IRF_HD void put_token(uint32_t len, uint32_t dist) {   // the literals since the previous match, then the match (len = 0: literals only)
    uint32_t lit = op - last_end;
    while (lit > 255) { tok[ntok++] = tok_make(255, 0, 1); lit -= 255; }   // a literal run longer than a token can say (rare)
    if (lit | len) tok[ntok++] = tok_make(lit, len, dist);
    op += len; last_end = op;
  }
  IRF_HD void finish() {
    put_token(0, 1);                                  // the literals behind the last match
    if (sh) { tok[TOK_CAP - 1 - nlw] = acc; nlw++; }
    status = op == cap ? INF_OK : INF_ERR_ISIZE; state = LS_DONE;
  }

  IRF_HD void header() {
    uint8_t lens[320];
    br.refill();
    bfinal = (int)br.take(1); int btype = (int)br.take(2);
    if (btype == 0) {  // stored block: rare in BAM; its bytes are literals
      br.drop(br.bc & 7); br.refill();
      uint32_t len = br.take(16); br.refill(); uint32_t nlen = br.take(16);
      if ((len ^ nlen) != 0xffffu) return fail(INF_ERR_STORED);
      if (op + len > cap) return fail(INF_ERR_OVERRUN_OUT);
      for (uint32_t i = 0; i < len; i++) { br.refill(); put_literal(br.take(8)); if (br.overrun()) return fail(INF_ERR_OVERRUN_IN); }
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
      static_assert(DIST_BITS <= 7, "the code-length code borrows the distance LUT (byte entries: at most 7 bits)");
      Huff<LANES, DIST_BITS, 19, uint8_t, 1> hc; hc.symbol = hd.symbol; hc.lut.base = hd.lut.base; hc.aux.base = hd.aux.base;
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

  // One decode iteration of a lane in LS_SYMBOL: a literal (and a second one when the next code is a LUT-resolved literal too),
  // or a match, or the end of the DEFLATE block.  Errors only set `bad`; nothing is stored outside the block whatever the input.
  IRF_HD void iter() {
    br.refill();  // >= 33 bits: two codes, or a 15-bit length code + 5 extra bits
    uint32_t e = hl.lut.get(br.peek(LIT_BITS)); uint32_t l = e & 15u, sym = e >> 4;
    if (l) br.drop((int)l);
    else { int idx = hl.slow_index(br, hl.head15(br)); if (idx < 0) { bad = INF_ERR_SYMBOL; return; } sym = hl.symbol[idx]; }   // code longer than the LUT: canonical limits + the lane's symbol table
    if (sym < 256) {
      if (op >= cap) { bad = INF_ERR_OVERRUN_OUT; return; }
      put_literal(sym);
      e = hl.lut.get(br.peek(LIT_BITS)); l = e & 15u;
      if (!FUSE) {
        if (l && e < (256u << 4) && op < cap) { br.drop((int)l); put_literal(e >> 4); }
        return;
      }
      // bits: the first code took <= 15, the LUT codes that follow <= LIT_BITS each, the length's extra bits <= 5: <= 33 since the refill
      static_assert(!FUSE || 15 + 2 * LIT_BITS + 5 <= 33, "the fused iteration reads two more LUT codes before it refills");
      if (!l) return;                                    // a long code is next: it opens the next iteration
      sym = e >> 4;
      if (sym < 256) {
        if (op >= cap) return;                           // (the next iteration reports the overrun)
        br.drop((int)l); put_literal(sym);
        e = hl.lut.get(br.peek(LIT_BITS)); l = e & 15u; sym = e >> 4;
        if (!l || sym < 256) return;                     // a long code or a third literal: next iteration
      }
      br.drop((int)l);                                   // a match (or the end of the block) straight behind the literals
    }
    if (sym == 256) { state = bfinal ? LS_FINISH : LS_HEADER; return; }
    // ---- match: length, then the distance code
    const uint32_t s = sym - 257u;                       // 0..28 for valid symbols
    const uint32_t xb = (s < 8u || s >= 28u) ? 0u : (s - 4u) >> 2;
    uint32_t len = s < 8u ? s + 3u : (s >= 28u ? 258u : ((4u + (s & 3u)) << xb) + 3u + br.take((int)xb));
    br.refill();
    const uint8_t ed = hd.lut.get(br.peek(DIST_BITS)); const uint32_t ld = (uint32_t)HD::len_of(ed); uint32_t ds = (uint32_t)HD::sym_of(ed);
    if (ld) br.drop((int)ld);
    else { int idx = hd.slow_index(br, hd.head15(br)); if (idx < 0) { bad = INF_ERR_SYMBOL; return; } ds = hd.symbol[idx]; }
    const uint32_t xd = ds < 4u ? 0u : (ds >> 1) - 1u;
    const uint32_t dist = ds < 4u ? ds + 1u : ((2u + (ds & 1u)) << xd) + 1u + br.take((int)(xd & 15u));
    if ((s > 28u) | (ds > 29u)) { bad = INF_ERR_SYMBOL; return; }
    if (dist > op) { bad = INF_ERR_DIST; return; }
    if (op + len > cap) { bad = INF_ERR_OVERRUN_OUT; return; }
    put_token(len, dist);
  }
  // once per vote of the driver loop: turn a noticed error into a failure, catch a bit reader that ran past its input
  IRF_HD void check() {
    if (state == LS_DONE) return;
    if (bad) return fail((int)bad);
    if (br.overrun()) return fail(INF_ERR_OVERRUN_IN);
    if (state == LS_FINISH) finish();
  }
  // scalar driver step (host emulation): one table build or one decode iteration
  IRF_HD void step() {
    if (state == LS_HEADER) header(); else if (state == LS_SYMBOL) iter();
    check();
  }
};

// Scalar resolve of one block (the reference semantics of the warp-wide resolve in inflate_tok.cu; used by tests/emu): literals from the
// stream at the top of the region, matches from the output itself.
IRF_HD bool tok_resolve_scalar(uint8_t* out, uint32_t isize, const uint32_t* region, uint32_t ntok) {
  uint32_t pos = 0, q = 0;
  for (uint32_t i = 0; i < ntok; i++) {
    uint32_t t = region[i], lit = tok_lit(t), ml = tok_mlen(t), d = tok_dist(t);
    if (pos + lit + ml > isize || (ml && d > pos + lit)) return false;
    for (uint32_t k = 0; k < lit; k++, q++) out[pos + k] = (uint8_t)(region[TOK_CAP - 1 - (q >> 2)] >> (8 * (q & 3)));
    pos += lit;
    for (uint32_t k = 0; k < ml; k++) out[pos + k] = out[pos + k - d];
    pos += ml;
  }
  return pos == isize;
}
template <int LANES, int LIT_BITS, int DIST_BITS, int FUSE = (LIT_BITS <= 6 ? 1 : 0)>
IRF_HD int inflate_block_tok(const uint8_t* cdata, uint32_t clen, uint8_t* out, uint32_t isize, uint16_t* lit_lut_base, uint16_t* dist_lut_base, int lane, uint32_t* tok_region, uint32_t* ntok_out = nullptr) {
  TokLane<LANES, LIT_BITS, DIST_BITS, FUSE> L; uint16_t sym_store[320];
  L.start(cdata, clen, isize, lit_lut_base, dist_lut_base, lane, sym_store, tok_region);
  while (!L.done()) L.step();
  if (ntok_out) *ntok_out = L.ntok + L.nlw;
  if (L.status) return L.status;
  if (L.ntok + L.nlw > TOK_CAP) return INF_ERR_OVERRUN_OUT;
  return tok_resolve_scalar(out, isize, tok_region, L.ntok) ? INF_OK : INF_ERR_DIST;
}

}  // namespace irf
