// inflate_tok.cu — the two-phase SM inflate kernel (see inflate_tok.cuh for the format and the reasons).
//
// One persistent launch, 148 SMs x resident CTAs; every warp claims 32 consecutive BGZF blocks from a global counter and runs
//   phase A  lane per block: Huffman decode into the lane's scratch region: one 4-byte token per LZ77 match (literal count, length,
//            distance) from the bottom, the literal bytes packed four to a word from the top; nothing is written to the output;
//   phase B  warp per block, for each of the 32 blocks in turn: tokens are read 32 at a time (coalesced), two warp scans turn their
//            lengths into output and literal-stream positions, the literals are placed, and the matches are copied — in parallel
//            when their source bytes are final, in order otherwise (a match may read what an earlier match of the same batch writes);
// then claims the next 32.  Replaces `gzip -cd` of the reference pipeline; RFC 1951.
#include <atomic>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include "common.cuh"
#include "inflate_tok.cuh"
#include "kernels.h"

namespace irf {

__device__ __forceinline__ const void* shfl_ptr(const void* p, int src) {
  unsigned long long v = (unsigned long long)p;
  unsigned lo = __shfl_sync(0xffffffffu, (unsigned)v, src), hi = __shfl_sync(0xffffffffu, (unsigned)(v >> 32), src);
  return (const void*)((unsigned long long)hi << 32 | lo);
}

// developer counters (IRF_TOK_PROF=1 prints them after a launch): cycles in phase A / phase B summed over warps, decode votes,
// table-build events, resolve batches and rounds
__device__ unsigned long long g_tok_prof[8];

// Phase B for one batch of up to 32 tokens of one block: lane i holds token i (ml = 0: nothing to copy), the output offsets
// [mstart, mend) of its match and the distance d.  Literals are final already, so a match may be copied as soon as no unfinished
// match of the batch overlaps its source bytes.  Holes are ordered like the tokens, so the matches that can overlap lane i's
// source [src, need) form a range of lanes [lo, hi) found by two binary searches over the warp (shuffles); a round then costs one
// ballot and one AND per lane, and the number of rounds is the depth of the dependency chains, not the number of dependent matches.
constexpr uint32_t TOK_LONG = 16;   // matches longer than this are copied by the whole warp (4 / 8 / 16 / 32 / 64 measured: 184 / 143 / 131 / 132 / 136 ms)
constexpr uint32_t LIT_LONG = 4;    // literal runs longer than this are placed by the whole warp
// lstart / lit / lq: output offset, length and literal-stream offset of the literal run in front of lane i's match; lits: the block's
// literal stream (word j at lits[-j], bytes in little-endian order)
__device__ __forceinline__ uint8_t lit_byte(const uint32_t* lits, uint32_t q) { return (uint8_t)(lits[-(int)(q >> 2)] >> (8 * (q & 3))); }
// lwin / wq: the 32 words of the literal stream from word wq on, one per lane, loaded a batch ahead (the short runs of a batch are
// served from this window by shuffles instead of dependent loads from the scratch region)
__device__ __forceinline__ void resolve_batch(uint8_t* ob, const uint32_t* lits, uint32_t lwin, uint32_t wq, uint32_t lstart, uint32_t lit, uint32_t lq, uint32_t mstart, uint32_t ml, uint32_t d, int lane, uint32_t& rounds) {
  // ---- literals first: they depend on nothing, and the matches of this batch may read them
  {
    const uint32_t w0 = (lq >> 2) - wq;                  // window word that holds the run's first byte (garbage for lanes without a short run)
    const uint32_t a = __shfl_sync(0xffffffffu, lwin, (int)(w0 & 31u)), b = __shfl_sync(0xffffffffu, lwin, (int)((w0 + 1) & 31u));
    if (lit && lit <= LIT_LONG) {
      uint32_t v;
      if (w0 < 31u) v = __funnelshift_r(a, b, (lq & 3u) * 8u);
      else { v = 0; for (uint32_t k = 0; k < lit; k++) v |= (uint32_t)lit_byte(lits, lq + k) << (8 * k); }   // run beyond the window (a batch with many literals)
      uint8_t* dp = ob + lstart;
      dp[0] = (uint8_t)v; if (lit > 1) dp[1] = (uint8_t)(v >> 8); if (lit > 2) dp[2] = (uint8_t)(v >> 16); if (lit > 3) dp[3] = (uint8_t)(v >> 24);
    }
  }
  unsigned ll = __ballot_sync(0xffffffffu, lit > LIT_LONG);
  while (ll) {                                           // long runs: 32 consecutive bytes per trip (warp-uniform loop)
    const int f = __ffs(ll) - 1; ll &= ll - 1;
    const uint32_t fs = __shfl_sync(0xffffffffu, lstart, f), fn = __shfl_sync(0xffffffffu, lit, f), fq = __shfl_sync(0xffffffffu, lq, f);
    for (uint32_t k = lane; k < fn; k += 32) ob[fs + k] = lit_byte(lits, fq + k);
  }
  __syncwarp();
  const uint32_t mend = mstart + ml, src = mstart - d, need = src + min(ml, d);   // source bytes [src, need) must be final before this match may be copied
  uint32_t lo = 0, hi = 0;   // lo = lanes whose hole ends at or before src; hi = lanes whose hole starts before need
#pragma unroll
  for (int step = 16; step; step >>= 1) {
    const uint32_t a = __shfl_sync(0xffffffffu, mend, (int)(lo + step - 1)), b = __shfl_sync(0xffffffffu, mstart, (int)(hi + step - 1));
    if (a <= src) lo += step;
    if (b < need) hi += step;
  }
  hi = min(hi, (uint32_t)lane);
  const uint32_t dep = hi > lo ? ((1u << hi) - 1u) & ~((1u << lo) - 1u) : 0u;   // earlier lanes of the batch whose output this match reads
  bool pending = ml != 0;
  for (;;) {
    const unsigned pm = __ballot_sync(0xffffffffu, pending);
    if (!pm) break;
    rounds++;
    const bool ready = pending && !(pm & dep);           // the first pending lane has no pending lane below it: every round makes progress
    unsigned rl = __ballot_sync(0xffffffffu, ready && ml > TOK_LONG);
    while (rl) {                                         // long matches: all 32 lanes copy one match (warp-uniform loop)
      const int f = __ffs(rl) - 1; rl &= rl - 1;
      const uint32_t fstart = __shfl_sync(0xffffffffu, mstart, f), fml = __shfl_sync(0xffffffffu, ml, f), fd = __shfl_sync(0xffffffffu, d, f);
      uint8_t* dp = ob + fstart; const uint8_t* sp = dp - fd;
      if (fd >= 32) {                                    // a 32-byte slab never reads what the same slab writes; earlier slabs are ordered by the barrier
        for (uint32_t k0 = 0; k0 < fml; k0 += 32) { if (k0 + lane < fml) dp[k0 + lane] = sp[k0 + lane]; __syncwarp(); }
      } else {                                           // period shorter than a slab: every byte is a copy of one of the fd final bytes before the match
        for (uint32_t k = lane; k < fml; k += 32) dp[k] = sp[k % fd];
      }
    }
    if (ready && ml <= TOK_LONG) {
      uint8_t* dp = ob + mstart; const uint8_t* sp = dp - d; uint32_t w = 0, rem = ml;
      do {                                               // 4 bytes per trip (most matches are 3 or 4 long): loads first, then stores; the source index
        const uint32_t k = rem < 4 ? rem : 4; uint32_t v[4];   // wraps inside the period, so run-length matches only ever read the d final bytes
#pragma unroll
        for (int j = 0; j < 4; j++) { if ((uint32_t)j < k) v[j] = sp[w]; w++; if (w == d) w = 0; }
#pragma unroll
        for (int j = 0; j < 4; j++) if ((uint32_t)j < k) dp[j] = (uint8_t)v[j];
        dp += k; rem -= k;
      } while (rem);
    }
    if (ready) pending = false;
    __syncwarp();   // orders this round's stores before the next round's loads (and before the next batch)
  }
}

template <int WARPS, int LB, int DB, int PATIENCE, int MINB, int FUSE>
__global__ void __launch_bounds__(WARPS * 32, MINB) inflate_tok_kernel(const uint8_t* __restrict__ comp, const u64* __restrict__ c_off, const u32* __restrict__ c_len,
                                                                  uint8_t* U, const u64* __restrict__ u_off, const u32* __restrict__ u_len, int nblk,
                                                                  int* __restrict__ status, int* __restrict__ err, unsigned int* __restrict__ next_block, const unsigned long long* watermark,
                                                                  const u32* __restrict__ list, uint32_t* tok_scratch) {
  extern __shared__ uint16_t smem_tables[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // per warp: [lit LUT, 16-bit entries][dist LUT, byte entries][canonical-limit entries of both], interleaved by lane
  constexpr size_t WARP_U16 = TokLane<32, LB, DB>::TABLE_U16;
  uint16_t* lit = smem_tables + (size_t)warp * WARP_U16;
  uint16_t* dst = lit + (1 << LB) * 32;
  uint16_t sym_store[320];
  uint32_t* my_tok = tok_scratch + ((size_t)(blockIdx.x * WARPS + warp) * 32 + lane) * TOK_CAP;
  for (;;) {
    unsigned int base = 0;
    if (lane == 0) base = atomicAdd(next_block, 32u);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= (unsigned int)nblk) break;
    if (watermark) {   // streaming input: wait until the host->device copy has passed the last byte these 32 blocks need
      int lastb = min((int)base + 31, nblk - 1); if (list) lastb = (int)list[lastb]; u64 need = c_off[lastb] + c_len[lastb] + 160; bool starved = false;
      if (lane == 0) {
        long long t0 = clock64();
        while (*(volatile const unsigned long long*)watermark < need) { __nanosleep(500); if (clock64() - t0 > 40000000000ll) { starved = true; break; } }   // ~20 s without data: give up instead of hanging the GPU
        __threadfence();
      }
      starved = __shfl_sync(0xffffffffu, (int)starved, 0) != 0;
      if (starved) { if (lane == 0) atomicOr(err, DERR_STARVED); break; }
    }
    int b = (int)base + lane; const bool in_range = b < nblk; if (list && in_range) b = (int)list[b];
    TokLane<32, LB, DB, FUSE> L;
    const bool work = in_range && !(u_len[b] == 0 && c_len[b] <= 2);  // skip the 28-byte EOF marker / empty blocks
    if (work) L.start(comp + c_off[b], c_len[b], u_len[b], lit, dst, lane, sym_store, my_tok); else L.idle();
    uint8_t* const my_out = work ? U + u_off[b] : nullptr;
    // ---- phase A.  One vote per four decode iterations.  A table build (thousands of instructions that every lane of the warp
    // pays for) starts only when no lane is left decoding, half the warp is waiting for one, or the waiting lanes have waited
    // PATIENCE votes; zlib closes a DEFLATE block after a fixed number of symbols, so lanes mostly arrive together anyway.
    int waited = 0; const long long tA = clock64(); uint32_t votes = 0, builds = 0, batches = 0, rounds = 0;
    for (;;) {
      const unsigned active = __ballot_sync(0xffffffffu, !L.done());
      if (!active) break;
      const bool at_header = L.state == LS_HEADER;
      const unsigned mh = __ballot_sync(0xffffffffu, at_header);
      if (mh) {
        const bool allow = (active & ~mh) == 0 || __popc(mh) >= 16 || waited >= PATIENCE; waited = allow ? 0 : waited + 1; builds += allow;
        if (allow && at_header) L.header();
        __syncwarp();
      }
      votes++;
#pragma unroll 1
      for (int k = 0; k < 4; k++) { if (L.state == LS_SYMBOL) L.iter(); __syncwarp(); }
      L.check();
    }
    if (work && L.status) { status[b] = L.status; atomicOr(err, DERR_INFLATE); }
    __syncwarp(); const long long tB = clock64();
    // ---- phase B: the warp resolves the matches of its 32 blocks, one block at a time
    for (int j = 0; j < 32; j++) {
      const uint32_t n = __shfl_sync(0xffffffffu, L.ntok, j); const int st = __shfl_sync(0xffffffffu, L.status, j);
      if (n == 0 || st != 0) continue;
      const uint32_t* tp = (const uint32_t*)shfl_ptr(my_tok, j); uint8_t* ob = (uint8_t*)shfl_ptr(my_out, j);
      const uint32_t* lits = tp + (TOK_CAP - 1);
      uint32_t pos = 0, qpos = 0;
      uint32_t t_next = lane < n ? tp[lane] : 0;
      const uint32_t nlw = __shfl_sync(0xffffffffu, L.nlw, j);               // literal words of this block
      uint32_t lwin = (uint32_t)lane < nlw ? lits[-lane] : 0u;
      for (uint32_t tb = 0; tb < n; tb += 32) {
        const uint32_t t = t_next, win = lwin, wq = qpos >> 2;
        t_next = tb + 32 + lane < n ? tp[tb + 32 + lane] : 0;   // the next batch's tokens are on their way while this one is resolved
        const uint32_t ml = tok_mlen(t), nl = tok_lit(t); uint32_t incl = nl + ml, linc = nl;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o), w = __shfl_up_sync(0xffffffffu, linc, o); if (lane >= o) { incl += v; linc += w; } }
        const uint32_t mstart = pos + incl - ml, lstart = mstart - nl, lq = qpos + linc - nl;
        pos += __shfl_sync(0xffffffffu, incl, 31); qpos += __shfl_sync(0xffffffffu, linc, 31);
        { const uint32_t nw = (qpos >> 2) + (uint32_t)lane; lwin = nw < nlw ? lits[-(int)nw] : 0u; }   // ... and so is the next batch's window of the literal stream
        resolve_batch(ob, lits, win, wq, lstart, nl, lq, mstart, ml, tok_dist(t), lane, rounds); batches++;
      }
    }
    __syncwarp();
    if (lane == 0) { const long long tC = clock64(); atomicAdd(&g_tok_prof[0], (unsigned long long)(tB - tA)); atomicAdd(&g_tok_prof[1], (unsigned long long)(tC - tB)); atomicAdd(&g_tok_prof[2], votes);
      atomicAdd(&g_tok_prof[3], builds); atomicAdd(&g_tok_prof[4], batches); atomicAdd(&g_tok_prof[5], rounds); atomicAdd(&g_tok_prof[6], 1ull); }
  }
}

// One geometry record per kernel instantiation and device (the shared-memory opt-in and the occupancy are per device).
struct TokGeom { int ctas_per_sm, sms; size_t smem; int warps; };
template <int WARPS, int LB, int DB, int PATIENCE, int MINB, int FUSE>
static TokGeom tok_geometry() {
  static int ctas[64] = {0}, sms[64] = {0};
  const size_t smem = (size_t)WARPS * TokLane<32, LB, DB>::TABLE_U16 * sizeof(uint16_t);
  int dev = 0; IRF_CUDA(cudaGetDevice(&dev)); if (dev < 0 || dev >= 64) throw std::string("device ordinal beyond 63 is not supported");
  if (!ctas[dev]) {
    auto kern = inflate_tok_kernel<WARPS, LB, DB, PATIENCE, MINB, FUSE>;
    IRF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    IRF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas[dev], kern, WARPS * 32, smem));
    IRF_CUDA(cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev));
    if (ctas[dev] < 1) ctas[dev] = 1;
  }
  return TokGeom{ctas[dev], sms[dev], smem, WARPS};
}
template <int WARPS, int LB, int DB, int PATIENCE, int MINB, int FUSE>
static void tok_launch(const InflateArgs& a, cudaStream_t s) {
  const TokGeom g = tok_geometry<WARPS, LB, DB, PATIENCE, MINB, FUSE>();
  const int need = (a.nblk + WARPS * 32 - 1) / (WARPS * 32), grid = std::min(need, g.ctas_per_sm * g.sms);
  if (grid <= 0) return;
  if (!a.tok_scratch || a.tok_scratch_bytes < (size_t)grid * WARPS * 32 * TOK_CAP * sizeof(uint32_t)) throw std::string("inflate: token scratch missing or too small");
  const bool prof = getenv("IRF_TOK_PROF") != nullptr;
  if (prof) { unsigned long long z[8] = {0}; IRF_CUDA(cudaMemcpyToSymbol(g_tok_prof, z, sizeof z)); }
  IRF_CUDA(cudaMemsetAsync(a.next_block, 0, sizeof(unsigned int), s));
  inflate_tok_kernel<WARPS, LB, DB, PATIENCE, MINB, FUSE><<<grid, WARPS * 32, g.smem, s>>>(a.comp, a.c_off, a.c_len, a.U, a.u_off, a.u_len, a.nblk, a.status, a.err, a.next_block, a.watermark, a.list, a.tok_scratch);
  note_launch();
  if (prof) { IRF_CUDA(cudaStreamSynchronize(s)); unsigned long long z[8]; IRF_CUDA(cudaMemcpyFromSymbol(z, g_tok_prof, sizeof z));
    fprintf(stderr, "[tok] <%d warps, %d/%d-bit LUTs, fuse %d> grid %d x %d CTAs/SM: phase A %.1f Mcycles/claim, phase B %.1f Mcycles/claim, %llu claims; per claim: %.0f votes, %.1f table builds, %.0f batches, %.2f rounds/batch\n", WARPS, LB, DB, FUSE, grid, g.ctas_per_sm,
            z[0] / 1e6 / std::max(1ull, z[6]), z[1] / 1e6 / std::max(1ull, z[6]), z[6], (double)z[2] / std::max(1ull, z[6]), (double)z[3] / std::max(1ull, z[6]), (double)z[4] / std::max(1ull, z[6]), (double)z[5] / std::max(1ull, z[4])); }
}
// variants: fused iteration, warps per SM (registers per thread follow from MINB), LUT width (shared memory per warp = TokLane::TABLE_U16 * 2 bytes).
// What counts is the steady state of a full grid, blocks per cycle = resident warps / cycles per claim of 32 blocks (IRF_TOK_PROF=1), because
// the pipeline launches the kernel one grid-full of blocks at a time; the time for a whole file also depends on how its block count
// divides into grid-fulls (5651 claims are 1.91 rounds of 2960 warps but 1.36 rounds of 4144, which hid this for a round).
//   12: fused, 6/5 bits, 7 CTAs x 4 warps = 28 warps per SM (72 registers, 60 bytes spilled): 137.7 Mcycles per claim -> 0.203 claims per Mcycle and SM
//   17: fused, 24 warps (80 registers): 125.0 Mcycles -> 0.192        13: fused, 20 warps (95 registers, no spill): 117.8 Mcycles -> 0.170
//   16: not fused, 20 warps (the kernel as first committed in round 2): 123.0 Mcycles -> 0.163
//   14: 7/5 bits, 15 warps; 15: 8/6 bits, 9 warps (not fused: the wider LUT leaves too few bits per refill; both far slower)
#define IRF_TOK_DISPATCH(variant, CALL) \
  switch (variant) { case 13: CALL(4, 6, 5, 24, 5, 1); break; case 14: CALL(3, 7, 5, 24, 5, 0); break; case 15: CALL(3, 8, 6, 24, 3, 0); break; case 16: CALL(4, 6, 5, 24, 5, 0); break; \
                     case 17: CALL(4, 6, 5, 24, 6, 1); break; default: CALL(4, 6, 5, 24, 7, 1); break; }
static TokGeom tok_geom_of(int variant) {
  TokGeom g{};
#define IRF_TOK_GEOM(W, LB, DB, P, M, F) g = tok_geometry<W, LB, DB, P, M, F>()
  IRF_TOK_DISPATCH(variant, IRF_TOK_GEOM)
  return g;
}
// lanes (= BGZF blocks in flight) of a full grid, and the scratch a launch over nblk blocks needs
int inflate_tok_lanes(int variant) { const TokGeom g = tok_geom_of(variant); return g.ctas_per_sm * g.sms * g.warps * 32; }
size_t inflate_tok_scratch_bytes(int variant, int nblk) {
  const TokGeom g = tok_geom_of(variant);
  const int per = g.warps * 32, need = (nblk + per - 1) / per, grid = std::max(1, std::min(need, g.ctas_per_sm * g.sms));
  return (size_t)grid * per * TOK_CAP * sizeof(uint32_t);
}
static std::atomic<unsigned long long> g_tok_launches{0};
unsigned long long tok_launch_counter() { return g_tok_launches.load(std::memory_order_relaxed); }   // launches of the two-phase kernel by this process (tests check that the default path really runs it)
void launch_inflate_tok(const InflateArgs& a, int variant, cudaStream_t s) {
  g_tok_launches.fetch_add(1, std::memory_order_relaxed);
#define IRF_TOK_LAUNCH(W, LB, DB, P, M, F) tok_launch<W, LB, DB, P, M, F>(a, s)
  IRF_TOK_DISPATCH(variant, IRF_TOK_LAUNCH)
}

}  // namespace irf
