// inflate_tok.cu — the two-phase SM inflate kernel (see inflate_tok.cuh for the format and the reasons).
//
// One persistent launch, 148 SMs x resident CTAs; every warp claims 32 consecutive BGZF blocks from a global counter and runs
//   phase A  lane per block: Huffman decode, literals stored at their final place as aligned 32-bit words, one 4-byte token per
//            LZ77 match appended to the lane's scratch region;
//   phase B  warp per block, for each of the 32 blocks in turn: tokens are read 32 at a time (coalesced), a warp scan turns their
//            lengths into output positions, and the matches are copied — in parallel when their source bytes are final, in order
//            otherwise (a match may read what an earlier match of the same batch writes);
// then claims the next 32.  Replaces `gzip -cd` of the reference pipeline; RFC 1951.
#include <algorithm>
#include "common.cuh"
#include "inflate_tok.cuh"
#include "kernels.h"

namespace irf {

__device__ __forceinline__ const void* shfl_ptr(const void* p, int src) {
  unsigned long long v = (unsigned long long)p;
  unsigned lo = __shfl_sync(0xffffffffu, (unsigned)v, src), hi = __shfl_sync(0xffffffffu, (unsigned)(v >> 32), src);
  return (const void*)((unsigned long long)hi << 32 | lo);
}

// Phase B for one batch of up to 32 tokens of one block: lane i holds token i (mlen = 0: nothing to copy) and the output offset
// `mstart` where its match begins.  Everything below the first unfinished match is final, which gives the readiness test.
constexpr uint32_t TOK_LONG = 24;   // matches longer than this are copied by the whole warp
__device__ __forceinline__ void resolve_batch(uint8_t* ob, uint32_t mstart, uint32_t ml, uint32_t d, int lane) {
  bool pending = ml != 0;
  const uint32_t need = mstart - d + min(ml, d);   // source bytes [mstart - d, need) must be final before this match may be copied
  for (;;) {
    const unsigned pm = __ballot_sync(0xffffffffu, pending);
    if (!pm) break;
    const int f = __ffs(pm) - 1;                     // first unfinished match: all output below its start is final
    const uint32_t fstart = __shfl_sync(0xffffffffu, mstart, f), fml = __shfl_sync(0xffffffffu, ml, f);
    if (fml > TOK_LONG) {                            // warp-uniform: all 32 lanes copy match f
      const uint32_t fd = __shfl_sync(0xffffffffu, d, f);
      uint8_t* dp = ob + fstart; const uint8_t* sp = dp - fd;
      if (fd >= 32) {                                // a 32-byte slab never reads what the same slab writes; earlier slabs are ordered by the barrier
        for (uint32_t k0 = 0; k0 < fml; k0 += 32) { if (k0 + lane < fml) dp[k0 + lane] = sp[k0 + lane]; __syncwarp(); }
      } else {                                       // period shorter than a slab: every byte is a copy of one of the fd final bytes before the match
        for (uint32_t k = lane; k < fml; k += 32) dp[k] = sp[k % fd];
      }
      if (lane == f) pending = false;
      __syncwarp();
      continue;
    }
    const bool ready = pending && ml <= TOK_LONG && (lane == f || need <= fstart);
    if (ready) {
      uint8_t* dp = ob + mstart; const uint8_t* sp = dp - d; uint32_t w = 0, rem = ml;
      while (rem) {                                  // 8 bytes per round: loads first, then stores; the source index wraps inside the period,
        const uint32_t k = rem < 8 ? rem : 8; uint32_t v[8];   // so overlapping (run-length) matches only ever read the d final bytes
#pragma unroll
        for (int j = 0; j < 8; j++) { if ((uint32_t)j < k) v[j] = sp[w]; w++; if (w == d) w = 0; }
#pragma unroll
        for (int j = 0; j < 8; j++) if ((uint32_t)j < k) dp[j] = (uint8_t)v[j];
        dp += k; rem -= k;
      }
      pending = false;
    }
    __syncwarp();   // orders this round's stores before the next round's loads (and before the next batch)
  }
}

template <int WARPS, int LB, int DB, int PATIENCE>
__global__ void __launch_bounds__(WARPS * 32) inflate_tok_kernel(const uint8_t* __restrict__ comp, const u64* __restrict__ c_off, const u32* __restrict__ c_len,
                                                                  uint8_t* U, const u64* __restrict__ u_off, const u32* __restrict__ u_len, int nblk,
                                                                  int* __restrict__ status, int* __restrict__ err, unsigned int* __restrict__ next_block, const unsigned long long* watermark,
                                                                  const u32* __restrict__ list, uint32_t* tok_scratch) {
  extern __shared__ uint16_t smem_tables[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // per warp: [lit LUT][dist LUT][2 x 32 canonical-limit entries], 16-bit entries interleaved by lane
  constexpr size_t WARP_U16 = (size_t)((1 << LB) + (1 << DB) + 64) * 32;
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
    TokLane<32, LB, DB> L;
    const bool work = in_range && !(u_len[b] == 0 && c_len[b] <= 2);  // skip the 28-byte EOF marker / empty blocks
    if (work) L.start(comp + c_off[b], c_len[b], U + u_off[b], u_len[b], lit, dst, lane, sym_store, my_tok); else L.idle();
    // ---- phase A.  Votes once per four symbols; a table build (thousands of instructions that every lane of the warp pays for)
    // starts only when no lane is left decoding, half the warp is waiting for one, or the waiting lanes have waited PATIENCE votes.
    int waited = 0;
    for (;;) {
      const unsigned active = __ballot_sync(0xffffffffu, !L.done());
      if (!active) break;
      const bool at_header = L.state == LS_HEADER;
      const unsigned mh = __ballot_sync(0xffffffffu, at_header);
      bool allow = false;
      if (mh) { allow = (active & ~mh) == 0 || __popc(mh) >= 16 || waited >= PATIENCE; waited = allow ? 0 : waited + 1; }
      L.step(allow && at_header); __syncwarp();
#pragma unroll 1
      for (int k = 0; k < 3; k++) { L.step(false); __syncwarp(); }
    }
    if (work && L.status) { status[b] = L.status; atomicOr(err, DERR_INFLATE); }
    __syncwarp();
    // ---- phase B: the warp resolves the matches of its 32 blocks, one block at a time
    for (int j = 0; j < 32; j++) {
      const uint32_t n = __shfl_sync(0xffffffffu, L.ntok, j); const int st = __shfl_sync(0xffffffffu, L.status, j);
      if (n == 0 || st != 0) continue;
      const uint32_t* tp = (const uint32_t*)shfl_ptr(my_tok, j); uint8_t* ob = (uint8_t*)shfl_ptr(L.out, j);
      uint32_t pos = 0;
      uint32_t t_next = lane < n ? tp[lane] : 0;
      for (uint32_t tb = 0; tb < n; tb += 32) {
        const uint32_t t = t_next;
        t_next = tb + 32 + lane < n ? tp[tb + 32 + lane] : 0;   // the next batch's tokens are on their way while this one is resolved
        const uint32_t ml = tok_mlen(t), tot = tok_lit(t) + ml; uint32_t incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
        const uint32_t mstart = pos + incl - ml;
        pos += __shfl_sync(0xffffffffu, incl, 31);
        resolve_batch(ob, mstart, ml, tok_dist(t), lane);
      }
    }
    __syncwarp();
  }
}

static int g_tok_ctas[64] = {0}, g_tok_sms[64] = {0};
template <int WARPS, int LB, int DB, int PATIENCE>
static void tok_geometry(int& ctas_per_sm, int& sms, size_t& smem) {
  smem = (size_t)WARPS * ((size_t)((1 << LB) + (1 << DB) + 64) * 32) * sizeof(uint16_t);
  int dev = 0; IRF_CUDA(cudaGetDevice(&dev)); if (dev < 0 || dev >= 64) throw std::string("device ordinal beyond 63 is not supported");
  if (!g_tok_ctas[dev]) {   // per device: the shared-memory opt-in and the occupancy
    auto kern = inflate_tok_kernel<WARPS, LB, DB, PATIENCE>;
    IRF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    IRF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_tok_ctas[dev], kern, WARPS * 32, smem));
    IRF_CUDA(cudaDeviceGetAttribute(&g_tok_sms[dev], cudaDevAttrMultiProcessorCount, dev));
    if (g_tok_ctas[dev] < 1) g_tok_ctas[dev] = 1;
  }
  ctas_per_sm = g_tok_ctas[dev]; sms = g_tok_sms[dev];
}
constexpr int TOK_WARPS = 4, TOK_LB = 6, TOK_DB = 5, TOK_PATIENCE = 24;

// lanes (= BGZF blocks in flight) of a full grid, and the scratch a launch over nblk blocks needs
int inflate_tok_lanes() { int c, s; size_t m; tok_geometry<TOK_WARPS, TOK_LB, TOK_DB, TOK_PATIENCE>(c, s, m); return c * s * TOK_WARPS * 32; }
size_t inflate_tok_scratch_bytes(int nblk) {
  int c, s; size_t m; tok_geometry<TOK_WARPS, TOK_LB, TOK_DB, TOK_PATIENCE>(c, s, m);
  const int need = (nblk + TOK_WARPS * 32 - 1) / (TOK_WARPS * 32), grid = std::max(1, std::min(need, c * s));
  return (size_t)grid * TOK_WARPS * 32 * TOK_CAP * sizeof(uint32_t);
}
void launch_inflate_tok(const InflateArgs& a, cudaStream_t s) {
  int c, sms; size_t smem; tok_geometry<TOK_WARPS, TOK_LB, TOK_DB, TOK_PATIENCE>(c, sms, smem);
  const int need = (a.nblk + TOK_WARPS * 32 - 1) / (TOK_WARPS * 32), grid = std::min(need, c * sms);
  if (grid <= 0) return;
  if (!a.tok_scratch || a.tok_scratch_bytes < (size_t)grid * TOK_WARPS * 32 * TOK_CAP * sizeof(uint32_t)) throw std::string("inflate: token scratch missing or too small");
  IRF_CUDA(cudaMemsetAsync(a.next_block, 0, sizeof(unsigned int), s));
  inflate_tok_kernel<TOK_WARPS, TOK_LB, TOK_DB, TOK_PATIENCE><<<grid, TOK_WARPS * 32, smem, s>>>(a.comp, a.c_off, a.c_len, a.U, a.u_off, a.u_len, a.nblk, a.status, a.err, a.next_block, a.watermark, a.list, a.tok_scratch);
  note_launch();
}

}  // namespace irf
