// kernels.cu — sm_100a kernels of the BAM -> intron-retention counting path.
//
// Stage map:  S1 inflate (inflate_tok.cu: the default two-phase kernel; inflate_kernel below: the one-phase variants 0-11) and
// crc_check_kernel  ->  S2 frame_guess / walk_count / frame_mismatch+mark / gather_records (record framing)  ->  S3 pair_records_kernel (skip /
// single / pair decision, mate pairing by device hash), process_pairs / process_singles (CIGAR -> blocks, FragmentBlocks formation and every
// ReadBlockProcessor's ProcessBlocks)  ->  S4 depth scan + intron_stats kernels; plus the table maintenance of the multi-GPU merge.
// Each device function that restates a reference routine names it (upstream file per SURVEY.md §8a; the
// mounted /root/reference has no source to cite by line).
#include <algorithm>
#include <atomic>
#include <cub/device/device_radix_sort.cuh>
#include "common.cuh"
#include "inflate_core.cuh"
#include "kernels.h"

namespace irf {

static std::atomic<unsigned long long> g_launches{0};  // kernels launched by this library (bench.py reports it as gpu_launches); contexts of several devices launch from their own host threads
unsigned long long launch_counter() { return g_launches.load(std::memory_order_relaxed); }
#define IRF_LAUNCHED() (g_launches.fetch_add(1, std::memory_order_relaxed))
void note_launch() { IRF_LAUNCHED(); }

// ------------------------------------------------------------------------------------------------
// small device helpers
__device__ __forceinline__ u32 ld32(const uint8_t* p) {  // unaligned little-endian load from two aligned words
  uintptr_t a = (uintptr_t)p; u32 sh = (u32)(a & 3) * 8; const u32* w = (const u32*)(a & ~(uintptr_t)3);
  u32 lo = w[0]; if (!sh) return lo;
  return __funnelshift_r(lo, w[1], sh);
}
__device__ __forceinline__ u64 mix64(u64 x) { x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33; return x; }

struct Core {  // the fields of a BAM record's fixed part that this path reads (SAM v1 §4.2)
  int block_size, refID, pos, l_seq, next_refID, next_pos; u32 l_read_name, n_cigar_op, flag;
};
__device__ __forceinline__ void load_core(const uint8_t* U, u64 off, Core& c) {
  const uint8_t* p = U + off; uintptr_t a = (uintptr_t)p; u32 sh = (u32)(a & 3) * 8; const u32* w = (const u32*)(a & ~(uintptr_t)3);
  u32 x[10];
#pragma unroll
  for (int i = 0; i < 10; i++) x[i] = w[i];
  u32 f[9];
#pragma unroll
  for (int i = 0; i < 9; i++) f[i] = sh ? __funnelshift_r(x[i], x[i + 1], sh) : x[i];
  c.block_size = (int)f[0]; c.refID = (int)f[1]; c.pos = (int)f[2]; c.l_read_name = f[3] & 255u; c.n_cigar_op = f[4] & 0xffffu; c.flag = f[4] >> 16;
  c.l_seq = (int)f[5]; c.next_refID = (int)f[6]; c.next_pos = (int)f[7];
}

// ------------------------------------------------------------------------------------------------
// S1: BGZF / DEFLATE inflate, one lane per BGZF block.  Persistent CTAs: every warp claims the next 32 blocks from a
// global counter until none are left, so a launch has one short tail instead of one per wave.
// D8: byte entries in the distance LUT (1 KB less shared memory per warp); MINB: CTAs per SM the register allocation must allow.
template <int WARPS, int LB, int DB, int HEADER_BATCH, int LANE_CLAIM, int PZ, int D8 = 0, int MINB = 0>
__global__ void __launch_bounds__(WARPS * 32, MINB) inflate_kernel(const uint8_t* __restrict__ comp, const u64* __restrict__ c_off, const u32* __restrict__ c_len,
                                                             uint8_t* __restrict__ U, const u64* __restrict__ u_off, const u32* __restrict__ u_len, int nblk,
                                                             int* __restrict__ status, int* __restrict__ err, unsigned int* __restrict__ next_block, const unsigned long long* watermark,
                                                             const u32* __restrict__ list) {   // list: optional indices of the blocks this kernel owns (the rest go to the decompression engine)
  extern __shared__ uint16_t smem_tables[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // per warp: [lit LUT][dist LUT][2 x 32 canonical-limit entries] (16-bit, lane-interleaved)
  constexpr size_t DIST_U16 = D8 ? (1 << DB) * 16 : (1 << DB) * 32;
  constexpr size_t WARP_U16 = (size_t)((1 << LB) + 64) * 32 + DIST_U16;
  uint16_t* lit = smem_tables + (size_t)warp * WARP_U16;
  uint16_t* dst = lit + (1 << LB) * 32;
  uint16_t sym_store[320];
  if (!LANE_CLAIM) {   // warp-granular claims: 32 consecutive blocks per warp, all lanes start together
    for (;;) {
      unsigned int base = 0;
      if (lane == 0) base = atomicAdd(next_block, 32u);
      base = __shfl_sync(0xffffffffu, base, 0);
      if (base >= (unsigned int)nblk) break;
      if (watermark) {   // streaming input: wait until the host->device copy has passed the last byte these 32 blocks need
        int lastb = min((int)base + 31, nblk - 1); if (list) lastb = (int)list[lastb]; u64 need = c_off[lastb] + c_len[lastb] + 160; bool starved = false;   // + the bit reader's look-ahead, rounded up to whole sectors:
                                                                               // no lane may pull a not-yet-written sector into L1
        if (lane == 0) {
          long long t0 = clock64();
          while (*(volatile const unsigned long long*)watermark < need) {
            __nanosleep(500);
            if (clock64() - t0 > 40000000000ll) { starved = true; break; }   // ~20 s without data: give up instead of hanging the GPU
          }
          __threadfence();
        }
        starved = __shfl_sync(0xffffffffu, (int)starved, 0) != 0;
        if (starved) { if (lane == 0) atomicOr(err, DERR_STARVED); break; }
      }
      int b = (int)base + lane; const bool in_range = b < nblk; if (list && in_range) b = (int)list[b];
      InflateLane<32, LB, DB, PZ, D8> L;
      bool work = in_range && !(u_len[b] == 0 && c_len[b] <= 2);  // skip the 28-byte EOF marker / empty blocks
      if (work) L.start(comp + c_off[b], c_len[b], U + u_off[b], u_len[b], lit, dst, lane, sym_store); else L.idle();
      while (__any_sync(0xffffffffu, !L.done())) L.step();   // the vote keeps the warp converged (see InflateLane)
      if (work && L.status) { status[b] = L.status; atomicOr(err, DERR_INFLATE); }
    }
    return;
  }
  InflateLane<32, LB, DB, PZ, D8> L; L.idle();
  int b = -1; bool exhausted = false;
  // HEADER_BATCH: lanes that must be waiting for a block header before the warp pays for one
  for (;;) {
    // lanes that finished their block report it and claim the next one (warp-aggregated counter bump): no lane idles
    // while another lane of its warp still has half a block to go
    bool need = L.done() && !exhausted;
    if (need && b >= 0 && L.status) { status[b] = L.status; atomicOr(err, DERR_INFLATE); }
    unsigned int m = __ballot_sync(0xffffffffu, need);
    if (m) {
      unsigned int base = 0; int leader = __ffs(m) - 1;
      if (lane == leader) base = atomicAdd(next_block, (unsigned int)__popc(m));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (need) {
        unsigned int nb = base + (unsigned int)__popc(m & ((1u << lane) - 1u));
        if (nb < (unsigned int)nblk) { b = (int)nb; if (u_len[b] == 0 && c_len[b] <= 2) L.idle(); else L.start(comp + c_off[b], c_len[b], U + u_off[b], u_len[b], lit, dst, lane, sym_store); }
        else { b = -1; exhausted = true; }
      }
    }
    if (!__any_sync(0xffffffffu, !L.done() || !exhausted)) break;
    unsigned int mh = __ballot_sync(0xffffffffu, L.state == LS_HEADER);
    unsigned int mrun = __ballot_sync(0xffffffffu, !L.done() && L.state != LS_HEADER);
    const bool allow_header = __popc(mh) >= HEADER_BATCH || mrun == 0;
    // the bookkeeping votes above run once per 8 steps; inside the group a plain warp barrier keeps the lanes converged
#pragma unroll 1
    for (int k = 0; k < 8; k++) { L.step(allow_header); __syncwarp(); }
  }
}

static int g_inflate_lanes = 0, g_inflate_cta_cap = 0;
void set_inflate_cta_cap(int ctas_per_sm) { g_inflate_cta_cap = ctas_per_sm; }   // 0 = as many CTAs per SM as fit
int inflate_resident_lanes() { return g_inflate_lanes; }   // lanes (= BGZF blocks in flight) of the last inflate launch's full grid; 0 before the first launch
template <int WARPS, int LB, int DB, int HB = 1, int LC = 0, int PZ = 0, int D8 = 0, int MINB = 0>
static void launch_inflate_t(const InflateArgs& a, cudaStream_t s) {
  size_t smem = (size_t)WARPS * (((size_t)((1 << LB) + 64) * 32 + (D8 ? (1 << DB) * 16 : (1 << DB) * 32)) * sizeof(uint16_t));
  // the shared-memory opt-in and the occupancy are per device: one cache slot per device ordinal
  static int ctas_cache[64] = {0}, sms_cache[64] = {0};
  int dev = 0; IRF_CUDA(cudaGetDevice(&dev)); if (dev < 0 || dev >= 64) throw std::string("device ordinal beyond 63 is not supported");
  int& ctas_per_sm = ctas_cache[dev]; int& sms = sms_cache[dev];
  if (!ctas_per_sm) {
    IRF_CUDA(cudaFuncSetAttribute(inflate_kernel<WARPS, LB, DB, HB, LC, PZ, D8, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    IRF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, inflate_kernel<WARPS, LB, DB, HB, LC, PZ, D8, MINB>, WARPS * 32, smem));
    IRF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    if (ctas_per_sm < 1) ctas_per_sm = 1;
  }
  const int per_sm = g_inflate_cta_cap > 0 ? std::min(g_inflate_cta_cap, ctas_per_sm) : ctas_per_sm;
  g_inflate_lanes = sms * per_sm * WARPS * 32;
  int need = (a.nblk + WARPS * 32 - 1) / (WARPS * 32), grid = std::min(need, sms * per_sm);   // 148 SMs x resident CTAs
  if (grid > 0) {
    IRF_CUDA(cudaMemsetAsync(a.next_block, 0, sizeof(unsigned int), s));
    inflate_kernel<WARPS, LB, DB, HB, LC, PZ, D8, MINB><<<grid, WARPS * 32, smem, s>>>(a.comp, a.c_off, a.c_len, a.U, a.u_off, a.u_len, a.nblk, a.status, a.err, a.next_block, LC ? nullptr : a.watermark, LC ? nullptr : a.list); IRF_LAUNCHED();
  }
}
size_t inflate_scratch_bytes(int variant, int nblk) { return variant >= 12 && variant <= 17 && nblk > 0 ? inflate_tok_scratch_bytes(variant, nblk) : 0; }
void launch_inflate(const InflateArgs& a, int variant, cudaStream_t s) {
  if (variant >= 12 && variant <= 17) { launch_inflate_tok(a, variant, s); g_inflate_lanes = inflate_tok_lanes(variant); return; }   // two-phase kernel: token decode per lane, then warp-wide resolve (inflate_tok.cu)
  switch (variant) {   // one-phase kernel: <warps per CTA, literal LUT bits, distance LUT bits, header batch, lane-granular claims, pre-written sectors, byte distance LUT, CTAs per SM>
    case 2: launch_inflate_t<4, 6, 5, 1, 0, 1>(a, s); break;   // pre-written output sectors
    case 3: launch_inflate_t<4, 6, 6, 1, 0, 1>(a, s); break;
    case 4: launch_inflate_t<4, 6, 6, 1, 0, 0>(a, s); break;
    case 5: launch_inflate_t<4, 7, 6, 1, 0, 1>(a, s); break;
    case 7: launch_inflate_t<4, 7, 6, 1, 0, 0>(a, s); break;
    case 8: launch_inflate_t<4, 6, 5, 4, 1, 0>(a, s); break;    // lane-granular claims + header batching (measured slower, kept for A/B)
    case 10: launch_inflate_t<4, 6, 5, 1, 0, 0, 1, 6>(a, s); break;   // byte distance LUT: 36 KB per CTA and <= 80 registers, six CTAs per SM
    case 11: launch_inflate_t<4, 6, 5, 1, 0, 0, 1, 0>(a, s); break;   // byte distance LUT only (five CTAs; isolates the table change from the occupancy change)
    case 0: launch_inflate_t<4, 6, 5, 1, 0, 0>(a, s); break;    // the round-1 default
    default: throw std::string("unknown inflate_variant ") + std::to_string(variant) + " (12-15: two-phase kernel, 0, 2-5, 7, 8, 10, 11: one-phase kernel)";
  }
}

// ------------------------------------------------------------------------------------------------
// S1b: CRC-32 (IEEE, reflected, as in RFC 1952) of each inflated block.  One warp per block: the block is cut from its END
// into 2 KiB slices, lane i takes slice i (lanes before the start of a short block stay empty), each lane runs the
// table-driven register update over its slice (slicing-by-4 tables in shared memory), and the 32 registers are merged in a
// 5-level tree: combine(A, B) = A * x^(8|B|) mod P  xor  B, where |B| is always 2^k * 2 KiB so the five multipliers are constants.
constexpr u32 CRC_POLY = 0xEDB88320u; constexpr u32 CRC_SLICE = 2048;
__host__ __device__ inline u32 crc_mulmod(u32 a, u32 b) {   // product of two polynomials mod P, bit 31 = x^0 (zlib's multmodp)
  u32 p = 0;
  for (int i = 0; i < 32; i++) { if (a & (0x80000000u >> i)) p ^= b; b = (b & 1u) ? (b >> 1) ^ CRC_POLY : b >> 1; }
  return p;
}
struct CrcMul { u32 x[5]; };
// Tables: slicing-by-4, one private copy per LANE in shared memory (entry (k, i) of lane l at ((k * 256 + i) * 32 + l): bank = lane),
// so the 32 data-dependent look-ups of a warp never collide.  With one shared copy the look-ups of a warp hit ~3.5 entries per bank
// on average and the kernel ran at 1.45 TB/s (20 ms per 50M-pair step, profiles/README.md round 1); the look-up, one per byte and
// lane, is the unit of work here, so conflict-free look-ups are what lets the kernel follow HBM.  128 KB of tables: one 512-thread
// CTA per SM, persistent, filled once per CTA.
constexpr int CRC_THREADS = 512;
__global__ void __maxnreg__(64) crc_check_kernel(const uint8_t* __restrict__ U, const u64* __restrict__ u_off, const u32* __restrict__ u_len, const u32* __restrict__ want,
                                                                   int nblk, CrcMul M, int* __restrict__ status, int* __restrict__ err, const u32* __restrict__ list) {
  extern __shared__ u32 crc_tab[];   // [4][256][32]
  for (int i = threadIdx.x; i < 256; i += blockDim.x) { u32 c = (u32)i; for (int k = 0; k < 8; k++) c = (c & 1u) ? (c >> 1) ^ CRC_POLY : c >> 1; crc_tab[i * 32] = c; }
  __syncthreads();
  for (int i = threadIdx.x; i < 256; i += blockDim.x) { u32 c = crc_tab[i * 32]; for (int t = 1; t < 4; t++) { c = crc_tab[(c & 255u) * 32] ^ (c >> 8); crc_tab[(t * 256 + i) * 32] = c; } }
  __syncthreads();
  for (int e = threadIdx.x; e < 1024 * 31; e += blockDim.x) { const int ent = e / 31, l = e % 31 + 1; crc_tab[ent * 32 + l] = crc_tab[ent * 32]; }   // the other 31 copies
  __syncthreads();
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  const u32* T0 = crc_tab + lane; const u32* T1 = T0 + 256 * 32; const u32* T2 = T1 + 256 * 32; const u32* T3 = T2 + 256 * 32;
  auto word = [&](u32 w, u32 r) -> u32 { w ^= r; return T3[(w & 255u) * 32] ^ T2[((w >> 8) & 255u) * 32] ^ T1[((w >> 16) & 255u) * 32] ^ T0[(w >> 24) * 32]; };
  for (int bi = blockIdx.x * wpb + (threadIdx.x >> 5); bi < nblk; bi += gridDim.x * wpb) {
    const int b = list ? (int)list[bi] : bi;   // optional list of block indices (the blocks one engine produced)
    u32 len = u_len[b]; if (!len) continue;
    const uint8_t* base = U + u_off[b];
    long long hi = (long long)len - (long long)(31 - lane) * CRC_SLICE, lo = hi - CRC_SLICE;   // this lane's slice [lo, hi) of the block
    if (lo < 0) lo = 0;
    u32 r = 0;
    if (hi > 0) {
      if (lo == 0) r = 0xFFFFFFFFu;   // the lane holding the first byte carries the initial register
      const uint8_t* p = base + lo; const uint8_t* e = base + hi;
      while (p < e && ((uintptr_t)p & 15)) { r = T0[((r ^ *p++) & 255u) * 32] ^ (r >> 8); }
      for (; p + 64 <= e; p += 64) {   // four 128-bit loads in flight per lane before the first is consumed
        const uint4 q0 = *(const uint4*)p, q1 = *(const uint4*)(p + 16), q2 = *(const uint4*)(p + 32), q3 = *(const uint4*)(p + 48);
        r = word(q0.x, r); r = word(q0.y, r); r = word(q0.z, r); r = word(q0.w, r);
        r = word(q1.x, r); r = word(q1.y, r); r = word(q1.z, r); r = word(q1.w, r);
        r = word(q2.x, r); r = word(q2.y, r); r = word(q2.z, r); r = word(q2.w, r);
        r = word(q3.x, r); r = word(q3.y, r); r = word(q3.z, r); r = word(q3.w, r);
      }
      for (; p + 16 <= e; p += 16) { const uint4 q = *(const uint4*)p; r = word(q.x, r); r = word(q.y, r); r = word(q.z, r); r = word(q.w, r); }
      while (p < e) { r = T0[((r ^ *p++) & 255u) * 32] ^ (r >> 8); }
    }
#pragma unroll 1
    for (int k = 0; k < 5; k++) {
      u32 v = __shfl_xor_sync(0xffffffffu, r, 1 << k); bool upper = (lane >> k) & 1;
      u32 a = upper ? v : r, bb = upper ? r : v;   // a = the earlier bytes, bb = the later 2^k slices
      r = crc_mulmod(M.x[k], a) ^ bb;
    }
    if (lane == 0 && (~r) != want[b]) { status[b] = 100; atomicOr(err, DERR_CRC); }
  }
}
__global__ void de_check_kernel(const u32* __restrict__ list, int n, const u32* __restrict__ act, const u32* __restrict__ u_len, int* __restrict__ status, int* __restrict__ err) {
  int i = blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
  u32 b = list[i]; if (act[b] != u_len[b]) { status[b] = INF_ERR_ISIZE; atomicOr(err, DERR_INFLATE); }
}
// Small tables (block offsets, lengths, CRCs, work lists) are fetched by a kernel from page-locked host memory instead of
// being copied by the DMA engine: the engine's queue is full of the BGZF bytes' 256 MB chunk copies at that time and a
// table queued behind them would arrive after the whole file.
__global__ void host_fetch_kernel(uint4* __restrict__ dst, const uint4* __restrict__ src_host, size_t n16) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = src_host[i];
}
void launch_host_fetch(void* dst_dev, const void* src_host_mapped, size_t nbytes, cudaStream_t s) {
  size_t n16 = (nbytes + 15) / 16; if (!n16) return;
  int grid = (int)std::min<size_t>((n16 + 255) / 256, 148 * 2);
  host_fetch_kernel<<<grid, 256, 0, s>>>((uint4*)dst_dev, (const uint4*)src_host_mapped, n16); IRF_LAUNCHED();
}
void launch_de_check(const u32* list, int n, const u32* act, const u32* u_len, int* status, int* err, cudaStream_t s) {
  if (n > 0) { de_check_kernel<<<(n + 255) / 256, 256, 0, s>>>(list, n, act, u_len, status, err); IRF_LAUNCHED(); }
}
void launch_crc_check(const uint8_t* U, const u64* u_off, const u32* u_len, const u32* crc_expected, int nblk, int* status, int* err, cudaStream_t s, const u32* list) {
  static int sms_of[64] = {0};
  static const CrcMul M = [] { CrcMul m; u32 p = 0x40000000u;   // x^1; square up to x^(8 * 2048) = x^(2^14), then keep squaring
    for (int i = 0; i < 14; i++) p = crc_mulmod(p, p);
    for (int k = 0; k < 5; k++) { m.x[k] = p; p = crc_mulmod(p, p); } return m; }();
  if (nblk <= 0) return;
  constexpr int smem = 4 * 256 * 32 * 4;
  int dev = 0; IRF_CUDA(cudaGetDevice(&dev)); if (dev < 0 || dev >= 64) throw std::string("device ordinal beyond 63 is not supported");
  if (!sms_of[dev]) { IRF_CUDA(cudaFuncSetAttribute(crc_check_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); IRF_CUDA(cudaDeviceGetAttribute(&sms_of[dev], cudaDevAttrMultiProcessorCount, dev)); }
  const int wpb = CRC_THREADS / 32, grid = std::min((nblk + wpb - 1) / wpb, sms_of[dev]);
  crc_check_kernel<<<grid, CRC_THREADS, smem, s>>>(U, u_off, u_len, crc_expected, nblk, M, status, err, list); IRF_LAUNCHED();
}

// ------------------------------------------------------------------------------------------------
// S2: record framing.  Records straddle BGZF blocks and each starts with its own block_size, so record
// starts form a serial chain.  We guess each block's entry (first record start at
// or after the block's first inflated byte) from the record layout, walk all blocks in parallel, then check
// exit[b] == entry[b+1]; with entry[0] exact this proves the whole chain by induction.  Wrong or missing
// guesses are repaired from the predecessor's exit and re-walked until none remain.
__device__ bool plausible_record(const uint8_t* U, u64 off, u64 Uend, int n_ref, const int32_t* ref_len, u64& next) {
  if (off + 36 > Uend) return false;
  Core c; load_core(U, off, c);
  if (c.block_size < 32 || c.block_size >= (1 << 26)) return false;
  next = off + 4 + (u64)c.block_size;
  if (next > Uend) return false;
  if (c.refID < -1 || c.refID >= n_ref || c.next_refID < -1 || c.next_refID >= n_ref) return false;
  if (c.pos < -1 || c.next_pos < -1 || c.l_seq < 0 || c.l_read_name < 1) return false;
  if (c.refID >= 0 && c.pos > ref_len[c.refID]) return false;
  u64 need = 32ull + c.l_read_name + 4ull * c.n_cigar_op + (u64)(c.l_seq + 1) / 2 + (u64)c.l_seq;
  if (need > (u64)c.block_size) return false;
  const uint8_t* nm = U + off + 36;
  if (nm[c.l_read_name - 1] != 0) return false;
  for (u32 i = 0; i + 1 < c.l_read_name; i++) { uint8_t ch = nm[i]; if (ch < 33 || ch > 126) return false; }
  const uint8_t* cg = nm + c.l_read_name;
  for (u32 i = 0; i < c.n_cigar_op && i < 8; i++) if ((ld32(cg + 4 * i) & 15u) > 8u) return false;
  return true;
}

__global__ void frame_guess_kernel(const uint8_t* __restrict__ U, const u64* __restrict__ u_off, const u32* __restrict__ u_len, int nblk, u64 Uend, u64 rec0, int rec0_valid,
                                   int n_ref, const int32_t* __restrict__ ref_len, u64* __restrict__ entry, int* __restrict__ first) {
  int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (b >= nblk) return;
  u64 o = u_off[b], e = o + u_len[b];
  if (rec0_valid && o <= rec0) { if (lane == 0) { entry[b] = rec0; atomicMin(first, b); } return; }
  u64 found = ~0ull;
  for (u64 base = o; base < e; base += 32) {
    u64 c = base + lane, nx = 0, nx2 = 0;
    bool ok = c < e && plausible_record(U, c, Uend, n_ref, ref_len, nx);
    if (ok && nx != Uend) ok = plausible_record(U, nx, Uend, n_ref, ref_len, nx2);
    u32 m = __ballot_sync(0xffffffffu, ok);
    if (m) { found = base + (u64)(__ffs(m) - 1); break; }
  }
  if (lane == 0) { entry[b] = found; if (found != ~0ull) atomicMin(first, b); }
}

// Walks the records that start inside block b; exit = first record start at or beyond the block end.
__device__ void walk_count_block(const uint8_t* U, u64 e, u64 blk_end, u64 Uend, u64& exit_off, u32& n_rec, u32& n_cand, int& ferr, u64* slots) {
  n_rec = 0; n_cand = 0; ferr = 0;
  while (e < blk_end) {
    if (e + 36 > Uend) { ferr = DERR_TRUNCATED; e = Uend; break; }
    Core c; load_core(U, e, c);
    if (c.block_size < 32 || e + 4 + (u64)c.block_size > Uend) { ferr = c.block_size < 32 ? DERR_BAD_RECORD : DERR_TRUNCATED; e = Uend; break; }
    if (slots && n_rec < REC_SLOTS) slots[n_rec] = e;   // the record starts of this block, gathered into the dense list afterwards (no second walk)
    n_rec++;
    if (!(c.flag & 0x4) && c.refID >= 0 && (c.flag & 0x1) && !(c.flag & 0x8) && c.next_refID == c.refID) n_cand++;
    e += 4 + (u64)c.block_size;
  }
  exit_off = e;
}
__global__ void walk_count_kernel(const uint8_t* __restrict__ U, const u64* __restrict__ u_off, const u32* __restrict__ u_len, int nblk, u64 Uend,
                                  const u64* __restrict__ entry, u64* __restrict__ exit_off, u32* __restrict__ n_rec, u32* __restrict__ n_cand,
                                  const uint8_t* __restrict__ mask, uint8_t* __restrict__ ferr, u64* __restrict__ slots) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nblk || (mask && !mask[b])) return;
  u64 e = entry[b];
  if (e == ~0ull) { exit_off[b] = ~0ull; n_rec[b] = 0; n_cand[b] = 0; ferr[b] = 0; return; }
  u64 x; u32 nr, nc; int fe; walk_count_block(U, e, u_off[b] + u_len[b], Uend, x, nr, nc, fe, slots ? slots + (size_t)b * REC_SLOTS : nullptr);
  exit_off[b] = x; n_rec[b] = nr; n_cand[b] = nc; ferr[b] = (uint8_t)fe;
}
// Repair step, in two kernels so that no thread reads an entry another thread is rewriting.  Block b is "consistent" when
// its entry equals its predecessor's exit (blocks up to the first one that has an entry at all start the chain and count
// as consistent).  An inconsistent block is repaired (entry := predecessor's exit, re-walked) only when the predecessor is
// consistent: an exit computed from an entry that is itself unconfirmed is garbage, and copying it forward overwrites the
// next block's (usually correct) guess -- one wrong guess then chases a wave of garbage through every later block, one
// block per round (seen: 5500 rounds in a 50M-pair file).  The first inconsistent block always has a consistent
// predecessor, so every round makes progress, and isolated wrong guesses anywhere in the file are repaired in the same round.
__global__ void frame_mismatch_kernel(int nblk, const u64* __restrict__ entry, const u64* __restrict__ exit_off, const int* __restrict__ first, uint8_t* __restrict__ mism) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nblk) return;
  uint8_t m = 0;
  if (b > *first) { u64 want = exit_off[b - 1]; m = want == ~0ull || entry[b] != want; }
  mism[b] = m;
}
__global__ void frame_mark_kernel(int nblk, u64* __restrict__ entry, const u64* __restrict__ exit_off, const uint8_t* __restrict__ mism, uint8_t* __restrict__ mask, u32* __restrict__ n_bad) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nblk) return;
  uint8_t m = 0;
  if (b >= 1 && mism[b]) { atomicAdd(n_bad, 1u); if (!mism[b - 1]) { entry[b] = exit_off[b - 1]; m = 1; } }
  mask[b] = m;
}
__global__ void frame_totals_kernel(int nblk, const u32* __restrict__ n_rec, const u32* __restrict__ n_cand, const uint8_t* __restrict__ ferr, const u64* __restrict__ exit_off, int* __restrict__ err, u64* __restrict__ totals) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  u64 r = b < nblk ? n_rec[b] : 0, c = b < nblk ? n_cand[b] : 0;
  if (b < nblk && ferr[b]) atomicOr(err, (int)ferr[b]);
  if (b < nblk && exit_off[b] == ~0ull) atomicOr(err, DERR_BAD_RECORD);  // chain never reached this block
  for (int o = 16; o; o >>= 1) { r += __shfl_down_sync(0xffffffffu, r, o); c += __shfl_down_sync(0xffffffffu, c, o); }
  if ((threadIdx.x & 31) == 0) { if (r) atomicAdd(&totals[0], r); if (c) atomicAdd(&totals[1], c); }
}

void launch_frame_guess(const FrameArgs& a, cudaStream_t s) {
  long threads = (long)a.nblk * 32; int grid = (int)((threads + 255) / 256);
  if (grid) { IRF_CUDA(cudaMemsetAsync(a.first, 0x7f, sizeof(int), s)); frame_guess_kernel<<<grid, 256, 0, s>>>(a.U, a.u_off, a.u_len, a.nblk, a.Uend, a.rec0, a.rec0_valid, a.n_ref, a.ref_len, a.entry, a.first); IRF_LAUNCHED(); }
}
void launch_walk_count(const FrameArgs& a, const uint8_t* mask, cudaStream_t s) {
  int grid = (a.nblk + 63) / 64;   // 64-thread CTAs: a chunk has only a few thousand blocks, spread them over all SMs
  if (grid) { walk_count_kernel<<<grid, 64, 0, s>>>(a.U, a.u_off, a.u_len, a.nblk, a.Uend, a.entry, a.exit_off, a.n_rec, a.n_cand, mask, a.ferr, a.slots); IRF_LAUNCHED(); }
}
// Dense record list from the per-block slot regions the walk filled: one warp per BGZF block, coalesced on both sides.
__global__ void gather_records_kernel(const u64* __restrict__ slots, const u32* __restrict__ n_rec, const u64* __restrict__ rec_base, int nblk, u64* __restrict__ rec_off) {
  const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (b >= nblk) return;
  const u32 n = min(n_rec[b], REC_SLOTS); const u64 at = rec_base[b]; const u64* sp = slots + (size_t)b * REC_SLOTS;
  for (u32 i = lane; i < n; i += 32) rec_off[at + i] = sp[i];
}
void launch_gather_records(const FrameArgs& a, const u64* rec_base, u64* rec_off, cudaStream_t s) {
  long threads = (long)a.nblk * 32; int grid = (int)((threads + 255) / 256);
  if (grid) { gather_records_kernel<<<grid, 256, 0, s>>>(a.slots, a.n_rec, rec_base, a.nblk, rec_off); IRF_LAUNCHED(); }
}
void launch_frame_mark(const FrameArgs& a, uint8_t* mism, uint8_t* mask, u32* n_bad, cudaStream_t s) {
  int grid = (a.nblk + 255) / 256;
  if (grid) { frame_mismatch_kernel<<<grid, 256, 0, s>>>(a.nblk, a.entry, a.exit_off, a.first, mism); IRF_LAUNCHED();
    frame_mark_kernel<<<grid, 256, 0, s>>>(a.nblk, a.entry, a.exit_off, mism, mask, n_bad); IRF_LAUNCHED(); }
}
void launch_frame_totals(const FrameArgs& a, u64* totals, cudaStream_t s) {
  int grid = (a.nblk + 255) / 256;
  if (grid) { frame_totals_kernel<<<grid, 256, 0, s>>>(a.nblk, a.n_rec, a.n_cand, a.ferr, a.exit_off, a.err, totals); IRF_LAUNCHED(); }
}

// ------------------------------------------------------------------------------------------------
// S3: per-fragment processing
__device__ __forceinline__ u32 ht_hash(u64 k) { return (u32)(mix64(k) >> 17); }
// insert-or-find in a keys-only open-addressing table; returns the slot or 0xffffffff when full
__device__ u32 ht_upsert(u64* keys, u32 mask, u64 key, int* err) {
  u32 s = ht_hash(key) & mask;
  for (u32 probe = 0; probe < 4096; probe++) {
    u64 k = keys[s];
    if (k == key) return s;
    if (k == JKEY_EMPTY) { u64 old = atomicCAS(&keys[s], JKEY_EMPTY, key); if (old == JKEY_EMPTY || old == key) return s; }
    s = (s + 1) & mask;
  }
  atomicOr(err, DERR_JUNC_FULL); return 0xffffffffu;
}
__device__ u32 ht_find(const u64* keys, u32 mask, u64 key) {
  u32 s = ht_hash(key) & mask;
  for (u32 probe = 0; probe < 4096; probe++) { u64 k = keys[s]; if (k == key) return s; if (k == JKEY_EMPTY) return 0xffffffffu; s = (s + 1) & mask; }
  return 0xffffffffu;
}

// cigar2block as a streaming iterator (upstream BAM2blocks::cigar2block): M,D,=,X extend the
// current block; N closes it and skips; I,S,H,P are ignored.  Yields exactly the (start,len) pairs the
// reference's rStarts/rLens vectors would hold, including a zero-length first/last block around a leading/trailing N.
struct BlockIter {
  const uint8_t* cig; u32 n, i; int relpos, cur_start, cur_len; bool inBlock, done;
  __device__ void init(const uint8_t* c, u32 n_) { cig = c; n = n_; i = 0; relpos = 0; cur_start = 0; cur_len = 0; inBlock = true; done = false; }
  __device__ bool next(int& s, int& l) {
    if (done) return false;
    while (i < n) {
      u32 c = ld32(cig + 4ull * i); i++; int op = (int)(c & 15u), len = (int)(c >> 4);
      if (op == 0 || op == 2 || op == 7 || op == 8) { cur_len += len; relpos += len; inBlock = true; }
      else if (op == 3) {
        relpos += len;
        if (inBlock) { s = cur_start; l = cur_len; cur_start = relpos; cur_len = 0; inBlock = false; return true; }
        cur_start = relpos;
      }
    }
    done = true; s = cur_start; l = cur_len; return true;
  }
};
__device__ void cigar_scan(const uint8_t* cig, u32 n, int& genome_len, int& first_len) {
  BlockIter it; it.init(cig, n); int s, l; bool first = true; first_len = 0;
  while (it.next(s, l)) { if (first) { first_len = l; first = false; } }
  genome_len = it.relpos;
}

// Receives the blocks of one FragmentBlocks value in order and applies every processor's ProcessBlocks.
struct FragSink {
  const DevTables& T; int chr, dir; u32 read_start; bool have_prev; u32 prev_end; int prev_len; u64 nuc;
  u32* dump_base; u32 dump_n, dump_nb[2]; int dump_read;
  int32_t* tile; unsigned long long* tile_base;   // the CTA's shared-memory depth tile ([2][DEPTH_TILE]) and the table index each strand's window starts at; null = no tile
  __device__ FragSink(const DevTables& t) : T(t), nuc(0), dump_base(nullptr), tile(nullptr), tile_base(nullptr) {}
  __device__ void begin_fragment(int chr_, int dir_, u32 reserve_blocks) {
    chr = chr_; dir = dir_; dump_base = nullptr;
    if (T.dump) {
      u64 words = 6 + 2ull * reserve_blocks; u64 at = atomicAdd(T.dump_cursor, words);
      if (at + words <= T.dump_cap) { dump_base = T.dump + at; dump_base[0] = (u32)words; dump_n = 6; dump_nb[0] = dump_nb[1] = 0; dump_read = -1; for (u64 w = 1; w < words; w++) dump_base[w] = 0; }
      else atomicOr(T.err, DERR_DUMP_FULL);
    }
  }
  __device__ __forceinline__ u64 bucket_of(u32 pos) const { u32 b = pos >> BUCKET_SHIFT, nb = T.bucket_n[chr]; return T.bucket_base[chr] + (b < nb ? b : nb - 1); }
  __device__ void begin_read(u32 rs) { read_start = rs; have_prev = false; dump_read++; }
  __device__ void end_fragment(int readCount) { if (dump_base) { dump_base[1] = (u32)chr; dump_base[2] = (u32)dir; dump_base[3] = (u32)readCount; dump_base[4] = dump_nb[0]; dump_base[5] = dump_nb[1]; } }

  // JunctionCount::ProcessBlocks (upstream ReadBlockProcessor.cpp)
  __device__ void junction(u32 left, u32 right) {
    u64 base = T.chr_goff[chr]; u64 gl = base + left, gr = base + right; u64 gap = (u64)right - (u64)left;
    if (right < left || gap >= (1ull << JKEY_GAP_BITS) || gr >= (1ull << 36)) { atomicOr(T.err, DERR_JKEY_RANGE); return; }
    // neighbouring fragments of a coordinate-sorted file cross the same junctions: the lanes of a warp that are here with the same
    // key and strand elect one lane, which does the table look-up and adds the whole group's count (one atomic instead of up to 32)
    const u64 key = jkey_make(gl, gap);
    const u32 peers = __match_any_sync(__activemask(), key ^ ((u64)dir << 63)); const u32 n = (u32)__popc(peers);
    if ((int)(threadIdx.x & 31) != __ffs(peers) - 1) return;
    u32 s = ht_upsert(T.jc_keys, T.jc_mask, key, T.err); if (s != 0xffffffffu) atomicAdd(&T.jc_cnt[2ull * s + dir], n);
    s = ht_upsert(T.jl_keys, T.jl_mask, gl, T.err); if (s != 0xffffffffu) atomicAdd(&T.jl_cnt[2ull * s + dir], n);
    s = ht_upsert(T.jr_keys, T.jr_mask, gr, T.err); if (s != 0xffffffffu) atomicAdd(&T.jr_cnt[2ull * s + dir], n);
  }
  // SpansPoint::ProcessBlocks (row a8): points p with start < p < end
  __device__ void spans(u32 start, u32 end) {
    Range r = T.sp_range[chr]; if (r.begin == r.end) return;
    u32 lo = T.sp_bidx[bucket_of(start)];                                   // first point >= the bucket start
    while (lo < r.end && T.sp_pos[lo] <= start) lo++;                        // upper_bound(start)
    for (; lo < r.end && T.sp_pos[lo] < end; lo++) atomicAdd(&T.sp_count[(u64)dir * T.n_sp + lo], 1u);
  }
  // CoverageBlocks::ProcessBlocks + CoverageBlock::RecordCover (row a9): +1 at start, -1 at end, clipped per measured segment
  __device__ void coverage(u32 start, u32 end) {
    Range r = T.seg_range[chr]; if (r.begin == r.end) return;
    u32 lo = T.seg_bidx[bucket_of(start)];                                  // first segment with end > the bucket start
    while (lo < r.end && T.seg_end[lo] <= start) lo++;                       // first segment with end > start
    for (; lo < r.end; lo++) {
      u32 ss = T.seg_start[lo]; if (ss >= end) break;
      u32 a = max(start, ss), b = min(end, T.seg_end[lo]);
      if (a < b) { const u64 g = (u64)dir * T.P + T.seg_coff[lo]; depth_add(g + (a - ss), 1); depth_add(g + (b - ss), -1); }
    }
  }
  // One +1 / -1 of the difference table (index g into the [2][P] table).  With a tile, the first counter a CTA touches on a strand fixes that
  // strand's window (the 2048-aligned block of the table around it: the input is coordinate-sorted, so the CTA's other fragments land in or
  // next to it); counters inside the window are summed in shared memory and reach the table once, in the flush at the end of the kernel,
  // counters outside go to the table directly.
  __device__ __forceinline__ void depth_add(u64 g, int v) {
    if (tile) {
      const u64 G = g & ~(u64)(DEPTH_TILE - 1); unsigned long long cur = tile_base[dir];
      if (cur == ~0ull) { const unsigned long long old = atomicCAS(&tile_base[dir], ~0ull, (unsigned long long)G); cur = old == ~0ull ? G : old; }
      if (cur == G) { atomicAdd(&tile[dir * DEPTH_TILE + (int)(g & (DEPTH_TILE - 1))], v); return; }
    }
    atomicAdd(T.diff + g, v);
  }
  __device__ __noinline__ void block(int s, int l) {   // one body shared by every call site: the fully inlined version thrashed the instruction cache
    u32 a = read_start + (u32)s, e = a + (u32)l;
    nuc += (u32)l;
    if (have_prev && prev_len >= T.rules.junc_min_flank && l >= T.rules.junc_min_flank) junction(prev_end, a);
    have_prev = true; prev_end = e; prev_len = l;
    if ((u32)l > T.rules.span_left + T.rules.span_right) spans(a + T.rules.span_left, e - T.rules.span_right);
    if (l > 0) coverage(a, e);
    if (dump_base && dump_n + 2 <= dump_base[0]) { dump_base[dump_n++] = a; dump_base[dump_n++] = e; dump_nb[dump_read]++; }
  }
  // FragmentsInChr / FragmentsInROI ::ProcessBlocks (row a13)
  __device__ void fragment_counts(u32 fs, u32 fe) {
    { u32 key = (u32)dir * (u32)T.n_ref + (u32)chr; u32 act = __activemask(); int first = __ffs(act) - 1;   // neighbouring records share chromosome and strand:
      u32 key0 = __shfl_sync(act, key, first); u32 same = __ballot_sync(act, key == key0);                  // one atomic for everybody who agrees with the first lane
      if (key != key0) atomicAdd(&T.chr_count[key], 1u); else if ((int)(threadIdx.x & 31) == first) atomicAdd(&T.chr_count[key], (u32)__popc(same)); }
    Range r = T.roi_range[chr]; if (r.begin == r.end) return;
    u32 lo = T.roi_bidx[bucket_of(fs)];
    while (lo < r.end && T.roi_end[lo] <= fs) lo++;                          // first ROI with end > fs
    if (lo < r.end && T.roi_start[lo] <= fs && fe <= T.roi_end[lo]) atomicAdd(&T.roi_count[(u64)dir * T.n_roi + lo], 1u);
  }
};

struct LocalStats { u32 single, paired, shortp, inter, longp, skipped, error; };

// BAM2blocks::processSingle (row a4)
__device__ void process_single(const DevTables& T, const uint8_t* U, u64 off, const Core& c, FragSink& sink) {
  int dir = !(c.flag & 0x10);
  const uint8_t* cig = U + off + 36 + c.l_read_name;
  sink.begin_fragment(c.refID, dir, c.n_cigar_op + 1); sink.begin_read((u32)c.pos);
  BlockIter it; it.init(cig, c.n_cigar_op); int s, l;
  while (it.next(s, l)) sink.block(s, l);
  sink.end_fragment(1);
  sink.fragment_counts((u32)c.pos, (u32)(c.pos + it.relpos));
}

// BAM2blocks::processPair (row a4; merge rule [C-10]).  Record 1 is the left mate (or level, earlier in the file).
__device__ void process_pair(const DevTables& T, const uint8_t* U, u64 off1, const Core& c1, u64 off2, const Core& c2, FragSink& sink, LocalStats& st) {
  int dir = (c1.flag & 0x40) ? !(c1.flag & 0x10) : !(c1.flag & 0x20);
  const uint8_t* cig1 = U + off1 + 36 + c1.l_read_name; const uint8_t* cig2 = U + off2 + 36 + c2.l_read_name;
  int len1, len2, fb1, fb2; cigar_scan(cig1, c1.n_cigar_op, len1, fb1); cigar_scan(cig2, c2.n_cigar_op, len2, fb2);
  int r1pos = c1.pos, r2pos = c2.pos; int readCount; int merge_at = -1;
  if (r1pos + len1 < r2pos) { st.longp++; readCount = 2; }
  else if (r1pos + len1 >= r2pos + len2) { st.shortp++; readCount = 1; }
  else {
    st.inter++; readCount = 1;
    BlockIter it; it.init(cig1, c1.n_cigar_op); int s, l, i = 0;
    while (it.next(s, l)) { if (r1pos + s + l >= r2pos) { if (r1pos + s <= r2pos) merge_at = i; else readCount = 2; break; } i++; }
  }
  sink.begin_fragment(c1.refID, dir, c1.n_cigar_op + c2.n_cigar_op + 2);
  sink.begin_read((u32)r1pos);
  BlockIter it; it.init(cig1, c1.n_cigar_op); int s, l, i = 0;
  if (merge_at < 0) { while (it.next(s, l)) sink.block(s, l); }
  else {
    while (it.next(s, l)) { if (i == merge_at) { sink.block(s, r2pos - r1pos - s + fb2); break; } sink.block(s, l); i++; }
    BlockIter it2; it2.init(cig2, c2.n_cigar_op); bool first = true;
    while (it2.next(s, l)) { if (first) { first = false; continue; } sink.block(s + r2pos - r1pos, l); }
    len1 = r2pos - r1pos + len2;
  }
  u32 fe = (u32)(r1pos + len1);
  if (readCount == 2) {
    sink.begin_read((u32)r2pos);
    BlockIter it2; it2.init(cig2, c2.n_cigar_op);
    while (it2.next(s, l)) sink.block(s, l);
    if ((u32)(r2pos + len2) > fe) fe = (u32)(r2pos + len2);
  }
  sink.end_fragment(readCount);
  sink.fragment_counts((u32)r1pos, fe);
}

// Mate pairing by device hash keyed on read name (+ reference and the two mate positions so multi-mapped
// templates keep their loci apart).  One 64-bit word per waiting record: tag(24) | record offset + 1 (40).
constexpr u64 PAIR_EMPTY = 0ull, PAIR_TOMB = ~0ull;
__device__ u64 pair_key(const uint8_t* U, u64 off, const Core& c) {
  u32 lo = (u32)min(c.pos, c.next_pos), hi = (u32)max(c.pos, c.next_pos);
  u64 h = mix64(((u64)lo << 32 | hi) ^ ((u64)(u32)c.refID * 0x9E3779B97F4A7C15ull));
  const uint8_t* nm = U + off + 36; u32 n = c.l_read_name - 1, i = 0;
  for (; i + 4 <= n; i += 4) h = (h ^ ld32(nm + i)) * 0x100000001b3ull + 0x632be59bd9b4e019ull;
  u32 tail = 0; for (u32 j = 0; i + j < n; j++) tail |= (u32)nm[i + j] << (8 * j);
  return mix64(h ^ tail ^ ((u64)n << 56));
}
__device__ bool is_mate(const uint8_t* U, u64 off, const Core& c, u64 off2, Core& c2) {
  load_core(U, off2, c2);
  if (c2.refID != c.refID || c2.l_read_name != c.l_read_name || c2.pos != c.next_pos || c2.next_pos != c.pos) return false;
  const uint8_t* a = U + off + 36; const uint8_t* b = U + off2 + 36;
  for (u32 i = 0; i + 1 < c.l_read_name; i++) if (a[i] != b[i]) return false;
  return true;
}
// Returns true and the mate's offset when this record completes a pair; otherwise the record is left waiting.
__device__ bool pair_lookup_or_wait(const DevTables& T, u64* ptab, u32 pmask, const uint8_t* U, u64 off, const Core& c, u64& mate_off, Core& mate) {
  u64 h = pair_key(U, off, c); u64 mine = ((h >> 40) << 40) | (off + 1); if (mine == PAIR_TOMB) mine ^= 1ull << 40;
  u32 s = (u32)h & pmask;
  for (u32 probe = 0; probe < (1u << 20); probe++) {
    u64 cur = ptab[s];
    for (;;) {
      if (cur == PAIR_EMPTY) { u64 old = atomicCAS(&ptab[s], PAIR_EMPTY, mine); if (old == PAIR_EMPTY) return false; cur = old; continue; }
      if (cur != PAIR_TOMB && (cur >> 40) == (mine >> 40)) {
        u64 o2 = (cur & ((1ull << 40) - 1)) - 1;
        if (o2 != off && is_mate(U, off, c, o2, mate)) { u64 old = atomicCAS(&ptab[s], cur, PAIR_TOMB); if (old == cur) { mate_off = o2; return true; } cur = old; continue; }
      }
      break;
    }
    s = (s + 1) & pmask;
  }
  atomicOr(T.err, DERR_PAIR_FULL); return false;
}

// Second framing pass: every lane re-walks its BGZF block and lists the record offsets at rec_base[b] + i.
__global__ void walk_list_kernel(const uint8_t* __restrict__ U, const u64* __restrict__ u_off, const u32* __restrict__ u_len, int nblk, const u64* __restrict__ entry,
                                 const u64* __restrict__ rec_base, u64* __restrict__ rec_off) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nblk) return;
  u64 e = entry[b], blk_end = u_off[b] + u_len[b], at = rec_base[b];
  while (e < blk_end) { rec_off[at++] = e; e += 4 + (u64)ld32(U + e); }
}
// exclusive scan of per-block record counts (single CTA; a few hundred thousand values at most)
__global__ void __launch_bounds__(1024) scan_counts_kernel(const u32* __restrict__ in, u64* __restrict__ out, int n) {
  __shared__ u64 sh[1024]; __shared__ u64 carry; if (threadIdx.x == 0) carry = 0; __syncthreads();
  for (int base = 0; base < n; base += 1024) {
    int i = base + threadIdx.x; u64 v = i < n ? in[i] : 0; sh[threadIdx.x] = v; __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) { u64 t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0; __syncthreads(); sh[threadIdx.x] += t; __syncthreads(); }
    u64 incl = sh[threadIdx.x], c = carry; if (i < n) out[i] = c + incl - v; __syncthreads();
    if (threadIdx.x == 1023) carry = c + incl; __syncthreads();
  }
}

// BAM2blocks::processAll (row a2) in two passes so that every warp does one kind of work:
//   pass 1 (lane per record): skip / single / pair decision and mate pairing; completed pairs and singles are appended to
//           dense work lists (warp-aggregated appends);
//   pass 2 (lane per fragment): BAM2blocks::processPair / processSingle and every processor's ProcessBlocks.
// (The first version did both in one kernel: half the lanes of a warp parked a mate and idled while the other half
//  processed pairs, and the large divergent body thrashed the instruction cache — profiles/r1.)
__device__ __forceinline__ u64 warp_append(u64* counter, bool want) {
  u32 m = __ballot_sync(0xffffffffu, want); if (!m) return 0;
  u32 lane = threadIdx.x & 31; int leader = __ffs(m) - 1; u64 base = 0;
  if ((int)lane == leader) base = atomicAdd(counter, (u64)__popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  return base + (u64)__popc(m & ((1u << lane) - 1u));
}
__device__ __forceinline__ void warp_add_stat(u64* slot, u64 v) {
  for (int o = 16; o; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0 && v) atomicAdd(slot, v);
}
// carry_base: records at offsets >= carry_base are waiting mates carried over from the previous segment of a segmented (streamed) run;
// they come EARLIER in the file than every record of this segment (the tie-break for level mates) and were counted when first seen
// (count_stats = 0 for the launch that parks them again).
// Adjacent-mate path (name-collated input, STAR's pipe).  L(i) = records i and i+1 are both pairing candidates and mates of
// each other.  Record i is the LEFT half of an adjacent pair when L(i) && !L(i-1), the RIGHT half when L(i-1) && !L(i-2); the two rules
// name the same pairs from both sides, so the left lane emits the pair and the right lane does nothing, and neither touches the hash
// table.  Everything else (mates apart, the third and later records of a run of mutual mates) pairs through the table as before, so the
// result does not depend on whether this path is on: it is a matter of speed only.
__device__ __forceinline__ bool pair_candidate(const DevTables& T, const Core& c) {
  return !(c.flag & 0x4) && c.refID >= 0 && !(c.flag & T.rules.skip_flags) && c.refID < T.n_ref && (c.flag & 0x1) && !(c.flag & 0x8) && c.next_refID == c.refID;
}
__device__ bool adjacent_mates(const DevTables& T, const uint8_t* U, const u64* rec_off, u64 nrec, u64 i) {   // L(i); i may be "negative" (wrapped)
  if (i >= nrec || i + 1 >= nrec) return false;
  const u64 a = rec_off[i], b = rec_off[i + 1]; Core ca, cb; load_core(U, a, ca);
  if (!pair_candidate(T, ca)) return false;
  return is_mate(U, a, ca, b, cb) && pair_candidate(T, cb);
}
__global__ void __launch_bounds__(256) pair_records_kernel(DevTables T, const uint8_t* __restrict__ U, const u64* __restrict__ rec_off, u64 nrec, u64* __restrict__ ptab, u32 pmask,
                                                            u64* __restrict__ pair_list, u64* __restrict__ single_list, u64* __restrict__ list_counts, u64 carry_base, int count_stats, int adjacent) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  bool is_single = false, is_pair = false; u64 e = 0, o1 = 0, o2 = 0; u32 skipped = 0, cand = 0, adj = 0;
  bool Li = false, Lm1 = false, Lm2 = false;
  if (adjacent) {   // (uniform across the grid) L of this lane's record; the two to the left come from the neighbouring lanes
    const u32 lane = threadIdx.x & 31;
    Li = adjacent_mates(T, U, rec_off, nrec, i);
    bool a = false, b = false; if (lane == 0) { a = adjacent_mates(T, U, rec_off, nrec, i - 1); b = adjacent_mates(T, U, rec_off, nrec, i - 2); }   // i - 1 / i - 2 wrap for i < 2: out of range, false
    const bool up1 = __shfl_up_sync(0xffffffffu, Li, 1), up2 = __shfl_up_sync(0xffffffffu, Li, 2), a0 = __shfl_sync(0xffffffffu, a, 0);
    Lm1 = lane == 0 ? a : up1; Lm2 = lane == 0 ? b : lane == 1 ? a0 : up2;
  }
  if (i < nrec) {
    e = rec_off[i]; Core c; load_core(U, e, c);
    if ((c.flag & 0x4) || c.refID < 0 || (c.flag & T.rules.skip_flags)) skipped = 1;   // [C-6] unmapped records, and what rule_skip_flags adds
    else if (c.refID >= T.n_ref) atomicOr(T.err, DERR_REFID);
    else if (!(c.flag & 0x1) || (c.flag & 0x8) || c.next_refID != c.refID) is_single = true;   // [C-7]
    else if (Li && !Lm1) {   // left half of an adjacent pair: the mate is the next record (later in the file: it wins only with the smaller pos)
      cand = 1; adj = 1; is_pair = true; const u64 mo = rec_off[i + 1]; Core m; load_core(U, mo, m);
      const bool first_is_me = c.pos <= m.pos; o1 = first_is_me ? e : mo; o2 = first_is_me ? mo : e;
    }
    else if (Lm1 && !Lm2) cand = 1;   // right half: emitted by the lane to the left
    else {
      u64 mo; Core m; cand = 1;
      if (pair_lookup_or_wait(T, ptab, pmask, U, e, c, mo, m)) {
        is_pair = true;
        const bool e_carry = e >= carry_base, m_carry = mo >= carry_base;
        const bool e_earlier = e_carry != m_carry ? e_carry : e < mo;
        bool first_is_me = c.pos < m.pos || (c.pos == m.pos && e_earlier);    // read 1 = left mate; level starts: the earlier record
        o1 = first_is_me ? e : mo; o2 = first_is_me ? mo : e;
      }
    }
  }
  u64 at = warp_append(&list_counts[0], is_pair); if (is_pair) { pair_list[2 * at] = o1; pair_list[2 * at + 1] = o2; }
  at = warp_append(&list_counts[1], is_single); if (is_single) single_list[at] = e;
  if (!count_stats) return;
  warp_add_stat(&T.stats[ST_SKIPPED], skipped); warp_add_stat(&T.stats[ST_CAND], cand); if (adjacent) warp_add_stat(&T.stats[ST_ADJ], adj);
  if (threadIdx.x == 0) { u64 base = (u64)blockIdx.x * blockDim.x; u64 n = nrec > base ? min((u64)blockDim.x, nrec - base) : 0; if (n) atomicAdd(&T.stats[ST_RECORDS], n); }
}
// Shared-memory depth tile of a counting CTA (DevTables::depth_tile): zeroed at the start, flushed at the end - with one red.global per non-zero
// counter, or by the TMA unit (one bulk reduce per strand: the "TMA-staged spill" of the difference array).
struct DepthTile {
  int32_t* tile; unsigned long long* base;
  __device__ void begin(const DevTables& T, int32_t* sh_tile, unsigned long long* sh_base) {
    tile = T.depth_tile ? sh_tile : nullptr; base = sh_base;
    if (!tile) return;
    for (int i = threadIdx.x; i < 2 * DEPTH_TILE; i += blockDim.x) tile[i] = 0;
    if (threadIdx.x < 2) base[threadIdx.x] = ~0ull;
    __syncthreads();
  }
  __device__ void attach(FragSink& s) const { s.tile = tile; s.tile_base = base; }
  __device__ void flush(const DevTables& T) {
    if (!tile) return;
    __syncthreads();
    if (T.depth_tile == 2) {
      if (threadIdx.x == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the tile was written through the generic proxy; the bulk reduce reads it through the async proxy
        for (int d = 0; d < 2; d++) {
          const unsigned long long G = base[d]; if (G == ~0ull) continue;
          const u64 left = 2 * T.P - G; const uint32_t bytes = (uint32_t)(((left < (u64)DEPTH_TILE ? left : (u64)DEPTH_TILE) * 4 + 15) & ~15ull);   // the table has 256 bytes of slack behind it
          const uint32_t src = (uint32_t)__cvta_generic_to_shared(tile + d * DEPTH_TILE);
          asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.s32 [%0], [%1], %2;" :: "l"(T.diff + G), "r"(src), "r"(bytes) : "memory");
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the shared memory may go once the TMA unit has read it
      }
    } else {
      for (int d = 0; d < 2; d++) {
        const unsigned long long G = base[d]; if (G == ~0ull) continue;
        for (int i = threadIdx.x; i < DEPTH_TILE; i += blockDim.x) { const int v = tile[d * DEPTH_TILE + i]; if (v) atomicAdd(T.diff + G + i, v); }
      }
    }
  }
};
// fr: fragments per thread (a CTA with a depth tile takes several consecutive 128-fragment slices so that zeroing and flushing the tile pay off)
__global__ void __launch_bounds__(128) process_pairs_kernel(DevTables T, const uint8_t* __restrict__ U, const u64* __restrict__ pair_list, u64 npairs, const u64* __restrict__ npairs_dev, int fr) {
  extern __shared__ __align__(16) int32_t sh_tile[];   // [2][DEPTH_TILE] when the depth tile is on, nothing otherwise (16 KB of static shared memory cost the default path 10 %)
  __shared__ unsigned long long sh_base[2];
  if (npairs_dev) npairs = min(npairs, *npairs_dev);   // the grid was sized from an upper bound; the pairing kernel left the actual count on the device
  if ((u64)blockIdx.x * fr * blockDim.x >= npairs) return;
  DepthTile dt; dt.begin(T, sh_tile, sh_base);
  FragSink sink(T); dt.attach(sink); LocalStats st = {0, 0, 0, 0, 0, 0, 0};
  for (int k = 0; k < fr; k++) {
    const u64 i = ((u64)blockIdx.x * fr + k) * blockDim.x + threadIdx.x;
    if (i < npairs) {
      u64 o1 = pair_list[2 * i], o2 = pair_list[2 * i + 1]; Core c1, c2; load_core(U, o1, c1); load_core(U, o2, c2);
      st.paired++; process_pair(T, U, o1, c1, o2, c2, sink, st);
    }
  }
  dt.flush(T);
  warp_add_stat(&T.stats[ST_PAIRED], st.paired); warp_add_stat(&T.stats[ST_SHORT], st.shortp); warp_add_stat(&T.stats[ST_INTERSECT], st.inter);
  warp_add_stat(&T.stats[ST_LONG], st.longp); warp_add_stat(&T.stats[ST_NUCLEOTIDES], sink.nuc);
}
__global__ void __launch_bounds__(128) process_singles_kernel(DevTables T, const uint8_t* __restrict__ U, const u64* __restrict__ single_list, u64 nsingles, const u64* __restrict__ nsingles_dev, int fr) {
  extern __shared__ __align__(16) int32_t sh_tile[]; __shared__ unsigned long long sh_base[2];
  if (nsingles_dev) nsingles = min(nsingles, *nsingles_dev);
  if ((u64)blockIdx.x * fr * blockDim.x >= nsingles) return;
  DepthTile dt; dt.begin(T, sh_tile, sh_base);
  FragSink sink(T); dt.attach(sink); u32 n = 0;
  for (int k = 0; k < fr; k++) {
    const u64 i = ((u64)blockIdx.x * fr + k) * blockDim.x + threadIdx.x;
    if (i < nsingles) { u64 e = single_list[i]; Core c; load_core(U, e, c); n++; process_single(T, U, e, c, sink); }
  }
  dt.flush(T);
  warp_add_stat(&T.stats[ST_SINGLE], n); warp_add_stat(&T.stats[ST_NUCLEOTIDES], sink.nuc);
}
void launch_pair_records(const DevTables& T, const uint8_t* U, const u64* rec_off, u64 nrec, u64* ptab, u32 pmask, u64* pair_list, u64* single_list, u64* list_counts, cudaStream_t s, u64 carry_base, int count_stats, int adjacent) {
  u64 grid = (nrec + 255) / 256;
  if (grid) { pair_records_kernel<<<(unsigned)grid, 256, 0, s>>>(T, U, rec_off, nrec, ptab, pmask, pair_list, single_list, list_counts, carry_base, count_stats, adjacent); IRF_LAUNCHED(); }
}
__global__ void shift_offsets_kernel(u64* off, u64 n, u64 delta) { u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; if (i < n) off[i] += delta; }
void launch_shift_offsets(u64* off, u64 n, u64 delta, cudaStream_t s) { u64 g = (n + 255) / 256; if (g) { shift_offsets_kernel<<<(unsigned)g, 256, 0, s>>>(off, n, delta); IRF_LAUNCHED(); } }
void launch_process_pairs(const DevTables& T, const uint8_t* U, const u64* pair_list, u64 npairs, cudaStream_t s, const u64* npairs_dev) {
  const int fr = T.depth_tile ? 4 : 1; u64 grid = (npairs + 128ull * fr - 1) / (128ull * fr);   // npairs_dev given: npairs is an upper bound and the kernel reads the count itself
  if (grid) { process_pairs_kernel<<<(unsigned)grid, 128, T.depth_tile ? 2 * DEPTH_TILE * sizeof(int32_t) : 0, s>>>(T, U, pair_list, npairs, npairs_dev, fr); IRF_LAUNCHED(); }
}
void launch_process_singles(const DevTables& T, const uint8_t* U, const u64* single_list, u64 nsingles, cudaStream_t s, const u64* nsingles_dev) {
  const int fr = T.depth_tile ? 4 : 1; u64 grid = (nsingles + 128ull * fr - 1) / (128ull * fr);
  if (grid) { process_singles_kernel<<<(unsigned)grid, 128, T.depth_tile ? 2 * DEPTH_TILE * sizeof(int32_t) : 0, s>>>(T, U, single_list, nsingles, nsingles_dev, fr); IRF_LAUNCHED(); }
}
void launch_walk_list(const FrameArgs& a, const u64* rec_base, u64* rec_off, cudaStream_t s) {
  int grid = (a.nblk + 127) / 128;
  if (grid) { walk_list_kernel<<<grid, 128, 0, s>>>(a.U, a.u_off, a.u_len, a.nblk, a.entry, rec_base, rec_off); IRF_LAUNCHED(); }
}
void launch_scan_counts(const u32* in, u64* out, int n, cudaStream_t s) { if (n) { scan_counts_kernel<<<1, 1024, 0, s>>>(in, out, n); IRF_LAUNCHED(); } }

// Records still waiting in the pair table never met a mate: error reads on one GPU, halo candidates when sharded.
__global__ void pair_leftover_kernel(const u64* __restrict__ ptab, u64 nslots, u64* __restrict__ n_left, u64* __restrict__ left_off, u64 left_cap) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nslots) return;
  u64 v = ptab[i];
  if (v != PAIR_EMPTY && v != PAIR_TOMB) { u64 at = atomicAdd(n_left, 1ull); if (left_off && at < left_cap) left_off[at] = (v & ((1ull << 40) - 1)) - 1; }
}
void launch_pair_leftover(const u64* ptab, u64 nslots, u64* n_left, u64* left_off, u64 left_cap, cudaStream_t s) {
  u64 grid = (nslots + 255) / 256; if (grid) { pair_leftover_kernel<<<(unsigned)grid, 256, 0, s>>>(ptab, nslots, n_left, left_off, left_cap); IRF_LAUNCHED(); }
}

// Copies each leftover record's fixed part + name + CIGAR into a dense halo buffer: [u32 bytes][record bytes padded to 4].
__global__ void halo_pack_kernel(const uint8_t* __restrict__ U, const u64* __restrict__ left_off, u64 n, uint8_t* __restrict__ out, u64* __restrict__ rec_off, u64* __restrict__ cursor, u64 cap) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
  u64 off = left_off[i]; Core c; load_core(U, off, c);
  u32 keep = 36 + c.l_read_name + 4 * c.n_cigar_op; u32 padded = (keep + 3) & ~3u;
  u64 at = atomicAdd(cursor, (u64)padded); rec_off[i] = at; if (at + padded > cap) return;
  for (u32 k = 0; k < keep; k++) out[at + k] = U[off + k];
  for (u32 k = keep; k < padded; k++) out[at + k] = 0;
  int bs = (int)padded - 4; out[at] = (uint8_t)bs; out[at + 1] = (uint8_t)(bs >> 8); out[at + 2] = (uint8_t)(bs >> 16); out[at + 3] = (uint8_t)(bs >> 24);  // shrink block_size: seq/qual/tags dropped
  // l_seq must stay consistent with the shortened record
  out[at + 20] = out[at + 21] = out[at + 22] = out[at + 23] = 0;
}
void launch_halo_pack(const uint8_t* U, const u64* left_off, u64 n, uint8_t* out, u64* rec_off, u64* cursor, u64 cap, cudaStream_t s) {
  u64 grid = (n + 127) / 128; if (grid) { halo_pack_kernel<<<(unsigned)grid, 128, 0, s>>>(U, left_off, n, out, rec_off, cursor, cap); IRF_LAUNCHED(); }
}

// Halo pass: the gathered leftovers of all shards form one small record stream; every rank sees the same stream and
// processes the pairs whose key hash it owns, so each cross-shard pair is counted exactly once.
__global__ void halo_process_kernel(DevTables T, const uint8_t* __restrict__ H, u64 Hend, const u64* __restrict__ rec_off, u64 nrec, u64* __restrict__ ptab, u32 pmask, u32 owner, u32 nowners) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; if (i >= nrec) return;
  u64 e = rec_off[i]; Core c; load_core(H, e, c);
  FragSink sink(T); LocalStats st = {0, 0, 0, 0, 0, 0, 0};
  u64 mo; Core m;
  if (pair_lookup_or_wait(T, ptab, pmask, H, e, c, mo, m)) {
    if ((u32)(pair_key(H, e, c) >> 8) % nowners == owner) {
      st.paired++;
      bool first_is_me = c.pos < m.pos || (c.pos == m.pos && e < mo);
      u64 o1 = first_is_me ? e : mo, o2 = first_is_me ? mo : e; Core c1 = first_is_me ? c : m, c2 = first_is_me ? m : c;
      process_pair(T, H, o1, c1, o2, c2, sink, st);
    }
  }
  if (st.paired) atomicAdd(&T.stats[ST_PAIRED], (u64)st.paired);
  if (st.shortp) atomicAdd(&T.stats[ST_SHORT], (u64)st.shortp);
  if (st.inter) atomicAdd(&T.stats[ST_INTERSECT], (u64)st.inter);
  if (st.longp) atomicAdd(&T.stats[ST_LONG], (u64)st.longp);
  if (sink.nuc) atomicAdd(&T.stats[ST_NUCLEOTIDES], sink.nuc);
}
void launch_halo_process(const DevTables& T, const uint8_t* H, u64 Hend, const u64* rec_off, u64 nrec, u64* ptab, u32 pmask, u32 owner, u32 nowners, cudaStream_t s) {
  u64 grid = (nrec + 127) / 128; if (grid) { halo_process_kernel<<<(unsigned)grid, 128, 0, s>>>(T, H, Hend, rec_off, nrec, ptab, pmask, owner, nowners); IRF_LAUNCHED(); }
}

// ------------------------------------------------------------------------------------------------
// JunctionCount table maintenance
__global__ void ht_insert_keys_kernel(u64* keys, u32 mask, const u64* __restrict__ in, u64 n, int* err) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; if (i < n) ht_upsert(keys, mask, in[i], err);
}
__global__ void ht_add_kernel(u64* keys, u32* cnt, u32 mask, const u64* __restrict__ in_keys, const u32* __restrict__ in_cnt, u64 n, int* err) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
  u32 s = ht_upsert(keys, mask, in_keys[i], err); if (s == 0xffffffffu) return;
  if (in_cnt[2 * i]) atomicAdd(&cnt[2ull * s], in_cnt[2 * i]); if (in_cnt[2 * i + 1]) atomicAdd(&cnt[2ull * s + 1], in_cnt[2 * i + 1]);
}
__global__ void ht_compact_kernel(const u64* __restrict__ keys, const u32* __restrict__ cnt, u64 nslots, u64* out_keys, u32* out_cnt, u64* cursor, u64 cap, int only_counted) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; if (i >= nslots) return;
  u64 k = keys[i]; if (k == JKEY_EMPTY) return;
  u32 a = cnt[2 * i], b = cnt[2 * i + 1]; if (only_counted && !a && !b) return;
  u64 at = atomicAdd(cursor, 1ull); if (at < cap) { out_keys[at] = k; out_cnt[2 * at] = a; out_cnt[2 * at + 1] = b; }
}
// Sorts the exported (key, counts) list by key so the host can emit JuncCount rows without sorting: keys order like (refID, start, end).
// CUB's radix sort is used for this small auxiliary step (a few hundred thousand entries once per run).
size_t junction_sort_temp_bytes(u64 n) { size_t b = 0; cub::DeviceRadixSort::SortPairs(nullptr, b, (const u64*)nullptr, (u64*)nullptr, (const u64*)nullptr, (u64*)nullptr, (int)n); return b; }
void launch_junction_sort(void* temp, size_t temp_bytes, const u64* keys_in, u64* keys_out, const u64* cnt_in, u64* cnt_out, u64 n, cudaStream_t s) {
  if (n) { cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, cnt_in, cnt_out, (int)n, 0, 64, s); IRF_LAUNCHED(); }
}
// Sorted (key, counts) pairs -> the rows the host works with {refID, start, end, count[0], count[1]} (the JuncEntry / IRF_T_JUNCTIONS layout).
struct JuncRow { int refid; u32 start, end, c0, c1; };
__global__ void junction_rows_kernel(const u64* __restrict__ keys, const u64* __restrict__ cnt, u64 n, const u64* __restrict__ chr_goff, int n_ref, JuncRow* __restrict__ out) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
  const u64 k = keys[i], gl = k >> JKEY_GAP_BITS, gap = k & ((1ull << JKEY_GAP_BITS) - 1), c = cnt[i];
  int lo = 0, hi = n_ref;                                   // last reference whose offset is <= gl
  while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (chr_goff[mid] <= gl) lo = mid; else hi = mid; }
  out[i] = JuncRow{lo, (u32)(gl - chr_goff[lo]), (u32)(gl - chr_goff[lo] + gap), (u32)c, (u32)(c >> 32)};
}
void launch_junction_rows(const u64* keys, const u64* cnt, u64 n, const u64* chr_goff, int n_ref, void* out, cudaStream_t s) {
  u64 g = (n + 255) / 256; if (g) { junction_rows_kernel<<<(unsigned)g, 256, 0, s>>>(keys, cnt, n, chr_goff, n_ref, (JuncRow*)out); IRF_LAUNCHED(); }
}
void launch_ht_insert_keys(u64* keys, u32 mask, const u64* in, u64 n, int* err, cudaStream_t s) { u64 g = (n + 255) / 256; if (g) { ht_insert_keys_kernel<<<(unsigned)g, 256, 0, s>>>(keys, mask, in, n, err); IRF_LAUNCHED(); } }
void launch_ht_add(u64* keys, u32* cnt, u32 mask, const u64* in_keys, const u32* in_cnt, u64 n, int* err, cudaStream_t s) { u64 g = (n + 255) / 256; if (g) { ht_add_kernel<<<(unsigned)g, 256, 0, s>>>(keys, cnt, mask, in_keys, in_cnt, n, err); IRF_LAUNCHED(); } }
void launch_ht_compact(const u64* keys, const u32* cnt, u64 nslots, u64* out_keys, u32* out_cnt, u64* cursor, u64 cap, int only_counted, cudaStream_t s) {
  u64 g = (nslots + 255) / 256; if (g) { ht_compact_kernel<<<(unsigned)g, 256, 0, s>>>(keys, cnt, nslots, out_keys, out_cnt, cursor, cap, only_counted); IRF_LAUNCHED(); }
}

// ------------------------------------------------------------------------------------------------
// S4a: difference counters -> per-base depth.  In-place inclusive scan over both strands at once (every
// segment's counters sum to zero at its sentinel slot, so no segment boundaries need handling).
constexpr int SCAN_THREADS = 256, SCAN_ITEMS = 16, SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;
__global__ void __launch_bounds__(SCAN_THREADS) scan_tile_sums_kernel(const int32_t* __restrict__ d, u64 n, int32_t* __restrict__ sums) {
  u64 base = (u64)blockIdx.x * SCAN_TILE; int acc = 0;
  for (int k = 0; k < SCAN_ITEMS; k++) { u64 i = base + (u64)k * SCAN_THREADS + threadIdx.x; if (i < n) acc += d[i]; }
  __shared__ int sh[SCAN_THREADS / 32];
  for (int o = 16; o; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) { int t = 0; for (int w = 0; w < SCAN_THREADS / 32; w++) t += sh[w]; sums[blockIdx.x] = t; }
}
__global__ void __launch_bounds__(1024) scan_sums_kernel(int32_t* sums, u64 nt) {  // exclusive scan, single CTA
  __shared__ int sh[1024]; __shared__ int carry; if (threadIdx.x == 0) carry = 0; __syncthreads();
  for (u64 base = 0; base < nt; base += 1024) {
    u64 i = base + threadIdx.x; int v = i < nt ? sums[i] : 0; sh[threadIdx.x] = v; __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) { int t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0; __syncthreads(); sh[threadIdx.x] += t; __syncthreads(); }
    int incl = sh[threadIdx.x], c = carry; if (i < nt) sums[i] = c + incl - v; __syncthreads();
    if (threadIdx.x == 1023) carry = c + incl; __syncthreads();
  }
}
// P / dmax: when given, the largest depth of either strand (elements below / from P) is folded into this pass (dmax[0], dmax[1]); their sum
// bounds the largest combined depth, which only sizes the per-intron histograms (a separate 1.1 ms pass over the table in round 1).
__global__ void __launch_bounds__(SCAN_THREADS) scan_apply_kernel(int32_t* __restrict__ d, u64 n, const int32_t* __restrict__ sums, u64 P, u32* __restrict__ dmax) {
  // each thread owns SCAN_ITEMS consecutive elements (vector loads), warp + CTA prefix over thread totals
  u64 base = (u64)blockIdx.x * SCAN_TILE + (u64)threadIdx.x * SCAN_ITEMS; int v[SCAN_ITEMS]; int tot = 0;
  if (base + SCAN_ITEMS <= n) { const int4* p = (const int4*)(d + base);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS / 4; k++) { int4 q = p[k]; v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w; } }
  else { for (int k = 0; k < SCAN_ITEMS; k++) v[k] = base + k < n ? d[base + k] : 0; }
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { tot += v[k]; v[k] = tot; }
  int incl = tot; const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
  __shared__ int wsum[SCAN_THREADS / 32];
  if (lane == 31) wsum[warp] = incl; __syncthreads();
  int off = sums[blockIdx.x] + incl - tot; for (int w = 0; w < warp; w++) off += wsum[w];
  if (base + SCAN_ITEMS <= n) { int4* p = (int4*)(d + base);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS / 4; k++) p[k] = make_int4(v[4 * k] + off, v[4 * k + 1] + off, v[4 * k + 2] + off, v[4 * k + 3] + off); }
  else { for (int k = 0; k < SCAN_ITEMS; k++) if (base + k < n) d[base + k] = v[k] + off; }
  if (dmax) {
    int m0 = 0, m1 = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) if (base + k < n) { const int x = v[k] + off; if (base + k < P) m0 = max(m0, x); else m1 = max(m1, x); }
    for (int o = 16; o; o >>= 1) { m0 = max(m0, __shfl_down_sync(0xffffffffu, m0, o)); m1 = max(m1, __shfl_down_sync(0xffffffffu, m1, o)); }
    if (lane == 0) { if (m0 > 0) atomicMax(&dmax[0], (u32)m0); if (m1 > 0) atomicMax(&dmax[1], (u32)m1); }
  }
}
void launch_depth_scan(int32_t* d, u64 n, int32_t* tile_sums, cudaStream_t s, u64 P, u32* dmax) {
  if (!n) return; u64 nt = (n + SCAN_TILE - 1) / SCAN_TILE;
  { scan_tile_sums_kernel<<<(unsigned)nt, SCAN_THREADS, 0, s>>>(d, n, tile_sums); IRF_LAUNCHED(); }
  { scan_sums_kernel<<<1, 1024, 0, s>>>(tile_sums, nt); IRF_LAUNCHED(); }
  { scan_apply_kernel<<<(unsigned)nt, SCAN_THREADS, 0, s>>>(d, n, tile_sums, P, dmax); IRF_LAUNCHED(); }
}
u64 depth_scan_tiles(u64 n) { return (n + SCAN_TILE - 1) / SCAN_TILE; }

__global__ void max_depth_kernel(const int32_t* __restrict__ depth, u64 P, u32* __restrict__ out) {
  u32 m = 0;
  for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < P; i += (u64)gridDim.x * blockDim.x) m = max(m, (u32)(depth[i] + depth[P + i]));
  for (int o = 16; o; o >>= 1) m = max(m, __shfl_down_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}
void launch_max_depth(const int32_t* depth, u64 P, u32* out, cudaStream_t s) { if (P) { max_depth_kernel<<<148 * 8, 256, 0, s>>>(depth, P, out); IRF_LAUNCHED(); } }

// ------------------------------------------------------------------------------------------------
// S4b: CoverageBlocks::fillHist + *FromHist + CoverageBlocksIRFinder::WriteOutput's numeric part (rows a10-a12).
// One CTA per ref-cover.bed record.  The depth histogram lives in shared memory when the intron's maximum depth
// fits, otherwise in the CTA's slice of a global scratch histogram.
constexpr int IS_THREADS = 256, IS_SMEM_BINS = 8192;
__device__ __forceinline__ u32 depth_at(const int32_t* depth, u64 P, u64 c, int sel) { return sel < 0 ? (u32)(depth[c] + depth[P + c]) : (u32)depth[(u64)sel * P + c]; }

// coverageFromHist / percentileFromHist / trimmedMeanFromHist in one ascending walk ([C-2][C-12]), the edge-window means, SpansPoint::lookup and
// JunctionCount::lookup{Left,Right,}, IRratio and the warning: everything of one IRFinder-IR row that follows the depth histogram.  One thread.
__device__ void intron_row_from_hist(const DevTables& T, const IntronDesc& in, const u32* hist, u32 mx, u64 npos, const u64* sh_edge, int sel, u32 trim_center, IntronRow* out) {
      IntronRow r; r.positions = npos; r.excluded = (long long)(in.end - in.start) - (long long)npos;
      // coverageFromHist / percentileFromHist / trimmedMeanFromHist, one ascending walk ([C-2][C-12])
      u64 n = npos, lo = n * (100 - trim_center) / 200, hi = n - lo, cum = 0, total = 0; double pct[3] = {0, 0, 0}; int pstate[3] = {0, 0, 0}; u32 pdepth[3] = {0, 0, 0};
      const u32 pp[3] = {25, 50, 75};
      for (u32 d = 0; d <= mx && n; d++) {
        u32 h = hist[d]; if (!h) continue;
        for (int q = 0; q < 3; q++) if (pstate[q] == 1) { pct[q] = ((double)pdepth[q] + (double)d) / 2.0; pstate[q] = 2; }
        u64 a = max(cum, lo), b = min(cum + h, hi); if (b > a) total += (b - a) * d;
        cum += h;
        for (int q = 0; q < 3; q++) if (pstate[q] == 0) { if (cum * 100 > n * pp[q]) { pct[q] = d; pstate[q] = 2; } else if (cum * 100 == n * pp[q]) { pdepth[q] = d; pct[q] = d; pstate[q] = T.rules.pct_tie_midpoint ? 1 : 2; } }
      }
      r.coverage = n ? (double)(n - hist[0]) / (double)n : 0.0;
      r.depth = (n && hi > lo) ? (double)total / (double)(hi - lo) : 0.0;
      r.p25 = pct[0]; r.p50 = pct[1]; r.p75 = pct[2];
      r.first50 = sh_edge[1] ? (double)sh_edge[0] / (double)sh_edge[1] : 0.0;
      r.last50 = sh_edge[3] ? (double)sh_edge[2] / (double)sh_edge[3] : 0.0;
      // SpansPoint::lookup and JunctionCount::lookup{Left,Right,}
      auto pick = [&](const u32* cnt, u64 stride2, u64 i) -> u32 { return sel < 0 ? cnt[i] + cnt[i + stride2] : cnt[i + (sel ? stride2 : 0)]; };
      r.sp_left = in.sp_left >= 0 ? pick(T.sp_count, T.n_sp, (u64)in.sp_left) : 0; r.sp_right = in.sp_right >= 0 ? pick(T.sp_count, T.n_sp, (u64)in.sp_right) : 0;
      r.jc_left = r.jc_right = r.jc_exact = 0;
      if (in.refid >= 0) {
        u64 base = T.chr_goff[in.refid];
        auto jl = [&](const u64* keys, const u32* cnt, u32 mask, u64 key) -> u32 { u32 s = ht_find(keys, mask, key); if (s == 0xffffffffu) return 0; return sel < 0 ? cnt[2ull * s] + cnt[2ull * s + 1] : cnt[2ull * s + sel]; };
        r.jc_left = jl(T.jl_keys, T.jl_cnt, T.jl_mask, base + in.start); r.jc_right = jl(T.jr_keys, T.jr_cnt, T.jr_mask, base + in.end);
        if (in.end >= in.start && (u64)(in.end - in.start) < (1ull << JKEY_GAP_BITS)) r.jc_exact = jl(T.jc_keys, T.jc_cnt, T.jc_mask, jkey_make(base + in.start, in.end - in.start));
      }
      u32 jmax = max(r.jc_left, r.jc_right);
      r.irratio = (r.depth == 0 && jmax == 0) ? 0.0 : r.depth / (r.depth + (double)jmax);
      // warnings [C-14]
      double spmax = max(r.sp_left, r.sp_right), spmin = min(r.sp_left, r.sp_right);
      const Rules& R = T.rules;
      if ((double)r.jc_exact + r.depth < R.warn_lowcover) r.warning = W_LOWCOVER;
      else if ((double)r.jc_exact < R.warn_lowsplice) r.warning = W_LOWSPLICING;
      else if ((double)r.jc_exact < (double)jmax * R.warn_minor_ratio) r.warning = W_MINORISOFORM;
      else if ((spmax > r.depth + R.warn_nonuniform_abs && spmax > r.depth * R.warn_nonuniform_ratio) || (spmin + R.warn_nonuniform_abs < r.depth && spmin * R.warn_nonuniform_ratio < r.depth)) r.warning = W_NONUNIFORM;
      else r.warning = W_NONE;
  *out = r;
}


__global__ void __launch_bounds__(IS_THREADS) intron_stats_kernel(DevTables T, const int32_t* __restrict__ depth, const IntronDesc* __restrict__ introns, const IntronBlock* __restrict__ blocks,
                                                                   u32 n_introns, int directionality, u32 trim_center, u32* __restrict__ scratch, u32 scratch_bins, IntronRow* __restrict__ rows,
                                                                   const u32* __restrict__ list, const u32* __restrict__ n_dev) {   // list: optional indices of the introns to compute (owned introns, or the fast path's overflow list); n_dev: their number when only the device knows it
  __shared__ u32 sh_hist[IS_SMEM_BINS]; __shared__ u32 sh_red[IS_THREADS / 32]; __shared__ u32 sh_max; __shared__ u64 sh_edge[4];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  u32* ghist = scratch + (u64)blockIdx.x * scratch_bins;
  if (n_dev) n_introns = min(n_introns, *n_dev);
  for (u32 li = blockIdx.x; li < n_introns; li += gridDim.x) {
    const u32 it = list ? list[li] : li;
    const IntronDesc in = introns[it];
    int sel = -1;
    if (directionality != 0) { bool measureDir = in.plus != 0; if (directionality < 0) measureDir = !measureDir; sel = measureDir ? 1 : 0; }
    // pass 1: maximum depth and number of positions
    u32 mx = 0; u64 npos = 0;
    for (u32 b = in.blk_begin; b < in.blk_main_end; b++) { IntronBlock bl = blocks[b]; npos += bl.len; for (u32 p = tid; p < bl.len; p += IS_THREADS) mx = max(mx, depth_at(depth, T.P, bl.c0 + p, sel)); }
    for (int o = 16; o; o >>= 1) mx = max(mx, __shfl_down_sync(0xffffffffu, mx, o));
    if (lane == 0) sh_red[warp] = mx; __syncthreads();
    if (tid == 0) { u32 m = 0; for (int w = 0; w < IS_THREADS / 32; w++) m = max(m, sh_red[w]); sh_max = m; } __syncthreads();
    mx = sh_max; const bool in_smem = mx < IS_SMEM_BINS; u32* hist = in_smem ? sh_hist : ghist;
    if (!in_smem && mx >= scratch_bins) { if (tid == 0) atomicOr(T.err, DERR_BAD_RECORD); __syncthreads(); continue; }
    for (u32 i = tid; i <= mx; i += IS_THREADS) hist[i] = 0;
    __syncthreads();
    // pass 2: histogram
    for (u32 b = in.blk_begin; b < in.blk_main_end; b++) { IntronBlock bl = blocks[b]; for (u32 p = tid; p < bl.len; p += IS_THREADS) atomicAdd(&hist[depth_at(depth, T.P, bl.c0 + p, sel)], 1u); }
    // edge windows (mean depth over measured positions): warps 1 and 2
    if (warp == 1 || warp == 2) {
      u32 b0 = warp == 1 ? in.blk_main_end : in.blk_first_end, b1 = warp == 1 ? in.blk_first_end : in.blk_last_end; u64 tot = 0, cnt = 0;
      for (u32 b = b0; b < b1; b++) { IntronBlock bl = blocks[b]; cnt += bl.len; for (u32 p = lane; p < bl.len; p += 32) tot += depth_at(depth, T.P, bl.c0 + p, sel); }
      for (int o = 16; o; o >>= 1) tot += __shfl_down_sync(0xffffffffu, tot, o);
      if (lane == 0) { sh_edge[(warp - 1) * 2] = tot; sh_edge[(warp - 1) * 2 + 1] = cnt; }
    }
    __syncthreads();
    if (tid == 0) intron_row_from_hist(T, in, hist, mx, npos, sh_edge, sel, trim_center, &rows[it]);
    __syncthreads();
  }
}
// Fast path of the same: ONE pass over the intron's depth values (round 1 read them twice: maximum first, then histogram) into a
// 2048-bin shared-memory histogram; an intron with a deeper position is appended to `overflow` and done by intron_stats_kernel.
constexpr int ISF_BINS = 2048;
__global__ void __launch_bounds__(IS_THREADS) intron_stats_fast_kernel(DevTables T, const int32_t* __restrict__ depth, const IntronDesc* __restrict__ introns, const IntronBlock* __restrict__ blocks,
                                                                        u32 n_introns, int directionality, u32 trim_center, IntronRow* __restrict__ rows, const u32* __restrict__ list,
                                                                        u32* __restrict__ overflow, u32* __restrict__ n_overflow) {
  __shared__ u32 sh_hist[ISF_BINS]; __shared__ u32 sh_max, sh_over; __shared__ u64 sh_edge[4];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const u32 li = blockIdx.x; if (li >= n_introns) return;
  const u32 it = list ? list[li] : li;
  const IntronDesc in = introns[it];
  int sel = -1;
  if (directionality != 0) { bool measureDir = in.plus != 0; if (directionality < 0) measureDir = !measureDir; sel = measureDir ? 1 : 0; }
  for (int i = tid; i < ISF_BINS; i += IS_THREADS) sh_hist[i] = 0;
  if (tid == 0) { sh_max = 0; sh_over = 0; }
  __syncthreads();
  u32 mx = 0; u64 npos = 0; bool over = false;
  for (u32 b = in.blk_begin; b < in.blk_main_end; b++) { const IntronBlock bl = blocks[b]; npos += bl.len;
    for (u32 p = tid; p < bl.len; p += IS_THREADS) { const u32 d = depth_at(depth, T.P, bl.c0 + p, sel); if (d < (u32)ISF_BINS) { atomicAdd(&sh_hist[d], 1u); mx = max(mx, d); } else over = true; } }
  for (int o = 16; o; o >>= 1) mx = max(mx, __shfl_down_sync(0xffffffffu, mx, o));
  if (lane == 0 && mx) atomicMax(&sh_max, mx);
  if (over) sh_over = 1;
  if (warp == 1 || warp == 2) {   // edge windows (mean depth over measured positions)
    u32 b0 = warp == 1 ? in.blk_main_end : in.blk_first_end, b1 = warp == 1 ? in.blk_first_end : in.blk_last_end; u64 tot = 0, cnt = 0;
    for (u32 b = b0; b < b1; b++) { const IntronBlock bl = blocks[b]; cnt += bl.len; for (u32 p = lane; p < bl.len; p += 32) tot += depth_at(depth, T.P, bl.c0 + p, sel); }
    for (int o = 16; o; o >>= 1) tot += __shfl_down_sync(0xffffffffu, tot, o);
    if (lane == 0) { sh_edge[(warp - 1) * 2] = tot; sh_edge[(warp - 1) * 2 + 1] = cnt; }
  }
  __syncthreads();
  if (tid == 0) {
    if (sh_over) overflow[atomicAdd(n_overflow, 1u)] = it;
    else intron_row_from_hist(T, in, sh_hist, sh_max, npos, sh_edge, sel, trim_center, &rows[it]);
  }
}
void launch_intron_stats(const DevTables& T, const int32_t* depth, const IntronDesc* introns, const IntronBlock* blocks, u32 n_introns, int directionality, u32 trim_center,
                         u32* scratch, u32 scratch_bins, int grid, IntronRow* rows, cudaStream_t s, const u32* list, const u32* n_dev) {
  if (n_introns) { intron_stats_kernel<<<grid, IS_THREADS, 0, s>>>(T, depth, introns, blocks, n_introns, directionality, trim_center, scratch, scratch_bins, rows, list, n_dev); IRF_LAUNCHED(); }
}
void launch_intron_stats_fast(const DevTables& T, const int32_t* depth, const IntronDesc* introns, const IntronBlock* blocks, u32 n_introns, int directionality, u32 trim_center,
                              IntronRow* rows, cudaStream_t s, const u32* list, u32* overflow, u32* n_overflow) {
  if (n_introns) { intron_stats_fast_kernel<<<n_introns, IS_THREADS, 0, s>>>(T, depth, introns, blocks, n_introns, directionality, trim_center, rows, list, overflow, n_overflow); IRF_LAUNCHED(); }
}

// ------------------------------------------------------------------------------------------------
// Region-owned depth (multi-GPU, SURVEY.md §8e "Depth needs no reduction (region-owned)").  Every measured segment has one
// owning rank; a rank's difference counters outside its own segments are almost all zero (a coordinate-sorted shard covers
// its own region plus the reads that reach over the cut), so instead of all-reducing the whole table each rank lists its
// non-zero foreign counters as (index, value) and the owners add them.
// tile_state: one byte per 1024 consecutive slots of the compact space: 1 = every slot owned by this rank (skipped), 2 = none owned
// (every non-zero counter is listed), 0 = mixed (the owner of a non-zero slot is looked up).  The table is streamed once at HBM speed;
// the first version walked segment by segment (one warp per segment, ~60 dependent loads each: 4 ms per step at 8 GPUs).
constexpr int FOREIGN_TILE = 1024;
__global__ void __launch_bounds__(256) foreign_depth_kernel(const int32_t* __restrict__ diff, u64 P, const u64* __restrict__ seg_coff, const uint8_t* __restrict__ seg_owner, u32 nseg, u32 me,
                                                             const uint8_t* __restrict__ tile_state, u64 ntiles, u64* __restrict__ out_idx, int32_t* __restrict__ out_val, u64* __restrict__ cursor, u64 cap) {
  const u32 lane = threadIdx.x & 31;
  for (u64 t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const uint8_t st = tile_state[t]; if (st == 1) continue;
    for (int dir = 0; dir < 2; dir++) {
#pragma unroll
      for (int k = 0; k < FOREIGN_TILE / 256; k++) {
        const u64 c = t * FOREIGN_TILE + (u64)k * 256 + threadIdx.x; const int32_t v = c < P ? diff[(u64)dir * P + c] : 0;
        bool emit = v != 0;
        if (emit && st == 0) { u32 lo = 0, hi = nseg; while (hi - lo > 1) { u32 mid = (lo + hi) >> 1; if (seg_coff[mid] <= c) lo = mid; else hi = mid; } emit = seg_owner[lo] != me; }
        const u32 m = __ballot_sync(0xffffffffu, emit); if (!m) continue;
        u64 at = 0; if (lane == (u32)(__ffs(m) - 1)) at = atomicAdd(cursor, (u64)__popc(m)); at = __shfl_sync(0xffffffffu, at, __ffs(m) - 1) + __popc(m & ((1u << lane) - 1u));
        if (emit && at < cap) { out_idx[at] = (u64)dir * P + c; out_val[at] = v; }
      }
    }
  }
}
__global__ void import_depth_kernel(int32_t* __restrict__ diff, u64 P, const u64* __restrict__ seg_coff, const uint8_t* __restrict__ seg_owner, u32 nseg, u32 me,
                                    const u64* __restrict__ idx, const int32_t* __restrict__ val, u64 n) {
  u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; if (i >= n) return;
  const u64 g = idx[i], c = g >= P ? g - P : g;
  u32 lo = 0, hi = nseg;                                  // last segment whose first slot is <= c
  while (hi - lo > 1) { u32 mid = (lo + hi) >> 1; if (seg_coff[mid] <= c) lo = mid; else hi = mid; }
  if (seg_owner[lo] == me) atomicAdd(&diff[g], val[i]);
}
void launch_foreign_depth(const DevTables& T, u32 nseg, const uint8_t* seg_owner, u32 me, const uint8_t* tile_state, u64* out_idx, int32_t* out_val, u64* cursor, u64 cap, cudaStream_t s) {
  const u64 ntiles = (T.P + FOREIGN_TILE - 1) / FOREIGN_TILE;
  if (nseg && ntiles) { foreign_depth_kernel<<<(unsigned)std::min<u64>(ntiles, 148 * 16), 256, 0, s>>>(T.diff, T.P, T.seg_coff, seg_owner, nseg, me, tile_state, ntiles, out_idx, out_val, cursor, cap); IRF_LAUNCHED(); }
}
void launch_import_depth(const DevTables& T, u32 nseg, const uint8_t* seg_owner, u32 me, const u64* idx, const int32_t* val, u64 n, cudaStream_t s) {
  u64 g = (n + 255) / 256; if (g && nseg) { import_depth_kernel<<<(unsigned)g, 256, 0, s>>>(T.diff, T.P, T.seg_coff, seg_owner, nseg, me, idx, val, n); IRF_LAUNCHED(); }
}

}  // namespace irf
