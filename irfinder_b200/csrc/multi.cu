// In-process multi-device run (SURVEY.md §8b "one ctx drives all its GPUs internally", §8e region sharding).
//
// The reference is one process with one set of processors; `irf_create` with n_devices > 1 keeps that shape: one irf_ctx per
// device, one host thread per context while a phase runs, and the merge protocol of §8e between the phases.  The protocol is the
// one irfinder_b200/distributed.py speaks over NCCL between processes; here every context lives in this process, so the
// exchanges are plain copies: the small ones (halo records, junction lists, span / ROI / chromosome / read counters, foreign
// depth entries: KBs to MBs) go through host memory, the row tables go device to device (unified addressing: the driver picks
// the path between two GPUs) into rank 0, which alone writes files.
// This file is a client of the C ABI (include/irfinder_b200.h) plus the three context hooks in multi.h.
#include <cuda_runtime.h>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include "../../include/irfinder_b200.h"
#include "common.cuh"
#include "multi.h"

namespace irf {
namespace {

using std::string; using std::vector;
constexpr size_t LOOKAHEAD = 1u << 18;            // bytes of the right neighbour a shard may read into (a record straddling the cut)
constexpr u64 FOREIGN_DENSE_LIMIT = 1ull << 27;   // foreign depth entries (all ranks) above which the depth tables are summed whole

__global__ void add_words_kernel(long long* __restrict__ dst, const long long* __restrict__ src, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] += src[i];
}
__global__ void add_i32_kernel(int* __restrict__ dst, const int* __restrict__ src, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] += src[i];
}

// runs f(r) for every rank on its own host thread with the rank's device current; the first failure is rethrown
template <class F> void each_rank(const vector<irf_ctx*>& R, F f) {
  const int n = (int)R.size(); vector<string> errs(n); vector<std::thread> th;
  auto body = [&](int r) { try { IRF_CUDA(cudaSetDevice(ctx_device(R[r]))); f(r); } catch (const string& e) { errs[r] = e.empty() ? string("unknown failure") : e; } catch (const std::exception& e) { errs[r] = e.what(); } };
  for (int r = 1; r < n; r++) th.emplace_back(body, r);
  body(0); for (auto& t : th) t.join();
  for (int r = 0; r < n; r++) if (!errs[r].empty()) throw n > 1 ? errs[r] + " [device " + std::to_string(ctx_device(R[r])) + ", shard " + std::to_string(r) + " of " + std::to_string(n) + "]" : errs[r];
}
inline void ck(irf_ctx* c, int rc) { if (rc != IRF_OK) { const char* m = irf_last_error(c); throw string(m && *m ? m : "call failed"); } }

// device scratch of one rank for imported lists
struct Scratch { void* p = nullptr; size_t cap = 0; ~Scratch() { if (p) cudaFree(p); } void* get(size_t n) { if (p) { cudaFree(p); p = nullptr; } cap = std::max<size_t>(n, 256); IRF_CUDA(cudaMalloc(&p, cap)); return p; }
  void* grow(size_t n) { return n <= cap ? p : get(n + n / 4); } };

bool block_at(const uint8_t* p, size_t n, size_t off, size_t& bsize) {   // BGZF block header with a BC subfield at p[off] (SAM v1 §4.1)
  if (off + 18 > n || p[off] != 31 || p[off + 1] != 139 || p[off + 2] != 8 || p[off + 3] != 4) return false;
  const u32 xlen = p[off + 10] | (u32)p[off + 11] << 8;
  for (u32 x = 0; x + 4 <= xlen && off + 12 + x + 6 <= n;) { const uint8_t* sf = p + off + 12 + x; const u32 slen = sf[2] | (u32)sf[3] << 8;
    if (sf[0] == 'B' && sf[1] == 'C' && slen == 2) { bsize = (size_t)(sf[4] | (u32)sf[5] << 8) + 1; return true; } x += 4 + slen; }
  return false;
}
// first block start at or after `approx`, confirmed by following three BSIZE links
size_t find_block_start(const uint8_t* p, size_t n, size_t approx) {
  for (size_t off = approx; off + 18 <= n; off++) {
    const void* q = memchr(p + off, 31, n - off); if (!q) return n; off = (size_t)((const uint8_t*)q - p);
    size_t cur = off, bs = 0; bool ok = true;
    for (int h = 0; h < 3 && cur < n; h++) { if (!block_at(p, n, cur, bs)) { ok = false; break; } cur += bs; }
    if (ok) return off;
  }
  return n;
}

// IRF_DEBUG_TIMING=1: wall-clock of every phase on stderr
struct PhaseClock { bool on = getenv("IRF_DEBUG_TIMING") != nullptr; std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now(); string line;
  void mark(const char* name) { if (!on) return; auto now = std::chrono::steady_clock::now(); char b[64]; snprintf(b, sizeof b, " %s %.1f", name, std::chrono::duration<double, std::milli>(now - t).count()); line += b; t = now; }
  ~PhaseClock() { if (on) fprintf(stderr, "[multi] ms:%s\n", line.c_str()); } };

template <class T> void d2h(vector<T>& dst, const void* dev, size_t n) { dst.resize(n); if (n) IRF_CUDA(cudaMemcpy(dst.data(), dev, n * sizeof(T), cudaMemcpyDeviceToHost)); }

}  // namespace

void multi_run(const vector<irf_ctx*>& all_ranks, const uint8_t* base, size_t n, int fd) {
  if (all_ranks.empty()) throw string("no device");
  if (ctx_require_eof(all_ranks[0])) {   // the same refusal as a single-device run: the 28-byte empty block that ends every BGZF file
    size_t bs = 0; if (n < 28 || !block_at(base, n, n - 28, bs) || bs != 28 || base[n - 4] || base[n - 3] || base[n - 2] || base[n - 1]) throw string("BGZF EOF marker missing (truncated file?)");
  }
  PhaseClock clk;
  vector<Scratch> S0(all_ranks.size()), S1(all_ranks.size());   // per-rank device scratch of the exchanges, grown as needed, freed when the run ends
  // the header first, on rank 0: no cut may fall inside it (it can span many blocks)
  auto read_header = [&](irf_ctx* c) { size_t probe = std::min<size_t>(n, 1u << 20);
    for (;;) { int rc = irf_read_header(c, base, probe); if (rc == IRF_OK) break; if (probe >= n) ck(c, rc); probe = std::min(n, probe * 4); } };
  IRF_CUDA(cudaSetDevice(ctx_device(all_ranks[0]))); read_header(all_ranks[0]);
  const size_t first_cut = (size_t)ctx_records_offset(all_ranks[0]);
  // cuts: BGZF block starts that balance the compressed bytes; the look-ahead of a shard ends on a block start as well.  A file with
  // fewer blocks than devices uses fewer devices: every shard holds at least one block besides the EOF marker.
  vector<size_t> cut{0};
  for (int r = 1, w = (int)all_ranks.size(); r < w; r++) { const size_t c = find_block_start(base, n, std::max(n / w * r, first_cut)); if (c > cut.back() && c + 28 < n) cut.push_back(c); }
  cut.push_back(n);
  const int W = (int)cut.size() - 1; const vector<irf_ctx*> R(all_ranks.begin(), all_ranks.begin() + W);
  clk.mark("header+cuts");
  // ---- count: every rank streams and counts its own byte range
  each_rank(R, [&](int r) {
    if (r) read_header(R[r]);
    const size_t lo = cut[r], hi = cut[r + 1], look = r == W - 1 ? hi : std::min(n, find_block_start(base, n, std::min(n, hi + LOOKAHEAD)));
    ck(R[r], ctx_count_shard(R[r], base + lo, hi - lo, look - lo, r == 0, fd, (uint64_t)lo));
  });
  clk.mark("count");
  if (W > 1) {
    // ---- frame hand-shake: a shard's first record starts where its left neighbour's last record ends; a recount moves the exit a
    // rank reports, so the chain is checked again until nobody has to move
    bool agreed = false;
    for (int round = 0; round <= W && !agreed; round++) {
      vector<uint64_t> e0(W), ex(W); for (int r = 0; r < W; r++) ck(R[r], irf_frame_info(R[r], &e0[r], &ex[r]));
      vector<char> bad(W, 0); agreed = true; for (int r = 1; r < W; r++) if (e0[r] != ex[r - 1]) { bad[r] = 1; agreed = false; }
      if (!agreed) each_rank(R, [&](int r) { if (!bad[r]) return; ck(R[r], irf_set_option(R[r], "entry0", std::to_string(ex[r - 1]).c_str())); ck(R[r], irf_reset(R[r])); ck(R[r], irf_count(R[r])); });
    }
    if (!agreed) throw string("record framing of neighbouring shards does not agree");
    clk.mark("handshake");
    // ---- 1. halo: records whose mate lies in another shard; every rank sees all of them and counts the pairs it owns by hash
    { vector<vector<uint8_t>> hb(W); vector<vector<u64>> ho(W);
      each_rank(R, [&](int r) { void *b, *o; size_t nb, nr; ck(R[r], irf_export_halo(R[r], &b, &nb, &o, &nr)); d2h(hb[r], b, nb); d2h(ho[r], o, nr); });
      vector<uint8_t> all; vector<u64> offs; for (int r = 0; r < W; r++) { const u64 s = all.size(); all.insert(all.end(), hb[r].begin(), hb[r].end()); for (u64 o : ho[r]) offs.push_back(o + s); }
      const size_t nbytes = all.size(); all.resize(nbytes + 64, 0);
      each_rank(R, [&](int r) { Scratch sb, so; void* db = sb.get(all.size()); IRF_CUDA(cudaMemcpy(db, all.data(), all.size(), cudaMemcpyHostToDevice)); void* d_off = nullptr;
        if (!offs.empty()) { d_off = so.get(offs.size() * 8); IRF_CUDA(cudaMemcpy(d_off, offs.data(), offs.size() * 8, cudaMemcpyHostToDevice)); }
        IRF_CUDA(cudaDeviceSynchronize()); ck(R[r], irf_import_halo(R[r], db, nbytes, d_off, offs.size(), r, W)); IRF_CUDA(cudaDeviceSynchronize()); }); }
    clk.mark("halo");
    // ---- 2. junction tables: every rank adds the other ranks' entries (finalize reads the merged tables on every rank).  The exported
    // lists stay where they are; each rank pulls the others' lists into its own scratch, device to device.
    { struct Exp { void *k, *c; size_t n; }; vector<Exp> E[3]; for (auto& e : E) e.resize(W);
      each_rank(R, [&](int r) { for (int which = 0; which < 3; which++) { Exp& e = E[which][r]; ck(R[r], irf_export_junctions(R[r], which, &e.k, &e.c, &e.n)); } IRF_CUDA(cudaDeviceSynchronize()); });
      each_rank(R, [&](int r) { for (int which = 0; which < 3; which++) {
        size_t total = 0; for (int q = 0; q < W; q++) if (q != r) total += E[which][q].n; if (!total) continue;
        uint8_t* dk = (uint8_t*)S0[r].grow(total * 8); uint8_t* dc = (uint8_t*)S1[r].grow(total * 8); size_t at = 0;
        for (int q = 0; q < W; q++) { const Exp& e = E[which][q]; if (q == r || !e.n) continue;
          IRF_CUDA(cudaMemcpy(dk + at * 8, e.k, e.n * 8, cudaMemcpyDefault)); IRF_CUDA(cudaMemcpy(dc + at * 8, e.c, e.n * 8, cudaMemcpyDefault)); at += e.n; }
        IRF_CUDA(cudaDeviceSynchronize()); ck(R[r], irf_import_junctions(R[r], which, dk, dc, total)); } }); }
    clk.mark("junctions");
    // ---- 3. small additive tables: summed on the host, written back to every rank
    { const int codes[4] = {IRF_T_SPANS, IRF_T_ROI, IRF_T_CHR, IRF_T_STATS}; vector<void*> dp[4]; size_t bytes[4]; int eb[4]; vector<vector<uint8_t>> H[4];
      for (int t = 0; t < 4; t++) { dp[t].resize(W); H[t].resize(W); size_t cnt = 0; eb[t] = 4;
        for (int r = 0; r < W; r++) { size_t c2; ck(R[r], irf_get_device_table(R[r], codes[t], &dp[t][r], &c2, &eb[t])); if (r && c2 != cnt) throw string("internal: table sizes differ between devices"); cnt = c2; }
        bytes[t] = cnt * (size_t)eb[t]; }
      each_rank(R, [&](int r) { for (int t = 0; t < 4; t++) d2h(H[t][r], dp[t][r], bytes[t]); });
      for (int t = 0; t < 4; t++) for (int r = 1; r < W; r++) {
        if (eb[t] == 8) { u64* a = (u64*)H[t][0].data(); const u64* b = (const u64*)H[t][r].data(); for (size_t i = 0; i < bytes[t] / 8; i++) a[i] += b[i]; }
        else { u32* a = (u32*)H[t][0].data(); const u32* b = (const u32*)H[t][r].data(); for (size_t i = 0; i < bytes[t] / 4; i++) a[i] += b[i]; } }
      each_rank(R, [&](int r) { for (int t = 0; t < 4; t++) if (bytes[t]) IRF_CUDA(cudaMemcpy(dp[t][r], H[t][0].data(), bytes[t], cudaMemcpyHostToDevice)); IRF_CUDA(cudaDeviceSynchronize()); }); }
    clk.mark("small_tables");
    // ---- 4. depth: region-owned.  Owners come from the first record of every shard; only counters outside a rank's own region travel.
    { vector<int32_t> rid(W), pos32(W); vector<int64_t> pos(W);
      for (int r = 0; r < W; r++) { ck(R[r], irf_first_record(R[r], &rid[r], &pos32[r])); pos[r] = pos32[r]; }
      vector<vector<u64>> FI(W); vector<vector<int32_t>> FV(W);
      each_rank(R, [&](int r) { ck(R[r], irf_owner_plan(R[r], rid.data(), pos.data(), W, r)); void *ip, *vp; size_t nf; ck(R[r], irf_export_foreign_depth(R[r], &ip, &vp, &nf)); d2h(FI[r], ip, nf); d2h(FV[r], vp, nf); });
      u64 total = 0; for (auto& v : FI) total += v.size();
      if (total > FOREIGN_DENSE_LIMIT) {
        // shards that are not genomic regions (unsorted input): sum the whole tables into rank 0 piece by piece, then hand the sum to everyone
        vector<void*> dp(W); size_t cnt = 0; int eb = 0; for (int r = 0; r < W; r++) ck(R[r], irf_get_device_table(R[r], IRF_T_DIFF, &dp[r], &cnt, &eb));
        IRF_CUDA(cudaSetDevice(ctx_device(R[0]))); const size_t piece = 16u << 20; Scratch tmp; int* t = (int*)tmp.get(piece * 4);
        for (int r = 1; r < W; r++) for (size_t lo = 0; lo < cnt; lo += piece) { const size_t k = std::min(piece, cnt - lo);
          IRF_CUDA(cudaMemcpy(t, (const int*)dp[r] + lo, k * 4, cudaMemcpyDefault)); add_i32_kernel<<<148 * 4, 256>>>((int*)dp[0] + lo, t, k); IRF_CUDA(cudaDeviceSynchronize()); }
        for (int r = 1; r < W; r++) IRF_CUDA(cudaMemcpy(dp[r], dp[0], cnt * 4, cudaMemcpyDefault));
        IRF_CUDA(cudaDeviceSynchronize());
      } else if (total) {
        vector<u64> ai; vector<int32_t> av; ai.reserve(total); av.reserve(total); for (int r = 0; r < W; r++) { ai.insert(ai.end(), FI[r].begin(), FI[r].end()); av.insert(av.end(), FV[r].begin(), FV[r].end()); }
        each_rank(R, [&](int r) { Scratch si, sv; void* di = si.get(ai.size() * 8); void* dv = sv.get(av.size() * 4);
          IRF_CUDA(cudaMemcpy(di, ai.data(), ai.size() * 8, cudaMemcpyHostToDevice)); IRF_CUDA(cudaMemcpy(dv, av.data(), av.size() * 4, cudaMemcpyHostToDevice)); IRF_CUDA(cudaDeviceSynchronize());
          ck(R[r], irf_import_depth(R[r], di, dv, ai.size())); IRF_CUDA(cudaDeviceSynchronize()); });
      }
    }
  }
  if (W > 1) clk.mark("depth_exchange");
  // ---- finalize: scan + statistics (of the owned introns when sharded)
  each_rank(R, [&](int r) { ck(R[r], irf_finalize(R[r])); IRF_CUDA(cudaDeviceSynchronize()); });
  clk.mark("finalize");
  if (W > 1) {
    // ---- 5. rows of the owned introns -> rank 0 (64-bit integer sum: exactly one rank holds a non-zero row), which writes the files
    IRF_CUDA(cudaSetDevice(ctx_device(R[0])));
    for (int which : {IRF_T_ROWS_NONDIR, IRF_T_ROWS_DIR}) {
      if (which == IRF_T_ROWS_DIR && irf_directionality(R[0]) == 0) continue;
      vector<void*> dp(W); size_t cnt = 0; int eb = 0; for (int r = 0; r < W; r++) ck(R[r], irf_get_device_table(R[r], which, &dp[r], &cnt, &eb));
      if (!cnt) continue;
      Scratch tmp; long long* t = (long long*)tmp.get(cnt * 8);
      for (int r = 1; r < W; r++) { IRF_CUDA(cudaMemcpy(t, dp[r], cnt * 8, cudaMemcpyDefault)); add_words_kernel<<<148 * 4, 256>>>((long long*)dp[0], t, cnt); IRF_CUDA(cudaDeviceSynchronize()); }
    }
    ck(R[0], irf_finalize_rows(R[0]));
    clk.mark("rows");
  }
}

}  // namespace irf
