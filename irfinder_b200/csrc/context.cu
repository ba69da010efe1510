// context.cu — irf_ctx and the C ABI declared in include/irfinder_b200.h.
//
// Host side of the path: BGZF block scan, pinned/async host->device streaming in batches that overlap the
// inflate kernels, BAM header parse (BAM2blocks::readBamHeader) from device-inflated bytes, ChrMapUpdate
// (binding name-keyed reference tables to BAM refIDs), kernel orchestration and result download.
// There is no CPU implementation of any counting step in this library: without a CUDA device every entry
// point fails with IRF_ECUDA.
#ifndef _GNU_SOURCE
#define _GNU_SOURCE
#endif
#include <cuda.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "../../include/irfinder_b200.h"
#include "common.cuh"
#include "kernels.h"
#include "multi.h"
#include "refs.h"

using namespace irf;
using std::string;
using std::vector;

namespace {

struct DevBuf {
  void* p = nullptr; size_t cap = 0;
  void ensure(size_t n) { n += 256; if (n > cap) { if (p) cudaFree(p); p = nullptr; cap = 0; IRF_CUDA(cudaMalloc(&p, n)); cap = n; } }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return (T*)p; }
};
template <class T> static void upload(DevBuf& b, const vector<T>& v, cudaStream_t s) {
  b.ensure(v.size() * sizeof(T)); if (!v.empty()) IRF_CUDA(cudaMemcpyAsync(b.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s));
}
// page-locked, device-mapped host memory for tables the GPU fetches itself (launch_host_fetch)
struct HostArena {
  void* p = nullptr; void* dp = nullptr; size_t cap = 0;
  void* get(size_t n) { n += 64; if (n > cap) { release(); IRF_CUDA(cudaHostAlloc(&p, n + n / 4, cudaHostAllocMapped)); cap = n + n / 4; IRF_CUDA(cudaHostGetDevicePointer(&dp, p, 0)); } return p; }
  const void* dev(const void* host) const { return (const uint8_t*)dp + ((const uint8_t*)host - (const uint8_t*)p); }
  void release() { if (p) cudaFreeHost(p); p = nullptr; dp = nullptr; cap = 0; }
};
struct Blk { u64 c_off; u32 c_len; u32 isize; u64 file_off; u32 crc; };
static double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
typedef CUresult (*pfn_decompress)(CUmemDecompressParams*, size_t, unsigned int, size_t*, CUstream);
typedef CUresult (*pfn_devattr)(int*, CUdevice_attribute, CUdevice);
typedef CUresult (*pfn_ctxdev)(CUdevice*);
struct DriverApi {   // the few driver-API calls of the decompression-engine path, resolved through the runtime so that the
  pfn_decompress decompress = nullptr; pfn_devattr devattr = nullptr; pfn_ctxdev ctxdev = nullptr; bool tried = false;   // library carries no link-time dependency on libcuda
  void load() {
    if (tried) return; tried = true; cudaDriverEntryPointQueryResult q;
    void* f = nullptr;
    if (cudaGetDriverEntryPoint("cuMemBatchDecompressAsync", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) decompress = (pfn_decompress)f;
    f = nullptr; if (cudaGetDriverEntryPoint("cuDeviceGetAttribute", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) devattr = (pfn_devattr)f;
    f = nullptr; if (cudaGetDriverEntryPoint("cuCtxGetDevice", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) ctxdev = (pfn_ctxdev)f;
    cudaGetLastError();
  }
  bool deflate_engine() { load(); if (!decompress || !devattr || !ctxdev) return false; CUdevice d; int mask = 0; if (ctxdev(&d) != CUDA_SUCCESS) return false;
    if (devattr(&mask, CU_DEVICE_ATTRIBUTE_MEM_DECOMPRESS_ALGORITHM_MASK, d) != CUDA_SUCCESS) return false; return (mask & CU_MEM_DECOMPRESS_ALGORITHM_DEFLATE) != 0; }
};
static DriverApi g_drv;
static u32 next_pow2(u64 v) { u64 p = 1; while (p < v) p <<= 1; return (u32)std::min<u64>(p, 1ull << 31); }

}  // namespace

struct irf_ctx {
  int device = 0; cudaStream_t s_compute = nullptr, s_copy = nullptr, s_de = nullptr, s_chk = nullptr;
  // options
  u32 trim_center = 40; int dump_fragments = 0, inflate_variant = 12, junc_cap_log2 = 24, require_eof = 1; long long entry0_override = -1; int stage_lazy = 0, verify_crc = 1, de_percent = 70, de_percent_stream = 100, planned_percent = -1, de_percent_overlap = 71, resident_pipeline = 1;
  // references (host)
  CoverageRef cov; JunctionRef jref; SpansPointRef sp; RoiRef roi; bool refs_loaded = false;
  // BAM header
  vector<string> chr_names; vector<int32_t> chr_lens; vector<u64> chr_goff; bool have_header = false, bound = false; u64 rec0 = 0;
  int adjacent_pairs = -1; bool hdr_coordinate_sorted = false;   // option "adjacent_pairs": -1 = from the header's @HD SO: tag (on unless SO:coordinate), 0 off, 1 on
  int adjacent_on() const { return adjacent_pairs < 0 ? !hdr_coordinate_sorted : adjacent_pairs != 0; }
  u64 records_coff = 0;   // irf_read_header: file offset of the first BGZF block that starts at or after the first record (no header bytes from there on)
  // device reference tables
  DevBuf d_ref_len, d_chr_goff, d_sp_pos, d_sp_range, d_sp_count, d_seg_start, d_seg_end, d_seg_coff, d_seg_range, d_diff, d_roi_start, d_roi_end, d_roi_range, d_roi_count, d_chr_count;
  DevBuf d_jc_keys, d_jc_cnt, d_jl_keys, d_jl_cnt, d_jr_keys, d_jr_cnt, d_annot_keys; u64 n_annot = 0; u32 jc_slots = 0, jside_slots = 0;
  DevBuf d_bucket_base, d_bucket_n, d_sp_bidx, d_seg_bidx, d_roi_bidx;
  DevBuf d_stats, d_err, d_introns, d_iblocks, d_rows, d_scratch, d_tile_sums, d_small, d_dump, d_dump_cursor; u64 dump_cap = 0;
  DevBuf d_exp_keys[3], d_exp_cnt[3], d_sort_tmp, d_sort_keys, d_sort_cnt, d_junc_rows, d_overflow;
  // staged input
  DevBuf d_tok, d_slots, d_sm_list, d_de_list, d_de_act, d_crc, d_comp, d_U, d_c_off, d_c_len, d_u_off, d_u_len, d_status, d_entry, d_exit, d_nrec, d_ncand, d_ferr, d_mask, d_mism, d_ptab, d_left, d_halo, d_halo_off, d_hptab, d_rec_base, d_rec_off, d_pair_list, d_single_list;
  const uint8_t* h_src = nullptr; size_t src_total = 0; bool resident = false;
  vector<Blk> blks; int nblk_own = 0; u64 Uend = 0; bool staged = false, staged_has_header = false, inflated_input = false;
  u64 pslots = 0, n_cand = 0, n_rec = 0, frame_entry0 = 0, frame_exit_rel = 0; bool sharded = false; u64 halo_leftover = 0;
  void* pinned = nullptr; size_t pinned_cap = 0;
  // results
  vector<u64> h_stats; vector<u32> h_sp, h_roi, h_chr, h_dump; vector<JuncEntry> h_junc; vector<IntronRow> rows_nondir, rows_dir; vector<int32_t> h_depth;
  int directionality = 0; string stdout_text, dir_log; double timing[14] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}; unsigned long long launches0 = 0;
  bool counted = false, finalized = false;
  string err;
  vector<irf_ctx*> all_ranks() { vector<irf_ctx*> v{this}; v.insert(v.end(), peers.begin(), peers.end()); return v; }
  vector<irf_ctx*> peers;   // irf_create with n_devices > 1: the contexts of device_ids[1..]; this context is shard 0 and holds the merged tables after a run (multi.cu)

  DevTables T{}; Rules rules;

  // ---------------------------------------------------------------- set-up
  void init(int dev) {
    int n = 0; cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) throw string("no CUDA device: this library has no CPU path (") + cudaGetErrorString(e) + ")";
    device = dev; IRF_CUDA(cudaSetDevice(dev));
    { int lo_p = 0, hi_p = 0; IRF_CUDA(cudaDeviceGetStreamPriorityRange(&lo_p, &hi_p));   // counting kernels take freed SM slots ahead of the long inflate launch on the side stream
      IRF_CUDA(cudaStreamCreateWithPriority(&s_compute, cudaStreamNonBlocking, hi_p)); IRF_CUDA(cudaStreamCreateWithPriority(&s_copy, cudaStreamNonBlocking, lo_p)); } IRF_CUDA(cudaStreamCreateWithFlags(&s_de, cudaStreamNonBlocking)); IRF_CUDA(cudaStreamCreateWithFlags(&s_chk, cudaStreamNonBlocking));
    d_stats.ensure(ST_SLOTS * 8); d_err.ensure(4); d_small.ensure(256);
  }
  void sync() { IRF_CUDA(cudaStreamSynchronize(s_compute)); }

  void load_ref(const char* cover, const char* sj, const char* rc, const char* roi_path) {
    { std::ifstream f(cover); if (!f) throw string("cannot open ") + cover; cov.loadRef(f); }
    { std::ifstream f(sj); if (!f) throw string("cannot open ") + sj; jref.loadRef(f); }
    { std::ifstream f(rc); if (!f) throw string("cannot open ") + rc; sp.loadRef(f); }   // SpansPoint::setSpanLength(5,4) is compiled in (common.cuh)
    roi = RoiRef();
    if (roi_path && strcmp(roi_path, "NULL") != 0 && roi_path[0]) { std::ifstream f(roi_path); if (!f) throw string("cannot open ") + roi_path; roi.loadRef(f); }
    refs_loaded = true; bound = false; owner_key.clear();
    // chromosome-independent device tables
    upload(d_sp_pos, sp.pos, s_compute); upload(d_seg_start, cov.seg_start, s_compute); upload(d_seg_end, cov.seg_end, s_compute); upload(d_seg_coff, cov.seg_coff, s_compute);
    vector<u32> rs, re; for (auto& r : roi.rois) { rs.push_back(r.start); re.push_back(r.end); }
    upload(d_roi_start, rs, s_compute); upload(d_roi_end, re, s_compute);
    d_sp_count.ensure(sp.pos.size() * 8); d_roi_count.ensure(roi.rois.size() * 8); d_diff.ensure(cov.P * 8); d_tile_sums.ensure(depth_scan_tiles(cov.P * 2) * 4);
    jc_slots = 1u << junc_cap_log2; jside_slots = jc_slots >> 1;
    if ((u64)jc_slots < 4 * jref.size()) jc_slots = next_pow2(4 * jref.size()), jside_slots = jc_slots >> 1;
    d_jc_keys.ensure((u64)jc_slots * 8); d_jc_cnt.ensure((u64)jc_slots * 8); d_jl_keys.ensure((u64)jside_slots * 8); d_jl_cnt.ensure((u64)jside_slots * 8);
    d_jr_keys.ensure((u64)jside_slots * 8); d_jr_cnt.ensure((u64)jside_slots * 8);
    sync();
  }

  // every processor's ChrMapUpdate: bind name-keyed tables to the BAM's refID order
  void chr_map_update() {
    if (!refs_loaded) throw string("irf_load_ref must be called before a run");
    int n = (int)chr_names.size();
    chr_goff.assign(n + 1, 0); for (int i = 0; i < n; i++) chr_goff[i + 1] = chr_goff[i] + (u64)std::max(chr_lens[i], 0) + 1;
    if (chr_goff[n] >= (1ull << 36)) throw string("genome larger than 2^36 bases is not supported by the junction key");
    vector<Range> spr(n), sgr(n), ror(n);
    for (int i = 0; i < n; i++) {
      auto a = sp.chr_range.find(chr_names[i]); spr[i] = a == sp.chr_range.end() ? Range{0, 0} : a->second;
      auto b = cov.chr_range.find(chr_names[i]); sgr[i] = b == cov.chr_range.end() ? Range{0, 0} : b->second;
      auto c = roi.chr_range.find(chr_names[i]); ror[i] = c == roi.chr_range.end() ? Range{0, 0} : c->second;
    }
    upload(d_ref_len, chr_lens, s_compute); upload(d_chr_goff, chr_goff, s_compute); upload(d_sp_range, spr, s_compute); upload(d_seg_range, sgr, s_compute); upload(d_roi_range, ror, s_compute);
    d_chr_count.ensure((size_t)n * 8);
    // coordinate-bucket indices (common.cuh BUCKET_SHIFT)
    { vector<u64> bbase(n); vector<u32> bn(n); u64 tot = 0; for (int i = 0; i < n; i++) { bbase[i] = tot; bn[i] = (u32)((u64)std::max(chr_lens[i], 0) >> BUCKET_SHIFT) + 2; tot += bn[i]; }
      vector<u32> spb(tot), sgb(tot), rob(tot);
      for (int i = 0; i < n; i++) {
        u32 a = spr[i].begin, b = sgr[i].begin, c = ror[i].begin;
        for (u32 k = 0; k < bn[i]; k++) {
          u64 start = (u64)k << BUCKET_SHIFT;
          while (a < spr[i].end && sp.pos[a] < start) a++;                 // first point >= bucket start
          while (b < sgr[i].end && cov.seg_end[b] <= start) b++;           // first segment with end > bucket start
          while (c < ror[i].end && roi.rois[c].end <= start) c++;
          spb[bbase[i] + k] = a; sgb[bbase[i] + k] = b; rob[bbase[i] + k] = c;
        }
      }
      upload(d_bucket_base, bbase, s_compute); upload(d_bucket_n, bn, s_compute); upload(d_sp_bidx, spb, s_compute); upload(d_seg_bidx, sgb, s_compute); upload(d_roi_bidx, rob, s_compute); sync(); }
    // annotated junction keys (JunctionCount::loadRef entries on chromosomes present in the BAM)
    std::map<string, int> id; for (int i = n - 1; i >= 0; i--) id[chr_names[i]] = i;
    vector<u64> keys;
    for (auto& c : jref.chr_junc) { auto it = id.find(c.first); if (it == id.end()) continue; u64 base = chr_goff[it->second];
      for (auto& j : c.second) { u64 gap = (u64)j.first.second - j.first.first; if (j.first.second >= j.first.first && gap < (1ull << JKEY_GAP_BITS)) keys.push_back(jkey_make(base + j.first.first, gap)); } }
    n_annot = keys.size(); upload(d_annot_keys, keys, s_compute);
    // intron descriptors
    vector<IntronDesc> descs; vector<IntronBlock> ib; descs.reserve(cov.BEDrecords.size());
    for (auto& rec : cov.BEDrecords) {
      IntronDesc d; d.blk_begin = (u32)ib.size();
      for (auto& b : rec.blocks) cov.pieces(rec.chrName, b.first, b.second, ib);
      d.blk_main_end = (u32)ib.size();
      const u32 EDGE_WINDOW_OFFSET = rules.edge_offset, EDGE_WINDOW_LEN = rules.edge_len;   // [C-13]
      u32 ws = rec.start + EDGE_WINDOW_OFFSET, we = rec.end > EDGE_WINDOW_OFFSET ? rec.end - EDGE_WINDOW_OFFSET : 0;
      cov.pieces(rec.chrName, ws, std::min(ws + EDGE_WINDOW_LEN, rec.end), ib); d.blk_first_end = (u32)ib.size();
      cov.pieces(rec.chrName, (we > EDGE_WINDOW_LEN && we - EDGE_WINDOW_LEN > rec.start) ? we - EDGE_WINDOW_LEN : rec.start, we, ib); d.blk_last_end = (u32)ib.size();
      d.start = rec.start; d.end = rec.end; auto it = id.find(rec.chrName); d.refid = it == id.end() ? -1 : it->second;
      d.sp_left = sp.lookup_index(rec.chrName, rec.start); d.sp_right = sp.lookup_index(rec.chrName, rec.end); d.plus = rec.direction ? 1 : 0;
      descs.push_back(d);
    }
    upload(d_introns, descs, s_compute); upload(d_iblocks, ib, s_compute); d_rows.ensure(descs.size() * sizeof(IntronRow));
    if (dump_fragments) { dump_cap = 1ull << 26; d_dump.ensure(dump_cap * 4); d_dump_cursor.ensure(8); }
    T.n_ref = n; T.ref_len = d_ref_len.as<int32_t>(); T.chr_goff = d_chr_goff.as<u64>();
    T.bucket_base = d_bucket_base.as<u64>(); T.bucket_n = d_bucket_n.as<u32>(); T.sp_bidx = d_sp_bidx.as<u32>(); T.seg_bidx = d_seg_bidx.as<u32>(); T.roi_bidx = d_roi_bidx.as<u32>();
    T.sp_pos = d_sp_pos.as<u32>(); T.sp_range = d_sp_range.as<Range>(); T.sp_count = d_sp_count.as<u32>(); T.n_sp = (u32)sp.pos.size();
    T.seg_start = d_seg_start.as<u32>(); T.seg_end = d_seg_end.as<u32>(); T.seg_coff = d_seg_coff.as<u64>(); T.seg_range = d_seg_range.as<Range>(); T.diff = d_diff.as<int32_t>(); T.P = cov.P;
    T.roi_start = d_roi_start.as<u32>(); T.roi_end = d_roi_end.as<u32>(); T.roi_range = d_roi_range.as<Range>(); T.roi_count = d_roi_count.as<u32>(); T.n_roi = (u32)roi.rois.size();
    T.chr_count = d_chr_count.as<u32>();
    T.jc_keys = d_jc_keys.as<u64>(); T.jc_cnt = d_jc_cnt.as<u32>(); T.jc_mask = jc_slots - 1;
    T.jl_keys = d_jl_keys.as<u64>(); T.jl_cnt = d_jl_cnt.as<u32>(); T.jl_mask = jside_slots - 1;
    T.jr_keys = d_jr_keys.as<u64>(); T.jr_cnt = d_jr_cnt.as<u32>(); T.jr_mask = jside_slots - 1;
    T.stats = d_stats.as<u64>(); T.err = d_err.as<int>();
    T.rules = rules;
    T.dump = dump_fragments ? d_dump.as<u32>() : nullptr; T.dump_cursor = dump_fragments ? d_dump_cursor.as<u64>() : nullptr; T.dump_cap = dump_cap;
    bound = true; owner_key.clear(); sync();
    reset(true);   // in the middle of a count: the inflate / CRC checks of this very run may already have raised the error flag
  }

  // keep_err: zero the tables but leave the device error flag alone.  chr_map_update() runs inside count() after the first
  // bytes have been inflated, while the CRC-32 and length checks of this run are queued or running on the side streams; clearing
  // the flag there would lose a mismatch they report (the flag is zeroed by the API-level irf_reset / at the start of a count on
  // an unbound context instead).
  void reset(bool keep_err = false) {
    if (!keep_err) join_crc();
    if (!bound) { counted = finalized = false; return; }
    cudaStream_t s = s_compute;
    IRF_CUDA(cudaMemsetAsync(d_sp_count.p, 0, sp.pos.size() * 8, s)); IRF_CUDA(cudaMemsetAsync(d_roi_count.p, 0, roi.rois.size() * 8, s));
    IRF_CUDA(cudaMemsetAsync(d_chr_count.p, 0, chr_names.size() * 8, s)); IRF_CUDA(cudaMemsetAsync(d_diff.p, 0, cov.P * 8, s));
    IRF_CUDA(cudaMemsetAsync(d_jc_keys.p, 0xff, (u64)jc_slots * 8, s)); IRF_CUDA(cudaMemsetAsync(d_jc_cnt.p, 0, (u64)jc_slots * 8, s));
    IRF_CUDA(cudaMemsetAsync(d_jl_keys.p, 0xff, (u64)jside_slots * 8, s)); IRF_CUDA(cudaMemsetAsync(d_jl_cnt.p, 0, (u64)jside_slots * 8, s));
    IRF_CUDA(cudaMemsetAsync(d_jr_keys.p, 0xff, (u64)jside_slots * 8, s)); IRF_CUDA(cudaMemsetAsync(d_jr_cnt.p, 0, (u64)jside_slots * 8, s));
    IRF_CUDA(cudaMemsetAsync(d_stats.p, 0, ST_SLOTS * 8, s)); if (!keep_err) IRF_CUDA(cudaMemsetAsync(d_err.p, 0, 4, s));
    if (dump_fragments) IRF_CUDA(cudaMemsetAsync(d_dump_cursor.p, 0, 8, s));
    launch_ht_insert_keys(T.jc_keys, T.jc_mask, d_annot_keys.as<u64>(), n_annot, T.err, s);
    counted = finalized = false; sharded = false; halo_leftover = 0; owner_mode = false; rows_pending = false; n_carry = 0; carry_bytes = 0; launches0 = launch_counter();
  }

  // ---------------------------------------------------------------- BGZF scan (SAM v1 §4.1)
  // Parses one block header at `off`; returns false when the bytes there cannot be a complete block inside [0, n_total).
  static bool parse_block(const uint8_t* p, size_t off, size_t n_total, Blk& b, size_t& total, string* why) {
    if (off + 18 > n_total) { if (why) *why = "truncated BGZF block header"; return false; }
    const uint8_t* h = p + off;
    if (h[0] != 31 || h[1] != 139 || h[2] != 8 || !(h[3] & 4)) { if (why) *why = string("not a BGZF block at offset ") + std::to_string(off); return false; }
    u32 xlen = h[10] | (u32)h[11] << 8; if (off + 12 + xlen > n_total) { if (why) *why = "truncated BGZF extra field"; return false; }
    int bsize = -1; for (u32 x = 0; x + 4 <= xlen;) { const uint8_t* sf = h + 12 + x; u32 slen = sf[2] | (u32)sf[3] << 8;
      if (x + 4 + slen > xlen) { if (why) *why = string("BGZF extra field with an inconsistent subfield at offset ") + std::to_string(off); return false; }   // the payload must lie inside the extra field
      if (sf[0] == 'B' && sf[1] == 'C' && slen == 2) { bsize = sf[4] | (int)sf[5] << 8; break; } x += 4 + slen; }
    if (bsize < 0) { if (why) *why = string("BGZF block without BC subfield at offset ") + std::to_string(off); return false; }
    total = (size_t)bsize + 1; if (total < 12 + xlen + 8) { if (why) *why = "bad BGZF BSIZE"; return false; }
    if (off + total > n_total) { if (why) *why = "truncated BGZF block (file cut short)"; return false; }
    b.file_off = off; b.c_off = off + 12 + xlen; b.c_len = (u32)(total - 12 - xlen - 8); memcpy(&b.isize, p + off + total - 4, 4); memcpy(&b.crc, p + off + total - 8, 4);
    if (b.isize > 65536) { if (why) *why = "BGZF ISIZE > 64 KiB"; return false; }
    return true;
  }
  // Serial walk of the BSIZE chain from `off` to `stop` (exclusive); throws on a malformed block unless it lies in the look-ahead.
  void scan_range(const uint8_t* p, size_t off, size_t stop, size_t n_own, size_t n_total, vector<Blk>& out, size_t& end_off) {
    while (off < stop) {
      Blk b; size_t total = 0; string why;
      if (!parse_block(p, off, n_total, b, total, &why)) { if (off >= n_own && why.find("truncated") != string::npos) break; throw why; }
      out.push_back(b); off += total;
    }
    end_off = off;
  }
  // The chain is serial (each header gives the next offset) and latency-bound on a 10+ GB buffer, so large inputs are cut into
  // segments: each worker finds a block start by magic + 3 verified hops, walks its segment, and the joins are checked
  // (a worker must land exactly on the next worker's start); any disagreement falls back to the serial walk.
  void scan_bgzf(const uint8_t* p, size_t n_own, size_t n_total) {
    blks.clear(); nblk_own = 0;
    unsigned nt = std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
    bool done = false;
    if (n_total >= (256ull << 20) && nt > 1) {
      vector<size_t> start(nt + 1, n_total); start[0] = 0; bool ok = true;
      for (unsigned t = 1; t < nt && ok; t++) {
        size_t off = n_total / nt * t, lim = std::min(n_total, off + (1u << 20)); bool found = false;
        for (; off + 18 <= lim; off++) {
          if (p[off] != 31 || p[off + 1] != 139 || p[off + 2] != 8 || !(p[off + 3] & 4)) continue;
          size_t cur = off; int hops = 0; for (; hops < 3 && cur < n_total; hops++) { Blk b; size_t total; if (!parse_block(p, cur, n_total, b, total, nullptr)) break; cur += total; }
          if (hops == 3 || cur >= n_total) { found = true; break; }
        }
        if (!found) ok = false; else start[t] = off;
      }
      if (ok) {
        vector<vector<Blk>> parts(nt); vector<size_t> ends(nt, 0); vector<string> errs(nt); vector<std::thread> th;
        for (unsigned t = 0; t < nt; t++) th.emplace_back([&, t]() { try { scan_range(p, start[t], start[t + 1], n_own, n_total, parts[t], ends[t]); } catch (const string& e) { errs[t] = e; } });
        for (auto& x : th) x.join();
        for (unsigned t = 0; t < nt && ok; t++) { if (!errs[t].empty()) ok = false; else if (t + 1 < nt && ends[t] != start[t + 1]) ok = false; }
        if (ok) { size_t tot = 0; for (auto& v : parts) tot += v.size(); blks.reserve(tot); for (auto& v : parts) blks.insert(blks.end(), v.begin(), v.end()); done = true; }
      }
    }
    if (!done) { blks.clear(); size_t e = 0; scan_range(p, 0, n_total, n_own, n_total, blks, e); }
    for (auto& b : blks) if (b.file_off < n_own) nblk_own++;
  }

  void stage(const void* src, size_t n_own, size_t n_total, int has_header, bool copy_now) {
    if (n_own > n_total) throw string("nbytes_own > nbytes_total");
    join_crc(); IRF_CUDA(cudaStreamSynchronize(s_copy)); drop_pre_copies();
    double t0 = now_ms();
    inflated_input = false;
    feed_stop();
    if (feed_fd >= 0 && !copy_now && ((has_header && n_own == n_total) || have_header) && n_total > 0 && feed_capable()) { d_comp.ensure(n_total + 64); feed_start(n_total); }
    else if (!copy_now && stream_pipeline && has_header && n_own == n_total && n_total > 0 && de_percent_stream == 100 && inflate_variant != 8 && g_drv.deflate_engine()) {
      // end-to-end path: the bytes can start crossing PCIe before the host has even walked the BGZF block chain
      d_comp.ensure(n_total + 64);
      // Every chunk is queued now when the references are already bound to this header (any run after a context's first):
      // nothing else needs the DMA engine then, the block tables are fetched by a kernel.  A first run still has the bound
      // reference tables to upload once the header is known, and those would queue behind the whole file: two chunks only.
      const size_t max_pre = bound ? (size_t)-1 : 2;
      for (size_t lo = 0; lo < n_total && pre_ev_copy.size() < max_pre; lo += chunk_bytes) { size_t hi = std::min(n_total, lo + chunk_bytes);
        IRF_CUDA(cudaMemcpyAsync(d_comp.as<uint8_t>() + lo, (const uint8_t*)src + lo, hi - lo, cudaMemcpyHostToDevice, s_copy));
        cudaEvent_t ev; IRF_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)); IRF_CUDA(cudaEventRecord(ev, s_copy)); pre_ev_copy.push_back(ev); }
      pre_copied_bytes = std::min(n_total, pre_ev_copy.size() * chunk_bytes);
    }
    scan_bgzf((const uint8_t*)src, n_own, n_total);
    if (require_eof && has_header && n_own == n_total) { if (blks.empty() || blks.back().isize != 0) throw string("BGZF EOF marker missing (truncated file?)"); }
    timing[6] = now_ms() - t0;
    h_src = (const uint8_t*)src; src_total = blks.empty() ? 0 : (size_t)(blks.back().c_off + blks.back().c_len + 8);
    // block tables: written into the page-locked arena and fetched by a kernel (see launch_host_fetch)
    { const size_t nbk = blks.size(), a8 = (nbk * 8 + 15) & ~(size_t)15, a4 = (nbk * 4 + 15) & ~(size_t)15; uint8_t* base = (uint8_t*)h_tab.get(2 * a8 + 3 * a4);
      u64* c_off = (u64*)base; u64* u_off = (u64*)(base + a8); u32* c_len = (u32*)(base + 2 * a8); u32* u_len = (u32*)(base + 2 * a8 + a4); u32* crcs = (u32*)(base + 2 * a8 + 2 * a4); u64 u = 0;
      for (size_t i = 0; i < nbk; i++) { c_off[i] = blks[i].c_off; c_len[i] = blks[i].c_len; u_off[i] = u; u_len[i] = blks[i].isize; crcs[i] = blks[i].crc; u += blks[i].isize; }
      Uend = u;
      auto fetch = [&](DevBuf& b, const void* src, size_t bytes) { b.ensure(bytes + 16); launch_host_fetch(b.p, h_tab.dev(src), bytes, s_compute); };
      fetch(d_c_off, c_off, nbk * 8); fetch(d_u_off, u_off, nbk * 8); fetch(d_c_len, c_len, nbk * 4); fetch(d_u_len, u_len, nbk * 4); fetch(d_crc, crcs, nbk * 4); }
    if (pre_ev_copy.empty() && !feed_active) d_comp.ensure(src_total + 64);
    d_U.ensure(Uend + 64 + (seg_mode ? carry_bytes + 256 : 0));
    size_t nb = blks.size(); d_status.ensure(nb * 4); d_entry.ensure(nb * 8); d_exit.ensure(nb * 8); d_nrec.ensure(nb * 4); d_ncand.ensure(nb * 4); d_ferr.ensure(nb); d_mask.ensure(nb); d_mism.ensure(nb);
    sync();  // allocation changes above settle before anything is queued against the buffers
    resident = false; staged = true; staged_has_header = has_header != 0; engines_planned = false;
    if (copy_now) { if (src_total) IRF_CUDA(cudaMemcpyAsync(d_comp.p, h_src, src_total, cudaMemcpyHostToDevice, s_compute)); sync(); resident = true; h_src = nullptr; }
  }
  // Reference stdin mode with an already-inflated stream (`gzip -cd | irfinder ... /dev/stdin`): the bytes are cut into 64 KiB
  // pseudo-blocks so that record framing works as for BGZF input.  n_own < n_total: a segment of a longer stream with look-ahead
  // bytes (n_own a multiple of 64 KiB); has_header = 0: a later segment.
  void stage_inflated(const void* src, size_t n_own, size_t n_total, bool has_header) {
    inflated_input = true; blks.clear(); const u32 CH = 65536; const size_t n = n_total; size_t nb = (n + CH - 1) / CH; vector<u64> u_off(nb); vector<u32> u_len(nb);
    if (n_own < n_total && n_own % CH) throw string("internal: a segment of an inflated stream must end on a 64 KiB boundary");
    for (size_t i = 0; i < nb; i++) { u_off[i] = (u64)i * CH; u_len[i] = (u32)std::min<size_t>(CH, n - i * CH); blks.push_back(Blk{0, 0, u_len[i], (u64)i * CH}); }
    nblk_own = (int)((n_own + CH - 1) / CH); Uend = n; upload(d_u_off, u_off, s_compute); upload(d_u_len, u_len, s_compute); d_U.ensure(n + 64 + (seg_mode ? carry_bytes + 256 : 0));
    IRF_CUDA(cudaMemcpyAsync(d_U.p, src, n, cudaMemcpyHostToDevice, s_compute)); IRF_CUDA(cudaMemsetAsync(d_U.as<uint8_t>() + n, 0, 64, s_compute));
    d_status.ensure(nb * 4); d_entry.ensure(nb * 8); d_exit.ensure(nb * 8); d_nrec.ensure(nb * 4); d_ncand.ensure(nb * 4); d_ferr.ensure(nb); d_mask.ensure(nb); d_mism.ensure(nb);
    sync(); staged = true; staged_has_header = has_header; resident = true;
  }

  // S1: one persistent inflate launch.  When the bytes are still on the host the copy runs beside it in 256 MB chunks on
  // the copy stream, each followed by a tiny stream-ordered "watermark" copy that tells the kernel how many bytes have
  // landed; warps whose next blocks lie beyond the watermark wait for it.  (Batched launches were tried first: every launch
  // ends with tens of milliseconds of partly idle SMs because one lane needs ~40 ms per 64 KiB block.)
  // Which blocks go to the hardware decompression engine (DE) and which to the SM kernel.  The file is cut into the same
  // 256 MB chunks the streamed copy uses; inside every chunk the first de_percent % of the blocks are the DE's, the rest
  // the kernel's, so both engines always have landed data to work on.
  // mode 1 (resident overlap): the SM kernel takes the leading chunks whole, the DE the last de_percent % of the chunks (k_de of them).
  size_t k_de = 0;
  double sm_byte_share() const {   // fraction of the algorithmic bytes (compressed in + inflated out) that inflate_kernel handles under the current plan
    double all = 0, sm = 0; for (auto& b : blks) all += (double)b.c_len + b.isize;
    if (de_list.empty()) return all > 0 ? 1.0 : 0.0;
    for (u32 i : sm_list) sm += (double)blks[i].c_len + blks[i].isize;
    return all > 0 ? sm / all : 0.0;
  }
  // BGZF blocks one full grid of the SM inflate kernel holds: the two-phase kernel's geometry for this device (148 SMs x 7 CTAs x 128 lanes), else what the last launch used
  size_t overlap_lanes() const { if (inflate_variant >= 12 && inflate_variant <= 17) return (size_t)inflate_tok_lanes(inflate_variant); int l = inflate_resident_lanes(); return l > 0 ? (size_t)l : 94720; }
  void plan_engines(int de_percent, int mode = 0) {
    sync();   // a fetch of the previous plan's lists may still be queued against the arena rewritten below
    const size_t nb = blks.size(), chunk = chunk_bytes;
    sm_list.clear(); de_list.clear(); de_chunk_end.clear(); sm_chunk_end.clear(); chunk_blk_end.clear();
    const bool de_ok = de_percent > 0 && inflate_variant != 8 && g_drv.deflate_engine();
    const size_t nch = std::max<size_t>((src_total + chunk - 1) / chunk, 1);
    for (size_t c = 0, b1 = 0; c < nch; c++) { size_t hi = std::min(src_total, (c + 1) * chunk);
      while (b1 < nb && (blks[b1].c_off + blks[b1].c_len + 8 <= hi || c + 1 >= nch)) b1++;
      chunk_blk_end.push_back(b1); }
    size_t k_sm = 0;
    if (mode == 1 && de_ok) {   // leading chunks for the kernel: its share of the blocks (count_streaming launches it one grid-full of blocks at a time)
      size_t target = nb * (size_t)(100 - de_percent) / 100;
      { const size_t lanes = overlap_lanes(); if (target >= lanes) target = target / lanes * lanes; }   // whole grid-fulls: a launch over a fraction of the grid takes as long as a full one
      while (k_sm + 1 < nch && chunk_blk_end[k_sm] <= target) k_sm++;
    }
    k_de = nch - k_sm;
    for (size_t c = 0, b0 = 0; c < nch; c++) {
      size_t b1 = chunk_blk_end[c];
      size_t nde = !de_ok ? 0 : mode == 1 ? (c >= k_sm ? b1 - b0 : 0) : (b1 - b0) * (size_t)de_percent / 100;
      for (size_t b = b0; b < b1; b++) { if (blks[b].isize == 0) continue; if (b < b0 + nde) de_list.push_back((u32)b); else sm_list.push_back((u32)b); }
      de_chunk_end.push_back(de_list.size()); sm_chunk_end.push_back(sm_list.size()); b0 = b1;
    }
    de_params.resize(de_list.size()); d_de_act.ensure(nb * 4);
    for (size_t i = 0; i < de_list.size(); i++) { const Blk& k = blks[de_list[i]]; CUmemDecompressParams& q = de_params[i]; memset(&q, 0, sizeof q);
      q.srcNumBytes = k.c_len; q.dstNumBytes = k.isize; q.dstActBytes = d_de_act.as<cuuint32_t>() + de_list[i]; q.src = d_comp.as<uint8_t>() + k.c_off; q.algo = CU_MEM_DECOMPRESS_ALGORITHM_DEFLATE; }
    { u64 u = 0; size_t j = 0; for (size_t bI = 0; bI < nb && j < de_list.size(); bI++) { if (de_list[j] == bI) { de_params[j].dst = d_U.as<uint8_t>() + u; j++; } u += blks[bI].isize; } }
    { const size_t a1 = (sm_list.size() * 4 + 15) & ~(size_t)15, a2 = (de_list.size() * 4 + 15) & ~(size_t)15; uint8_t* base = (uint8_t*)h_lists.get(a1 + a2);
      if (!sm_list.empty()) memcpy(base, sm_list.data(), sm_list.size() * 4); if (!de_list.empty()) memcpy(base + a1, de_list.data(), de_list.size() * 4);
      d_sm_list.ensure(a1 + 16); d_de_list.ensure(a2 + 16); launch_host_fetch(d_sm_list.p, h_lists.dev(base), sm_list.size() * 4, s_compute); launch_host_fetch(d_de_list.p, h_lists.dev(base + a1), de_list.size() * 4, s_compute); }
    engines_planned = true; planned_percent = de_percent + 1000 * mode;
  }
  HostArena h_tab, h_lists;
  vector<cudaEvent_t> pre_ev_copy; size_t pre_copied_bytes = 0;
  void drop_pre_copies() { for (auto e : pre_ev_copy) cudaEventDestroy(e); pre_ev_copy.clear(); pre_copied_bytes = 0; }
  bool crc_pending = false;
  void join_crc() { if (crc_pending) { IRF_CUDA(cudaStreamSynchronize(s_de)); crc_pending = false; } }
  size_t chunk_bytes = 256ull << 20; vector<size_t> chunk_blk_end; int stream_pipeline = 1;
  vector<u32> sm_list, de_list; vector<size_t> de_chunk_end, sm_chunk_end; vector<CUmemDecompressParams> de_params; bool engines_planned = false;

  // S1: BGZF inflate on two engines at once.  The Blackwell decompression engine (cuMemBatchDecompressAsync, raw DEFLATE,
  // no SM involved) takes its share of every chunk; one persistent launch of inflate_kernel takes the rest.  When the bytes
  // are still on the host the copy runs beside both on the copy stream in 256 MB chunks: the DE batch of a chunk is
  // stream-ordered behind that chunk's copy, and the kernel follows an 8-byte "watermark" copied after every chunk.
  void inflate_all() {
    if (inflated_input) return;
    size_t nb = blks.size(); if (!nb) return;
    // resident bytes: the two engines share the work (70/30 measured best); bytes streamed from the host arrive at PCIe speed,
    // which the decompression engine alone keeps up with, so the SMs stay out of the way (measured: 288 ms vs 333 ms mixed)
    const int want_percent = resident ? de_percent : de_percent_stream;
    if (!engines_planned || planned_percent != want_percent) plan_engines(want_percent);
    tok_scratch(de_list.empty() ? (int)nb : (int)sm_list.size());   // allocated (if it has to grow) before anything is queued
    IRF_CUDA(cudaMemsetAsync(d_U.as<uint8_t>() + Uend, 0, 64, s_compute));
    unsigned long long* wm = (unsigned long long*)(d_small.as<uint8_t>() + 224);
    const bool stream_in = !resident && inflate_variant != 8;
    if (!resident && !stream_in) { IRF_CUDA(cudaMemcpyAsync(d_comp.p, h_src, src_total, cudaMemcpyHostToDevice, s_compute)); }
    if (stream_in) IRF_CUDA(cudaMemsetAsync(wm, 0, 8, s_compute));
    const bool listed = !de_list.empty();
    InflateArgs a{d_comp.as<uint8_t>(), d_c_off.as<u64>(), d_c_len.as<u32>(), d_U.as<uint8_t>(), d_u_off.as<u64>(), d_u_len.as<u32>(), listed ? (int)sm_list.size() : (int)nb, d_status.as<int>(),
                  T.err ? T.err : d_err.as<int>(), (unsigned int*)(d_small.as<uint8_t>() + 160), stream_in ? wm : nullptr, listed ? d_sm_list.as<u32>() : nullptr};
    a.tok_scratch = tok_scratch(a.nblk); a.tok_scratch_bytes = d_tok.cap;
    cudaEvent_t k0, k1, ready, de_done; IRF_CUDA(cudaEventCreate(&k0)); IRF_CUDA(cudaEventCreate(&k1));
    IRF_CUDA(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming)); IRF_CUDA(cudaEventCreateWithFlags(&de_done, cudaEventDisableTiming));
    IRF_CUDA(cudaEventRecord(k0, s_compute));
    IRF_CUDA(cudaEventRecord(ready, s_compute)); IRF_CUDA(cudaStreamWaitEvent(s_copy, ready, 0)); IRF_CUDA(cudaStreamWaitEvent(s_de, ready, 0));   // tables + zeroed watermark first
    auto submit_de = [&](size_t from, size_t to) {
      if (to <= from) return; size_t bad = 0;
      CUresult r = g_drv.decompress(de_params.data() + from, to - from, 0, &bad, (CUstream)s_de);
      if (r != CUDA_SUCCESS) throw string("cuMemBatchDecompressAsync failed with CUresult ") + std::to_string((int)r) + " (cuda)";
    };
    vector<cudaEvent_t> chunk_ev;
    if (stream_in) {
      // Copies (and the DE batches behind them) are queued BEFORE the kernel is launched: if the driver maps two of these
      // streams onto one hardware queue the kernel then simply runs after them instead of waiting for ever.
      const size_t chunk = chunk_bytes; size_t nch = (src_total + chunk - 1) / chunk;
      if (wm_host_cap < nch) { if (wm_host) cudaFreeHost(wm_host); IRF_CUDA(cudaMallocHost((void**)&wm_host, (nch + 8) * 8)); wm_host_cap = nch + 8; }
      for (size_t i = 0; i < nch; i++) {
        size_t lo = i * chunk, hi = std::min(src_total, lo + chunk);
        IRF_CUDA(cudaMemcpyAsync(d_comp.as<uint8_t>() + lo, h_src + lo, hi - lo, cudaMemcpyHostToDevice, s_copy));
        wm_host[i] = (i + 1 == nch) ? hi + 4096 : hi;   // the last mark also covers the kernel's look-ahead margin
        IRF_CUDA(cudaMemcpyAsync(wm, &wm_host[i], 8, cudaMemcpyHostToDevice, s_copy));
        if (listed) { cudaEvent_t ev; IRF_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)); IRF_CUDA(cudaEventRecord(ev, s_copy)); IRF_CUDA(cudaStreamWaitEvent(s_de, ev, 0)); chunk_ev.push_back(ev);
          size_t f0 = i ? de_chunk_end[i - 1] : 0, f1 = de_chunk_end[i]; submit_de(f0, f1);
          if (verify_crc && f1 > f0) launch_crc_check(d_U.as<uint8_t>(), d_u_off.as<u64>(), d_u_len.as<u32>(), d_crc.as<u32>(), (int)(f1 - f0), d_status.as<int>(), a.err, s_de, d_de_list.as<u32>() + f0); }
      }
    } else if (listed) { const size_t cap = 32768;   // very large batches (> ~150k operations) were measured 2.5x slower per byte
      for (size_t i = 0; i < de_list.size(); i += cap) submit_de(i, std::min(de_list.size(), i + cap)); }
    launch_inflate(a, inflate_variant, s_compute);
    if (listed) { launch_de_check(d_de_list.as<u32>(), (int)de_list.size(), d_de_act.as<u32>(), d_u_len.as<u32>(), d_status.as<int>(), a.err, s_de);
      IRF_CUDA(cudaEventRecord(de_done, s_de)); IRF_CUDA(cudaStreamWaitEvent(s_compute, de_done, 0)); }
    IRF_CUDA(cudaEventRecord(k1, s_compute));
    if (verify_crc) {   // CRC-32 of what is not checked yet, on the side stream: it overlaps record framing and counting and is joined in count()
      IRF_CUDA(cudaEventRecord(ready, s_compute)); IRF_CUDA(cudaStreamWaitEvent(s_de, ready, 0));
      if (!listed) launch_crc_check(d_U.as<uint8_t>(), d_u_off.as<u64>(), d_u_len.as<u32>(), d_crc.as<u32>(), (int)nb, d_status.as<int>(), a.err, s_de);
      else { if (!stream_in) launch_crc_check(d_U.as<uint8_t>(), d_u_off.as<u64>(), d_u_len.as<u32>(), d_crc.as<u32>(), (int)de_list.size(), d_status.as<int>(), a.err, s_de, d_de_list.as<u32>());
             launch_crc_check(d_U.as<uint8_t>(), d_u_off.as<u64>(), d_u_len.as<u32>(), d_crc.as<u32>(), (int)sm_list.size(), d_status.as<int>(), a.err, s_de, d_sm_list.as<u32>()); }
      crc_pending = true;
    }
    if (stream_in) IRF_CUDA(cudaStreamSynchronize(s_copy));
    IRF_CUDA(cudaGetLastError()); sync();
    float ms = 0; IRF_CUDA(cudaEventElapsedTime(&ms, k0, k1)); timing[1] = ms; n_inflate_launches = 1; timing[10] = ms; timing[11] = sm_byte_share();
    if (getenv("IRF_DEBUG_TIMING")) fprintf(stderr, "[inflate] %.1f ms: %zu blocks on the decompression engine, %zu on the SM kernel%s\n", ms, de_list.size(), listed ? sm_list.size() : nb, stream_in ? " (streamed from the host)" : "");
    cudaEventDestroy(k0); cudaEventDestroy(k1); cudaEventDestroy(ready); cudaEventDestroy(de_done); for (auto e : chunk_ev) cudaEventDestroy(e);
    resident = true;
  }
  // match-token scratch of the two-phase inflate kernel (inflate_variant 12): TOK_CAP words per block in flight
  uint32_t* tok_scratch(int nblk_kernel) { size_t need = inflate_scratch_bytes(inflate_variant, nblk_kernel); if (!need) return nullptr; if (need + 256 > d_tok.cap) { sync(); cudaStreamSynchronize(s_copy); d_tok.ensure(need); } return d_tok.as<uint32_t>(); }
  unsigned long long* wm_host = nullptr; size_t wm_host_cap = 0;
  int n_inflate_launches = 0;

  int dev_err() { int e = 0; IRF_CUDA(cudaMemcpyAsync(&e, d_err.p, 4, cudaMemcpyDeviceToHost, s_compute)); sync(); return e; }
  void check_dev_err(const char* stage_name) { int e = dev_err(); if (e) throw err_text(e, stage_name); }
  static string err_text(int e, const char* stage_name) {
    string m = string(stage_name) + ":";
    if (e & DERR_INFLATE) m += " corrupt DEFLATE data in a BGZF block;"; if (e & DERR_TRUNCATED) m += " truncated BAM record stream;"; if (e & DERR_BAD_RECORD) m += " malformed BAM record / unframeable stream;";
    if (e & DERR_JUNC_FULL) m += " junction table full (raise junc_capacity_log2);"; if (e & DERR_JKEY_RANGE) m += " junction outside the supported key range;";
    if (e & DERR_PAIR_FULL) m += " pairing table full;"; if (e & DERR_REFID) m += " record refID beyond the header's reference list;"; if (e & DERR_DUMP_FULL) m += " fragment dump buffer full;"; if (e & DERR_STARVED) m += " inflate kernel starved of input (host copy stalled);"; if (e & DERR_CRC) m += " CRC-32 of an inflated BGZF block does not match its footer;";
    return m;
  }

  // BAM2blocks::readBamHeader on device-inflated bytes
  void parse_header_from_device() {
    size_t have = 0; vector<uint8_t> h;
    auto fetch = [&](size_t want) { want = std::min<size_t>(want, Uend); if (want <= have) return; h.resize(want); IRF_CUDA(cudaMemcpyAsync(h.data() + have, d_U.as<uint8_t>() + have, want - have, cudaMemcpyDeviceToHost, s_compute)); sync(); have = want; };
    fetch(1 << 16);
    auto need = [&](size_t n) { if (n > have) fetch(std::max(n, have * 2)); if (n > have) throw string("BAM header truncated"); };
    need(12); if (memcmp(h.data(), "BAM\1", 4) != 0) throw string("not a BAM stream (bad magic)");
    int32_t l_text; memcpy(&l_text, &h[4], 4); if (l_text < 0) throw string("bad BAM header");
    size_t p = 8 + (size_t)l_text; need(p + 4);
    { const char* t = (const char*)h.data() + 8; const size_t first_line = (size_t)(std::find(t, t + l_text, '\n') - t); const string hd(t, first_line);   // @HD	VN:1.6	SO:coordinate
      hdr_coordinate_sorted = hd.compare(0, 3, "@HD") == 0 && hd.find("SO:coordinate") != string::npos; } int32_t n_ref; memcpy(&n_ref, &h[p], 4); p += 4; if (n_ref < 0) throw string("bad BAM header");
    vector<string> names; vector<int32_t> lens;
    for (int i = 0; i < n_ref; i++) { need(p + 4); int32_t l_name; memcpy(&l_name, &h[p], 4); p += 4; if (l_name < 1) throw string("bad BAM header"); need(p + l_name + 4);
      names.emplace_back((const char*)&h[p], (size_t)l_name - 1); p += l_name; int32_t l_ref; memcpy(&l_ref, &h[p], 4); p += 4; lens.push_back(l_ref); }
    rec0 = p;
    if (!have_header || names != chr_names || lens != chr_lens || !bound) { chr_names = names; chr_lens = lens; have_header = true; chr_map_update(); }
  }

  // ---- file input: a ring of page-locked buffers instead of a page-locked copy of the whole file ----------------------
  // irf_run_bam maps the file (for the BGZF block scan, which touches two pages per block) and hands its descriptor to
  // the feeder: a few threads pread() 16 MB pieces straight into free ring slots and queue each slot's host->device copy
  // in file order; a chunk's event is recorded behind its last piece.  Page-locking 14 GB for one run cost ~4 s (and as
  // long again to release); the ring is 128 MB, allocated once per context.
  struct Piece { size_t lo, hi; int chunk; bool last; };
  static constexpr int FEED_SLOTS = 24; static constexpr size_t FEED_SLOT_BYTES = 16u << 20; int feed_threads_n = 16;   // round 1: 8 slots, 4 threads -> ~17 GB/s from the page cache
  int feed_fd = -1; int file_ring = 1;   // option "file_ring": 0 never, 1 files above 64 MB, 2 always
  u64 feed_base = 0;                     // file offset of the staged bytes' first byte (a shard of an in-process multi-device run)
  vector<Piece> feed_pieces; vector<cudaEvent_t> feed_ev_chunk; vector<char> feed_chunk_queued; cudaEvent_t feed_slot_ev[FEED_SLOTS] = {};
  uint8_t* feed_ring = nullptr; std::mutex feed_mu; std::condition_variable feed_cv; size_t feed_next = 0, feed_queued_upto = 0; bool feed_abort = false, feed_active = false; string feed_error; vector<std::thread> feed_threads;
  bool feed_capable() { return file_ring && stream_pipeline && de_percent_stream == 100 && inflate_variant != 8 && g_drv.deflate_engine(); }
  void feed_worker() {
    cudaSetDevice(device);
    for (;;) {
      size_t i; { std::lock_guard<std::mutex> g(feed_mu); if (feed_abort) return; i = feed_next++; }
      if (i >= feed_pieces.size()) return;
      const Piece pc = feed_pieces[i]; const int slot = (int)(i % FEED_SLOTS); uint8_t* buf = feed_ring + (size_t)slot * FEED_SLOT_BYTES;
      { std::unique_lock<std::mutex> lk(feed_mu); feed_cv.wait(lk, [&] { return feed_abort || feed_queued_upto + FEED_SLOTS > i; }); if (feed_abort) return; }   // piece i - FEED_SLOTS has been queued
      bool ok = true;
      if (i >= (size_t)FEED_SLOTS) ok = cudaEventSynchronize(feed_slot_ev[slot]) == cudaSuccess;                                                                  // ... and its copy has left the slot
      for (size_t off = pc.lo; ok && off < pc.hi;) { ssize_t k = pread(feed_fd, buf + (off - pc.lo), pc.hi - off, (off_t)(feed_base + off)); if (k <= 0) ok = false; else off += (size_t)k; }
      std::unique_lock<std::mutex> lk(feed_mu); feed_cv.wait(lk, [&] { return feed_abort || feed_queued_upto == i; }); if (feed_abort) return;
      if (ok) ok = cudaMemcpyAsync(d_comp.as<uint8_t>() + pc.lo, buf, pc.hi - pc.lo, cudaMemcpyHostToDevice, s_copy) == cudaSuccess && cudaEventRecord(feed_slot_ev[slot], s_copy) == cudaSuccess;
      if (ok && pc.last) ok = cudaEventRecord(feed_ev_chunk[pc.chunk], s_copy) == cudaSuccess;
      if (!ok) { feed_error = "read error (file input ring)"; feed_abort = true; std::fill(feed_chunk_queued.begin(), feed_chunk_queued.end(), 1); lk.unlock(); feed_cv.notify_all(); return; }
      if (pc.last) feed_chunk_queued[pc.chunk] = 1;
      feed_queued_upto = i + 1; lk.unlock(); feed_cv.notify_all();
    }
  }
  void feed_start(size_t total) {
    feed_stop();
    if (!feed_ring) { IRF_CUDA(cudaMallocHost((void**)&feed_ring, FEED_SLOTS * FEED_SLOT_BYTES)); for (int k = 0; k < FEED_SLOTS; k++) IRF_CUDA(cudaEventCreateWithFlags(&feed_slot_ev[k], cudaEventDisableTiming)); }
    const size_t nch = std::max<size_t>((total + chunk_bytes - 1) / chunk_bytes, 1), sub = std::min(FEED_SLOT_BYTES, chunk_bytes);
    feed_pieces.clear();
    for (size_t c = 0; c < nch; c++) { size_t lo = c * chunk_bytes, hi = std::min(total, lo + chunk_bytes); if (hi <= lo) { feed_pieces.push_back(Piece{lo, lo, (int)c, true}); continue; }
      for (size_t o = lo; o < hi; o += sub) feed_pieces.push_back(Piece{o, std::min(hi, o + sub), (int)c, o + sub >= hi}); }
    feed_ev_chunk.resize(nch); for (auto& e : feed_ev_chunk) IRF_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    feed_chunk_queued.assign(nch, 0); feed_next = 0; feed_queued_upto = 0; feed_abort = false; feed_error.clear(); feed_active = true;
    unsigned nt = std::min<unsigned>((unsigned)std::min(feed_threads_n, FEED_SLOTS - 2), std::max(1u, std::thread::hardware_concurrency()));
    for (unsigned t = 0; t < nt; t++) feed_threads.emplace_back([this]() { feed_worker(); });
  }
  cudaEvent_t feed_wait_chunk(size_t c) {   // returns once every copy of chunk c has been queued on the copy stream
    std::unique_lock<std::mutex> lk(feed_mu); feed_cv.wait(lk, [&] { return feed_abort || feed_chunk_queued[c]; });
    if (!feed_error.empty()) throw feed_error;
    return feed_ev_chunk[c];
  }
  void feed_stop() {
    if (!feed_active) return;
    { std::lock_guard<std::mutex> g(feed_mu); if (feed_queued_upto < feed_pieces.size()) feed_abort = true; } feed_cv.notify_all();
    for (auto& t : feed_threads) t.join(); feed_threads.clear();
    cudaStreamSynchronize(s_copy); for (auto e : feed_ev_chunk) cudaEventDestroy(e); feed_ev_chunk.clear(); feed_active = false;
  }

  // S1-S3 as one pipeline for bytes that are still on the host (the end-to-end path).  The file is cut into chunks; every
  // chunk is copied, inflated by the decompression engine and CRC-checked on the side streams (all queued up front, chained
  // by events), and the compute stream frames, pairs and counts chunk k as soon as chunk k+1 has been inflated (a record that
  // starts in chunk k may end in chunk k+1).  Mate pairing is order-free, so counting chunk by chunk gives the same tables;
  // the PCIe transfer (the bound of the end-to-end path) then hides framing and counting as well as the inflate.
  bool can_pipeline() { return !resident && !inflated_input && stream_pipeline && (staged_has_header || have_header) && nblk_own > 0 && !blks.empty() && de_percent_stream == 100 && inflate_variant != 8 && g_drv.deflate_engine(); }
  // The same pipeline for bytes that are already resident (repeated irf_count on staged bytes, the device-only figure): no
  // copies; ONE persistent inflate_kernel launch on the side stream takes the leading chunks of the file while the
  // decompression engine inflates the last de_percent_overlap % of the chunks one by one; the compute stream counts the
  // kernel's chunks when that launch is done and then the DE's chunks as they appear, so the DE never waits for the SMs.
  bool can_overlap() { return resident && !inflated_input && resident_pipeline && (staged_has_header || have_header) && nblk_own > 0 && !blks.empty() && de_percent > 0 && de_percent_overlap > 0
                              && inflate_variant != 8 && g_drv.deflate_engine() && (src_total + chunk_bytes - 1) / chunk_bytes >= 8
                              && (resident_pipeline == 2 || blks.size() * (size_t)(100 - de_percent_overlap) / 100 * 10 >= overlap_lanes() * 9); }   // pays once the kernel's share fills its grid
  void count_streaming(bool from_host = true) {
    cudaEvent_t e0, e3; IRF_CUDA(cudaEventCreate(&e0)); IRF_CUDA(cudaEventCreate(&e3)); IRF_CUDA(cudaEventRecord(e0, s_compute));
    if (!bound) IRF_CUDA(cudaMemsetAsync(d_err.p, 0, 4, s_compute));
    if (from_host) { if (!engines_planned || planned_percent != 100) plan_engines(100); }
    else if (!engines_planned || planned_percent != 1000 + de_percent_overlap) plan_engines(de_percent_overlap, 1);
    if (!from_host && !sm_list.empty()) tok_scratch((int)std::min(sm_list.size(), overlap_lanes()));
    const size_t nb = blks.size(), nch = chunk_blk_end.size(); const bool sm_used = !from_host && !sm_list.empty(); const size_t ksm = sm_used ? nch - k_de : 0;   // chunks [0, ksm) are the SM kernel's
    cudaEvent_t k0 = nullptr, k1 = nullptr;
    int* errp = T.err ? T.err : d_err.as<int>();
    IRF_CUDA(cudaMemsetAsync(d_U.as<uint8_t>() + Uend, 0, 64, s_compute));
    cudaEvent_t ready; IRF_CUDA(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming)); IRF_CUDA(cudaEventRecord(ready, s_compute));
    IRF_CUDA(cudaStreamWaitEvent(s_copy, ready, 0)); IRF_CUDA(cudaStreamWaitEvent(s_de, ready, 0));
    vector<cudaEvent_t> ev_copy(nch, nullptr), ev_inf(nch, nullptr); vector<u64> inf_end(nch);   // inflated offset where each chunk ends
    { u64 u = 0; size_t b = 0; for (size_t c = 0; c < nch; c++) { for (; b < chunk_blk_end[c]; b++) u += blks[b].isize; inf_end[c] = u; } }
    const size_t npre = std::min(pre_ev_copy.size(), nch);   // chunks whose copy stage() already queued (same chunking)
    const bool fed = from_host && feed_active;                 // file input: the feeder threads queue the copies (same chunking)
    if (fed && feed_ev_chunk.size() < nch) throw string("internal: file ring and pipeline disagree on the chunk count");
    // Copy (unless queued already), engine batch and checks of chunk c.  Everything is queued up front, except for file input: there a
    // chunk can only be submitted once the reader threads have queued its copies, so submission is interleaved with the counting loop
    // (submit_upto) - round 1 waited here for the whole file to be read before the first chunk was counted.
    size_t submitted = 0;
    auto submit_chunk = [&](size_t c) {
      size_t lo = c * chunk_bytes, hi = std::min(src_total, lo + chunk_bytes);
      if (!from_host) { IRF_CUDA(cudaEventCreateWithFlags(&ev_copy[c], cudaEventDisableTiming)); }
      else if (fed) { ev_copy[c] = feed_wait_chunk(c); }
      else if (c < npre) { ev_copy[c] = pre_ev_copy[c]; }
      else { if (hi > lo) IRF_CUDA(cudaMemcpyAsync(d_comp.as<uint8_t>() + lo, h_src + lo, hi - lo, cudaMemcpyHostToDevice, s_copy));
        IRF_CUDA(cudaEventCreateWithFlags(&ev_copy[c], cudaEventDisableTiming)); IRF_CUDA(cudaEventRecord(ev_copy[c], s_copy)); }
      if (from_host) IRF_CUDA(cudaStreamWaitEvent(s_de, ev_copy[c], 0));
      size_t f0 = c ? de_chunk_end[c - 1] : 0, f1 = de_chunk_end[c];
      for (size_t i = f0; i < f1; i += 32768) { size_t bad = 0; CUresult r = g_drv.decompress(de_params.data() + i, std::min(f1, i + 32768) - i, 0, &bad, (CUstream)s_de);
        if (r != CUDA_SUCCESS) throw string("cuMemBatchDecompressAsync failed with CUresult ") + std::to_string((int)r) + " (cuda)"; }
      IRF_CUDA(cudaEventCreate(&ev_inf[c])); IRF_CUDA(cudaEventRecord(ev_inf[c], s_de));
      // the length and CRC-32 checks run on their own stream: on the DE's stream they would hold the next batch back whenever
      // the SMs are busy; a mismatch raises the error flag, which is read after the last chunk
      if (f1 > f0) { IRF_CUDA(cudaStreamWaitEvent(s_chk, ev_inf[c], 0)); launch_de_check(d_de_list.as<u32>() + f0, (int)(f1 - f0), d_de_act.as<u32>(), d_u_len.as<u32>(), d_status.as<int>(), errp, s_chk);
        if (verify_crc) launch_crc_check(d_U.as<uint8_t>(), d_u_off.as<u64>(), d_u_len.as<u32>(), d_crc.as<u32>(), (int)(f1 - f0), d_status.as<int>(), errp, s_chk, d_de_list.as<u32>() + f0); }
    };
    auto chunk_ready = [&](size_t c) { if (!fed) return true; std::lock_guard<std::mutex> g(feed_mu); return feed_abort || feed_chunk_queued[c] != 0; };
    auto submit_upto = [&](size_t last, bool block) { for (; submitted < nch && submitted <= last; submitted++) { if (!block && !chunk_ready(submitted)) return; submit_chunk(submitted); } };
    if (!fed) submit_upto(nch - 1, true);
    for (size_t c = npre; c < pre_ev_copy.size(); c++) cudaEventDestroy(pre_ev_copy[c]);
    pre_ev_copy.clear(); pre_copied_bytes = 0;   // ownership of the first npre events moved to ev_copy
    // The SM kernel's share, beside everything else: one launch per grid-full of blocks (a launch whose blocks do not fill the grid
    // runs as long as a full one - each warp works through its 32 blocks at its own pace - so a share of 1.4 grids goes out as a full launch
    // and a 0.4 one; while the second runs, the SM slots it leaves empty go to the counting kernels of the first launch's chunks).
    vector<cudaEvent_t> k_ev; vector<size_t> k_chunk_end;   // event after launch g; chunks [k_chunk_end[g-1], k_chunk_end[g]) belong to it
    if (sm_used) {
      IRF_CUDA(cudaEventCreate(&k0)); IRF_CUDA(cudaEventRecord(k0, s_copy));
      const size_t lanes = overlap_lanes();
      for (size_t c0 = 0; c0 < ksm;) {
        size_t c1 = c0 + 1; const size_t s0 = c0 ? sm_chunk_end[c0 - 1] : 0;
        while (c1 < ksm && sm_chunk_end[c1] - s0 <= lanes) c1++;
        const size_t s1 = sm_chunk_end[c1 - 1];
        InflateArgs a{d_comp.as<uint8_t>(), d_c_off.as<u64>(), d_c_len.as<u32>(), d_U.as<uint8_t>(), d_u_off.as<u64>(), d_u_len.as<u32>(), (int)(s1 - s0), d_status.as<int>(),
                      errp, (unsigned int*)(d_small.as<uint8_t>() + 160), nullptr, d_sm_list.as<u32>() + s0};
        a.tok_scratch = tok_scratch((int)(s1 - s0)); a.tok_scratch_bytes = d_tok.cap;
        launch_inflate(a, inflate_variant, s_copy);
        cudaEvent_t ev; IRF_CUDA(cudaEventCreate(&ev)); IRF_CUDA(cudaEventRecord(ev, s_copy)); k_ev.push_back(ev); k_chunk_end.push_back(c1);
        if (verify_crc && s1 > s0) { IRF_CUDA(cudaStreamWaitEvent(s_chk, ev, 0)); launch_crc_check(d_U.as<uint8_t>(), d_u_off.as<u64>(), d_u_len.as<u32>(), d_crc.as<u32>(), (int)(s1 - s0), d_status.as<int>(), errp, s_chk, d_sm_list.as<u32>() + s0); }
        c0 = c1;
      }
      k1 = k_ev.back();
    }
    auto cleanup = [&]() { if (fed) feed_stop(); else for (auto e : ev_copy) if (e) cudaEventDestroy(e); for (auto e : ev_inf) if (e) cudaEventDestroy(e); cudaEventDestroy(ready); cudaEventDestroy(e0); cudaEventDestroy(e3); if (k0) cudaEventDestroy(k0); for (auto e : k_ev) cudaEventDestroy(e); };
    try {
      // BAM2blocks::readBamHeader as soon as the first chunk is there (a header larger than one chunk waits for more)
      if (sm_used && staged_has_header) IRF_CUDA(cudaEventSynchronize(k_ev[0]));   // the header lies in the kernel's first launch
      for (size_t c = 0; staged_has_header; c++) {
        submit_upto(c, true);
        IRF_CUDA(cudaEventSynchronize(ev_inf[c])); check_dev_err("inflate");
        u64 full = Uend; Uend = inf_end[c]; bool ok = true;
        try { parse_header_from_device(); } catch (const string& e) { ok = false; if (e.find("truncated") == string::npos || c + 1 == nch) { Uend = full; throw; } }
        Uend = full; if (ok) break;
      }
      if (!bound) chr_map_update();
      pslots = next_pow2(std::max<u64>(1024, Uend / 80 + 2 * n_carry)); d_ptab.ensure(pslots * 8); IRF_CUDA(cudaMemsetAsync(d_ptab.p, 0, pslots * 8, s_compute));
      park_carry();
      u32* n_bad = d_small.as<u32>(); u64* totals = (u64*)(d_small.as<uint8_t>() + 16); u64* list_counts = (u64*)(d_small.as<uint8_t>() + 192);
      // A shard of a larger file (multi-GPU) has no header and ends with look-ahead blocks that are inflated but not counted
      // here: its first record start is a guess unless the caller passed one ("entry0"), exactly as in count().
      { size_t mx = 0, p0 = 0; for (size_t c = 0; c < nch; c++) { size_t p1 = std::min<size_t>(chunk_blk_end[c], (size_t)nblk_own); if (p1 > p0) { mx = std::max(mx, p1 - p0); p0 = p1; } }
        if (mx * REC_SLOTS * 8 + 256 > d_slots.cap) { sync(); d_slots.ensure(mx * REC_SLOTS * 8); } }
      u64 entry_prev = rec0; int entry_valid = 1; n_rec = 0; n_cand = 0; frame_rounds = 0; size_t b0 = 0; u64 fe0 = rec0; bool first = true;
      if (!staged_has_header) { entry_valid = entry0_override >= 0; entry_prev = entry_valid ? (u64)entry0_override : 0; entry0_override = -1; }
      for (size_t c = 0; c < nch; c++) {
        size_t b1 = std::min<size_t>(chunk_blk_end[c], (size_t)nblk_own); if (b1 <= b0) continue;
        const int nbk = (int)(b1 - b0); const bool last_own = b1 == (size_t)nblk_own;
        // chunk c is counted once chunk c+1 is inflated; the last counted chunk waits for everything (the look-ahead)
        // (with the default 256 MB chunks that is the next chunk; tiny test chunks wait for 8 MB of look-ahead so that any
        // realistic record, however many small chunks it spans, is complete)
        size_t need = std::min(c + 1, nch - 1); while (need + 1 < nch && inf_end[need] - inf_end[c] < (8u << 20)) need++;
        if (last_own) need = nch - 1;
        submit_upto(need, true); submit_upto(need + 3, false);   // file input: what this chunk needs (waiting for the readers if necessary), and whatever else has been read by now
        IRF_CUDA(cudaStreamWaitEvent(s_compute, ev_inf[need], 0));
        if (sm_used && c < ksm) { size_t g = 0; while (k_chunk_end[g] <= std::min(need, ksm - 1)) g++; IRF_CUDA(cudaStreamWaitEvent(s_compute, k_ev[g], 0)); }   // the launch that inflates the last chunk this one needs
        const u64 visible = inf_end[need];
        FrameArgs fa{d_U.as<uint8_t>(), d_u_off.as<u64>() + b0, d_u_len.as<u32>() + b0, nbk, visible, entry_prev, entry_valid, T.n_ref, T.ref_len,
                     d_entry.as<u64>() + b0, d_exit.as<u64>() + b0, d_nrec.as<u32>() + b0, d_ncand.as<u32>() + b0, d_ferr.as<uint8_t>() + b0, T.err, (int*)(d_small.as<uint8_t>() + 48)};
        fa.slots = d_slots.as<u64>();   // record starts per block, written by the walk (chunk-local block index)
        launch_frame_guess(fa, s_compute); launch_walk_count(fa, nullptr, s_compute);
        // ONE host round trip per chunk in the usual case: the consistency check, the totals and the chunk's exit are queued together and
        // read back at once; only a chunk with a wrong first-record guess (n_bad != 0) goes through repair rounds, each with its own read-back.
        u64 ht[3] = {0, 0, 0}, ex = 0; u32 hb = 0;
        for (int rounds = 0;; rounds++) {
          IRF_CUDA(cudaMemsetAsync(n_bad, 0, 4, s_compute)); launch_frame_mark(fa, d_mism.as<uint8_t>() + b0, d_mask.as<uint8_t>() + b0, n_bad, s_compute);
          IRF_CUDA(cudaMemcpyAsync(&hb, n_bad, 4, cudaMemcpyDeviceToHost, s_compute));
          // the totals of a chain that may still be repaired are provisional: their error bits go to a word of their own and count only once the chain holds
          IRF_CUDA(cudaMemsetAsync(totals, 0, 24, s_compute)); { FrameArgs ft = fa; ft.err = (int*)(totals + 2); launch_frame_totals(ft, totals, s_compute); }
          IRF_CUDA(cudaMemcpyAsync(ht, totals, 24, cudaMemcpyDeviceToHost, s_compute)); IRF_CUDA(cudaMemcpyAsync(&ex, d_exit.as<u64>() + (b1 - 1), 8, cudaMemcpyDeviceToHost, s_compute));
          if (first) IRF_CUDA(cudaMemcpyAsync(&fe0, d_entry.as<u64>() + b0, 8, cudaMemcpyDeviceToHost, s_compute));
          sync();
          if (!hb) break; if (rounds > nbk + 2) throw string("record framing did not converge"); frame_rounds++;
          if (getenv("IRF_DEBUG_FRAMING") && (rounds < 3 || rounds % 100 == 0)) fprintf(stderr, "[framing] chunk %zu (%d blocks from %zu): round %d repairs %u entries\n", c, nbk, b0, rounds, hb);
          launch_walk_count(fa, d_mask.as<uint8_t>() + b0, s_compute);
        }
        if ((int)ht[2]) throw err_text((int)ht[2], "record framing");
        check_dev_err("record framing");
        entry_prev = ex; entry_valid = 1; first = false; n_rec += ht[0]; n_cand += ht[1];
        if (ht[0]) {
          // the lists keep their size from chunk to chunk (generous first allocation), so no allocation and no host wait separates the
          // chunks: the pairing kernel leaves its counts on the device and the two counting kernels are launched over upper bounds
          d_rec_base.ensure((size_t)nbk * 8); d_rec_off.ensure((ht[0] + ht[0] / 2) * 8); d_pair_list.ensure((ht[1] + ht[1] / 2 + 16) * 16); d_single_list.ensure((ht[0] + ht[0] / 2 + 16) * 8);
          launch_scan_counts(d_nrec.as<u32>() + b0, d_rec_base.as<u64>(), nbk, s_compute); launch_gather_records(fa, d_rec_base.as<u64>(), d_rec_off.as<u64>(), s_compute);
          IRF_CUDA(cudaMemsetAsync(list_counts, 0, 16, s_compute));
          launch_pair_records(T, d_U.as<uint8_t>(), d_rec_off.as<u64>(), ht[0], d_ptab.as<u64>(), (u32)(pslots - 1), d_pair_list.as<u64>(), d_single_list.as<u64>(), list_counts, s_compute, carry_base(), 1, adjacent_on());
          launch_process_pairs(T, d_U.as<uint8_t>(), d_pair_list.as<u64>(), ht[1] + 1, s_compute, list_counts);   // every pair is completed by one candidate record of this chunk (its mate may have waited since an earlier chunk)
          launch_process_singles(T, d_U.as<uint8_t>(), d_single_list.as<u64>(), ht[0] - ht[1] + 1, s_compute, list_counts + 1);
        }
        b0 = b1;
      }
      submit_upto(nch - 1, true);
      if (nblk_own == (int)nb && entry_prev != Uend) throw string("record framing: last record does not end at the end of the stream (truncated BAM)");
      IRF_CUDA(cudaEventRecord(e3, s_compute)); IRF_CUDA(cudaStreamSynchronize(s_copy)); IRF_CUDA(cudaStreamSynchronize(s_de)); IRF_CUDA(cudaStreamSynchronize(s_chk)); sync();
      check_dev_err("fragment counting");
      float tot = 0, inf = 0; IRF_CUDA(cudaEventElapsedTime(&tot, e0, e3)); IRF_CUDA(cudaEventElapsedTime(&inf, e0, ev_inf[nch - 1]));
      timing[10] = 0; timing[11] = 0;
      if (sm_used) { float t_sm = 0; IRF_CUDA(cudaEventElapsedTime(&t_sm, e0, k1)); inf = std::max(inf, t_sm); IRF_CUDA(cudaEventElapsedTime(&t_sm, k0, k1)); timing[10] = t_sm; timing[11] = sm_byte_share(); }
      if (getenv("IRF_DEBUG_TIMING")) { float t_de = 0, t_sm = 0; IRF_CUDA(cudaEventElapsedTime(&t_de, e0, ev_inf[nch - 1])); if (sm_used) IRF_CUDA(cudaEventElapsedTime(&t_sm, k0, k1));
        fprintf(stderr, "[pipeline] %s: %zu chunks (%zu DE), DE done at %.1f ms, SM inflate %.1f ms, all counted at %.1f ms\n", from_host ? "streamed" : "resident", nch, nch - ksm, t_de, t_sm, tot); }
      timing[0] = inf; timing[1] = inf; timing[2] = 0; timing[3] = tot - inf; timing[5] = tot; n_inflate_launches = sm_used ? (int)k_ev.size() : 0;
      { u64 own_end = 0; for (int i = 0; i < nblk_own; i++) own_end += blks[i].isize; frame_entry0 = fe0; frame_exit_rel = entry_prev >= own_end ? entry_prev - own_end : 0; }
    } catch (...) { cudaStreamSynchronize(s_copy); cudaStreamSynchronize(s_de); cudaStreamSynchronize(s_chk); cudaStreamSynchronize(s_compute); cleanup(); throw; }
    cleanup();
    resident = true; counted = true; finalized = false;
    collect_carry();
  }

  // S2 + S3
  void count() {
    if (!staged) throw string("nothing staged");
    if (feed_active && !can_pipeline()) { for (size_t c2 = 0; c2 < feed_chunk_queued.size(); c2++) feed_wait_chunk(c2); feed_stop(); resident = true; }   // everything is on the GPU: count stage by stage
    if (can_pipeline()) { count_streaming(true); return; }
    if (can_overlap()) { count_streaming(false); return; }
    if (!pre_ev_copy.empty()) { IRF_CUDA(cudaStreamSynchronize(s_copy)); if (pre_copied_bytes >= src_total) resident = true; drop_pre_copies(); }
    cudaEvent_t e0, e1, e2, e3; IRF_CUDA(cudaEventCreate(&e0)); IRF_CUDA(cudaEventCreate(&e1)); IRF_CUDA(cudaEventCreate(&e2)); IRF_CUDA(cudaEventCreate(&e3));
    IRF_CUDA(cudaEventRecord(e0, s_compute));
    if (!bound) IRF_CUDA(cudaMemsetAsync(d_err.p, 0, 4, s_compute));
    inflate_all();
    check_dev_err("inflate");
    IRF_CUDA(cudaEventRecord(e1, s_compute));
    if (staged_has_header) parse_header_from_device();
    else if (!have_header) throw string("no BAM header: call irf_read_header before staging a shard without one");
    if (!bound) chr_map_update();
    int nb = (int)blks.size();
    FrameArgs fa{d_U.as<uint8_t>(), d_u_off.as<u64>(), d_u_len.as<u32>(), nb, Uend, rec0, staged_has_header ? 1 : 0, T.n_ref, T.ref_len,
                 d_entry.as<u64>(), d_exit.as<u64>(), d_nrec.as<u32>(), d_ncand.as<u32>(), d_ferr.as<uint8_t>(), T.err, (int*)(d_small.as<uint8_t>() + 48)};
    if (!staged_has_header && entry0_override >= 0) { fa.rec0 = (u64)entry0_override; fa.rec0_valid = 1; entry0_override = -1; }
    launch_frame_guess(fa, s_compute); launch_walk_count(fa, nullptr, s_compute);
    u32* n_bad = d_small.as<u32>(); int rounds = 0;
    for (;;) {
      IRF_CUDA(cudaMemsetAsync(n_bad, 0, 4, s_compute)); launch_frame_mark(fa, d_mism.as<uint8_t>(), d_mask.as<uint8_t>(), n_bad, s_compute);
      u32 hb = 0; IRF_CUDA(cudaMemcpyAsync(&hb, n_bad, 4, cudaMemcpyDeviceToHost, s_compute)); sync();
      if (!hb) break; if (++rounds > nb + 2) throw string("record framing did not converge");
      launch_walk_count(fa, d_mask.as<uint8_t>(), s_compute);
    }
    frame_rounds = rounds;
    u64* totals = (u64*)(d_small.as<uint8_t>() + 16); IRF_CUDA(cudaMemsetAsync(totals, 0, 16, s_compute));
    FrameArgs own = fa; own.nblk = nblk_own; launch_frame_totals(own, totals, s_compute);
    u64 ht[2] = {0, 0}; IRF_CUDA(cudaMemcpyAsync(ht, totals, 16, cudaMemcpyDeviceToHost, s_compute));
    u64 fe[2] = {0, 0}; if (nblk_own > 0) { IRF_CUDA(cudaMemcpyAsync(&fe[0], d_entry.as<u64>(), 8, cudaMemcpyDeviceToHost, s_compute)); IRF_CUDA(cudaMemcpyAsync(&fe[1], d_exit.as<u64>() + (nblk_own - 1), 8, cudaMemcpyDeviceToHost, s_compute)); }
    sync(); n_rec = ht[0]; n_cand = ht[1];
    u64 own_end = 0; for (int i = 0; i < nblk_own; i++) own_end += blks[i].isize;
    frame_entry0 = fe[0]; frame_exit_rel = fe[1] >= own_end ? fe[1] - own_end : 0;
    check_dev_err("record framing");
    if (nblk_own == nb && nb > 0 && fe[1] != Uend) throw string("record framing: last record does not end at the end of the stream (truncated BAM)");
    IRF_CUDA(cudaEventRecord(e2, s_compute));
    pslots = next_pow2(std::max<u64>(1024, n_cand + n_cand / 2 + 2 * n_carry)); d_ptab.ensure(pslots * 8); IRF_CUDA(cudaMemsetAsync(d_ptab.p, 0, pslots * 8, s_compute));
    park_carry();
    d_rec_base.ensure((size_t)nblk_own * 8); d_rec_off.ensure(n_rec * 8);
    launch_scan_counts(d_nrec.as<u32>(), d_rec_base.as<u64>(), nblk_own, s_compute);
    launch_walk_list(own, d_rec_base.as<u64>(), d_rec_off.as<u64>(), s_compute);
    d_pair_list.ensure((n_cand / 2 + 1) * 16); d_single_list.ensure((n_rec - n_cand + 1) * 8);
    u64* list_counts = (u64*)(d_small.as<uint8_t>() + 192); IRF_CUDA(cudaMemsetAsync(list_counts, 0, 16, s_compute));
    launch_pair_records(T, d_U.as<uint8_t>(), d_rec_off.as<u64>(), n_rec, d_ptab.as<u64>(), (u32)(pslots - 1), d_pair_list.as<u64>(), d_single_list.as<u64>(), list_counts, s_compute, carry_base(), 1, adjacent_on());
    u64 lc[2] = {0, 0}; IRF_CUDA(cudaMemcpyAsync(lc, list_counts, 16, cudaMemcpyDeviceToHost, s_compute)); sync();
    launch_process_pairs(T, d_U.as<uint8_t>(), d_pair_list.as<u64>(), lc[0], s_compute);
    launch_process_singles(T, d_U.as<uint8_t>(), d_single_list.as<u64>(), lc[1], s_compute);
    IRF_CUDA(cudaEventRecord(e3, s_compute)); IRF_CUDA(cudaGetLastError());
    join_crc();
    check_dev_err("fragment counting");
    float a, b, c; IRF_CUDA(cudaEventElapsedTime(&a, e0, e1)); IRF_CUDA(cudaEventElapsedTime(&b, e1, e2)); IRF_CUDA(cudaEventElapsedTime(&c, e2, e3));
    timing[0] = a; timing[2] = b; timing[3] = c; timing[5] = a + b + c;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2); cudaEventDestroy(e3);
    counted = true; finalized = false;
    collect_carry();
  }
  int frame_rounds = 0;

  // ---------------------------------------------------------------- region-owned depth (multi-GPU)
  // Every measured segment gets one owning rank from the genomic position of each rank's first record (a coordinate-sorted
  // shard is a genomic region); segments that share a ref-cover record are kept together, so all blocks of an intron have one
  // owner.  Each rank lists its non-zero counters in foreign segments (export_foreign_depth), the owners add them
  // (import_depth), and irf_finalize computes the statistics of the owned introns only; the rows of the other introns stay
  // zero on the device, so the caller merges the row tables with an integer SUM and then calls finalize_rows().
  bool owner_mode = false, rows_pending = false; int owner_me = 0, owner_world = 1; u32 n_own_introns = 0;
  DevBuf d_seg_owner, d_tile_state, d_own_introns, d_fidx, d_fval, d_rows_dir; u64 foreign_cap = 1ull << 20; vector<int64_t> owner_key;
  void first_record(int32_t& refid, int32_t& pos) {
    refid = -2; pos = 0; if (!counted) throw string("irf_count has not run");
    if (n_rec == 0 || frame_entry0 + 12 > Uend) return;
    int32_t w[2]; IRF_CUDA(cudaMemcpyAsync(w, d_U.as<uint8_t>() + frame_entry0 + 4, 8, cudaMemcpyDeviceToHost, s_compute)); sync();
    refid = w[0]; pos = w[1];
  }
  void owner_plan(const int32_t* cut_refid, const int64_t* cut_pos, int world, int me) {
    if (!bound) throw string("irf_count has not run");
    if (world < 1 || world > 255 || me < 0 || me >= world) throw string("bad owner plan");
    const u32 nseg = (u32)cov.seg_start.size(); const int n = (int)chr_names.size();
    { vector<int64_t> key; key.push_back(world); key.push_back(me); for (int r = 0; r < world; r++) { key.push_back(cut_refid[r]); key.push_back(cut_pos[r]); }
      if (key == owner_key && d_seg_owner.p) { owner_mode = true; owner_me = me; owner_world = world; return; }   // same shards as last time (repeated runs on staged bytes): the plan stands
      owner_key = key; }
    std::map<string, int> id; for (int i = n - 1; i >= 0; i--) id[chr_names[i]] = i;
    // cut keys (refID, pos); a rank without records (refID < 0) owns nothing: its cut is the next rank's
    vector<std::pair<int64_t, int64_t>> cut(world); const std::pair<int64_t, int64_t> INF{1ll << 40, 0};
    for (int r = world - 1; r >= 0; r--) cut[r] = cut_refid[r] >= 0 ? std::make_pair((int64_t)cut_refid[r], cut_pos[r]) : (r + 1 < world ? cut[r + 1] : INF);
    cut[0] = {-1, 0}; for (int r = 1; r < world; r++) cut[r] = std::max(cut[r], cut[r - 1]);
    vector<std::pair<int64_t, int64_t>> key(nseg); vector<uint8_t> want(nseg, 0);
    for (auto& c : cov.chr_range) { auto it = id.find(c.first); const int64_t rid = it == id.end() ? -1 : it->second;
      for (u32 sgi = c.second.begin; sgi < c.second.end; sgi++) { key[sgi] = {rid, cov.seg_start[sgi]}; int r = 0; if (rid >= 0) { r = (int)(std::upper_bound(cut.begin(), cut.end(), key[sgi]) - cut.begin()) - 1; if (r < 0) r = 0; } want[sgi] = (uint8_t)r; } }
    // segments joined by a ref-cover record stay together (union-find); the group goes to the smallest wanted owner
    vector<u32> parent(nseg); for (u32 i = 0; i < nseg; i++) parent[i] = i;
    std::function<u32(u32)> find = [&](u32 x) { while (parent[x] != x) { parent[x] = parent[parent[x]]; x = parent[x]; } return x; };
    vector<u32> first_seg(cov.BEDrecords.size(), 0xffffffffu);
    for (size_t k = 0; k < cov.BEDrecords.size(); k++) { auto& rec = cov.BEDrecords[k]; auto c = cov.chr_range.find(rec.chrName); if (c == cov.chr_range.end()) continue;
      for (auto& b : rec.blocks) { if (b.first >= b.second) continue;
        u32 sgi = (u32)(std::upper_bound(cov.seg_end.begin() + c->second.begin, cov.seg_end.begin() + c->second.end, b.first) - cov.seg_end.begin()); if (sgi >= c->second.end) continue;
        if (first_seg[k] == 0xffffffffu) first_seg[k] = sgi; else { u32 a = find(first_seg[k]), bb = find(sgi); if (a != bb) parent[bb] = a; } } }
    vector<uint8_t> group(nseg, 255); for (u32 i = 0; i < nseg; i++) { u32 r = find(i); group[r] = std::min(group[r], want[i]); }
    vector<uint8_t> owner(nseg); for (u32 i = 0; i < nseg; i++) owner[i] = group[find(i)];
    vector<u32> mine; for (size_t k = 0; k < cov.BEDrecords.size(); k++) { int o = first_seg[k] == 0xffffffffu ? 0 : owner[first_seg[k]]; if (o == me) mine.push_back((u32)k); }
    // ownership per 1024-slot tile of the compact space, for the kernel that lists the foreign counters
    { const u64 ntiles = (cov.P + 1023) / 1024; vector<uint8_t> ts(ntiles, 3);   // 3 = nothing seen yet
      for (u32 i = 0; i < nseg; i++) { const u64 c0 = cov.seg_coff[i], c1 = c0 + (u64)(cov.seg_end[i] - cov.seg_start[i]); const uint8_t v = owner[i] == me ? 1 : 2;
        for (u64 t = c0 / 1024; t <= c1 / 1024; t++) ts[t] = ts[t] == 3 ? v : (ts[t] == v ? v : 0); }
      for (auto& x : ts) if (x == 3) x = 1;
      upload(d_tile_state, ts, s_compute); }
    upload(d_seg_owner, owner, s_compute); upload(d_own_introns, mine, s_compute); sync();
    n_own_introns = (u32)mine.size(); owner_mode = true; owner_me = me; owner_world = world;
  }
  u64 export_foreign_depth() {
    if (!owner_mode) throw string("irf_owner_plan must be called before the depth exchange");
    u64* cur = (u64*)(d_small.as<uint8_t>() + 104);
    for (;;) {
      d_fidx.ensure(foreign_cap * 8); d_fval.ensure(foreign_cap * 4); IRF_CUDA(cudaMemsetAsync(cur, 0, 8, s_compute));
      launch_foreign_depth(T, (u32)cov.seg_start.size(), d_seg_owner.as<uint8_t>(), (u32)owner_me, d_tile_state.as<uint8_t>(), d_fidx.as<u64>(), d_fval.as<int32_t>(), cur, foreign_cap, s_compute);
      u64 n = 0; IRF_CUDA(cudaMemcpyAsync(&n, cur, 8, cudaMemcpyDeviceToHost, s_compute)); sync();
      if (n <= foreign_cap) return n;
      foreign_cap = n + n / 4;   // rare: grow and list again
    }
  }
  void import_depth(const u64* idx, const int32_t* val, u64 n) {
    if (!owner_mode) throw string("irf_owner_plan must be called before the depth exchange");
    launch_import_depth(T, (u32)cov.seg_start.size(), d_seg_owner.as<uint8_t>(), (u32)owner_me, idx, val, n, s_compute); sync();
  }
  void finalize_rows() {
    if (!rows_pending) throw string("no owner-mode finalize to complete");
    const u32 n_in = (u32)cov.BEDrecords.size();
    pin_rows(rows_nondir, n_in); if (n_in) IRF_CUDA(cudaMemcpyAsync(rows_nondir.data(), d_rows.p, n_in * sizeof(IntronRow), cudaMemcpyDeviceToHost, s_compute));
    if (directionality != 0) { pin_rows(rows_dir, n_in); if (n_in) IRF_CUDA(cudaMemcpyAsync(rows_dir.data(), d_rows_dir.p, n_in * sizeof(IntronRow), cudaMemcpyDeviceToHost, s_compute)); }
    sync(); rows_pending = false; finalized = true;
  }

  // ---------------------------------------------------------------- segmented (streamed) runs
  // A stream of unknown length (stdin behind STAR or `gzip -cd`) is counted segment by segment: each segment is
  // staged and counted like a shard (own bytes + look-ahead, first record handed over from the previous segment's last one) on the
  // same context without a reset in between.  Mates that are still waiting when a segment ends are trimmed to fixed part + name +
  // CIGAR (the halo format) and carried into the next segment, where they are parked in the fresh pair table ahead of its records.
  bool seg_mode = false; u64 n_carry = 0, carry_bytes = 0; DevBuf d_carry, d_carry_off; size_t stream_segment_bytes = 1ull << 30;
  u64 carry_base() const { return seg_mode && n_carry ? ((Uend + 64 + 63) & ~63ull) : ~0ull; }
  void park_carry() {   // called after the pair table has been cleared, before the segment's own records are paired
    if (!seg_mode || !n_carry) return;
    const u64 cb = carry_base();
    IRF_CUDA(cudaMemcpyAsync(d_U.as<uint8_t>() + cb, d_carry.p, carry_bytes, cudaMemcpyDeviceToDevice, s_compute));
    IRF_CUDA(cudaMemsetAsync(d_U.as<uint8_t>() + cb + carry_bytes, 0, 64, s_compute));
    launch_shift_offsets(d_carry_off.as<u64>(), n_carry, cb, s_compute);
    u64* lc = (u64*)(d_small.as<uint8_t>() + 192); IRF_CUDA(cudaMemsetAsync(lc, 0, 16, s_compute));
    d_pair_list.ensure((n_carry + 16) * 16); d_single_list.ensure((n_carry + 16) * 8);
    launch_pair_records(T, d_U.as<uint8_t>(), d_carry_off.as<u64>(), n_carry, d_ptab.as<u64>(), (u32)(pslots - 1), d_pair_list.as<u64>(), d_single_list.as<u64>(), lc, s_compute, cb, 0);
  }
  void collect_carry() {   // after a segment's count: the records still waiting for a mate
    if (!seg_mode) return;
    u64 n = count_leftovers(nullptr, 0); d_left.ensure(n * 8 + 8); if (n) count_leftovers(d_left.as<u64>(), n);
    u64 cap = n * 320 + 1024; u64 used = 0;
    for (;;) {
      d_carry.ensure(cap + 64); d_carry_off.ensure(n * 8 + 8); u64* cur = (u64*)(d_small.as<uint8_t>() + 128); IRF_CUDA(cudaMemsetAsync(cur, 0, 8, s_compute));
      launch_halo_pack(d_U.as<uint8_t>(), d_left.as<u64>(), n, d_carry.as<uint8_t>(), d_carry_off.as<u64>(), cur, cap, s_compute);
      IRF_CUDA(cudaMemcpyAsync(&used, cur, 8, cudaMemcpyDeviceToHost, s_compute)); sync();
      if (used <= cap) break; cap = used + 1024;
    }
    n_carry = n; carry_bytes = used;
  }

  u64 count_leftovers(u64* fill, u64 cap) {
    u64* n_left = (u64*)(d_small.as<uint8_t>() + 64); IRF_CUDA(cudaMemsetAsync(n_left, 0, 8, s_compute));
    launch_pair_leftover(d_ptab.as<u64>(), pslots, n_left, fill, cap, s_compute);
    u64 h = 0; IRF_CUDA(cudaMemcpyAsync(&h, n_left, 8, cudaMemcpyDeviceToHost, s_compute)); sync(); return h;
  }

  u64 export_table(int which) {   // compact the counted entries of one junction table into dense device arrays
    u64* keys = which == 0 ? T.jc_keys : which == 1 ? T.jl_keys : T.jr_keys; u32* cnt = which == 0 ? T.jc_cnt : which == 1 ? T.jl_cnt : T.jr_cnt; u64 slots = which == 0 ? jc_slots : jside_slots;
    u64* cur = (u64*)(d_small.as<uint8_t>() + 96);
    IRF_CUDA(cudaMemsetAsync(cur, 0, 8, s_compute)); launch_ht_compact(keys, cnt, slots, nullptr, nullptr, cur, 0, 1, s_compute);
    u64 n = 0; IRF_CUDA(cudaMemcpyAsync(&n, cur, 8, cudaMemcpyDeviceToHost, s_compute)); sync();
    d_exp_keys[which].ensure(n * 8); d_exp_cnt[which].ensure(n * 8);
    IRF_CUDA(cudaMemsetAsync(cur, 0, 8, s_compute)); launch_ht_compact(keys, cnt, slots, d_exp_keys[which].as<u64>(), d_exp_cnt[which].as<u32>(), cur, n, 1, s_compute); sync();
    return n;
  }

  // result rows are read back into page-locked memory (a pageable read-back of 20 MB costs ~10 ms on this host)
  void pin_rows(vector<IntronRow>& v, size_t n) {
    if (v.size() == n) return;
    if (!v.empty()) cudaHostUnregister(v.data());
    v.clear(); v.resize(n);
    if (n && cudaHostRegister(v.data(), n * sizeof(IntronRow), cudaHostRegisterDefault) != cudaSuccess) cudaGetLastError();   // pageable still works, only slower
  }
  // S4 + output-side statistics
  void finalize() {
    if (!counted) throw string("irf_count has not run");
    if (finalized || rows_pending) return;
    cudaEvent_t e0, e1; IRF_CUDA(cudaEventCreate(&e0)); IRF_CUDA(cudaEventCreate(&e1)); IRF_CUDA(cudaEventRecord(e0, s_compute));
    const bool dbg = getenv("IRF_DEBUG_TIMING") != nullptr; double tdbg = now_ms(); auto mark = [&](const char* what) { if (dbg) { sync(); cudaStreamSynchronize(s_copy); double t = now_ms(); fprintf(stderr, "[finalize] %-22s %8.3f ms\n", what, t - tdbg); tdbg = t; } };
    // Two streams: the compute stream exports and sorts the junction tables while the host turns them into rows and runs
    // Directional(); the side stream takes the per-intron statistics (maximum depth, histogram kernel, row read-back), which
    // need the scanned depth and the junction hash tables but not the exported rows.
    cudaEvent_t ev_scan, ev_side; IRF_CUDA(cudaEventCreateWithFlags(&ev_scan, cudaEventDisableTiming)); IRF_CUDA(cudaEventCreateWithFlags(&ev_side, cudaEventDisableTiming));
    u32* dmax = d_small.as<u32>() + 10; u32 hm[2] = {0, 0};   // bytes 40..47 of the scratch words: the largest depth of each strand, found by the scan itself
    IRF_CUDA(cudaMemsetAsync(dmax, 0, 8, s_compute));
    launch_depth_scan(T.diff, cov.P * 2, d_tile_sums.as<int32_t>(), s_compute, cov.P, dmax); IRF_CUDA(cudaEventRecord(ev_scan, s_compute));
    IRF_CUDA(cudaStreamWaitEvent(s_copy, ev_scan, 0));
    IRF_CUDA(cudaMemcpyAsync(hm, dmax, 8, cudaMemcpyDeviceToHost, s_copy));
    mark("scan");
    IRF_CUDA(cudaStreamSynchronize(s_copy));
    const u32 hmax = hm[0] + hm[1];   // an upper bound of the largest combined depth: it only chooses the histogram layout
    // shared-memory histograms (the usual case): one short-lived CTA per intron, so the export kernels of the other stream
    // get SM slots as CTAs retire; deep data needs a global scratch histogram per CTA, hence a fixed number of looping CTAs
    u32 n_in = (u32)cov.BEDrecords.size(); u32 bins = hmax >= 8192 ? hmax + 1 : 1;
    // region-owned run: this rank computes the rows of its own introns only (the others stay zero and are merged by the caller)
    const u32 n_do = owner_mode ? n_own_introns : n_in; const u32* do_list = owner_mode ? d_own_introns.as<u32>() : nullptr;
    // one-pass fast kernel for every intron (2048-bin shared-memory histogram); the few introns with a deeper position land in an
    // overflow list and go through the general two-pass kernel (looping CTAs; global scratch histograms when the depth exceeds 8192)
    int grid = (int)std::min<u32>(std::max<u32>(n_do, 1), bins == 1 ? 148 * 8 : 148 * 4);
    d_scratch.ensure((u64)grid * bins * 4); d_overflow.ensure((u64)n_do * 4 + 4);
    u32* n_over = d_small.as<u32>() + 14;
    auto stats = [&](int dir_mode, IntronRow* out) {
      IRF_CUDA(cudaMemsetAsync(n_over, 0, 4, s_copy));
      launch_intron_stats_fast(T, T.diff, d_introns.as<IntronDesc>(), d_iblocks.as<IntronBlock>(), n_do, dir_mode, trim_center, out, s_copy, do_list, d_overflow.as<u32>(), n_over);
      launch_intron_stats(T, T.diff, d_introns.as<IntronDesc>(), d_iblocks.as<IntronBlock>(), n_do, dir_mode, trim_center, d_scratch.as<u32>(), bins, grid, out, s_copy, d_overflow.as<u32>(), n_over);
    };
    if (owner_mode) { d_rows_dir.ensure(n_in * sizeof(IntronRow)); IRF_CUDA(cudaMemsetAsync(d_rows.p, 0, n_in * sizeof(IntronRow), s_copy)); IRF_CUDA(cudaMemsetAsync(d_rows_dir.p, 0, n_in * sizeof(IntronRow), s_copy)); }
    stats(0, d_rows.as<IntronRow>());
    if (!owner_mode) { pin_rows(rows_nondir, n_in); if (n_in) IRF_CUDA(cudaMemcpyAsync(rows_nondir.data(), d_rows.p, n_in * sizeof(IntronRow), cudaMemcpyDeviceToHost, s_copy)); }
    // junctions -> host (JunctionCount map contents)
    u64 nj = export_table(0);
    static_assert(sizeof(JuncEntry) == 20, "JuncEntry is the 20-byte row the device writes and IRF_T_JUNCTIONS exposes");
    if (h_junc.size() != nj) { if (!h_junc.empty()) cudaHostUnregister(h_junc.data()); h_junc.clear(); h_junc.resize(nj); if (nj && cudaHostRegister(h_junc.data(), nj * sizeof(JuncEntry), cudaHostRegisterDefault) != cudaSuccess) cudaGetLastError(); }
    if (nj) {   // sort by key on the device (key order == (refID, start, end) order), turn the pairs into rows there, one read-back into page-locked memory
      size_t tb = junction_sort_temp_bytes(nj); d_sort_tmp.ensure(tb); d_sort_keys.ensure(nj * 8); d_sort_cnt.ensure(nj * 8); d_junc_rows.ensure(nj * sizeof(JuncEntry));
      launch_junction_sort(d_sort_tmp.p, tb, d_exp_keys[0].as<u64>(), d_sort_keys.as<u64>(), d_exp_cnt[0].as<u64>(), d_sort_cnt.as<u64>(), nj, s_compute);
      launch_junction_rows(d_sort_keys.as<u64>(), d_sort_cnt.as<u64>(), nj, T.chr_goff, T.n_ref, d_junc_rows.p, s_compute);
      IRF_CUDA(cudaMemcpyAsync(h_junc.data(), d_junc_rows.p, nj * sizeof(JuncEntry), cudaMemcpyDeviceToHost, s_compute)); sync();
    }
    mark("junction export+sort"); dir_log.clear(); directionality = directional(jref, chr_names, h_junc, dir_log); mark("directional");
    IRF_CUDA(cudaStreamSynchronize(s_copy)); mark("intron stats nondir (side stream)");
    if (directionality == 0 && !rows_dir.empty()) { cudaHostUnregister(rows_dir.data()); vector<IntronRow>().swap(rows_dir); }
    if (directionality != 0) {
      stats(directionality, owner_mode ? d_rows_dir.as<IntronRow>() : d_rows.as<IntronRow>());
      if (!owner_mode) { pin_rows(rows_dir, n_in); if (n_in) IRF_CUDA(cudaMemcpyAsync(rows_dir.data(), d_rows.p, n_in * sizeof(IntronRow), cudaMemcpyDeviceToHost, s_copy)); } IRF_CUDA(cudaStreamSynchronize(s_copy));
    }
    IRF_CUDA(cudaEventRecord(ev_side, s_copy)); IRF_CUDA(cudaStreamWaitEvent(s_compute, ev_side, 0)); cudaEventDestroy(ev_scan); cudaEventDestroy(ev_side);
    mark("intron stats dir"); h_stats.assign(ST_SLOTS, 0); h_sp.assign(sp.pos.size() * 2, 0); h_roi.assign(roi.rois.size() * 2, 0); h_chr.assign(chr_names.size() * 2, 0);
    IRF_CUDA(cudaMemcpyAsync(h_stats.data(), d_stats.p, ST_SLOTS * 8, cudaMemcpyDeviceToHost, s_compute));
    if (!h_sp.empty()) IRF_CUDA(cudaMemcpyAsync(h_sp.data(), d_sp_count.p, h_sp.size() * 4, cudaMemcpyDeviceToHost, s_compute));
    if (!h_roi.empty()) IRF_CUDA(cudaMemcpyAsync(h_roi.data(), d_roi_count.p, h_roi.size() * 4, cudaMemcpyDeviceToHost, s_compute));
    if (!h_chr.empty()) IRF_CUDA(cudaMemcpyAsync(h_chr.data(), d_chr_count.p, h_chr.size() * 4, cudaMemcpyDeviceToHost, s_compute));
    IRF_CUDA(cudaEventRecord(e1, s_compute)); sync();
    check_dev_err("finalize"); mark("tables d2h");
    // orphans (pair candidates whose mate never came) = candidates - 2 x pairs; a sharded run counted them during the halo exchange
    h_stats[ST_ERROR] += sharded ? halo_leftover : h_stats[ST_CAND] - 2 * h_stats[ST_PAIRED];
    timing[13] = (double)h_stats[ST_ADJ];
    h_stats.resize(ST_COUNT);
    float ms; IRF_CUDA(cudaEventElapsedTime(&ms, e0, e1)); timing[4] = ms; cudaEventDestroy(e0); cudaEventDestroy(e1);
    h_dump.clear(); h_depth.clear();
    if (dump_fragments) { u64 n = 0; IRF_CUDA(cudaMemcpy(&n, d_dump_cursor.p, 8, cudaMemcpyDeviceToHost)); n = std::min(n, dump_cap); h_dump.resize(n); if (n) IRF_CUDA(cudaMemcpy(h_dump.data(), d_dump.p, n * 4, cudaMemcpyDeviceToHost)); }
    // the reference's stdout block
    stdout_text = dir_log + "Directionality:\t" + std::to_string(directionality) + "\n";
    auto line = [&](const char* key, u64 v) { stdout_text += key; stdout_text += '\t'; stdout_text += std::to_string(v); stdout_text += '\n'; };
    line("Total reads processed", h_stats[ST_SINGLE] + h_stats[ST_PAIRED]); line("Total nucleotides", h_stats[ST_NUCLEOTIDES]); line("Total singles processed", h_stats[ST_SINGLE]);
    line("Total pairs processed", h_stats[ST_PAIRED]); line("Short pairs", h_stats[ST_SHORT]); line("Intersect pairs", h_stats[ST_INTERSECT]); line("Long pairs", h_stats[ST_LONG]);
    line("Skipped reads", h_stats[ST_SKIPPED]); line("Error reads", h_stats[ST_ERROR]);
    if (owner_mode) rows_pending = true; else finalized = true;   // region-owned run: the caller merges the row tables, then finalize_rows()
  }

  TextTable t_ir, t_dir, t_junc, t_sp;
  void write_outputs(const char* outdir) {
    if (!finalized) throw string("no finished run to write");
    double t0 = now_ms(); string d = outdir;
    // every table is cut into independent parts; ONE pool of threads formats the parts of all tables, then the files are written side by side
    const bool dbg = getenv("IRF_DEBUG_TIMING") != nullptr;
    vector<std::function<void()>> tasks;   // the tables themselves are members: their buffers are reused from run to run
    { vector<std::function<void()>> prep;   // cutting the tables into parts is itself worth a thread each (junction bucketing, span names)
      prep.push_back([&]() { prepare_ir_table(t_ir, cov, rows_nondir); if (directionality != 0) prepare_ir_table(t_dir, cov, rows_dir); });
      prep.push_back([&]() { prepare_junctions(t_junc, jref, chr_names, h_junc); }); prep.push_back([&]() { prepare_spans(t_sp, sp, h_sp); });
      run_tasks(prep, 3); }
    if (directionality == 0) { t_dir.parts.clear(); t_dir.tasks.clear(); }
    for (TextTable* t : {&t_ir, &t_dir, &t_junc, &t_sp}) for (auto& f : t->tasks) tasks.push_back(f);
    string s_chr, s_roi; tasks.push_back([&]() { s_chr = format_chr(chr_names, h_chr); if (roi.present) s_roi = format_roi(roi, h_roi); });
    double t1 = now_ms(); run_tasks(tasks, format_threads()); double t2 = now_ms();
    // one writer per file (parallel pwrite of one file's parts was measured no faster: page-cache allocation is the cost)
    auto putv = [&](const char* name, const vector<string>& parts) { string p = d + "/" + name; FILE* f = fopen(p.c_str(), "wb"); if (!f) throw string("cannot write ") + p; bool ok = true;
      for (auto& x : parts) ok &= fwrite(x.data(), 1, x.size(), f) == x.size(); if (fclose(f) || !ok) throw string("short write to ") + p; };
    vector<std::function<void()>> writers;
    writers.push_back([&]() { putv("IRFinder-IR-nondir.txt", t_ir.parts); });
    if (directionality != 0) writers.push_back([&]() { putv("IRFinder-IR-dir.txt", t_dir.parts); });
    writers.push_back([&]() { putv("IRFinder-JuncCount.txt", t_junc.parts); });
    writers.push_back([&]() { putv("IRFinder-SpansPoint.txt", t_sp.parts); });
    writers.push_back([&]() { putv("IRFinder-ChrCoverage.txt", vector<string>{s_chr}); if (roi.present) putv("IRFinder-ROI.txt", vector<string>{s_roi}); });
    run_tasks(writers, (unsigned)writers.size());
    if (dbg) fprintf(stderr, "[output] prepare %.2f ms, format %zu parts %.2f ms, write %.2f ms\n", t1 - t0, tasks.size(), t2 - t1, now_ms() - t2);
    timing[7] = now_ms() - t0;
  }

  void run_host_buffer(const void* p, size_t n) {
    const bool dbg = getenv("IRF_DEBUG_TIMING") != nullptr; double t0 = now_ms(), t1, t2, t3;
    stage(p, n, n, 1, false); t1 = now_ms();
    if (bound) reset(); t2 = now_ms();
    count(); t3 = now_ms(); finalize();
    if (dbg) fprintf(stderr, "[run] stage %.1f ms, reset %.1f ms, count %.1f ms, finalize %.1f ms\n", t1 - t0, t2 - t1, t3 - t2, now_ms() - t3);
  }
  // A CUDA failure while blocks were with the hardware decompression engine is, in practice, malformed DEFLATE data: the engine
  // reports it as a device-side fault that leaves this process's CUDA context unusable (every later call fails too).
  string cuda_hint() const {
    if (de_list.empty()) return "";
    return " [blocks were being inflated by the hardware decompression engine, which reports malformed DEFLATE data as a CUDA fault: the input is probably corrupt;"
           " this process's CUDA context is lost; rerun with option safe_inflate=1 (IRF_SAFE_INFLATE=1 for the irfinder binary) to have the SM kernel name the damaged block]";
  }
  // irf_run_stream: a pipe of unknown length, counted segment by segment with two page-locked buffers (a reader thread fills one
  // while the GPU counts the other).  A segment = own bytes (ending on a BGZF block boundary, or a 64 KiB boundary of an inflated
  // stream) + look-ahead bytes, which are also the first bytes of the next segment.
  void run_stream(int fd, bool inflated) {
    const size_t SEG = std::max<size_t>(stream_segment_bytes, 1u << 20), LA = std::min<size_t>(SEG, 8u << 20), SLOT = SEG + LA + (1u << 17);
#ifdef F_SETPIPE_SZ
    fcntl(fd, F_SETPIPE_SZ, 1 << 20);   // a pipe: the largest buffer an unprivileged process may ask for (best effort; fewer wake-ups per GB)
#endif
    struct Slot { uint8_t* p = nullptr; size_t own = 0, total = 0, filled = 0; bool last = false, ready = false; string err; } slot[2];
    for (auto& sl : slot) IRF_CUDA(cudaMallocHost((void**)&sl.p, SLOT + 64));
    std::mutex mu; std::condition_variable cv; bool stop = false;
    auto boundary = [&](const uint8_t* p, size_t filled, size_t limit, bool eof) -> size_t {   // largest segment end <= limit
      if (inflated) return eof && filled <= limit ? filled : (std::min(limit, filled) / 65536) * 65536;
      size_t off = 0; Blk b; size_t tot = 0;
      while (off + 18 <= filled && parse_block(p, off, filled, b, tot, nullptr) && off + tot <= limit) off += tot;
      return off;
    };
    std::thread reader([&]() {
      size_t keep = 0; int cur = 0; bool eof = false;
      while (!eof) {
        Slot& sl = slot[cur];
        { std::unique_lock<std::mutex> lk(mu); cv.wait(lk, [&] { return stop || !sl.ready; }); if (stop) return; }
        size_t filled = keep;   // the tail of the previous slot (its look-ahead and what followed) is already at the front
        while (filled < SEG + LA) { ssize_t k = read(fd, sl.p + filled, std::min<size_t>(SEG + LA - filled, 1u << 26)); if (k < 0) { sl.err = "read error"; eof = true; break; } if (k == 0) { eof = true; break; } filled += (size_t)k; }
        sl.filled = filled; sl.last = eof;
        if (eof) { sl.total = boundary(sl.p, filled, filled, true); sl.own = sl.total; if (sl.err.empty() && sl.total != filled) sl.err = inflated ? "internal" : "truncated BGZF block at the end of the stream"; }
        else { sl.own = boundary(sl.p, filled, SEG, false); sl.total = boundary(sl.p, filled, filled, false);
          if (!sl.own) { sl.err = "no BGZF block boundary inside a segment (not a BGZF stream?)"; eof = true; }
          else { Slot& nx = slot[cur ^ 1]; { std::unique_lock<std::mutex> lk(mu); cv.wait(lk, [&] { return stop || !nx.ready; }); if (stop) return; }
            keep = filled - sl.own; memcpy(nx.p, sl.p + sl.own, keep); } }
        { std::lock_guard<std::mutex> g(mu); sl.ready = true; } cv.notify_all();
        cur ^= 1;
      }
    });
    auto finish = [&]() { { std::lock_guard<std::mutex> g(mu); stop = true; } cv.notify_all(); reader.join(); for (auto& sl : slot) cudaFreeHost(sl.p); seg_mode = false; };
    try {
      if (bound) reset();
      seg_mode = true; n_carry = 0; carry_bytes = 0; entry0_override = -1;
      bool first = true; u64 n_rec_all = 0; double t_count = 0;
      for (int cur = 0;; cur ^= 1) {
        Slot& sl = slot[cur];
        { std::unique_lock<std::mutex> lk(mu); cv.wait(lk, [&] { return sl.ready; }); }
        if (!sl.err.empty()) throw sl.err;
        if (first && sl.total == 0) throw string(inflated ? "empty input stream" : "not a BGZF stream (empty input)");
        if (sl.total) {
          if (inflated) stage_inflated(sl.p, sl.own, sl.total, first);
          else { if (sl.last && require_eof) { Blk b; size_t tot = 0, off = 0, lastsz = 1; while (off < sl.total && parse_block(sl.p, off, sl.total, b, tot, nullptr)) { lastsz = b.isize; off += tot; } if (lastsz != 0) throw string("BGZF EOF marker missing (truncated stream?)"); }
            bool req = require_eof; require_eof = false; try { stage(sl.p, sl.own, sl.total, first ? 1 : 0, false); } catch (...) { require_eof = req; throw; } require_eof = req; }
          if (first && !bound && false) {}
          count(); n_rec_all += n_rec; t_count += timing[5];
          entry0_override = (long long)frame_exit_rel;   // where the next segment's first record starts, relative to its first inflated byte
        }
        const bool last = sl.last; first = false;
        { std::lock_guard<std::mutex> g(mu); sl.ready = false; } cv.notify_all();
        if (last) break;
      }
      if (n_carry) { /* mates that never came: counted as error reads through the candidate count in finalize */ }
      n_rec = n_rec_all; timing[5] = t_count;
      finish();
      finalize();
    } catch (...) { finish(); throw; }
  }
  void* pinned_buf(size_t n) { if (n > pinned_cap) { if (pinned) cudaFreeHost(pinned); pinned = nullptr; IRF_CUDA(cudaMallocHost(&pinned, n + 64)); pinned_cap = n; } return pinned; }
  size_t read_fd_all(int fd, size_t hint) {
    size_t cap = std::max<size_t>(hint, 1 << 20), n = 0; vector<uint8_t> tmp; uint8_t* dst = (uint8_t*)pinned_buf(cap);
    for (;;) {
      if (n == cap) { tmp.assign(dst, dst + n); cap *= 2; dst = (uint8_t*)pinned_buf(cap); memcpy(dst, tmp.data(), n); }
      ssize_t k = read(fd, dst + n, std::min<size_t>(cap - n, 1u << 30)); if (k < 0) throw string("read error"); if (k == 0) break; n += (size_t)k;
    }
    return n;
  }
};

// ------------------------------------------------------------------------------------------------ C ABI
#define IRF_TRY(ctx, ...) do { if (!(ctx)) return IRF_EINVAL; try { cudaSetDevice((ctx)->device); __VA_ARGS__; return IRF_OK; } catch (const string& e_) { (ctx)->err = e_; int rc_ = classify(e_); if (rc_ == IRF_ECUDA) (ctx)->err += (ctx)->cuda_hint(); return rc_; } catch (const std::exception& e_) { (ctx)->err = e_.what(); return IRF_EINVAL; } } while (0)
static int classify(const string& e) {
  if (e.find("cuda") != string::npos || e.find("CUDA") != string::npos) return IRF_ECUDA;
  if (e.find("cannot open") != string::npos || e.find("cannot write") != string::npos || e.find("read error") != string::npos || e.find("short write") != string::npos) return IRF_EIO;
  if (e.find("full") != string::npos) return IRF_ECAPACITY;
  if (e.find("must be called") != string::npos || e.find("has not run") != string::npos || e.find("nothing staged") != string::npos || e.find("no finished run") != string::npos || e.find("no BAM header") != string::npos) return IRF_ESTATE;
  return IRF_EFORMAT;
}

extern "C" {

int irf_create(irf_ctx** out, const int* device_ids, int n_devices) {
  if (!out || n_devices < 1 || n_devices > 64 || !device_ids) return IRF_EINVAL;
  *out = nullptr; vector<irf_ctx*> made;
  for (int k = 0; k < n_devices; k++) {   // one context per entry; an ordinal may repeat (several shards on one GPU)
    irf_ctx* c = new irf_ctx();
    try { c->init(device_ids[k]); } catch (const string& e) { fprintf(stderr, "irf_create: device %d: %s\n", device_ids[k], e.c_str()); delete c; for (auto m : made) irf_destroy(m); return IRF_ECUDA; }
    made.push_back(c);
  }
  made[0]->peers.assign(made.begin() + 1, made.end());
  *out = made[0]; return IRF_OK;
}
void irf_destroy(irf_ctx* c) {
  if (!c) return; for (auto p : c->peers) irf_destroy(p); c->peers.clear();
  cudaSetDevice(c->device); cudaDeviceSynchronize();
  DevBuf* all[] = {&c->d_ref_len, &c->d_chr_goff, &c->d_sp_pos, &c->d_sp_range, &c->d_sp_count, &c->d_seg_start, &c->d_seg_end, &c->d_seg_coff, &c->d_seg_range, &c->d_diff, &c->d_roi_start, &c->d_roi_end,
                   &c->d_roi_range, &c->d_roi_count, &c->d_chr_count, &c->d_jc_keys, &c->d_jc_cnt, &c->d_jl_keys, &c->d_jl_cnt, &c->d_jr_keys, &c->d_jr_cnt, &c->d_annot_keys, &c->d_bucket_base, &c->d_bucket_n, &c->d_sp_bidx, &c->d_seg_bidx, &c->d_roi_bidx, &c->d_stats, &c->d_err, &c->d_introns,
                   &c->d_iblocks, &c->d_rows, &c->d_scratch, &c->d_tile_sums, &c->d_small, &c->d_dump, &c->d_dump_cursor, &c->d_tok, &c->d_slots, &c->d_sm_list, &c->d_de_list, &c->d_de_act, &c->d_crc, &c->d_comp, &c->d_U, &c->d_c_off, &c->d_c_len, &c->d_u_off, &c->d_u_len, &c->d_status,
                   &c->d_entry, &c->d_exit, &c->d_nrec, &c->d_ncand, &c->d_ferr, &c->d_mask, &c->d_mism, &c->d_ptab, &c->d_left, &c->d_halo, &c->d_halo_off, &c->d_hptab, &c->d_rec_base, &c->d_rec_off, &c->d_pair_list, &c->d_single_list, &c->d_exp_keys[0], &c->d_exp_keys[1], &c->d_exp_keys[2],
                   &c->d_exp_cnt[0], &c->d_exp_cnt[1], &c->d_exp_cnt[2], &c->d_seg_owner, &c->d_tile_state, &c->d_own_introns, &c->d_fidx, &c->d_fval, &c->d_rows_dir, &c->d_carry, &c->d_carry_off, &c->d_sort_tmp, &c->d_sort_keys, &c->d_sort_cnt, &c->d_junc_rows, &c->d_overflow};
  for (auto b : all) b->release();
  if (!c->h_junc.empty()) cudaHostUnregister(c->h_junc.data());
  if (!c->rows_nondir.empty()) cudaHostUnregister(c->rows_nondir.data()); if (!c->rows_dir.empty()) cudaHostUnregister(c->rows_dir.data());
  if (c->pinned) cudaFreeHost(c->pinned);
  if (c->wm_host) cudaFreeHost(c->wm_host);
  c->h_tab.release(); c->h_lists.release();
  c->feed_stop(); if (c->feed_ring) { cudaFreeHost(c->feed_ring); for (int k = 0; k < irf_ctx::FEED_SLOTS; k++) cudaEventDestroy(c->feed_slot_ev[k]); }
  if (c->s_compute) cudaStreamDestroy(c->s_compute); if (c->s_copy) cudaStreamDestroy(c->s_copy); if (c->s_de) cudaStreamDestroy(c->s_de); if (c->s_chk) cudaStreamDestroy(c->s_chk);
  delete c;
}
const char* irf_last_error(irf_ctx* c) { return c ? c->err.c_str() : "null context"; }

int irf_set_option(irf_ctx* c, const char* key, const char* value) {
  if (!c || !key || !value) return IRF_EINVAL; string k = key; long long v = atoll(value);
  if (k == "trim_center") { if (v < 1 || v > 100) return IRF_EINVAL; c->trim_center = (u32)v; }
  else if (k == "dump_fragments") { c->dump_fragments = v != 0; c->bound = false; }
  else if (k == "inflate_variant") c->inflate_variant = (int)v;
  else if (k == "junc_capacity_log2") { if (v < 10 || v > 30) return IRF_EINVAL; c->junc_cap_log2 = (int)v; }
  else if (k == "require_eof") c->require_eof = v != 0;
  else if (k == "entry0") c->entry0_override = v;
  else if (k == "stage_lazy") c->stage_lazy = v != 0;
  else if (k == "verify_crc") c->verify_crc = v != 0;
  else if (k == "de_percent") { if (v < 0 || v > 100) return IRF_EINVAL; c->de_percent = (int)v; c->de_percent_stream = v == 0 ? 0 : c->de_percent_stream; c->engines_planned = false; }
  else if (k == "safe_inflate") { if (v) { c->de_percent = 0; c->de_percent_stream = 0; } else { c->de_percent = 70; c->de_percent_stream = 100; } c->engines_planned = false; }   // untrusted input: every block through the SM kernel, which answers malformed DEFLATE with IRF_EFORMAT and stays usable
  else if (k.rfind("rule_", 0) == 0) {   // SURVEY.md Appendix C rules (see Rules in common.cuh); floating thresholds are given in millionths
    Rules& R = c->rules; const string r = k.substr(5);
    if (r == "junc_min_flank") R.junc_min_flank = (int)v; else if (r == "span_left") R.span_left = (u32)v; else if (r == "span_right") R.span_right = (u32)v;
    else if (r == "skip_flags") R.skip_flags = (u32)v; else if (r == "pct_tie_midpoint") R.pct_tie_midpoint = v != 0;
    else if (r == "edge_offset") R.edge_offset = (u32)v; else if (r == "edge_len") R.edge_len = (u32)v;
    else if (r == "warn_lowcover_e6") R.warn_lowcover = v / 1e6; else if (r == "warn_lowsplice_e6") R.warn_lowsplice = v / 1e6; else if (r == "warn_minor_ratio_e6") R.warn_minor_ratio = v / 1e6;
    else if (r == "warn_nonuniform_abs_e6") R.warn_nonuniform_abs = v / 1e6; else if (r == "warn_nonuniform_ratio_e6") R.warn_nonuniform_ratio = v / 1e6;
    else { c->err = "unknown option " + k; return IRF_EINVAL; }
    c->bound = false;   // the device tables carry the rules (and the edge windows are cut on the host): bind again
  }
  else if (k == "adjacent_pairs") { if (v < -1 || v > 1) return IRF_EINVAL; c->adjacent_pairs = (int)v; }
  else if (k == "depth_tile") { if (v < 0 || v > 2) return IRF_EINVAL; c->T.depth_tile = (int)v; }
  else if (k == "stream_segment_mb") { if (v < 1 || v > 65536) return IRF_EINVAL; c->stream_segment_bytes = (size_t)v << 20; }
  else if (k == "stream_chunk_kb") { if (v < 64) return IRF_EINVAL; c->chunk_bytes = (size_t)v << 10; c->engines_planned = false; }
  else if (k == "inflate_ctas") { if (v < 0 || v > 32) return IRF_EINVAL; set_inflate_cta_cap((int)v); }   // process-wide, for A/B measurements
  else if (k == "file_ring_threads") { if (v < 1 || v > 64) return IRF_EINVAL; c->feed_threads_n = (int)v; }
  else if (k == "file_ring") { if (v < 0 || v > 2) return IRF_EINVAL; c->file_ring = (int)v; }
  else if (k == "stream_pipeline") c->stream_pipeline = v != 0;
  else if (k == "resident_pipeline") { if (v < 0 || v > 2) return IRF_EINVAL; c->resident_pipeline = (int)v; }
  else if (k == "de_percent_overlap") { if (v < 1 || v > 100) return IRF_EINVAL; c->de_percent_overlap = (int)v; }
  else if (k == "de_percent_stream") { if (v < 0 || v > 100) return IRF_EINVAL; c->de_percent_stream = (int)v; c->engines_planned = false; }
  else { c->err = "unknown option " + k; return IRF_EINVAL; }
  for (auto p : c->peers) { int rc = irf_set_option(p, key, value); if (rc != IRF_OK) { c->err = p->err; return rc; } }   // every device of a multi-device context runs with the same options
  return IRF_OK;
}
int irf_load_ref(irf_ctx* c, const char* cover, const char* sj, const char* rc, const char* roi) {
  if (!cover || !sj || !rc) return IRF_EINVAL;
  if (c && !c->peers.empty()) {   // the references are loaded on every device, side by side
    vector<irf_ctx*> all = c->all_ranks(); vector<int> rcs(all.size(), IRF_OK); vector<std::thread> th;
    for (size_t k = 1; k < all.size(); k++) th.emplace_back([&, k]() { rcs[k] = irf_load_ref(all[k], cover, sj, rc, roi); });
    irf_ctx* self = c; auto one = [&]() -> int { IRF_TRY(self, self->load_ref(cover, sj, rc, roi)); }; rcs[0] = one();
    for (auto& t : th) t.join();
    for (size_t k = 0; k < all.size(); k++) if (rcs[k] != IRF_OK) { if (k) c->err = all[k]->err; return rcs[k]; }
    return IRF_OK;
  }
  IRF_TRY(c, c->load_ref(cover, sj, rc, roi));
}

int irf_run_buffer(irf_ctx* c, const void* bgzf, size_t n) {
  if (!bgzf) return IRF_EINVAL;
  if (c && !c->peers.empty()) IRF_TRY(c, irf::multi_run(c->all_ranks(), (const uint8_t*)bgzf, n, -1));
  IRF_TRY(c, c->run_host_buffer(bgzf, n));
}
int irf_run_bam(irf_ctx* c, const char* path, const char* /*bai_path*/) {
  if (!path) return IRF_EINVAL;
  IRF_TRY(c, {
    int fd = open(path, O_RDONLY); if (fd < 0) throw string("cannot open ") + path;
    struct stat st; size_t hint = 0; if (fstat(fd, &st) == 0 && st.st_size > 0) hint = (size_t)st.st_size + 1;
    const size_t size = hint ? hint - 1 : 0;
    if (!c->peers.empty()) {   // several devices: the file is mapped, cut into one byte range per device, and every device reads its own range
      void* map = size && S_ISREG(st.st_mode) ? mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0) : MAP_FAILED;
      if (map != MAP_FAILED) {
        madvise(map, size, MADV_RANDOM);
        try { irf::multi_run(c->all_ranks(), (const uint8_t*)map, size, fd); } catch (...) { munmap(map, size); close(fd); throw; }
        munmap(map, size); close(fd); return IRF_OK;
      }
      size_t n = 0; try { n = c->read_fd_all(fd, hint); } catch (...) { close(fd); throw; } close(fd);
      irf::multi_run(c->all_ranks(), (const uint8_t*)c->pinned, n, -1); return IRF_OK;
    }
    if (size && S_ISREG(st.st_mode) && c->feed_capable() && (c->file_ring == 2 || size > (64u << 20))) {
      // regular file: mapped for the BGZF block scan, its bytes go to the GPU through the ring of page-locked buffers
      void* map = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
      if (map != MAP_FAILED) {   // (a file system that cannot map files takes the page-locked copy below)
        madvise(map, size, MADV_RANDOM);   // the scan reads two small pieces of every block; the bulk goes through pread
        c->feed_fd = fd;
        try { c->run_host_buffer(map, size); } catch (...) { c->feed_stop(); c->feed_fd = -1; c->h_src = nullptr; munmap(map, size); close(fd); throw; }
        c->feed_stop(); c->feed_fd = -1; c->h_src = nullptr; munmap(map, size); close(fd);
        return IRF_OK;
      }
    }
    size_t n = 0;
    try {
      if (hint > (64u << 20)) {   // regular file of known size: several threads pread disjoint slices into the pinned buffer
        uint8_t* dst = (uint8_t*)c->pinned_buf(size); unsigned nt = std::min(8u, std::max(1u, std::thread::hardware_concurrency())); vector<std::thread> th; vector<int> bad(nt, 0);
        for (unsigned t = 0; t < nt; t++) th.emplace_back([&, t]() { size_t lo = size * t / nt, hi = size * (t + 1) / nt; while (lo < hi) { ssize_t k = pread(fd, dst + lo, std::min<size_t>(hi - lo, 64u << 20), (off_t)lo); if (k <= 0) { bad[t] = 1; break; } lo += (size_t)k; } });
        for (auto& x : th) x.join(); for (int b : bad) if (b) throw string("read error");
        n = size;
      } else n = c->read_fd_all(fd, hint);
    } catch (...) { close(fd); throw; } close(fd);
    c->run_host_buffer(c->pinned, n);
  });
}
int irf_run_stream(irf_ctx* c, int fd, int inflated) { IRF_TRY(c, c->run_stream(fd, inflated != 0)); }
int irf_write_outputs(irf_ctx* c, const char* outdir) { if (!outdir) return IRF_EINVAL; IRF_TRY(c, c->write_outputs(outdir)); }
const char* irf_stdout_text(irf_ctx* c) { return c ? c->stdout_text.c_str() : ""; }
int irf_directionality(irf_ctx* c) { return c ? c->directionality : 0; }

int irf_get_table(irf_ctx* c, int which, const void** p, size_t* n) {
  if (!c || !p || !n) return IRF_EINVAL;
  IRF_TRY(c, {
    if (which == IRF_T_TIMING) { c->timing[8] = (double)(launch_counter() - c->launches0); c->timing[9] = c->n_inflate_launches; c->timing[12] = (double)tok_launch_counter(); *p = c->timing; *n = 14; }
    else {
      if (!c->finalized) throw string("no finished run to read tables from");
      switch (which) {
        case IRF_T_STATS: *p = c->h_stats.data(); *n = c->h_stats.size(); break;
        case IRF_T_SPANS: *p = c->h_sp.data(); *n = c->h_sp.size(); break;
        case IRF_T_ROI: *p = c->h_roi.data(); *n = c->h_roi.size(); break;
        case IRF_T_CHR: *p = c->h_chr.data(); *n = c->h_chr.size(); break;
        case IRF_T_JUNCTIONS: *p = c->h_junc.data(); *n = c->h_junc.size(); break;
        case IRF_T_ROWS_NONDIR: *p = c->rows_nondir.data(); *n = c->rows_nondir.size(); break;
        case IRF_T_ROWS_DIR: *p = c->rows_dir.data(); *n = c->rows_dir.size(); break;
        case IRF_T_FRAGMENTS: *p = c->h_dump.data(); *n = c->h_dump.size(); break;
        case IRF_T_DEPTH:
          if (c->h_depth.empty() && c->cov.P) { c->h_depth.resize(c->cov.P * 2); IRF_CUDA(cudaMemcpy(c->h_depth.data(), c->d_diff.p, c->cov.P * 8, cudaMemcpyDeviceToHost)); }
          *p = c->h_depth.data(); *n = c->h_depth.size(); break;
        default: throw string("unknown table code");
      }
    }
  });
}

int irf_reset(irf_ctx* c) { IRF_TRY(c, c->reset()); }
int irf_read_header(irf_ctx* c, const void* prefix, size_t n) {
  if (!prefix) return IRF_EINVAL;
  IRF_TRY(c, {
    // inflate just the leading blocks that hold the header, through the same device path as a run
    bool req = c->require_eof; c->require_eof = false;
    try { c->stage(prefix, 0, n, 1, false); } catch (...) { c->require_eof = req; throw; } c->require_eof = req;
    IRF_CUDA(cudaMemsetAsync(c->d_err.p, 0, 4, c->s_compute));   // a failed earlier run must not fail this one
    c->inflate_all(); c->join_crc(); c->check_dev_err("inflate"); c->parse_header_from_device(); c->staged = false;
    { u64 u = 0; c->records_coff = c->blks.empty() ? 0 : c->blks.back().c_off + c->blks.back().c_len + 8; for (auto& b : c->blks) { if (u >= c->rec0) { c->records_coff = b.file_off; break; } u += b.isize; } }
  });
}
int irf_stage(irf_ctx* c, const void* bgzf, size_t own, size_t total, int has_header) { if (!bgzf) return IRF_EINVAL; IRF_TRY(c, c->stage(bgzf, own, total, has_header, !c->stage_lazy)); }
int irf_count(irf_ctx* c) { IRF_TRY(c, c->count()); }
int irf_frame_info(irf_ctx* c, uint64_t* entry0, uint64_t* exit_rel) { if (!c || !c->counted) return IRF_ESTATE; if (entry0) *entry0 = c->frame_entry0; if (exit_rel) *exit_rel = c->frame_exit_rel; return IRF_OK; }
int irf_finalize(irf_ctx* c) { IRF_TRY(c, c->finalize()); }

int irf_get_device_table(irf_ctx* c, int which, void** dp, size_t* n, int* eb) {
  if (!c || !dp || !n || !eb || !c->bound) return c ? IRF_ESTATE : IRF_EINVAL;
  switch (which) {
    case IRF_T_STATS: *dp = c->d_stats.p; *n = ST_COUNT; *eb = 8; break;
    case IRF_T_SPANS: *dp = c->d_sp_count.p; *n = c->sp.pos.size() * 2; *eb = 4; break;
    case IRF_T_ROI: *dp = c->d_roi_count.p; *n = c->roi.rois.size() * 2; *eb = 4; break;
    case IRF_T_CHR: *dp = c->d_chr_count.p; *n = c->chr_names.size() * 2; *eb = 4; break;
    case IRF_T_DIFF: *dp = c->d_diff.p; *n = c->cov.P * 2; *eb = 4; break;
    case IRF_T_ROWS_NONDIR: *dp = c->d_rows.p; *n = c->cov.BEDrecords.size() * sizeof(IntronRow) / 8; *eb = 8; break;   // as 64-bit words, for the integer SUM merge of a region-owned run
    case IRF_T_ROWS_DIR: if (!c->rows_pending) return IRF_ESTATE; *dp = c->d_rows_dir.p; *n = c->cov.BEDrecords.size() * sizeof(IntronRow) / 8; *eb = 8; break;
    default: return IRF_EINVAL;
  }
  return IRF_OK;
}
int irf_export_junctions(irf_ctx* c, int which, void** keys, void** counts, size_t* n) {
  if (!keys || !counts || !n || which < 0 || which > 2) return IRF_EINVAL;
  IRF_TRY(c, { if (!c->counted) throw string("irf_count has not run"); u64 k = c->export_table(which); *keys = c->d_exp_keys[which].p; *counts = c->d_exp_cnt[which].p; *n = (size_t)k; });
}
int irf_import_junctions(irf_ctx* c, int which, const void* keys, const void* counts, size_t n) {
  if (which < 0 || which > 2) return IRF_EINVAL;
  IRF_TRY(c, {
    if (!c->bound) throw string("irf_count has not run");
    u64* tk = which == 0 ? c->T.jc_keys : which == 1 ? c->T.jl_keys : c->T.jr_keys; u32* tc = which == 0 ? c->T.jc_cnt : which == 1 ? c->T.jl_cnt : c->T.jr_cnt; u32 mask = which == 0 ? c->T.jc_mask : c->T.jl_mask;
    launch_ht_add(tk, tc, mask, (const u64*)keys, (const u32*)counts, n, c->T.err, c->s_compute); c->sync(); c->check_dev_err("junction merge");
  });
}
int irf_first_record(irf_ctx* c, int32_t* refid, int32_t* pos) { if (!refid || !pos) return IRF_EINVAL; IRF_TRY(c, c->first_record(*refid, *pos)); }
int irf_owner_plan(irf_ctx* c, const int32_t* cut_refid, const int64_t* cut_pos, int n_ranks, int my_rank) { if (!cut_refid || !cut_pos) return IRF_EINVAL; IRF_TRY(c, c->owner_plan(cut_refid, cut_pos, n_ranks, my_rank)); }
int irf_export_foreign_depth(irf_ctx* c, void** dev_index, void** dev_value, size_t* n) {
  if (!dev_index || !dev_value || !n) return IRF_EINVAL;
  IRF_TRY(c, { u64 k = c->export_foreign_depth(); *dev_index = c->d_fidx.p; *dev_value = c->d_fval.p; *n = (size_t)k; });
}
int irf_import_depth(irf_ctx* c, const void* dev_index, const void* dev_value, size_t n) { IRF_TRY(c, c->import_depth((const u64*)dev_index, (const int32_t*)dev_value, n)); }
int irf_finalize_rows(irf_ctx* c) { IRF_TRY(c, c->finalize_rows()); }
int irf_export_halo(irf_ctx* c, void** bytes, size_t* nbytes, void** offsets, size_t* nrec) {
  if (!bytes || !nbytes || !offsets || !nrec) return IRF_EINVAL;
  IRF_TRY(c, {
    if (!c->counted) throw string("irf_count has not run");
    u64 n = c->count_leftovers(nullptr, 0); c->d_left.ensure(n * 8); if (n) c->count_leftovers(c->d_left.as<u64>(), n);
    // each trimmed record is at most 4 + 32 + 255 + 4 * 65535 bytes; typical ones are < 100
    u64 cap = n * 320 + 1024; for (;;) {
      c->d_halo.ensure(cap + 64); c->d_halo_off.ensure(n * 8); u64* cur = (u64*)(c->d_small.as<uint8_t>() + 128); IRF_CUDA(cudaMemsetAsync(cur, 0, 8, c->s_compute));
      launch_halo_pack(c->d_U.as<uint8_t>(), c->d_left.as<u64>(), n, c->d_halo.as<uint8_t>(), c->d_halo_off.as<u64>(), cur, cap, c->s_compute);
      u64 used = 0; IRF_CUDA(cudaMemcpyAsync(&used, cur, 8, cudaMemcpyDeviceToHost, c->s_compute)); c->sync();
      if (used <= cap) { *nbytes = (size_t)used; break; } cap = used + 1024;
    }
    *bytes = c->d_halo.p; *offsets = c->d_halo_off.p; *nrec = (size_t)n; c->sharded = true;
  });
}
int irf_import_halo(irf_ctx* c, const void* bytes, size_t nbytes, const void* offsets, size_t nrec, int owner, int n_owners) {
  if (n_owners < 1 || owner < 0 || owner >= n_owners) return IRF_EINVAL;
  IRF_TRY(c, {
    if (!c->bound) throw string("irf_count has not run");
    c->sharded = true; c->halo_leftover = 0; if (!nrec) return IRF_OK;
    u64 slots = next_pow2(std::max<u64>(1024, 2 * nrec)); c->d_hptab.ensure(slots * 8); IRF_CUDA(cudaMemsetAsync(c->d_hptab.p, 0, slots * 8, c->s_compute));
    launch_halo_process(c->T, (const uint8_t*)bytes, nbytes, (const u64*)offsets, nrec, c->d_hptab.as<u64>(), (u32)(slots - 1), (u32)owner, (u32)n_owners, c->s_compute);
    u64* n_left = (u64*)(c->d_small.as<uint8_t>() + 64); IRF_CUDA(cudaMemsetAsync(n_left, 0, 8, c->s_compute));
    launch_pair_leftover(c->d_hptab.as<u64>(), slots, n_left, nullptr, 0, c->s_compute);
    u64 h = 0; IRF_CUDA(cudaMemcpyAsync(&h, n_left, 8, cudaMemcpyDeviceToHost, c->s_compute)); c->sync(); c->check_dev_err("halo pairing");
    c->halo_leftover = h;   // templates whose mate is in no shard: error reads (every rank sees the same gathered set, so the same number)
  });
}

}  // extern "C"

// ---- hooks of the in-process multi-device run (multi.h)
namespace irf {
int ctx_device(const irf_ctx* c) { return c->device; }
bool ctx_require_eof(const irf_ctx* c) { return c->require_eof != 0; }
uint64_t ctx_records_offset(const irf_ctx* c) { return c->records_coff; }
int ctx_count_shard(irf_ctx* c, const uint8_t* src, size_t own, size_t total, bool has_header, int fd, uint64_t file_off) {
  IRF_TRY(c, {
    c->reset();
    c->feed_fd = fd >= 0 && c->feed_capable() && (c->file_ring == 2 || total > (64u << 20)) ? fd : -1; c->feed_base = file_off;
    const bool req = c->require_eof; c->require_eof = false;   // the EOF block is checked once for the whole file (multi_run), not per shard
    try { c->stage(src, own, total, has_header ? 1 : 0, false); c->count(); } catch (...) { c->require_eof = req; c->feed_stop(); c->feed_fd = -1; c->feed_base = 0; c->h_src = nullptr; throw; }
    c->require_eof = req; c->feed_stop(); c->feed_fd = -1; c->feed_base = 0; c->h_src = nullptr;
  });
}
}  // namespace irf
