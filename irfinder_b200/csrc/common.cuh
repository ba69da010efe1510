// common.cuh — shared types of the B200 BAM->IR counting path.
// Names follow the reference's domain: FragmentBlocks, JunctionCount, SpansPoint,
// CoverageBlocks, FragmentsInChr/ROI, ChrMapUpdate.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>

namespace irf {

typedef unsigned long long u64;
typedef uint32_t u32;

// ---- disputed behaviours: every rule the recollected specification is unsure about is a RUN-TIME setting, on
// this side (irf_set_option "rule_*") and, independently, in the oracle (environment IRF_RULE_*), so that the day the reference source
// can be read a wrong guess is a configuration change and not a kernel edit.  Defaults = the resolutions listed in DESIGN.md §1.
struct Rules {
  int junc_min_flank = 5;                  // [C-3]  both blocks flanking a junction must be at least this long
  u32 span_left = 5, span_right = 4;       // [C-4]  a span point p is counted when blockStart + left < p < blockEnd - right
  u32 skip_flags = 0;                      // [C-6]  extra flag bits whose records are skipped (0x100 secondary, 0x800 supplementary); unmapped always are
  int pct_tie_midpoint = 1;                // [C-12] cumulative count exactly on the percentile boundary: 1 = midpoint with the next depth, 0 = this depth
  u32 edge_offset = 5, edge_len = 50;      // [C-13] first/last windows of an intron
  double warn_lowcover = 10, warn_lowsplice = 4, warn_minor_ratio = 0.33333, warn_nonuniform_abs = 4, warn_nonuniform_ratio = 1.5;   // [C-14]
};
constexpr int BUCKET_SHIFT = 10;
constexpr int DEPTH_TILE = 2048;            // counters per strand in a CTA's shared-memory depth tile (a 2048-aligned window of the [2][P] difference table)
constexpr u32 REC_SLOTS = 65536 / 36 + 2;   // records that can START inside one BGZF block (a record has at least 36 bytes)

// ---- junction key: 36-bit concatenated-genome coordinate of the left end, 28-bit gap (CIGAR op lengths are 28-bit)
constexpr int JKEY_GAP_BITS = 28;
constexpr u64 JKEY_EMPTY = ~0ull;
__host__ __device__ inline u64 jkey_make(u64 gleft, u64 gap) { return (gleft << JKEY_GAP_BITS) | gap; }

// ---- error bits set by kernels in DevTables::err
enum DevErr : int {
  DERR_INFLATE = 1, DERR_TRUNCATED = 2, DERR_BAD_RECORD = 4, DERR_JUNC_FULL = 8, DERR_JKEY_RANGE = 16,
  DERR_PAIR_FULL = 32, DERR_REFID = 64, DERR_DUMP_FULL = 128, DERR_STARVED = 256, DERR_CRC = 512,
};

// ---- counters (the reference prints these on stdout; SURVEY.md §5 "Metrics")
enum Stat : int { ST_SINGLE = 0, ST_PAIRED, ST_SHORT, ST_INTERSECT, ST_LONG, ST_SKIPPED, ST_ERROR, ST_NUCLEOTIDES, ST_RECORDS, ST_COUNT };
constexpr int ST_CAND = ST_COUNT, ST_ADJ = ST_COUNT + 1, ST_SLOTS = ST_COUNT + 2;   // internal: pair-candidate records seen (orphans = candidates - 2 x pairs) and pairs formed by the adjacent-mate path; not part of the public table

struct Range { u32 begin, end; };

// Device-side view of every processor's tables after ChrMapUpdate (indexed by BAM refID).
struct DevTables {
  int n_ref;
  const int32_t* ref_len;   // [n_ref]
  const u64* chr_goff;      // [n_ref + 1] offset of each BAM reference in the concatenated coordinate space
  // coordinate buckets (1 << BUCKET_SHIFT bases): first table index whose key is >= / > the bucket start, so every
  // upper_bound becomes one load + a short forward scan instead of a ~20-step binary search
  const u64* bucket_base;   // [n_ref] first bucket of each reference
  const u32* bucket_n;      // [n_ref] number of buckets of each reference
  const u32* sp_bidx; const u32* seg_bidx; const u32* roi_bidx;
  // SpansPoint
  const u32* sp_pos; const Range* sp_range; u32* sp_count; u32 n_sp;          // sp_count[dir * n_sp + i]
  // CoverageBlocks: measured segments (sorted, disjoint), compact depth space with one sentinel slot per segment
  const u32* seg_start; const u32* seg_end; const u64* seg_coff; const Range* seg_range; int32_t* diff; u64 P;  // diff[dir * P + c]
  // FragmentsInROI
  const u32* roi_start; const u32* roi_end; const Range* roi_range; u32* roi_count; u32 n_roi;  // roi_count[dir * n_roi + i]
  // FragmentsInChr
  u32* chr_count;           // [dir * n_ref + refID]
  // JunctionCount: open-addressing tables, keys + counts[slot * 2 + dir]
  u64* jc_keys; u32* jc_cnt; u32 jc_mask;
  u64* jl_keys; u32* jl_cnt; u32 jl_mask;   // left ends (key = concatenated coordinate)
  u64* jr_keys; u32* jr_cnt; u32 jr_mask;   // right ends
  u64* stats;               // [ST_SLOTS]
  int* err;
  // optional fragment dump (tests): words appended at dump[atomicAdd(dump_cursor, n)]
  u32* dump; u64* dump_cursor; u64 dump_cap;
  Rules rules;
  int depth_tile;           // CoverageBlocks counters of a CTA's fragments are first summed in a shared-memory tile: 0 = off (global atomics), 1 = tile flushed with
                            // one red.global per non-zero counter, 2 = tile flushed by one TMA bulk reduce per strand (cp.reduce.async.bulk ... add.s32)
};

// One row of IRFinder-IR-*.txt in numeric form; text formatting happens on the host.
struct IntronRow {
  long long excluded; u64 positions;
  double coverage, depth, p25, p50, p75, first50, last50, irratio;
  u32 sp_left, sp_right, jc_left, jc_right, jc_exact; int warning;
};
enum Warning : int { W_NONE = 0, W_LOWCOVER, W_LOWSPLICING, W_MINORISOFORM, W_NONUNIFORM };

// Device description of one ref-cover.bed record.
struct IntronDesc {
  u32 blk_begin, blk_main_end, blk_first_end, blk_last_end;  // [begin,main_end) measured blocks, then first-50 pieces, then last-50 pieces
  u32 start, end; int refid;                                  // refid = -1 when the chromosome is absent from the BAM
  int sp_left, sp_right;                                      // indices into sp arrays or -1
  u32 plus;                                                   // annotated strand
};
struct IntronBlock { u64 c0; u32 len; u32 pad; };            // compact range [c0, c0+len)

#define IRF_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) throw std::string(#call) + ": " + cudaGetErrorString(e_); } while (0)

}  // namespace irf
