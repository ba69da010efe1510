/* irfinder_b200.h — C ABI of the B200-native IRFinder BAM -> intron-retention counting core.
 *
 * The reference exposes this path only as a process boundary:
 *     irfinder <outdir> <ref-cover.bed> <ref-sj.ref> <ref-read-continues.ref> <ref-ROI.bed|NULL> <bam|/dev/stdin>
 * (upstream src/irfinder/irfinder.cpp main(); the mounted /root/reference holds only README.md:1-7, so no
 * file:line inside it can be cited — SURVEY.md §0).  Each entry point below names the reference routine it
 * stands in for.  Conventions: every call returns 0 or a negative IRF_E* code; no C++ exception crosses
 * the ABI; the caller owns every pointer it passes; memory returned by irf_get_table belongs to the
 * context until the next run or irf_destroy; a context is used from one host thread at a time.  Multi-GPU:
 * either one context created over several devices (irf_create with n_devices > 1: the library shards, merges
 * and finalizes inside the process), or one process and one context per GPU driven over NCCL.
 */
#ifndef IRFINDER_B200_H
#define IRFINDER_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct irf_ctx irf_ctx;

enum { IRF_OK = 0, IRF_EINVAL = -1, IRF_EIO = -2, IRF_EFORMAT = -3, IRF_ECUDA = -4, IRF_ESTATE = -5, IRF_ECAPACITY = -6 };

/* which-codes of irf_get_table / irf_get_device_table */
enum {
  IRF_T_STATS = 0,        /* uint64[9]: singles, pairs, short, intersect, long, skipped, error, nucleotides, records (reference: stdout counters) */
  IRF_T_SPANS = 1,        /* uint32[2][n_points]: [0]=neg/dir0, [1]=pos/dir1, chromosome-name order (SpansPoint counts) */
  IRF_T_ROI = 2,          /* uint32[2][n_roi] (FragmentsInROI counts) */
  IRF_T_CHR = 3,          /* uint32[2][n_ref], BAM header order (FragmentsInChr counts) */
  IRF_T_JUNCTIONS = 4,    /* struct {int32 refid; uint32 start, end, neg, pos;}[n] counted junctions (JunctionCount map) */
  IRF_T_DEPTH = 5,        /* int32[2][P] per-base depth over the compact measured space (CoverageBlock counters after scan); after a region-owned
                             sharded run (irf_owner_plan) only the segments this rank owns hold the merged depth */
  IRF_T_ROWS_NONDIR = 6,  /* irf_intron_row[n_introns], ref-cover.bed order (IRFinder-IR-nondir.txt, numeric) */
  IRF_T_ROWS_DIR = 7,     /* same for IRFinder-IR-dir.txt; n = 0 when the library is unstranded */
  IRF_T_FRAGMENTS = 8,    /* uint32 words of the FragmentBlocks dump (option dump_fragments=1): per fragment
                             [n_words, chr, dir, readCount, nblk0, nblk1, (start,end)*] */
  IRF_T_TIMING = 9,       /* double[14] of the last run: ms h2d+inflate, ms inflate kernels, ms framing, ms counting, ms finalize, ms count total, ms host BGZF scan, ms output formatting, kernels launched since irf_reset, inflate launches, ms of the SM inflate launch(es) alone, share of the algorithmic bytes they inflated, launches of the two-phase inflate kernel by this process so far, pairs formed by the adjacent-mate path */
  IRF_T_DIFF = 10         /* device only: int32[2][P] difference counters before irf_finalize */
};

typedef struct irf_intron_row {
  long long excluded; unsigned long long positions;
  double coverage, depth, p25, p50, p75, first50, last50, irratio;
  uint32_t sp_left, sp_right, jc_left, jc_right, jc_exact; int32_t warning; /* 0 -, 1 LowCover, 2 LowSplicing, 3 MinorIsoform, 4 NonUniformIntronCover */
} irf_intron_row;

/* Creates a context.  Stands in for process start of `irfinder`.
 * n_devices == 1: everything runs on device_ids[0].
 * n_devices > 1 (at most 64; an ordinal may repeat = several shards on one GPU): irf_run_bam and irf_run_buffer cut the BAM into one
 * BGZF byte range per entry (never inside the header; fewer ranges when the file has fewer blocks), count the ranges side by side
 * (one host thread per device), exchange what the reference's single set of processors would have seen together - frame hand-shake,
 * cross-range mates, junction tables, span / ROI / chromosome / read counters, region-owned depth - and leave the complete, finalized
 * tables in this context: irf_write_outputs, irf_get_table, irf_stdout_text work as after a single-device run and give the same
 * bytes.  irf_set_option and irf_load_ref apply to every device; irf_run_stream (a pipe) and the staged calls below use
 * device_ids[0] only.  (csrc/multi.cu; the irfinder binary takes the list from IRF_DEVICES=0,1,... or 0-7.) */
int irf_create(irf_ctx** out, const int* device_ids, int n_devices);
void irf_destroy(irf_ctx* ctx);
const char* irf_last_error(irf_ctx* ctx);

/* Options (strings): "trim_center" (default 40), "dump_fragments" (0/1), "inflate_variant" (A/B variants of the SM inflate kernel: 12 = two-phase
 * token decode + warp-wide match resolve, 0-11 = the one-phase lane-per-block kernel and its table/occupancy variants),
 * "safe_inflate" (1 = untrusted input: every block through the SM kernel, which answers malformed DEFLATE with IRF_EFORMAT and leaves the context usable),
 * "de_percent" (default 70) / "de_percent_stream" (default 100): share of the BGZF blocks inflated by the Blackwell hardware
 * decompression engine when the bytes are resident / streamed from the host; the rest go to the SM kernel; 0 = SM kernel only.
 * "resident_pipeline" (default 1): large resident inputs are counted by the overlapped chunk pipeline, in which the engine takes the
 * trailing "de_percent_overlap" % (default 71; the kernel's share is rounded down to whole grid-fulls of blocks) of the file's chunks and the kernel the leading part; 0 = always stage by stage, 2 = always
 * pipelined when structurally possible.  "file_ring" (default 1): irf_run_bam sends files above 64 MB through a ring of page-locked
 * buffers instead of page-locking a copy of the file; 0 = never, 2 = always.  "stream_pipeline" (default 1) / "stream_chunk_kb" (default 262144): chunk pipeline for host bytes.
 * NOTE: the engine answers malformed DEFLATE data with a sticky CUDA error (the process must restart), the SM kernel with IRF_EFORMAT.
 * "junc_capacity_log2" (default 24), "require_eof" (default 1), "verify_crc" (default 1: check every block's CRC-32 as gzip does), "stage_lazy" (default 0; 1 = irf_stage only records the host
 * buffer and irf_count streams it to the GPU beside the inflate kernel: the buffer must then stay valid until irf_count returns).
 * "depth_tile" (default 0): 1 / 2 = CoverageBlocks counters of a CTA are summed in a shared-memory tile and flushed with red.global / by one TMA
 * bulk reduce per strand (measured slower than the plain global atomics of 0 on the benchmark shapes; kept selectable).
 * "adjacent_pairs" (default -1): mates that follow each other in the file (name-collated input, STAR's pipe) are paired without the hash table;
 * -1 = on unless the header says @HD SO:coordinate, 0 = off, 1 = on.  A matter of speed only: the pairs are the same either way.
 * "stream_segment_mb" (default 1024): irf_run_stream reads its pipe in segments of this size (two page-locked buffers of it).
 * "rule_*": the behaviours SURVEY.md Appendix C marks as uncertain, as run-time settings (defaults = DESIGN.md section 1): rule_junc_min_flank (5),
 * rule_span_left (5), rule_span_right (4), rule_skip_flags (0; e.g. 0x900 skips secondary and supplementary records), rule_pct_tie_midpoint (1),
 * rule_edge_offset (5), rule_edge_len (50), and in millionths rule_warn_lowcover_e6 (10000000), rule_warn_lowsplice_e6 (4000000),
 * rule_warn_minor_ratio_e6 (333330), rule_warn_nonuniform_abs_e6 (4000000), rule_warn_nonuniform_ratio_e6 (1500000).
 * Must be set before irf_load_ref / the first run (the rule_* and safe_inflate options may also be changed between runs). */
int irf_set_option(irf_ctx* ctx, const char* key, const char* value);

/* Replaces CoverageBlocksIRFinder::loadRef, JunctionCount::loadRef, SpansPoint::setSpanLength(5,4)+loadRef and
 * FragmentsInROI::loadRef (irfinder.cpp main).  roi_bed may be NULL or "NULL". */
int irf_load_ref(irf_ctx* ctx, const char* cover_bed, const char* sj_ref, const char* read_continues_ref, const char* roi_bed);

/* Replaces BAM2blocks::openFile + processAll over a whole BGZF BAM file, followed by the output-side statistics
 * (everything of main() between loadRef and the file writers).  bai_path is accepted for CLI compatibility and ignored:
 * record framing needs no index.  Mates may be in any order (coordinate-sorted or mate-adjacent). */
int irf_run_bam(irf_ctx* ctx, const char* bam_path, const char* bai_path);
/* Same, from a host buffer holding the BGZF bytes (pinned memory copies fastest). */
int irf_run_buffer(irf_ctx* ctx, const void* bgzf, size_t nbytes);
/* Reference-compatible stdin mode (irfinder.cpp main reading /dev/stdin behind `gzip -cd` or STAR): fd delivers BGZF/gzip bytes (inflated = 0) or an
 * already-inflated BAM stream (inflated = 1).  The stream is read and counted in bounded segments ("stream_segment_mb"), so its length is not limited
 * by host memory; a record must fit into the 8 MB of look-ahead that follow a segment. */
int irf_run_stream(irf_ctx* ctx, int fd, int inflated);

/* Replaces the six file writers at the end of main(): IRFinder-IR-nondir.txt, IRFinder-IR-dir.txt (stranded libraries only),
 * IRFinder-JuncCount.txt, IRFinder-SpansPoint.txt, IRFinder-ChrCoverage.txt, IRFinder-ROI.txt (when an ROI file was loaded). */
int irf_write_outputs(irf_ctx* ctx, const char* outdir);
/* The stdout block of the reference (directionality evidence + read counters) as one NUL-terminated string. */
const char* irf_stdout_text(irf_ctx* ctx);
/* +1 / 0 / -1 as returned by JunctionCount::Directional. */
int irf_directionality(irf_ctx* ctx);

/* Raw tables for tests and embedding (host copies). */
int irf_get_table(irf_ctx* ctx, int which, const void** host_ptr, size_t* n_elems);

/* ---- staged operation: the same pipeline in separately timed / separately sharded steps ---- */
/* Zeroes every counter (keeps the loaded references). */
int irf_reset(irf_ctx* ctx);
/* Parses the BAM header (BAM2blocks::readBamHeader + every ChrMapUpdate) from the first bytes of a BGZF BAM. */
int irf_read_header(irf_ctx* ctx, const void* bgzf_prefix, size_t nbytes);
/* Copies BGZF bytes to the GPU.  Records that START in the first nbytes_own bytes' blocks are this context's to count;
 * the bytes up to nbytes_total are look-ahead so a record straddling the shard end can be read.  has_header = 1 when the
 * buffer starts at the beginning of the file.  The staged bytes stay resident until the next irf_stage, so irf_reset +
 * irf_count + irf_finalize can be repeated on them. */
int irf_stage(irf_ctx* ctx, const void* bgzf, size_t nbytes_own, size_t nbytes_total, int has_header);
/* Inflate + record framing + per-fragment counting of the staged bytes, device only. */
int irf_count(irf_ctx* ctx);
/* Framing hand-shake between neighbouring shards: entry = inflated offset of the first record counted here,
 * exit = inflated offset (relative to the end of the own bytes) where the next shard's first record starts. */
int irf_frame_info(irf_ctx* ctx, uint64_t* entry0, uint64_t* exit_rel);
/* Depth scan, junction merge, directionality and per-intron statistics on the (possibly all-reduced) tables. */
int irf_finalize(irf_ctx* ctx);
/* Device pointers of the additive tables, for an NCCL all-reduce by the caller (IRF_T_SPANS, _ROI, _CHR, _STATS, _DIFF; in owner mode
 * also IRF_T_ROWS_NONDIR / IRF_T_ROWS_DIR as 64-bit words). */
int irf_get_device_table(irf_ctx* ctx, int which, void** dev_ptr, size_t* n_elems, int* elem_bytes);
/* Junction tables as dense device arrays (keys uint64, counts uint32[n][2]); which = 0 exact, 1 left ends, 2 right ends. */
int irf_export_junctions(irf_ctx* ctx, int which, void** dev_keys, void** dev_counts, size_t* n);
int irf_import_junctions(irf_ctx* ctx, int which, const void* dev_keys, const void* dev_counts, size_t n);
/* Records whose mate was not seen in this shard, trimmed to fixed part + name + CIGAR, and their byte offsets. */
int irf_export_halo(irf_ctx* ctx, void** dev_bytes, size_t* nbytes, void** dev_offsets, size_t* nrec);
/* Pairs the gathered halo records of all shards; this context counts the pairs whose key hash % n_owners == owner. */
int irf_import_halo(irf_ctx* ctx, const void* dev_bytes, size_t nbytes, const void* dev_offsets, size_t nrec, int owner, int n_owners);

/* ---- region-owned depth for sharded runs ("Depth needs no reduction (region-owned) ... per-intron stats computed by
 * the owner; rows gathered to rank 0").  The reference has one process and one CoverageBlocks object; these calls split that object's
 * counters (CoverageBlocks::ProcessBlocks / CoverageBlocksIRFinder::WriteOutput, ReadBlockProcessor_CoverageBlocks.cpp) by genomic region. ---- */
/* refID / pos of the first record this context counted (refid = -2 when its shard held no record). */
int irf_first_record(irf_ctx* ctx, int32_t* refid, int32_t* pos);
/* Gives every measured segment an owning rank from the first-record coordinates of all ranks (same arrays on every rank); segments
 * that share a ref-cover record stay together.  Switches the context to owner mode until the next irf_reset. */
int irf_owner_plan(irf_ctx* ctx, const int32_t* first_refid, const int64_t* first_pos, int n_ranks, int my_rank);
/* This rank's non-zero difference counters in segments it does not own: device arrays of uint64 table indices and int32 values. */
int irf_export_foreign_depth(irf_ctx* ctx, void** dev_index, void** dev_value, size_t* n);
/* Adds the listed counters that fall into segments this rank owns (entries for other owners are skipped, so every rank can be handed
 * the concatenation of all ranks' lists). */
int irf_import_depth(irf_ctx* ctx, const void* dev_index, const void* dev_value, size_t n);
/* In owner mode irf_finalize computes the rows of the owned introns only and leaves both row tables on the device with all other rows
 * zero: the caller adds the tables of all ranks as 64-bit integers (irf_get_device_table IRF_T_ROWS_NONDIR / IRF_T_ROWS_DIR; exactly one
 * rank contributes a non-zero row, so the sum is exact) and then calls irf_finalize_rows, which brings them to the host. */
int irf_finalize_rows(irf_ctx* ctx);

/* No counterpart in the reference (it has no GPU path and therefore no allocation question).
 * Depth scan for callers that must not allocate GPU memory: out[i] = in[0] + ... + in[i] over `n` int32 difference
 * counters, the same operation irf_finalize applies to the CoverageBlocks counters.  `diff_dev` and `depth_dev` are the
 * caller's device buffers (they may be the same buffer); `cuda_stream` is a cudaStream_t or NULL; the work is queued on
 * that stream and the call returns without waiting for it.  The call performs no cudaMalloc/cudaFree: its scratch is
 * a 128 KB array that is part of the library image on every GPU.  Consequences the caller must know:
 *   - at most 2^27 counters (512 MB) per call; a larger n is refused with IRF_ECAPACITY and an explanatory message;
 *   - all calls on a GPU share that array, so calls that overlap in time on one GPU corrupt each other: serialise them;
 *   - the first call made with a given GPU current looks the array's address up in the CUDA runtime; make that first
 *     call during set-up, before any stream capture starts.
 * `errbuf` (optional) receives a NUL-terminated message on failure.  Returns IRF_OK or a negative IRF_E* code. */
int irf_depth_scan_static(const int32_t* diff_dev, int32_t* depth_dev, size_t n, void* cuda_stream, char* errbuf, size_t errbuf_len);

#ifdef __cplusplus
}
#endif
#endif
