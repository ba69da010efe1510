// kernels.h — host-callable launchers of kernels.cu (all asynchronous on the given stream).
#pragma once
#include "common.cuh"

namespace irf {

struct InflateArgs {
  const uint8_t* comp; const u64* c_off; const u32* c_len;   // compressed bytes, per-block CDATA offset / length
  uint8_t* U; const u64* u_off; const u32* u_len;            // inflated bytes, per-block offset / ISIZE
  int nblk; int* status; int* err; unsigned int* next_block;   // next_block: device work counter, zeroed by the launcher
  const unsigned long long* watermark;   // optional: bytes of `comp` that have landed so far (streaming from the host); null = all resident
  const u32* list;                       // optional: nblk indices into the block arrays (blocks not listed are inflated elsewhere)
  uint32_t* tok_scratch = nullptr; size_t tok_scratch_bytes = 0;   // two-phase kernel (variant 12): match-token scratch, inflate_scratch_bytes() bytes
};
struct FrameArgs {
  const uint8_t* U; const u64* u_off; const u32* u_len; int nblk; u64 Uend; u64 rec0; int rec0_valid;  // rec0: inflated offset of the first record when known exactly
  int n_ref; const int32_t* ref_len;
  u64* entry; u64* exit_off; u32* n_rec; u32* n_cand; uint8_t* ferr; int* err;  // ferr: per-block framing error of the latest walk
  int* first = nullptr;   // device scalar: index of the first block that has an entry (set by launch_frame_guess, read by launch_frame_mark)
  u64* slots = nullptr;   // optional: REC_SLOTS record starts per block, filled by the walk (launch_gather_records turns them into the dense list)
};

void launch_inflate(const InflateArgs& a, int variant, cudaStream_t s);
size_t inflate_scratch_bytes(int variant, int nblk);   // device scratch the chosen kernel needs for a launch over nblk blocks (0 for the one-phase variants)
void launch_inflate_tok(const InflateArgs& a, int variant, cudaStream_t s);   // inflate_tok.cu; variants 12-15 differ in LUT width / warps per SM
size_t inflate_tok_scratch_bytes(int variant, int nblk);
int inflate_tok_lanes(int variant);
unsigned long long tok_launch_counter();
void note_launch();                                     // bumps the library's kernel-launch counter (launchers outside kernels.cu)
int inflate_resident_lanes();
void set_inflate_cta_cap(int ctas_per_sm);
// CRC-32 of every inflated BGZF block against the value in its footer (what `gzip -cd` checks for the reference)
void launch_host_fetch(void* dst_dev, const void* src_host_mapped, size_t nbytes, cudaStream_t s);   // nbytes is rounded up to 16: both buffers must be padded
void launch_de_check(const u32* list, int n, const u32* act, const u32* u_len, int* status, int* err, cudaStream_t s);
void launch_crc_check(const uint8_t* U, const u64* u_off, const u32* u_len, const u32* crc_expected, int nblk, int* status, int* err, cudaStream_t s, const u32* list = nullptr);
void launch_frame_guess(const FrameArgs& a, cudaStream_t s);
void launch_walk_count(const FrameArgs& a, const uint8_t* mask, cudaStream_t s);
void launch_frame_mark(const FrameArgs& a, uint8_t* mism_scratch, uint8_t* mask, u32* n_bad, cudaStream_t s);   // both byte arrays hold a.nblk entries
void launch_frame_totals(const FrameArgs& a, u64* totals, cudaStream_t s);
void launch_walk_list(const FrameArgs& a, const u64* rec_base, u64* rec_off, cudaStream_t s);
void launch_scan_counts(const u32* in, u64* out, int n, cudaStream_t s);
void launch_pair_records(const DevTables& T, const uint8_t* U, const u64* rec_off, u64 nrec, u64* ptab, u32 pmask, u64* pair_list, u64* single_list, u64* list_counts, cudaStream_t s, u64 carry_base = ~0ull, int count_stats = 1, int adjacent = 0);
void launch_shift_offsets(u64* off, u64 n, u64 delta, cudaStream_t s);
void launch_process_pairs(const DevTables& T, const uint8_t* U, const u64* pair_list, u64 npairs, cudaStream_t s, const u64* npairs_dev = nullptr);
void launch_process_singles(const DevTables& T, const uint8_t* U, const u64* single_list, u64 nsingles, cudaStream_t s, const u64* nsingles_dev = nullptr);
void launch_gather_records(const FrameArgs& a, const u64* rec_base, u64* rec_off, cudaStream_t s);
void launch_pair_leftover(const u64* ptab, u64 nslots, u64* n_left, u64* left_off, u64 left_cap, cudaStream_t s);
void launch_halo_pack(const uint8_t* U, const u64* left_off, u64 n, uint8_t* out, u64* rec_off, u64* cursor, u64 cap, cudaStream_t s);
unsigned long long launch_counter();
void launch_halo_process(const DevTables& T, const uint8_t* H, u64 Hend, const u64* rec_off, u64 nrec, u64* ptab, u32 pmask, u32 owner, u32 nowners, cudaStream_t s);
void launch_ht_insert_keys(u64* keys, u32 mask, const u64* in, u64 n, int* err, cudaStream_t s);
void launch_ht_add(u64* keys, u32* cnt, u32 mask, const u64* in_keys, const u32* in_cnt, u64 n, int* err, cudaStream_t s);
void launch_ht_compact(const u64* keys, const u32* cnt, u64 nslots, u64* out_keys, u32* out_cnt, u64* cursor, u64 cap, int only_counted, cudaStream_t s);
size_t junction_sort_temp_bytes(u64 n);
void launch_junction_rows(const u64* keys, const u64* cnt, u64 n, const u64* chr_goff, int n_ref, void* out_rows, cudaStream_t s);   // 20-byte rows {refID, start, end, count[0], count[1]}
void launch_junction_sort(void* temp, size_t temp_bytes, const u64* keys_in, u64* keys_out, const u64* cnt_in, u64* cnt_out, u64 n, cudaStream_t s);
void launch_depth_scan(int32_t* d, u64 n, int32_t* tile_sums, cudaStream_t s, u64 P = 0, u32* dmax = nullptr);   // dmax: two words, zeroed by the caller, receive the largest depth below / from element P
u64 depth_scan_tiles(u64 n);
void launch_max_depth(const int32_t* depth, u64 P, u32* out, cudaStream_t s);
void launch_intron_stats(const DevTables& T, const int32_t* depth, const IntronDesc* introns, const IntronBlock* blocks, u32 n_introns, int directionality, u32 trim_center,
                         u32* scratch, u32 scratch_bins, int grid, IntronRow* rows, cudaStream_t s, const u32* list = nullptr, const u32* n_dev = nullptr);
// one-pass fast path (depth < 2048 everywhere in the intron); deeper introns are appended to overflow[*n_overflow++] for launch_intron_stats
void launch_intron_stats_fast(const DevTables& T, const int32_t* depth, const IntronDesc* introns, const IntronBlock* blocks, u32 n_introns, int directionality, u32 trim_center,
                              IntronRow* rows, cudaStream_t s, const u32* list, u32* overflow, u32* n_overflow);
void launch_foreign_depth(const DevTables& T, u32 nseg, const uint8_t* seg_owner, u32 me, const uint8_t* tile_state, u64* out_idx, int32_t* out_val, u64* cursor, u64 cap, cudaStream_t s);   // tile_state: per 1024 slots, 1 all mine / 2 none mine / 0 mixed
void launch_import_depth(const DevTables& T, u32 nseg, const uint8_t* seg_owner, u32 me, const u64* idx, const int32_t* val, u64 n, cudaStream_t s);

}  // namespace irf
