// In-process multi-device run: one context per device, all driven from one irf_create ("one ctx ... drives all
// its GPUs internally").  Internal to the library; the C ABI in include/irfinder_b200.h is the public surface.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

struct irf_ctx;

namespace irf {

// context.cu: what multi.cu needs from a context beyond the C ABI
int ctx_device(const irf_ctx* c);
bool ctx_require_eof(const irf_ctx* c);
// after irf_read_header: file offset of the first BGZF block that starts at or after the first alignment record; a shard cut must not lie before it
uint64_t ctx_records_offset(const irf_ctx* c);
// reset + stage [src, src + total) (own bytes first, look-ahead after) streamed beside the inflate + count.  With fd >= 0 the
// bytes are read from the file at file_off through the context's ring of page-locked buffers (src is the mapping of the same
// bytes, used for the BGZF block scan only).  Returns an IRF_* code, message in irf_last_error.
int ctx_count_shard(irf_ctx* c, const uint8_t* src, size_t own, size_t total, bool has_header, int fd, uint64_t file_off);

// multi.cu: counts [base, base + n) (a whole BGZF BAM) over ranks[0..], merges, finalizes; ranks[0] then holds the complete
// tables.  Throws std::string.  fd >= 0: `base` is a mapping of that file.
void multi_run(const std::vector<irf_ctx*>& ranks, const uint8_t* base, size_t n, int fd);

}  // namespace irf
