// Host-side emulation harness (TEST ONLY): runs irfinder_b200/csrc/inflate_core.cuh — the exact routine the
// GPU lanes execute — on the CPU over every BGZF block of a file and compares each block with zlib.
// usage: inflate_emu <file.bam> [LIT_BITS variant: 8|9|10]
#include <zlib.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../irfinder_b200/csrc/inflate_core.cuh"
#include "../../irfinder_b200/csrc/inflate_tok.cuh"

template <int LB, int DB, int PZ = 0, int D8 = 0>
static int run(const std::vector<uint8_t>& f) {
  std::vector<uint16_t> lit(1 << LB), dist((1 << DB) + 64);  // dist LUT + the two 32-entry aux tables behind it
  size_t off = 0; long nblk = 0, bad = 0; std::vector<uint8_t> a(65536 + 64), b(65536 + 64);
  while (off + 18 <= f.size()) {
    const uint8_t* h = &f[off];
    if (h[0] != 31 || h[1] != 139 || h[12] != 'B' || h[13] != 'C') { fprintf(stderr, "bad BGZF header at %zu\n", off); return 2; }
    uint32_t bsize = (h[16] | h[17] << 8) + 1u, clen = bsize - 26, isize; memcpy(&isize, &f[off + bsize - 4], 4);
    std::fill(a.begin(), a.end(), 0xAA);
    int st = irf::inflate_block<1, LB, DB, PZ, D8>(&f[off + 18], clen, a.data() + (nblk % 37), isize, lit.data(), dist.data(), 0);  // vary output alignment
    z_stream zs; memset(&zs, 0, sizeof zs); inflateInit2(&zs, -15); zs.next_in = (Bytef*)&f[off + 18]; zs.avail_in = clen; zs.next_out = b.data(); zs.avail_out = 65536;
    int zr = inflate(&zs, Z_FINISH); inflateEnd(&zs);
    if (st != 0 || zr != Z_STREAM_END || zs.total_out != isize || memcmp(a.data() + (nblk % 37), b.data(), isize) != 0 || a[(nblk % 37) + isize] != 0xAA || (nblk % 37 && a[(nblk % 37) - 1] != 0xAA)) {
      if (bad < 5) fprintf(stderr, "block %ld at %zu: status %d zlib %d isize %u\n", nblk, off, st, zr, isize); bad++; }
    off += bsize; nblk++;
  }
  printf("blocks %ld bad %ld\n", nblk, bad); return bad ? 1 : 0;
}
// The two-phase decoder (inflate_tok.cuh): token decode by the lane routine, then the scalar reference resolve.  The output buffer
// starts as 0xAA so that an unwritten literal, an unfilled match or a byte written outside the block shows up.
template <int LB, int DB, int FUSE = (LB <= 6 ? 1 : 0)>
static int run_tok(const std::vector<uint8_t>& f) {
  std::vector<uint16_t> lit(1 << LB), dist((1 << DB) + 64); std::vector<uint32_t> tok(irf::TOK_CAP);
  size_t off = 0; long nblk = 0, bad = 0; unsigned long long ntok_total = 0, out_total = 0; uint32_t ntok_max = 0; std::vector<uint8_t> a(65536 + 64), b(65536 + 64);
  while (off + 18 <= f.size()) {
    const uint8_t* h = &f[off];
    if (h[0] != 31 || h[1] != 139 || h[12] != 'B' || h[13] != 'C') { fprintf(stderr, "bad BGZF header at %zu\n", off); return 2; }
    uint32_t bsize = (h[16] | h[17] << 8) + 1u, clen = bsize - 26, isize; memcpy(&isize, &f[off + bsize - 4], 4);
    std::fill(a.begin(), a.end(), 0xAA); uint32_t nt = 0; const int sk = 8 + (int)(nblk % 37);
    int st = irf::inflate_block_tok<1, LB, DB, FUSE>(&f[off + 18], clen, a.data() + sk, isize, lit.data(), dist.data(), 0, tok.data(), &nt);
    z_stream zs; memset(&zs, 0, sizeof zs); inflateInit2(&zs, -15); zs.next_in = (Bytef*)&f[off + 18]; zs.avail_in = clen; zs.next_out = b.data(); zs.avail_out = 65536;
    int zr = inflate(&zs, Z_FINISH); inflateEnd(&zs);
    bool outside = a[sk + isize] != 0xAA || a[sk - 1] != 0xAA || a[sk + isize + 3] != 0xAA || a[sk - 4] != 0xAA;
    if (st != 0 || zr != Z_STREAM_END || zs.total_out != isize || memcmp(a.data() + sk, b.data(), isize) != 0 || outside || nt > irf::TOK_CAP) {
      if (bad < 5) fprintf(stderr, "block %ld at %zu: status %d zlib %d isize %u tokens %u outside %d\n", nblk, off, st, zr, isize, nt, (int)outside); bad++; }
    ntok_total += nt; out_total += isize; if (nt > ntok_max) ntok_max = nt;
    off += bsize; nblk++;
  }
  printf("blocks %ld bad %ld tokens/KB %.1f max tokens %u\n", nblk, bad, out_total ? 1024.0 * ntok_total / out_total : 0.0, ntok_max); return bad ? 1 : 0;
}
int main(int argc, char** argv) {
  FILE* fp = fopen(argv[1], "rb"); if (!fp) return 2; std::vector<uint8_t> f; uint8_t buf[1 << 16]; size_t n;
  while ((n = fread(buf, 1, sizeof buf, fp)) > 0) f.insert(f.end(), buf, buf + n); fclose(fp); f.reserve(f.size() + 64);   // look-ahead loads read a few bytes past the last block
  int v = argc > 2 ? atoi(argv[2]) : 9;
  if (v == 406) return run_tok<6, 5>(f); if (v == 407) return run_tok<7, 5>(f); if (v == 408) return run_tok<8, 6>(f); if (v == 409) return run_tok<9, 6>(f); if (v == 416) return run_tok<6, 5, 0>(f);   // 406: fused iteration (the default), 416: one symbol kind per iteration
  if (v == 8) return run<8, 6>(f); if (v == 10) return run<10, 7>(f); if (v == 7) return run<7, 5>(f); if (v == 6) return run<6, 5>(f);
  if (v == 306) return run<6, 5, 0, 1>(f); if (v == 307) return run<7, 7, 0, 1>(f);   // byte entries in the distance / code-length LUT
  if (v == 206) return run<6, 5, 1>(f); return run<9, 6>(f);
}
