// irfinder — drop-in for the reference's counting binary (upstream src/irfinder/irfinder.cpp main):
//   irfinder <outdir> <ref-cover.bed> <ref-sj.ref> <ref-read-continues.ref> <ref-ROI.bed|NULL> <bam|/dev/stdin|->
// Same six positional arguments, same output files, counters and directionality on stdout, non-zero exit on error.
// Extras go in the environment so the reference invocation works unchanged:
//   IRF_DEVICE (GPU ordinal, default 0) or IRF_DEVICES (a list such as 0,1,2,3 or 0-7: the BAM file is cut into one byte range per device and
//   counted by all of them inside this process; an ordinal may repeat; stdin input uses the first device only), IRF_TRIM_CENTER, IRF_SAFE_INFLATE=1 (no hardware decompression engine), IRF_STDIN_INFLATED=1 (stdin carries inflated BAM
//   bytes, as the reference's `gzip -cd |` pipe does; default: auto-detect from the gzip magic).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <unistd.h>
#include <vector>
#include "irfinder_b200.h"

int main(int argc, char** argv) {
  if (argc != 7) { fprintf(stderr, "Usage: irfinder <outputDirectory> <ref-cover.bed> <ref-sj.ref> <ref-read-continues.ref> <ref-ROI.bed|NULL> <bam-file|/dev/stdin>\n"); return 1; }
  std::vector<int> devs;
  if (const char* l = getenv("IRF_DEVICES")) {   // "0,1,2" and ranges "0-7", mixed freely
    for (const char* p = l; *p;) {
      char* e; long a = strtol(p, &e, 10); if (e == p || a < 0) { fprintf(stderr, "irfinder: cannot parse IRF_DEVICES=%s\n", l); return 1; }
      long b = a; if (*e == '-') { const char* q = e + 1; b = strtol(q, &e, 10); if (e == q || b < a) { fprintf(stderr, "irfinder: cannot parse IRF_DEVICES=%s\n", l); return 1; } }
      for (long d = a; d <= b && devs.size() < 64; d++) devs.push_back((int)d);
      p = e; if (*p == ',') p++; else if (*p) { fprintf(stderr, "irfinder: cannot parse IRF_DEVICES=%s\n", l); return 1; }
    }
  }
  if (devs.empty()) devs.push_back(getenv("IRF_DEVICE") ? atoi(getenv("IRF_DEVICE")) : 0);
  irf_ctx* ctx = nullptr;
  if (irf_create(&ctx, devs.data(), (int)devs.size()) != IRF_OK) { fprintf(stderr, "irfinder: cannot initialise CUDA device %d%s\n", devs[0], devs.size() > 1 ? " (or another of IRF_DEVICES)" : ""); return 2; }
  if (const char* t = getenv("IRF_TRIM_CENTER")) irf_set_option(ctx, "trim_center", t);
  if (const char* t = getenv("IRF_SAFE_INFLATE")) irf_set_option(ctx, "safe_inflate", t);   // untrusted input: SM inflate kernel only (recoverable errors that name the problem)
  int rc = irf_load_ref(ctx, argv[2], argv[3], argv[4], argv[5]);
  if (rc == IRF_OK) {
    bool is_stdin = !strcmp(argv[6], "-") || !strcmp(argv[6], "/dev/stdin");
    if (!is_stdin) rc = irf_run_bam(ctx, argv[6], nullptr);
    else {
      int inflated = -1; if (const char* e = getenv("IRF_STDIN_INFLATED")) inflated = atoi(e) != 0;
      if (inflated < 0) {  // peek two bytes: 1f 8b => gzip/BGZF, 'B' 'A' => inflated BAM
        // stdin may be a pipe: read into a temp file-less buffer is the library's job, so sniff via a tiny relay pipe
        unsigned char b[2]; ssize_t k = read(0, b, 2); int p[2];
        if (k != 2 || pipe(p) != 0) { fprintf(stderr, "irfinder: cannot read stdin\n"); return 1; }
        inflated = !(b[0] == 0x1f && b[1] == 0x8b);
        if (fork() == 0) { close(p[0]); if (write(p[1], b, 2) != 2) _exit(1); char buf[1 << 16]; ssize_t n; while ((n = read(0, buf, sizeof buf)) > 0) { char* q = buf; while (n > 0) { ssize_t w = write(p[1], q, n); if (w <= 0) _exit(1); q += w; n -= w; } } _exit(0); }
        close(p[1]); rc = irf_run_stream(ctx, p[0], inflated);
      } else rc = irf_run_stream(ctx, 0, inflated);
    }
  }
  if (rc == IRF_OK) rc = irf_write_outputs(ctx, argv[1]);
  if (rc != IRF_OK) { fprintf(stderr, "irfinder: %s\n", irf_last_error(ctx)); irf_destroy(ctx); return 1; }
  fputs(irf_stdout_text(ctx), stdout);
  irf_destroy(ctx);
  return 0;
}
