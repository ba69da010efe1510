// de_probe — does this GPU expose the Blackwell hardware decompression engine for raw DEFLATE (cuMemBatchDecompressAsync),
// and how fast does it inflate the BGZF blocks of a BAM?  Developer probe; not part of the library.
//   nvcc -O2 -std=c++17 -o de_probe de_probe.cu -lcuda -lz ; ./de_probe file.bam [batch]
#include <cuda.h>
#include <cuda_runtime.h>
#include <zlib.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#define CK(x) do { CUresult r_ = (x); if (r_ != CUDA_SUCCESS) { const char* s_; cuGetErrorString(r_, &s_); printf("%s -> %d %s\n", #x, (int)r_, s_ ? s_ : "?"); return 1; } } while (0)
int main(int argc, char** argv) {
  if (argc < 2) return 2;
  size_t batch = argc > 2 ? atol(argv[2]) : 65536; int params_on_device = argc > 3 ? atoi(argv[3]) : 0;
  cudaFree(0); CUdevice dev; CK(cuCtxGetDevice(&dev));
  int mask = 0, maxlen = 0; CK(cuDeviceGetAttribute(&mask, CU_DEVICE_ATTRIBUTE_MEM_DECOMPRESS_ALGORITHM_MASK, dev)); CK(cuDeviceGetAttribute(&maxlen, CU_DEVICE_ATTRIBUTE_MEM_DECOMPRESS_MAXIMUM_LENGTH, dev));
  printf("decompress algorithm mask %d (deflate=1 snappy=2 lz4=4), max length %d\n", mask, maxlen);
  if (!(mask & 1)) { printf("DEFLATE not supported by the hardware engine on this device\n"); return 0; }
  FILE* f = fopen(argv[1], "rb"); if (!f) return 2; fseek(f, 0, SEEK_END); size_t n = ftell(f); fseek(f, 0, SEEK_SET);
  std::vector<unsigned char> h(n); if (fread(h.data(), 1, n, f) != n) return 2; fclose(f);
  struct B { size_t coff, clen, uoff, ulen; }; std::vector<B> bl; size_t off = 0, u = 0;
  while (off + 18 <= n) { size_t bs = (h[off + 16] | h[off + 17] << 8) + 1; unsigned isz; memcpy(&isz, &h[off + bs - 4], 4); if (isz) bl.push_back({off + 18, bs - 26, u, isz}); u += isz; off += bs; }
  printf("%zu blocks, %zu compressed, %zu inflated bytes\n", bl.size(), n, u);
  if (getenv("DE_CORRUPT") && bl.size() > 3) { for (size_t k = 2000; k < 2064 && k < bl[3].clen; k++) h[bl[3].coff + k] ^= 0x5A; printf("corrupted 64 payload bytes of block 3\n"); }
  unsigned char *dc, *du; unsigned* dact; cudaMalloc(&dc, n + 64); cudaMalloc(&du, u + 64); cudaMalloc(&dact, bl.size() * 4); cudaMemcpy(dc, h.data(), n, cudaMemcpyHostToDevice); cudaMemset(du, 0, u);
  std::vector<CUmemDecompressParams> P(bl.size());
  for (size_t i = 0; i < bl.size(); i++) { memset(&P[i], 0, sizeof P[i]); P[i].srcNumBytes = bl[i].clen; P[i].dstNumBytes = bl[i].ulen; P[i].dstActBytes = dact + i; P[i].src = dc + bl[i].coff; P[i].dst = du + bl[i].uoff; P[i].algo = CU_MEM_DECOMPRESS_ALGORITHM_DEFLATE; }
  CUmemDecompressParams* Pd = nullptr; if (params_on_device) { cudaMalloc(&Pd, P.size() * sizeof P[0]); cudaMemcpy(Pd, P.data(), P.size() * sizeof P[0], cudaMemcpyHostToDevice); }
  cudaStream_t s; cudaStreamCreate(&s); cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  // DE_STREAMS=k: batches go round-robin to k streams (does the engine overlap batches of different streams?)
  int ns = getenv("DE_STREAMS") ? atoi(getenv("DE_STREAMS")) : 1; if (ns < 1) ns = 1; std::vector<cudaStream_t> ss(ns); std::vector<cudaEvent_t> done(ns);
  for (int k = 0; k < ns; k++) { cudaStreamCreateWithFlags(&ss[k], cudaStreamNonBlocking); cudaEventCreateWithFlags(&done[k], cudaEventDisableTiming); }
  for (int rep = 0; rep < 3; rep++) {
    cudaEventRecord(e0, s); for (int k = 0; k < ns; k++) cudaStreamWaitEvent(ss[k], e0, 0);
    int k = 0;
    for (size_t i = 0; i < bl.size(); i += batch, k = (k + 1) % ns) { size_t c = std::min(batch, bl.size() - i), ei = 0; CUresult r = cuMemBatchDecompressAsync((params_on_device ? Pd : P.data()) + i, c, 0, &ei, (CUstream)ss[k]);
      if (r != CUDA_SUCCESS) { const char* es; cuGetErrorString(r, &es); printf("cuMemBatchDecompressAsync -> %d %s (index %zu)\n", (int)r, es ? es : "?", ei); return 1; } }
    for (int j = 0; j < ns; j++) { cudaEventRecord(done[j], ss[j]); cudaStreamWaitEvent(s, done[j], 0); }
    cudaEventRecord(e1, s); cudaError_t ce = cudaStreamSynchronize(s); if (ce != cudaSuccess) { printf("sync -> %s\n", cudaGetErrorString(ce)); return 1; }
    float ms; cudaEventElapsedTime(&ms, e0, e1); printf("streams %d batch %zu rep %d: %.2f ms, %.1f GB/s (compressed + inflated), %.1f GB/s inflated\n", ns, batch, rep, ms, (n + u) / ms / 1e6, u / ms / 1e6);
  }
  std::vector<unsigned char> out(u); cudaMemcpy(out.data(), du, u, cudaMemcpyDeviceToHost); std::vector<unsigned> act(bl.size()); cudaMemcpy(act.data(), dact, bl.size() * 4, cudaMemcpyDeviceToHost);
  if (getenv("DE_CORRUPT")) { printf("block 3: expected %zu bytes, engine reported %u; block 4: expected %zu, reported %u\n", bl[3].ulen, act[3], bl[4].ulen, act[4]); return 0; }
  size_t bad = 0; std::vector<unsigned char> ref(65536);
  for (size_t i = 0; i < bl.size(); i += std::max<size_t>(1, bl.size() / 500)) { z_stream z; memset(&z, 0, sizeof z); inflateInit2(&z, -15); z.next_in = &h[bl[i].coff]; z.avail_in = bl[i].clen; z.next_out = ref.data(); z.avail_out = 65536; inflate(&z, Z_FINISH); inflateEnd(&z);
    if (z.total_out != bl[i].ulen || act[i] != bl[i].ulen || memcmp(ref.data(), &out[bl[i].uoff], bl[i].ulen)) bad++; }
  printf("verified a sample of blocks against zlib: %zu mismatches\n", bad);
  return bad ? 1 : 0;
}
