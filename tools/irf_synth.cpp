// irf_synth — seeded synthetic workload generator for the IRFinder BAM->IR counting path.
//
// Writes (a) a BuildRef-format annotation directory (ref-cover.bed, ref-sj.ref,
// ref-read-continues.ref, ref-ROI.bed; formats per SURVEY.md Appendix B.1) and (b) the same
// simulated RNA-seq alignments twice: `sorted.bam` (coordinate-sorted BGZF BAM, the input of the
// B200 path) and `collated.bam` (mates adjacent, first-in-pair first: the order the reference's
// BAM2blocks needs).  Wire formats follow the SAM/BAM specification v1 §4.
//
// SEQ/QUAL bytes are never read by the counting path but are filled with seeded random bases and
// skewed Phred values so DEFLATE sees realistic entropy.
//
// Determinism: the output bytes depend on (--seed, --chunks, all model options) and NOT on the
// number of worker threads: the genome is cut into `chunks` contiguous unit groups, each group is
// an independent BGZF stream produced from an RNG seeded by (seed, unit index).
//
// This is workload tooling (used by tests and bench.py); it is not part of the product library.
#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include <sys/stat.h>

using std::string;
using std::vector;
typedef uint64_t u64;
typedef uint32_t u32;

// ---------------------------------------------------------------- RNG
struct Rng {
  u64 s;
  explicit Rng(u64 seed) : s(seed * 0x9E3779B97F4A7C15ull + 0xD1B54A32D192ED03ull) { next(); next(); }
  inline u64 next() {  // splitmix64
    u64 z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  inline double uni() { return (next() >> 11) * (1.0 / 9007199254740992.0); }
  inline u64 below(u64 n) { return n ? (u64)(((unsigned __int128)next() * n) >> 64) : 0; }
  inline long range(long lo, long hi) { return lo + (long)below((u64)(hi - lo + 1)); }  // inclusive
  inline double normal() {
    double u1 = uni(), u2 = uni();
    if (u1 < 1e-300) u1 = 1e-300;
    return std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586 * u2);
  }
  inline double lognormal(double median, double sigma) { return median * std::exp(sigma * normal()); }
};

// ---------------------------------------------------------------- options
struct Opt {
  string out = "synth_out";
  u64 seed = 1;
  string genome = "chr22";   // chr22 | grch38
  long introns = 5000;       // target number of constitutive introns
  long pairs = 100000;       // fragments (pairs, or reads when single)
  int readlen = 100;
  bool single = false;
  int stranded = 0;          // 0 unstranded, -1 first mate antisense (dUTP), +1 first mate sense
  double intronic = 0.05;    // fraction of fragments drawn from unspliced pre-mRNA
  long hot_introns = 0;      // >0: intronic fragments are concentrated in the K longest introns
  int level = 1;             // zlib level
  int threads = 0;
  int chunks = 256;
  bool write_sorted = true, write_collated = true;
  bool align_records = false;  // htslib-style: never split a record across BGZF blocks
  bool refs_only = false;
  int chunk_lo = 0, chunk_hi = -1;  // --chunk-range A:B writes only the BGZF streams of chunks [A,B): byte range of the whole file
};

static void die(const char* m) { fprintf(stderr, "irf_synth: %s\n", m); exit(2); }

// ---------------------------------------------------------------- annotation model
struct Gene {
  int chr;
  long start, end;       // span
  bool plus;
  double weight;         // expression weight
  vector<std::pair<long, long>> exons;  // genomic [s,e), sorted
  int skip_from = -1;    // annotated exon-skipping isoform removes exon skip_from+1
  bool roi = false;      // pseudo-unit: intergenic region of interest (no introns)
  string name;
};
struct Intron {
  int chr; long start, end; bool plus; string name;
  vector<std::pair<long, long>> blocks;  // measured (non-excluded) sub-intervals
};

static const char* GRCH38_NAMES[25] = {"chr1","chr2","chr3","chr4","chr5","chr6","chr7","chr8","chr9","chr10",
  "chr11","chr12","chr13","chr14","chr15","chr16","chr17","chr18","chr19","chr20","chr21","chr22","chrX","chrY","chrM"};
static const long GRCH38_LENS[25] = {248956422,242193529,198295559,190214555,181538259,170805979,159345973,
  145138636,138394717,133797422,135086622,133275309,114364328,107043718,101991189,90338345,83257441,80373285,
  58617616,64444167,46709983,50818468,156040895,57227415,16569};

struct Annotation {
  vector<string> chr_names; vector<long> chr_lens;
  vector<Gene> units;      // genes and ROI pseudo-units, genome order
  vector<Intron> introns;  // ref-cover.bed records
};

static long clampl(long v, long lo, long hi) { return v < lo ? lo : (v > hi ? hi : v); }

static void build_annotation(const Opt& o, Annotation& A) {
  if (o.genome == "chr22") { A.chr_names = {"chr22"}; A.chr_lens = {50000000}; }
  else if (o.genome == "grch38") { for (int i = 0; i < 25; i++) { A.chr_names.push_back(GRCH38_NAMES[i]); A.chr_lens.push_back(GRCH38_LENS[i]); } }
  else if (o.genome == "tiny") { A.chr_names = {"chrA", "chrB", "chrEmpty"}; A.chr_lens = {400000, 300000, 50000}; }
  else die("unknown --genome (chr22|grch38|tiny)");
  Rng g(o.seed * 7919 + 17);
  double total_len = 0; for (size_t c = 0; c < A.chr_lens.size(); c++) if (A.chr_names[c] != "chrM" && A.chr_names[c] != "chrEmpty") total_len += A.chr_lens[c];
  long gene_id = 0;
  for (size_t c = 0; c < A.chr_lens.size(); c++) {
    if (A.chr_names[c] == "chrM" || A.chr_names[c] == "chrEmpty") continue;
    long want = std::max<long>(1, std::lround(o.introns * (A.chr_lens[c] / total_len)));
    long made = 0; long cursor = 10000; long L = A.chr_lens[c];
    // expected gene span ~ 10 introns * ~4 kb + exons; spread genes evenly with random gaps
    long ngenes = std::max<long>(1, want / 9);
    long gap_mean = std::max<long>(2000, (L - 20000) / ngenes / 3);
    while (made < want && cursor < L - 20000) {
      Gene ge; ge.chr = (int)c; ge.plus = g.next() & 1; ge.weight = g.lognormal(1.0, 2.0);
      int nex = 2 + (int)g.below(17);  // 2..18 exons
      if ((long)nex - 1 > want - made) nex = (int)(want - made) + 1;
      long p = cursor + (long)g.below(gap_mean * 2 + 1);
      if (p >= L - 20000) break;
      ge.start = p;
      bool ok = true;
      for (int e = 0; e < nex; e++) {
        long el = clampl(std::lround(g.lognormal(130, 0.6)), 30, 3000);
        if (e == 0 || e == nex - 1) el = clampl(el * 3, 100, 5000);  // UTR exons longer
        ge.exons.push_back({p, p + el}); p += el;
        if (e + 1 < nex) { long il = clampl(std::lround(g.lognormal(1500, 1.2)), 70, 200000); p += il; }
        if (p >= L - 5000) { ok = false; break; }
      }
      if (!ok) { if (ge.exons.size() < 2) break; /* truncate gene */ while (ge.exons.back().second >= L - 5000) ge.exons.pop_back(); if (ge.exons.size() < 2) break; }
      ge.end = ge.exons.back().second;
      if (ge.exons.size() >= 4 && g.uni() < 0.25) ge.skip_from = (int)g.below(ge.exons.size() - 2);
      char nm[64]; snprintf(nm, sizeof nm, "GENE%06ld/ENSG%011ld", gene_id, gene_id); ge.name = nm; gene_id++;
      made += (long)ge.exons.size() - 1;
      cursor = ge.end + 500;
      A.units.push_back(ge);
      // an intergenic ROI pseudo-unit after ~8 % of genes
      if (g.uni() < 0.08 && cursor + 6000 < L - 20000) {
        Gene r; r.chr = (int)c; r.plus = true; r.roi = true; r.weight = g.lognormal(0.3, 1.0);
        r.start = cursor + 500; r.end = r.start + g.range(800, 4000); r.exons.push_back({r.start, r.end});
        char rn[64]; snprintf(rn, sizeof rn, "ROI%06ld/%s", gene_id, (g.next() & 1) ? "rRNA" : "NonPolyA"); r.name = rn;
        cursor = r.end + 500; A.units.push_back(r);
      }
    }
  }
  // introns + measured blocks
  for (auto& ge : A.units) {
    if (ge.roi) continue;
    auto add_intron = [&](long s, long e, const char* cls, long hole_s, long hole_e) {
      Intron in; in.chr = ge.chr; in.start = s; in.end = e; in.plus = ge.plus; in.name = ge.name + "/" + cls;
      long ms = s + 5, me = e - 5;  // 5 bp margin inside each splice site
      vector<std::pair<long, long>> ex;  // exclusions, sorted
      if (hole_s >= 0) ex.push_back({hole_s, hole_e});
      if (me - ms > 200 && g.uni() < 0.3) {
        int k = 1 + (int)g.below(4);
        for (int i = 0; i < k; i++) { long len = g.range(20, std::max<long>(21, (me - ms) / 12)); long xs = g.range(ms + 1, me - len - 1); ex.push_back({xs, xs + len}); }
        std::sort(ex.begin(), ex.end());
      }
      long cur = ms;
      for (auto& x : ex) { if (x.first > cur) in.blocks.push_back({cur, std::min(x.first, me)}); cur = std::max(cur, x.second); if (cur >= me) break; }
      if (cur < me) in.blocks.push_back({cur, me});
      if (in.blocks.empty()) in.blocks.push_back({ms, std::min(ms + 10, me)});
      A.introns.push_back(in);
    };
    for (size_t e = 0; e + 1 < ge.exons.size(); e++) add_intron(ge.exons[e].second, ge.exons[e + 1].first, "clean", -1, -1);
    if (ge.skip_from >= 0) {
      size_t e = ge.skip_from;
      add_intron(ge.exons[e].second, ge.exons[e + 2].first, "known-exon", ge.exons[e + 1].first - 10, ge.exons[e + 1].second + 10);
    }
  }
  std::stable_sort(A.introns.begin(), A.introns.end(), [](const Intron& a, const Intron& b) {
    if (a.chr != b.chr) return a.chr < b.chr; if (a.start != b.start) return a.start < b.start; return a.end < b.end; });
}

static void write_refs(const Opt& o, const Annotation& A) {
  string d = o.out + "/ref"; mkdir(d.c_str(), 0777); d += "/IRFinder"; mkdir(d.c_str(), 0777);
  FILE* f = fopen((d + "/ref-cover.bed").c_str(), "w"); if (!f) die("cannot write ref-cover.bed");
  for (auto& in : A.introns) {
    fprintf(f, "%s\t%ld\t%ld\t%s\t0\t%c\t%ld\t%ld\t255,0,0\t%zu\t", A.chr_names[in.chr].c_str(), in.start, in.end, in.name.c_str(),
            in.plus ? '+' : '-', in.start, in.end, in.blocks.size());
    for (size_t i = 0; i < in.blocks.size(); i++) fprintf(f, "%s%ld", i ? "," : "", in.blocks[i].second - in.blocks[i].first);
    fputc('\t', f);
    for (size_t i = 0; i < in.blocks.size(); i++) fprintf(f, "%s%ld", i ? "," : "", in.blocks[i].first - in.start);
    fputc('\n', f);
  }
  fclose(f);
  // ref-sj.ref: unique (chr,start,end); ref-read-continues.ref: unique (chr,pos)
  f = fopen((d + "/ref-sj.ref").c_str(), "w");
  FILE* f2 = fopen((d + "/ref-read-continues.ref").c_str(), "w");
  vector<std::tuple<int, long, char>> pts;
  for (size_t i = 0; i < A.introns.size(); i++) {
    auto& in = A.introns[i];
    if (i && A.introns[i - 1].chr == in.chr && A.introns[i - 1].start == in.start && A.introns[i - 1].end == in.end) continue;
    fprintf(f, "%s\t%ld\t%ld\t%c\n", A.chr_names[in.chr].c_str(), in.start, in.end, in.plus ? '+' : '-');
    pts.push_back({in.chr, in.start, in.plus ? '+' : '-'}); pts.push_back({in.chr, in.end, in.plus ? '+' : '-'});
  }
  std::sort(pts.begin(), pts.end());
  pts.erase(std::unique(pts.begin(), pts.end(), [](auto& a, auto& b) { return std::get<0>(a) == std::get<0>(b) && std::get<1>(a) == std::get<1>(b); }), pts.end());
  for (auto& p : pts) fprintf(f2, "%s\t%ld\t%c\n", A.chr_names[std::get<0>(p)].c_str(), std::get<1>(p), std::get<2>(p));
  fclose(f); fclose(f2);
  f = fopen((d + "/ref-ROI.bed").c_str(), "w");
  for (auto& u : A.units) if (u.roi) fprintf(f, "%s\t%ld\t%ld\t%s\n", A.chr_names[u.chr].c_str(), u.start, u.end, u.name.c_str());
  fclose(f);
}

// ---------------------------------------------------------------- BGZF writer
struct Bgzf {
  vector<uint8_t> out;     // compressed stream
  uint8_t buf[65280]; int n = 0; int level; bool align;
  u64 inflated = 0; z_stream zs;
  Bgzf(int lvl, bool al) : level(lvl), align(al) {}
  void flush_block() {
    if (!n) return;
    uint8_t cbuf[70000];
    memset(&zs, 0, sizeof zs);
    if (deflateInit2(&zs, level, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) die("deflateInit2");
    zs.next_in = buf; zs.avail_in = n; zs.next_out = cbuf; zs.avail_out = sizeof cbuf;
    if (deflate(&zs, Z_FINISH) != Z_STREAM_END) die("deflate");
    int clen = (int)zs.total_out; deflateEnd(&zs);
    int bsize = clen + 25;  // total block size - 1
    uint8_t h[18] = {31, 139, 8, 4, 0, 0, 0, 0, 0, 255, 6, 0, 'B', 'C', 2, 0, (uint8_t)(bsize & 255), (uint8_t)(bsize >> 8)};
    out.insert(out.end(), h, h + 18); out.insert(out.end(), cbuf, cbuf + clen);
    u32 crc = (u32)crc32(crc32(0, nullptr, 0), buf, n); u32 tail[2] = {crc, (u32)n};
    out.insert(out.end(), (uint8_t*)tail, (uint8_t*)tail + 8);
    inflated += n; n = 0;
  }
  void write(const uint8_t* p, int len) {
    if (align && n && n + len > (int)sizeof buf) flush_block();
    while (len > 0) {
      int k = std::min<int>(len, (int)sizeof buf - n);
      memcpy(buf + n, p, k); n += k; p += k; len -= k;
      if (n == (int)sizeof buf) flush_block();
    }
  }
};
static const uint8_t BGZF_EOF[28] = {0x1f,0x8b,8,4,0,0,0,0,0,0xff,6,0,0x42,0x43,2,0,0x1b,0,3,0,0,0,0,0,0,0,0,0};

// ---------------------------------------------------------------- read model
struct Rec {            // one alignment, packed later
  long pos; u32 order;  // order = emission order in the collated stream (stable tie-break)
  vector<uint8_t> bytes;
};
static int reg2bin(long beg, long end) {
  --end;
  if (beg >> 14 == end >> 14) return (int)(((1 << 15) - 1) / 7 + (beg >> 14));
  if (beg >> 17 == end >> 17) return (int)(((1 << 12) - 1) / 7 + (beg >> 17));
  if (beg >> 20 == end >> 20) return (int)(((1 << 9) - 1) / 7 + (beg >> 20));
  if (beg >> 23 == end >> 23) return (int)(((1 << 6) - 1) / 7 + (beg >> 23));
  if (beg >> 26 == end >> 26) return (int)(((1 << 3) - 1) / 7 + (beg >> 26));
  return 0;
}
struct Aln { long pos; long end; vector<u32> cigar; };  // cigar: len<<4|op

// Map transcript interval [t0,t0+len) through exons into an alignment (M and N ops).
static Aln map_read(const vector<std::pair<long, long>>& ex, long t0, long len) {
  Aln a; a.pos = -1; long t = 0; long remaining = len; long last_end = -1;
  for (size_t i = 0; i < ex.size() && remaining > 0; i++) {
    long el = ex[i].second - ex[i].first;
    if (t0 < t + el) {
      long off = std::max<long>(0, t0 - t); long take = std::min(el - off, remaining);
      long gs = ex[i].first + off;
      if (a.pos < 0) a.pos = gs; else a.cigar.push_back((u32)((gs - last_end) << 4 | 3));
      a.cigar.push_back((u32)(take << 4 | 0)); last_end = gs + take; remaining -= take; t0 += take;
    }
    t += el;
  }
  a.end = last_end; return a;
}

struct Worker {
  const Opt& o; const Annotation& A; const vector<long>& unit_frags; const vector<u64>& unit_first_id;
  Worker(const Opt& o_, const Annotation& A_, const vector<long>& uf, const vector<u64>& ui) : o(o_), A(A_), unit_frags(uf), unit_first_id(ui) {}
  u64 n_records = 0, n_pairs = 0, n_singles = 0, n_unmapped = 0;

  void pack(Rec& r, int refid, const Aln& a, int flag, int mrefid, long mpos, long tlen, const string& name, int lseq, Rng& g) {
    int ncig = (int)a.cigar.size(); int lname = (int)name.size() + 1;
    int bs = 32 + lname + 4 * ncig + (lseq + 1) / 2 + lseq + 16;
    r.bytes.resize(4 + bs); uint8_t* p = r.bytes.data();
    auto w32 = [&](int v) { memcpy(p, &v, 4); p += 4; };
    w32(bs); w32(refid); w32((int)a.pos);
    long end = a.end > a.pos ? a.end : a.pos + 1;
    u32 bmq = ((u32)reg2bin(a.pos < 0 ? 0 : a.pos, a.pos < 0 ? 1 : end) << 16) | ((flag & 4 ? 0u : 255u) << 8) | (u32)lname; w32((int)bmq);
    u32 fn = ((u32)flag << 16) | (u32)ncig; w32((int)fn);
    w32(lseq); w32(mrefid); w32((int)mpos); w32((int)tlen);
    memcpy(p, name.c_str(), lname); p += lname;
    for (u32 c : a.cigar) { memcpy(p, &c, 4); p += 4; }
    static const uint8_t nib[4] = {1, 2, 4, 8};
    for (int i = 0; i < (lseq + 1) / 2; i++) { u64 x = g.next(); *p++ = (uint8_t)(nib[x & 3] << 4 | nib[(x >> 2) & 3]); }
    for (int i = 0; i < lseq; i++) { u64 x = g.next(); *p++ = (uint8_t)((x & 15) < 11 ? 37 : 2 + ((x >> 4) % 39)); }
    const uint8_t tags[16] = {'N','H','C',1,'H','I','C',1,'A','S','C',(uint8_t)(lseq > 127 ? 250 : lseq * 2 - 2),'n','M','C',0};
    memcpy(p, tags, 16);
  }

  // apply small edits to an alignment whose read length is `L` (soft clip / insertion / deletion)
  void edit(Aln& a, int L, Rng& g) {
    double u = g.uni();
    if (u < 0.01) {  // soft clip
      int c = (int)g.range(1, 10);
      if ((g.next() & 1) && (long)(a.cigar[0] >> 4) > c + 8) {  // left clip: shift start
        a.cigar[0] = ((a.cigar[0] >> 4) - c) << 4; a.pos += c; a.cigar.insert(a.cigar.begin(), (u32)(c << 4 | 4));
      } else if ((long)(a.cigar.back() >> 4) > c + 8) {
        a.cigar.back() = ((a.cigar.back() >> 4) - c) << 4; a.end -= c; a.cigar.push_back((u32)(c << 4 | 4));
      }
    } else if (u < 0.015) {  // insertion in last M: aM 2I bM, consumes 2 fewer reference bases
      long m = a.cigar.back() >> 4;
      if (m > 24) { long x = g.range(8, m - 12); a.cigar.back() = (u32)(x << 4); a.cigar.push_back(2 << 4 | 1); a.cigar.push_back((u32)((m - x - 2) << 4)); a.end -= 2; }
    } else if (u < 0.02) {   // deletion in last M: aM 3D bM, consumes 3 more reference bases
      long m = a.cigar.back() >> 4;
      if (m > 24) { long x = g.range(8, m - 12); a.cigar.back() = (u32)(x << 4); a.cigar.push_back(3 << 4 | 2); a.cigar.push_back((u32)((m - x) << 4)); a.end += 3; }
    }
    (void)L;
  }

  // Generate all records of one unit; `sorted` receives them coordinate-ordered, `coll` in pair order.
  void gen_unit(size_t ui, Bgzf* sorted, Bgzf* coll, vector<const vector<Intron>*>*) {
    const Gene& ge = A.units[ui]; long nf = unit_frags[ui]; if (!nf) return;
    Rng g(o.seed * 1000003ull + ui * 7ull + 1);
    vector<Rec> recs; recs.reserve(nf * 2);
    u64 fid = unit_first_id[ui]; u32 order = 0;
    vector<std::pair<long, long>> whole = {{ge.start, ge.end}};
    // hot introns inside this gene (for --hot-introns): precomputed flag via name suffix is avoided; use long introns
    for (long f = 0; f < nf; f++, fid++) {
      char nmb[32]; snprintf(nmb, sizeof nmb, "r%09llu", (unsigned long long)fid); string name = nmb;
      const vector<std::pair<long, long>>* ex = &ge.exons; vector<std::pair<long, long>> tmp;
      bool unspliced = ge.roi || g.uni() < o.intronic;
      if (unspliced && !ge.roi && o.hot_introns > 0) {
        // concentrate in this gene's longest intron
        long best = -1, bl = 0; for (size_t e = 0; e + 1 < ge.exons.size(); e++) { long l = ge.exons[e + 1].first - ge.exons[e].second; if (l > bl) { bl = l; best = (long)e; } }
        if (best >= 0 && bl > 3 * o.readlen) { tmp = {{ge.exons[best].second, ge.exons[best + 1].first}}; ex = &tmp; }
        else ex = &whole;
      } else if (unspliced) ex = &whole;
      else {
        double u = g.uni();
        if (ge.skip_from >= 0 && u < 0.3) { tmp = ge.exons; tmp.erase(tmp.begin() + ge.skip_from + 1); ex = &tmp; }
        else if (u > 0.98 && ge.exons.size() >= 2) {  // novel junction: shift one donor site into the intron
          tmp = ge.exons; size_t j = g.below(tmp.size() - 1); long k = g.range(1, 10);
          if (tmp[j + 1].first - tmp[j].second > k + 20) tmp[j].second += k; ex = &tmp;
        }
      }
      long TL = 0; for (auto& e : *ex) TL += e.second - e.first;
      int L = o.readlen;
      if (o.single) {
        if (TL < 20) continue;
        long len = std::min<long>(L, TL); long t0 = g.range(0, TL - len);
        Aln a = map_read(*ex, t0, len); edit(a, (int)len, g);
        bool rev = g.next() & 1;
        if (o.stranded) rev = (o.stranded < 0) ? ge.plus : !ge.plus;
        Rec r; r.order = order++; r.pos = a.pos; pack(r, ge.chr, a, rev ? 0x10 : 0, -1, -1, 0, name, (int)len, g);
        recs.push_back(std::move(r)); n_singles++; continue;
      }
      long I;
      if (g.uni() < 0.02) I = g.range(40, L); else I = std::max<long>(L, std::lround(300 + 50 * g.normal()));
      if (I > TL) I = TL; if (I < 20) continue;
      long len = std::min<long>(L, I); long f0 = g.range(0, TL - I);
      Aln a1 = map_read(*ex, f0, len), a2 = map_read(*ex, f0 + I - len, len);
      edit(a1, (int)len, g); edit(a2, (int)len, g);
      if (a2.pos < a1.pos) std::swap(a1, a2);
      // left read forward, right read reverse (FR). Which one is first-in-pair?
      bool first_is_left = g.next() & 1;
      if (o.stranded) { bool first_rev = (o.stranded < 0) ? ge.plus : !ge.plus; first_is_left = !first_rev; }
      int fl = 0x1 | 0x2 | 0x20 | (first_is_left ? 0x40 : 0x80);
      int fr = 0x1 | 0x2 | 0x10 | (first_is_left ? 0x80 : 0x40);
      long tlen = std::max(a1.end, a2.end) - a1.pos;
      double uo = g.uni();
      Rec rl, rr; rl.pos = a1.pos; rr.pos = a2.pos;
      if (uo < 0.005) {          // orphan: right mate unmapped and present in the file
        fl = (fl & ~0x22) | 0x8; Aln un; un.pos = a1.pos; un.end = a1.pos;
        pack(rl, ge.chr, a1, fl, ge.chr, a1.pos, 0, name, (int)len, g);
        int fu = 0x1 | 0x4 | (first_is_left ? 0x80 : 0x40); rr.pos = a1.pos;
        pack(rr, ge.chr, un, fu, ge.chr, a1.pos, 0, name, (int)len, g); n_unmapped++; n_singles++;
      } else if (uo < 0.01) {    // orphan: mate absent from the file
        fl = (fl & ~0x22) | 0x8; pack(rl, ge.chr, a1, fl, -1, -1, 0, name, (int)len, g);
        rl.order = order++; recs.push_back(std::move(rl)); n_singles++; continue;
      } else {
        pack(rl, ge.chr, a1, fl, ge.chr, a2.pos, tlen, name, (int)len, g);
        pack(rr, ge.chr, a2, fr, ge.chr, a1.pos, -tlen, name, (int)len, g); n_pairs++;
      }
      // collated order: first-in-pair first
      if (first_is_left) { rl.order = order++; rr.order = order++; } else { rr.order = order++; rl.order = order++; }
      recs.push_back(std::move(rl)); recs.push_back(std::move(rr));
    }
    n_records += recs.size();
    if (coll) {
      vector<u32> idx(recs.size()); for (u32 i = 0; i < idx.size(); i++) idx[recs[i].order] = i;
      for (u32 i : idx) coll->write(recs[i].bytes.data(), (int)recs[i].bytes.size());
    }
    if (sorted) {
      vector<u32> idx(recs.size()); for (u32 i = 0; i < idx.size(); i++) idx[i] = i;
      std::stable_sort(idx.begin(), idx.end(), [&](u32 x, u32 y) { if (recs[x].pos != recs[y].pos) return recs[x].pos < recs[y].pos; return recs[x].order < recs[y].order; });
      for (u32 i : idx) sorted->write(recs[i].bytes.data(), (int)recs[i].bytes.size());
    }
  }
};

static vector<uint8_t> bam_header(const Annotation& A, bool sorted) {
  string text = sorted ? "@HD\tVN:1.6\tSO:coordinate\n" : "@HD\tVN:1.6\tSO:unsorted\tGO:query\n";
  for (size_t i = 0; i < A.chr_names.size(); i++) text += "@SQ\tSN:" + A.chr_names[i] + "\tLN:" + std::to_string(A.chr_lens[i]) + "\n";
  text += "@PG\tID:irf_synth\tPN:irf_synth\n";
  vector<uint8_t> h; auto w32 = [&](int v) { h.insert(h.end(), (uint8_t*)&v, (uint8_t*)&v + 4); };
  h.insert(h.end(), {'B', 'A', 'M', 1}); w32((int)text.size()); h.insert(h.end(), text.begin(), text.end());
  w32((int)A.chr_names.size());
  for (size_t i = 0; i < A.chr_names.size(); i++) { w32((int)A.chr_names[i].size() + 1); h.insert(h.end(), A.chr_names[i].begin(), A.chr_names[i].end()); h.push_back(0); w32((int)A.chr_lens[i]); }
  return h;
}

int main(int argc, char** argv) {
  Opt o;
  for (int i = 1; i < argc; i++) {
    string a = argv[i]; auto val = [&]() { if (i + 1 >= argc) die("missing value"); return string(argv[++i]); };
    if (a == "--out") o.out = val(); else if (a == "--seed") o.seed = std::stoull(val());
    else if (a == "--genome") o.genome = val(); else if (a == "--introns") o.introns = std::stol(val());
    else if (a == "--pairs" || a == "--reads") o.pairs = std::stol(val()); else if (a == "--readlen") o.readlen = std::stoi(val());
    else if (a == "--single") o.single = true; else if (a == "--stranded") o.stranded = std::stoi(val());
    else if (a == "--intronic") o.intronic = std::stod(val()); else if (a == "--hot-introns") o.hot_introns = std::stol(val());
    else if (a == "--level") o.level = std::stoi(val()); else if (a == "--threads") o.threads = std::stoi(val());
    else if (a == "--chunks") o.chunks = std::stoi(val()); else if (a == "--no-sorted") o.write_sorted = false;
    else if (a == "--no-collated") o.write_collated = false; else if (a == "--align-records") o.align_records = true;
    else if (a == "--refs-only") o.refs_only = true;
    else if (a == "--chunk-range") { string v = val(); size_t c = v.find(':'); if (c == string::npos) die("--chunk-range A:B"); o.chunk_lo = std::stoi(v.substr(0, c)); o.chunk_hi = std::stoi(v.substr(c + 1)); }
    else { fprintf(stderr, "unknown option %s\n", a.c_str()); return 2; }
  }
  if (o.threads <= 0) o.threads = std::max(1u, std::thread::hardware_concurrency());
  mkdir(o.out.c_str(), 0777);
  Annotation A; build_annotation(o, A); write_refs(o, A);
  if (o.refs_only) return 0;

  // fragments per unit: proportional to weight * transcript length (largest-remainder rounding)
  size_t U = A.units.size(); vector<double> w(U); double wt = 0;
  for (size_t i = 0; i < U; i++) { long tl = 0; for (auto& e : A.units[i].exons) tl += e.second - e.first; w[i] = A.units[i].weight * tl; wt += w[i]; }
  vector<long> nf(U); long assigned = 0; vector<std::pair<double, size_t>> rem(U);
  for (size_t i = 0; i < U; i++) { double x = o.pairs * (w[i] / wt); nf[i] = (long)x; assigned += nf[i]; rem[i] = {-(x - nf[i]), i}; }
  std::sort(rem.begin(), rem.end()); for (long k = 0; k < o.pairs - assigned; k++) nf[rem[k % U].second]++;
  vector<u64> first_id(U); u64 acc = 0; for (size_t i = 0; i < U; i++) { first_id[i] = acc; acc += nf[i]; }

  // chunk boundaries: contiguous unit ranges with ~equal fragment counts
  int C = std::max(1, std::min<int>(o.chunks, (int)U)); vector<size_t> cb(C + 1, U); cb[0] = 0;
  { u64 run = 0; int c = 1; for (size_t i = 0; i < U && c < C; i++) { run += nf[i]; if (run >= (u64)o.pairs * c / C) cb[c++] = i + 1; } }
  for (int c = 1; c <= C; c++) if (cb[c] < cb[c - 1]) cb[c] = cb[c - 1];
  cb[C] = U;

  int clo = std::max(0, o.chunk_lo), chi = o.chunk_hi < 0 ? C : std::min(C, o.chunk_hi);
  vector<Bgzf*> S(C, nullptr), K(C, nullptr); std::atomic<int> next(clo);
  vector<Worker> workers; for (int t = 0; t < o.threads; t++) workers.emplace_back(o, A, nf, first_id);
  vector<std::thread> th;
  for (int t = 0; t < o.threads; t++) th.emplace_back([&, t]() {
    for (;;) { int c = next.fetch_add(1); if (c >= chi) break;
      if (o.write_sorted) S[c] = new Bgzf(o.level, o.align_records); if (o.write_collated) K[c] = new Bgzf(o.level, o.align_records);
      for (size_t u = cb[c]; u < cb[c + 1]; u++) workers[t].gen_unit(u, S[c], K[c], nullptr);
      if (S[c]) S[c]->flush_block(); if (K[c]) K[c]->flush_block(); }
  });
  for (auto& t : th) t.join();
  u64 nrec = 0, npairs = 0, nsing = 0, nunm = 0; for (auto& wk : workers) { nrec += wk.n_records; npairs += wk.n_pairs; nsing += wk.n_singles; nunm += wk.n_unmapped; }

  u64 infl_sorted = 0, comp_sorted = 0;
  for (int pass = 0; pass < 2; pass++) {
    bool sorted = pass == 0; if (sorted ? !o.write_sorted : !o.write_collated) continue;
    string path = o.out + (sorted ? "/sorted.bam" : "/collated.bam"); FILE* f = fopen(path.c_str(), "wb"); if (!f) die("cannot open bam for write");
    Bgzf hb(o.level, false); auto h = bam_header(A, sorted); hb.write(h.data(), (int)h.size()); hb.flush_block();
    u64 infl = 0, comp = 0;
    if (clo == 0) { fwrite(hb.out.data(), 1, hb.out.size(), f); infl = hb.inflated; comp = hb.out.size(); }
    else { FILE* hf = fopen((path + ".header").c_str(), "wb"); if (hf) { fwrite(hb.out.data(), 1, hb.out.size(), hf); fclose(hf); } }  // a later part: header kept beside it
    for (int c = 0; c < C; c++) { Bgzf* b = sorted ? S[c] : K[c]; if (!b) continue; fwrite(b->out.data(), 1, b->out.size(), f); infl += b->inflated; comp += b->out.size(); delete b; }
    if (chi == C) { fwrite(BGZF_EOF, 1, 28, f); comp += 28; }
    fclose(f);
    if (sorted || !o.write_sorted) { infl_sorted = infl; comp_sorted = comp; }
  }
  FILE* m = fopen((o.out + "/meta.json").c_str(), "w");
  fprintf(m, "{\"seed\": %llu, \"genome\": \"%s\", \"introns\": %zu, \"units\": %zu, \"fragments\": %llu, \"records\": %llu, \"pairs\": %llu, "
             "\"singles\": %llu, \"unmapped\": %llu, \"readlen\": %d, \"single_end\": %s, \"stranded\": %d, \"inflated_bytes\": %llu, \"compressed_bytes\": %llu, \"level\": %d, \"chunks\": %d, \"chunk_lo\": %d, \"chunk_hi\": %d}\n",
          (unsigned long long)o.seed, o.genome.c_str(), A.introns.size(), U, (unsigned long long)acc, (unsigned long long)nrec, (unsigned long long)npairs,
          (unsigned long long)nsing, (unsigned long long)nunm, o.readlen, o.single ? "true" : "false", o.stranded, (unsigned long long)infl_sorted,
          (unsigned long long)comp_sorted, o.level, C, clo, chi);
  fclose(m);
  return 0;
}
