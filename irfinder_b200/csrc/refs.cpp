// refs.cpp — loadRef parsers and WriteOutput formatters (host side of the reference's processors).
#include "refs.h"

#include <algorithm>
#include <atomic>
#include <charconv>
#include <cmath>
#include <cstring>
#include <memory>
#include <sstream>
#include <thread>

namespace irf {

static std::vector<std::string> split_tabs(const std::string& line) {
  std::vector<std::string> f; size_t p = 0;
  for (;;) { size_t q = line.find('\t', p); if (q == std::string::npos) { f.push_back(line.substr(p)); break; } f.push_back(line.substr(p, q - p)); p = q + 1; }
  if (!f.empty() && !f.back().empty() && f.back().back() == '\r') f.back().pop_back();
  return f;
}

// Reference files are written by BuildRef (Perl); a malformed line is reported with its file and line number instead of
// being skipped, because a silently dropped intron or junction changes every downstream table.
static u32 parse_u32(const std::string& f, const char* file, size_t line, const char* what) {
  if (f.empty() || f.size() > 10) throw std::string(file) + " line " + std::to_string(line) + ": bad " + what + " '" + f + "'";
  unsigned long long v = 0; for (char ch : f) { if (ch < '0' || ch > '9') throw std::string(file) + " line " + std::to_string(line) + ": bad " + what + " '" + f + "'"; v = v * 10 + (unsigned)(ch - '0'); }
  if (v > 0xffffffffull) throw std::string(file) + " line " + std::to_string(line) + ": " + what + " out of range";
  return (u32)v;
}
static void bad_line(const char* file, size_t line, const std::string& why) { throw std::string(file) + " line " + std::to_string(line) + ": " + why; }

// SpansPoint::loadRef — "chr pos strand"; the strand column is not used at lookup time [C-9]
void SpansPointRef::loadRef(std::istream& in) {
  std::map<std::string, std::vector<u32>> by_chr; std::string line; size_t ln = 0;
  while (std::getline(in, line)) { ln++; if (line.empty()) continue; auto f = split_tabs(line); if (f.size() < 2) bad_line("ref-read-continues.ref", ln, "expected 'chr<TAB>pos[<TAB>strand]'");
    by_chr[f[0]].push_back(parse_u32(f[1], "ref-read-continues.ref", ln, "position")); }
  pos.clear(); chr_range.clear();
  for (auto& c : by_chr) { std::sort(c.second.begin(), c.second.end()); Range r; r.begin = (u32)pos.size(); pos.insert(pos.end(), c.second.begin(), c.second.end()); r.end = (u32)pos.size(); chr_range[c.first] = r; }
}
int SpansPointRef::lookup_index(const std::string& chr, u32 p) const {
  auto c = chr_range.find(chr); if (c == chr_range.end()) return -1;
  auto b = pos.begin() + c->second.begin, e = pos.begin() + c->second.end; auto it = std::lower_bound(b, e, p);
  return (it == e || *it != p) ? -1 : (int)(it - pos.begin());
}

// JunctionCount::loadRef — "chr start end strand"
void JunctionRef::loadRef(std::istream& in) {
  std::string line; size_t ln = 0;
  while (std::getline(in, line)) {
    ln++; if (line.empty()) continue; auto f = split_tabs(line); if (f.size() < 4) bad_line("ref-sj.ref", ln, "expected 'chr<TAB>start<TAB>end<TAB>strand'");
    u32 js = parse_u32(f[1], "ref-sj.ref", ln, "start"), je = parse_u32(f[2], "ref-sj.ref", ln, "end"); if (je < js) bad_line("ref-sj.ref", ln, "junction end before start");
    u32& bits = chr_junc[f[0]][std::make_pair(js, je)];
    if (f[3] == "-") bits |= 1; else if (f[3] == "+") bits |= 2;
  }
  flat.clear();
  for (auto& c : chr_junc) { auto& v = flat[c.first]; v.reserve(c.second.size()); for (auto& j : c.second) v.push_back(JRow{j.first.first, j.first.second, j.second, 0, 0}); }
}

// FragmentsInROI::loadRef — "chr start end name", kept sorted by end
void RoiRef::loadRef(std::istream& in) {
  std::map<std::string, std::vector<Roi>> by_chr; std::string line; size_t ln = 0;
  while (std::getline(in, line)) { ln++; if (line.empty()) continue; auto f = split_tabs(line); if (f.size() < 4) bad_line("ref-ROI.bed", ln, "expected 'chr<TAB>start<TAB>end<TAB>name'");
    Roi r{parse_u32(f[1], "ref-ROI.bed", ln, "start"), parse_u32(f[2], "ref-ROI.bed", ln, "end"), f[3]}; if (r.end < r.start) bad_line("ref-ROI.bed", ln, "end before start"); by_chr[f[0]].push_back(r); }
  rois.clear(); chr_range.clear(); present = true;
  for (auto& c : by_chr) {
    std::stable_sort(c.second.begin(), c.second.end(), [](const Roi& a, const Roi& b) { return a.end < b.end; });
    Range r; r.begin = (u32)rois.size(); rois.insert(rois.end(), c.second.begin(), c.second.end()); r.end = (u32)rois.size(); chr_range[c.first] = r;
  }
}

// CoverageBlocks::loadRef — BED12; measured segments are the union of every record's blocks
void CoverageRef::loadRef(std::istream& in) {
  std::map<std::string, std::vector<std::pair<u32, u32>>> raw; std::string line; size_t ln = 0; const char* F = "ref-cover.bed";
  BEDrecords.clear();
  while (std::getline(in, line)) {
    ln++; if (line.empty()) continue; auto f = split_tabs(line); if (f.size() < 12) bad_line(F, ln, "expected 12 BED columns, found " + std::to_string(f.size()));
    BEDrecord r; r.chrName = f[0]; r.start = parse_u32(f[1], F, ln, "start"); r.end = parse_u32(f[2], F, ln, "end"); r.name = f[3]; r.direction = (f[5] == "+");
    if (r.end < r.start) bad_line(F, ln, "end before start");
    u32 nblocks = parse_u32(f[9], F, ln, "blockCount");
    std::istringstream ls(f[10]), os(f[11]); std::string a, b;
    while (std::getline(ls, a, ',') && std::getline(os, b, ',')) {
      if (a.empty() || b.empty()) break; u32 len = parse_u32(a, F, ln, "block size"), off = parse_u32(b, F, ln, "block start");
      if ((u64)r.start + off + len > r.end) bad_line(F, ln, "block reaches beyond the record's end");
      if (!r.blocks.empty() && r.start + off < r.blocks.back().second) bad_line(F, ln, "blocks overlap or are not in ascending order");
      r.blocks.push_back({r.start + off, r.start + off + len}); raw[r.chrName].push_back(r.blocks.back());
    }
    if (r.blocks.size() != nblocks) bad_line(F, ln, "blockCount says " + std::to_string(nblocks) + " but " + std::to_string(r.blocks.size()) + " blocks are listed");
    BEDrecords.push_back(std::move(r));
  }
  seg_start.clear(); seg_end.clear(); seg_coff.clear(); chr_range.clear(); P = 0;
  for (auto& c : raw) {
    std::sort(c.second.begin(), c.second.end()); Range rg; rg.begin = (u32)seg_start.size();
    u32 cs = 0, ce = 0; bool open = false;
    auto close = [&]() { seg_start.push_back(cs); seg_end.push_back(ce); seg_coff.push_back(P); P += (u64)(ce - cs) + 1; };
    for (auto& b : c.second) { if (b.first >= b.second) continue; if (open && b.first <= ce) ce = std::max(ce, b.second); else { if (open) close(); cs = b.first; ce = b.second; open = true; } }
    if (open) close();
    rg.end = (u32)seg_start.size(); chr_range[c.first] = rg;
  }
}
void CoverageRef::pieces(const std::string& chr, u32 s, u32 e, std::vector<IntronBlock>& out) const {
  auto c = chr_range.find(chr); if (c == chr_range.end() || s >= e) return;
  u32 lo = c->second.begin, hi = c->second.end;
  u32 i = (u32)(std::upper_bound(seg_end.begin() + lo, seg_end.begin() + hi, s) - seg_end.begin());  // first segment with end > s
  for (; i < hi && seg_start[i] < e; i++) { u32 a = std::max(s, seg_start[i]), b = std::min(e, seg_end[i]); if (a < b) out.push_back(IntronBlock{seg_coff[i] + (a - seg_start[i]), b - a, 0}); }
}

// ---------------------------------------------------------------------------------------------
// formatting: std::ostream << (default precision 6) == printf("%g") == to_chars(general, 6).
// All writers work on raw char pointers into buffers sized from a per-row bound; tables are cut into independent parts so
// that one pool of threads can format every table of a run at once (TextTable in refs.h).
static inline char* put_u(char* p, unsigned long long v) { return std::to_chars(p, p + 24, v).ptr; }
static inline char* put_i(char* p, long long v) { return std::to_chars(p, p + 24, v).ptr; }
static inline void put_u(std::string& s, unsigned long long v) { char b[24]; s.append(b, put_u(b, v)); }
static inline char* put_s(char* p, const std::string& s) { memcpy(p, s.data(), s.size()); return p + s.size(); }

// %g with 6 significant digits for finite v in [1e-5, 1e15), exactly as printf rounds (round-half-even on the exact binary
// value): v = m * 2^e is scaled by a power of ten in 128-bit integer arithmetic, so quotient and remainder are exact.
// Everything outside that range (and inf/nan) goes to std::to_chars, which is slower (~5x) but rare in these tables.
static const unsigned long long P10[] = {1ull, 10ull, 100ull, 1000ull, 10000ull, 100000ull, 1000000ull, 10000000ull, 100000000ull, 1000000000ull, 10000000000ull, 100000000000ull};
static const double D10[] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15};
static const double DM10[] = {1e0, 1e-1, 1e-2, 1e-3, 1e-4, 1e-5};
static bool six_digits(unsigned long long m, int e, int X, unsigned long long& D) {   // D = round(v / 10^(X-5)); false when out of the supported shifts
  typedef unsigned __int128 u128; int k = 5 - X;
  if (e >= 0) return false;
  int sh = -e; u128 q, r, half_or_den;
  if (k >= 0) { if (k > 11 || sh > 120) return false; u128 N = (u128)m * P10[k]; q = N >> sh; r = N & (((u128)1 << sh) - 1); half_or_den = (u128)1 << (sh - 1);
    if (r > half_or_den || (r == half_or_den && (q & 1))) q++; }
  else { int j = -k; if (j > 11 || sh > 60) return false; u128 den = (u128)P10[j] << sh; q = (u128)m / den; r = (u128)m % den;
    if (2 * r > den || (2 * r == den && (q & 1))) q++; }
  D = (unsigned long long)q; return true;
}
static char* put_g_general(char* p, double v) { return std::to_chars(p, p + 32, v, std::chars_format::general, 6).ptr; }
static inline char* put_g(char* p, double v) {
  // integral values below 1e6 print as plain integers; that covers most cells (0, 1, integer percentiles)
  if (v >= 0 && v < 1e6) { unsigned long long u = (unsigned long long)v; if ((double)u == v && !(u == 0 && std::signbit(v))) return put_u(p, u); }
  double a = v < 0 ? -v : v;
  if (!(a >= 1e-5 && a < 1e15)) return put_g_general(p, v);
  unsigned long long b; memcpy(&b, &a, 8); int be = (int)(b >> 52) & 0x7ff; if (be == 0 || be == 0x7ff) return put_g_general(p, v);
  unsigned long long m = (b & ((1ull << 52) - 1)) | (1ull << 52); int e = be - 1075;
  int X; if (a >= 1.0) { X = 0; while (X < 14 && a >= D10[X + 1]) X++; } else { X = -1; while (X > -5 && a < DM10[-X]) X--; }
  unsigned long long D = 0;
  for (int tries = 0;; tries++) {
    if (tries > 2 || !six_digits(m, e, X, D)) return put_g_general(p, v);
    if (D < 100000ull) { X--; continue; }
    if (D > 1000000ull) { X++; continue; }
    if (D == 1000000ull) { D = 100000ull; X++; }
    break;
  }
  if (v < 0) *p++ = '-';
  char d[6]; { unsigned long long t = D; for (int i = 5; i >= 0; i--) { d[i] = (char)('0' + t % 10); t /= 10; } }
  int nd = 6; while (nd > 1 && d[nd - 1] == '0') nd--;
  if (X < -4 || X >= 6) {   // d.ddddde+XX
    *p++ = d[0]; if (nd > 1) { *p++ = '.'; for (int i = 1; i < nd; i++) *p++ = d[i]; }
    *p++ = 'e'; int ax = X; if (ax < 0) { *p++ = '-'; ax = -ax; } else *p++ = '+';
    if (ax < 10) *p++ = '0'; return put_u(p, (unsigned long long)ax);
  }
  if (X >= 0) { for (int i = 0; i <= X; i++) *p++ = d[i]; if (nd > X + 1) { *p++ = '.'; for (int i = X + 1; i < nd; i++) *p++ = d[i]; } return p; }
  *p++ = '0'; *p++ = '.'; for (int i = 0; i < -X - 1; i++) *p++ = '0'; for (int i = 0; i < nd; i++) *p++ = d[i];
  return p;
}
std::string format_g6(double v) { char b[48]; return std::string(b, put_g(b, v)); }   // for the unit test of the formatter

static const char* HDR_IR = "Chr\tStart\tEnd\tName\tNull\tStrand\tExcludedBases\tCoverage\tIntronDepth\tIntronDepth25Percentile\tIntronDepth50Percentile\tIntronDepth75Percentile\tExonToIntronReadsLeft\tExonToIntronReadsRight\tIntronDepthFirst50bp\tIntronDepthLast50bp\tSpliceLeft\tSpliceRight\tSpliceExact\tIRratio\tWarnings\n";
static const char* WARN[5] = {"-", "LowCover", "LowSplicing", "MinorIsoform", "NonUniformIntronCover"};

// a part is formatted into the thread's scratch buffer (sized from an upper bound of its rows, reused from part to part) and
// then copied into the part's string at its exact size; a TextTable that is reused keeps its strings' capacity, so a
// long-lived caller does not fault fresh pages in on every run
template <class F> static void fill_part(std::string& s, size_t bound, F f) {
  thread_local std::vector<char> scratch; if (scratch.size() < bound) scratch.resize(bound + bound / 4);
  char* b = scratch.data(); char* e = f(b); s.assign(b, (size_t)(e - b));
}

static void format_ir_rows(const CoverageRef& cov, const std::vector<IntronRow>& rows, size_t a, size_t b, std::string& s) {
  size_t bound = 0; for (size_t i = a; i < b; i++) bound += cov.BEDrecords[i].chrName.size() + cov.BEDrecords[i].name.size() + 360;   // 3 x 20 + 8 x 24 + 5 x 10 + tabs + warning
  fill_part(s, bound, [&](char* p) {
    for (size_t i = a; i < b; i++) {
      const auto& rec = cov.BEDrecords[i]; const IntronRow& r = rows[i];
      p = put_s(p, rec.chrName); *p++ = '\t'; p = put_u(p, rec.start); *p++ = '\t'; p = put_u(p, rec.end); *p++ = '\t'; p = put_s(p, rec.name); *p++ = '\t'; *p++ = '0'; *p++ = '\t'; *p++ = rec.direction ? '+' : '-'; *p++ = '\t';
      p = put_i(p, r.excluded); *p++ = '\t'; p = put_g(p, r.coverage); *p++ = '\t'; p = put_g(p, r.depth); *p++ = '\t'; p = put_g(p, r.p25); *p++ = '\t'; p = put_g(p, r.p50); *p++ = '\t'; p = put_g(p, r.p75); *p++ = '\t';
      p = put_u(p, r.sp_left); *p++ = '\t'; p = put_u(p, r.sp_right); *p++ = '\t'; p = put_g(p, r.first50); *p++ = '\t'; p = put_g(p, r.last50); *p++ = '\t';
      p = put_u(p, r.jc_left); *p++ = '\t'; p = put_u(p, r.jc_right); *p++ = '\t'; p = put_u(p, r.jc_exact); *p++ = '\t';
      if (r.depth == 0 && std::max(r.jc_left, r.jc_right) == 0) *p++ = '0'; else p = put_g(p, r.irratio);
      *p++ = '\t'; const char* w = WARN[r.warning]; size_t wl = strlen(w); memcpy(p, w, wl); p += wl; *p++ = '\n';
    }
    return p; });
}

void run_tasks(std::vector<std::function<void()>>& tasks, unsigned threads) {
  if (tasks.empty()) return;
  threads = std::max(1u, std::min<unsigned>(threads, (unsigned)tasks.size()));
  if (threads == 1) { for (auto& t : tasks) t(); return; }
  std::atomic<size_t> next(0); std::vector<std::string> errs(threads); std::vector<std::thread> th;
  for (unsigned k = 0; k < threads; k++) th.emplace_back([&, k]() { try { for (;;) { size_t i = next.fetch_add(1); if (i >= tasks.size()) break; tasks[i](); } } catch (const std::string& e) { errs[k] = e; } catch (const std::exception& e) { errs[k] = e.what(); } });
  for (auto& t : th) t.join();
  for (auto& e : errs) if (!e.empty()) throw e;
}
unsigned format_threads() { return std::max(1u, std::min(16u, std::thread::hardware_concurrency())); }
static std::string join_parts(const std::vector<std::string>& parts) { size_t tot = 0; for (auto& p : parts) tot += p.size(); std::string out; out.reserve(tot); for (auto& p : parts) out += p; return out; }
static unsigned parts_for(size_t rows, size_t rows_per_part) { return (unsigned)std::max<size_t>(1, std::min<size_t>(64, rows / rows_per_part)); }

void prepare_ir_table(TextTable& t, const CoverageRef& cov, const std::vector<IntronRow>& rows) {
  size_t n = cov.BEDrecords.size(); unsigned np = parts_for(n, 4000);
  t.tasks.clear(); t.parts.resize(np + 1); t.parts[0] = HDR_IR;
  for (unsigned k = 0; k < np; k++) t.tasks.push_back([&t, &cov, &rows, n, np, k]() { format_ir_rows(cov, rows, n * k / np, n * (k + 1) / np, t.parts[k + 1]); });
}

// per-chromosome rows in (start,end) order: annotated junctions (possibly uncounted) merged with counted ones
static void merge_and_format_junctions(const std::string& chr, const std::vector<JRow>* base, std::vector<JRow>& v, std::string& s) {
  auto less = [](const JRow& a, const JRow& b) { return a.start != b.start ? a.start < b.start : a.end < b.end; };
  if (!std::is_sorted(v.begin(), v.end(), less)) std::sort(v.begin(), v.end(), less);
  static const std::vector<JRow> none; const std::vector<JRow>& B = base ? *base : none;
  fill_part(s, (B.size() + v.size()) * (chr.size() + 90), [&](char* p) {
    auto emit = [&](const JRow& r) { p = put_s(p, chr); *p++ = '\t'; p = put_u(p, r.start); *p++ = '\t'; p = put_u(p, r.end); *p++ = '\t'; *p++ = r.bits == 1 ? '-' : r.bits == 2 ? '+' : '.'; *p++ = '\t';
      p = put_u(p, (unsigned long long)r.c0 + r.c1); *p++ = '\t'; p = put_u(p, r.c1); *p++ = '\t'; p = put_u(p, r.c0); *p++ = '\n'; };
    size_t i = 0, j = 0; JRow cur{}; bool have = false;   // same-key rows (annotation + counts, or repeated counted keys) collapse into one
    auto push = [&](const JRow& r) { if (have && cur.start == r.start && cur.end == r.end) { cur.c0 += r.c0; cur.c1 += r.c1; cur.bits |= r.bits; } else { if (have) emit(cur); cur = r; have = true; } };
    while (i < B.size() || j < v.size()) {
      bool takeA = j >= v.size() || (i < B.size() && (B[i].start < v[j].start || (B[i].start == v[j].start && B[i].end <= v[j].end)));
      if (takeA) push(B[i++]); else push(v[j++]);
    }
    if (have) emit(cur);
    return p; });
}
void prepare_junctions(TextTable& t, const JunctionRef& ref, const std::vector<std::string>& chr_names, const std::vector<JuncEntry>& counted) {
  auto extra = std::make_shared<std::map<std::string, std::vector<JRow>>>();
  { const std::string* last_name = nullptr; std::vector<JRow>* last = nullptr; int last_id = -1;   // counted entries arrive grouped by refid
    for (auto& e : counted) { if (e.refid < 0 || e.refid >= (int)chr_names.size()) continue;
      if (e.refid != last_id) { last_id = e.refid; last_name = &chr_names[e.refid]; last = &(*extra)[*last_name]; }
      last->push_back(JRow{e.start, e.end, 0, e.cnt[0], e.cnt[1]}); } }
  std::map<std::string, std::pair<const std::vector<JRow>*, std::vector<JRow>*>> chrs;
  for (auto& c : ref.flat) chrs[c.first].first = &c.second;
  for (auto& c : *extra) chrs[c.first].second = &c.second;
  t.tasks.clear(); t.parts.resize(chrs.size() + 1); t.parts[0] = "JC_seqname\tstart\tend\tstrand\ttotal\tpos\tneg\n";
  auto none = std::make_shared<std::vector<JRow>>();
  size_t k = 1; for (auto& c : chrs) { const std::string name = c.first; auto base = c.second.first; auto v = c.second.second; size_t idx = k++;
    t.tasks.push_back([&t, name, base, v, idx, extra, none]() { std::vector<JRow> empty; merge_and_format_junctions(name, base, v ? *v : empty, t.parts[idx]); }); }
}
int directional(const JunctionRef& ref, const std::vector<std::string>& chr_names, const std::vector<JuncEntry>& counted, std::string& log) {
  // only counted junctions can pass the evidence threshold, so walk those and look the annotation up (no merged map needed)
  unsigned long long same = 0, diff = 0, ev = 0, nev = 0, evk = 0, nevk = 0;
  std::vector<const std::vector<JRow>*> by_id(chr_names.size(), nullptr);
  for (size_t i = 0; i < chr_names.size(); i++) { auto it = ref.flat.find(chr_names[i]); if (it != ref.flat.end()) by_id[i] = &it->second; }
  // `counted` is ordered by (refid, start, end) and so is every chromosome's annotation: one forward cursor per chromosome
  int cur_id = -1; size_t cur = 0;
  for (auto& e : counted) {  // thresholds [C-11]
    u32 c0 = e.cnt[0], c1 = e.cnt[1]; if (c0 + c1 <= 8) continue;
    u32 bits = 0;
    if (e.refid >= 0 && e.refid < (int)by_id.size() && by_id[e.refid]) { auto& v = *by_id[e.refid];
      if (e.refid != cur_id) { cur_id = e.refid; cur = 0; }
      while (cur < v.size() && (v[cur].start < e.start || (v[cur].start == e.start && v[cur].end < e.end))) cur++;
      if (cur < v.size() && v[cur].start == e.start && v[cur].end == e.end) bits = v[cur].bits; }
    if (c0 > c1 * 4) { ev++; if (bits) evk++; if (bits == 1) same++; else if (bits == 2) diff++; }
    else if (c1 > c0 * 4) { ev++; if (bits) evk++; if (bits == 2) same++; else if (bits == 1) diff++; }
    else { nev++; if (bits) nevk++; }
  }
  auto line = [&](const char* k, unsigned long long v) { log += k; put_u(log, v); log += '\n'; };
  line("Directionality: Dir evidence:\t", ev); line("Directionality: Nondir evidence:\t", nev);
  line("Directionality: Dir evidence known junctions:\t", evk); line("Directionality: Nondir evidence known junctions:\t", nevk);
  line("Directionality: Dir matches ref:\t", same); line("Directionality: Dir opposed to ref:\t", diff);
  line("Directionality: Dir score all (0-10000):\t", ev * 10000 / (ev + nev + 1)); line("Directionality: Dir score known junctions (0-10000):\t", evk * 10000 / (evk + nevk + 1));
  if (same > diff * 100 && evk > nevk * 100) return 1;
  if (diff > same * 100 && evk > nevk * 100) return -1;
  return 0;
}
void prepare_spans(TextTable& t, const SpansPointRef& ref, const std::vector<u32>& count) {
  size_t n = ref.pos.size(); unsigned np = parts_for(n, 20000);
  auto name_of = std::make_shared<std::vector<const std::string*>>(n); for (auto& c : ref.chr_range) for (u32 i = c.second.begin; i < c.second.end; i++) (*name_of)[i] = &c.first;
  size_t longest = 0; for (auto& c : ref.chr_range) longest = std::max(longest, c.first.size());
  t.tasks.clear(); t.parts.resize(np + 1); t.parts[0] = "SP_seqname\tcoord\ttotal\tpos\tneg\n";
  for (unsigned k = 0; k < np; k++) t.tasks.push_back([&t, &ref, &count, name_of, n, np, k, longest]() {
    size_t a = n * k / np, b = n * (k + 1) / np;
    fill_part(t.parts[k + 1], (b - a) * (longest + 60), [&](char* p) {
      for (size_t i = a; i < b; i++) { u32 neg = count[i], pos = count[n + i]; p = put_s(p, *(*name_of)[i]); *p++ = '\t'; p = put_u(p, ref.pos[i]); *p++ = '\t'; p = put_u(p, (unsigned long long)neg + pos); *p++ = '\t'; p = put_u(p, pos); *p++ = '\t'; p = put_u(p, neg); *p++ = '\n'; }
      return p; }); });
}
static std::vector<std::string> finish(TextTable& t) { run_tasks(t.tasks, format_threads()); return std::move(t.parts); }
std::vector<std::string> format_ir_table_parts(const CoverageRef& cov, const std::vector<IntronRow>& rows) { TextTable t; prepare_ir_table(t, cov, rows); return finish(t); }
std::vector<std::string> format_junctions_parts(const JunctionRef& ref, const std::vector<std::string>& chr_names, const std::vector<JuncEntry>& counted) { TextTable t; prepare_junctions(t, ref, chr_names, counted); return finish(t); }
std::vector<std::string> format_spans_parts(const SpansPointRef& ref, const std::vector<u32>& count) { TextTable t; prepare_spans(t, ref, count); return finish(t); }
std::string format_ir_table(const CoverageRef& cov, const std::vector<IntronRow>& rows) { return join_parts(format_ir_table_parts(cov, rows)); }
std::string format_junctions(const JunctionRef& ref, const std::vector<std::string>& chr_names, const std::vector<JuncEntry>& counted) { return join_parts(format_junctions_parts(ref, chr_names, counted)); }
std::string format_spans(const SpansPointRef& ref, const std::vector<u32>& count) { return join_parts(format_spans_parts(ref, count)); }
std::string format_roi(const RoiRef& ref, const std::vector<u32>& count) {
  std::string s = "ROI_Name\ttotal\tpos\tneg\n"; size_t n = ref.rois.size();
  for (auto& c : ref.chr_range) for (u32 i = c.second.begin; i < c.second.end; i++) {
    u32 neg = count[i], pos = count[n + i]; s += ref.rois[i].name; s += '\t'; put_u(s, (unsigned long long)neg + pos); s += '\t'; put_u(s, pos); s += '\t'; put_u(s, neg); s += '\n'; }
  return s;
}
std::string format_chr(const std::vector<std::string>& chr_names, const std::vector<u32>& count) {
  std::map<std::string, std::pair<unsigned long long, unsigned long long>> m; size_t n = chr_names.size();
  for (size_t i = 0; i < n; i++) { auto& e = m[chr_names[i]]; e.first += count[i]; e.second += count[n + i]; }
  std::string s = "ChrCoverage_RName\ttotal\tpos\tneg\n";
  for (auto& c : m) { s += c.first; s += '\t'; put_u(s, c.second.first + c.second.second); s += '\t'; put_u(s, c.second.second); s += '\t'; put_u(s, c.second.first); s += '\n'; }
  return s;
}

}  // namespace irf
