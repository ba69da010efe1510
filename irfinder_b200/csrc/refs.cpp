// refs.cpp — loadRef parsers and WriteOutput formatters (host side of the reference's processors).
#include "refs.h"

#include <algorithm>
#include <atomic>
#include <charconv>
#include <cstring>
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
// formatting: std::ostream << (default precision 6) == printf("%g") == to_chars(general, 6)
static inline void put_u(std::string& s, unsigned long long v) { char b[24]; auto r = std::to_chars(b, b + sizeof b, v); s.append(b, r.ptr); }
static inline void put_i(std::string& s, long long v) { char b[24]; auto r = std::to_chars(b, b + sizeof b, v); s.append(b, r.ptr); }
static inline void put_g(std::string& s, double v) {
  // %g prints integral values below 1e6 as plain integers; that covers most cells (0, 1, integer percentiles) without the general path
  if (v >= 0 && v < 1e6) { unsigned long long u = (unsigned long long)v; if ((double)u == v) { put_u(s, u); return; } }
  char b[40]; auto r = std::to_chars(b, b + sizeof b, v, std::chars_format::general, 6); s.append(b, r.ptr);
}

static const char* HDR_IR = "Chr\tStart\tEnd\tName\tNull\tStrand\tExcludedBases\tCoverage\tIntronDepth\tIntronDepth25Percentile\tIntronDepth50Percentile\tIntronDepth75Percentile\tExonToIntronReadsLeft\tExonToIntronReadsRight\tIntronDepthFirst50bp\tIntronDepthLast50bp\tSpliceLeft\tSpliceRight\tSpliceExact\tIRratio\tWarnings\n";
static const char* WARN[5] = {"-", "LowCover", "LowSplicing", "MinorIsoform", "NonUniformIntronCover"};

static void format_ir_rows(const CoverageRef& cov, const std::vector<IntronRow>& rows, size_t a, size_t b, std::string& s) {
  for (size_t i = a; i < b; i++) {
    const auto& rec = cov.BEDrecords[i]; const IntronRow& r = rows[i];
    s += rec.chrName; s += '\t'; put_u(s, rec.start); s += '\t'; put_u(s, rec.end); s += '\t'; s += rec.name; s += "\t0\t"; s += rec.direction ? '+' : '-'; s += '\t';
    put_i(s, r.excluded); s += '\t'; put_g(s, r.coverage); s += '\t'; put_g(s, r.depth); s += '\t'; put_g(s, r.p25); s += '\t'; put_g(s, r.p50); s += '\t'; put_g(s, r.p75); s += '\t';
    put_u(s, r.sp_left); s += '\t'; put_u(s, r.sp_right); s += '\t'; put_g(s, r.first50); s += '\t'; put_g(s, r.last50); s += '\t';
    put_u(s, r.jc_left); s += '\t'; put_u(s, r.jc_right); s += '\t'; put_u(s, r.jc_exact); s += '\t';
    if (r.depth == 0 && std::max(r.jc_left, r.jc_right) == 0) s += '0'; else put_g(s, r.irratio);
    s += '\t'; s += WARN[r.warning]; s += '\n';
  }
}
static unsigned fmt_threads(size_t n, unsigned cap = 6) { unsigned nt = std::max(1u, std::min(cap, std::thread::hardware_concurrency())); return n < 20000 ? 1 : nt; }
template <class F> static void run_parts(unsigned nt, F f) {   // f(t) for t in [0, nt), on nt threads
  if (nt == 1) { f(0); return; }
  std::vector<std::thread> th; for (unsigned t = 0; t < nt; t++) th.emplace_back([&, t]() { f(t); }); for (auto& t : th) t.join();
}
static std::string join_parts(const std::vector<std::string>& parts) { size_t tot = 0; for (auto& p : parts) tot += p.size(); std::string out; out.reserve(tot); for (auto& p : parts) out += p; return out; }

std::vector<std::string> format_ir_table_parts(const CoverageRef& cov, const std::vector<IntronRow>& rows) {
  size_t n = cov.BEDrecords.size(); unsigned nt = fmt_threads(n, 16);
  std::vector<std::string> parts(nt + 1); parts[0] = HDR_IR;
  run_parts(nt, [&](unsigned t) { size_t a = n * t / nt, b = n * (t + 1) / nt; parts[t + 1].reserve((b - a) * 160); format_ir_rows(cov, rows, a, b, parts[t + 1]); });
  return parts;
}
std::string format_ir_table(const CoverageRef& cov, const std::vector<IntronRow>& rows) { return join_parts(format_ir_table_parts(cov, rows)); }

// per-chromosome rows in (start,end) order: annotated junctions (possibly uncounted) merged with counted ones
static std::map<std::string, std::vector<JRow>> merged_junctions(const JunctionRef& ref, const std::vector<std::string>& chr_names, const std::vector<JuncEntry>& counted) {
  std::map<std::string, std::vector<JRow>> out = ref.flat;
  std::map<std::string, std::vector<JRow>> extra;
  for (auto& e : counted) { if (e.refid < 0 || e.refid >= (int)chr_names.size()) continue; extra[chr_names[e.refid]].push_back(JRow{e.start, e.end, 0, e.cnt[0], e.cnt[1]}); }
  for (auto& x : extra) {
    auto& v = x.second; std::sort(v.begin(), v.end(), [](const JRow& a, const JRow& b) { return a.start != b.start ? a.start < b.start : a.end < b.end; });
    auto& base = out[x.first]; std::vector<JRow> m; m.reserve(base.size() + v.size()); size_t i = 0, j = 0;
    while (i < base.size() || j < v.size()) {
      bool takeA = j >= v.size() || (i < base.size() && (base[i].start < v[j].start || (base[i].start == v[j].start && base[i].end <= v[j].end)));
      if (takeA) { JRow r = base[i++]; if (j < v.size() && v[j].start == r.start && v[j].end == r.end) { r.c0 += v[j].c0; r.c1 += v[j].c1; j++; } m.push_back(r); }
      else { if (!m.empty() && m.back().start == v[j].start && m.back().end == v[j].end) { m.back().c0 += v[j].c0; m.back().c1 += v[j].c1; } else m.push_back(v[j]); j++; }
    }
    base.swap(m);
  }
  return out;
}
static void format_junction_rows(const std::string& chr, const std::vector<JRow>& rows, std::string& s) {
  s.reserve(rows.size() * 40);
  for (auto& r : rows) {
    s += chr; s += '\t'; put_u(s, r.start); s += '\t'; put_u(s, r.end); s += '\t'; s += r.bits == 1 ? '-' : r.bits == 2 ? '+' : '.'; s += '\t';
    put_u(s, (unsigned long long)r.c0 + r.c1); s += '\t'; put_u(s, r.c1); s += '\t'; put_u(s, r.c0); s += '\n';
  }
}
std::vector<std::string> format_junctions_parts(const JunctionRef& ref, const std::vector<std::string>& chr_names, const std::vector<JuncEntry>& counted) {
  auto m = merged_junctions(ref, chr_names, counted);
  std::vector<std::pair<const std::string*, const std::vector<JRow>*>> chrs; for (auto& c : m) chrs.push_back({&c.first, &c.second});
  std::vector<std::string> parts(chrs.size() + 1); parts[0] = "JC_seqname\tstart\tend\tstrand\ttotal\tpos\tneg\n";
  size_t total = 0; for (auto& c : chrs) total += c.second->size();
  unsigned nt = fmt_threads(total); std::atomic<size_t> next(0);
  run_parts(nt, [&](unsigned) { for (;;) { size_t i = next.fetch_add(1); if (i >= chrs.size()) break; format_junction_rows(*chrs[i].first, *chrs[i].second, parts[i + 1]); } });
  return parts;
}
std::string format_junctions(const JunctionRef& ref, const std::vector<std::string>& chr_names, const std::vector<JuncEntry>& counted) { return join_parts(format_junctions_parts(ref, chr_names, counted)); }
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
std::vector<std::string> format_spans_parts(const SpansPointRef& ref, const std::vector<u32>& count) {
  size_t n = ref.pos.size(); unsigned nt = fmt_threads(n);
  std::vector<const std::string*> name_of(n); for (auto& c : ref.chr_range) for (u32 i = c.second.begin; i < c.second.end; i++) name_of[i] = &c.first;
  std::vector<std::string> parts(nt + 1); parts[0] = "SP_seqname\tcoord\ttotal\tpos\tneg\n";
  run_parts(nt, [&](unsigned t) { size_t a = n * t / nt, b = n * (t + 1) / nt; std::string& s = parts[t + 1]; s.reserve((b - a) * 32);
    for (size_t i = a; i < b; i++) { u32 neg = count[i], pos = count[n + i]; s += *name_of[i]; s += '\t'; put_u(s, ref.pos[i]); s += '\t'; put_u(s, (unsigned long long)neg + pos); s += '\t'; put_u(s, pos); s += '\t'; put_u(s, neg); s += '\n'; } });
  return parts;
}
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
