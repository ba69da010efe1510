// refs.h — host-side mirrors of the reference's processors' loadRef / ChrMapUpdate / WriteOutput
// (upstream ReadBlockProcessor.cpp, ReadBlockProcessor_CoverageBlocks.cpp).
// The per-record work (ProcessBlocks) lives on the device; these classes own the name-keyed reference
// data, bind it to BAM refIDs, and turn finished device tables into the reference's text files.
#pragma once
#include <cstdint>
#include <istream>
#include <functional>
#include <map>
#include <ostream>
#include <string>
#include <vector>
#include "common.cuh"

namespace irf {

// Flat, chromosome-name-ordered (std::map order == the reference's output order) reference tables.
struct SpansPointRef {
  std::map<std::string, Range> chr_range;   // into pos
  std::vector<u32> pos;                     // sorted within each chromosome
  void loadRef(std::istream& in);
  int lookup_index(const std::string& chr, u32 p) const;  // -1 when absent
};
struct JRow { u32 start, end, bits, c0, c1; };
struct JunctionRef {
  std::map<std::string, std::map<std::pair<u32, u32>, u32>> chr_junc;  // value: annotated strand bits (1 '-', 2 '+')
  std::map<std::string, std::vector<JRow>> flat;                        // the same, as (start,end)-sorted rows with zero counts
  void loadRef(std::istream& in);
  size_t size() const { size_t n = 0; for (auto& c : chr_junc) n += c.second.size(); return n; }
};
struct RoiRef {
  struct Roi { u32 start, end; std::string name; };
  bool present = false;
  std::map<std::string, Range> chr_range;
  std::vector<Roi> rois;                    // sorted by end within each chromosome
  void loadRef(std::istream& in);
};
struct CoverageRef {
  struct BEDrecord { std::string chrName, name; u32 start, end; bool direction; std::vector<std::pair<u32, u32>> blocks; };
  std::vector<BEDrecord> BEDrecords;        // file order == output order
  std::map<std::string, Range> chr_range;   // into seg_*
  std::vector<u32> seg_start, seg_end;      // merged measured segments, sorted, disjoint
  std::vector<u64> seg_coff;                // compact offset; each segment owns len+1 slots
  u64 P = 0;                                // total compact slots
  void loadRef(std::istream& in);
  // compact ranges of [s,e) intersected with the chromosome's measured segments
  void pieces(const std::string& chr, u32 s, u32 e, std::vector<IntronBlock>& out) const;
};

struct JuncEntry { int refid; u32 start, end; u32 cnt[2]; };

// text writers (byte layout per SURVEY.md B.7 / B.8).  Large tables come back as consecutive pieces (formatted by several
// threads) so the caller can write them without another copy.
// A table is prepared as parts (part 0 = header line) plus one task per remaining part; the tasks are independent of each
// other and of other tables' tasks, so a caller can pool the tasks of all its tables (run_tasks) and then write the parts.
// The TextTable must stay where it is until its tasks have run (they point into it), as must the inputs.
struct TextTable { std::vector<std::string> parts; std::vector<std::function<void()>> tasks; };
void prepare_ir_table(TextTable& t, const CoverageRef& cov, const std::vector<IntronRow>& rows);
void prepare_junctions(TextTable& t, const JunctionRef& ref, const std::vector<std::string>& chr_names, const std::vector<JuncEntry>& counted);
void prepare_spans(TextTable& t, const SpansPointRef& ref, const std::vector<u32>& count);
void run_tasks(std::vector<std::function<void()>>& tasks, unsigned threads);   // rethrows the first failure as std::string
unsigned format_threads();
std::string format_g6(double v);   // one cell as the tables print it (printf("%g")); exposed for tests
std::vector<std::string> format_ir_table_parts(const CoverageRef& cov, const std::vector<IntronRow>& rows);
std::vector<std::string> format_junctions_parts(const JunctionRef& ref, const std::vector<std::string>& chr_names, const std::vector<JuncEntry>& counted);
std::vector<std::string> format_spans_parts(const SpansPointRef& ref, const std::vector<u32>& count);
std::string format_ir_table(const CoverageRef& cov, const std::vector<IntronRow>& rows);
std::string format_junctions(const JunctionRef& ref, const std::vector<std::string>& chr_names, const std::vector<JuncEntry>& counted);
std::string format_spans(const SpansPointRef& ref, const std::vector<u32>& count /* [2][n] */);
std::string format_roi(const RoiRef& ref, const std::vector<u32>& count /* [2][n] */);
std::string format_chr(const std::vector<std::string>& chr_names, const std::vector<u32>& count /* [2][n_ref] */);
// JunctionCount::Directional; appends the evidence lines to `log`
int directional(const JunctionRef& ref, const std::vector<std::string>& chr_names, const std::vector<JuncEntry>& counted, std::string& log);

}  // namespace irf
