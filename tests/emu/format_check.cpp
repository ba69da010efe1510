// Host check of the table writers in csrc/refs.cpp: every cell must print exactly as printf("%g") (= std::ostream with its
// default precision, which is what the reference's writers use), and a junction table built from shuffled, repeated
// counted entries must equal one built row by row.  Test infrastructure; built and run by tests/test_format_cpu.py.
#include "refs.h"
#include <cmath>
#include <cstdio>
#include <cstring>
#include <random>
#include <sstream>
using namespace irf;
static std::string want(double v) { char b[64]; int n = snprintf(b, 64, "%g", v); return std::string(b, n); }
int main(int argc, char** argv) {
  size_t N = argc > 1 ? (size_t)atol(argv[1]) : 2000000; std::mt19937_64 g(7); size_t bad = 0, n = 0;
  auto chk = [&](double v) { n++; std::string a = format_g6(v), b = want(v); if (a != b && bad++ < 20) printf("MISMATCH %.17g: got %s want %s\n", v, a.c_str(), b.c_str()); };
  const double specials[] = {0.0, 1.0, 0.5, 100000.5, 100001.5, 999999.5, 999999.4999, 1e6, 1e-5, 9.99999e-6, 9.999995e-5, 0.1, 0.2, 0.3, 1e15, 9.999995e14, 1e-4, 1e-310, 123456.5, 1234565, 12345650, 0.0001234565,
                            2.5, 1e22, 1e23, 5e-324, 1.7976931348623157e308, 0.999999499999, 0.9999995, 99999.95, 9.9999949999999999, 1234567.0, 12345675.0, 1.5e-5, 0.000099999949, INFINITY};
  for (double v : specials) { chk(v); chk(-v); }
  for (size_t i = 0; i < N / 4; i++) {   // exact ties at the sixth digit (round-half-even), scaled by powers of two so they stay exact
    unsigned long long d = 100000 + g() % 900000; double v = (double)(d * 10 + 5); int sc = (int)(g() % 5); for (int k = 0; k < sc; k++) v /= 2; chk(v); chk(v / 1024 / 1024); }
  for (size_t i = 0; i < N; i++) {   // ratios of integers as the tables produce them, and random bit patterns over the whole range
    double a = (double)(g() % 100000000), b = (double)(1 + g() % 100000); chk(a / b); chk((double)(g() % 1000) / (double)(1 + g() % 1000)); chk(a / (b * 1000.0) + 0.5);
    unsigned long long bits = g(); double v; memcpy(&v, &bits, 8); if (v == v) chk(v);
    chk(std::ldexp((double)(g() >> 11), -53 + (int)(g() % 80) - 20));
  }
  // junction table: annotation + counted entries in random order with repeated keys
  JunctionRef jr; { std::istringstream in("chrB\t100\t200\t+\nchrB\t100\t300\t-\nchrA\t5\t9\t+\nchrA\t5\t9\t-\nchrC\t1\t2\t+\n"); jr.loadRef(in); }
  std::vector<std::string> names = {"chrB", "chrA", "chrZ"};
  std::vector<JuncEntry> counted = {{0, 100, 300, {1, 2}}, {2, 7, 8, {0, 4}}, {0, 50, 60, {3, 0}}, {1, 5, 9, {10, 20}}, {0, 100, 300, {5, 5}}, {0, 100, 250, {0, 1}}, {5, 1, 2, {9, 9}}};
  std::string got = format_junctions(jr, names, counted);
  std::string exp = "JC_seqname\tstart\tend\tstrand\ttotal\tpos\tneg\nchrA\t5\t9\t.\t30\t20\t10\nchrB\t50\t60\t.\t3\t0\t3\nchrB\t100\t200\t+\t0\t0\t0\nchrB\t100\t250\t.\t1\t1\t0\nchrB\t100\t300\t-\t13\t7\t6\nchrC\t1\t2\t+\t0\t0\t0\nchrZ\t7\t8\t.\t4\t4\t0\n";
  if (got != exp) { bad++; printf("JUNCTION TABLE MISMATCH\n--- got\n%s--- want\n%s", got.c_str(), exp.c_str()); }
  printf("%zu cells checked, bad %zu\n", n, bad);
  return bad ? 1 : 0;
}
