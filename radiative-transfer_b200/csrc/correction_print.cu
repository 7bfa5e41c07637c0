// correction_print.cu -- the text report of the reference's Correction::Print() (correction.cpp:404-603) from the DEVICE's
// per-group tables: energy grid, Planck integrals, opacities, edge opacities, the energy differences and cor1..3.
//
// No driver of the reference calls Print(); it is the class's debugging view and the drop-in replaces
// correction.cpp, so it is offered here: rt_correction_print() formats into a caller-owned buffer, host/solver.cpp and
// the Python mirror put it on a stream.  Same titles, column widths, `left << setw` padding and default ostream number
// format (6 significant digits) as the reference, so the two texts can be compared token by token
// (tests/golden/correction_print_*.txt come from the unmodified reference, tests/test_correction_print.py).
//
// Three columns print storage the reference never fills: ukappa, ckappa and emis_spec are resized in the constructor
// (correction.cpp:306-308) and only generate_multigroup_opacities would write them, which is commented out
// (correction.cpp:376) -- with real Eigen the reference prints indeterminate memory there.  This report prints 0, which is
// also what the reference built on the zero-filling Eigen stand-in of oracle/ prints.
#include <cmath>
#include <cstring>
#include <iomanip>
#include <sstream>
#include <string>
#include <vector>

#include "rt_internal.h"

namespace
{
struct Column
{
   const char *title, *rule;
   int width;           // of the column in the rule line and in the rows
   int title_width;     // of the title (the reference's last column is sometimes padded wider in the title line only)
   const double *value; // NULL: the row index
};

// one table: title line, rule line, `rows` rows, then the reference's `"\n" << endl`
void table(std::ostream &o, const std::vector<Column> &cols, int rows)
{
   for (const Column &c : cols) o << std::left << std::setw(c.title_width) << c.title;
   o << '\n';
   for (const Column &c : cols) o << std::left << std::setw(c.width) << c.rule;
   o << '\n';
   for (int r = 0; r < rows; ++r)
   {
      for (const Column &c : cols)
      {
         o << std::left << std::setw(c.width);
         if (c.value) o << c.value[r];
         else o << r;
      }
      o << '\n';
   }
   o << "\n\n";
}
} // namespace

extern "C" int rt_correction_print(rt_solver *s, char *text, size_t capacity, size_t *length)
{
   if (!s) return rt_internal_fail("rt_correction_print: null handle");
   rt_print_tables t;
   if (rt_internal_print_tables(s, &t)) return 1;
   const int G = t.G;
   std::vector<double> e_ave(G), de_ave(G), zero(G, 0.0);
   for (int g = 0; g < G; ++g) // generate_group_averages, solver.cpp:22-32
   {
      e_ave[g] = 0.5 * (t.e_edge[g] + t.e_edge[g + 1]);
      de_ave[g] = t.e_edge[g + 1] - t.e_edge[g];
   }
   std::ostringstream o; // default format: general, precision 6 -- what cout has in the reference
   o << "========== Correction::Print() ==========\n"
     << "Contants:\n" // sic
     << "\tT: " << t.T << '\n'
     << "\tkappa_grey: " << t.kappa_grey << '\n'
     << "\tG: " << G << '\n';
   const double *E = e_ave.data();
   table(o, {{"Group Index", "-----------", 13, 13, nullptr}, {"Average Energy", "(keV)---------", 16, 16, E},
             {"Upper Energy", "(keV)-------", 14, 14, t.e_edge.data() + 1}, {"Group Width", "(keV)------", 13, 13, de_ave.data()}}, G);
   table(o, {{"Group", "-----", 7, 7, nullptr}, {"EG Min", "(keV)-", 14, 14, t.e_edge.data()}, {"EG Max", "(keV)-", 14, 14, t.e_edge.data() + 1},
             {"B", "(jk/cm^2-sh)", 14, 14, t.B.data()}, {"dBdT", "(jk/cm^2-sh-keV)", 18, 22, t.dBdT.data()}}, G);
   table(o, {{"Group", "-----", 7, 7, nullptr}, {"Energy", "(keV)-", 14, 14, E},
             {"Planck-Averaged Opacity", "(cm^2/g)---------------", 25, 25, zero.data()},
             {"Group-Centered Opacity", "(cm^2/g)--------------", 24, 24, zero.data()}}, G);
   table(o, {{"Group", "-----", 7, 7, nullptr}, {"Energy", "(keV)-", 14, 14, E}, {"Final Opacity", "(cm^2/g)-----", 15, 15, t.kappa.data()},
             {"Emission Spec", "(jk-g-sh)----", 15, 15, zero.data()}}, G);
   table(o, {{"Group", "-----", 7, 7, nullptr}, {"Energy", "------", 14, 14, E}, {"dEB", "-----", 14, 14, t.dEB.data()}}, G);
   table(o, {{"Edge Index", "---------", 12, 12, nullptr}, {"Energy", "------", 14, 14, t.e_edge.data()},
             {"Opacity", "-------", 14, 14, t.kappa_edge.data()}}, G + 1);
   double sum = 0.0, sumabs = 0.0; // the zero-sum check of the differences (correction.cpp:545-556)
   for (int g = 0; g < G; ++g)
   {
      sum += t.dkapEB[g];
      sumabs += std::fabs(t.dkapEB[g]);
   }
   o << "Sum dkapEB/Sum |dkapEB| = " << sum / sumabs << "\n\n";
   table(o, {{"Group", "-----", 7, 7, nullptr}, {"Energy", "------", 14, 14, E}, {"dsigEdE", "-------", 14, 14, t.dsigEdE.data()}}, G);
   table(o, {{"Group", "-----", 7, 7, nullptr}, {"Energy", "------", 14, 14, E}, {"cor1", "----", 14, 14, t.cor1.data()},
             {"cor2", "----", 14, 14, t.cor2.data()}, {"cor3", "----", 14, 14, t.cor3.data()}}, G);
   const std::string out = o.str();
   if (length) *length = out.size();
   if (text && capacity > 0)
   {
      const size_t n = out.size() < capacity - 1 ? out.size() : capacity - 1;
      std::memcpy(text, out.data(), n);
      text[n] = '\0';
      if (n < out.size()) return rt_internal_fail("rt_correction_print: buffer too small (call with text = NULL for the length)");
   }
   return 0;
}
