// b200_styles.h — shared by the LAMMPS-side suffix styles (`-sf b200`) that bind libb200md.so.
// These files are compiled AGAINST the reference's headers (they subclass its style classes);
// they contain no reference code.  See INTEGRATION.md.
#ifndef LMP_B200_STYLES_H
#define LMP_B200_STYLES_H

#include "b200md.h"

#include <cstddef>
#include <vector>

namespace LAMMPS_NS {

// a pair style that can hand its coefficient tables to the device library
class B200PairStyle {
 public:
  virtual ~B200PairStyle() = default;
  virtual int b200_push(b200md_ctx *ctx) = 0;   // returns the library's status code
  // Pair::eatom / Pair::vatom of the owned atoms for the flags of this step (Pair::ev_setup allocates
  // and clears them, src/pair.cpp:893-985); rows[k] = host index of the device's atom k
  virtual int b200_peratom(b200md_ctx *ctx, int eflag, int vflag, const int *rows, int n) = 0;
};

// a bond style that can hand its coefficients to the device library
class B200BondStyle {
 public:
  virtual ~B200BondStyle() = default;
  virtual int b200_push_bond(b200md_ctx *ctx) = 0;
  virtual int b200_peratom_bond(b200md_ctx *ctx, int eflag, int vflag, const int *rows, int n) = 0;   // Bond::eatom / vatom
};

// device per-atom tallies (b200md_peratom: ghost contributions already returned to the owners) ->
// the style's host arrays, device row k to host row rows[k]
inline int b200_fetch_peratom(b200md_ctx *ctx, int which, int eflag_atom, int vflag_atom, double *eatom,
                              double **vatom, const int *rows, int n)
{
  if (!eflag_atom && !vflag_atom) return 0;
  std::vector<double> e(eflag_atom ? n : 0), v(vflag_atom ? 6 * (std::size_t) n : 0);
  if (b200md_peratom(ctx, eflag_atom ? 1 : 0, vflag_atom ? 1 : 0, which, eflag_atom ? e.data() : nullptr,
                     vflag_atom ? v.data() : nullptr))
    return 1;
  for (int k = 0; k < n; k++) {
    const int i = rows[k];
    if (eflag_atom) eatom[i] = e[k];
    if (vflag_atom)
      for (int c = 0; c < 6; c++) vatom[i][c] = v[6 * (std::size_t) k + c];
  }
  return 0;
}

// a fix that acts inside the device step loop
class B200FixStyle {
 public:
  virtual ~B200FixStyle() = default;
  virtual int b200_push_fix(b200md_ctx *ctx) = 0;      // once the atoms are uploaded, before the set-up evaluation
  virtual int b200_begin_chunk(b200md_ctx *ctx) = 0;   // before every b200md_run of a run
};

// the one device context of this LAMMPS instance (owned by VerletB200)
struct B200Shared {
  static b200md_ctx *ctx;
};

}    // namespace LAMMPS_NS
#endif
