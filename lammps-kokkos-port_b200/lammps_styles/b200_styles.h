// b200_styles.h — shared by the LAMMPS-side suffix styles (`-sf b200`) that bind libb200md.so.
// These files are compiled AGAINST the reference's headers (they subclass its style classes);
// they contain no reference code.  See INTEGRATION.md.
#ifndef LMP_B200_STYLES_H
#define LMP_B200_STYLES_H

#include "b200md.h"

namespace LAMMPS_NS {

// a pair style that can hand its coefficient tables to the device library
class B200PairStyle {
 public:
  virtual ~B200PairStyle() = default;
  virtual int b200_push(b200md_ctx *ctx) = 0;   // returns the library's status code
};

// a bond style that can hand its coefficients to the device library
class B200BondStyle {
 public:
  virtual ~B200BondStyle() = default;
  virtual int b200_push_bond(b200md_ctx *ctx) = 0;
};

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
