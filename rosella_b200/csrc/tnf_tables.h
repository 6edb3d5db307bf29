// tnf_tables.h — lookup tables shared by the TNF host packer and the counting kernels.
#pragma once
#include <cstdint>

namespace rg {
constexpr uint8_t CLS_INVALID = 4;   // kept by needletail normalize(false) but not A/C/G/T: breaks a window
constexpr uint8_t CLS_SKIP = 5;      // removed by normalize(false): space, tab, CR, LF

// lut_be: 8-bit code with the first base in the high bits -> bin;  lut_le: first base in the low bits
// (the order of the packed words) -> bin;  names: 136 x 4 ASCII.  Any pointer may be null.
void build_bin_tables(uint8_t *lut_be, uint8_t *lut_le, uint8_t *names);
void build_class_table(uint8_t *cls);
}  // namespace rg
