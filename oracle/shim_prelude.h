// TEST INFRASTRUCTURE ONLY (oracle build).  Force-included into the two
// sed-patched shadow copies of reference sources so that GCC accepts the
// designated initialisers that clang accepted: `.member { 0 }` through an
// explicit constructor becomes `.member = hs_oracle_zero_init{}`.
#pragma once

struct hs_oracle_zero_init {
    template <typename T>
    operator T() const { return T(0); }
};
