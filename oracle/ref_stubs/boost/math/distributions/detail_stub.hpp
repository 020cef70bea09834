// Boost.Math stand-in (see ../../../README.md).  TEST INFRASTRUCTURE.
#pragma once
#include "../../../../special_functions.hpp"
#include "../../multiprecision/cpp_dec_float.hpp"

namespace boost { namespace math {

template <class Dist, class Real>
struct complemented2_type {
    const Dist& dist;
    const Real& param;
};
template <class Dist, class Real>
inline complemented2_type<Dist, Real> complement(const Dist& d, const Real& r) { return {d, r}; }

}}  // namespace boost::math
