// Boost.Math stand-in: students_t + complemented cdf.  TEST INFRASTRUCTURE (see ../../../README.md).
#pragma once
#include "detail_stub.hpp"

namespace boost { namespace math {

template <class RealType = double>
class students_t_distribution {
public:
    students_t_distribution(RealType df) : df_(df) {}
    RealType degrees_of_freedom() const { return df_; }
private:
    RealType df_;
};
typedef students_t_distribution<double> students_t;

// upper tail for t >= 0 (the only use: src/stats_test.cpp:484 passes std::abs(t))
inline double cdf(const complemented2_type<students_t_distribution<double>, double>& c) {
    return orc_math::student_t_upper_tail(c.dist.degrees_of_freedom(), c.param);
}

}}  // namespace boost::math
