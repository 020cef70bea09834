// Boost.Math stand-in: chi_squared_distribution + cdf.  TEST INFRASTRUCTURE (see ../../../README.md).
#pragma once
#include "detail_stub.hpp"

namespace boost { namespace math {

template <class RealType = double>
class chi_squared_distribution {
public:
    chi_squared_distribution(RealType df) : df_(df) {}
    RealType degrees_of_freedom() const { return df_; }
private:
    RealType df_;
};
typedef chi_squared_distribution<double> chi_squared;

// cdf(chi_squared(df), x) = P(df/2, x/2), rounded to double
inline double cdf(const chi_squared_distribution<double>& d, const double& x) {
    long double p, q;
    orc_math::gamma_pq(0.5L * d.degrees_of_freedom(), 0.5L * (long double)x, p, q);
    return double(p);
}
// the 50-digit instantiation: the result remembers Q so that `1.0 - cdf` is the true upper tail
inline multiprecision::cpp_dec_float_50 cdf(const chi_squared_distribution<multiprecision::cpp_dec_float_50>& d,
                                            const multiprecision::cpp_dec_float_50& x) {
    long double p, q;
    orc_math::gamma_pq(0.5L * d.degrees_of_freedom().value(), 0.5L * x.value(), p, q);
    return multiprecision::cpp_dec_float_50::with_complement(p, q);
}

}}  // namespace boost::math
