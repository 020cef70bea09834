// Boost.Math stand-in: normal_distribution + cdf.  TEST INFRASTRUCTURE (see ../../../README.md).
#pragma once
#include "detail_stub.hpp"

namespace boost { namespace math {

template <class RealType = double>
class normal_distribution {
public:
    normal_distribution(RealType mean = 0, RealType sd = 1) : mean_(mean), sd_(sd) {}
    RealType mean() const { return mean_; }
    RealType standard_deviation() const { return sd_; }
private:
    RealType mean_, sd_;
};
typedef normal_distribution<double> normal;

inline double cdf(const normal_distribution<double>& d, const double& x) {
    return orc_math::normal_cdf((x - d.mean()) / d.standard_deviation());
}

}}  // namespace boost::math
