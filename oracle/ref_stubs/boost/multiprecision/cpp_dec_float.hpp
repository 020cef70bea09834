// Boost.Multiprecision stand-in.  TEST INFRASTRUCTURE (see ../../README.md).
//
// The reference uses cpp_dec_float_50 for one thing: `1.0 - cdf(chi_squared<cpp_dec_float_50>, stat)` when
// stat > 85 (src/stats_test.cpp:205-208, 246-252), where 50 decimal digits make the subtraction the true upper tail,
// and to print that value (src/utils.cpp:17-29).  This stand-in is a long double that carries its complement when it
// came out of a cdf, so the same expression yields Q(df/2, stat/2) without cancellation.
#pragma once
#include <cmath>
#include <cstdlib>
#include <ostream>
#include <string>

namespace boost { namespace multiprecision {

class cpp_dec_float_50 {
public:
    cpp_dec_float_50() : v_(0), c_(1), has_c_(false) {}
    cpp_dec_float_50(double v) : v_(v), c_(0), has_c_(false) {}
    cpp_dec_float_50(long double v) : v_(v), c_(0), has_c_(false) {}
    cpp_dec_float_50(int v) : v_(v), c_(0), has_c_(false) {}
    cpp_dec_float_50(unsigned long v) : v_((long double)v), c_(0), has_c_(false) {}
    explicit cpp_dec_float_50(const char* s) : v_(strtold(s, nullptr)), c_(0), has_c_(false) {}
    static cpp_dec_float_50 with_complement(long double v, long double c) {
        cpp_dec_float_50 r(v);
        r.c_ = c;
        r.has_c_ = true;
        return r;
    }
    long double value() const { return v_; }
    template <class T> T convert_to() const { return static_cast<T>(v_); }

    friend cpp_dec_float_50 operator-(const cpp_dec_float_50& a, const cpp_dec_float_50& b) {
        if (a.v_ == 1.0L && b.has_c_) return with_complement(b.c_, b.v_);
        return cpp_dec_float_50(a.v_ - b.v_);
    }
    friend bool operator<(const cpp_dec_float_50& a, const cpp_dec_float_50& b) { return a.v_ < b.v_; }
    friend bool operator!=(const cpp_dec_float_50& a, const cpp_dec_float_50& b) { return a.v_ != b.v_; }
    friend bool operator==(const cpp_dec_float_50& a, const cpp_dec_float_50& b) { return a.v_ == b.v_; }
    friend cpp_dec_float_50 abs(const cpp_dec_float_50& a) { return cpp_dec_float_50(fabsl(a.v_)); }
    friend std::ostream& operator<<(std::ostream& os, const cpp_dec_float_50& a) { return os << a.v_; }

private:
    long double v_, c_;
    bool has_c_;
};

}}  // namespace boost::multiprecision
