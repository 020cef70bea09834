// special_functions.hpp — TEST INFRASTRUCTURE (shared by oracle/stoat_oracle.cpp and the Boost stand-in headers
// under oracle/ref_stubs/boost/ that oracle/_ref is compiled against).  Not product code.
//
// Boost.Math is a system package the reference links (CMakeLists.txt:51, unpinned; Boost 1.74 in the published
// container, Dockerfile:14) and is absent here.  These are the published definitions of the functions behind the
// reference's call sites — chi_squared cdf (src/stats_test.cpp:207-210,249-255), students_t complement (:476,484),
// normal cdf (:23-24) — evaluated in long double, validated against mpmath at 60 digits (tests/test_oracle_tails.py)
// and pinned by the reference's golden strings (tests/golden/).
#pragma once
#include <cmath>
#include <cstdint>

namespace orc_math {
// ----------------------------------------------------------------------------------------------
// Distribution tails (Boost.Math stand-ins, see header).  long double throughout.
// ----------------------------------------------------------------------------------------------
using ld = long double;

// Q(a, x) = Gamma(a, x) / Gamma(a): Legendre continued fraction (modified Lentz), x > a + 1
inline ld gamma_q_cf(ld a, ld x) {
    const ld tiny = 1e-4000L;
    ld b = x + 1.0L - a;
    ld c = 1.0L / tiny;
    ld d = 1.0L / b;
    ld h = d;
    for (int i = 1; i < 100000; ++i) {
        const ld an = -ld(i) * (ld(i) - a);
        b += 2.0L;
        d = an * d + b;
        if (fabsl(d) < tiny) d = tiny;
        c = b + an / c;
        if (fabsl(c) < tiny) c = tiny;
        d = 1.0L / d;
        const ld delta = d * c;
        h *= delta;
        if (fabsl(delta - 1.0L) < 1e-20L) break;
    }
    return expl(-x + a * logl(x) - lgammal(a)) * h;
}

// P(a, x) = gamma(a, x) / Gamma(a): power series, x < a + 1
inline ld gamma_p_series(ld a, ld x) {
    ld term = 1.0L / a;
    ld sum = term;
    ld ap = a;
    for (int i = 0; i < 100000; ++i) {
        ap += 1.0L;
        term *= x / ap;
        sum += term;
        if (fabsl(term) < fabsl(sum) * 1e-21L) break;
    }
    return sum * expl(-x + a * logl(x) - lgammal(a));
}

// upper and lower regularized incomplete gamma, each computed without cancellation where it is
// the small one
inline void gamma_pq(ld a, ld x, ld& p, ld& q) {
    if (x <= 0.0L) {
        p = 0.0L;
        q = 1.0L;
        return;
    }
    if (x < a + 1.0L) {
        p = gamma_p_series(a, x);
        q = 1.0L - p;
    } else {
        q = gamma_q_cf(a, x);
        p = 1.0L - q;
    }
}

// The reference's chi-squared p-value — src/stats_test.cpp:205-210 (2x2) and :246-256 (2xN):
//   stat <= 85 : double  1.0 - boost::math::cdf(chi_squared(df), stat)
//   stat  > 85 : the same subtraction carried out in cpp_dec_float_50, i.e. the true upper tail.
// boost::math::cdf(chi_squared) = gamma_p(df/2, stat/2) rounded to double; the subtraction from
// 1.0 then quantises p to the double grid around 1 (this is why golden rows print "0").
inline double chi2_pvalue_like_reference(double stat, uint64_t df) {
    ld p, q;
    gamma_pq(ld(df) * 0.5L, ld(stat) * 0.5L, p, q);
    if (stat > 85.0) return double(q);
    const double cdf = double(p);  // what boost hands back as a double
    return 1.0 - cdf;
}

// continued fraction of the regularized incomplete beta (Lentz), valid for x < (a+1)/(a+b+2)
inline ld ibeta_cf(ld a, ld b, ld x) {
    const ld tiny = 1e-4000L;
    const ld qab = a + b, qap = a + 1.0L, qam = a - 1.0L;
    ld c = 1.0L;
    ld d = 1.0L - qab * x / qap;
    if (fabsl(d) < tiny) d = tiny;
    d = 1.0L / d;
    ld h = d;
    for (int m = 1; m < 100000; ++m) {
        const ld m2 = 2.0L * m;
        ld aa = m * (b - m) * x / ((qam + m2) * (a + m2));
        d = 1.0L + aa * d;
        if (fabsl(d) < tiny) d = tiny;
        c = 1.0L + aa / c;
        if (fabsl(c) < tiny) c = tiny;
        d = 1.0L / d;
        h *= d * c;
        aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
        d = 1.0L + aa * d;
        if (fabsl(d) < tiny) d = tiny;
        c = 1.0L + aa / c;
        if (fabsl(c) < tiny) c = tiny;
        d = 1.0L / d;
        const ld delta = d * c;
        h *= delta;
        if (fabsl(delta - 1.0L) < 1e-20L) break;
    }
    return h;
}

// I_x(a, b) and its complement, each without cancellation on its small side
inline void ibeta_pq(ld a, ld b, ld x, ld& ix, ld& icx) {
    if (x <= 0.0L) {
        ix = 0.0L;
        icx = 1.0L;
        return;
    }
    if (x >= 1.0L) {
        ix = 1.0L;
        icx = 0.0L;
        return;
    }
    const ld lbeta = lgammal(a) + lgammal(b) - lgammal(a + b);
    const ld front = expl(a * logl(x) + b * log1pl(-x) - lbeta);
    if (x < (a + 1.0L) / (a + b + 2.0L)) {
        ix = front * ibeta_cf(a, b, x) / a;
        icx = 1.0L - ix;
    } else {
        icx = front * ibeta_cf(b, a, 1.0L - x) / b;
        ix = 1.0L - icx;
    }
}

// boost::math::cdf(complement(students_t(df), t)) for t >= 0 — src/stats_test.cpp:484.
// Upper tail of Student's t: 0.5 * I_{df/(df+t^2)}(df/2, 1/2).
inline double student_t_upper_tail(double df, double t) {
    if (std::isinf(t)) return 0.0;
    const ld t2 = ld(t) * ld(t);
    ld ix, icx;
    if (ld(df) > 2.0L * t2) {
        // z = t^2/(df+t^2) is small: tail = 0.5 * (1 - I_z(1/2, df/2))
        ibeta_pq(0.5L, ld(df) * 0.5L, t2 / (ld(df) + t2), ix, icx);
        return double(0.5L * icx);
    }
    ibeta_pq(ld(df) * 0.5L, 0.5L, ld(df) / (ld(df) + t2), ix, icx);
    return double(0.5L * ix);
}

// Phi(z): boost::math::cdf(normal_distribution<>(0, 1), z) — src/stats_test.cpp:23-24
inline double normal_cdf(double z) { return double(0.5L * erfcl(-ld(z) / sqrtl(2.0L))); }

}  // namespace orc_math
