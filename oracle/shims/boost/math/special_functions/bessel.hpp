// Build shim for the oracle only: the reference includes boost's Bessel header,
// boost is not installed here, and the only symbol used (cyl_bessel_k, in the
// Matern kernel, off this path) exists in C++17 <cmath>.
#pragma once
#include <cmath>
namespace boost { namespace math {
template <class T1, class T2> inline double cyl_bessel_k(T1 v, T2 x) { return std::cyl_bessel_k((double)v, (double)x); }
}}
