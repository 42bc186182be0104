// Build shim for the oracle only (boost absent): erf from <cmath>; erf_inv is
// only reached from GaussianProcess CDF code that is off the sparse-conv path.
#pragma once
#include <cmath>
#include <limits>
namespace boost { namespace math {
template <class T> inline T erf(T x) { return std::erf(x); }
template <class T> inline T erf_inv(T x) {
    // Giles' single-precision-quality approximation refined by two Newton steps.
    if (x <= T(-1)) return -std::numeric_limits<T>::infinity();
    if (x >= T(1)) return std::numeric_limits<T>::infinity();
    double w = -std::log((1.0 - (double)x) * (1.0 + (double)x)), p;
    if (w < 5.0) { w -= 2.5; p = 2.81022636e-08; p = 3.43273939e-07 + p*w; p = -3.5233877e-06 + p*w; p = -4.39150654e-06 + p*w; p = 0.00021858087 + p*w; p = -0.00125372503 + p*w; p = -0.00417768164 + p*w; p = 0.246640727 + p*w; p = 1.50140941 + p*w; }
    else { w = std::sqrt(w) - 3.0; p = -0.000200214257; p = 0.000100950558 + p*w; p = 0.00134934322 + p*w; p = -0.00367342844 + p*w; p = 0.00573950773 + p*w; p = -0.0076224613 + p*w; p = 0.00943887047 + p*w; p = 1.00167406 + p*w; p = 2.83297682 + p*w; }
    double y = p * (double)x;
    for (int i = 0; i < 2; ++i) y -= (std::erf(y) - (double)x) / (1.1283791670955126 * std::exp(-y*y));
    return (T)y;
}
}}
