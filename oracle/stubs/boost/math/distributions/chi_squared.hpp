// TEST-ONLY stub of the one boost facility the reference path touches
// (sdmm/include/sdmm/opt/init.h:6,41-44: quantile(chi_squared(3), 0.90)).  boost 1.71 is a
// conan dependency of the reference (sdmm/conanfile.txt:6) and is absent from this image; the
// quantile is the published constant chi2_3(0.90) = 6.251388631170325.
#pragma once
#include <stdexcept>
namespace boost { namespace math {
struct chi_squared {
    double dof;
    explicit chi_squared(double d) : dof(d) {}
};
inline double quantile(const chi_squared& d, double p) {
    if (d.dof == 3.0 && p > 0.8999 && p < 0.9001) return 6.251388631170325;
    throw std::runtime_error("chi_squared stub: only quantile(chi2(3), 0.90) is provided");
}
}}  // namespace boost::math
