// ck::utils::check_err (library/include/ck/library/utility/check_err.hpp:26-324): criterion
// |o - r| > atol + rtol*|r| or non-finite (:54); defaults float 1e-5/3e-6, half/bhalf 1e-3/1e-3, integers exact.
#pragma once
#include <cmath>
#include <iomanip>
#include <iostream>
#include <string>
#include <type_traits>

#include "ck/utility/data_type.hpp"

namespace ck {
namespace utils {
namespace detail {
template <typename T>
inline double as_double(const T& v)
{
    if constexpr(std::is_same_v<T, bhalf_t>) return double(type_convert<float>(v));
    else return double(v);
}
template <typename T>
struct tol { static constexpr double r = 1e-5, a = 3e-6; };
template <> struct tol<half_t> { static constexpr double r = 1e-3, a = 1e-3; };
template <> struct tol<bhalf_t> { static constexpr double r = 1e-3, a = 1e-3; };
} // namespace detail

template <typename Range, typename RefRange>
bool check_err(const Range& out, const RefRange& ref, const std::string& msg = "Error: Incorrect results!", double rtol = -1,
               double atol = -1)
{
    using T = std::remove_cv_t<std::remove_reference_t<decltype(*std::begin(out))>>;
    constexpr bool exact = std::is_integral_v<T> && !std::is_same_v<T, bhalf_t>;
    if(rtol < 0) rtol = exact ? 0 : detail::tol<T>::r;
    if(atol < 0) atol = exact ? 0 : detail::tol<T>::a;
    if(out.size() != ref.size())
    {
        std::cerr << msg << " out.size() != ref.size(), :" << out.size() << " != " << ref.size() << std::endl;
        return false;
    }
    bool res = true;
    int err_count = 0;
    double max_err = 0;
    auto io = std::begin(out);
    auto ir = std::begin(ref);
    for(std::size_t i = 0; i < ref.size(); ++i, ++io, ++ir)
    {
        const double o = detail::as_double(*io), r = detail::as_double(*ir);
        const double err = std::abs(o - r);
        if(err > atol + rtol * std::abs(r) || !std::isfinite(o) || !std::isfinite(r))
        {
            max_err = std::max(max_err, err);
            if(++err_count < 5)
                std::cerr << msg << std::setw(12) << std::setprecision(7) << " out[" << i << "] != ref[" << i << "]: " << o << " != " << r << std::endl;
            res = false;
        }
    }
    if(!res)
        std::cerr << "max err: " << max_err << ", number of errors: " << err_count << ", "
                  << 100.f * float(err_count) / float(out.size()) << "% wrong values" << std::endl;
    return res;
}
} // namespace utils
} // namespace ck
