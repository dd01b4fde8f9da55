// B200 cudapoa engine -- helper templates the public batch header relies on.
// API-compatible with GenomeWorks common/base/include/claraparabricks/genomeworks/utils/signed_integer_utils.hpp:23-48.
#pragma once

#include <cassert>
#include <limits>
#include <stdexcept>
#include <type_traits>

namespace claraparabricks
{
namespace genomeworks
{

/// Size of a container as the signed counterpart of its size_type.
template <class Container>
typename std::make_signed<typename Container::size_type>::type get_size(const Container& c)
{
    using signed_t = typename std::make_signed<typename Container::size_type>::type;
    assert(c.size() <= static_cast<typename Container::size_type>(std::numeric_limits<signed_t>::max()));
    return static_cast<signed_t>(c.size());
}

/// Size of a container converted to an explicitly requested integer type.
template <class Integer, class Container>
Integer get_size(const Container& c)
{
    assert(c.size() <= static_cast<typename Container::size_type>(std::numeric_limits<Integer>::max()));
    return static_cast<Integer>(c.size());
}

/// Returns x unchanged, or throws std::invalid_argument(message) when x < 0.
template <class T>
T throw_on_negative(T x, const char* message)
{
    static_assert(std::is_arithmetic<T>::value, "throw_on_negative needs an arithmetic type");
    if (x < T(0))
    {
        throw std::invalid_argument(message);
    }
    return x;
}

} // namespace genomeworks
} // namespace claraparabricks
