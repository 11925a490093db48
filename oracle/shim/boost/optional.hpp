/* oracle/shim: map boost::optional to std::optional (C++17) for crypto/equihash.cpp. */
#pragma once
#include <optional>
namespace boost {
template <class T> using optional = std::optional<T>;
static const std::nullopt_t none = std::nullopt;
}
