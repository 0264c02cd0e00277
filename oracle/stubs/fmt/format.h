// Test-infrastructure shim (oracle/_ref build only): no-op stand-in for {fmt}.
// The reference only uses fmt::format for diagnostics; the image's fmt v12 is too new
// for the reference's implicit ostream formatting, so formatting is dropped.
#pragma once
#include <string>
namespace fmt {
template <typename... Args>
inline std::string format(const char*, Args&&...) { return std::string(); }
}  // namespace fmt
