// Test-infrastructure shim (oracle/_ref build only): no-op stand-in for spdlog.
// <mutex>/<iostream> are pulled in because the reference relies on spdlog's transitive
// includes (sdmm/core/utils.h MutexWrapper, spatio_directional.h std::cerr).
#pragma once
#include <iostream>
#include <mutex>
#include <string>
namespace spdlog {
template <typename... Args> inline void info(Args&&...) {}
template <typename... Args> inline void warn(Args&&...) {}
template <typename... Args> inline void debug(Args&&...) {}
template <typename... Args> inline void error(Args&&...) {}
}  // namespace spdlog
