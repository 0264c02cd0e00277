// Test-infrastructure shim (oracle/_ref build only); see fmt/format.h.
#pragma once
#include "format.h"
