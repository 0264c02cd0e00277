// Test-infrastructure shim (oracle/_ref build only): Tracy profiler zones compiled out.
#pragma once
#define ZoneScoped
#define ZoneScopedN(x)
#define FrameMarkStart(x)
#define FrameMarkEnd(x)
#define FrameMark
