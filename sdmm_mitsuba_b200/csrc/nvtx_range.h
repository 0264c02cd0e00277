// nvtx_range.h -- NVTX ranges around the phases of the hot path (bind / E / M / push / query), so that a timeline
// (Nsight Systems, or ncu --nvtx) shows them the way the reference's Tracy zones do (ZoneScoped in
// mitsuba/src/integrators/sdmm/sdmm_proc.cpp:287, 312, 360 and volpath_sdmm.cpp:245).  The NVTX library is resolved
// lazily by the header-only nvtx3 layer: without a profiler attached a range costs one predicted branch.
#pragma once
#include <nvtx3/nvToolsExt.h>

namespace sdmmb {
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};
}  // namespace sdmmb
#define SDMMB_NVTX(name) ::sdmmb::NvtxRange _nvtx_range_(name)
