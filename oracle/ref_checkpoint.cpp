// oracle/ref_checkpoint.cpp -- TEST-ONLY checker, never linked into the product.
// The reference driver (ref_driver.cpp, unmodified reference headers) compiled once more with a
// nlohmann_json that has NLOHMANN_DEFINE_TYPE_INTRUSIVE (the reference pins nlohmann_json/3.9.1 in
// sdmm/conanfile.txt:9; the copy vendored under sdmm/ext/jmm is 3.7.3 and lacks the macro), so that the
// reference's own save_json / from_json of the STree accelerator -- the `.asdmm` checkpoint of
// mitsuba/src/integrators/sdmm/volpath_sdmm.cpp:120-130 -- can be run here.  The newer single header
// comes from the image (cudnn_frontend/thirdparty, 3.11); see oracle/Makefile (JSON_INC).
#include <nlohmann/json.hpp>
#if NLOHMANN_JSON_VERSION_MAJOR * 100 + NLOHMANN_JSON_VERSION_MINOR < 309
#error "ref_checkpoint.cpp needs nlohmann_json >= 3.9 ahead of the reference's vendored copy on the include path"
#endif
#include <fstream>
#define SDMM_REF_WITH_JSON 1
#include "ref_driver.cpp"
