// ref_product.cpp -- TEST INFRASTRUCTURE: oracle/_ref/libsdmm_ref_product.so = the reference driver plus the
// product-guiding entry points of ref_product.inc, as its OWN shared object.  Kept out of libsdmm_ref.so on
// purpose: adding the sdmm_product.h instantiations to that translation unit changes gcc's inlining of the
// reference's EM (-funsafe-math-optimizations -ffp-contract=fast), which moves samples sitting on the FTZ
// underflow edge and with them the EM fixtures pinned to that build.  RefCell handles created by libsdmm_ref.so
// are passed in as plain pointers (same headers, same flags, same layout).
#define SDMM_REF_WITH_PRODUCT 1
#include "ref_driver.cpp"
