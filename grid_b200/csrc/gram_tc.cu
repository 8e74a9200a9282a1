// placeholder until the tcgen05 kernel lands (next commit)
#include "common.cuh"
int gridb200_gram_tc(const void*, int64_t, int64_t, const void*, int64_t, int64_t, int64_t, int32_t, int64_t*, int64_t,
                     int, void*) {
  g2::set_error("tcgen05 Gram kernel not built");
  return GRIDB200_ENODEV;
}
