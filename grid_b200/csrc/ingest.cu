// placeholder until the device text parser lands
#include "common.cuh"
extern "C" size_t gridb200_ingest_ws_bytes(int64_t nbytes) { return (size_t)(nbytes > 0 ? nbytes : 0) * 8 + 65536; }
extern "C" int gridb200_ingest_text(const char*, int64_t, const char*, int32_t, int64_t, int64_t, const uint32_t*, int64_t,
                                    int64_t, int64_t, float*, int64_t*, void*, size_t, void*) {
  g2::set_error("device ingest not built");
  return GRIDB200_ENODEV;
}
