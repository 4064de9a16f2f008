// C entry point over host/image_io.h, so that the readers can be exercised without the C++ classes (tests/test_image_io.py).
#include "image_io.h"

extern "C" int vo_host_read_frame(const char* path, uint8_t* out, int64_t cap, int32_t* width, int32_t* height, char* err,
                                  int32_t err_cap) {
  std::vector<uint8_t> grey;
  std::string why;
  if (!path || !width || !height) return 1;
  if (!vo_io::read_frame(path, grey, width, height, &why)) {
    if (err && err_cap > 0) snprintf(err, (size_t)err_cap, "%s", why.c_str());
    return 2;
  }
  if ((int64_t)grey.size() > cap || !out) return 3;
  memcpy(out, grey.data(), grey.size());
  return 0;
}
