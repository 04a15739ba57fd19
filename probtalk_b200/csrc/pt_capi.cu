// probtalk_b200 -- library-level entry points (version, last error).
#include "pt_host.h"
#include "../../include/probtalk_b200.h"

namespace pt {
char* error_buffer() {
  static thread_local char buf[kErrorBufferSize] = {0};
  return buf;
}
}  // namespace pt

extern "C" int pt_version(void) { return 100; }
extern "C" const char* pt_last_error(void) { return pt::error_buffer(); }
