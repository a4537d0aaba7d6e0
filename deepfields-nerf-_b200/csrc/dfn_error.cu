// Thread-local error string for the C-ABI (include/deepfields_b200.h: dfn_last_error).
#include "dfn_common.cuh"
#include <stdarg.h>

static thread_local char g_err[512] = "";

void dfn_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* dfn_last_error() { return g_err; }
extern "C" int dfn_abi_version() { return 1; }
