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

// SM budget of the MLP launches of the calling thread (0 = the whole device).  The persistent MLP kernels size their grids to
// the SM count; a caller that runs two of them CONCURRENTLY on two streams (the coarse network's backward beside the fine
// network's forward, train.FusedTrainer) gives each a share of the device so that both are resident.
static thread_local int g_sm_limit = 0;
extern "C" int dfn_set_sm_limit(int n_sms) {
  g_sm_limit = n_sms > 0 ? n_sms : 0;
  return 0;
}
int dfn_mlp_sms() {
  const int n = dfn_num_sms();
  return (g_sm_limit > 0 && g_sm_limit < n) ? g_sm_limit : n;
}
