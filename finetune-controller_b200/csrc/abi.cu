// Library identity and per-thread error string for the C ABI (include/b200ft.h).
#include <stdarg.h>
#include <stdlib.h>
#include "common.cuh"
#include "../../include/b200ft.h"

static thread_local char g_err[512] = "";

void b200ft_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

bool b200ft_pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("B200FT_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}

extern "C" const char* b200ft_version(void) { return "b200ft 0.1.0 (sm_100a)"; }
extern "C" const char* b200ft_last_error_string(void) { return g_err; }
