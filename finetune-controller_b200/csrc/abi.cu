// Library identity and per-thread error string for the C ABI (include/b200ft.h).
#include <stdarg.h>
#include <stdlib.h>
#include <atomic>
#include "common.cuh"
#include "../../include/b200ft.h"

static thread_local char g_err[512] = "";

void b200ft_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static std::atomic<int> g_pdl_runtime{1};

bool b200ft_pdl_enabled() {
  static const bool env_on = [] {
    const char* e = getenv("B200FT_PDL");
    return !(e && e[0] == '0');
  }();
  return env_on && g_pdl_runtime.load(std::memory_order_relaxed) != 0;
}

extern "C" int b200ft_set_dependent_launch(int enabled) { return g_pdl_runtime.exchange(enabled ? 1 : 0); }

extern "C" const char* b200ft_version(void) { return "b200ft 0.1.0 (sm_100a)"; }
extern "C" const char* b200ft_last_error_string(void) { return g_err; }
