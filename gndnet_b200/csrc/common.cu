// common.cu -- error text and launch accounting shared by every translation unit.
#include "common.cuh"
#include <string.h>

namespace gnd {

thread_local int g_launches = 0;
static thread_local char t_err[512] = "";

char *err_buf() { return t_err; }

int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof(t_err), fmt, ap);
    va_end(ap);
    return code;
}

}  // namespace gnd

extern "C" const char *gnd_last_error_string(void) { return gnd::err_buf(); }
extern "C" int gnd_version(void) { return 100; }
