// common.cu -- error text and launch accounting shared by every translation unit.
#include "common.cuh"
#include <stdlib.h>
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

static int env_int(const char *name, int dflt) {
    const char *e = getenv(name);
    return (e && *e) ? atoi(e) : dflt;
}

const Tuning &tuning() {
    static const Tuning t = {env_int("GNDNET_DEBUG", 0) || (getenv("GNDNET_DEBUG") != nullptr), getenv("GNDNET_VOX_SERIAL") != nullptr,
                             env_int("GNDNET_VOX_K", 0), env_int("GNDNET_VOX_G", 0), env_int("GNDNET_PFN_GSHIFT", -1),
                             env_int("GNDNET_H2_SMALL_NT", 64), env_int("GNDNET_H2_MINSB", 6)};
    return t;
}

}  // namespace gnd

extern "C" const char *gnd_last_error_string(void) { return gnd::err_buf(); }
extern "C" int gnd_version(void) { return 200; }
