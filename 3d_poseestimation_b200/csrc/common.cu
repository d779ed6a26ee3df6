#include "common.cuh"
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

namespace lifter {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

const char* get_error() { return g_error; }

int num_sms() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

bool pdl_enabled() {
    static int cached = -1;
    if (cached < 0) {
        const char* e = getenv("LIFTER_NO_PDL");
        cached = (e != nullptr && e[0] == '1') ? 0 : 1;
    }
    return cached == 1;
}

}  // namespace lifter
