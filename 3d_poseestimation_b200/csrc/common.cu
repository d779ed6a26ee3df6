#include "common.cuh"
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <string>
#include <vector>

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

// ---- debug trace -------------------------------------------------------------------------------------
namespace {
struct TraceRec { std::string name; cudaEvent_t ev; };
std::vector<TraceRec> g_trace;
bool g_trace_on = false;
}  // namespace

static std::atomic<long long> g_launches{0};
long long launch_count() { return g_launches.load(); }

// Peer waits of the data-parallel kernels (SyncBN exchange, optimiser flag barriers) give up after this many SM clocks
// and trap -- a peer that died must not hang the box for good -- but a healthy peer may be late by whole seconds
// (checkpoint on rank 0, a DataLoader respawn).  LIFTER_DP_TIMEOUT_S: seconds (default 300, like a collective
// watchdog; 0 = wait forever).
long long dp_spin_limit_clocks() {
    static long long cached = -2;
    if (cached == -2) {
        double s = 300.0;
        const char* e = getenv("LIFTER_DP_TIMEOUT_S");
        if (e != nullptr && e[0] != 0) s = atof(e);
        cached = s <= 0.0 ? 0 : static_cast<long long>(s * 2.0e9);
    }
    return cached;
}

void trace_mark(const char* name, cudaStream_t st) {
    g_launches.fetch_add(1, std::memory_order_relaxed);     // every launch of the library passes through here
    if (!g_trace_on) return;
    TraceRec r;
    r.name = name;
    if (cudaEventCreate(&r.ev) != cudaSuccess) return;
    cudaEventRecord(r.ev, st);
    g_trace.push_back(r);
}

int trace_begin(cudaStream_t st) {
    for (auto& r : g_trace) cudaEventDestroy(r.ev);
    g_trace.clear();
    g_trace_on = true;
    trace_mark("begin", st);
    return LIFTER_OK;
}

// writes "name us\n" per launch since trace_begin (time from the previous mark to this launch's completion)
int trace_end(char* buf, int len) {
    g_trace_on = false;
    if (buf == nullptr || len <= 0) return LIFTER_ERR_INVALID;
    int off = 0;
    buf[0] = 0;
    for (size_t i = 1; i < g_trace.size(); ++i) {
        if (cudaEventSynchronize(g_trace[i].ev) != cudaSuccess) return LIFTER_ERR_CUDA;
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, g_trace[i - 1].ev, g_trace[i].ev);
        const int n = snprintf(buf + off, len - off, "%s %.3f\n", g_trace[i].name.c_str(), ms * 1e3);
        if (n < 0 || n >= len - off) break;
        off += n;
    }
    for (auto& r : g_trace) cudaEventDestroy(r.ev);
    g_trace.clear();
    return off;
}

}  // namespace lifter
