// Per-kernel-family device timing behind kernels.h:prof_* (measurement aid for bench.py's roofline line).
#include "kernels.h"
#include <cstdlib>

#include <mutex>
#include <vector>

namespace ace {

namespace {
struct Rec {
    cudaEvent_t a, b;
    int cat;
    double flops, bytes;
};
bool g_on = false;
std::vector<Rec> g_recs;
std::vector<cudaEvent_t> g_pool;
std::mutex g_mu;

cudaEvent_t get_event() {
    if (!g_pool.empty()) {
        cudaEvent_t e = g_pool.back();
        g_pool.pop_back();
        return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}
}  // namespace

bool prof_enabled() { return g_on; }

void prof_enable(bool on) {
    std::lock_guard<std::mutex> lk(g_mu);
    for (auto& r : g_recs) {
        g_pool.push_back(r.a);
        g_pool.push_back(r.b);
    }
    g_recs.clear();
    g_on = on;
}

void prof_begin(int cat, double flops, double bytes, cudaStream_t st) {
    std::lock_guard<std::mutex> lk(g_mu);
    Rec r{get_event(), get_event(), cat, flops, bytes};
    cudaEventRecord(r.a, st);
    g_recs.push_back(r);
}

void prof_end(cudaStream_t st) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_recs.empty()) cudaEventRecord(g_recs.back().b, st);
}

void prof_collect(ProfTotals& out) {
    std::lock_guard<std::mutex> lk(g_mu);
    for (int c = 0; c < PROF_NCAT; ++c) out.ms[c] = out.flops[c] = out.bytes[c] = 0, out.launches[c] = 0;
    for (auto& r : g_recs) {
        cudaEventSynchronize(r.b);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
            out.ms[r.cat] += ms;
            out.flops[r.cat] += r.flops;
            out.bytes[r.cat] += r.bytes;
            out.launches[r.cat] += 1;
        }
    }
}


bool pdl_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = std::getenv("ACE_B200_PDL");
        v = (e && e[0] == '0') ? 0 : 1;
    }
    return v == 1;
}

}  // namespace ace
