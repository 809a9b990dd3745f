// Launch counter and optional per-kernel-family CUDA-event timing (no nsys on this image: this is how bench.py
// attributes the step time to kernels, live, on the launching stream).
#include <algorithm>
#include <map>
#include <mutex>
#include <string>
#include <vector>
#include "../../include/mrisr.h"
#include "common.cuh"

namespace mrisr {
namespace {
struct Rec { const char* tag; cudaEvent_t e0, e1; };
std::mutex g_mu;
bool g_on = false;
std::vector<Rec> g_recs;
std::vector<cudaEvent_t> g_pool;
long long g_launches = 0;

cudaEvent_t get_event() {
    if (!g_pool.empty()) { cudaEvent_t e = g_pool.back(); g_pool.pop_back(); return e; }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}
}  // namespace

void count_launch() { ++g_launches; }
bool prof_on() { return g_on; }
const char* prof_intern(const char* tag) {
    static std::map<std::string, std::string> table;
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = table.find(tag);
    if (it == table.end()) it = table.emplace(tag, tag).first;
    return it->second.c_str();
}

ProfScope::ProfScope(const char* tag, cudaStream_t stream) : slot(-1), s(stream) {
    if (!g_on) return;
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(stream, &st);
    if (st != cudaStreamCaptureStatusNone) return;      // events with timing cannot be captured
    std::lock_guard<std::mutex> lk(g_mu);
    Rec r{tag, get_event(), get_event()};
    cudaEventRecord(r.e0, stream);
    slot = (int)g_recs.size();
    g_recs.push_back(r);
}
ProfScope::~ProfScope() {
    if (slot < 0) return;
    std::lock_guard<std::mutex> lk(g_mu);
    cudaEventRecord(g_recs[slot].e1, s);
}
}  // namespace mrisr

using namespace mrisr;

extern "C" {

long long mrisr_launch_count(void) { return g_launches; }

int mrisr_prof_enable(int on) {
    std::lock_guard<std::mutex> lk(g_mu);
    g_on = on != 0;
    return 0;
}

// Synchronises the device, then writes one line per tag: "<tag> <calls> <total_ms>\n" (sorted by time, descending)
// into buf (NUL-terminated, truncated to cap) and clears the records.  Returns the number of tags.
int mrisr_prof_report(char* buf, size_t cap) {
    cudaDeviceSynchronize();
    std::lock_guard<std::mutex> lk(g_mu);
    std::map<std::string, std::pair<int, double>> agg;
    for (auto& r : g_recs) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) {
            auto& a = agg[r.tag];
            a.first += 1;
            a.second += ms;
        }
        g_pool.push_back(r.e0);
        g_pool.push_back(r.e1);
    }
    g_recs.clear();
    std::vector<std::pair<double, std::string>> order;
    for (auto& kv : agg) order.push_back({-kv.second.second, kv.first});
    std::sort(order.begin(), order.end());
    std::string out;
    for (auto& o : order) {
        char line[256];
        snprintf(line, sizeof(line), "%s %d %.6f\n", o.second.c_str(), agg[o.second].first, agg[o.second].second);
        out += line;
    }
    if (buf && cap) {
        const size_t n = out.size() < cap - 1 ? out.size() : cap - 1;
        memcpy(buf, out.data(), n);
        buf[n] = 0;
    }
    return (int)agg.size();
}

}  // extern "C"
