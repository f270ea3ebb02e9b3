#include "prof.h"

#include <atomic>
#include <mutex>
#include <vector>

namespace ab {
namespace k {
namespace {

struct Sample { cudaEvent_t a, b; int family; double flops, bytes; };
std::atomic<bool> g_on{false};
std::mutex g_mu;
std::vector<Sample> g_samples;
std::vector<cudaEvent_t> g_free;

const char* kNames[F_COUNT] = {"conv3x3", "conv1x1", "up_cont", "wgrad", "bias_grad", "bn_mish", "max_pool", "loss", "argmax",
                               "blend", "input", "optimizer", "pack", "conv_tc", "wgrad_tc", "ew_f16"};

cudaEvent_t get_event()
{
    if (!g_free.empty()) { cudaEvent_t e = g_free.back(); g_free.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}

}  // namespace

bool prof_enabled() { return g_on.load(std::memory_order_relaxed); }

void prof_enable(bool on)
{
    std::lock_guard<std::mutex> lk(g_mu);
    for (Sample& s : g_samples) { g_free.push_back(s.a); g_free.push_back(s.b); }
    g_samples.clear();
    g_on.store(on);
}

void prof_collect(FamilyStat out[F_COUNT])
{
    cudaDeviceSynchronize();
    std::lock_guard<std::mutex> lk(g_mu);
    for (int f = 0; f < F_COUNT; ++f) out[f] = FamilyStat{kNames[f], 0, 0, 0, 0};
    for (const Sample& s : g_samples) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, s.a, s.b) != cudaSuccess) { cudaGetLastError(); continue; }
        FamilyStat& o = out[s.family];
        ++o.launches; o.ms += ms; o.flops += s.flops; o.bytes += s.bytes;
    }
}

ProfScope::ProfScope(int family, double flops, double bytes, cudaStream_t st_) : slot(-1), st(st_)
{
    if (!g_on.load(std::memory_order_relaxed)) return;
    std::lock_guard<std::mutex> lk(g_mu);
    Sample s{get_event(), get_event(), family, flops, bytes};
    cudaEventRecord(s.a, st);
    slot = (int)g_samples.size();
    g_samples.push_back(s);
}

ProfScope::~ProfScope()
{
    if (slot < 0) return;
    std::lock_guard<std::mutex> lk(g_mu);
    if (slot < (int)g_samples.size()) cudaEventRecord(g_samples[slot].b, st);
}

}  // namespace k
}  // namespace ab
