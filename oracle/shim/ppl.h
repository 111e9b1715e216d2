// Linux shim for MSVC <ppl.h> (RayTracingWeekend.cpp:15-16, :191). Test infrastructure only.
//
// concurrency::parallel_for(first, last, step, f) on std::thread with an atomic work counter.
// Nested calls run serially on the calling worker (a thread_local flag), which matches how the
// reference's 3-deep nest (rows, columns, samples) is effectively scheduled.  Thread count comes from
// RTW_PPL_THREADS (default: hardware_concurrency; 1 = fully serial and therefore deterministic).
#pragma once
#include <atomic>
#include <cstdlib>
#include <thread>
#include <vector>

namespace concurrency {

inline int shim_thread_count() {
    static int n = [] {
        const char* e = std::getenv("RTW_PPL_THREADS");
        int v = e ? std::atoi(e) : (int)std::thread::hardware_concurrency();
        return v < 1 ? 1 : v;
    }();
    return n;
}

inline bool& shim_in_parallel() {
    static thread_local bool flag = false;
    return flag;
}

template <typename Index, typename Function>
void parallel_for(Index first, Index last, Index step, const Function& f) {
    int nt = shim_thread_count();
    if (shim_in_parallel() || nt == 1) {
        for (Index i = first; i < last; i += step) f(i);
        return;
    }
    long long count = last > first ? ((long long)(last - first) + step - 1) / step : 0;
    std::atomic<long long> next(0);
    auto worker = [&] {
        shim_in_parallel() = true;
        for (;;) {
            long long k = next.fetch_add(1);
            if (k >= count) break;
            f((Index)(first + k * step));
        }
        shim_in_parallel() = false;
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < nt; t++) pool.emplace_back(worker);
    worker();
    for (auto& t : pool) t.join();
}

}  // namespace concurrency
namespace Concurrency = concurrency;
