// fxcu_hostpar.h -- the host-thread helper of the cache builders that need the HOST's libm (coordinate search, fxcu_kdtree.cu
// and fxcu_nn.cu): split [0, n) over the host cores.
#pragma once

#include <algorithm>
#include <cstddef>
#include <thread>
#include <vector>

namespace fxcu {

inline unsigned host_threads()
{
    unsigned hw = std::thread::hardware_concurrency();
    return hw == 0 ? 4u : std::min(hw, 64u);
}

template <class F> void parallel_ranges(size_t n, size_t grain, F&& fn)
{
    const size_t parts = std::min<size_t>(host_threads(), (n + grain - 1) / grain);
    if (parts <= 1) {
        fn(size_t(0), n);
        return;
    }
    std::vector<std::thread> th;
    th.reserve(parts - 1);
    const size_t step = (n + parts - 1) / parts;
    for (size_t k = 1; k < parts; ++k)
        th.emplace_back([&fn, n, step, k]() { fn(std::min(n, k * step), std::min(n, (k + 1) * step)); });
    fn(size_t(0), std::min(n, step));
    for (auto& t : th)
        t.join();
}

} // namespace fxcu
