// Minimal stand-in for the slice of cpp-taskflow 2.1.0 that the integrator concept touches
// (integrator.h:19-25: construct, let the integrator emplace work, dispatch() → shared_future).
// The reference vendors the real library; it is not part of this product.  When
// gpu_path_integrator.h is dropped into the reference tree the real <taskflow/taskflow.hpp>
// is used instead — the integrator only calls emplace().
#pragma once
#include <algorithm>
#include <atomic>
#include <exception>
#include <functional>
#include <memory>
#include <mutex>
#include <future>
#include <thread>
#include <vector>

namespace tf {
class Taskflow
{
public:
    explicit Taskflow(unsigned workers = std::thread::hardware_concurrency()) : workers_(workers ? workers : 1) {}
    ~Taskflow() { if (runner_.joinable()) runner_.join(); }
    template <class F>
    void emplace(F &&f) { tasks_.emplace_back(std::forward<F>(f)); }
    // Runs the emplaced tasks (independent of each other) and returns a future for all of them.
    std::shared_future<void> dispatch()
    {
        auto batch = std::make_shared<std::vector<std::function<void()>>>(std::move(tasks_));
        tasks_.clear();
        auto done = std::make_shared<std::promise<void>>();
        std::shared_future<void> fut = done->get_future().share();
        if (runner_.joinable()) runner_.join();
        runner_ = std::thread([batch, done, n = workers_] {
            std::vector<std::thread> pool;
            std::atomic<size_t> next{0};
            std::mutex guard;
            std::exception_ptr failure;
            const unsigned k = (unsigned)std::min<size_t>(n, batch->size());
            for (unsigned w = 0; w < k; ++w)
                pool.emplace_back([&] {
                    for (size_t i; (i = next++) < batch->size();)
                        try { (*batch)[i](); }
                        catch (...) { std::lock_guard<std::mutex> g(guard); if (!failure) failure = std::current_exception(); }
                });
            for (auto &t : pool) t.join();
            if (failure) done->set_exception(failure); else done->set_value();
        });
        return fut;
    }
private:
    unsigned workers_;
    std::vector<std::function<void()>> tasks_;
    std::thread runner_;
};
} // namespace tf
