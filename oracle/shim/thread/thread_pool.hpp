// ORACLE / TEST INFRASTRUCTURE — not part of the product.
//
// Stand-in for ma-libs' header-only cpp/utils/thread/thread_pool.hpp (absent submodule).  Interface from
// the only call sites, src/nn_engine_trainer.cpp:460 (default construction, once per generation), :533
// (push_task with a void() callable) and :538 (wait_for_tasks).  Worker count =
// std::thread::hardware_concurrency(), overridable with LL_REF_THREADS so the CPU baseline can report
// the core count it really used.
#ifndef ORACLE_SHIM_THREAD_POOL_HPP_
#define ORACLE_SHIM_THREAD_POOL_HPP_

#include <condition_variable>
#include <cstdlib>
#include <functional>
#include <mutex>
#include <queue>
#include <thread>
#include <vector>

namespace tp
{
    class thread_pool
    {
      public:
        static unsigned default_threads()
        {
            if (const char* e = std::getenv("LL_REF_THREADS"))
            {
                int n = std::atoi(e);
                if (n > 0) return static_cast<unsigned>(n);
            }
            unsigned n = std::thread::hardware_concurrency();
            return n ? n : 1u;
        }

        explicit thread_pool(unsigned n = 0)
        {
            if (n == 0) n = default_threads();
            for (unsigned i = 0; i < n; ++i) workers_.emplace_back([this] { run(); });
        }
        ~thread_pool()
        {
            {
                std::lock_guard<std::mutex> lk(m_);
                stop_ = true;
            }
            cv_work_.notify_all();
            for (auto& t : workers_) t.join();
        }
        thread_pool(const thread_pool&)            = delete;
        thread_pool& operator=(const thread_pool&) = delete;

        template <typename F> void push_task(F&& f)
        {
            {
                std::lock_guard<std::mutex> lk(m_);
                q_.emplace(std::forward<F>(f));
                ++pending_;
            }
            cv_work_.notify_one();
        }

        void wait_for_tasks()
        {
            std::unique_lock<std::mutex> lk(m_);
            cv_done_.wait(lk, [this] { return pending_ == 0; });
        }

      private:
        void run()
        {
            for (;;)
            {
                std::function<void()> job;
                {
                    std::unique_lock<std::mutex> lk(m_);
                    cv_work_.wait(lk, [this] { return stop_ || !q_.empty(); });
                    if (q_.empty()) return;
                    job = std::move(q_.front());
                    q_.pop();
                }
                job();
                {
                    std::lock_guard<std::mutex> lk(m_);
                    if (--pending_ == 0) cv_done_.notify_all();
                }
            }
        }

        std::vector<std::thread> workers_;
        std::queue<std::function<void()>> q_;
        std::mutex m_;
        std::condition_variable cv_work_, cv_done_;
        size_t pending_ = 0;
        bool stop_      = false;
    };
} // namespace tp

#endif
