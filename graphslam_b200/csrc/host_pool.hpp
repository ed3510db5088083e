// A process-wide pool of host threads for the library's data-parallel host loops (structure build, aggregation, staging
// of the structure upload, pose conversions).  Creating a std::thread costs ~20 us; an incremental update runs some
// twenty such loops over a few tens of thousands of items each, and an end-to-end LM step four over a million — with
// fresh threads per loop the spawning was a quarter of the structure phase of a 50,000-pose update.
// Standard C++ only (structure.hpp includes this and stays compilable without CUDA).
#pragma once
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

namespace gs {

class HostPool {
 public:
  // never destroyed: the workers are detached and may still be parked in wait() when static destructors run
  static HostPool& get() {
    static HostPool* p = new HostPool;
    return *p;
  }

  // fn(part) for every part in [0, parts), on the calling thread and up to parts - 1 workers; returns when all parts are
  // done.  Calls from different threads take turns.  fn must not call run() itself.
  template <typename F>
  void run(int parts, F&& fn) {
    if (parts <= 1) {
      if (parts == 1) fn(0);
      return;
    }
    std::lock_guard<std::mutex> turn(run_mu_);
    auto job = std::make_shared<Job>();
    job->fn = std::ref(fn);
    job->parts = parts;
    {
      std::lock_guard<std::mutex> lk(mu_);
      const int want = std::min(parts - 1, kMaxWorkers);
      while ((int)workers_ < want) {
        std::thread([this] { work(); }).detach();
        ++workers_;
      }
      cur_ = job;
      ++epoch_;
    }
    cv_work_.notify_all();
    for (int i; (i = job->next.fetch_add(1)) < parts;) {
      fn(i);
      job->done.fetch_add(1);
    }
    std::unique_lock<std::mutex> lk(mu_);
    cv_done_.wait(lk, [&] { return job->done.load() == parts; });
    cur_.reset();
  }

 private:
  static constexpr int kMaxWorkers = 15;
  struct Job {
    std::function<void(int)> fn;
    int parts = 0;
    std::atomic<int> next{0}, done{0};
  };
  void work() {
    uint64_t seen = 0;
    for (;;) {
      std::shared_ptr<Job> job;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_work_.wait(lk, [&] { return epoch_ != seen; });
        seen = epoch_;
        job = cur_;
      }
      if (!job) continue;
      for (int i; (i = job->next.fetch_add(1)) < job->parts;) {
        job->fn(i);
        if (job->done.fetch_add(1) + 1 == job->parts) {
          std::lock_guard<std::mutex> lk(mu_);   // the caller checks its predicate under mu_: no lost wake-up
          cv_done_.notify_all();
        }
      }
    }
  }
  std::mutex run_mu_, mu_;
  std::condition_variable cv_work_, cv_done_;
  std::shared_ptr<Job> cur_;
  uint64_t epoch_ = 0;
  size_t workers_ = 0;
};

}  // namespace gs
