// A small persistent thread pool for the host-buffer envs (csrc/host_env.cu, csrc/tfe_host_env.cu): run(fn) executes
// fn(worker) on every worker, the calling thread being worker 0, and returns when all are done.
#pragma once
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace pulse_host {

class Pool {
 public:
  explicit Pool(int n) : n_(n < 1 ? 1 : n) {
    for (int i = 1; i < n_; ++i) th_.emplace_back([this, i] { loop(i); });
  }
  ~Pool() {
    { std::lock_guard<std::mutex> lk(m_); stop_ = true; ++gen_; }
    cv_.notify_all();
    for (auto& t : th_) t.join();
  }
  int size() const { return n_; }
  // fn(worker) on every worker, the caller being worker 0; returns when all are done
  void run(const std::function<void(int)>& fn) {
    if (n_ == 1) { fn(0); return; }
    { std::lock_guard<std::mutex> lk(m_); fn_ = &fn; pending_ = n_ - 1; ++gen_; }
    cv_.notify_all();
    fn(0);
    std::unique_lock<std::mutex> lk(m_);
    done_cv_.wait(lk, [this] { return pending_ == 0; });
  }

 private:
  void loop(int id) {
    unsigned long long seen = 0;
    for (;;) {
      const std::function<void(int)>* fn;
      {
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [&] { return gen_ != seen; });
        seen = gen_;
        if (stop_) return;
        fn = fn_;
      }
      (*fn)(id);
      { std::lock_guard<std::mutex> lk(m_); if (--pending_ == 0) done_cv_.notify_one(); }
    }
  }
  int n_;
  std::vector<std::thread> th_;
  std::mutex m_;
  std::condition_variable cv_, done_cv_;
  const std::function<void(int)>* fn_ = nullptr;
  unsigned long long gen_ = 0;
  int pending_ = 0;
  bool stop_ = false;
};


}  // namespace pulse_host
