// workqueue.h — the queue between watch events and the reconcilers.
//
// The guarantees the reference's reconcilers rely on come from controller-runtime's
// workqueue (sigs.k8s.io/controller-runtime v0.17.0 over client-go's workqueue; neither is
// vendored): a key waits in the queue at most once however often it
// is added; a key that is being reconciled is not handed out again — an Add in the
// meantime marks it dirty and it comes back when the reconcile is Done; `RequeueAfter`
// (culling_controller.go:143,168) re-adds a key when its time has come; an error re-adds it
// with the per-key back-off 5 ms * 2^failures, capped at 1000 s, which Forget resets.
// The clock is the caller's (seconds, any origin): tests and simulations drive it.
//
// Batching: Take hands out up to `max` ready keys at once — the reconcilers here work on
// batches — and every one of them is "processing" until Done.
#pragma once
#include <algorithm>
#include <deque>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <vector>

#include "notebook_reconciler.h"

namespace nbrm_host {

class WorkQueue {
 public:
  void Add(const Request& r) {
    std::lock_guard<std::mutex> l(mu_);
    add_locked(key(r));
  }
  void AddAfter(const Request& r, double now, double seconds) {
    std::lock_guard<std::mutex> l(mu_);
    if (seconds <= 0) return add_locked(key(r));
    waiting_.insert({now + seconds, key(r)});
  }
  // the back-off a failed reconcile gets; returns the delay used
  double AddRateLimited(const Request& r, double now) {
    std::lock_guard<std::mutex> l(mu_);
    const int n = failures_[key(r)]++;
    double d = 0.005;
    for (int k = 0; k < n && d < 1000.0; ++k) d *= 2;
    d = std::min(d, 1000.0);
    waiting_.insert({now + d, key(r)});
    return d;
  }
  void Forget(const Request& r) {
    std::lock_guard<std::mutex> l(mu_);
    failures_.erase(key(r));
  }
  int NumRequeues(const Request& r) {
    std::lock_guard<std::mutex> l(mu_);
    auto it = failures_.find(key(r));
    return it == failures_.end() ? 0 : it->second;
  }
  // Up to `max` keys that are ready at `now`; each is processing until Done.
  size_t Take(size_t max, double now, std::vector<Request>* out) {
    std::lock_guard<std::mutex> l(mu_);
    while (!waiting_.empty() && waiting_.begin()->first <= now) {
      add_locked(waiting_.begin()->second);
      waiting_.erase(waiting_.begin());
    }
    size_t n = 0;
    while (n < max && !queue_.empty()) {
      const std::string k = queue_.front();
      queue_.pop_front();
      dirty_.erase(k);
      processing_.insert(k);
      out->push_back(unkey(k));
      ++n;
    }
    return n;
  }
  void Done(const Request& r) {
    std::lock_guard<std::mutex> l(mu_);
    const std::string k = key(r);
    processing_.erase(k);
    if (dirty_.count(k)) queue_.push_back(k);  // added again while it was being reconciled
  }
  size_t Len() {
    std::lock_guard<std::mutex> l(mu_);
    return queue_.size();
  }
  size_t Waiting() {
    std::lock_guard<std::mutex> l(mu_);
    return waiting_.size();
  }
  // when the earliest delayed key becomes ready (< 0: none waits)
  double NextReady() {
    std::lock_guard<std::mutex> l(mu_);
    return waiting_.empty() ? -1.0 : waiting_.begin()->first;
  }

 private:
  static std::string key(const Request& r) { return r.ns + "/" + r.name; }
  static Request unkey(const std::string& k) {
    const size_t s = k.find('/');
    return Request{k.substr(0, s), k.substr(s + 1)};
  }
  void add_locked(const std::string& k) {
    if (dirty_.count(k)) return;  // already waiting to be handed out
    dirty_.insert(k);
    if (processing_.count(k)) return;  // comes back at Done
    queue_.push_back(k);
  }
  std::mutex mu_;
  std::deque<std::string> queue_;
  std::set<std::string> dirty_, processing_;
  std::multiset<std::pair<double, std::string>> waiting_;
  std::map<std::string, int> failures_;
};

}  // namespace nbrm_host
