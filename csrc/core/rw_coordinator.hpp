// rw_coordinator.hpp — named reader/writer locks served by the master, used by the data-layer cache so that one
// process builds a cached data set while the others wait and then read it.
//
// Role of the reference's rwCoordinator actor (master/internal/rw_coordinator.go:1-174: a WebSocket per lock holder
// on /ws/data-layer/{resource}?read_lock=..., FIFO fairness, lock released when the socket closes). Here a lock is a
// LEASE obtained over HTTP long-poll: acquire blocks until granted (or the poll times out), the holder releases it
// explicitly, and a lease that is neither released nor renewed within its TTL is reclaimed — the equivalent of the
// socket dropping when a holder dies.
#pragma once
#include <chrono>
#include <condition_variable>
#include <deque>
#include <map>
#include <mutex>
#include <string>

namespace det {

class RWCoordinator {
 public:
  struct Grant {
    bool granted = false;
    std::string lease;
  };

  // Blocks up to wait_ms for a read (shared) or write (exclusive) lock on `resource`. FIFO: a reader that arrives
  // after a waiting writer queues behind it, so writers are not starved.
  Grant acquire(const std::string& resource, bool read_lock, int wait_ms, int ttl_ms) {
    using clock = std::chrono::steady_clock;
    std::unique_lock<std::mutex> lk(mu_);
    Res& r = res_[resource];
    const long long ticket = next_ticket_++;
    r.queue.push_back({ticket, read_lock});
    const auto deadline = clock::now() + std::chrono::milliseconds(wait_ms);
    for (;;) {
      reap(r);
      if (can_enter(r, ticket, read_lock)) {
        remove_ticket(r, ticket);
        Grant g;
        g.granted = true;
        g.lease = resource + "#" + std::to_string(ticket);
        r.holders[g.lease] = {read_lock, clock::now() + std::chrono::milliseconds(ttl_ms)};
        if (read_lock) ++r.readers; else r.writer = true;
        cv_.notify_all();  // the next reader in the queue may be admissible too
        return g;
      }
      if (cv_.wait_until(lk, std::min(deadline, clock::now() + std::chrono::milliseconds(200))) == std::cv_status::timeout &&
          clock::now() >= deadline) {
        remove_ticket(r, ticket);
        cv_.notify_all();
        return Grant{};
      }
    }
  }

  bool renew(const std::string& lease, int ttl_ms) {
    std::lock_guard<std::mutex> lk(mu_);
    for (auto& kv : res_) {
      auto it = kv.second.holders.find(lease);
      if (it != kv.second.holders.end()) {
        it->second.expires = std::chrono::steady_clock::now() + std::chrono::milliseconds(ttl_ms);
        return true;
      }
    }
    return false;
  }

  bool release(const std::string& lease) {
    std::lock_guard<std::mutex> lk(mu_);
    for (auto& kv : res_) {
      auto it = kv.second.holders.find(lease);
      if (it != kv.second.holders.end()) {
        drop(kv.second, it);
        cv_.notify_all();
        return true;
      }
    }
    return false;
  }

  // {readers, writer(0/1), waiting}
  void stats(const std::string& resource, int& readers, int& writer, int& waiting) {
    std::lock_guard<std::mutex> lk(mu_);
    Res& r = res_[resource];
    reap(r);
    readers = r.readers; writer = r.writer ? 1 : 0; waiting = static_cast<int>(r.queue.size());
  }

 private:
  struct Holder { bool read; std::chrono::steady_clock::time_point expires; };
  struct Waiter { long long ticket; bool read; };
  struct Res {
    int readers = 0;
    bool writer = false;
    std::deque<Waiter> queue;
    std::map<std::string, Holder> holders;
  };

  static bool can_enter(const Res& r, long long ticket, bool read_lock) {
    if (r.writer) return false;
    if (!read_lock) return r.readers == 0 && !r.queue.empty() && r.queue.front().ticket == ticket;
    for (auto& w : r.queue) {         // a reader may pass only other readers ahead of it
      if (w.ticket == ticket) return true;
      if (!w.read) return false;
    }
    return false;
  }
  static void remove_ticket(Res& r, long long ticket) {
    for (auto it = r.queue.begin(); it != r.queue.end(); ++it)
      if (it->ticket == ticket) { r.queue.erase(it); return; }
  }
  static void drop(Res& r, std::map<std::string, Holder>::iterator it) {
    if (it->second.read) --r.readers; else r.writer = false;
    r.holders.erase(it);
  }
  void reap(Res& r) {
    const auto now = std::chrono::steady_clock::now();
    for (auto it = r.holders.begin(); it != r.holders.end();) {
      auto cur = it++;
      if (cur->second.expires <= now) drop(r, cur);
    }
  }

  std::mutex mu_;
  std::condition_variable cv_;
  std::map<std::string, Res> res_;
  long long next_ticket_ = 1;
};

}  // namespace det
