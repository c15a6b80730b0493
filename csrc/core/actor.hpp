// actor.hpp — hierarchical actor runtime: every master / agent component is an actor with a mailbox; one message
// at a time per actor (so actor state needs no locks), Tell / Ask, children with supervision callbacks, timers,
// stop notifications. Messages are (type, JSON body) pairs.
// Parity: reference master/pkg/actor/{system.go:10-110, ref.go, context.go, actors/timer.go, actors/stop.go}.
#pragma once
#include <atomic>
#include <cstdio>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <functional>
#include <future>
#include <map>
#include <memory>
#include <mutex>
#include <queue>
#include <string>
#include <thread>
#include <vector>

#include "json.hpp"

namespace det {

class ActorSystem;
class ActorCell;
using ActorRef = std::shared_ptr<ActorCell>;

struct Message {
  std::string type;
  Json body;
  ActorRef sender;
  std::shared_ptr<std::promise<Json>> reply;  // set for Ask
  void respond(const Json& j) const {
    if (reply) {
      try { reply->set_value(j); } catch (...) {}
    }
  }
};

class Context;

class Actor {
 public:
  virtual ~Actor() = default;
  virtual void pre_start(Context&) {}
  virtual void receive(Context& ctx, const Message& msg) = 0;
  virtual void post_stop(Context&) {}
  virtual void child_stopped(Context&, const ActorRef& /*child*/) {}
  virtual void child_failed(Context&, const ActorRef& /*child*/, const std::string& /*error*/) {}
};

class ActorCell : public std::enable_shared_from_this<ActorCell> {
 public:
  ActorCell(ActorSystem* sys, std::string address, std::unique_ptr<Actor> actor, ActorRef parent)
      : system_(sys), address_(std::move(address)), actor_(std::move(actor)), parent_(std::move(parent)),
        registered_(std::chrono::steady_clock::now()) {}
  const std::string& address() const { return address_; }
  std::chrono::steady_clock::time_point registered_time() const { return registered_; }
  bool stopped() const { return stopped_.load(); }
  ActorRef parent() const { return parent_; }

 private:
  friend class ActorSystem;
  friend class Context;
  ActorSystem* system_;
  std::string address_;
  std::unique_ptr<Actor> actor_;
  ActorRef parent_;
  std::chrono::steady_clock::time_point registered_;
  std::mutex mu_;
  std::deque<Message> mailbox_;
  bool scheduled_ = false;
  bool started_ = false;
  std::atomic<bool> stopped_{false};
  std::map<std::string, ActorRef> children_;
  std::vector<std::function<void()>> on_stop_;
};

class ActorSystem {
 public:
  explicit ActorSystem(int threads = 4) {
    for (int i = 0; i < threads; ++i) workers_.emplace_back([this] { worker(); });
    timer_thread_ = std::thread([this] { timer_loop(); });
  }
  ~ActorSystem() { shutdown(); }

  void shutdown() {
    {
      std::lock_guard<std::mutex> g(mu_);
      if (done_) return;
      done_ = true;
    }
    cv_.notify_all();
    timer_cv_.notify_all();
    for (auto& t : workers_) if (t.joinable()) t.join();
    if (timer_thread_.joinable()) timer_thread_.join();
  }

  ActorRef actor_of(const std::string& address, std::unique_ptr<Actor> actor, ActorRef parent = nullptr) {
    auto cell = std::make_shared<ActorCell>(this, address, std::move(actor), parent);
    {
      std::lock_guard<std::mutex> g(mu_);
      registry_[address] = cell;
    }
    if (parent) {
      std::lock_guard<std::mutex> g(parent->mu_);
      parent->children_[address] = cell;
    }
    tell(cell, Message{"__start__", Json(), nullptr, nullptr});
    return cell;
  }
  ActorRef get(const std::string& address) {
    std::lock_guard<std::mutex> g(mu_);
    auto it = registry_.find(address);
    return it == registry_.end() ? nullptr : it->second;
  }
  std::vector<ActorRef> children_of(const std::string& prefix) {
    std::lock_guard<std::mutex> g(mu_);
    std::vector<ActorRef> out;
    for (auto& kv : registry_)
      if (kv.first.size() > prefix.size() && kv.first.compare(0, prefix.size(), prefix) == 0) out.push_back(kv.second);
    return out;
  }

  void tell(const ActorRef& to, Message msg) {
    if (!to || to->stopped()) { msg.respond(Json()); return; }
    bool schedule = false;
    {
      std::lock_guard<std::mutex> g(to->mu_);
      to->mailbox_.push_back(std::move(msg));
      if (!to->scheduled_) { to->scheduled_ = true; schedule = true; }
    }
    if (schedule) {
      { std::lock_guard<std::mutex> g(mu_); run_queue_.push_back(to); }
      cv_.notify_one();
    }
  }
  void tell(const ActorRef& to, const std::string& type, const Json& body = Json(), ActorRef sender = nullptr) {
    tell(to, Message{type, body, std::move(sender), nullptr});
  }
  // Ask: returns the actor's response (null Json on timeout / stopped actor).
  Json ask(const ActorRef& to, const std::string& type, const Json& body = Json(), int timeout_ms = 10000) {
    if (!to || to->stopped()) return Json();
    auto p = std::make_shared<std::promise<Json>>();
    auto f = p->get_future();
    tell(to, Message{type, body, nullptr, p});
    if (f.wait_for(std::chrono::milliseconds(timeout_ms)) != std::future_status::ready) return Json();
    return f.get();
  }
  // Deliver `type` to `to` after `delay_ms` (actors.NotifyAfter).
  void notify_after(const ActorRef& to, int delay_ms, const std::string& type, const Json& body = Json()) {
    {
      std::lock_guard<std::mutex> g(timer_mu_);
      timers_.push(Timer{std::chrono::steady_clock::now() + std::chrono::milliseconds(delay_ms), to, type, body});
    }
    timer_cv_.notify_one();
  }
  void stop(const ActorRef& a) { tell(a, Message{"__stop__", Json(), nullptr, nullptr}); }

  // Tracing (role of the reference's OpenTracing spans around every Receive, master/pkg/actor/trace.go:26-106, which
  // it enables at link time): DET_ACTOR_TRACE=<file> or enable_trace(file) at run time makes every Receive a complete
  // event ("ph":"X") in Chrome trace-event JSON lines — load the file in chrome://tracing or Perfetto. Timer ticks are
  // filtered like the reference filters its noisy timers.
  void enable_trace(const std::string& path) {
    std::lock_guard<std::mutex> g(trace_mu_);
    if (trace_) fclose(trace_);
    trace_ = path.empty() ? nullptr : fopen(path.c_str(), "a");
    tracing_.store(trace_ != nullptr, std::memory_order_release);
    trace_origin_ = std::chrono::steady_clock::now();
  }
  // [{address, mailbox, children, stopped}] for /debug/actors
  Json describe_actors() {
    std::lock_guard<std::mutex> g(mu_);
    Json out = Json::array();
    for (auto& kv : registry_) {
      Json a = Json::object();
      a["address"] = kv.first;
      {
        std::lock_guard<std::mutex> g2(kv.second->mu_);
        a["mailbox"] = static_cast<long long>(kv.second->mailbox_.size());
        a["children"] = static_cast<long long>(kv.second->children_.size());
      }
      a["stopped"] = kv.second->stopped();
      out.push_back(a);
    }
    return out;
  }

 private:
  friend class Context;
  struct Timer {
    std::chrono::steady_clock::time_point at;
    ActorRef to;
    std::string type;
    Json body;
    bool operator<(const Timer& o) const { return at > o.at; }
  };
  void worker();
  void timer_loop() {
    std::unique_lock<std::mutex> lk(timer_mu_);
    while (true) {
      { std::lock_guard<std::mutex> g(mu_); if (done_) return; }
      if (timers_.empty()) { timer_cv_.wait_for(lk, std::chrono::milliseconds(200)); continue; }
      auto at = timers_.top().at;
      if (std::chrono::steady_clock::now() < at) { timer_cv_.wait_until(lk, at); continue; }
      Timer t = timers_.top();
      timers_.pop();
      lk.unlock();
      tell(t.to, t.type, t.body);
      lk.lock();
    }
  }
  void finish_stop(const ActorRef& cell);
  void trace_receive(const std::string& actor, const std::string& type, const std::string& sender,
                     std::chrono::steady_clock::time_point t0) {
    if (type == "tick" || type == "poll_timeout") return;   // periodic timers would drown everything else
    using namespace std::chrono;
    const auto t1 = steady_clock::now();
    Json e = Json::object();
    e["name"] = type; e["cat"] = "actor"; e["ph"] = "X"; e["pid"] = 1;
    e["tid"] = actor;
    e["ts"] = static_cast<long long>(duration_cast<microseconds>(t0 - trace_origin_).count());
    e["dur"] = static_cast<long long>(duration_cast<microseconds>(t1 - t0).count());
    Json args = Json::object();
    args["sender"] = sender;
    e["args"] = args;
    std::lock_guard<std::mutex> g(trace_mu_);
    if (trace_) { fputs((e.dump() + "\n").c_str(), trace_); fflush(trace_); }
  }

 public:
  // Test knob (det-selftest): a worker that has just given a cell up sleeps this long before it would look at the cell
  // again — which it must not do. 0 in production.
  static inline std::atomic<int> test_release_delay_us{0};

 private:
  std::mutex trace_mu_;
  FILE* trace_ = nullptr;                 // guarded by trace_mu_
  std::atomic<bool> tracing_{false};      // what the workers look at without the lock
  std::chrono::steady_clock::time_point trace_origin_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<ActorRef> run_queue_;
  std::map<std::string, ActorRef> registry_;
  std::vector<std::thread> workers_;
  bool done_ = false;
  std::mutex timer_mu_;
  std::condition_variable timer_cv_;
  std::priority_queue<Timer> timers_;
  std::thread timer_thread_;
};

class Context {
 public:
  Context(ActorSystem* s, ActorRef self) : system_(s), self_(std::move(self)) {}
  ActorSystem& system() { return *system_; }
  const ActorRef& self() const { return self_; }
  ActorRef actor_of(const std::string& name, std::unique_ptr<Actor> a) {
    return system_->actor_of(self_->address() + "/" + name, std::move(a), self_);
  }
  ActorRef child(const std::string& name) {
    std::lock_guard<std::mutex> g(self_->mu_);
    auto it = self_->children_.find(self_->address() + "/" + name);
    return it == self_->children_.end() ? nullptr : it->second;
  }
  std::vector<ActorRef> children() {
    std::lock_guard<std::mutex> g(self_->mu_);
    std::vector<ActorRef> out;
    for (auto& kv : self_->children_) out.push_back(kv.second);
    return out;
  }
  void tell(const ActorRef& to, const std::string& type, const Json& body = Json()) { system_->tell(to, type, body, self_); }
  void stop_self() { system_->stop(self_); }

 private:
  ActorSystem* system_;
  ActorRef self_;
};

inline void ActorSystem::worker() {
  while (true) {
    ActorRef cell;
    {
      std::unique_lock<std::mutex> lk(mu_);
      cv_.wait(lk, [this] { return done_ || !run_queue_.empty(); });
      if (done_ && run_queue_.empty()) return;
      cell = run_queue_.front();
      run_queue_.pop_front();
    }
    // drain a bounded batch so one chatty actor cannot starve the rest
    bool released = false;   // this worker gave the cell up (cleared scheduled_): from then on tell() owns its scheduling
    for (int n = 0; n < 64; ++n) {
      Message msg;
      {
        std::lock_guard<std::mutex> g(cell->mu_);
        if (cell->mailbox_.empty()) { cell->scheduled_ = false; released = true; break; }
        msg = std::move(cell->mailbox_.front());
        cell->mailbox_.pop_front();
      }
      Context ctx(this, cell);
      try {
        if (msg.type == "__start__") {
          cell->started_ = true;
          cell->actor_->pre_start(ctx);
        } else if (msg.type == "__stop__") {
          finish_stop(cell);
          break;
        } else if (msg.type == "__child_stopped__") {
          auto c = get(msg.body.as_string());
          cell->actor_->child_stopped(ctx, msg.sender);
        } else if (msg.type == "__child_failed__") {
          cell->actor_->child_failed(ctx, msg.sender, msg.body.as_string());
        } else if (!cell->stopped()) {
          if (tracing_.load(std::memory_order_acquire)) {
            const auto t0 = std::chrono::steady_clock::now();
            cell->actor_->receive(ctx, msg);
            trace_receive(cell->address(), msg.type, msg.sender ? msg.sender->address() : std::string(), t0);
          } else {
            cell->actor_->receive(ctx, msg);
          }
        } else {
          msg.respond(Json());
        }
      } catch (const std::exception& e) {
        msg.respond(Json());
        if (cell->parent_) tell(cell->parent_, Message{"__child_failed__", Json(std::string(e.what())), cell, nullptr});
        finish_stop(cell);
        break;
      }
    }
    // Re-queue only a cell this worker still owns (batch limit reached, stop, failure). After `released` a concurrent
    // tell() may already have set scheduled_ and queued the cell — touching it here again would put it on the run queue
    // twice and let two workers run the same actor at once (found by ThreadSanitizer under the chaos search: two threads in
    // one TrialActor::receive; it showed up as a rare "double free or corruption" of the master).
    if (released) {
      const int d = test_release_delay_us.load(std::memory_order_relaxed);    // selftest only: holds the window open
      if (d > 0) std::this_thread::sleep_for(std::chrono::microseconds(d));
    }
    bool again = false;
    if (!released) {
      std::lock_guard<std::mutex> g(cell->mu_);
      if (!cell->mailbox_.empty() && !cell->stopped()) { cell->scheduled_ = true; again = true; }
      else cell->scheduled_ = false;
    }
    if (again) {
      { std::lock_guard<std::mutex> g(mu_); run_queue_.push_back(cell); }
      cv_.notify_one();
    }
  }
}

inline void ActorSystem::finish_stop(const ActorRef& cell) {
  if (cell->stopped_.exchange(true)) return;
  Context ctx(this, cell);
  std::vector<ActorRef> kids;
  {
    std::lock_guard<std::mutex> g(cell->mu_);
    for (auto& kv : cell->children_) kids.push_back(kv.second);
  }
  for (auto& k : kids) finish_stop(k);
  try { cell->actor_->post_stop(ctx); } catch (...) {}
  std::deque<Message> leftover;
  {
    std::lock_guard<std::mutex> g(cell->mu_);
    leftover.swap(cell->mailbox_);
  }
  for (auto& m : leftover) m.respond(Json());
  {
    std::lock_guard<std::mutex> g(mu_);
    registry_.erase(cell->address());
  }
  if (cell->parent_) {
    { std::lock_guard<std::mutex> g(cell->parent_->mu_); cell->parent_->children_.erase(cell->address()); }
    tell(cell->parent_, Message{"__child_stopped__", Json(cell->address()), cell, nullptr});
  }
}

}  // namespace det
