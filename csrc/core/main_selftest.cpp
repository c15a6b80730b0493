// det-selftest — native concurrency self-test of the control-plane building blocks, meant to be run under
// ThreadSanitizer (`python -m determined_b200._build --tsan` builds build/det-selftest-tsan). The reference has no race
// detection at all (its Go tests run without -race); a C++ actor runtime does not get the Go runtime's
// safety net, so this binary hammers exactly the shared-state paths: mailboxes under concurrent tell/ask, timers,
// actor stop while messages are in flight, the reader/writer lease table, and the journal store.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#include "actor.hpp"
#include "rw_coordinator.hpp"
#include "scheduler.hpp"
#include "store.hpp"

using namespace det;

static int failures = 0;
#define EXPECT(cond)                                                              \
  do {                                                                            \
    if (!(cond)) { std::fprintf(stderr, "FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond); ++failures; } \
  } while (0)

class Counter : public Actor {
 public:
  void receive(Context& ctx, const Message& msg) override {
    if (msg.type == "inc") { total_ += msg.body.as_int(1); }
    else if (msg.type == "get") msg.respond(Json(total_));
    else if (msg.type == "spawn") {
      auto child = ctx.system().actor_of(ctx.self()->address() + "/c" + std::to_string(n_children_++), std::unique_ptr<Actor>(new Counter()), ctx.self());
      ctx.system().tell(child, "inc", Json(static_cast<long long>(1)));
      msg.respond(Json(true));
    } else if (msg.type == "timer") { ++timer_hits_; }
    else if (msg.type == "timer_hits") msg.respond(Json(timer_hits_));
  }
 private:
  long long total_ = 0, timer_hits_ = 0;
  int n_children_ = 0;
};

// An actor must never be inside receive() on two threads. Senders that let the mailbox run empty between messages hit the
// window where a worker has just given the cell up while a tell() re-queues it (the double-scheduling bug fixed in
// ActorSystem::worker: a cell queued twice ran on two workers at once).
class Serial : public Actor {
 public:
  void receive(Context&, const Message& msg) override {
    if (msg.type == "work") {
      if (busy_.exchange(true)) ++overlaps_;
      for (volatile int i = 0; i < 3000; ++i) {}      // long enough for a second worker to arrive while this one is inside
      ++seen_;
      busy_ = false;
    } else if (msg.type == "report") {
      Json r = Json::object();
      r["overlaps"] = static_cast<long long>(overlaps_.load());
      r["seen"] = seen_;
      msg.respond(r);
    }
  }
 private:
  std::atomic<bool> busy_{false};
  std::atomic<long long> overlaps_{0};
  long long seen_ = 0;
};

static void test_actor_is_never_run_concurrently() {
  ActorSystem::test_release_delay_us = 30;     // hold the give-up window open: a re-queue from inside it is then certain
  ActorSystem sys(8);
  auto a = sys.actor_of("serial", std::unique_ptr<Actor>(new Serial()));
  const int T = 8, N = 1500;
  std::vector<std::thread> th;
  for (int t = 0; t < T; ++t)
    th.emplace_back([&, t] {
      for (int i = 0; i < N; ++i) {
        sys.tell(a, "work");
        for (volatile int k = 0; k < 2000 + 700 * ((i + t) % 5); ++k) {}   // let the mailbox run empty between messages
      }
    });
  for (auto& t : th) t.join();
  Json r;
  for (int i = 0; i < 200; ++i) {
    r = sys.ask(a, "report");
    if (r["seen"].as_int() == static_cast<long long>(T) * N) break;
    std::this_thread::sleep_for(std::chrono::milliseconds(5));
  }
  EXPECT(r["seen"].as_int() == static_cast<long long>(T) * N);
  EXPECT(r["overlaps"].as_int() == 0);
  sys.shutdown();
  ActorSystem::test_release_delay_us = 0;
}

static void test_actors() {
  ActorSystem sys(8);
  auto a = sys.actor_of("counter", std::unique_ptr<Actor>(new Counter()));
  const int T = 8, N = 5000;
  std::vector<std::thread> th;
  for (int t = 0; t < T; ++t)
    th.emplace_back([&] {
      for (int i = 0; i < N; ++i) {
        sys.tell(a, "inc", Json(static_cast<long long>(1)));
        if (i % 500 == 0) EXPECT(sys.ask(a, "get").as_int(-1) >= 0);
      }
    });
  for (auto& t : th) t.join();
  EXPECT(sys.ask(a, "get").as_int() == static_cast<long long>(T) * N);   // every message delivered exactly once, in a serial Receive
  for (int i = 0; i < 20; ++i) sys.notify_after(a, 1 + i % 5, "timer");
  for (int i = 0; i < 50; ++i) EXPECT(sys.ask(a, "spawn").as_bool());
  std::this_thread::sleep_for(std::chrono::milliseconds(100));
  EXPECT(sys.ask(a, "timer_hits").as_int() == 20);
  EXPECT(sys.children_of("counter/").size() == 50);
  // stop the parent while other threads keep sending: no crash, asks come back (null) instead of hanging
  std::atomic<bool> go{true};
  std::thread spam([&] { while (go) { sys.tell(a, "inc", Json(static_cast<long long>(1))); sys.ask(a, "get", Json(), 50); } });
  std::this_thread::sleep_for(std::chrono::milliseconds(20));
  sys.stop(a);
  std::this_thread::sleep_for(std::chrono::milliseconds(50));
  go = false;
  spam.join();
  EXPECT(sys.get("counter") == nullptr);
  EXPECT(sys.children_of("counter/").empty());
  sys.shutdown();
}

static void test_rw() {
  RWCoordinator rw;
  std::atomic<int> readers{0}, writers{0}, violations{0}, grants{0};
  std::vector<std::thread> th;
  for (int t = 0; t < 12; ++t)
    th.emplace_back([&, t] {
      for (int i = 0; i < 200; ++i) {
        const bool read = (t % 3) != 0;
        auto g = rw.acquire("res", read, 2000, 60000);
        if (!g.granted) continue;
        ++grants;
        if (read) { ++readers; if (writers.load() != 0) ++violations; }
        else { if (++writers != 1 || readers.load() != 0) ++violations; }
        std::this_thread::yield();
        if (read) --readers; else --writers;
        EXPECT(rw.release(g.lease));
      }
    });
  for (auto& t : th) t.join();
  EXPECT(violations.load() == 0);
  EXPECT(grants.load() == 12 * 200);
  int r, w, q;
  rw.stats("res", r, w, q);
  EXPECT(r == 0 && w == 0 && q == 0);
}

static void test_store(const std::string& dir) {
  {
    Store s(dir);
    std::vector<std::thread> th;
    for (int t = 0; t < 6; ++t)
      th.emplace_back([&, t] {
        for (int i = 0; i < 300; ++i) {
          const long long id = s.next_id("rows");
          Json row = Json::object();
          row["t"] = static_cast<long long>(t); row["i"] = static_cast<long long>(i);
          s.put("rows", id, row);
          Json patch = Json::object();
          patch["seen"] = true;
          s.update("rows", id, patch);
          EXPECT(s.get("rows", id)["seen"].as_bool(false));
        }
      });
    for (auto& t : th) t.join();
  }
  Store again(dir);   // replay of the journal
  int n = 0;
  for (long long id = 1; id <= 6 * 300; ++id)
    if (again.get("rows", id)["seen"].as_bool(false)) ++n;
  EXPECT(n == 6 * 300);
}

int main(int argc, char** argv) {
  const std::string dir = argc > 1 ? argv[1] : "/tmp/det-selftest-store";
  test_actors();
  test_actor_is_never_run_concurrently();
  test_rw();
  test_store(dir);
  // a deterministic piece, to make sure the binary links the library properly
  EXPECT(md5_low64("determined") == md5_low64("determined"));
  if (failures) { std::fprintf(stderr, "%d failure(s)\n", failures); return 1; }
  std::puts("selftest ok");
  return 0;
}
