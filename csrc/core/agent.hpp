// agent.hpp — node-agent machinery: device detection (nvidia-smi, CPU fallback, artificial slots for tests), a
// process-based "container" manager (fork/exec per task with injected DET_* environment and GPU pinning through
// CUDA_VISIBLE_DEVICES, stdout/stderr capture with [rank=N] parsing, signal / kill, state reporting
// Assigned→Pulling→Starting→Running→Terminated), and an HTTP readiness checker.
// Parity: reference agent/internal/{detect.go:20-115, nvidia.go:17-41, containers.go:145-213, container.go:29-238,
// docker.go:28-314, fluent.go (log shipping), check.go:20-112}. Docker is replaced by plain processes: on one
// 8-GPU box the isolation unit is the process + its GPU mask, which removes image pulls and container start-up from
// the trial launch path.
#pragma once
#include <sys/stat.h>
#include <limits.h>
#include <fcntl.h>
#include <grp.h>
#include <signal.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <cstdio>
#include <cstring>
#include <functional>
#include <map>
#include <memory>
#include <condition_variable>
#include <mutex>
#include <regex>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "json.hpp"

namespace det {

struct Device {
  int id = 0;
  std::string brand, uuid, type;  // type: "gpu" | "cpu"
  Json to_json() const {
    Json j = Json::object();
    j["id"] = id; j["brand"] = brand; j["uuid"] = uuid; j["type"] = type;
    return j;
  }
};

inline std::string run_capture(const std::string& cmd) {
  std::string out;
  FILE* p = popen((cmd + " 2>/dev/null").c_str(), "r");
  if (!p) return out;
  char buf[4096];
  size_t n;
  while ((n = fread(buf, 1, sizeof buf, p)) > 0) out.append(buf, n);
  pclose(p);
  return out;
}

inline std::string nvidia_driver_version() {
  std::string v = run_capture("nvidia-smi --query-gpu=driver_version --format=csv,noheader");
  auto nl = v.find('\n');
  return nl == std::string::npos ? v : v.substr(0, nl);
}

// slot_type: "auto" | "gpu" | "cpu" | "none"; visible_gpus: comma list of indices/UUIDs ("" = all);
// artificial_slots > 0 fabricates that many slots (scheduling tests without hardware).
inline std::vector<Device> detect_devices(const std::string& slot_type = "auto", const std::string& visible_gpus = "", int artificial_slots = 0) {
  std::vector<Device> out;
  if (artificial_slots > 0) {
    for (int i = 0; i < artificial_slots; ++i) out.push_back(Device{i, "Artificial", "artificial-" + std::to_string(i), "gpu"});
    return out;
  }
  if (slot_type == "none") return out;
  if (slot_type == "auto" || slot_type == "gpu") {
    std::string cmd = "nvidia-smi --query-gpu=index,name,uuid --format=csv,noheader";
    if (!visible_gpus.empty()) cmd += " --id=" + visible_gpus;
    std::istringstream ss(run_capture(cmd));
    std::string line;
    while (std::getline(ss, line)) {
      std::vector<std::string> f;
      std::stringstream ls(line);
      std::string item;
      while (std::getline(ls, item, ',')) {
        while (!item.empty() && item.front() == ' ') item.erase(item.begin());
        while (!item.empty() && (item.back() == ' ' || item.back() == '\r')) item.pop_back();
        f.push_back(item);
      }
      if (f.size() == 3) out.push_back(Device{atoi(f[0].c_str()), f[1], f[2], "gpu"});
    }
    if (!out.empty() || slot_type == "gpu") return out;
  }
  // CPU fallback: one slot standing for the host's cores
  std::string model = "cpu";
  std::istringstream ci(run_capture("grep -m1 'model name' /proc/cpuinfo"));
  std::string l;
  if (std::getline(ci, l)) { auto c = l.find(':'); if (c != std::string::npos) model = l.substr(c + 2); }
  out.push_back(Device{0, model, "cpu-0", "cpu"});
  return out;
}

struct ContainerSpec {
  std::string container_id, task_id, description;
  std::vector<std::string> argv;
  std::map<std::string, std::string> env;
  std::string cwd;
  std::vector<Device> devices;
};

struct ContainerEvent {
  std::string container_id, state;  // ASSIGNED | PULLING | STARTING | RUNNING | TERMINATED
  int exit_code = 0;
  std::string failure;  // "" on clean exit
};

struct LogLine {
  std::string container_id, stdtype, message;
  int rank = -1;
};

// bind_mounts of a process "container" (see task_spec.hpp): every `host|container|mode` line makes the container path point
// at the host path. Runs in the forked child, in the task's working directory, right before exec.
inline bool apply_bind_mounts(const std::string& lines) {
  size_t pos = 0;
  while (pos < lines.size()) {
    size_t nl = lines.find('\n', pos);
    if (nl == std::string::npos) nl = lines.size();
    const std::string line = lines.substr(pos, nl - pos);
    pos = nl + 1;
    const size_t a = line.find('|'), b = line.rfind('|');
    if (line.empty() || a == std::string::npos || b == a) continue;
    const std::string host = line.substr(0, a), target = line.substr(a + 1, b - a - 1);
    char have[4096], want[4096];
    if (realpath(host.c_str(), want) == nullptr) { fprintf(stderr, "bind mount: host path %s does not exist\n", host.c_str()); return false; }
    if (realpath(target.c_str(), have) != nullptr) {
      if (strcmp(have, want) == 0) continue;                       // already the same place
      fprintf(stderr, "bind mount: %s exists and is not %s\n", target.c_str(), host.c_str());
      return false;
    }
    const size_t slash = target.find_last_of('/');
    if (slash != std::string::npos && slash > 0) {
      std::string dir = target.substr(0, slash);
      for (size_t i = 1; i <= dir.size(); ++i)
        if (i == dir.size() || dir[i] == '/') ::mkdir(dir.substr(0, i).c_str(), 0755);
    }
    ::unlink(target.c_str());   // a dangling link left by an earlier run
    if (::symlink(want, target.c_str()) != 0) { perror(("bind mount: symlink " + target).c_str()); return false; }
  }
  return true;
}

class ContainerManager {
 public:
  using EventFn = std::function<void(const ContainerEvent&)>;
  using LogFn = std::function<void(const LogLine&)>;
  ContainerManager(EventFn on_event, LogFn on_log) : on_event_(std::move(on_event)), on_log_(std::move(on_log)) {}
  ~ContainerManager() { kill_all(); }

  // Warm workers: a Python process per slot that has already imported torch and the package, created its CUDA context and
  // loaded libdet_kernels.so, and now waits on stdin for a task. A single-slot trial assigned to that slot is handed to it
  // (environment, working directory) instead of being forked cold — trial start latency is then the trial's own set-up.
  // The reference pays container start + pip install + sshd per trial (master/static/srv/entrypoint.sh:1-47).
  // mode: "off"; "gpu" (default: a GPU slot gets its warm worker after its first cold start); "lazy" (the same for CPU /
  // artificial slots too); "eager" (as lazy, and the owner calls prewarm() for every slot up front). The interpreter and
  // PYTHONPATH are given, or learnt from the first eligible container spec.
  void enable_warm(const std::string& mode, const std::string& python, const std::string& pythonpath) {
    std::lock_guard<std::mutex> g(mu_);
    warm_mode_ = mode; warm_python_ = python; warm_path_ = pythonpath;
  }
  // Byte-code cache for every Python process this manager starts. Images that set PYTHONDONTWRITEBYTECODE and ship no
  // __pycache__ make each process compile torch's sources again (measured: `import torch` 6.0 s, 1.9 s from a cache).
  // How many slots this manager's node runs tasks on: task processes get DET_AGENT_SLOT_COUNT and size their CPU thread
  // pools to cores / slots (eight trials x one OpenMP team per core each, spinning between parallel regions, is what made a
  // 0.3 s workload take 7 s on the 8-GPU box).
  void set_slot_count(int n) {
    std::lock_guard<std::mutex> g(mu_);
    slot_count_ = n;
  }
  void set_pycache_dir(const std::string& dir) {
    std::lock_guard<std::mutex> g(mu_);
    pycache_dir_ = dir;
  }
  // Environment every task on this node starts with unless its own spec sets the variable (the agent's http_proxy /
  // https_proxy / ftp_proxy / no_proxy options, reference agent/internal/options.go).
  void set_task_env(const std::map<std::string, std::string>& env) {
    std::lock_guard<std::mutex> g(mu_);
    task_env_ = env;
  }
  void prewarm(const std::string& key) {
    {
      std::lock_guard<std::mutex> g(mu_);
      if (warm_mode_ == "off" || warm_mode_.empty() || warm_python_.empty()) return;
      if (warm_live_[key] > 0) return;            // this slot already has a worker, idle or serving a trial
      ++warm_live_[key];
    }
    int in_pipe[2], out_pipe[2], err_pipe[2], ctl_pipe[2];
    if (pipe(in_pipe) != 0 || pipe(out_pipe) != 0 || pipe(err_pipe) != 0 || pipe(ctl_pipe) != 0) {
      std::lock_guard<std::mutex> g(mu_);
      --warm_live_[key];
      return;
    }
    pid_t pid = fork();
    if (pid < 0) {
      std::lock_guard<std::mutex> g(mu_);
      --warm_live_[key];
      return;
    }
    if (pid == 0) {
      setpgid(0, 0);
      dup2(in_pipe[0], 0); dup2(out_pipe[1], 1); dup2(err_pipe[1], 2);
      if (ctl_pipe[1] != 3) dup2(ctl_pipe[1], 3);         // control channel: "ready", "done <rc>" (exec/warm_worker.py)
      for (int fd : {in_pipe[0], in_pipe[1], out_pipe[0], out_pipe[1], err_pipe[0], err_pipe[1], ctl_pipe[0], ctl_pipe[1]})
        if (fd > 3) close(fd);
      setenv("CUDA_VISIBLE_DEVICES", key == "cpu" ? "" : key.c_str(), 1);
      setenv("PYTHONUNBUFFERED", "1", 1);
      setenv("DET_WARM_CONTROL_FD", "3", 1);
      if (slot_count_ > 0) setenv("DET_AGENT_SLOT_COUNT", std::to_string(slot_count_).c_str(), 1);
      apply_pycache_env();
      const char* pp = getenv("PYTHONPATH");
      if (!warm_path_.empty()) setenv("PYTHONPATH", (warm_path_ + (pp ? std::string(":") + pp : "")).c_str(), 1);
      execlp(warm_python_.c_str(), warm_python_.c_str(), "-m", "determined_b200.exec.warm_worker", (char*)nullptr);
      _exit(127);
    }
    close(in_pipe[0]); close(out_pipe[1]); close(err_pipe[1]); close(ctl_pipe[1]);
    for (int fd : {in_pipe[1], out_pipe[0], err_pipe[0], ctl_pipe[0]}) fcntl(fd, F_SETFD, FD_CLOEXEC);
    auto w = std::make_shared<WarmWorker>();
    w->pid = pid; w->in_fd = in_pipe[1]; w->key = key;
    {
      std::lock_guard<std::mutex> g(mu_);
      warm_[key] = w;
      workers_.push_back(w);
    }
    std::thread([this, w, fd = out_pipe[0]] { pump_worker(w, fd, "stdout"); }).detach();
    std::thread([this, w, fd = err_pipe[0]] { pump_worker(w, fd, "stderr"); }).detach();
    std::thread([this, w, fd = ctl_pipe[0]] { control_loop(w, fd); }).detach();
  }
  size_t warm_count() {
    std::lock_guard<std::mutex> g(mu_);
    return warm_.size();
  }

  bool start(const ContainerSpec& spec) {
    on_event_({spec.container_id, "ASSIGNED", 0, ""});
    on_event_({spec.container_id, "PULLING", 0, ""});   // no image to pull: state kept for protocol parity
    on_event_({spec.container_id, "STARTING", 0, ""});
    std::map<std::string, std::string> env = spec.env;
    {
      std::lock_guard<std::mutex> g(mu_);
      for (auto& kv : task_env_) env.emplace(kv.first, kv.second);      // the spec's own value wins
    }
    env["DET_CONTAINER_ID"] = spec.container_id;
    std::string slot_ids = "[", gpu_mask;
    bool use_gpu = false;
    for (size_t i = 0; i < spec.devices.size(); ++i) {
      slot_ids += (i ? "," : "") + std::to_string(spec.devices[i].id);
      if (spec.devices[i].type == "gpu" && spec.devices[i].brand != "Artificial") { use_gpu = true; gpu_mask += (gpu_mask.empty() ? "" : ",") + std::to_string(spec.devices[i].id); }
    }
    env["DET_SLOT_IDS"] = slot_ids + "]";
    env["DET_USE_GPU"] = use_gpu ? "true" : "false";
    if (slot_count_ > 0) env["DET_AGENT_SLOT_COUNT"] = std::to_string(slot_count_);
    if (use_gpu) env["CUDA_VISIBLE_DEVICES"] = gpu_mask;
    else if (!spec.devices.empty() || env.count("DET_FORCE_CPU")) env["CUDA_VISIBLE_DEVICES"] = "";
    {
      struct timespec ts;
      clock_gettime(CLOCK_REALTIME, &ts);
      env["DET_CONTAINER_START_TS"] = std::to_string(static_cast<double>(ts.tv_sec) + ts.tv_nsec * 1e-9);
    }
    const std::string warm_key = use_gpu ? gpu_mask : "cpu";
    const bool warmable = warm_eligible(spec, env);
    pid_t pid = -1;
    int out_pipe[2] = {-1, -1}, err_pipe[2] = {-1, -1};
    if (warmable) {
      std::shared_ptr<WarmWorker> w;
      {
        std::lock_guard<std::mutex> g(mu_);
        auto it = warm_.find(warm_key);
        if (it != warm_.end()) {
          w = it->second;
          warm_.erase(it);
          if (::kill(w->pid, 0) != 0) w.reset();      // died while idle; its control loop cleans up
        }
        if (w) {
          std::lock_guard<std::mutex> gw(w->mu);
          w->cid = spec.container_id;
          pids_[spec.container_id] = w->pid;
        }
      }
      if (w) {
        Json task = Json::object();
        Json e = Json::object();
        for (auto& kv : env) e[kv.first] = kv.second;
        task["env"] = e;
        task["cwd"] = spec.cwd;
        std::string line = task.dump() + "\n";
        size_t off = 0;
        bool ok = true;
        while (off < line.size()) {
          ssize_t n = write(w->in_fd, line.data() + off, line.size() - off);
          if (n <= 0) { ok = false; break; }
          off += static_cast<size_t>(n);
        }
        if (ok) {
          // the worker's own pump and control threads already run: its output is attributed to this container from now
          // on, and "done" / process exit produce the TERMINATED event
          on_event_({spec.container_id, "RUNNING", 0, ""});
          return true;
        }
        {
          std::lock_guard<std::mutex> g(mu_);
          std::lock_guard<std::mutex> gw(w->mu);
          w->cid.clear();
          pids_.erase(spec.container_id);
        }
        ::kill(-w->pid, SIGKILL);
      }
    }
    if (pid < 0) {
    if (pipe(out_pipe) != 0 || pipe(err_pipe) != 0) { on_event_({spec.container_id, "TERMINATED", 127, "pipe() failed"}); return false; }
    pid = fork();
    if (pid < 0) { on_event_({spec.container_id, "TERMINATED", 127, "fork() failed"}); return false; }
    if (pid == 0) {
      setpgid(0, 0);
      dup2(out_pipe[1], 1);
      dup2(err_pipe[1], 2);
      close(out_pipe[0]); close(out_pipe[1]); close(err_pipe[0]); close(err_pipe[1]);
      if (!spec.cwd.empty() && chdir(spec.cwd.c_str()) != 0) { perror("chdir"); _exit(126); }
      if (!env.count("PYTHONPYCACHEPREFIX") && !env.count("PYTHONDONTWRITEBYTECODE")) apply_pycache_env();
      for (auto& kv : env) setenv(kv.first.c_str(), kv.second.c_str(), 1);
      {
        auto bm = env.find("DET_BIND_MOUNTS");
        if (bm != env.end() && !apply_bind_mounts(bm->second)) _exit(125);
      }
      if (!drop_to_task_user(env)) _exit(124);
      std::vector<char*> argv;
      for (auto& a : spec.argv) argv.push_back(const_cast<char*>(a.c_str()));
      argv.push_back(nullptr);
      execvp(argv[0], argv.data());
      perror("execvp");
      _exit(127);
    }
    close(out_pipe[1]);
    close(err_pipe[1]);
    }
    if (warmable) std::thread([this, warm_key] { prewarm(warm_key); }).detach();   // the NEXT trial on this slot
    {
      std::lock_guard<std::mutex> g(mu_);
      pids_[spec.container_id] = pid;
    }
    on_event_({spec.container_id, "RUNNING", 0, ""});
    const std::string cid = spec.container_id;
    // The exit is reported after the task's last output: the reaper waits for both pumps to reach end-of-file (bounded — a
    // grandchild that kept the pipe open must not hold the exit back) so "TERMINATED" never overtakes the final log lines.
    auto drain = std::make_shared<Drain>();
    std::thread([this, cid, drain, fd = out_pipe[0]] { pump(cid, fd, "stdout"); drain->done(); }).detach();
    std::thread([this, cid, drain, fd = err_pipe[0]] { pump(cid, fd, "stderr"); drain->done(); }).detach();
    std::thread([this, cid, pid, drain] {
      int status = 0;
      waitpid(pid, &status, 0);
      drain->wait_ms(2000);
      ContainerEvent ev{cid, "TERMINATED", 0, ""};
      if (WIFEXITED(status)) { ev.exit_code = WEXITSTATUS(status); if (ev.exit_code) ev.failure = "container failed with non-zero exit code: " + std::to_string(ev.exit_code); }
      else if (WIFSIGNALED(status)) { ev.exit_code = 128 + WTERMSIG(status); ev.failure = std::string("container killed by signal ") + strsignal(WTERMSIG(status)); }
      { std::lock_guard<std::mutex> g(mu_); pids_.erase(cid); }
      on_event_(ev);
    }).detach();
    return true;
  }
  void signal(const std::string& container_id, int sig) {
    std::lock_guard<std::mutex> g(mu_);
    auto it = pids_.find(container_id);
    if (it != pids_.end()) ::kill(-it->second, sig);  // whole process group (the launcher's worker ranks too)
  }
  void kill_all() {
    std::vector<std::shared_ptr<WarmWorker>> ws;
    {
      std::lock_guard<std::mutex> g(mu_);
      for (auto& kv : pids_) ::kill(-kv.second, SIGKILL);
      warm_mode_ = "off";
      ws = workers_;
      warm_.clear();
    }
    for (auto& w : ws) ::kill(-w->pid, SIGKILL);
  }
  size_t running() {
    std::lock_guard<std::mutex> g(mu_);
    return pids_.size();
  }

 private:
  struct Drain {
    std::mutex mu;
    std::condition_variable cv;
    int open = 2;
    void done() { std::lock_guard<std::mutex> g(mu); if (--open == 0) cv.notify_all(); }
    void wait_ms(int ms) { std::unique_lock<std::mutex> g(mu); cv.wait_for(g, std::chrono::milliseconds(ms), [this] { return open == 0; }); }
  };
  void pump(const std::string& cid, int fd, const std::string& stdtype) {
    static const std::regex rank_re("^\\[rank=(\\d+)\\] ?(.*)$");
    std::string pending;
    char buf[4096];
    ssize_t n;
    while ((n = read(fd, buf, sizeof buf)) > 0) {
      pending.append(buf, static_cast<size_t>(n));
      size_t nl;
      while ((nl = pending.find('\n')) != std::string::npos) {
        std::string line = pending.substr(0, nl);
        pending.erase(0, nl + 1);
        LogLine l{cid, stdtype, line, -1};
        std::smatch m;
        if (std::regex_match(line, m, rank_re)) { l.rank = atoi(m[1].str().c_str()); l.message = m[2].str(); }
        on_log_(l);
      }
    }
    if (!pending.empty()) on_log_(LogLine{cid, stdtype, pending, -1});
    close(fd);
  }
  // A pre-started harness process bound to one slot. It serves one trial at a time and, when the trial ends cleanly,
  // goes back to the pool with its interpreter, imports, CUDA context and kernel library intact.
  struct WarmWorker {
    pid_t pid = -1;
    int in_fd = -1;
    std::string key;
    std::mutex mu;
    std::string cid;      // container being served ("" while idle)
  };
  void emit_log(const std::string& cid, const std::string& stdtype, const std::string& line) {
    static const std::regex rank_re("^\\[rank=(\\d+)\\] ?(.*)$");
    LogLine l{cid, stdtype, line, -1};
    std::smatch m;
    if (std::regex_match(line, m, rank_re)) { l.rank = atoi(m[1].str().c_str()); l.message = m[2].str(); }
    on_log_(l);
  }
  void pump_worker(std::shared_ptr<WarmWorker> w, int fd, const std::string& stdtype) {
    std::string pending;
    char buf[4096];
    ssize_t n;
    while ((n = read(fd, buf, sizeof buf)) > 0) {
      pending.append(buf, static_cast<size_t>(n));
      size_t nl;
      while ((nl = pending.find('\n')) != std::string::npos) {
        std::string line = pending.substr(0, nl);
        pending.erase(0, nl + 1);
        std::string cid;
        { std::lock_guard<std::mutex> g(w->mu); cid = w->cid; }
        if (!cid.empty()) emit_log(cid, stdtype, line);     // output while idle (preload chatter) belongs to no container
      }
    }
    close(fd);
  }
  void control_loop(std::shared_ptr<WarmWorker> w, int fd) {
    FILE* f = fdopen(fd, "r");
    char line[256];
    while (f && fgets(line, sizeof line, f)) {
      if (strncmp(line, "done", 4) != 0) continue;           // "ready" needs no action: a task sent early just waits in the pipe
      const int rc = atoi(line + 4);
      usleep(50 * 1000);                                     // let the pumps drain what the trial wrote before its last line
      std::string cid;
      { std::lock_guard<std::mutex> g(w->mu); cid = w->cid; w->cid.clear(); }
      if (cid.empty()) continue;
      {
        std::lock_guard<std::mutex> g(mu_);
        pids_.erase(cid);
        if (rc == 0 && warm_mode_ != "off") warm_[w->key] = w;   // back in the pool BEFORE the event: the next trial finds it
      }
      ContainerEvent ev{cid, "TERMINATED", rc, ""};
      if (rc) ev.failure = "container failed with non-zero exit code: " + std::to_string(rc);
      on_event_(ev);
    }
    if (f) fclose(f); else close(fd);
    int status = 0;
    waitpid(w->pid, &status, 0);
    std::string cid;
    { std::lock_guard<std::mutex> g(w->mu); cid = w->cid; w->cid.clear(); }
    {
      std::lock_guard<std::mutex> g(mu_);
      if (warm_live_[w->key] > 0) --warm_live_[w->key];
      auto it = warm_.find(w->key);
      if (it != warm_.end() && it->second == w) warm_.erase(it);
      if (!cid.empty()) pids_.erase(cid);
      workers_.erase(std::remove(workers_.begin(), workers_.end(), w), workers_.end());
    }
    close(w->in_fd);
    if (!cid.empty()) {                                        // died (or was killed) while serving a trial
      ContainerEvent ev{cid, "TERMINATED", 0, ""};
      if (WIFEXITED(status)) { ev.exit_code = WEXITSTATUS(status); if (ev.exit_code) ev.failure = "container failed with non-zero exit code: " + std::to_string(ev.exit_code); }
      else if (WIFSIGNALED(status)) { ev.exit_code = 128 + WTERMSIG(status); ev.failure = std::string("container killed by signal ") + strsignal(WTERMSIG(status)); }
      on_event_(ev);
    }
  }
  void apply_pycache_env() {   // in a forked child, before exec
    if (pycache_dir_.empty()) return;
    unsetenv("PYTHONDONTWRITEBYTECODE");
    setenv("PYTHONPYCACHEPREFIX", pycache_dir_.c_str(), 1);
  }
  // Only a single-slot harness start in a one-container trial goes to a warm worker, and only when nothing in its
  // environment has to be in place before the interpreter / CUDA start (loader paths, CUDA_*, NCCL_*, OMP_*, bind mounts).
  // The task owner's linked agent user (`det user link-with-agent-user`; reference master/pkg/model/user.go AgentUserGroup,
  // applied as the container's user): the forked child gives up its supplementary groups and takes the GID, then the UID, after
  // the working directory and bind mounts are set up and right before exec. An agent that is not root cannot do this and
  // fails the task rather than run it under the wrong identity.
  static bool drop_to_task_user(const std::map<std::string, std::string>& env) {
    auto u = env.find("DET_RUN_AS_UID");
    if (u == env.end()) return true;
    auto g = env.find("DET_RUN_AS_GID");
    const uid_t uid = static_cast<uid_t>(strtoul(u->second.c_str(), nullptr, 10));
    const gid_t gid = g == env.end() ? static_cast<gid_t>(uid) : static_cast<gid_t>(strtoul(g->second.c_str(), nullptr, 10));
    if (geteuid() == uid && getegid() == gid) return true;
    if (setgroups(0, nullptr) != 0 || setgid(gid) != 0 || setuid(uid) != 0) {
      fprintf(stderr, "cannot run the task as uid %u gid %u: %s (the agent must run as root to use linked agent users)\n",
              static_cast<unsigned>(uid), static_cast<unsigned>(gid), strerror(errno));
      return false;
    }
    auto name = env.find("DET_RUN_AS_USER");
    if (name != env.end() && !name->second.empty()) { setenv("USER", name->second.c_str(), 1); setenv("LOGNAME", name->second.c_str(), 1); }
    return true;
  }
  bool warm_eligible(const ContainerSpec& spec, const std::map<std::string, std::string>& env) {
    if (spec.devices.size() != 1) return false;
    if (env.count("DET_RUN_AS_UID")) return false;      // a pre-started worker runs as the agent's own user
    // the harness command, bare or behind the entrypoint shell that sources startup-hook.sh (task_spec.hpp); a model
    // definition that ships such a hook needs the shell, so it starts cold
    const size_t n = spec.argv.size();
    if (n < 3 || spec.argv[n - 2] != "-m" || spec.argv[n - 1] != "determined_b200.exec.harness") return false;
    if (n != 3 && !(n == 7 && spec.argv[0] == "sh" && spec.argv[3] == "det-entrypoint")) return false;
    if (n == 7 && access((spec.cwd + "/startup-hook.sh").c_str(), F_OK) == 0) return false;
    const std::string& python = spec.argv[n - 3];
    {
      std::lock_guard<std::mutex> g(mu_);
      if (warm_mode_ == "off" || warm_mode_.empty()) return false;
      const bool gpu = spec.devices[0].type == "gpu" && spec.devices[0].brand != "Artificial";
      if (warm_mode_ == "gpu" && !gpu) return false;
      if (warm_python_.empty()) {
        warm_python_ = python;
        auto pp = env.find("PYTHONPATH");
        if (pp != env.end()) warm_path_ = pp->second;
      }
      if (python != warm_python_) return false;
    }
    auto nc = env.find("DET_NUM_CONTAINERS");
    if (nc != env.end() && nc->second != "1") return false;
    for (auto& kv : env) {
      const std::string& k = kv.first;
      if (k == "CUDA_VISIBLE_DEVICES" || k == "PYTHONPATH") continue;
      if (k == "DET_BIND_MOUNTS" || k.rfind("LD_", 0) == 0 || k.rfind("CUDA_", 0) == 0 || k.rfind("NCCL_", 0) == 0 || k.rfind("OMP_", 0) == 0 ||
          k.rfind("PYTHON", 0) == 0 || k.rfind("TORCH_", 0) == 0 || k.rfind("PYTORCH_", 0) == 0)
        return false;
    }
    return true;
  }
  EventFn on_event_;
  LogFn on_log_;
  std::map<std::string, std::string> task_env_;
  std::mutex mu_;
  std::map<std::string, pid_t> pids_;
  std::map<std::string, std::shared_ptr<WarmWorker>> warm_;   // idle workers by slot key
  std::map<std::string, int> warm_live_;                       // workers alive (idle or serving) by slot key
  std::vector<std::shared_ptr<WarmWorker>> workers_;
  std::string warm_mode_, warm_python_, warm_path_, pycache_dir_;
  int slot_count_ = 0;
};

}  // namespace det
