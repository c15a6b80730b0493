// task_spec.hpp — the part of a task's container spec that comes from the experiment / command configuration:
// environment variables, bind mounts, shared-memory size, ports, the start-up hook.
//
// Reference: master/pkg/tasks/task.go:172-331 (TrialDockerMounts, TrialEnvVars, startContainer), ports.go:12-38,
// master/static/srv/entrypoint.sh:1-47 (startup-hook.sh is sourced before the harness starts). There the spec becomes a
// Docker container; here a task is a process group on the agent's host, so:
//   * `environment.environment_variables` — a list of "K=V" strings, a map, or {cpu: [...], gpu: [...]} keyed by the slot
//     type of the allocation — is merged into the task's environment;
//   * `bind_mounts` [{host_path, container_path, read_only}] — the host path is reachable as it is; the container path
//     is made to point at it (a symbolic link, relative paths under the task's working directory), so code written for
//     the containerised layout finds its data. An existing different target is an error, not silently shadowed;
//   * `resources.shm_size` — /dev/shm is the host's; the requested size is checked against what is free there;
//   * `environment.ports` {name: port} — exported as DET_PORTS, each port shifted by the trial's unique offset
//     (DET_TRIAL_UNIQUE_PORT_OFFSET) exactly as the rendezvous ports are;
//   * `startup-hook.sh` in the model definition is sourced by the entrypoint shell before the task command runs.
#pragma once
#include <sys/statvfs.h>

#include <map>
#include <string>
#include <vector>

#include "json.hpp"

namespace det {

// "K=V" list, plain map, or per-slot-type map of either
inline void merge_environment_variables(const Json& ev, const std::string& slot_type, std::map<std::string, std::string>& env) {
  auto add_list = [&](const Json& list) {
    for (auto& item : list.arr()) {
      if (!item.is_string()) continue;
      const std::string s = item.as_string();
      const auto eq = s.find('=');
      if (eq == std::string::npos || eq == 0) continue;
      env[s.substr(0, eq)] = s.substr(eq + 1);
    }
  };
  if (ev.is_array()) { add_list(ev); return; }
  if (!ev.is_object()) return;
  const bool by_type = ev.contains("cpu") || ev.contains("gpu");
  if (by_type) {
    const Json& sel = ev[slot_type == "gpu" ? "gpu" : "cpu"];
    if (sel.is_array()) add_list(sel);
    else if (sel.is_object()) for (auto& kv : sel.obj()) env[kv.first] = kv.second.is_string() ? kv.second.as_string() : kv.second.dump();
    return;
  }
  for (auto& kv : ev.obj()) env[kv.first] = kv.second.is_string() ? kv.second.as_string() : kv.second.dump();
}

// One line per mount: host|container|ro — read by the agent right before exec (agent.hpp: apply_bind_mounts).
inline std::string bind_mount_lines(const Json& mounts, std::string* err) {
  std::string out;
  if (!mounts.is_array()) return out;
  for (auto& m : mounts.arr()) {
    const std::string host = m["host_path"].str_or(""), cont = m["container_path"].str_or("");
    if (host.empty() || cont.empty()) { if (err) *err = "bind_mounts entries need host_path and container_path"; return ""; }
    if (host[0] != '/') { if (err) *err = "bind_mounts.host_path must be absolute: " + host; return ""; }
    if (host.find('|') != std::string::npos || cont.find('|') != std::string::npos || host.find('\n') != std::string::npos) {
      if (err) *err = "bind_mounts paths must not contain '|' or newlines";
      return "";
    }
    out += host + "|" + cont + "|" + (m["read_only"].as_bool(false) ? "ro" : "rw") + "\n";
  }
  return out;
}

inline long long parse_size_bytes(const Json& v) {
  if (v.is_null()) return 0;
  if (!v.is_string()) return v.as_int(0);
  const std::string s = v.as_string();
  char* end = nullptr;
  double n = strtod(s.c_str(), &end);
  std::string unit = end ? end : "";
  for (auto& c : unit) c = static_cast<char>(tolower(c));
  while (!unit.empty() && unit.front() == ' ') unit.erase(unit.begin());
  if (unit == "k" || unit == "kb" || unit == "kib") n *= 1024.0;
  else if (unit == "m" || unit == "mb" || unit == "mib") n *= 1024.0 * 1024;
  else if (unit == "g" || unit == "gb" || unit == "gib") n *= 1024.0 * 1024 * 1024;
  return static_cast<long long>(n);
}

// Applies the configuration-derived parts to a task's env / argv. Returns "" or a reason the spec is invalid.
inline std::string apply_task_config(const Json& config, const std::string& slot_type, int port_offset, bool source_startup_hook,
                                     std::map<std::string, std::string>& env, std::vector<std::string>& argv, std::string* warning) {
  merge_environment_variables(config["environment"]["environment_variables"], slot_type, env);
  std::string err;
  const std::string lines = bind_mount_lines(config["bind_mounts"], &err);
  if (!err.empty()) return err;
  if (!lines.empty()) env["DET_BIND_MOUNTS"] = lines;
  const long long shm = parse_size_bytes(config["resources"]["shm_size"]);
  if (shm > 0) {
    env["DET_SHM_SIZE"] = std::to_string(shm);
    struct statvfs st {};
    if (statvfs("/dev/shm", &st) == 0) {
      const long long avail = static_cast<long long>(st.f_bavail) * static_cast<long long>(st.f_frsize);
      if (avail < shm && warning) *warning = "resources.shm_size asks for " + std::to_string(shm) + " bytes of shared memory, /dev/shm has " + std::to_string(avail) + " free";
    }
  }
  const Json& ports = config["environment"]["ports"];
  if (ports.is_object() && !ports.obj().empty()) {
    Json shifted = Json::object();
    for (auto& kv : ports.obj()) shifted[kv.first] = kv.second.as_int(0) + port_offset;
    env["DET_PORTS"] = shifted.dump();
  }
  if (source_startup_hook) {
    // entrypoint shell: the model definition's startup-hook.sh runs in the task's own shell, so what it exports reaches the
    // harness; a failing hook fails the task
    std::vector<std::string> wrapped = {"sh", "-c", "if [ -f startup-hook.sh ]; then . ./startup-hook.sh || exit $?; fi; exec \"$@\"", "det-entrypoint"};
    wrapped.insert(wrapped.end(), argv.begin(), argv.end());
    argv.swap(wrapped);
  }
  return "";
}

}  // namespace det
