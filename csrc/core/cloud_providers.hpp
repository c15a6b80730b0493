// cloud_providers.hpp — instance providers behind the provisioner's scale decider.
//
// Parity map (reference → here): master/internal/provisioner/aws.go (on-demand: RunInstances with user data that starts
// the agent, DescribeInstances by tag, TerminateInstances), aws_spot.go (spot market option + max price), gcp.go /
// gcp_operation_tracker.go (instances.insert / list by label / delete). The reference links the cloud SDKs; here the
// providers drive the vendors' own command-line clients (`aws`, `gcloud`), which carry credentials, retries and
// request signing, so the master needs no TLS stack. Each operation is a pure function building an argv (unit-tested
// through det_core_cloud_command) plus a runner that executes it and parses the JSON it prints.
#pragma once
#include <sys/wait.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "json.hpp"

namespace det {

struct CloudInstance {
  std::string id;
  std::string state;   // starting | running | stopped
};

// argv -> stdout (stderr discarded), exit status in *rc.
inline std::string run_argv_capture(const std::vector<std::string>& argv, int* rc) {
  int fds[2];
  if (pipe(fds) != 0) { if (rc) *rc = -1; return ""; }
  pid_t pid = fork();
  if (pid < 0) { close(fds[0]); close(fds[1]); if (rc) *rc = -1; return ""; }
  if (pid == 0) {
    dup2(fds[1], 1);
    close(fds[0]);
    close(fds[1]);
    std::vector<char*> av;
    for (auto& a : argv) av.push_back(const_cast<char*>(a.c_str()));
    av.push_back(nullptr);
    execvp(av[0], av.data());
    _exit(127);
  }
  close(fds[1]);
  std::string out;
  char buf[4096];
  ssize_t n;
  while ((n = read(fds[0], buf, sizeof buf)) > 0) out.append(buf, static_cast<size_t>(n));
  close(fds[0]);
  int st = 0;
  waitpid(pid, &st, 0);
  if (rc) *rc = WIFEXITED(st) ? WEXITSTATUS(st) : -1;
  return out;
}

struct CloudProviderConfig {
  std::string type;                 // aws | gcp
  std::string cli;                  // executable (default "aws" / "gcloud")
  std::string cluster_id;           // value of the tag / label that marks this master's instances
  std::string master_host;
  int master_port = 8080;
  std::string agent_command = "det-agent";   // what the startup script runs on the instance
  int slots_per_instance = 8;
  Json raw;

  std::string tag_key() const { return raw["tag_key"].str_or(raw["label_key"].str_or("determined-cluster")); }
  std::string tag_value() const { return raw["tag_value"].str_or(raw["label_value"].str_or(cluster_id)); }
};

// The boot script of an agent instance: the instance's own id becomes the agent id, so instances and agents pair up
// without any extra bookkeeping (reference aws.go: user data runs the agent container with DET_AGENT_ID=instance id).
inline std::string agent_startup_script(const CloudProviderConfig& c, const std::string& id_expr) {
  std::string s = "#!/bin/bash\n";
  if (!c.raw["startup_script"].str_or("").empty()) s += c.raw["startup_script"].as_string() + "\n";
  s += "exec " + c.agent_command + " --master-host " + c.master_host + " --master-port " + std::to_string(c.master_port) +
       " --agent-id " + id_expr;
  if (!c.raw["agent_label"].str_or("").empty()) s += " --label " + c.raw["agent_label"].as_string();
  return s + "\n";
}

// ---- AWS ------------------------------------------------------------------------------------------------------------
inline std::vector<std::string> aws_prefix(const CloudProviderConfig& c) {
  std::vector<std::string> a = {c.cli.empty() ? "aws" : c.cli};
  if (!c.raw["region"].str_or("").empty()) { a.push_back("--region"); a.push_back(c.raw["region"].as_string()); }
  a.push_back("--output");
  a.push_back("json");
  a.push_back("ec2");
  return a;
}
inline std::vector<std::string> aws_launch_argv(const CloudProviderConfig& c, int count) {
  auto a = aws_prefix(c);
  a.insert(a.end(), {"run-instances", "--image-id", c.raw["image_id"].str_or(""), "--instance-type", c.raw["instance_type"].str_or("p5.48xlarge"),
                     "--count", std::to_string(count)});
  auto opt = [&](const char* key, const char* flag) {
    if (!c.raw[key].str_or("").empty()) { a.push_back(flag); a.push_back(c.raw[key].as_string()); }
  };
  opt("ssh_key_name", "--key-name");
  opt("security_group_id", "--security-group-ids");
  opt("subnet_id", "--subnet-id");
  if (!c.raw["iam_instance_profile_arn"].str_or("").empty()) { a.push_back("--iam-instance-profile"); a.push_back("Arn=" + c.raw["iam_instance_profile_arn"].as_string()); }
  if (c.raw["root_volume_size"].as_int(0) > 0) {
    a.push_back("--block-device-mappings");
    a.push_back("DeviceName=/dev/sda1,Ebs={VolumeSize=" + std::to_string(c.raw["root_volume_size"].as_int()) + ",DeleteOnTermination=true}");
  }
  if (c.raw["spot"].as_bool(false)) {   // reference aws_spot.go: one-time spot requests, optional price ceiling
    std::string m = "MarketType=spot";
    if (!c.raw["spot_max_price"].str_or("").empty()) m += ",SpotOptions={MaxPrice=" + c.raw["spot_max_price"].as_string() + ",SpotInstanceType=one-time}";
    a.push_back("--instance-market-options");
    a.push_back(m);
  }
  a.push_back("--tag-specifications");
  a.push_back("ResourceType=instance,Tags=[{Key=" + c.tag_key() + ",Value=" + c.tag_value() + "},{Key=Name,Value=" +
              c.raw["instance_name"].str_or("determined-agent") + "}]");
  a.push_back("--user-data");
  a.push_back(agent_startup_script(c, "$(curl -s http://169.254.169.254/latest/meta-data/instance-id)"));
  return a;
}
inline std::vector<std::string> aws_list_argv(const CloudProviderConfig& c) {
  auto a = aws_prefix(c);
  a.insert(a.end(), {"describe-instances", "--filters", "Name=tag:" + c.tag_key() + ",Values=" + c.tag_value(),
                     "Name=instance-state-name,Values=pending,running,stopping,stopped,shutting-down"});
  return a;
}
inline std::vector<std::string> aws_terminate_argv(const CloudProviderConfig& c, const std::vector<std::string>& ids) {
  auto a = aws_prefix(c);
  a.push_back("terminate-instances");
  a.push_back("--instance-ids");
  a.insert(a.end(), ids.begin(), ids.end());
  return a;
}
inline std::vector<CloudInstance> aws_parse_list(const Json& j) {
  std::vector<CloudInstance> out;
  if (!j["Reservations"].is_array()) return out;
  for (auto& r : j["Reservations"].arr())
    if (r["Instances"].is_array())
      for (auto& i : r["Instances"].arr()) {
        const std::string st = i["State"]["Name"].str_or("");
        if (st == "terminated") continue;
        out.push_back({i["InstanceId"].str_or(""), st == "pending" ? "starting" : (st == "running" ? "running" : "stopped")});
      }
  return out;
}

// ---- GCP ------------------------------------------------------------------------------------------------------------
inline std::vector<std::string> gcp_prefix(const CloudProviderConfig& c) {
  std::vector<std::string> a = {c.cli.empty() ? "gcloud" : c.cli, "compute", "instances"};
  return a;
}
inline void gcp_scope(const CloudProviderConfig& c, std::vector<std::string>& a, bool zone) {
  if (!c.raw["project"].str_or("").empty()) { a.push_back("--project"); a.push_back(c.raw["project"].as_string()); }
  if (zone && !c.raw["zone"].str_or("").empty()) { a.push_back("--zone"); a.push_back(c.raw["zone"].as_string()); }
  a.push_back("--format");
  a.push_back("json");
}
inline std::vector<std::string> gcp_launch_argv(const CloudProviderConfig& c, const std::vector<std::string>& names) {
  auto a = gcp_prefix(c);
  a.push_back("create");
  a.insert(a.end(), names.begin(), names.end());
  const Json& it = c.raw["instance_type"];
  a.push_back("--machine-type");
  a.push_back(it["machine_type"].str_or("a3-highgpu-8g"));
  if (it["gpu_num"].as_int(0) > 0) {
    a.push_back("--accelerator");
    a.push_back("type=" + it["gpu_type"].str_or("nvidia-b200") + ",count=" + std::to_string(it["gpu_num"].as_int()));
    a.push_back("--maintenance-policy");
    a.push_back("TERMINATE");
  }
  if (it["preemptible"].as_bool(false)) a.push_back("--preemptible");
  if (!c.raw["boot_disk_source_image"].str_or("").empty()) { a.push_back("--image"); a.push_back(c.raw["boot_disk_source_image"].as_string()); }
  if (c.raw["boot_disk_size"].as_int(0) > 0) { a.push_back("--boot-disk-size"); a.push_back(std::to_string(c.raw["boot_disk_size"].as_int()) + "GB"); }
  if (!c.raw["network_interface"]["network"].str_or("").empty()) { a.push_back("--network"); a.push_back(c.raw["network_interface"]["network"].as_string()); }
  if (!c.raw["network_interface"]["subnetwork"].str_or("").empty()) { a.push_back("--subnet"); a.push_back(c.raw["network_interface"]["subnetwork"].as_string()); }
  if (!c.raw["service_account"]["email"].str_or("").empty()) { a.push_back("--service-account"); a.push_back(c.raw["service_account"]["email"].as_string()); }
  a.push_back("--labels");
  a.push_back(c.tag_key() + "=" + c.tag_value());
  a.push_back("--metadata");
  a.push_back("startup-script=" + agent_startup_script(c, "$(hostname)"));
  gcp_scope(c, a, true);
  return a;
}
inline std::vector<std::string> gcp_list_argv(const CloudProviderConfig& c) {
  auto a = gcp_prefix(c);
  a.push_back("list");
  a.push_back("--filter");
  a.push_back("labels." + c.tag_key() + "=" + c.tag_value());
  gcp_scope(c, a, false);
  return a;
}
inline std::vector<std::string> gcp_terminate_argv(const CloudProviderConfig& c, const std::vector<std::string>& ids) {
  auto a = gcp_prefix(c);
  a.push_back("delete");
  a.insert(a.end(), ids.begin(), ids.end());
  a.push_back("--quiet");
  gcp_scope(c, a, true);
  return a;
}
inline std::vector<CloudInstance> gcp_parse_list(const Json& j) {
  std::vector<CloudInstance> out;
  if (!j.is_array()) return out;
  for (auto& i : j.arr()) {
    const std::string st = i["status"].str_or("");
    out.push_back({i["name"].str_or(""), (st == "PROVISIONING" || st == "STAGING") ? "starting" : (st == "RUNNING" ? "running" : "stopped")});
  }
  return out;
}

// ---- provider facade ---------------------------------------------------------------------------------------------------
class CloudCliProvider {
 public:
  explicit CloudCliProvider(CloudProviderConfig c) : c_(std::move(c)) {}
  const CloudProviderConfig& config() const { return c_; }

  bool list(std::vector<CloudInstance>* out, std::string* err) {
    int rc = 0;
    const std::string txt = run_argv_capture(c_.type == "aws" ? aws_list_argv(c_) : gcp_list_argv(c_), &rc);
    if (rc != 0) { if (err) *err = c_.type + " list failed with status " + std::to_string(rc); return false; }
    try {
      const Json j = Json::parse(txt.empty() ? (c_.type == "aws" ? "{}" : "[]") : txt);
      *out = c_.type == "aws" ? aws_parse_list(j) : gcp_parse_list(j);
    } catch (const std::exception& e) {
      if (err) *err = std::string("cannot parse ") + c_.type + " output: " + e.what();
      return false;
    }
    return true;
  }
  bool launch(int count, std::string* err) {
    if (count <= 0) return true;
    int rc = 0;
    if (c_.type == "aws") {
      run_argv_capture(aws_launch_argv(c_, count), &rc);
    } else {
      std::vector<std::string> names;
      for (int i = 0; i < count; ++i) names.push_back(c_.raw["name_prefix"].str_or("det-agent-") + c_.cluster_id.substr(0, 8) + "-" + std::to_string(++seq_) + "-" + std::to_string(getpid() % 10000));
      run_argv_capture(gcp_launch_argv(c_, names), &rc);
    }
    if (rc != 0 && err) *err = c_.type + " launch failed with status " + std::to_string(rc);
    return rc == 0;
  }
  bool terminate(const std::vector<std::string>& ids, std::string* err) {
    if (ids.empty()) return true;
    int rc = 0;
    run_argv_capture(c_.type == "aws" ? aws_terminate_argv(c_, ids) : gcp_terminate_argv(c_, ids), &rc);
    if (rc != 0 && err) *err = c_.type + " terminate failed with status " + std::to_string(rc);
    return rc == 0;
  }

 private:
  CloudProviderConfig c_;
  long long seq_ = 0;
};

inline CloudProviderConfig cloud_config_from_json(const Json& provider, const std::string& cluster_id, const std::string& master_host,
                                                  int master_port, int slots_per_instance) {
  CloudProviderConfig c;
  c.type = provider["type"].str_or("");
  c.cli = provider["cli"].str_or("");
  c.cluster_id = cluster_id;
  c.master_host = provider["master_url_host"].str_or(master_host);
  c.master_port = static_cast<int>(provider["master_url_port"].as_int(master_port));
  c.agent_command = provider["agent_command"].str_or(c.agent_command);
  c.slots_per_instance = slots_per_instance;
  c.raw = provider;
  return c;
}

}  // namespace det
