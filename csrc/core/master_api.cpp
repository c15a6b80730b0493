// master_api.cpp — REST API of the master (JSON over HTTP/1.1). Long-poll endpoints serve the harness where the
// reference uses WebSockets (/ws/trial/:exp/:trial/:container). Route groups follow the reference's
// master/internal/core.go:313-606 and api_*.go: experiments, trials, checkpoints, tasks, searcher preview, agents/slots,
// users/auth, templates, models, commands/notebooks/shells/tensorboards, master info/logs/config.
#include <sys/stat.h>
#include <algorithm>
#include <fstream>
#include <sstream>

#include "lttb.hpp"
#include "master.hpp"
#include "task_spec.hpp"

namespace det {

// A TCP port nobody is listening on right now (bind to 0, read it back, release it).
static int free_tcp_port() {
  int fd = ::socket(AF_INET, SOCK_STREAM | SOCK_CLOEXEC, 0);
  if (fd < 0) return 0;
  sockaddr_in addr{};
  addr.sin_family = AF_INET;
  addr.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
  addr.sin_port = 0;
  socklen_t len = sizeof addr;
  int port = 0;
  if (::bind(fd, reinterpret_cast<sockaddr*>(&addr), sizeof addr) == 0 && ::getsockname(fd, reinterpret_cast<sockaddr*>(&addr), &len) == 0)
    port = ntohs(addr.sin_port);
  ::close(fd);
  return port;
}


namespace {
Json obj(std::initializer_list<std::pair<const char*, Json>> kv) {
  Json j = Json::object();
  for (auto& p : kv) j[p.first] = p.second;
  return j;
}
Json arr(const std::vector<Json>& v) {
  Json a = Json::array();
  for (auto& x : v) a.push_back(x);
  return a;
}
HttpResponse ok(const Json& j) { return HttpResponse::json(j); }
HttpResponse not_found(const std::string& what) { return HttpResponse::error(404, what + " not found"); }

Json experiment_summary(Store& store, const Json& e) {
  Json out = e.deep_copy();
  const long long id = e["id"].as_int();
  long long n = 0;
  for (auto& t : store.list("trials", [&](const Json& r) { return r["experiment_id"].as_int() == id; })) { (void)t; ++n; }
  out["num_trials"] = n;
  return out;
}
Json trial_detail(Store& store, const Json& t, bool with_metrics) {
  Json out = t.deep_copy();
  const std::string key = std::to_string(t["id"].as_int());
  auto steps = store.series("steps", key);
  auto vals = store.series("validations", key);
  out["total_batches_processed"] = steps.empty() ? 0 : steps.back()["total_batches"].as_int();
  out["num_steps"] = static_cast<long long>(steps.size());
  if (with_metrics) { out["steps"] = arr(steps); out["validations"] = arr(vals); }
  Json ck = Json::array();
  for (auto& c : store.list("checkpoints", [&](const Json& r) { return r["trial_id"].as_int() == t["id"].as_int(); })) ck.push_back(c);
  out["checkpoints"] = ck;
  if (!vals.empty()) out["latest_validation"] = vals.back();
  return out;
}
}  // namespace

void Master::register_routes() {
  Master* m = this;
  http.set_middleware([m](const HttpRequest& req, HttpResponse& resp) {
    // Exempt are exact (method, route) pairs only: login, the harness's workload exchange and log shipping, the agents'
    // register / commands / events calls, the data-layer lease service used from inside tasks, and the public info pages.
    // Everything that changes cluster state from outside a task (trial kill, agent unregister, slot enable / disable,
    // master config, tracing) needs a session when auth is required.
    auto ends_with = [](const std::string& s, const char* suf) { const size_t n = strlen(suf); return s.size() >= n && s.compare(s.size() - n, n, suf) == 0; };
    const std::string& p = req.path;
    const bool get = req.method == "GET", post = req.method == "POST";
    if (post && p == "/api/v1/auth/login") return true;
    if (get && (p == "/info" || p == "/api/v1/master" || p.rfind("/ui", 0) == 0 || p == "/")) return true;
    if (p.rfind("/api/v1/trials/", 0) == 0 && ((get && ends_with(p, "/next_workload")) || (post && ends_with(p, "/workload_completed")))) return true;
    if (post && p == "/api/v1/trial_logs") return true;
    if (p.rfind("/api/v1/agents/", 0) == 0 && p.find("/slots/") == std::string::npos &&
        ((post && (ends_with(p, "/register") || ends_with(p, "/events"))) || (get && ends_with(p, "/commands")))) return true;
    if (p.rfind("/api/v1/rw_coordinator/", 0) == 0) return true;
    if (m->check_auth(req, nullptr)) return true;
    resp = HttpResponse::error(401, "authentication required: POST /api/v1/auth/login");
    return false;
  });

  // ---- master ----
  auto info = [m](const HttpRequest&) {
    return ok(obj({{"version", "0.13.10.dev0+b200.1"}, {"master_id", m->cluster_id}, {"cluster_id", m->cluster_id}, {"cluster_name", m->cfg.cluster_name},
                   {"uptime_s", static_cast<long long>(std::time(nullptr)) - m->started_at.load()}, {"scheduler", m->cfg.scheduler}}));
  };
  http.route("GET", "/info", info);
  // ---- debugging (role of the reference's /debug/pprof and actor tracing, master/internal/core.go:564-568) ----
  http.route("GET", "/debug/actors", [m](const HttpRequest&) { return ok(m->system.describe_actors()); });
  http.route("POST", "/debug/trace", [m](const HttpRequest& r) {   // {"file": "actors.trace"} (written under <data_dir>/traces/) or {"file": ""} to stop
    // the trace is written under <data_dir>/traces/: a caller chooses a file NAME, never a path
    std::string name = r.json()["file"].str_or("");
    if (name.empty()) { m->system.enable_trace(""); return ok(Json::object()); }
    const auto slash = name.find_last_of('/');
    if (slash != std::string::npos) name = name.substr(slash + 1);
    if (name.empty() || name == "." || name == "..") return HttpResponse::error(400, "file must be a plain file name");
    const std::string dir = m->cfg.data_dir + "/traces";
    ::mkdir(dir.c_str(), 0755);
    m->system.enable_trace(dir + "/" + name);
    return ok(obj({{"file", dir + "/" + name}}));
  });
  // ---- data-layer reader/writer leases (reference: /ws/data-layer/{resource}, master/internal/rw_coordinator.go) ----
  http.route("POST", "/api/v1/rw_coordinator/:resource/:verb", [m](const HttpRequest& r) {
    const std::string verb = r.params.at("verb");
    Json b = r.json();
    const int ttl = static_cast<int>(b["ttl_ms"].as_int(60000));
    if (verb == "read_lock" || verb == "write_lock") {
      auto g = m->rw.acquire(r.params.at("resource"), verb == "read_lock", static_cast<int>(b["wait_ms"].as_int(25000)), ttl);
      return ok(obj({{"granted", g.granted}, {"lease", g.lease}}));
    }
    if (verb == "renew") return ok(obj({{"ok", m->rw.renew(b["lease"].as_string(), ttl)}}));
    if (verb == "unlock") return ok(obj({{"ok", m->rw.release(b["lease"].as_string())}}));
    return HttpResponse::error(400, "verb must be read_lock, write_lock, renew or unlock");
  });
  http.route("GET", "/api/v1/rw_coordinator/:resource/:verb", [m](const HttpRequest& r) {
    int readers = 0, writer = 0, waiting = 0;
    m->rw.stats(r.params.at("resource"), readers, writer, waiting);
    return ok(obj({{"readers", static_cast<long long>(readers)}, {"writer", static_cast<long long>(writer)}, {"waiting", static_cast<long long>(waiting)}}));
  });
  http.route("GET", "/api/v1/provisioner", [m](const HttpRequest&) {
    if (!m->provisioner) return ok(obj({{"enabled", false}}));
    Json s = m->system.ask(m->provisioner, "summary");
    s["enabled"] = true;
    return ok(s);
  });
  // ---- reverse proxy to service tasks (reference: master/internal/proxy/proxy.go; HTTP replayed, Upgrade / WebSocket handshakes tunnelled) ----
  for (const char* method : {"GET", "POST", "PUT", "PATCH", "DELETE"}) {
    auto fwd = [m](const HttpRequest& r) {
      Json c = m->store.get("commands", r.params.at("id"));
      if (c.is_null()) return not_found("task");
      if (c["proxy_port"].as_int(0) <= 0) return HttpResponse::error(404, "task exposes no service");
      if (c["state"].str_or("") != "RUNNING") return HttpResponse::error(502, "task is " + c["state"].str_or("not running"));
      if (c["type"].str_or("") == "tensorboard") {    // keeps an idle TensorBoard from being reaped (tensorboard_timeout)
        const long long now_s = static_cast<long long>(std::time(nullptr));
        if (c["last_request_s"].as_int(0) != now_s) m->store.update("commands", r.params.at("id"), obj({{"last_request_s", now_s}}));
      }
      std::string path = "/";
      auto it = r.params.find("*");
      if (it != r.params.end()) path += it->second;
      if (c["spec"]["keep_prefix"].as_bool(false)) path = "/proxy/" + r.params.at("id") + path;
      if (!r.query_string.empty()) path += "?" + r.query_string;
      const std::string host = c["proxy_host"].str_or("127.0.0.1");
      const int port = static_cast<int>(c["proxy_port"].as_int());
      auto up = r.headers.find("upgrade");
      if (up != r.headers.end() && up->second == "det-tcp") {   // raw byte tunnel to the task's port (`det tunnel`)
        HttpResponse t;
        t.hijack = [host, port](int client) { tcp_tunnel(host, port, client); };
        return t;
      }
      if (up != r.headers.end() && !up->second.empty()) {
        HttpResponse t;
        t.hijack = [host, port, r, path](int client) { http_tunnel(host, port, r, path, client); };
        return t;
      }
      return http_forward(host, port, r, path);
    };
    http.route(method, "/proxy/:id", fwd);
    http.route(method, "/proxy/:id/*", fwd);
  }
  // ---- Kubernetes back-end: pods owned by this master (reference: master/internal/kubernetes/pods.go summaries) ----
  http.route("GET", "/api/v1/kubernetes/pods", [m](const HttpRequest&) {
    if (!m->kube) return ok(obj({{"enabled", false}}));
    return ok(obj({{"enabled", true}, {"pods", m->kube->summary()}}));
  });
  // ---- provisioner decision preview (reference: master/internal/provisioner/scale_decider.go) ----
  http.route("POST", "/api/v1/provisioner/decide", [](const HttpRequest& r) {
    Json b = r.json();
    return ok(scale_decide(ProvisionerConfig::from_json(b["config"]), b["snapshot"]));
  });
  // ---- minimal status page (the reference ships a React SPA; this is one static page over the same REST API) ----
  // ---- WebUI ----
  // determined_b200/master/webui/index.html (dashboard, experiments with metric charts, trials with logs, tasks, cluster,
  // master logs — the page set of the reference's React app, webui/react/src/pages) served from the repo root; a build
  // without that file answers with the embedded status page below.
  http.route("GET", "/", [](const HttpRequest&) {
    HttpResponse resp;
    resp.status = 200;
    resp.content_type = "text/html; charset=utf-8";
    resp.body = "<!doctype html><meta http-equiv=\"refresh\" content=\"0; url=/ui\"><a href=\"/ui\">WebUI</a>";
    return resp;
  });
  http.route("GET", "/ui", [m](const HttpRequest&) {
    HttpResponse resp;
    resp.status = 200;
    resp.content_type = "text/html; charset=utf-8";
    {
      std::ifstream f(m->cfg.repo_root + "/determined_b200/master/webui/index.html");
      if (f) {
        std::stringstream ss;
        ss << f.rdbuf();
        resp.body = ss.str();
        return resp;
      }
    }
    resp.body = R"HTML(<!doctype html><html><head><meta charset="utf-8"><title>determined_b200</title>
<style>body{font:14px system-ui,sans-serif;margin:2em;color:#222}table{border-collapse:collapse;margin:1em 0}
td,th{border:1px solid #ccc;padding:4px 10px;text-align:left}th{background:#f3f3f3}h2{margin-top:1.5em}.bar{background:#ddd;width:120px;height:10px}
.bar div{background:#2a7;height:10px}</style></head><body><h1>determined_b200 master</h1><div id="info"></div>
<h2>Experiments</h2><table id="exps"><tr><th>id</th><th>description</th><th>state</th><th>searcher</th><th>progress</th><th>trials</th><th>best</th></tr></table>
<h2>Slots</h2><table id="slots"><tr><th>agent</th><th>slot</th><th>device</th><th>enabled</th><th>container</th></tr></table>
<h2>Tasks</h2><table id="tasks"><tr><th>id</th><th>name</th><th>slots</th><th>state</th></tr></table>
<script>
async function j(u){const r=await fetch(u);return r.json()}
function row(t,cells){const tr=t.insertRow();cells.forEach(c=>{const td=tr.insertCell();if(c instanceof Node)td.appendChild(c);else td.textContent=c==null?'':c})}
function clear(t){while(t.rows.length>1)t.deleteRow(1)}
async function tick(){try{
 const m=await j('/api/v1/master');document.getElementById('info').textContent='cluster '+(m.cluster_name||'')+' — version '+(m.version||'')+' — scheduler '+(m.scheduler||'');
 const es=await j('/api/v1/experiments');const te=document.getElementById('exps');clear(te);
 (es.experiments||es).forEach(e=>{const b=document.createElement('div');b.className='bar';const f=document.createElement('div');f.style.width=Math.round(100*(e.progress||0))+'%';b.appendChild(f);
  row(te,[e.id,(e.config||{}).description,e.state,((e.config||{}).searcher||{}).name,b,e.num_trials,e.best_validation])});
 const ss=await j('/api/v1/slots');const ts=document.getElementById('slots');clear(ts);ss.forEach(s=>row(ts,[s.agent_id,s.id,(s.device||{}).brand,s.enabled,s.container_id]));
 const tk=await j('/api/v1/tasks');const tt=document.getElementById('tasks');clear(tt);tk.forEach(t=>row(tt,[t.id,t.name,t.slots_needed,t.state]));
}catch(e){}}
tick();setInterval(tick,2000);
</script></body></html>)HTML";
    return resp;
  });
  http.route("GET", "/api/v1/master", info);
  http.route("GET", "/api/v1/master/config", [m](const HttpRequest&) { return ok(m->cfg.to_json()); });
  http.route("GET", "/logs", [m](const HttpRequest& r) { return ok(arr(m->recent_logs(static_cast<size_t>(r.qi("limit", 1000))))); });
  // the last `limit` lines of the master's log, or — for followers — the lines after the one with id `after`
  http.route("GET", "/api/v1/master/logs", [m](const HttpRequest& r) {
    return ok(arr(m->recent_logs(static_cast<size_t>(std::max<long long>(0, r.qi("limit", 1000))), r.q("after").empty() ? -1 : r.qi("after", -1))));
  });

  // ---- auth / users ----
  http.route("POST", "/api/v1/auth/login", [m](const HttpRequest& r) {
    Json b = r.json();
    Json u = m->store.get("users", b["username"].as_string());
    if (u.is_null() || !u["active"].as_bool(true)) return HttpResponse::error(403, "invalid credentials");
    const std::string want = u["password_hash"].as_string();
    const std::string got = b["password"].as_string().empty() ? "" : md5_hex(u["salt"].as_string() + b["password"].as_string());
    if (want != got) return HttpResponse::error(403, "invalid credentials");
    const std::string token = random_hex(24);
    m->store.put("sessions", token, obj({{"username", u["username"]}, {"created", now_rfc3339()}}));
    return ok(obj({{"token", token}, {"user", obj({{"username", u["username"]}, {"admin", u["admin"]}})}}));
  });
  http.route("POST", "/api/v1/auth/logout", [m](const HttpRequest& r) {
    auto it = r.headers.find("authorization");
    if (it != r.headers.end() && it->second.rfind("Bearer ", 0) == 0) m->store.del("sessions", it->second.substr(7));
    return ok(Json::object());
  });
  http.route("GET", "/api/v1/auth/user", [m](const HttpRequest& r) { std::string u; m->check_auth(r, &u); return ok(obj({{"username", u}})); });
  http.route("GET", "/api/v1/users", [m](const HttpRequest&) {
    Json a = Json::array();
    for (auto& u : m->store.list("users")) a.push_back(obj({{"username", u["username"]}, {"admin", u["admin"]}, {"active", u["active"]}, {"agent_user_group", u["agent_user_group"]}}));
    return ok(a);
  });
  http.route("GET", "/api/v1/users/:name", [m](const HttpRequest& r) {
    Json u = m->store.get("users", r.params.at("name"));
    if (u.is_null()) return not_found("user");
    return ok(obj({{"username", u["username"]}, {"admin", u["admin"]}, {"active", u["active"]}, {"agent_user_group", u["agent_user_group"]}}));
  });
  http.route("GET", "/api/v1/master/telemetry", [m](const HttpRequest&) {
    const bool on = m->cfg.telemetry.is_object() && m->cfg.telemetry["enabled"].as_bool(false);
    return ok(obj({{"enabled", on}, {"segment_key", ""}, {"sink", on ? m->cfg.telemetry["sink"].str_or(m->cfg.data_dir + "/telemetry.jsonl") : std::string()}}));
  });
  // User management follows the reference (master/internal/user/service.go): creating users and changing anyone's
  // active / admin / agent-user fields or name is for administrators; a password may also be changed by its owner. Enforced
  // whenever the master requires authentication (without it every caller is the default user anyway).
  auto is_admin = [m](const HttpRequest& r, std::string* who) {
    m->check_auth(r, who);
    if (!m->cfg.auth_required) return true;
    return m->store.get("users", *who)["admin"].as_bool(false);
  };
  http.route("POST", "/api/v1/users", [m, is_admin](const HttpRequest& r) {
    std::string who;
    if (!is_admin(r, &who)) return HttpResponse::error(403, "only an administrator can create users");
    Json b = r.json();
    const std::string name = b["username"].as_string();
    if (name.empty()) return HttpResponse::error(400, "username required");
    if (m->store.exists("users", name)) return HttpResponse::error(409, "user exists");
    m->store.put("users", name, obj({{"username", name}, {"admin", b["admin"].as_bool(false)}, {"active", true}, {"password_hash", ""}, {"salt", ""}}));
    return ok(obj({{"username", name}}));
  });
  http.route("PATCH", "/api/v1/users/:name", [m, is_admin](const HttpRequest& r) {
    Json u = m->store.get("users", r.params.at("name"));
    if (u.is_null()) return not_found("user");
    Json b = r.json(), upd = Json::object();
    std::string who;
    if (!is_admin(r, &who)) {
      for (const char* k : {"active", "admin", "agent_user_group", "username"})
        if (b.contains(k)) return HttpResponse::error(403, std::string("only an administrator can change ") + k);
      if (who != r.params.at("name")) return HttpResponse::error(403, "only an administrator can change another user's password");
    }
    if (b.contains("password")) {
      const std::string salt = random_hex(8);
      upd["salt"] = salt;
      upd["password_hash"] = b["password"].as_string().empty() ? "" : md5_hex(salt + b["password"].as_string());
    }
    for (const char* k : {"active", "admin", "agent_user_group"}) if (b.contains(k)) upd[k] = b[k];
    m->store.update("users", r.params.at("name"), upd);
    const std::string renamed = b["username"].str_or("");
    if (!renamed.empty() && renamed != r.params.at("name")) {      // `det user rename`: the row moves to its new key
      if (m->store.exists("users", renamed)) return HttpResponse::error(409, "user exists");
      Json row = m->store.get("users", r.params.at("name"));
      row["username"] = renamed;
      m->store.put("users", renamed, row);
      m->store.del("users", r.params.at("name"));
      return ok(obj({{"username", renamed}}));
    }
    return ok(obj({{"username", r.params.at("name")}}));
  });

  // ---- experiments ----
  http.route("POST", "/api/v1/experiments", [m](const HttpRequest& r) {
    Json b = r.json();
    Json tmpl;
    if (b["template"].is_string()) {
      tmpl = m->store.get("templates", b["template"].as_string())["config"];
      if (tmpl.is_null()) return HttpResponse::error(404, "template " + b["template"].as_string() + " not found");
    }
    Json cfg;
    try {
      cfg = resolve_experiment_config(b["config"], m->cfg.checkpoint_storage.is_object() ? &m->cfg.checkpoint_storage : nullptr,
                                      tmpl.is_object() ? &tmpl : nullptr, static_cast<unsigned>(std::time(nullptr)));
    } catch (const std::exception& e) {
      return HttpResponse::error(400, std::string("invalid experiment configuration: ") + e.what());
    }
    auto errs = validate_experiment_config(cfg);
    if (!errs.empty()) {
      std::string all;
      for (auto& e : errs) all += (all.empty() ? "" : "; ") + e;
      return HttpResponse::error(400, "invalid experiment configuration: " + all);
    }
    if (b["validate_only"].as_bool(false)) return ok(obj({{"config", cfg}}));
    std::string user;
    m->check_auth(r, &user);
    const long long id = m->create_experiment(cfg, b["model_definition"], b["parent_id"].as_int(0), b["activate"].is_null() ? true : b["activate"].as_bool(true), user);
    if (b["git"].is_object()) {   // `det experiment create --git`: where the model definition came from (remote, commit, committer, date)
      Json g = Json::object();
      for (const char* k : {"git_remote", "git_commit", "git_committer", "git_commit_date"}) g[k] = b["git"][k].str_or("");
      m->store.update("experiments", std::to_string(id), g);
    }
    return HttpResponse::json(obj({{"id", id}, {"config", cfg}}), 201);
  });
  http.route("GET", "/api/v1/experiments", [m](const HttpRequest& r) {
    // reference GetExperiments: archived / states / users / labels / description filters, sort_by + order_by, offset + limit
    const bool all = r.q("all") == "true" || r.q("archived") == "true";
    auto set_of = [&](const char* a, const char* b2) {
      std::set<std::string> out;
      for (const char* key : {a, b2}) {
        std::stringstream ss(r.q(key));
        std::string item;
        while (std::getline(ss, item, ',')) if (!item.empty()) out.insert(item);
      }
      return out;
    };
    const auto states = set_of("state", "states"), users = set_of("user", "users"), labels = set_of("label", "labels");
    const std::string description = r.q("description");
    std::vector<Json> rows;
    for (auto& e : m->store.list("experiments", [&](const Json& row) {
           if (!all && row["archived"].as_bool(false)) return false;
           if (!states.empty() && !states.count(row["state"].str_or(""))) return false;
           if (!users.empty() && !users.count(row["owner"].str_or("determined"))) return false;
           if (!description.empty() && row["description"].str_or("").find(description) == std::string::npos) return false;
           if (!labels.empty()) {
             std::set<std::string> have;
             if (row["labels"].is_array()) for (auto& l : row["labels"].arr()) have.insert(l.str_or(""));
             for (auto& want : labels) if (!have.count(want)) return false;
           }
           return true;
         }))
      rows.push_back(experiment_summary(m->store, e));
    std::string sort_by = r.q("sort_by", "id");
    if (sort_by.rfind("SORT_BY_", 0) == 0) sort_by = sort_by.substr(8);
    for (auto& c : sort_by) c = static_cast<char>(tolower(c));
    if (sort_by == "user") sort_by = "owner";
    static const std::set<std::string> sortable = {"id", "description", "start_time", "end_time", "state", "progress", "owner", "num_trials"};
    if (!sortable.count(sort_by)) return HttpResponse::error(400, "sort_by must be one of id, description, start_time, end_time, state, progress, user, num_trials");
    std::string order = r.q("order_by", "asc");
    for (auto& c : order) c = static_cast<char>(tolower(c));
    const bool desc = order == "desc" || order == "order_by_desc";
    std::stable_sort(rows.begin(), rows.end(), [&](const Json& x, const Json& y) {
      const Json &a2 = desc ? y[sort_by] : x[sort_by], &b2 = desc ? x[sort_by] : y[sort_by];
      if (a2.is_number() || b2.is_number()) return a2.as_double(0) < b2.as_double(0);
      return a2.str_or("") < b2.str_or("");
    });
    const size_t offset = static_cast<size_t>(std::max<long long>(0, r.qi("offset", 0)));
    const long long limit = r.qi("limit", 0);
    Json a = Json::array();
    for (size_t i = offset; i < rows.size() && (limit <= 0 || a.size() < static_cast<size_t>(limit)); ++i) a.push_back(rows[i]);
    return ok(a);
  });
  http.route("GET", "/api/v1/experiments/:id", [m](const HttpRequest& r) {
    Json e = m->store.get("experiments", r.params.at("id"));
    if (e.is_null()) return not_found("experiment");
    Json out = experiment_summary(m->store, e);
    Json trials = Json::array();
    const long long id = e["id"].as_int();
    for (auto& t : m->store.list("trials", [&](const Json& row) { return row["experiment_id"].as_int() == id; })) trials.push_back(trial_detail(m->store, t, r.q("metrics") == "true"));
    out["trials"] = trials;
    return ok(out);
  });
  http.route("PATCH", "/api/v1/experiments/:id", [m](const HttpRequest& r) {
    const std::string id = r.params.at("id");
    Json e = m->store.get("experiments", id);
    if (e.is_null()) return not_found("experiment");
    Json b = r.json(), upd = Json::object();
    for (const char* k : {"description", "labels", "archived"}) if (b.contains(k)) upd[k] = b[k];
    if (b.contains("archived") && b["archived"].as_bool(false) && !is_terminal_state(e["state"].as_string()))
      return HttpResponse::error(400, "only experiments in a terminal state can be archived");
    if (upd.size()) m->store.update("experiments", id, upd);
    auto actor = m->experiment_actor(atoll(id.c_str()));
    if (b.contains("state")) {
      if (!actor) return HttpResponse::error(400, "experiment is in terminal state " + e["state"].as_string());
      m->system.ask(actor, "set_state", b["state"]);
    }
    if (b.contains("resources") && actor) m->system.ask(actor, "set_resources", b["resources"]);
    return ok(m->store.get("experiments", id));
  });
  auto state_route = [m, this](const char* verb, const char* state, bool kill) {
    http.route("POST", std::string("/api/v1/experiments/:id/") + verb, [m, state, kill](const HttpRequest& r) {
      const std::string id = r.params.at("id");
      Json e = m->store.get("experiments", id);
      if (e.is_null()) return not_found("experiment");
      auto actor = m->experiment_actor(atoll(id.c_str()));
      if (!actor) return is_terminal_state(e["state"].as_string()) ? ok(e) : HttpResponse::error(409, "experiment actor is gone");
      Json resp = kill ? m->system.ask(actor, "kill") : m->system.ask(actor, "set_state", Json(std::string(state)));
      return ok(obj({{"id", e["id"]}, {"state", resp["state"]}}));
    });
  };
  state_route("activate", "ACTIVE", false);
  state_route("pause", "PAUSED", false);
  state_route("cancel", "STOPPING_CANCELED", false);
  state_route("kill", "STOPPING_CANCELED", true);
  for (const char* verb : {"archive", "unarchive"}) {
    const bool val = std::string(verb) == "archive";
    http.route("POST", std::string("/api/v1/experiments/:id/") + verb, [m, val](const HttpRequest& r) {
      Json e = m->store.get("experiments", r.params.at("id"));
      if (e.is_null()) return not_found("experiment");
      if (val && !is_terminal_state(e["state"].as_string())) return HttpResponse::error(400, "only experiments in a terminal state can be archived");
      return ok(m->store.update("experiments", r.params.at("id"), obj({{"archived", val}})));
    });
  }
  http.route("DELETE", "/api/v1/experiments/:id", [m](const HttpRequest& r) {
    Json e = m->store.get("experiments", r.params.at("id"));
    if (e.is_null()) return not_found("experiment");
    if (!is_terminal_state(e["state"].as_string())) return HttpResponse::error(400, "only experiments in a terminal state can be deleted");
    const long long id = e["id"].as_int();
    for (auto& t : m->store.list("trials", [&](const Json& row) { return row["experiment_id"].as_int() == id; })) m->store.del("trials", std::to_string(t["id"].as_int()));
    m->store.del("experiments", r.params.at("id"));
    return ok(Json::object());
  });
  http.route("GET", "/api/v1/experiment/labels", [m](const HttpRequest&) {
    // every label in use with the number of experiments carrying it, most used first (reference GetExperimentLabels)
    std::map<std::string, long long> n;
    for (auto& e : m->store.list("experiments")) {
      // labels set after creation (PATCH) live on the row; before that the config's own list applies
      const Json& ls = e["labels"].is_array() ? e["labels"] : e["config"]["labels"];
      for (auto& l : ls.arr()) if (l.is_string()) ++n[l.as_string()];
    }
    std::vector<std::pair<std::string, long long>> v(n.begin(), n.end());
    std::stable_sort(v.begin(), v.end(), [](const auto& a, const auto& b) { return a.second > b.second; });
    Json labels = Json::array();
    for (auto& kv : v) labels.push_back(obj({{"label", kv.first}, {"count", kv.second}}));
    return ok(obj({{"labels", labels}}));
  });
  http.route("GET", "/api/v1/experiments/:id/trials", [m](const HttpRequest& r) {
    const long long id = atoll(r.params.at("id").c_str());
    if (m->store.get("experiments", id).is_null()) return not_found("experiment");
    const std::string state = r.q("state");
    Json trials = Json::array();
    for (auto& t : m->store.list("trials", [&](const Json& row) { return row["experiment_id"].as_int() == id && (state.empty() || row["state"].as_string() == state); }))
      trials.push_back(trial_detail(m->store, t, r.q("metrics") == "true"));
    return ok(trials);
  });
  http.route("GET", "/api/v1/experiments/:id/validation-history", [m](const HttpRequest& r) {
    // the experiment's record progression: every validation that improved on all earlier ones, in time order
    const long long id = atoll(r.params.at("id").c_str());
    Json e = m->store.get("experiments", id);
    if (e.is_null()) return not_found("experiment");
    const std::string metric = e["config"]["searcher"]["metric"].as_string();
    const bool smaller = e["config"]["searcher"]["smaller_is_better"].as_bool(true);
    std::vector<Json> all;
    for (auto& t : m->store.list("trials", [&](const Json& row) { return row["experiment_id"].as_int() == id; }))
      for (auto& v : m->store.series("validations", std::to_string(t["id"].as_int())))
        if (v["metrics"]["validation_metrics"][metric].is_number())
          all.push_back(obj({{"trial_id", t["id"]}, {"end_time", v["end_time"]}, {"total_batches", v["total_batches"]},
                             {"searcher_metric", v["metrics"]["validation_metrics"][metric]}}));
    std::stable_sort(all.begin(), all.end(), [](const Json& a, const Json& b) { return a["end_time"].as_string() < b["end_time"].as_string(); });
    Json hist = Json::array();
    bool have = false;
    double best = 0;
    for (auto& v : all) {
      const double x = v["searcher_metric"].as_double();
      if (!have || (smaller ? x < best : x > best)) { best = x; have = true; hist.push_back(v); }
    }
    return ok(obj({{"validation_history", hist}}));
  });
  http.route("GET", "/api/v1/experiments/:id/checkpoints", [m](const HttpRequest& r) {
    const long long id = atoll(r.params.at("id").c_str());
    return ok(arr(m->store.list("checkpoints", [&](const Json& row) { return row["experiment_id"].as_int() == id; })));
  });
  http.route("GET", "/api/v1/experiments/:id/model_def", [m](const HttpRequest& r) {
    // returns the stored files as the same [{path, content}] list that was uploaded
    const std::string dir = m->model_def_dir(atoll(r.params.at("id").c_str()));
    Json files = Json::array();
    std::istringstream ls(run_capture("cd '" + dir + "' && find . -type f | sort"));
    std::string rel;
    while (std::getline(ls, rel)) {
      if (rel.size() < 3) continue;
      std::ifstream f(dir + "/" + rel.substr(2), std::ios::binary);
      std::stringstream ss;
      ss << f.rdbuf();
      files.push_back(obj({{"path", rel.substr(2)}, {"content", ss.str()}}));
    }
    return ok(files);
  });
  http.route("GET", "/api/v1/experiments/:id/metrics", [m](const HttpRequest& r) {
    // metric series per trial, optionally LTTB-downsampled: ?name=loss&type=training|validation&max_points=N
    const long long id = atoll(r.params.at("id").c_str());
    const std::string name = r.q("name", "loss"), type = r.q("type", "training");
    const int max_points = static_cast<int>(r.qi("max_points", 0));
    Json out = Json::object();
    for (auto& t : m->store.list("trials", [&](const Json& row) { return row["experiment_id"].as_int() == id; })) {
      std::vector<std::pair<double, double>> pts;
      const std::string key = std::to_string(t["id"].as_int());
      if (type == "training") {
        for (auto& s : m->store.series("steps", key)) { const Json& v = s["metrics"]["avg_metrics"][name]; if (v.is_number()) pts.emplace_back(s["total_batches"].as_double(), v.as_double()); }
      } else {
        for (auto& s : m->store.series("validations", key)) { const Json& v = s["metrics"]["validation_metrics"][name]; if (v.is_number()) pts.emplace_back(s["total_batches"].as_double(), v.as_double()); }
      }
      if (max_points > 2) pts = lttb_downsample(pts, max_points);
      Json a = Json::array();
      for (auto& p : pts) { Json q = Json::array(); q.push_back(p.first); q.push_back(p.second); a.push_back(q); }
      out[key] = a;
    }
    return ok(out);
  });

  // ---- hyperparameter-search visualisation (reference api.proto MetricNames / MetricBatches / TrialsSnapshot / TrialsSample,
  // served there as polling streams; here each GET answers with the current state, and the gRPC gateway streams it) ----
  http.route("GET", "/api/v1/experiments/:id/metrics-stream/metric-names", [m](const HttpRequest& r) {
    const long long id = atoll(r.params.at("id").c_str());
    Json e = m->store.get("experiments", id);
    if (e.is_null()) return not_found("experiment");
    std::set<std::string> train, val;
    for (auto& t : m->store.list("trials", [&](const Json& row) { return row["experiment_id"].as_int() == id; })) {
      const std::string key = std::to_string(t["id"].as_int());
      for (auto& s : m->store.series("steps", key)) for (auto& kv : s["metrics"]["avg_metrics"].obj()) if (kv.second.is_number()) train.insert(kv.first);
      for (auto& s : m->store.series("validations", key)) for (auto& kv : s["metrics"]["validation_metrics"].obj()) if (kv.second.is_number()) val.insert(kv.first);
    }
    Json tr = Json::array(), va = Json::array();
    for (auto& n : train) tr.push_back(n);
    for (auto& n : val) va.push_back(n);
    return ok(obj({{"searcher_metric", e["config"]["searcher"]["metric"]}, {"training_metrics", tr}, {"validation_metrics", va}}));
  });
  // one trial's series of (total batches, value) of a metric; type = training | validation
  auto metric_series = [m](const Json& t, const std::string& name, const std::string& type) {
    std::vector<std::pair<double, double>> pts;
    const std::string key = std::to_string(t["id"].as_int());
    if (type == "validation") {
      for (auto& s : m->store.series("validations", key)) { const Json& v = s["metrics"]["validation_metrics"][name]; if (v.is_number()) pts.emplace_back(s["total_batches"].as_double(), v.as_double()); }
    } else {
      for (auto& s : m->store.series("steps", key)) { const Json& v = s["metrics"]["avg_metrics"][name]; if (v.is_number()) pts.emplace_back(s["total_batches"].as_double(), v.as_double()); }
    }
    return pts;
  };
  auto metric_type_of = [](const HttpRequest& r) {
    const std::string t = r.q("metric_type", r.q("type", "validation"));
    return (t == "training" || t == "METRIC_TYPE_TRAINING" || t == "1") ? std::string("training") : std::string("validation");
  };
  http.route("GET", "/api/v1/experiments/:id/metrics-stream/batches", [m, metric_series, metric_type_of](const HttpRequest& r) {
    const long long id = atoll(r.params.at("id").c_str());
    if (r.q("metric_name").empty()) return HttpResponse::error(400, "metric_name is required");
    std::set<long long> at;
    for (auto& t : m->store.list("trials", [&](const Json& row) { return row["experiment_id"].as_int() == id; }))
      for (auto& p : metric_series(t, r.q("metric_name"), metric_type_of(r))) at.insert(static_cast<long long>(p.first));
    Json b = Json::array();
    for (auto v : at) b.push_back(v);
    return ok(obj({{"batches", b}}));
  });
  http.route("GET", "/api/v1/experiments/:id/metrics-stream/trials-snapshot", [m, metric_series, metric_type_of](const HttpRequest& r) {
    const long long id = atoll(r.params.at("id").c_str());
    if (r.q("metric_name").empty()) return HttpResponse::error(400, "metric_name is required");
    const long long at = r.qi("batches_processed", -1);
    Json trials = Json::array();
    for (auto& t : m->store.list("trials", [&](const Json& row) { return row["experiment_id"].as_int() == id; }))
      for (auto& p : metric_series(t, r.q("metric_name"), metric_type_of(r)))
        if (static_cast<long long>(p.first) == at) { trials.push_back(obj({{"trial_id", t["id"]}, {"hparams", t["hparams"]}, {"metric", p.second}})); break; }
    return ok(obj({{"trials", trials}}));
  });
  http.route("GET", "/api/v1/experiments/:id/metrics-stream/trials-sample", [m, metric_series, metric_type_of](const HttpRequest& r) {
    // the `max_trials` best trials by their latest value of the metric (order from searcher.smaller_is_better), each series
    // cut to [start_batches, end_batches] and LTTB-downsampled to max_datapoints
    const long long id = atoll(r.params.at("id").c_str());
    if (r.q("metric_name").empty()) return HttpResponse::error(400, "metric_name is required");
    Json e = m->store.get("experiments", id);
    if (e.is_null()) return not_found("experiment");
    const bool smaller = e["config"]["searcher"]["smaller_is_better"].as_bool(true);
    const int max_trials = static_cast<int>(r.qi("max_trials", 25)), max_points = static_cast<int>(r.qi("max_datapoints", 1000));
    const long long lo = r.qi("start_batches", 0), hi = r.qi("end_batches", 0) > 0 ? r.qi("end_batches", 0) : (1LL << 60);
    struct Row { Json trial; std::vector<std::pair<double, double>> pts; double last; };
    std::vector<Row> rows;
    for (auto& t : m->store.list("trials", [&](const Json& row) { return row["experiment_id"].as_int() == id; })) {
      auto pts = metric_series(t, r.q("metric_name"), metric_type_of(r));
      if (pts.empty()) continue;
      rows.push_back(Row{t, pts, pts.back().second});
    }
    std::stable_sort(rows.begin(), rows.end(), [&](const Row& a, const Row& b) { return smaller ? a.last < b.last : a.last > b.last; });
    if (max_trials > 0 && static_cast<int>(rows.size()) > max_trials) rows.resize(static_cast<size_t>(max_trials));
    Json trials = Json::array(), promoted = Json::array();
    for (auto& row : rows) {
      std::vector<std::pair<double, double>> win;
      for (auto& p : row.pts) if (p.first >= static_cast<double>(lo) && p.first <= static_cast<double>(hi)) win.push_back(p);
      if (max_points > 2) win = lttb_downsample(win, max_points);
      Json data = Json::array();
      for (auto& p : win) data.push_back(obj({{"batches", static_cast<long long>(p.first)}, {"value", p.second}}));
      trials.push_back(obj({{"trial_id", row.trial["id"]}, {"hparams", row.trial["hparams"]}, {"data", data}}));
      promoted.push_back(row.trial["id"]);
    }
    return ok(obj({{"trials", trials}, {"promoted_trials", promoted}, {"demoted_trials", Json::array()}}));
  });

  // ---- trials ----
  http.route("GET", "/api/v1/trials/:id", [m](const HttpRequest& r) {
    Json t = m->store.get("trials", r.params.at("id"));
    if (t.is_null()) return not_found("trial");
    return ok(trial_detail(m->store, t, r.q("metrics") != "false"));
  });
  http.route("GET", "/api/v1/trials/:id/logs", [m](const HttpRequest& r) {
    // Filters as in the reference's TrialLogs API (master/internal/api_trials.go): agent / container / rank / stdtype / source take
    // comma-separated sets, level is a minimum, timestamps bound the window, search is a substring; the filters apply to the
    // whole log from `offset`, then head / tail / limit pick lines. Every line carries its position ("id"): a follower asks
    // for offset = last id + 1.
    m->system.ask(m->logger, "flush");
    const size_t offset = static_cast<size_t>(std::max<long long>(0, r.qi("offset", 0)));
    const long long head = r.qi("head", -1), tail = r.qi("tail", -1);
    size_t limit = static_cast<size_t>(std::max<long long>(0, r.qi("limit", 5000)));
    if (head >= 0) limit = std::min(limit, static_cast<size_t>(head));
    auto set_of = [&](const char* a, const char* b) {
      std::set<std::string> out;
      for (const char* key : {a, b}) {
        std::stringstream ss(r.q(key));
        std::string item;
        while (std::getline(ss, item, ',')) if (!item.empty()) out.insert(item);
      }
      return out;
    };
    const auto ranks = set_of("rank", "rank_ids"), stdtypes = set_of("stdtype", "stdtypes"), containers = set_of("container_id", "container_ids"),
               agents = set_of("agent_id", "agent_ids"), sources = set_of("source", "sources");
    static const std::map<std::string, int> order = {{"TRACE", 0}, {"DEBUG", 1}, {"INFO", 2}, {"WARNING", 3}, {"ERROR", 4}, {"CRITICAL", 5}};
    int min_level = -1;
    if (!r.q("level").empty()) {
      std::string lv = r.q("level");
      for (auto& c : lv) c = static_cast<char>(toupper(c));
      if (lv.rfind("LOG_LEVEL_", 0) == 0) lv = lv.substr(10);
      auto it = order.find(lv);
      if (it == order.end()) return HttpResponse::error(400, "level must be one of TRACE, DEBUG, INFO, WARNING, ERROR, CRITICAL");
      min_level = it->second;
    }
    const std::string before = r.q("timestamp_before"), after = r.q("timestamp_after"), search = r.q("search");
    std::vector<Json> out;
    size_t idx = offset;
    for (auto& l : m->store.read_logs(atoll(r.params.at("id").c_str()), offset, 1u << 30)) {
      Json row = l;
      if (!row["id"].is_number()) row["id"] = static_cast<long long>(idx);
      ++idx;
      if (!ranks.empty() && (!l["rank"].is_number() || !ranks.count(std::to_string(l["rank"].as_int())))) continue;
      if (!stdtypes.empty() && !stdtypes.count(l["stdtype"].str_or(""))) continue;
      if (!containers.empty() && !containers.count(l["container_id"].str_or(""))) continue;
      if (!agents.empty() && !agents.count(l["agent_id"].str_or(""))) continue;
      if (!sources.empty() && !sources.count(l["source"].str_or(""))) continue;
      if (min_level >= 0) {
        auto it = order.find(l["level"].str_or("INFO"));
        if ((it == order.end() ? 2 : it->second) < min_level) continue;
      }
      // RFC 3339 UTC timestamps of one format compare as strings
      if (!before.empty() && !(l["timestamp"].str_or("") < before)) continue;
      if (!after.empty() && !(l["timestamp"].str_or("") > after)) continue;
      if (!search.empty() && l["message"].str_or("").find(search) == std::string::npos) continue;
      out.push_back(row);
      if (tail < 0 && out.size() >= limit) break;
    }
    if (tail >= 0) {
      const size_t keep = std::min(static_cast<size_t>(tail), limit);
      if (out.size() > keep) out.erase(out.begin(), out.end() - static_cast<long>(keep));
    }
    return ok(arr(out));
  });
  http.route("GET", "/api/v1/trials/:id/logs/fields", [m](const HttpRequest& r) {
    // the distinct values the log filters can take for this trial (reference TrialLogsFields)
    m->system.ask(m->logger, "flush");
    std::set<std::string> agents, containers, stdtypes, sources;
    std::set<long long> ranks;
    for (auto& l : m->store.read_logs(atoll(r.params.at("id").c_str()), 0, 1u << 30)) {
      if (l["agent_id"].is_string()) agents.insert(l["agent_id"].as_string());
      if (l["container_id"].is_string()) containers.insert(l["container_id"].as_string());
      if (l["stdtype"].is_string()) stdtypes.insert(l["stdtype"].as_string());
      if (l["source"].is_string()) sources.insert(l["source"].as_string());
      if (l["rank"].is_number()) ranks.insert(l["rank"].as_int());
    }
    auto strs = [](const std::set<std::string>& v) { Json a = Json::array(); for (auto& x : v) a.push_back(x); return a; };
    Json rk = Json::array();
    for (auto x : ranks) rk.push_back(x);
    return ok(obj({{"agent_ids", strs(agents)}, {"container_ids", strs(containers)}, {"rank_ids", rk}, {"stdtypes", strs(stdtypes)}, {"sources", strs(sources)}}));
  });
  http.route("GET", "/api/v1/trials/:id/checkpoints", [m](const HttpRequest& r) {
    const long long id = atoll(r.params.at("id").c_str());
    if (m->store.get("trials", id).is_null()) return not_found("trial");
    const std::string state = r.q("state");
    return ok(arr(m->store.list("checkpoints", [&](const Json& row) { return row["trial_id"].as_int() == id && (state.empty() || row["state"].as_string() == state); })));
  });
  http.route("POST", "/api/v1/trials/:id/kill", [m](const HttpRequest& r) {
    Json t = m->store.get("trials", r.params.at("id"));
    if (t.is_null()) return not_found("trial");
    auto actor = m->system.get("experiments/" + std::to_string(t["experiment_id"].as_int()) + "/" + t["request_id"].as_string());
    if (actor) m->system.ask(actor, "kill");
    return ok(Json::object());
  });
  // harness side: long-poll for the next workload, report completion
  http.route("GET", "/api/v1/trials/:id/next_workload", [m](const HttpRequest& r) {
    Json t = m->store.get("trials", r.params.at("id"));
    if (t.is_null()) return not_found("trial");
    auto actor = m->system.get("experiments/" + std::to_string(t["experiment_id"].as_int()) + "/" + t["request_id"].as_string());
    if (!actor) return ok(obj({{"type", "TERMINATE"}}));
    Json resp = m->system.ask(actor, "next_workload", obj({{"container_id", r.q("container_id", "")}}), 30000);
    return ok(resp.is_null() ? obj({{"type", "WAIT"}}) : resp);
  });
  http.route("POST", "/api/v1/trials/:id/workload_completed", [m](const HttpRequest& r) {
    Json t = m->store.get("trials", r.params.at("id"));
    if (t.is_null()) return not_found("trial");
    auto actor = m->system.get("experiments/" + std::to_string(t["experiment_id"].as_int()) + "/" + t["request_id"].as_string());
    if (!actor) return HttpResponse::error(409, "trial is not running");
    m->system.ask(actor, "workload_completed", r.json(), 30000);
    return ok(Json::object());
  });
  http.route("POST", "/api/v1/trial_logs", [m](const HttpRequest& r) {
    Json b = r.json();
    for (auto& l : (b.is_array() ? b : b["logs"]).arr()) m->system.tell(m->logger, "log", l);
    return ok(Json::object());
  });

  // ---- checkpoints ----
  http.route("GET", "/api/v1/checkpoints/:uuid", [m](const HttpRequest& r) {
    Json c = m->store.get("checkpoints", r.params.at("uuid"));
    if (c.is_null()) return not_found("checkpoint");
    Json t = m->store.get("trials", c["trial_id"].as_int());
    Json e = m->store.get("experiments", c["experiment_id"].as_int());
    c["hparams"] = t["hparams"];
    c["experiment_config"] = e["config"];
    for (auto& v : m->store.series("validations", std::to_string(c["trial_id"].as_int())))
      if (v["total_batches"].as_int() == c["total_batches"].as_int()) c["validation"] = v;
    return ok(c);
  });
  http.route("POST", "/api/v1/checkpoints/:uuid/metadata", [m](const HttpRequest& r) {
    Json c = m->store.get("checkpoints", r.params.at("uuid"));
    if (c.is_null()) return not_found("checkpoint");
    Json md = c["metadata"].is_object() ? c["metadata"] : Json::object();
    Json b = r.json();
    for (auto& kv : b["metadata"].obj()) md[kv.first] = kv.second;
    for (auto& k : b["remove"].arr()) md.obj().erase(k.as_string());
    return ok(m->store.update("checkpoints", r.params.at("uuid"), obj({{"metadata", md}})));
  });

  // ---- cluster ----
  http.route("GET", "/api/v1/agents", [m](const HttpRequest&) { return ok(m->system.ask(m->rm, "summary")["agents"]); });
  http.route("GET", "/api/v1/slots", [m](const HttpRequest&) {
    Json out = Json::array();
    Json agents = m->system.ask(m->rm, "summary")["agents"];
    for (auto& a : agents.arr()) for (auto& s : a["slots"].arr()) { Json x = s.deep_copy(); x["agent_id"] = a["id"]; out.push_back(x); }
    return ok(out);
  });
  http.route("GET", "/api/v1/agents/:agent", [m](const HttpRequest& r) {
    const Json agents = m->system.ask(m->rm, "summary")["agents"];      // a named copy: a range-for over the temporary dangles
    for (auto& a : agents.arr()) if (a["id"].as_string() == r.params.at("agent")) return ok(a);
    return not_found("agent");
  });
  http.route("GET", "/api/v1/agents/:agent/slots/:slot", [m](const HttpRequest& r) {
    const Json agents = m->system.ask(m->rm, "summary")["agents"];
    for (auto& a : agents.arr()) {
      if (a["id"].as_string() != r.params.at("agent")) continue;
      for (auto& s : a["slots"].arr()) if (s["id"].dump() == r.params.at("slot") || s["id"].str_or("") == r.params.at("slot")) { Json x = s.deep_copy(); x["agent_id"] = a["id"]; return ok(x); }
    }
    return not_found("slot");
  });
  http.route("GET", "/api/v1/tasks", [m](const HttpRequest&) { return ok(m->system.ask(m->rm, "summary")["tasks"]); });
  http.route("POST", "/api/v1/agents/:agent/slots/:slot/:verb", [m](const HttpRequest& r) {
    const std::string verb = r.params.at("verb");
    if (verb != "enable" && verb != "disable") return HttpResponse::error(400, "verb must be enable or disable");
    Json ok_ = m->system.ask(m->rm, verb + "_slot", obj({{"agent", r.params.at("agent")}, {"slot", atoll(r.params.at("slot").c_str())}}));
    return ok_.as_bool(false) ? ok(Json::object()) : not_found("agent");
  });
  // det-agent's side of the container protocol: long-poll for start / signal commands, report state changes and logs
  http.route("GET", "/api/v1/agents/:agent/:verb", [m](const HttpRequest& r) {
    if (r.params.at("verb") != "commands") return HttpResponse::error(404, "unknown agent resource");
    return ok(obj({{"commands", m->take_agent_commands(r.params.at("agent"), static_cast<int>(r.qi("wait_ms", 20000)))}}));
  });
  http.route("POST", "/api/v1/agents/:agent/:verb", [m](const HttpRequest& r) {
    const std::string verb = r.params.at("verb");
    if (verb == "events") {
      Json b = r.json();
      for (auto& l : b["logs"].arr()) {
        LogLine ll;
        ll.container_id = l["container_id"].as_string(); ll.stdtype = l["stdtype"].str_or("stdout"); ll.message = l["message"].as_string();
        ll.rank = l["rank"].is_null() ? -1 : static_cast<int>(l["rank"].as_int(-1));
        m->route_log(ll);
      }
      for (auto& e : b["events"].arr())
        m->route_container_event(ContainerEvent{e["container_id"].as_string(), e["state"].as_string(), static_cast<int>(e["exit_code"].as_int(0)), e["failure"].str_or("")});
      return ok(Json::object());
    }
    if (verb == "register") {  // det-agent on another node announces its devices
      Json b = r.json();
      b["id"] = r.params.at("agent");
      b["registered_time"] = now_rfc3339();
      // the address other nodes reach this agent at is where its connection came from (host names need not resolve);
      // the name it reported stays as a second entry
      if (!r.remote_addr.empty()) {
        Json addrs = Json::array();
        addrs.push_back(Json(r.remote_addr));
        if (b["addresses"].is_array()) for (auto& a : b["addresses"].arr()) addrs.push_back(a);
        b["addresses"] = addrs;
      }
      m->system.ask(m->rm, "add_agent", b);
      return ok(obj({{"master_set_agent_options", obj({{"master_info", obj({{"cluster_id", m->cluster_id}})}})}}));
    }
    if (verb == "unregister") { m->system.ask(m->rm, "remove_agent", obj({{"id", r.params.at("agent")}})); return ok(Json::object()); }
    if (verb != "enable" && verb != "disable") return HttpResponse::error(400, "unknown agent verb");
    m->system.ask(m->rm, verb + "_slot", obj({{"agent", r.params.at("agent")}, {"slot", -1}}));
    return ok(Json::object());
  });
  http.route("POST", "/api/v1/searcher/preview", [](const HttpRequest& r) {
    Json b = r.json();
    SearcherConfig cfg = SearcherConfig::from_json(b["searcher"]);
    std::vector<std::string> errs;
    cfg.validate(errs);
    if (!errs.empty()) return HttpResponse::error(400, errs[0]);
    Searcher s(static_cast<uint32_t>(b["seed"].as_int(0)), new_search_method(cfg), hyperparameters_from_json(b["hyperparameters"]));
    ValidationFunction vf = [](std::mt19937_64& rng, int, int) { return std::uniform_real_distribution<double>(0, 1)(rng); };
    SimulationResult sim = simulate(s, b["seed"].as_int(0), vf, true, cfg.metric);
    return ok(obj({{"results", sim.summary()}, {"trials", static_cast<long long>(sim.results.size())}}));
  });

  // ---- templates / model registry ----
  http.route("GET", "/api/v1/templates", [m](const HttpRequest&) { return ok(arr(m->store.list("templates"))); });
  http.route("GET", "/api/v1/templates/:name", [m](const HttpRequest& r) { Json t = m->store.get("templates", r.params.at("name")); return t.is_null() ? not_found("template") : ok(t); });
  http.route("PUT", "/api/v1/templates/:name", [m](const HttpRequest& r) {
    m->store.put("templates", r.params.at("name"), obj({{"name", r.params.at("name")}, {"config", r.json()["config"]}}));
    return ok(obj({{"name", r.params.at("name")}}));
  });
  http.route("DELETE", "/api/v1/templates/:name", [m](const HttpRequest& r) { m->store.del("templates", r.params.at("name")); return ok(Json::object()); });
  http.route("GET", "/api/v1/models", [m](const HttpRequest&) { return ok(arr(m->store.list("models"))); });
  http.route("POST", "/api/v1/models", [m](const HttpRequest& r) {
    Json b = r.json();
    const std::string name = b["name"].as_string();
    if (name.empty()) return HttpResponse::error(400, "model name required");
    if (m->store.exists("models", name)) return HttpResponse::error(409, "model exists");
    m->store.put("models", name, obj({{"name", name}, {"description", b["description"]}, {"metadata", b["metadata"].is_object() ? b["metadata"] : Json::object()},
                                      {"creation_time", now_rfc3339()}, {"last_updated_time", now_rfc3339()}, {"num_versions", 0}}));
    return HttpResponse::json(m->store.get("models", name), 201);
  });
  http.route("GET", "/api/v1/models/:name", [m](const HttpRequest& r) { Json t = m->store.get("models", r.params.at("name")); return t.is_null() ? not_found("model") : ok(t); });
  http.route("PATCH", "/api/v1/models/:name", [m](const HttpRequest& r) {
    if (!m->store.exists("models", r.params.at("name"))) return not_found("model");
    Json b = r.json(), upd = obj({{"last_updated_time", now_rfc3339()}});
    for (const char* k : {"description", "metadata"}) if (b.contains(k)) upd[k] = b[k];
    return ok(m->store.update("models", r.params.at("name"), upd));
  });
  http.route("GET", "/api/v1/models/:name/versions", [m](const HttpRequest& r) { return ok(arr(m->store.series("model_versions", r.params.at("name")))); });
  http.route("GET", "/api/v1/models/:name/versions/:version", [m](const HttpRequest& r) {
    for (auto& v : m->store.series("model_versions", r.params.at("name"))) {
      if (v["version"].as_int(-1) != atoll(r.params.at("version").c_str())) continue;
      Json out = v.deep_copy();
      out["model"] = m->store.get("models", r.params.at("name"));
      out["checkpoint_uuid"] = v["checkpoint"]["uuid"];
      return ok(out);
    }
    return not_found("model version");
  });
  http.route("POST", "/api/v1/models/:name/versions", [m](const HttpRequest& r) {
    Json model = m->store.get("models", r.params.at("name"));
    if (model.is_null()) return not_found("model");
    Json b = r.json();
    Json c = m->store.get("checkpoints", b["checkpoint_uuid"].as_string());
    if (c.is_null()) return not_found("checkpoint");
    const long long v = model["num_versions"].as_int() + 1;
    Json ver = obj({{"version", v}, {"checkpoint", c}, {"creation_time", now_rfc3339()}, {"model", r.params.at("name")}});
    m->store.append("model_versions", r.params.at("name"), ver);
    m->store.update("models", r.params.at("name"), obj({{"num_versions", v}, {"last_updated_time", now_rfc3339()}}));
    return HttpResponse::json(ver, 201);
  });

  // ---- commands / notebooks / shells / tensorboards: all are command tasks with a type tag ----
  for (const char* kind : {"commands", "notebooks", "shells", "tensorboards"}) {
    const std::string type = std::string(kind).substr(0, strlen(kind) - 1);
    http.route("POST", std::string("/api/v1/") + kind, [m, type](const HttpRequest& r) {
      Json b = r.json();
      // task config: template <- submitted config (reference: command configs take a template the way experiments do);
      // resources.slots / description / entrypoint in it stand in for the top-level fields when those are absent
      Json config = Json::object();
      if (b["template"].is_string()) {
        Json tmpl = m->store.get("templates", b["template"].as_string())["config"];
        if (tmpl.is_null()) return HttpResponse::error(404, "template " + b["template"].as_string() + " not found");
        if (tmpl.is_object()) config = tmpl.deep_copy();
      }
      if (b["config"].is_object()) json_merge(config, b["config"]);
      else if (!b["config"].is_null()) return HttpResponse::error(400, "config must be a mapping");
      static const std::set<std::string> known = {"description", "entrypoint", "resources", "environment", "bind_mounts", "debug", "work_dir"};
      const Json& cc = config;
      for (auto& kv : cc.obj())
        if (!known.count(kv.first)) return HttpResponse::error(400, "unknown field in task config: '" + kv.first + "'");
      Json argv = b["argv"];
      if ((!argv.is_array() || argv.size() == 0) && cc["entrypoint"].is_array()) argv = cc["entrypoint"];
      if (!argv.is_array() || argv.size() == 0) return HttpResponse::error(400, "argv (list of strings) is required");
      for (auto& a : argv.arr())
        if (!a.is_string()) return HttpResponse::error(400, "argv (list of strings) is required");
      {
        std::string bad;
        bind_mount_lines(cc["bind_mounts"], &bad);
        if (!bad.empty()) return HttpResponse::error(400, bad);
      }
      const long long slots = b["slots"].is_number() ? b["slots"].as_int(0) : cc["resources"]["slots"].as_int(0);
      if (slots < 0) return HttpResponse::error(400, "resources.slots must be >= 0");
      const long long id = m->store.next_id("commands");
      std::string cwd = b["cwd"].str_or(cc["work_dir"].str_or(""));
      const bool has_context = b["context"].is_array() && b["context"].size() > 0;
      if (has_context || cwd.empty()) {   // every task gets a working directory of its own (relative bind mounts land in it); uploaded
        const std::string dir = m->command_context_dir(id);   // files are laid out there, and a startup-hook.sh among them runs first
        m->write_context_files(dir, b["context"]);
        if (cwd.empty()) cwd = dir;
      }
      Json spec = obj({{"type", type}, {"description", b["description"].str_or(cc["description"].str_or(type))}, {"slots", slots}, {"argv", argv},
                       {"env", b["env"].is_object() ? b["env"] : Json::object()}, {"cwd", cwd}, {"config", cc.size() ? config : Json()}, {"has_context", has_context}});
      const bool service = b.contains("proxy") ? b["proxy"].as_bool(false) : (type == "notebook" || type == "tensorboard");
      if (service) {
        spec["proxy_port"] = b["port"].as_int(0) > 0 ? b["port"].as_int() : static_cast<long long>(free_tcp_port());
        spec["ready_path"] = b["ready_path"].str_or("/");
        // Jupyter and TensorBoard are told their URL prefix (base_url / --path_prefix) and expect to see it on every request,
        // as behind the reference's proxy; other services get the path below /proxy/<id>/
        spec["keep_prefix"] = b.contains("keep_prefix") ? b["keep_prefix"].as_bool(false) : (type == "notebook" || type == "tensorboard");
      }
      if (type == "tensorboard" && (b["experiment_ids"].is_array() || b["trial_ids"].is_array())) {
        // where the event files of the requested experiments / trials live: their storage section and key prefix
        // (reference master/internal/command/tensorboard_manager.go builds the same list from the experiment configs)
        Json sources = Json::array();
        auto add = [&](long long exp_id, long long trial_id) {
          const Json e = m->store.get("experiments", exp_id);
          if (e.is_null()) return false;
          const Json& ec = e["config"];
          std::string prefix = m->cluster_id + "/tensorboard/experiment/" + std::to_string(exp_id);
          if (trial_id > 0) prefix += "/trial/" + std::to_string(trial_id);
          sources.push_back(obj({{"experiment_id", exp_id}, {"trial_id", trial_id > 0 ? Json(trial_id) : Json()}, {"prefix", prefix},
                                 {"storage", ec["tensorboard_storage"].is_object() ? ec["tensorboard_storage"] : ec["checkpoint_storage"]}}));
          return true;
        };
        if (b["experiment_ids"].is_array())
          for (auto& x : b["experiment_ids"].arr())
            if (!add(x.as_int(0), 0)) return HttpResponse::error(404, "experiment " + std::to_string(x.as_int(0)) + " not found");
        if (b["trial_ids"].is_array())
          for (auto& x : b["trial_ids"].arr()) {
            const Json t = m->store.get("trials", x.as_int(0));
            if (t.is_null() || !add(t["experiment_id"].as_int(0), x.as_int(0))) return HttpResponse::error(404, "trial " + std::to_string(x.as_int(0)) + " not found");
          }
        if (sources.size() == 0) return HttpResponse::error(400, "a TensorBoard needs at least one experiment or trial");
        spec["env"]["DET_TENSORBOARD_SOURCES"] = sources.dump();
      }
      std::string owner;
      m->check_auth(r, &owner);
      m->store.put("commands", id, obj({{"id", id}, {"type", type}, {"state", "PENDING"}, {"spec", spec}, {"start_time", now_rfc3339()}, {"owner", owner}}));
      m->system.actor_of("commands/" + std::to_string(id), make_command(m, id, spec), m->commands_root);
      return HttpResponse::json(m->store.get("commands", id), 201);
    });
    http.route("GET", std::string("/api/v1/") + kind, [m, type](const HttpRequest& r) {
      std::set<std::string> users;      // ?users=a,b: only these owners' tasks
      std::stringstream ss(r.q("users"));
      std::string item;
      while (std::getline(ss, item, ',')) if (!item.empty()) users.insert(item);
      return ok(arr(m->store.list("commands", [&](const Json& row) {
        return row["type"].as_string() == type && (users.empty() || users.count(row["owner"].str_or("determined")));
      })));
    });
    http.route("GET", std::string("/api/v1/") + kind + "/:id", [m](const HttpRequest& r) {
      Json c = m->store.get("commands", r.params.at("id"));
      if (c.is_null()) return not_found("task");
      c["logs"] = arr(m->store.series("command_logs", r.params.at("id")));
      return ok(c);
    });
    http.route("GET", std::string("/api/v1/") + kind + "/:id/logs", [m](const HttpRequest& r) {
      if (m->store.get("commands", r.params.at("id")).is_null()) return not_found("task");
      auto lines = m->store.series("command_logs", r.params.at("id"));
      const size_t off = static_cast<size_t>(std::max<long long>(0, r.qi("offset", 0)));
      const size_t lim = static_cast<size_t>(std::max<long long>(0, r.qi("limit", 5000)));
      Json out = Json::array();
      for (size_t i = off; i < lines.size() && out.size() < lim; ++i) out.push_back(lines[i]);
      return ok(out);
    });
    http.route("POST", std::string("/api/v1/") + kind + "/:id/kill", [m](const HttpRequest& r) {
      auto actor = m->system.get("commands/" + r.params.at("id"));
      if (actor) m->system.ask(actor, "kill");
      return ok(Json::object());
    });
  }
}

}  // namespace det
