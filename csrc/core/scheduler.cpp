// scheduler.cpp — see scheduler.hpp.
#include "scheduler.hpp"

#include <algorithm>
#include <cstring>
#include <set>
#include <stdexcept>

namespace det {

// ---- md5 (RFC 1321) for the fitting tie-break -------------------------------------------------------------------
namespace {
inline uint32_t rotl(uint32_t x, int c) { return (x << c) | (x >> (32 - c)); }
void md5(const std::string& msg, uint8_t out[16]) {
  static const uint32_t K[64] = {
      0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501, 0x698098d8, 0x8b44f7af,
      0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821, 0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa,
      0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8, 0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8,
      0x676f02d9, 0x8d2a4c8a, 0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
      0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665, 0xf4292244, 0x432aff97,
      0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1, 0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1,
      0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};
  static const int S[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 5, 9, 14, 20, 5, 9, 14, 20, 5, 9, 14, 20, 5, 9, 14, 20,
                            4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};
  uint32_t a0 = 0x67452301, b0 = 0xefcdab89, c0 = 0x98badcfe, d0 = 0x10325476;
  std::string m = msg;
  const uint64_t bitlen = static_cast<uint64_t>(msg.size()) * 8;
  m.push_back(static_cast<char>(0x80));
  while (m.size() % 64 != 56) m.push_back(0);
  for (int i = 0; i < 8; ++i) m.push_back(static_cast<char>((bitlen >> (8 * i)) & 0xff));
  for (size_t off = 0; off < m.size(); off += 64) {
    uint32_t M[16];
    for (int i = 0; i < 16; ++i) {
      const uint8_t* p = reinterpret_cast<const uint8_t*>(m.data()) + off + 4 * i;
      M[i] = static_cast<uint32_t>(p[0]) | (static_cast<uint32_t>(p[1]) << 8) | (static_cast<uint32_t>(p[2]) << 16) | (static_cast<uint32_t>(p[3]) << 24);
    }
    uint32_t A = a0, B = b0, C = c0, D = d0;
    for (int i = 0; i < 64; ++i) {
      uint32_t F;
      int g;
      if (i < 16) { F = (B & C) | (~B & D); g = i; }
      else if (i < 32) { F = (D & B) | (~D & C); g = (5 * i + 1) % 16; }
      else if (i < 48) { F = B ^ C ^ D; g = (3 * i + 5) % 16; }
      else { F = C ^ (B | ~D); g = (7 * i) % 16; }
      F = F + A + K[i] + M[g];
      A = D; D = C; C = B;
      B = B + rotl(F, S[i]);
    }
    a0 += A; b0 += B; c0 += C; d0 += D;
  }
  const uint32_t h[4] = {a0, b0, c0, d0};
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) out[4 * i + j] = static_cast<uint8_t>((h[i] >> (8 * j)) & 0xff);
}
}  // namespace

uint64_t md5_low64(const std::string& s) {
  uint8_t d[16];
  md5(s, d);
  uint64_t v = 0;
  for (int i = 7; i >= 0; --i) v = (v << 8) | d[i];  // little-endian first 8 bytes
  return v;
}

// ---- agents --------------------------------------------------------------------------------------------------------
AgentState AgentState::make(const std::string& id, int slots, const std::string& label, int max_zero) {
  AgentState a;
  a.id = id; a.label = label; a.num_slots = slots; a.max_zero_slot_containers = max_zero;
  for (int i = 0; i < slots; ++i) a.devices[i] = "";
  return a;
}
std::vector<int> AgentState::allocate_free_devices(int slots, const std::string& container_id) {
  std::vector<int> got;
  if (slots == 0) { zero_slot_containers[container_id] = true; return got; }
  for (auto& d : devices) {
    if (static_cast<int>(got.size()) == slots) break;
    if (d.second.empty()) { d.second = container_id; got.push_back(d.first); }
  }
  if (static_cast<int>(got.size()) != slots) throw std::runtime_error("agent " + id + " has too few free devices");
  return got;
}
void AgentState::release_container(const std::string& container_id) {
  zero_slot_containers.erase(container_id);
  for (auto& d : devices) if (d.second == container_id) d.second.clear();
}

// ---- fitting ---------------------------------------------------------------------------------------------------------
static double fit_score(FittingPolicy policy, const AllocateRequest& req, const AgentState& a) {
  if (a.num_used_slots() != 0 || req.slots_needed != 0)
    return policy == FittingPolicy::Best ? 1.0 / (1.0 + a.num_empty_slots()) : static_cast<double>(a.num_empty_slots()) / std::max(a.num_slots, 1);
  if (a.max_zero_slot_containers == 0) return 0.0;
  const int free_zero = a.max_zero_slot_containers - a.num_zero_slot_containers();
  return policy == FittingPolicy::Best ? 1.0 / (1.0 + free_zero) : static_cast<double>(free_zero) / a.max_zero_slot_containers;
}
static bool fit_less(const Fit& a, const Fit& b) {
  if (a.score != b.score) return a.score > b.score;
  if (a.hash_distance != b.hash_distance) return a.hash_distance < b.hash_distance;
  return a.agent_id < b.agent_id;
}
static Fit candidate(const AllocateRequest& req, const AgentState& a, FittingPolicy policy) {
  Fit f;
  f.agent_id = a.id;
  f.score = fit_score(policy, req, a);
  f.hash_distance = md5_low64(req.id) - md5_low64(a.id);
  return f;
}

std::vector<Fit> find_fits(const AllocateRequest& req, const std::map<std::string, AgentState>& agents, FittingPolicy policy) {
  // 1. one agent that can hold the whole task
  std::vector<Fit> cands;
  for (auto& kv : agents) {
    const AgentState& a = kv.second;
    if (!a.enabled || req.label != a.label) continue;
    if (req.slots_needed > a.num_empty_slots()) continue;
    if (req.slots_needed == 0 && (a.max_zero_slot_containers == 0 || a.num_zero_slot_containers() >= a.max_zero_slot_containers)) continue;
    cands.push_back(candidate(req, a, policy));
  }
  if (!cands.empty()) {
    std::sort(cands.begin(), cands.end(), fit_less);
    cands[0].slots = req.slots_needed;
    return {cands[0]};
  }
  if (req.single_agent || req.slots_needed <= 1) return {};
  // 2. several completely idle agents of equal size whose slot count divides the request
  std::map<int, std::vector<const AgentState*>> by_slots;
  for (auto& kv : agents) {
    const AgentState& a = kv.second;
    if (!a.enabled || req.label != a.label || a.num_used_slots() != 0) continue;
    by_slots[a.num_empty_slots()].push_back(&a);
  }
  int chosen = 0;
  for (auto it = by_slots.rbegin(); it != by_slots.rend(); ++it) {
    const int n = it->first;
    if (n == 0 || req.slots_needed % n != 0) continue;
    if (static_cast<int>(it->second.size()) * n >= req.slots_needed) { chosen = n; break; }
  }
  if (chosen == 0) return {};
  std::vector<Fit> fits;
  for (auto* a : by_slots[chosen]) fits.push_back(candidate(req, *a, policy));
  std::sort(fits.begin(), fits.end(), fit_less);
  const int containers = req.slots_needed / chosen;
  fits.resize(static_cast<size_t>(containers));
  for (auto& f : fits) f.slots = req.slots_needed / containers;
  return fits;
}

void add_task_to_agents(std::map<std::string, AgentState>& agents, const std::vector<Fit>& fits, const std::string& container_prefix) {
  int i = 0;
  for (auto& f : fits) agents.at(f.agent_id).allocate_free_devices(f.slots, container_prefix + "#" + std::to_string(i++));
}

// ---- fair share ----------------------------------------------------------------------------------------------------------
namespace {
struct GroupState {
  const Group* group = nullptr;
  bool disabled = false;
  int slot_demand = 0, active_slots = 0, presubscribed = 0, offered = 0;
  std::vector<const AllocateRequest*> reqs, pending, allocated;
};

double total_weight(const std::vector<GroupState*>& states) {
  double t = 0;
  for (auto* s : states)
    if (!s->disabled && s->offered < s->slot_demand) t += s->group->weight;
  return t;
}

void allocate_slot_offers(std::vector<GroupState*>& states, int capacity) {
  std::map<GroupState*, int> preoffers;
  for (auto* s : states) {
    if (s->presubscribed == 0) continue;
    s->offered = s->presubscribed;
    preoffers[s] = s->presubscribed;
    capacity -= s->presubscribed;
  }
  std::sort(states.begin(), states.end(), [](GroupState* a, GroupState* b) {
    if (a->slot_demand != b->slot_demand) return a->slot_demand < b->slot_demand;
    return a->group->registered < b->group->registered;
  });
  std::vector<GroupState*> by_time = states;
  std::sort(by_time.begin(), by_time.end(), [](GroupState* a, GroupState* b) { return a->group->registered > b->group->registered; });
  double tw = total_weight(states);
  for (int left = static_cast<int>(states.size()); left > 0;) {
    bool progress = false;
    const int start_capacity = capacity;
    for (auto* s : states) {
      if (s->disabled || s->offered == s->slot_demand) continue;
      const int fair = std::max(1, static_cast<int>(static_cast<double>(start_capacity) * s->group->weight / tw));
      progress = true;
      int offer = std::min(std::min(fair, capacity), s->slot_demand - s->offered);
      int& pre = preoffers[s];
      if (pre > 0) {  // slots already granted to non-preemptible tasks count against the offer first
        if (pre == offer) { pre = 0; offer = 0; }
        else if (pre > offer) { pre -= offer; offer = 0; }
        else { pre = 0; }
      }
      s->offered += offer;
      capacity -= offer;
      if (s->offered == s->slot_demand) { --left; tw = total_weight(states); }
    }
    if (capacity == 0) {
      // deadlock breaker: the youngest group whose smallest pending task cannot fit in its offer gives it back
      bool adjusted = false;
      for (auto* s : by_time) {
        const AllocateRequest* smallest = nullptr;
        for (auto* r : s->pending) if (!smallest || r->slots_needed < smallest->slots_needed) smallest = r;
        if (!s->disabled && s->offered != s->slot_demand && smallest && smallest->slots_needed > s->offered) {
          capacity += s->offered;
          s->offered = 0;
          s->disabled = true;
          adjusted = true;
          --left;
          tw = total_weight(states);
          break;
        }
      }
      if (!adjusted) return;
    } else if (!progress) {
      return;
    }
  }
}
}  // namespace

ScheduleDecision fairshare_schedule(const std::vector<AllocateRequest>& tasks, const std::map<std::string, Group>& groups,
                                    const std::map<std::string, AgentState>& agents, FittingPolicy policy) {
  ScheduleDecision d;
  for (auto& req : tasks)  // zero-slot tasks are scheduled independently of the shares
    if (req.slots_needed == 0 && !req.allocated() && !find_fits(req, agents, policy).empty()) d.to_allocate.push_back(req.id);
  std::map<std::string, int> capacity;
  for (auto& kv : agents) if (kv.second.enabled) capacity[kv.second.label] += kv.second.num_slots;
  std::map<std::string, std::vector<GroupState*>> by_label;
  std::map<std::string, GroupState> storage;
  for (auto& req : tasks) {
    if (req.slots_needed == 0 || req.slots_needed > capacity[req.label]) continue;
    auto git = groups.find(req.group);
    if (git == groups.end()) throw std::runtime_error("task " + req.id + " references unknown group " + req.group);
    auto it = storage.find(req.group);
    if (it == storage.end()) {
      it = storage.emplace(req.group, GroupState{}).first;
      it->second.group = &git->second;
      by_label[req.label].push_back(&it->second);
    }
    it->second.reqs.push_back(&req);
  }
  for (auto& kv : by_label) {
    for (auto* s : kv.second) {
      for (auto* req : s->reqs) {
        s->slot_demand += req->slots_needed;
        if (!req->allocated()) s->pending.push_back(req);
        else {
          if (req->non_preemptible) s->presubscribed += req->slots_needed;
          s->allocated.push_back(req);
          s->active_slots += req->slots_needed;
        }
      }
      if (s->group->has_max_slots) s->slot_demand = std::min(s->slot_demand, s->group->max_slots);
    }
    allocate_slot_offers(kv.second, capacity[kv.first]);
    std::map<std::string, AgentState> local = agents;  // fits consume devices so one pass cannot double-book
    for (auto* s : kv.second) {
      if (s->active_slots > s->offered) {
        for (auto* req : s->allocated) {
          if (req->non_preemptible) continue;
          d.to_release.push_back(req->id);
          s->active_slots -= req->slots_needed;
          if (s->active_slots <= s->offered) break;
        }
      } else if (s->active_slots < s->offered) {
        s->offered -= s->active_slots;
        for (auto* req : s->pending) {
          if (req->slots_needed > s->offered) continue;
          auto fits = find_fits(*req, local, policy);
          if (fits.empty()) continue;
          add_task_to_agents(local, fits, req->id);
          d.to_allocate.push_back(req->id);
          s->offered -= req->slots_needed;
        }
      }
    }
  }
  return d;
}

// ---- priority ----------------------------------------------------------------------------------------------------------------
namespace {
void remove_task_from_agents(std::map<std::string, AgentState>& agents, const AllocateRequest& req) {
  for (auto& al : req.allocations) {
    auto it = agents.find(al.agent_id);
    if (it == agents.end()) continue;
    if (al.devices.empty() && al.slots == 0) it->second.zero_slot_containers.erase(al.container_id);
    for (int dev : al.devices) {
      auto d = it->second.devices.find(dev);
      if (d != it->second.devices.end()) d->second.clear();
    }
  }
}

void priority_with_filter(const std::vector<AllocateRequest>& tasks, const std::map<std::string, Group>& groups,
                          const std::map<std::string, AgentState>& agents_for_label, const std::string& label, FittingPolicy policy,
                          bool preemption, bool zero_slot, ScheduleDecision& out, std::set<std::string>& released) {
  auto filter = [&](const AllocateRequest& r) { return zero_slot ? r.slots_needed == 0 : r.slots_needed > 0; };
  std::map<int, std::vector<const AllocateRequest*>> pending, scheduled;
  for (auto& req : tasks) {
    if (req.label != label || !filter(req)) continue;
    auto g = groups.find(req.group);
    const int prio = (g != groups.end() && g->second.has_priority) ? g->second.priority : kDefaultSchedulingPriority;
    (req.allocated() ? scheduled : pending)[prio].push_back(&req);
  }
  for (auto& kv : pending) std::stable_sort(kv.second.begin(), kv.second.end(), [](auto* a, auto* b) { return a->registered < b->registered; });
  for (auto& kv : scheduled) std::stable_sort(kv.second.begin(), kv.second.end(), [](auto* a, auto* b) { return a->registered > b->registered; });
  std::map<std::string, AgentState> local = agents_for_label;
  bool start_tasks = true;
  for (auto& kv : pending) {  // ascending priority number = most important first
    const int prio = kv.first;
    std::vector<const AllocateRequest*> failed;
    for (auto* req : kv.second) {
      auto fits = find_fits(*req, local, policy);
      if (fits.empty()) { failed.push_back(req); continue; }
      add_task_to_agents(local, fits, req->id);
      if (start_tasks) out.to_allocate.push_back(req->id);
    }
    if (failed.empty()) continue;
    start_tasks = false;
    if (!preemption) break;
    for (auto* want : failed) {
      auto fits = find_fits(*want, local, policy);
      if (!fits.empty()) { add_task_to_agents(local, fits, want->id); continue; }  // will fit once earlier preemptions land
      std::map<std::string, AgentState> trial_state = local;
      std::vector<std::string> victims;
      bool placed = false;
      for (int p = kMaxUserSchedulingPriority; p > prio && !placed; --p) {
        auto sit = scheduled.find(p);
        if (sit == scheduled.end()) continue;
        for (auto* cand : sit->second) {
          if (cand->non_preemptible || !filter(*cand) || released.count(cand->id)) continue;
          remove_task_from_agents(trial_state, *cand);
          victims.push_back(cand->id);
          auto f2 = find_fits(*want, trial_state, policy);
          if (!f2.empty()) { add_task_to_agents(trial_state, f2, want->id); placed = true; break; }
        }
      }
      if (placed) {
        local = trial_state;
        for (auto& v : victims) released.insert(v);
      }
    }
  }
}
}  // namespace

ScheduleDecision priority_schedule(const std::vector<AllocateRequest>& tasks, const std::map<std::string, Group>& groups,
                                   const std::map<std::string, AgentState>& agents, FittingPolicy policy, bool preemption) {
  ScheduleDecision d;
  std::map<std::string, std::map<std::string, AgentState>> by_label;
  for (auto& kv : agents) by_label[kv.second.label][kv.first] = kv.second;
  for (auto& kv : by_label) {
    std::set<std::string> released;
    priority_with_filter(tasks, groups, kv.second, kv.first, policy, preemption, true, d, released);
    priority_with_filter(tasks, groups, kv.second, kv.first, policy, preemption, false, d, released);
    for (auto& r : released) d.to_release.push_back(r);
  }
  return d;
}

// ---- round robin -----------------------------------------------------------------------------------------------------------------
ScheduleDecision round_robin_schedule(const std::vector<AllocateRequest>& tasks, const std::map<std::string, Group>& groups,
                                      const std::map<std::string, AgentState>& agents, FittingPolicy policy) {
  struct RR { const Group* g; int active = 0; std::vector<const AllocateRequest*> pending; };
  std::map<std::string, RR> by_group;
  std::vector<std::string> order;
  for (auto& req : tasks) {
    auto git = groups.find(req.group);
    if (git == groups.end()) throw std::runtime_error("task " + req.id + " references unknown group " + req.group);
    if (!by_group.count(req.group)) { by_group[req.group].g = &git->second; order.push_back(req.group); }
    if (req.allocated()) by_group[req.group].active += req.slots_needed; else by_group[req.group].pending.push_back(&req);
  }
  std::vector<RR*> states;
  for (auto& id : order) states.push_back(&by_group[id]);
  std::stable_sort(states.begin(), states.end(), [](RR* a, RR* b) {
    if (a->active != b->active) return a->active < b->active;
    return a->g->registered < b->g->registered;
  });
  ScheduleDecision d;
  std::map<std::string, AgentState> local = agents;
  while (!states.empty()) {
    std::vector<RR*> next;
    for (auto* s : states) {
      if (s->pending.empty()) continue;
      const AllocateRequest* req = s->pending.front();
      auto fits = find_fits(*req, local, policy);
      if (fits.empty()) continue;  // group drops out of this pass
      add_task_to_agents(local, fits, req->id);
      d.to_allocate.push_back(req->id);
      s->pending.erase(s->pending.begin());
      next.push_back(s);
    }
    states.swap(next);
  }
  return d;
}

// ---- JSON front door -------------------------------------------------------------------------------------------------------------------
Json schedule_json(const Json& in) {
  std::map<std::string, AgentState> agents;
  for (auto& a : in["agents"].arr()) {
    AgentState s = AgentState::make(a["id"].as_string(), static_cast<int>(a["slots"].as_int()), a["label"].str_or(""),
                                    static_cast<int>(a["max_zero_slot_containers"].as_int(100)));
    s.enabled = a["enabled"].is_null() ? true : a["enabled"].as_bool(true);
    agents[s.id] = s;
  }
  std::map<std::string, Group> groups;
  long long stamp = 0;
  for (auto& g : in["groups"].arr()) {
    Group gr;
    gr.id = g["id"].as_string();
    if (!g["max_slots"].is_null()) { gr.has_max_slots = true; gr.max_slots = static_cast<int>(g["max_slots"].as_int()); }
    gr.weight = g["weight"].as_double(1.0);
    if (!g["priority"].is_null()) { gr.has_priority = true; gr.priority = static_cast<int>(g["priority"].as_int()); }
    gr.registered = g["registered"].is_null() ? stamp : g["registered"].as_int();
    ++stamp;
    groups[gr.id] = gr;
  }
  std::vector<AllocateRequest> tasks;
  stamp = 0;
  for (auto& t : in["tasks"].arr()) {
    AllocateRequest r;
    r.id = t["id"].as_string();
    r.name = t["name"].str_or(r.id);
    r.group = t["group"].str_or(r.id);
    if (!groups.count(r.group)) { Group g; g.id = r.group; g.registered = 1000000 + stamp; groups[r.group] = g; }
    r.slots_needed = static_cast<int>(t["slots_needed"].as_int());
    r.label = t["label"].str_or("");
    r.non_preemptible = t["non_preemptible"].as_bool(false);
    r.single_agent = t["single_agent"].as_bool(false);
    r.registered = t["registered"].is_null() ? stamp : t["registered"].as_int();
    ++stamp;
    for (auto& al : t["allocations"].arr()) {
      Allocation a;
      a.agent_id = al["agent"].as_string();
      a.slots = static_cast<int>(al["slots"].as_int());
      a.container_id = al["container_id"].str_or(r.id + "#" + std::to_string(r.allocations.size()));
      auto ag = agents.find(a.agent_id);
      if (ag == agents.end()) throw std::runtime_error("allocation references unknown agent " + a.agent_id);
      a.devices = ag->second.allocate_free_devices(a.slots, a.container_id);
      r.allocations.push_back(a);
    }
    tasks.push_back(r);
  }
  const FittingPolicy policy = in["fitting_policy"].str_or("best") == "worst" ? FittingPolicy::Worst : FittingPolicy::Best;
  const std::string kind = in["scheduler"].str_or("fair_share");
  ScheduleDecision d;
  if (kind == "fair_share") d = fairshare_schedule(tasks, groups, agents, policy);
  else if (kind == "priority") d = priority_schedule(tasks, groups, agents, policy, in["preemption"].as_bool(false));
  else if (kind == "round_robin") d = round_robin_schedule(tasks, groups, agents, policy);
  else throw std::runtime_error("unknown scheduler '" + kind + "'");
  Json out = Json::object();
  out["to_allocate"] = Json::array();
  out["to_release"] = Json::array();
  out["fits"] = Json::object();
  std::map<std::string, AgentState> local = agents;
  for (auto& id : d.to_allocate) {
    out["to_allocate"].push_back(id);
    for (auto& t : tasks) {
      if (t.id != id) continue;
      auto fits = find_fits(t, local, policy);
      Json fj = Json::array();
      for (auto& f : fits) { Json o = Json::object(); o["agent"] = f.agent_id; o["slots"] = f.slots; fj.push_back(o); }
      if (!fits.empty()) add_task_to_agents(local, fits, id);
      out["fits"][id] = fj;
    }
  }
  for (auto& id : d.to_release) out["to_release"].push_back(id);
  return out;
}

}  // namespace det
